"""Aggregate an ncu report's warp-stall samples by CUDA source line (reads the .ncu-rep with the ncu CLI).
usage: python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top_n]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
agg = []; cur = None; hdr = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0] == "Line No": hdr = r; continue
    if len(r) > 8 and r[2] == "-" and r[0].isdigit():
        try: agg.append((int(r[4]), cur, int(r[0]), r[1][:90], int(r[7]), r))
        except ValueError: pass
tot = sum(a[0] for a in agg)
print("total samples", tot)
names = [h for h in hdr] if hdr else []
st = [i for i, h in enumerate(names) if h.startswith("stall_") and "Not Issued" not in h]
for a in sorted(agg, key=lambda t: -t[0])[:top]:
    r = a[5]
    top3 = sorted(((int(r[i]) if r[i].isdigit() else 0, names[i][6:]) for i in st), reverse=True)[:3]
    print(f"{a[0]:7d} {100*a[0]/tot:5.1f}% inst={a[4]:>10d} {a[1]}:{a[2]:<4d} {a[3][:70]:70s} {top3}")
