"""Diagnostic (GPU box): per-layer phase timeline of the fused sampling-loop kernel for one cluster / one step."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepmimic_diffusion_mujoco_b200 as dm
from deepmimic_diffusion_mujoco_b200 import _lib
from deepmimic_diffusion_mujoco_b200 import synthetic as O

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
H = int(sys.argv[2]) if len(sys.argv) > 2 else 64
steps = 6
ATTN = bool(os.environ.get("TRACE_ATTN"))
net = dm.TemporalUnet(H, 35, 35, attention=ATTN)
net.load_state_dict(O.seeded_state_dict(H=H, seed=0, attention=ATTN))
diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=steps, predict_epsilon=False).cuda()
plan = net.plan_for(B, H)
names = plan.layer_names()
tr = torch.zeros(len(names), 64, dtype=torch.int64, device="cuda")
_lib.check(_lib.lib().dm_plan_set_trace(plan.handle, tr.data_ptr(), 3, 4))
for _ in range(2):
    diff.p_sample_loop((B, H, 35), {}, verbose=False, seed=1)
torch.cuda.synchronize()
t = tr.cpu().numpy().astype("float64")
t0 = t[0, 0]
ev = ["start", "jobcp", "prod0", "A_iss", "prodE", "A_arr", "W_arr", "mmaIs", "epi0", "accF", "pass1", "pass2", "unitE", "sync", "p2loop", "p2bar"]
print("layer".ljust(28), " ".join(e.rjust(6) for e in ev), "  total_us")
tot = 0
for i, (n, C, L) in enumerate(names):
    r = t[i]
    base = r[0]
    vals = [(r[k] - base) / 1e3 if r[k] > 0 else float("nan") for k in range(16)]
    print(f"{n[:27]:28s}", " ".join(f"{v:6.1f}" for v in vals), f"  {(r[13]-r[0])/1e3:7.1f}")
    tot += r[13] - r[0]
if os.environ.get("TRACE_START"):
    print("layer start (us): mma warp at loop top | before run_unit | mma warp enters | descriptors ready | first A seen | first W seen | epilogue params loaded")
    for i, (n, C, L) in enumerate(names):
        r = t[i]
        print(f"{n[:27]:28s} " + " ".join(f"{(r[k]-r[0])/1e3:5.2f}" if r[k] > 0 else "  nan" for k in (22, 23, 20, 21, 5, 6, 8)))
if os.environ.get("TRACE_PASS1"):
    print("pass 1 split (us after accF barrier): own partials | +barrier | combine+barrier | table+barrier")
    for i, (n, C, L) in enumerate(names):
        r = t[i]
        if r[40] > 0:
            print(f"{n[:27]:28s} {(r[40]-r[9])/1e3:5.2f} {(r[41]-r[9])/1e3:5.2f} {(r[42]-r[9])/1e3:5.2f} {(r[10]-r[9])/1e3:5.2f}")
detail = os.environ.get("TRACE_DETAIL")
if detail:
    for i, (n, C, L) in enumerate(names):
        if any(d in n for d in detail.split(",")):
            r = t[i]
            print(n, "A chunks seen by MMA warp (us):", "  ".join(f"{(r[48+c]-r[0])/1e3:.2f}" for c in range(8) if r[48+c] > 0))
            print(n, "W loads: (issue, seen by MMA warp) us from layer start")
            print("  " + "  ".join(f"{(r[16+2*w]-r[0])/1e3:.2f}/{(r[17+2*w]-r[0])/1e3:.2f}" for w in range(24) if r[17+2*w] > 0))
print("effective SM clock MHz over the traced step:", (t[-1, 1] - t[0, 1]) / ((t[-1, 0] - t[0, 0]) / 1e3))
print("sum of layers us", tot / 1e3, " step span us", (t[-1, 13] - t[0, 0]) / 1e3)
if os.environ.get("TRACE_MMA_CLOCKS"):
    # DM_TRACE_MMA build: slots 17+2i / 16+2i hold clock64 when issuer warp A passed tap i's weight wait / finished issuing it
    for i, (n, C, L) in enumerate(names):
        if any(d in n for d in os.environ["TRACE_MMA_CLOCKS"].split(",")):
            r = tr.cpu().numpy()[i].astype("int64")
            seen = [int(r[17 + 2 * w]) for w in range(20) if r[17 + 2 * w] > 0]
            done = [int(r[16 + 2 * w]) for w in range(20) if r[16 + 2 * w] > 0]
            m = min(len(seen), len(done))
            print(n, "cycles: [wait passed -> issue loop done | issue done -> next wait passed]")
            print("  " + "  ".join(f"{done[w]-seen[w]}|{(seen[w+1]-done[w]) if w + 1 < m else 0}" for w in range(m)))
