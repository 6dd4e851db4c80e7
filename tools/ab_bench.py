"""Developer A/B (GPU box): time the fused sampling loop for several builds of libdm_b200.so in ONE gpurun call,
interleaved, so box-to-box clock variance cancels.  usage: python tools/ab_bench.py libA.so libB.so [B] [H] [reps]"""
import os, subprocess, sys, json
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch
    import deepmimic_diffusion_mujoco_b200 as dm
    from deepmimic_diffusion_mujoco_b200 import synthetic as O
    B, H, reps = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=0))
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=100, predict_epsilon=False).cuda()
    for _ in range(2):
        diff.p_sample_loop((B, H, 35), {}, verbose=False, seed=1)
    ts = []
    for r in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        out = diff.p_sample_loop((B, H, 35), {}, verbose=False, seed=2 + r)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    print(json.dumps({"ms": sorted(ts)[len(ts) // 2], "min": min(ts), "checksum": float(out.trajectories.double().sum())}))
    sys.exit(0)
libs = [a for a in sys.argv[1:] if ".so" in a]          # "lib.so" or "lib.so@ENVVAR" (sets ENVVAR=1 for that run)
rest = [a for a in sys.argv[1:] if ".so" not in a]
B = rest[0] if len(rest) > 0 else "1024"; H = rest[1] if len(rest) > 1 else "64"; reps = rest[2] if len(rest) > 2 else "5"
for rnd in range(2):
    for l in libs:
        path, _, flag = l.partition("@")
        env = dict(os.environ, DM_B200_LIB=os.path.abspath(path))
        if flag:
            env[flag] = "1"
        r = subprocess.run([sys.executable, __file__, "--one", B, H, reps], env=env, capture_output=True, text=True)
        print(rnd, l, r.stdout.strip()[-300:], r.stderr.strip()[-300:] if r.returncode else "", flush=True)
