"""GPU-box measurement of the scoring-side kernels against their roofline: K4 (DDPM update, standalone
launch), K5 (FK + tracking reward) and K6 (FID moments) at the configs[4] per-GPU size (8192 samples, H=64).
Prints one JSON line per kernel: algorithmic bytes / CUDA-event time vs MEASURED_PEAKS.json hbm_gbs."""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepmimic_diffusion_mujoco_b200 as dm
from deepmimic_diffusion_mujoco_b200 import _lib, scoring
from deepmimic_diffusion_mujoco_b200 import synthetic as O

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
H, T = 64, 35
peak = 6540.5
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, reps=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


x = torch.randn(N, H, T, device="cuda") * 0.3
x[:, :, 2] += 0.9
x[:, :, 3] += 1.0
ref = x[0].clone()
frames = N * H
# K5 reward only / with body positions
ms = timed(lambda: scoring.fk_reward(x, ref, 1 / 60))
b = frames * (140 + 4)
print(json.dumps({"kernel": "fk_reward_kernel (reward only)", "frames": frames, "ms": ms, "GBps": b / ms / 1e6, "frac_hbm": b / ms / 1e6 / peak}))
ms = timed(lambda: scoring.fk_reward(x, ref, 1 / 60, return_body_pos=True))
b = frames * (140 + 4 + 180)
print(json.dumps({"kernel": "fk_reward_kernel (+body positions)", "frames": frames, "ms": ms, "GBps": b / ms / 1e6, "frac_hbm": b / ms / 1e6 / peak}))
# K6
s = torch.zeros(H * T, device="cuda", dtype=torch.float64); sxx = torch.zeros(H * T, H * T, device="cuda", dtype=torch.float64)
ms = timed(lambda: scoring.fid_moments(x, s, sxx), reps=5)
D = H * T
print(json.dumps({"kernel": "fid_moments_kernel", "N": N, "D": D, "ms": ms, "TFLOPs_algorithmic": 2.0 * N * D * D / ms / 1e9,
                  "GBps_input": 4.0 * N * D / ms / 1e6}))
# K4 standalone
net = dm.TemporalUnet(H, T, T); net.load_state_dict(O.seeded_state_dict(H=H, seed=0))
diff = dm.GaussianDiffusion(net, H, T, n_timesteps=100, predict_epsilon=False).cuda()
nul = ctypes.POINTER(_lib.DmCond)()
for B in (1024, 16384):   # configs[1] batch (9 MB working set: launch-latency sized) and a batch far larger than L2
    plan = net.plan_for(B, H); plan.set_schedule(diff._schedule_bufs(), diff._schedule_id())
    xk = torch.randn(B, H, T, device="cuda"); mo = torch.randn(B, H, T, device="cuda"); z = torch.randn(B, H, T, device="cuda")
    for name, zp, bytes_per in [("ddpm_step_kernel (Philox noise)", None, 12 * T + 4 * 64), ("ddpm_step_kernel (injected noise)", z.data_ptr(), 16 * T + 4 * 64)]:
        ms = timed(lambda: _lib.check(_lib.lib().dm_ddpm_step(plan.handle, xk.data_ptr(), mo.data_ptr(), zp, 1, 50, 0, 0, nul, _lib.stream_ptr())))
        bb = B * H * bytes_per
        print(json.dumps({"kernel": name, "B": B, "frames": B * H, "ms": ms, "GBps": bb / ms / 1e6, "frac_hbm": bb / ms / 1e6 / peak,
                          "note": "bytes = x read+write, model_out (+noise) read, plus the 64-channel fp16 hi/lo operand re-pack"}))
    del plan, xk, mo, z
# SiFID window statistics (eval.py:59-77), HBM-bound on the [N, 350, 350] covariance output
Nw = 2048
xw = torch.randn(Nw, 64, T, device="cuda")
ms = timed(lambda: scoring.window_moments(xw, 10))
bw = Nw * (64 * T * 4 + 350 * 4 + 350 * 350 * 4)
print(json.dumps({"kernel": "window_moments_kernel", "N": Nw, "ms": ms, "GBps": bw / ms / 1e6, "frac_hbm": bw / ms / 1e6 / peak}))
