"""GPU-box timing of the other BASELINE.json configs at their per-GPU sizes and of the widened rows (attention=True,
DiffusionV4 on the dim=128 U-Net).  One JSON line per variant: one full sampling pass, CUDA events, median of 3."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepmimic_diffusion_mujoco_b200 as dm
from deepmimic_diffusion_mujoco_b200 import synthetic as O


def box(x):
    x[:, :, 13:16] = 0.0; x[:, :, 16] = 1.57; x[:, :, 17:20] = 0.0; x[:, :, 20] = 1.57
    return x


def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def gd(name, B, H, attention=False, cond=None, steps=100):
    net = dm.TemporalUnet(H, 35, 35, attention=attention)
    net.load_state_dict(O.seeded_state_dict(H=H, seed=0, attention=attention))
    diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=steps, predict_epsilon=False).cuda()
    kw = {"conditioning_fn": cond} if cond else {}
    ms = timed(lambda: diff.p_sample_loop((B, H, 35), {}, verbose=False, **kw))
    print(json.dumps({"variant": name, "B": B, "H": H, "steps": steps, "ms_per_pass": round(ms, 2),
                      "frames_per_s": round(B * H / ms * 1e3), "launches": diff.last_launch_count}), flush=True)


which = sys.argv[1:] or ["c2", "c3", "c4", "c5", "attn", "v4"]
if "c2" in which: gd("configs[1] B=1024 H=64 box inpaint", 1024, 64, cond=box)
if "c3" in which: gd("configs[2] per-GPU share of 8: B=512 H=256 arm limit", 512, 256, cond=box)
if "c4" in which: gd("configs[3] per-GPU share of 8: B=2048 H=128", 2048, 128)
if "c5" in which: gd("configs[4] per-GPU share of 8: B=8192 H=64", 8192, 64)
if "attn" in which: gd("attention=True B=1024 H=64", 1024, 64, attention=True, cond=box)
if "unfused" in which:
    net = dm.TemporalUnet(64, 35, 35); net.load_state_dict(O.seeded_state_dict(H=64, seed=0))
    diff = dm.GaussianDiffusion(net, 64, 35, n_timesteps=100, predict_epsilon=False).cuda()
    ms = timed(lambda: diff.p_sample_loop((1024, 64, 35), {}, verbose=False, fused=False))
    print(json.dumps({"variant": "configs[1] per-layer launches (fused=False)", "ms_per_pass": round(ms, 2), "launches": diff.last_launch_count}), flush=True)
if "v4" in which:
    for dim, T, B in ((128, 35, 1024), (128, 69, 1024), (32, 35, 1024)):
        H = 32
        net = dm.TemporalUnet(H, T, T, dim=dim)
        net.load_state_dict(O.seeded_state_dict(H=H, T=T, dim=dim, seed=0))
        net.cuda()
        d = dm.DiffusionV4(100, 1e-4, 0.02, T, H, device="cuda", predict_x0=True, schedule_type="cosine")
        ms = timed(lambda: d.sample(net, B))
        print(json.dumps({"variant": f"DiffusionV4.sample predict_x0 cosine, dim={dim} T={T}", "B": B, "H": H, "steps": 99,
                          "ms_per_pass": round(ms, 2), "frames_per_s": round(B * H / ms * 1e3), "launches": d.last_launch_count}), flush=True)
