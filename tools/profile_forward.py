"""Developer aid (GPU box): kernel durations of ONE per-layer-launch forward (torch profiler / CUPTI), grouped by kernel name in
launch order.  usage: python tools/profile_forward.py [B] [H] [attention 0|1]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import deepmimic_diffusion_mujoco_b200 as dm
from deepmimic_diffusion_mujoco_b200 import synthetic as O

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
H = int(sys.argv[2]) if len(sys.argv) > 2 else 64
attn = bool(int(sys.argv[3])) if len(sys.argv) > 3 else True
net = dm.TemporalUnet(H, 35, 35, attention=attn)
net.load_state_dict(O.seeded_state_dict(H=H, seed=0, attention=attn))
x = torch.randn(B, H, 35, device="cuda")
t = torch.randint(0, 100, (B,), device="cuda")
for _ in range(3):
    net(x, {}, t)
torch.cuda.synchronize()
names = [n for n, C, L in net.plan_for(B, H).layer_names()]
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    net(x, {}, t)
    torch.cuda.synchronize()
evs = sorted([e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA], key=lambda e: e.time_range.start)
ker = [e for e in evs if "conv_gemm" in e.name or "attention_stage" in e.name]
for n, e in zip(names, ker):
    print(f"{n:28s} {e.name[:40]:40s} {e.device_time:8.1f} us")
print("total us", sum(e.device_time for e in ker))
