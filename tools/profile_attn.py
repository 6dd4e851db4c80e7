"""One attention=True U-Net forward at B=1024, H=64 (for an ncu launch list of the attention kernels)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepmimic_diffusion_mujoco_b200 as dm
from deepmimic_diffusion_mujoco_b200 import synthetic as O
B, H = 1024, 64
net = dm.TemporalUnet(H, 35, 35, attention=True)
net.load_state_dict(O.seeded_state_dict(H=H, seed=0, attention=True))
x = torch.randn(B, H, 35, device="cuda"); t = torch.full((B,), 50.0, device="cuda")
for _ in range(2):
    y = net(x, {}, t)
torch.cuda.synchronize()
print("ok", float(y.abs().mean()))
