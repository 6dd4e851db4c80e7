"""Profiling driver: a few denoising steps at the bench shape, launched kernel by kernel (no graph)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepmimic_diffusion_mujoco_b200 as dm
from deepmimic_diffusion_mujoco_b200 import synthetic as O

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
H = int(sys.argv[2]) if len(sys.argv) > 2 else 64
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3


def box(x):
    x[:, :, 13:16] = 0.0
    x[:, :, 16] = 1.57
    x[:, :, 17:20] = 0.0
    x[:, :, 20] = 1.57
    return x


net = dm.TemporalUnet(H, 35, 35)
net.load_state_dict(O.seeded_state_dict(H=H, seed=0))
diff = dm.GaussianDiffusion(net, H, 35, n_timesteps=steps, predict_epsilon=False).cuda()
for _ in range(2):
    out = diff.p_sample_loop((B, H, 35), {}, verbose=False, conditioning_fn=box, fused=(os.environ.get('DM_FUSED','1')=='1'), seed=1)
torch.cuda.synchronize()
print("ok", out.trajectories.abs().mean().item(), diff.last_launch_count)
