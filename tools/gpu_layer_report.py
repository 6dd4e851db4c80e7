"""Diagnostic (run on the GPU box): per-layer max relative error of the CUDA U-Net vs the CPU oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepmimic_diffusion_mujoco_b200 as dm
from oracle import diffusion_oracle as O


def report(B=3, H=64, seed=0, verbose=True):
    sd = O.seeded_state_dict(H=H, seed=seed)
    net = dm.TemporalUnet(H, 35, 35)
    net.load_state_dict(sd)
    g = torch.Generator().manual_seed(11)
    x = torch.randn(B, H, 35, generator=g)
    t = torch.randint(0, 100, (B,), generator=g)
    rec = {}
    y_ref = O.unet_forward(sd, x, t, record=rec)
    y = net(x.cuda(), {}, t.cuda())
    torch.cuda.synchronize()
    plan = net.plan_for(B, H)
    worst = 0.0
    for i, (name, C, L) in enumerate(plan.layer_names()):
        if C == 0:
            continue
        got = plan.read_layer(i, C, L).cpu()
        ref = rec[name]
        err = ((got - ref).abs().max() / ref.abs().max()).item()
        worst = max(worst, err)
        if verbose:
            print(f"{i:3d} {name:32s} C={C:4d} L={L:4d} relerr={err:.3e}", flush=True)
    err = ((y.cpu() - y_ref).abs().max() / y_ref.abs().max()).item()
    if verbose:
        print(f"final output relerr={err:.3e}  launches={plan.last_launch_count()}")
    return max(worst, err)


if __name__ == "__main__":
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    H = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    print("worst", report(B, H))
