"""Generates tests/golden/*.npz by IMPORTING THE UNMODIFIED REFERENCE from /root/reference/diffusion.
Run in the build container only (the reference does not exist on the GPU box):
    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py
The fixtures pin oracle/diffusion_oracle.py (tests/test_oracle_golden.py), which in turn checks the CUDA path.
"""
import contextlib
import io
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference/diffusion")
sys.path.insert(0, ROOT)

with contextlib.redirect_stdout(io.StringIO()):
    from diffuser.models.diffusion import GaussianDiffusion  # noqa: E402
    from diffuser.models.temporal import TemporalUnet  # noqa: E402
from oracle.diffusion_oracle import seeded_state_dict  # noqa: E402


class InjectedNoise:
    """sample_fn plug point (diffusion.py:193,214): default_sample_fn with the randn_like draw replaced by a
    pre-generated tensor, so CUDA and reference consume identical noise."""

    def __init__(self, stack):
        self.stack, self.k = stack, 1

    def __call__(self, model, x, cond, t):
        mean, _, logv = model.p_mean_variance(x=x, cond=cond, t=t)
        noise = self.stack[self.k].clone()
        self.k += 1
        noise[t == 0] = 0
        return mean + torch.exp(0.5 * logv) * noise, torch.zeros(len(x))


def box_holding(x):  # sampling_config.py:23-32
    x[:, :, 13:16] = 0.0
    x[:, :, 16] = 1.57
    x[:, :, 17:20] = 0.0
    x[:, :, 20] = 1.57
    return x


def run(name, B, H, n_steps, pe, clip, cond=None, start=False, max_t=None, seed=0, attention=False):
    with contextlib.redirect_stdout(io.StringIO()):
        net = TemporalUnet(H, 35, 35, attention=attention)
    sd = seeded_state_dict(H=H, seed=seed, attention=attention)
    net.load_state_dict(sd)
    diff = GaussianDiffusion(net, H, 35, n_timesteps=n_steps, predict_epsilon=pe, clip_denoised=clip).eval()
    g = torch.Generator().manual_seed(1234 + seed)
    steps = n_steps if max_t is None else max_t
    stack = torch.randn(steps + 1, B, H, 35, generator=g)
    kw = {}
    if cond is not None:
        kw["conditioning_fn"] = cond
    if start:
        kw["starting_motion"] = stack[0].clone() * 0.3 + 0.1
    if max_t is not None:
        kw["max_timesteps"] = max_t
    torch.manual_seed(99)
    if not start:
        # the reference draws torch.randn(shape) itself (diffusion.py:205): make that draw equal stack[0]
        torch.manual_seed(4321)
        stack[0] = torch.randn(B, H, 35)
        torch.manual_seed(4321)
    out = diff.p_sample_loop((B, H, 35), {}, verbose=False, return_chain=True, sample_fn=InjectedNoise(stack), **kw)
    chain = out.chains.numpy()
    # one plain forward for the U-Net alone, non-uniform t
    xf = torch.randn(B, H, 35, generator=g)
    tf = torch.randint(0, n_steps, (B,), generator=g)
    with torch.no_grad():
        yf = net(xf, {}, tf)
    ck = float(sum(v.double().sum() for v in sd.values()))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), chain=chain, noise=stack.numpy(), fwd_x=xf.numpy(),
                        fwd_t=tf.numpy(), fwd_y=yf.numpy(), weight_checksum=ck, attention=np.array(int(attention)),
                        start=(kw["starting_motion"].numpy() if start else np.zeros(0, np.float32)),
                        meta=np.array([B, H, n_steps, int(pe), int(clip), int(cond is not None), int(start),
                                       -1 if max_t is None else max_t, seed]))
    print(name, chain.shape, "checksum", ck)


def v4_fixture():
    """DiffusionV4.sample (diffuser/models/diffusion_v4.py:82-149) run unmodified on CPU around the reference
    TemporalUnet (an adaptor maps the sampler's model(x, t, y) call onto forward(x, cond, time)); the noise it drew
    is recovered by re-seeding."""
    with contextlib.redirect_stdout(io.StringIO()):
        from diffuser.models.diffusion_v4 import DiffusionV4
    out = {}
    for tag, px0, sched, seed in (("x0_cosine", True, "cosine", 11), ("eps_linear", False, "linear", 12)):
        B, H, T = 2, 16, 10
        with contextlib.redirect_stdout(io.StringIO()):
            net = TemporalUnet(H, 35, 35)
        net.load_state_dict(seeded_state_dict(H=H, seed=seed))

        class Adaptor(torch.nn.Module):
            def forward(self, x, t, y):
                return net(x, {}, t)

        d = DiffusionV4(T, 1e-4, 0.02, 35, H, device="cpu", predict_x0=px0, schedule_type=sched)
        torch.manual_seed(555 + seed)
        x = d.sample(Adaptor(), B)
        torch.manual_seed(555 + seed)
        stack = [torch.randn(B, H, 35)]
        for i in reversed(range(1, T)):
            stack.append(torch.randn(B, H, 35) if i > 1 else torch.zeros(B, H, 35))
        out[tag + "_x"] = x.numpy()
        out[tag + "_noise"] = torch.stack(stack).numpy()
        out[tag + "_beta"] = d.beta.numpy()
        out[tag + "_meta"] = np.array([B, H, T, int(px0), seed])
    np.savez_compressed(os.path.join(HERE, "v4_reference.npz"), **out)
    print("v4_reference", {k: v.shape for k, v in out.items()})


def eval_metrics_fixture():
    """eval.py:15-89 metric functions.  eval.py itself cannot be imported (it needs diffuser.models.qna, which the
    reference does not ship), so the function definitions are compiled from its source, unmodified, at generation time."""
    import ast
    from scipy import linalg
    with contextlib.redirect_stdout(io.StringIO()):
        from metrics.fid_score import MotionFID
    src = open("/root/reference/diffusion/eval.py").read()
    want = {"slice_motion_sample", "CustomEvalWrapper", "calc_inter_diversity", "calc_sifid", "calc_intra_diversity"}
    body = [n for n in ast.parse(src).body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in want]
    ns = {"np": np, "linalg": linalg, "torch": torch, "MotionFID": MotionFID, "device": torch.device("cpu")}
    exec(compile(ast.Module(body=body, type_ignores=[]), "eval.py", "exec"), ns)
    g = torch.Generator().manual_seed(21)
    N, H, F, w = 6, 40, 35, 10
    gt = torch.randn(H, F, generator=g) * 0.4 + torch.linspace(0, 1, H)[:, None]
    gen = gt[None] + 0.3 * torch.randn(N, H, F, generator=g)
    wrap = ns["CustomEvalWrapper"](torch.nn.Identity())
    sifid = ns["calc_sifid"](wrap, gen, gt, window_size=w)
    inter = ns["calc_inter_diversity"](wrap, gen)
    np.random.seed(5)
    intra = ns["calc_intra_diversity"](wrap, gen, window_size=w)
    np.random.seed(5)
    offsets = np.stack([np.random.randint(H - w, size=2) for _ in range(N)])
    slices = ns["slice_motion_sample"](gt, w).numpy()
    np.savez_compressed(os.path.join(HERE, "eval_metrics_reference.npz"), gt=gt.numpy(), gen=gen.numpy(), sifid=np.array(sifid),
                        inter=np.array(inter), intra=np.array(intra), offsets=offsets, gt_slices=slices)
    print("eval_metrics_reference", sifid, inter, intra, slices.shape)


def fid_fixture():
    """metrics/fid_score.py:24-81 run unmodified on CPU: statistics of two seeded sample sets and their FID."""
    with contextlib.redirect_stdout(io.StringIO()):
        from metrics.fid_score import MotionFID
    g = torch.Generator().manual_seed(77)
    N, H = 600, 6
    a = torch.randn(N, H, 35, generator=g) * 0.3 + torch.randn(1, H, 35, generator=g)
    b = a[: N // 2] * 1.1 + 0.05 + 0.02 * torch.randn(N // 2, H, 35, generator=g)
    fid = MotionFID(None, None, device="cpu")
    mu1, s1 = fid.calculate_activation_statistics(a)
    mu2, s2 = fid.calculate_activation_statistics(b)
    val = fid.calculate_fid(mu1, s1, mu2, s2)
    half = fid.calculate_fid(*fid.calculate_activation_statistics(a[::2]), *fid.calculate_activation_statistics(a[1::2]))
    np.savez_compressed(os.path.join(HERE, "fid_reference.npz"), a=a.numpy(), b=b.numpy(), mu1=mu1, sigma1=s1, mu2=mu2,
                        sigma2=s2, fid=np.array(val), fid_halves=np.array(half))
    print("fid_reference", val, half)


if __name__ == "__main__":
    run("unet_chain_plain", B=2, H=64, n_steps=8, pe=False, clip=False)
    run("unet_chain_box_eps_clip", B=3, H=16, n_steps=12, pe=True, clip=True, cond=box_holding, seed=1)
    run("unet_chain_start_maxt", B=2, H=24, n_steps=10, pe=False, clip=False, cond=box_holding, start=True, max_t=4,
        seed=2)
    # attention=True variant (LinearAttention, helpers.py:75-131); H=24 exercises L=3 at the deepest level
    run("unet_chain_attention", B=2, H=24, n_steps=5, pe=False, clip=False, cond=box_holding, seed=3, attention=True)
    fid_fixture()
    v4_fixture()
    eval_metrics_fixture()
    # schedule buffers straight from the reference module
    with contextlib.redirect_stdout(io.StringIO()):
        net = TemporalUnet(8, 35, 35)
    for n in (100, 1000):
        d = GaussianDiffusion(net, 8, 35, n_timesteps=n)
        np.savez_compressed(os.path.join(HERE, f"schedule_{n}.npz"),
                            **{k: v.numpy() for k, v in d.state_dict().items() if not k.startswith("model.")})
    print("done")
