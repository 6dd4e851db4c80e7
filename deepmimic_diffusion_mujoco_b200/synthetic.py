"""Synthetic TemporalUnet weights with the reference's state_dict keys and shapes.

The reference ships no checkpoints, so benchmarks, tests, smoke() and the golden-fixture generator
all draw their weights from this seeded generator.  It is neutral ground: neither the oracle nor a kernel path.
Key / shape layout: diffuser/models/temporal.py:51-113, helpers.py:50-69,110-117 (under /root/reference/diffusion/).
"""
import math

import torch


def seeded_state_dict(H=64, T=35, dim=32, mults=(1, 2, 4, 8), seed=0, attention=False):
    """Synthetic weights with the reference's state_dict keys and shapes (no checkpoints ship with the
    reference).  GroupNorm affine and all biases are randomised so that bugs in them show."""
    g = torch.Generator().manual_seed(seed)
    dims = [T] + [dim * m for m in mults]
    sd = {}

    def lin(key, out_f, in_f):
        sd[key + ".weight"] = torch.randn(out_f, in_f, generator=g) / math.sqrt(in_f)
        sd[key + ".bias"] = 0.1 * torch.randn(out_f, generator=g)

    def conv(key, co, ci, k, transpose=False):
        shape = (ci, co, k) if transpose else (co, ci, k)
        sd[key + ".weight"] = torch.randn(*shape, generator=g) / math.sqrt(ci * k)
        sd[key + ".bias"] = 0.1 * torch.randn(co, generator=g)

    def block(key, ci, co):
        conv(key + ".block.0", co, ci, 5)
        sd[key + ".block.2.weight"] = 1 + 0.2 * torch.randn(co, generator=g)
        sd[key + ".block.2.bias"] = 0.1 * torch.randn(co, generator=g)

    def res(key, ci, co):
        block(key + ".blocks.0", ci, co)
        block(key + ".blocks.1", co, co)
        lin(key + ".time_mlp.1", co, dim)
        if ci != co:
            conv(key + ".residual_conv", co, ci, 1)

    def attn(key, c):
        if not attention:
            return
        sd[key + ".fn.fn.to_qkv.weight"] = torch.randn(384, c, 1, generator=g) / math.sqrt(c)
        sd[key + ".fn.fn.to_out.weight"] = torch.randn(c, 128, 1, generator=g) / math.sqrt(128)
        sd[key + ".fn.fn.to_out.bias"] = 0.1 * torch.randn(c, generator=g)
        sd[key + ".fn.norm.g"] = 1 + 0.2 * torch.randn(1, c, 1, generator=g)
        sd[key + ".fn.norm.b"] = 0.1 * torch.randn(1, c, 1, generator=g)

    lin("time_mlp.1", 4 * dim, dim)
    lin("time_mlp.3", dim, 4 * dim)
    n = len(mults)
    for i in range(n):
        res(f"downs.{i}.0", dims[i], dims[i + 1])
        res(f"downs.{i}.1", dims[i + 1], dims[i + 1])
        attn(f"downs.{i}.2", dims[i + 1])
        if i < n - 1:
            conv(f"downs.{i}.3.conv", dims[i + 1], dims[i + 1], 3)
    res("mid_block1", dims[n], dims[n])
    attn("mid_attn", dims[n])
    res("mid_block2", dims[n], dims[n])
    for u in range(n - 1):
        lev = n - 1 - u
        res(f"ups.{u}.0", 2 * dims[lev + 1], dims[lev])
        res(f"ups.{u}.1", dims[lev], dims[lev])
        attn(f"ups.{u}.2", dims[lev])
        conv(f"ups.{u}.3.conv", dims[lev], dims[lev], 4, transpose=True)
    block("final_conv.0", dim, dim)
    conv("final_conv.1", T, dim, 1)
    return sd
