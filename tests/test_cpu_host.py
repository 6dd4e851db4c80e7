"""CPU: host logic and the C-ABI surface (no compute calls without a GPU)."""
import ctypes
import os
import re

import pytest
import torch

import deepmimic_diffusion_mujoco_b200 as dm
from deepmimic_diffusion_mujoco_b200 import _lib
from oracle import diffusion_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"
    L = ctypes.CDLL(_lib.LIB_PATH)
    header = open(os.path.join(ROOT, "include", "dm_b200.h")).read()
    declared = set(re.findall(r"\b(dm_[a-z_0-9]+)\s*\(", header))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name


def test_state_dict_keys_and_shapes_match_reference_layout():
    net = dm.TemporalUnet(64, 35, 35)
    sd = O.seeded_state_dict()
    mine = net.state_dict()
    assert set(mine) == set(sd)
    for k in sd:
        assert tuple(mine[k].shape) == tuple(sd[k].shape), k
    assert sum(p.numel() for p in net.parameters()) == 3961027  # SURVEY.md 3.2 (measured on the reference)


@pytest.mark.skipif(not os.path.isdir("/root/reference/diffusion"), reason="reference only exists in the build container")
def test_seeded_init_equals_reference_init():
    import contextlib, io, sys
    sys.path.insert(0, "/root/reference/diffusion")
    with contextlib.redirect_stdout(io.StringIO()):
        from diffuser.models.temporal import TemporalUnet as Ref
        torch.manual_seed(0)
        ref = Ref(64, 35, 35)
    torch.manual_seed(0)
    mine = dm.TemporalUnet(64, 35, 35)
    rs, ms = ref.state_dict(), mine.state_dict()
    assert list(rs) == list(ms)
    for k in rs:
        assert torch.equal(rs[k], ms[k]), k


def test_diffusion_buffers_match_oracle_schedule():
    net = dm.TemporalUnet(64, 35, 35)
    d = dm.GaussianDiffusion(net, 64, 35, n_timesteps=100)
    s = O.make_schedule(100)
    for k, v in s.items():
        assert torch.equal(getattr(d, k), v), k
    assert d.predict_epsilon is True and d.clip_denoised is False  # reference ctor defaults (diffusion.py:52-53)


def test_trace_conditioning_forms():
    shape = (4, 16, 35)
    m, v = dm.trace_conditioning(dm.apply_conditioning, shape, "cpu")
    assert m is None and v is None
    m, v = dm.trace_conditioning(dm.apply_conditioning_motion_editing, shape, "cpu")
    assert m.shape == (16, 35) and m.sum() == 16 * 8 and v[0, 16] == 1.57 and v[0, 13] == 0.0
    ref_clip = torch.randn(1, 16, 35)

    def inbetween(x):  # 4_motion_inbetweening.ipynb cell 26 form
        x[:, 0:5, :] = ref_clip[:, 3:8, :]
        return x
    m, v = dm.trace_conditioning(inbetween, shape, "cpu")
    assert m.shape == (16, 35) and m[:5].all() and not m[5:].any() and torch.equal(v[:5], ref_clip[0, 3:8])

    def keyframes(x):  # 6_walking_trajectory_v2.ipynb cell 20 form
        for k in range(8):
            x[:, k, 0] = 0.1 * k
        return x
    m, v = dm.trace_conditioning(keyframes, shape, "cpu")
    assert m.sum() == 8

    per_sample = torch.randn(4, 16, 35)

    def batched(x):
        x[:, :, 0] = per_sample[:, :, 0]
        return x
    m, v = dm.trace_conditioning(batched, shape, "cpu")
    assert m.dim() == 3 and v.dim() == 3

    captured = torch.zeros(shape)

    def buggy(x):  # 3_5_long_projection.ipynb cell 25: writes into a captured tensor -> identity on x
        captured[:, :, 0] = 1.0
        return x
    m, v = dm.trace_conditioning(buggy, shape, "cpu")
    assert m is None
    with pytest.raises(dm.DmError):
        dm.trace_conditioning(lambda x: x * 2, shape, "cpu")


def test_no_cpu_fallback():
    net = dm.TemporalUnet(64, 35, 35)
    with pytest.raises(dm.DmError):
        net(torch.zeros(1, 64, 35), {}, torch.zeros(1))
    d = dm.GaussianDiffusion(net, 64, 35, n_timesteps=10)
    with pytest.raises(dm.DmError):
        d.p_sample_loop((1, 64, 35), {}, verbose=False)
    if not torch.cuda.is_available():
        h = ctypes.c_void_p()
        mults = (ctypes.c_int * 4)(1, 2, 4, 8)
        e = _lib.DmTensor(b"x", None, 1)
        rc = _lib.lib().dm_plan_create((_lib.DmTensor * 1)(e), 0, 1, 64, 35, 32, mults, 4, ctypes.byref(h))
        assert rc == -2 and b"no CPU fallback" in _lib.lib().dm_last_error()


def test_attention_state_dict_matches_reference_keys():
    """attention=True keeps the reference's parameter names/shapes (helpers.py:75-131, temporal.py:85,94,103)."""
    from oracle import diffusion_oracle as O
    net = dm.TemporalUnet(64, 35, 35, attention=True)
    sd = O.seeded_state_dict(H=64, attention=True)
    mine = net.state_dict()
    assert set(mine) == set(sd)
    assert all(tuple(mine[k].shape) == tuple(sd[k].shape) for k in sd)
    assert tuple(mine["mid_attn.fn.fn.to_qkv.weight"].shape) == (384, 256, 1)
    assert tuple(mine["downs.0.2.fn.norm.g"].shape) == (1, 32, 1)
    with pytest.raises(dm.DmError):                      # still no CPU path
        net(torch.zeros(1, 64, 35), {}, torch.zeros(1))


@pytest.mark.parametrize("case", ["default_h64", "attention_h32_l2", "wide_t69_dim128"])
def test_reference_checkpoint_state_dict_loads_strictly(case):
    """Trainer.save stores GaussianDiffusion.state_dict() under 'model' / 'ema' (utils/training.py:133-137): `model.`
    prefixed U-Net tensors, the schedule buffers and loss_fn.weights.  The fixture lists the keys / shapes / dtypes of
    the UNMODIFIED reference module (tests/golden/make_state_dict_golden.py); a checkpoint with exactly those entries
    must load with strict=True, and a bare U-Net state_dict is its `model.`-prefixed slice."""
    import json
    fx = json.load(open(os.path.join(ROOT, "tests", "golden", "state_dict_reference.json")))[case]
    net = dm.TemporalUnet(**fx["unet"])
    diff = dm.GaussianDiffusion(net, fx["unet"]["horizon"], fx["unet"]["transition_dim"], **fx["diffusion"])
    mine = diff.state_dict()
    assert list(mine) == list(fx["keys"])                       # same keys, same order
    for k, (shape, dtype) in fx["keys"].items():
        assert list(mine[k].shape) == shape and str(mine[k].dtype) == dtype, k
    assert torch.allclose(mine["loss_fn.weights"], torch.tensor(fx["loss_fn.weights"]), rtol=0, atol=0)
    g = torch.Generator().manual_seed(3)
    ckpt = {"step": 7, "model": {k: torch.randn(shape, generator=g).to(getattr(torch, dtype.split(".")[1]))
                                 for k, (shape, dtype) in fx["keys"].items()}}
    diff.load_state_dict(ckpt["model"])                          # strict
    assert torch.equal(diff.state_dict()["model.final_conv.1.weight"], ckpt["model"]["model.final_conv.1.weight"])
    unet_sd = {k[len("model."):]: v for k, v in ckpt["model"].items() if k.startswith("model.")}
    dm.TemporalUnet(**fx["unet"]).load_state_dict(unet_sd)       # strict
