"""GPU test of dm_fk_reward_noalloc (collected last on purpose: the file name sorts after every other test file).

The entry point has no counterpart in the reference; it was added for callers that
cannot allocate.  It must give what dm_fk_reward gives (same kernels, scratch in a static device array instead of a pool
allocation), on consecutive calls that reuse the shared scratch with different horizons, and reject horizons beyond
DM_FK_NOALLOC_MAX_H.  Called through the C ABI (ctypes), like every other GPU test."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_H = 1024   # DM_FK_NOALLOC_MAX_H in include/dm_b200.h


def _inputs(N, H, seed):
    g = torch.Generator().manual_seed(seed)
    ref = torch.zeros(H, 35)
    ref[:, 2] = 0.9
    ref[:, 3] = 1.0
    ref[:, 3:7] += 0.1 * torch.randn(H, 4, generator=g)
    ref[:, 7:] = 0.3 * torch.randn(H, 28, generator=g)
    q = ref[None] + 0.15 * torch.randn(N, H, 35, generator=g)
    q[0] = ref
    return q.cuda().contiguous(), ref.cuda().contiguous()


def _call(fn, q, ref, dt, body):
    from deepmimic_diffusion_mujoco_b200 import _lib
    N, H, _ = q.shape
    reward = torch.full((N, H), float("nan"), device="cuda")
    sums = torch.zeros(8, device="cuda", dtype=torch.float64)
    bp = torch.full((N, H, 15, 3), float("nan"), device="cuda") if body else None
    rc = fn(q.data_ptr(), ref.data_ptr(), N, H, ctypes.c_float(dt), bp.data_ptr() if body else None, reward.data_ptr(),
            sums.data_ptr(), _lib.stream_ptr())
    return rc, reward, sums, bp


def test_noalloc_matches_fk_reward_on_consecutive_calls():
    from deepmimic_diffusion_mujoco_b200 import _lib
    assert torch.cuda.is_available(), "GPU tests need a B200"
    L = _lib.lib()
    dt = 1 / 60
    # two calls in a row (then a third): the second reuses the static scratch with a SHORTER horizon, so stale rows of the
    # first call sit behind the rows it uses; the third goes back up to the maximum horizon
    for N, H, body, seed in [(37, 64, True, 0), (5, 40, False, 1), (3, MAX_H, True, 2)]:
        q, ref = _inputs(N, H, seed)
        rc0, r0, s0, b0 = _call(L.dm_fk_reward, q, ref, dt, body)
        rc1, r1, s1, b1 = _call(L.dm_fk_reward_noalloc, q, ref, dt, body)
        torch.cuda.synchronize()
        assert rc0 == 0 and rc1 == 0, L.dm_last_error()
        assert torch.isfinite(r1).all()
        assert (r1 - r0).abs().max().item() <= 1e-6, (N, H)             # same kernels: expected bit-identical
        assert abs(r1[0].mean().item() - 1.0) < 1e-4                     # a perfect tracker scores 1
        assert torch.allclose(s1, s0, rtol=1e-9, atol=1e-9), (s1, s0)    # fp64 atomics: order of addition may differ
        assert s1[6].item() == N * H
        if body:
            assert (b1 - b0).abs().max().item() <= 1e-6


def test_noalloc_rejects_horizons_beyond_the_static_scratch():
    from deepmimic_diffusion_mujoco_b200 import _lib
    L = _lib.lib()
    q, ref = _inputs(1, MAX_H + 8, 3)
    rc, reward, _, _ = _call(L.dm_fk_reward_noalloc, q, ref, 1 / 60, False)
    torch.cuda.synchronize()
    assert rc == -1 and b"DM_FK_NOALLOC_MAX_H" in L.dm_last_error()      # DM_ERR_ARG, loud, and nothing was launched
    assert torch.isnan(reward).all()
    rc, reward, _, _ = _call(L.dm_fk_reward, q, ref, 1 / 60, False)      # the allocating entry point still takes it
    torch.cuda.synchronize()
    assert rc == 0 and torch.isfinite(reward).all()
