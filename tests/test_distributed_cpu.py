"""world_size-2 gloo test of the path's one exchange step (scoring.allreduce_statistics) and of the sample sharding:
moments are sums, so the 2-rank result must equal the 1-rank result."""
import os
import socket

import numpy as np
import torch
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, X, out_path):
    import torch.distributed as dist
    from deepmimic_diffusion_mujoco_b200 import scoring
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = scoring.shard_range(X.shape[0], rank, world)
    Xr = X[lo:hi].double()
    s, sxx = Xr.sum(0), Xr.T @ Xr                      # per-rank partial sums (what dm_fid_moments returns on a GPU)
    rew = torch.tensor([float(Xr.sum()), float((Xr * Xr).sum())], dtype=torch.float64)
    n = scoring.allreduce_statistics(hi - lo, s, sxx, rew)
    mu, sigma = scoring.finalize_statistics(n, s, sxx)
    if rank == 0:
        torch.save({"n": n, "mu": mu, "sigma": sigma, "rew": rew}, out_path)
    dist.destroy_process_group()


def test_two_rank_statistics_equal_single_rank(tmp_path):
    g = torch.Generator().manual_seed(3)
    X = torch.randn(37, 24, generator=g)
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), X, out), nprocs=2, join=True)
    r = torch.load(out)
    Xd = X.double()
    assert r["n"] == 37
    mu = Xd.mean(0)
    sigma = (Xd - mu).T @ (Xd - mu) / (37 - 1)         # metrics/fid_score.py:35-40
    assert torch.allclose(r["mu"], mu, rtol=1e-12, atol=1e-12)
    assert torch.allclose(r["sigma"], sigma, rtol=1e-9, atol=1e-12)
    assert abs(r["rew"][0].item() - Xd.sum().item()) < 1e-9


def test_shard_range_partitions_contiguously():
    from deepmimic_diffusion_mujoco_b200 import scoring
    for n, w in [(65536, 8), (37, 2), (5, 8), (1024, 1)]:
        spans = [scoring.shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1
