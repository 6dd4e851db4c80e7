"""Sample scoring on the GPU: batched forward kinematics + DeepMimic tracking reward (K5) and motion-FID
statistics (K6), with the one collective of the path -- an all-reduce of the moment / reward sums.

Replaces: sim.forward() playback FK (mocap_player.py:67-73), MotionFID.calculate_activation_statistics and
calculate_fid (metrics/fid_score.py:24-81).  No CPU fallback: tensors must live on a B200.
"""
import ctypes

import torch

from . import _lib

N_BODIES = 15  # BODY_DEFS order (utils/mocap_util.py:17-19): 13 bodies + 2 wrist geoms


def _need_cuda(t, name):
    if not (torch.is_tensor(t) and t.is_cuda):
        raise _lib.DmError(f"{name} must be a CUDA tensor on a B200 (no CPU fallback)")


def fk_reward(qpos, ref, dt, return_body_pos=False):
    """qpos [N,H,35] fp32 CUDA, ref [H,35] -> dict(reward [N,H], sums double[8], body_pos [N,H,15,3] | None).
    sums = (sum r, sum r^2, sum pose term, vel term, end-effector term, com term, frame count, 0)."""
    _need_cuda(qpos, "qpos")
    qpos = qpos.to(torch.float32).contiguous()
    N, H, D = qpos.shape
    if D != 35 or tuple(ref.shape) != (H, 35):
        raise _lib.DmError("fk_reward expects qpos [N,H,35] and ref [H,35]")
    ref = ref.to(device=qpos.device, dtype=torch.float32).contiguous()
    reward = torch.empty(N, H, device=qpos.device, dtype=torch.float32)
    sums = torch.zeros(8, device=qpos.device, dtype=torch.float64)
    body = torch.empty(N, H, N_BODIES, 3, device=qpos.device, dtype=torch.float32) if return_body_pos else None
    with torch.cuda.device(qpos.device):
        _lib.check(_lib.lib().dm_fk_reward(qpos.data_ptr(), ref.data_ptr(), N, H, ctypes.c_float(dt),
                                           body.data_ptr() if body is not None else None, reward.data_ptr(),
                                           sums.data_ptr(), _lib.stream_ptr()))
    return {"reward": reward, "sums": sums, "body_pos": body}


def fid_moments(data, sum_out=None, sxx_out=None):
    """Accumulates sum_n x and X^T X (fp64) of flattened trajectories data [N, ...] into (sum [D], sxx [D,D])."""
    _need_cuda(data, "data")
    X = data.reshape(data.shape[0], -1).to(torch.float32).contiguous()
    N, D = X.shape
    if sum_out is None:
        sum_out = torch.zeros(D, device=X.device, dtype=torch.float64)
        sxx_out = torch.zeros(D, D, device=X.device, dtype=torch.float64)
    with torch.cuda.device(X.device):
        _lib.check(_lib.lib().dm_fid_moments(X.data_ptr(), N, D, sum_out.data_ptr(), sxx_out.data_ptr(),
                                             _lib.stream_ptr()))
    return sum_out, sxx_out


def shard_range(n, rank, world):
    """Contiguous split of n independent samples over `world` ranks (sizes differ by at most one)."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_statistics(count, *tensors):
    """The path's only exchange step: sum the packed statistics over ranks (NCCL over NVLink; gloo in CPU tests).
    count: python int or 0-d tensor; tensors: fp64 tensors reduced in place.  Returns the global count."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return int(count)
    dev = tensors[0].device
    flat = torch.cat([torch.tensor([float(count)], dtype=torch.float64, device=dev)] +
                     [t.reshape(-1).to(torch.float64) for t in tensors])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    off = 1
    for t in tensors:
        n = t.numel()
        t.copy_(flat[off:off + n].reshape(t.shape))
        off += n
    return int(round(flat[0].item()))


def finalize_statistics(count, s, sxx):
    """mu = S/N ; sigma = (Sxx - N mu mu^T)/(N-1)  == fid_score.py:35-40 up to rounding."""
    mu = s / count
    sigma = (sxx - count * torch.outer(mu, mu)) / (count - 1)
    return mu, sigma


class MotionFID:
    """Drop-in for metrics/fid_score.py:7-81 (statistics + the SVD-based FID formula)."""

    def __init__(self, real_dataset=None, model=None, device="cuda", batch_size=128, num_samples=512):
        self.real_dataset, self.model, self.device = real_dataset, model, device
        self.batch_size, self.num_samples = batch_size, num_samples

    def calculate_activation_statistics(self, data, distributed=False):
        data = data.to(self.device) if not data.is_cuda else data
        s, sxx = fid_moments(data)
        n = data.shape[0]
        if distributed:
            n = allreduce_statistics(n, s, sxx)
        mu, sigma = finalize_statistics(n, s, sxx)
        return mu.float().cpu().numpy(), sigma.float().cpu().numpy()

    def calculate_fid(self, mu1, sigma1, mu2, sigma2):
        # one D x D SVD per evaluation on rank 0; host-side torch by design
        mu1, mu2 = (torch.as_tensor(m, device=self.device, dtype=torch.float32) for m in (mu1, mu2))
        sigma1, sigma2 = (torch.as_tensor(m, device=self.device, dtype=torch.float32) for m in (sigma1, sigma2))
        diff = mu1 - mu2
        U, sv, Vh = torch.linalg.svd(sigma1 @ sigma2)
        covmean = U @ torch.diag(torch.sqrt(torch.clamp(sv, min=1e-6))) @ Vh
        fid = torch.dot(diff, diff) + torch.trace(sigma1) + torch.trace(sigma2) - 2 * torch.trace(covmean)
        return fid.cpu().item()
