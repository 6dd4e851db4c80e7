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


def slice_motion_sample(sample, window_size, step_size=10):
    """eval.py:15-21: windows of `window_size` frames at offsets 0, step, 2 step, ... -> [W, window_size, nfeats]."""
    return sample.unfold(0, window_size, step_size).permute(0, 2, 1).contiguous()


def window_moments(samples, window_size=10, step_size=10):
    """Per-sample mean [N,D] and covariance [N,D,D] (fp32) over each sample's windows -- the statistics calc_sifid
    (eval.py:59-77) computes one sample at a time -- in one launch."""
    _need_cuda(samples, "samples")
    X = samples.to(torch.float32).contiguous()
    N, H, F = X.shape
    D = window_size * F
    mu = torch.empty(N, D, device=X.device, dtype=torch.float32)
    sigma = torch.empty(N, D, D, device=X.device, dtype=torch.float32)
    with torch.cuda.device(X.device):
        _lib.check(_lib.lib().dm_window_moments(X.data_ptr(), N, H, F, window_size, step_size, mu.data_ptr(),
                                                sigma.data_ptr(), _lib.stream_ptr()))
    return mu, sigma


def calc_sifid(gen_samples, gt_sample, window_size=10, batch=64):
    """SiFID (eval.py:59-77): mean over generated samples of the FID between the window statistics of the ground-truth
    motion and of that sample.  Statistics: one dm_window_moments launch; the per-sample D x D SVDs
    (fid_score.py:64-67) are batched torch.linalg.svd calls (library code by design)."""
    _need_cuda(gen_samples, "gen_samples")
    dev = gen_samples.device
    gt = gt_sample.to(device=dev, dtype=torch.float32)
    gmu, gsig = window_moments(gt[None], window_size)
    gmu, gsig = gmu[0], gsig[0]
    mu, sig = window_moments(gen_samples, window_size)
    fids = []
    for i in range(0, mu.shape[0], batch):
        m, s = mu[i:i + batch], sig[i:i + batch]
        diff = gmu[None] - m
        U, sv, Vh = torch.linalg.svd(gsig[None] @ s)
        covmean_tr = ((U * torch.sqrt(torch.clamp(sv, min=1e-6))[:, None, :]) * Vh.transpose(1, 2)).sum(dim=(1, 2))
        fids.append((diff * diff).sum(1) + torch.trace(gsig) + s.diagonal(dim1=1, dim2=2).sum(1) - 2 * covmean_tr)
    return torch.cat(fids).mean().item()


def calc_inter_diversity(samples):
    """eval.py:54-57: distance between the first and second half of the sample set.  The 'unet' wrapper returns the
    samples themselves as embeddings (eval.py:31-35); the norm runs over axis 1 (frames), as in the reference."""
    _need_cuda(samples, "samples")
    h = samples.shape[0] // 2
    d = samples[:h].float() - samples[h:2 * h].float()
    return torch.linalg.norm(d, dim=1).mean().item()


def calc_intra_diversity(samples, window_size=10, offsets=None):
    """eval.py:80-89: per sample, the distance between two windows at random offsets.  offsets: optional int array
    [N,2]; None draws them with np.random.randint like the reference."""
    import numpy as np
    _need_cuda(samples, "samples")
    N, H, F = samples.shape
    if offsets is None:
        offsets = np.stack([np.random.randint(H - window_size, size=2) for _ in range(N)])
    off = torch.as_tensor(np.asarray(offsets), device=samples.device, dtype=torch.long)
    idx = off[:, :, None] + torch.arange(window_size, device=samples.device)[None, None, :]      # [N,2,w]
    win = torch.gather(samples.float()[:, None].expand(N, 2, H, F), 2, idx[..., None].expand(N, 2, window_size, F))
    return torch.linalg.norm((win[:, 0] - win[:, 1]).reshape(N, -1), dim=1).mean().item()


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
