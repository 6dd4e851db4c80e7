"""Drop-in for the reference's diffuser/models/diffusion.py sampling half (:16-41, :44-115, :147-240, :279-280).

GaussianDiffusion keeps the reference constructor, buffers and p_sample_loop / conditional_sample signatures; the
100-step loop itself is ONE call into libdm_b200.so (dm_sample_loop: one persistent cluster kernel that runs every
tcgen05 conv layer and the fused posterior+noise+inpaint update of every step).  Training (p_losses/loss, :242-277) is out of scope.
"""
import ctypes
from collections import namedtuple

import numpy as np
import torch
from torch import nn

from . import _lib
from .sampling_config import apply_conditioning, trace_conditioning_device

Sample = namedtuple("Sample", "trajectories values chains")


def cosine_beta_schedule(timesteps, s=0.008, dtype=torch.float32):
    """helpers.py:145-156 (float64 numpy, cast to fp32)."""
    steps = timesteps + 1
    x = np.linspace(0, steps, steps)
    ac = np.cos(((x / steps) + s) / (1 + s) * np.pi * 0.5) ** 2
    ac = ac / ac[0]
    betas = 1 - (ac[1:] / ac[:-1])
    return torch.tensor(np.clip(betas, a_min=0, a_max=0.999), dtype=dtype)


def default_sample_fn(model, x, cond, t):
    """Marker for the built-in sampler (diffusion.py:19-29).  The arithmetic runs inside dm_sample_loop with
    Philox noise; this function object is only compared by identity and is never called."""
    raise _lib.DmError("default_sample_fn is fused into the CUDA sampling loop and cannot be called directly")


class InjectedNoise:
    """sample_fn that replays pre-generated noise (the reference's own plug point, diffusion.py:193,214).
    stack: [timesteps+1, B, H, T]; slot 0 replaces the initial torch.randn draw (:205), slot k the k-th
    randn_like (:25)."""

    def __init__(self, stack):
        self.stack = stack


def sort_by_values(x, values):
    # diffusion.py:32-36.  values are all zero for the default sampler, so the reference's argsort returns an
    # arbitrary permutation; stable=True picks the identity one, which keeps rows aligned with their noise.
    inds = torch.argsort(values, descending=True, stable=True)
    return x[inds], values[inds]


def make_timesteps(batch_size, i, device):
    return torch.full((batch_size,), i, device=device, dtype=torch.long)


class _LossWeights(nn.Module):
    """State holder for the reference's loss module (helpers.py:163-167 WeightedLoss; :182-185 ValueLoss has no
    state).  Training is out of scope, but Trainer.save (utils/training.py:133-137) stores the GaussianDiffusion
    state_dict, which carries `loss_fn.weights`: the buffer must exist for those checkpoints to load strictly."""

    def __init__(self, weights, action_dim):
        super().__init__()
        if weights is not None:
            self.register_buffer("weights", weights)
        self.action_dim = action_dim

    def forward(self, pred, targ):
        raise NotImplementedError("training losses (helpers.py:169-214) are outside the accelerated hot path")


class GaussianDiffusion(nn.Module):
    def __init__(self, model, horizon, transition_dim, n_timesteps=1000, loss_type="l1", clip_denoised=False,
                 predict_epsilon=True, action_weight=1.0, loss_discount=1.0, loss_weights=None):
        super().__init__()
        self.horizon = horizon
        self.action_dim = transition_dim
        self.transition_dim = transition_dim
        self.model = model

        betas = cosine_beta_schedule(n_timesteps)
        alphas = 1.0 - betas
        alphas_cumprod = torch.cumprod(alphas, axis=0)
        alphas_cumprod_prev = torch.cat([torch.ones(1), alphas_cumprod[:-1]])
        self.n_timesteps = int(n_timesteps)
        self.clip_denoised = clip_denoised
        self.predict_epsilon = predict_epsilon

        reg = self.register_buffer
        reg("betas", betas)
        reg("alphas_cumprod", alphas_cumprod)
        reg("alphas_cumprod_prev", alphas_cumprod_prev)
        reg("sqrt_alphas_cumprod", torch.sqrt(alphas_cumprod))
        reg("sqrt_one_minus_alphas_cumprod", torch.sqrt(1.0 - alphas_cumprod))
        reg("log_one_minus_alphas_cumprod", torch.log(1.0 - alphas_cumprod))
        reg("sqrt_recip_alphas_cumprod", torch.sqrt(1.0 / alphas_cumprod))
        reg("sqrt_recipm1_alphas_cumprod", torch.sqrt(1.0 / alphas_cumprod - 1))
        posterior_variance = betas * (1.0 - alphas_cumprod_prev) / (1.0 - alphas_cumprod)
        reg("posterior_variance", posterior_variance)
        reg("posterior_log_variance_clipped", torch.log(torch.clamp(posterior_variance, min=1e-20)))
        reg("posterior_mean_coef1", betas * np.sqrt(alphas_cumprod_prev) / (1.0 - alphas_cumprod))
        reg("posterior_mean_coef2", (1.0 - alphas_cumprod_prev) * np.sqrt(alphas) / (1.0 - alphas_cumprod))
        # loss coefficients (diffusion.py:112-145): kept only so that the state_dict matches the reference's
        lw = self.get_loss_weights(action_weight, loss_discount, loss_weights)
        if loss_type not in ("l1", "l2", "value_l1", "value_l2"):
            raise KeyError(loss_type)   # Losses[loss_type] (helpers.py:233-238)
        self.loss_fn = _LossWeights(lw if loss_type in ("l1", "l2") else None, self.action_dim)
        self.seed = 1234
        self._calls = 0
        self._bufs = {}
        self.last_launch_count = 0

    def get_loss_weights(self, action_weight, discount, weights_dict):
        """diffusion.py:116-145: discount**t over the horizon (mean-normalised) x per-dimension weights, a0 weight set."""
        self.action_weight = action_weight
        dim_weights = torch.ones(self.transition_dim, dtype=torch.float32)
        for ind, w in (weights_dict or {}).items():
            dim_weights[self.action_dim + ind] *= w
        discounts = discount ** torch.arange(self.horizon, dtype=torch.float)
        discounts = discounts / discounts.mean()
        loss_weights = torch.einsum("h,t->ht", discounts, dim_weights)
        loss_weights[0, : self.action_dim] = action_weight
        return loss_weights

    def _schedule_bufs(self):
        return {k: getattr(self, k) for k in ("sqrt_recip_alphas_cumprod", "sqrt_recipm1_alphas_cumprod",
                                              "posterior_mean_coef1", "posterior_mean_coef2",
                                              "posterior_log_variance_clipped")}

    def _schedule_id(self):
        return tuple((b.data_ptr(), b._version) for b in self._schedule_bufs().values())

    @torch.no_grad()
    def p_sample_loop(self, shape, cond, verbose=True, return_chain=False, sample_fn=default_sample_fn,
                      conditioning_fn=apply_conditioning, starting_motion=None, max_timesteps=None,
                      fused=True, seed=None, sample_offset=0, **sample_kwargs):
        device = self.betas.device
        if device.type != "cuda":
            raise _lib.DmError("GaussianDiffusion.p_sample_loop needs the module on a B200 (no CPU fallback)")
        if sample_kwargs:
            raise _lib.DmError(f"unsupported sample_kwargs {sorted(sample_kwargs)}")
        B, H, T = (int(s) for s in shape)
        timesteps = self.n_timesteps if max_timesteps is None else int(max_timesteps)

        noise = None
        if isinstance(sample_fn, InjectedNoise):
            noise = sample_fn.stack.to(device=device, dtype=torch.float32).contiguous()
            if tuple(noise.shape) != (timesteps + 1, B, H, T):
                raise _lib.DmError(f"InjectedNoise stack must be {(timesteps + 1, B, H, T)}, got {tuple(noise.shape)}")
        elif sample_fn is not default_sample_fn:
            raise _lib.DmError("custom sample_fn callables cannot run inside the fused loop; use default_sample_fn "
                               "or InjectedNoise (no fallback)")

        with torch.cuda.device(device):
            # persistent device buffers: no allocator traffic on the sampling path
            x = self._persistent(("x", B, H, T), (B, H, T), torch.float32, device)
            if starting_motion is not None:
                if tuple(starting_motion.shape) != (B, H, T):
                    raise _lib.DmError("starting_motion shape does not match `shape`")
                x.copy_(starting_motion)
            traced = trace_conditioning_device(conditioning_fn, (B, H, T), device)
            mask = value = flags = None
            if traced is not None:
                mask = self._persistent(("mask", B, H, T), (B, H, T), torch.uint8, device)
                value = self._persistent(("value", B, H, T), (B, H, T), torch.float32, device)
                flags = self._persistent(("flags", B, H, T), (4,), torch.uint8, device)
                mask.copy_(traced[0])
                value.copy_(traced[1])
                flags.copy_(traced[2])
            plan = self.model.plan_for(B, H)
            plan.set_schedule(self._schedule_bufs(), self._schedule_id())
            # multi-GPU shards: rows [sample_offset, sample_offset + B) of one global batch (same seed on every rank)
            _lib.check(_lib.lib().dm_plan_set_sample_offset(plan.handle, int(sample_offset)))
            chain = torch.empty(B, timesteps + 1, H, T, device=device) if return_chain else None
            c = _lib.DmCond()
            if mask is not None:
                # full [B,H,T] tensors + device-resident structure flags: the kernel picks row 0 when batch-uniform
                c.mask_dev, c.value_dev, c.flags_dev = mask.data_ptr(), value.data_ptr(), flags.data_ptr()
                c.mask_batched = c.value_batched = 1
            if seed is None:
                seed = self.seed + 7919 * self._calls
                self._calls += 1
            _lib.check(_lib.lib().dm_sample_loop(
                plan.handle, x.data_ptr(), int(starting_motion is not None),
                noise.data_ptr() if noise is not None else None, ctypes.c_uint64(seed & (2 ** 64 - 1)), timesteps,
                int(self.predict_epsilon), int(self.clip_denoised), ctypes.byref(c),
                chain.data_ptr() if chain is not None else None, int(bool(fused)),
                _lib.stream_ptr()))
            self.last_launch_count = plan.last_launch_count()
            values = torch.zeros(B, device=device)
            if verbose:
                print(f"[ b200 sampler ] {timesteps} steps x {B} samples enqueued ({self.last_launch_count} launches)")
            # sort_by_values (diffusion.py:225): the default sampler's values are all zero and a stable descending
            # argsort of zeros is the identity, so the gather x[inds] is a plain copy out of the persistent buffer
            x = x.clone()
        return Sample(x, values, chain)

    def _persistent(self, key, shape, dtype, device):
        buf = self._bufs.get(key)
        if buf is None or buf.device != device:
            if len(self._bufs) >= 12:
                self._bufs.pop(next(iter(self._bufs)))
            buf = torch.empty(shape, dtype=dtype, device=device)
            self._bufs[key] = buf
        return buf

    @torch.no_grad()
    def conditional_sample(self, cond, horizon=None, **sample_kwargs):
        """conditions : [ (time, state), ... ] -- only len(cond[0]) is used, as in the reference (:236)."""
        batch_size = len(cond[0])
        horizon = horizon or self.horizon
        return self.p_sample_loop((batch_size, horizon, self.transition_dim), cond, **sample_kwargs)

    def forward(self, cond, *args, **kwargs):
        return self.conditional_sample(cond, *args, **kwargs)

    def loss(self, *args, **kwargs):
        raise NotImplementedError("training (diffusion.py:242-277) is outside the accelerated hot path")
