"""Drop-in for the sampling half of the reference's diffuser/models/diffusion_v4.py (DiffusionV4.__init__ :8-24,
prepare_noise_schedule :26-43, sample :82-149), the later-generation sampler that drives the same TemporalUnet
family (eval.py:189-197).

The reverse update of diffusion_v4.py:137-147 is affine in (x, model output, noise) for both parameterisations:
    predict noise: x' = x / sqrt(a) - (1 - a) / (sqrt(a) sqrt(1 - ah)) * out + sqrt(b) z
    predict x0   : x' = (1 - (1 - a) / (1 - ah)) / sqrt(a) * x + (1 - a) sqrt(ah) / ((1 - ah) sqrt(a)) * out + sqrt(b) z
so it runs on the SAME fused CUDA sampling loop as GaussianDiffusion (dm_sample_loop; K4 computes c2 x + c1 out + sigma z)
with a per-step coefficient table; the loop walks t = T-1 .. 1 (diffusion_v4.py:100) and the time embedding sees t
through dm_schedule_t.time_values.  No CPU fallback.
"""
import ctypes
import math

import torch

from . import _lib


class DiffusionV4:
    def __init__(self, noise_steps, beta_start, beta_end, joint_dim, frames, device="cuda", predict_x0=False,
                 schedule_type="linear", cosine_s=0.008, cfg_scale=3.0):
        self.noise_steps = noise_steps
        self.beta_start = beta_start
        self.beta_end = beta_end
        self.joint_dim = joint_dim
        self.frames = frames
        self.device = device
        self.predict_x0 = predict_x0
        self.cfg_scale = cfg_scale
        self.schedule_type = schedule_type
        self.cosine_s = cosine_s
        self.beta = self.prepare_noise_schedule().to(device)
        self.alpha = 1. - self.beta
        self.alpha_hat = torch.cumprod(self.alpha, dim=0)
        self.seed = 4321
        self._calls = 0
        self.last_launch_count = 0

    def prepare_noise_schedule(self):
        """diffusion_v4.py:26-43, same fp32 torch arithmetic."""
        if self.schedule_type == "linear":
            return torch.linspace(self.beta_start, self.beta_end, self.noise_steps)
        if self.schedule_type == "cosine":
            steps = self.noise_steps + 1
            x = torch.linspace(0, self.noise_steps, steps)
            ac = torch.cos(((x / self.noise_steps) + self.cosine_s) / (1 + self.cosine_s) * math.pi * 0.5) ** 2
            ac = ac / ac[0]
            betas = 1 - (ac[1:] / ac[:-1])
            return torch.clip(betas, self.beta_start, self.beta_end)
        raise ValueError(f"Unknown schedule type: {self.schedule_type}. Use 'linear' or 'cosine'.")

    def step_table(self):
        """Coefficients of the affine update for loop index j <-> reference timestep i = j + 1 (float64 from the
        reference's fp32 alpha / alpha_hat / beta, rounded once to fp32)."""
        a = self.alpha.detach().cpu().double()[1:]
        ah = self.alpha_hat.detach().cpu().double()[1:]
        b = self.beta.detach().cpu().double()[1:]
        if self.predict_x0:
            c2 = (1.0 - (1.0 - a) / (1.0 - ah)) / a.sqrt()
            c1 = (1.0 - a) * ah.sqrt() / ((1.0 - ah) * a.sqrt())
        else:
            c2 = 1.0 / a.sqrt()
            c1 = -(1.0 - a) / (a.sqrt() * (1.0 - ah).sqrt())
        n = a.numel()
        zeros = torch.zeros(n)
        return {"sqrt_recip_alphas_cumprod": zeros, "sqrt_recipm1_alphas_cumprod": zeros,
                "posterior_mean_coef1": c1.float(), "posterior_mean_coef2": c2.float(),
                "posterior_log_variance_clipped": b.log().float(),
                "time_values": torch.arange(1, n + 1, dtype=torch.float32)}

    @torch.no_grad()
    def sample(self, model, n, custom_frames=None, condition=None, y=None, noise=None, return_chain=False, seed=None,
               sample_offset=0, fused=True):
        """Reverse diffusion of n samples (diffusion_v4.py:82-149).  model: this package's TemporalUnet.
        noise: optional injected stack [noise_steps, n, frames, joint_dim] (slot 0 = the initial draw, slot k = the k-th
        randn_like; the reference zeroes the last one itself).  Returns x [n, frames, joint_dim]."""
        if y is not None:
            raise _lib.DmError("classifier-free guidance (y is not None) needs a label-conditioned model; "
                               "TemporalUnet takes no label (temporal_v2.py:118) -- no fallback")
        if not hasattr(model, "plan_for"):
            raise _lib.DmError("DiffusionV4.sample runs deepmimic_diffusion_mujoco_b200.TemporalUnet only (no fallback)")
        device = torch.device(self.device)
        if device.type != "cuda":
            raise _lib.DmError("DiffusionV4.sample needs a B200 device (no CPU fallback)")
        H = int(custom_frames) if custom_frames is not None else int(self.frames)
        B, T = int(n), int(self.joint_dim)
        steps = self.noise_steps - 1
        with torch.cuda.device(device):
            plan = model.plan_for(B, H)
            sched_id = ("v4", id(self), self.predict_x0, self.schedule_type, self.noise_steps,
                        float(self.beta_start), float(self.beta_end), self.beta.data_ptr(), self.beta._version)
            if plan.schedule_id != sched_id:   # the table costs device->host reads: build it only when the plan needs it
                plan.set_schedule(self.step_table(), sched_id)
            _lib.check(_lib.lib().dm_plan_set_sample_offset(plan.handle, int(sample_offset)))
            x = torch.empty(B, H, T, device=device, dtype=torch.float32)
            stack = None
            if noise is not None:
                stack = noise.to(device=device, dtype=torch.float32).contiguous()
                if tuple(stack.shape) != (steps + 1, B, H, T):
                    raise _lib.DmError(f"noise stack must be {(steps + 1, B, H, T)}, got {tuple(stack.shape)}")
            chain = torch.empty(B, steps + 1, H, T, device=device) if return_chain else None
            if seed is None:
                seed = self.seed + 7919 * self._calls
                self._calls += 1
            c = _lib.DmCond()
            _lib.check(_lib.lib().dm_sample_loop(
                plan.handle, x.data_ptr(), 0, stack.data_ptr() if stack is not None else None,
                ctypes.c_uint64(seed & (2 ** 64 - 1)), steps, 0, 0, ctypes.byref(c),
                chain.data_ptr() if chain is not None else None,
                int(bool(fused)), _lib.stream_ptr()))
            self.last_launch_count = plan.last_launch_count()
        return (x, chain) if return_chain else x
