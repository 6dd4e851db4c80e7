"""Drop-in for the reference's diffuser/models/sampling_config.py:6-32 plus the tracer that turns a
conditioning_fn callable into the declarative (mask, value) form the fused K4 kernel applies.

Every conditioning function the reference and its notebooks use overwrites a fixed set of (frame, channel)
entries with values that do not depend on x, i.e. x = where(mask, value, x).  The
tracer derives mask/value by calling the function on probe tensors; anything that is not of that form is a
hard error (there is no per-step Python fallback).
"""
import torch

from ._lib import DmError


# identity function for conditioning, ie no conditioning (sampling_config.py:6-7)
def apply_conditioning(x):
    return x


def apply_conditioning_in_betweening(x):   # sampling_config.py:11-13 (body lives in the notebook)
    return x


def apply_conditioning_motion_blending(x):  # sampling_config.py:17-19
    return x


def apply_conditioning_motion_editing(x):
    """Box holding: shoulders (13-15, 17-19) to 0, elbows (16, 20) to 90 degrees (sampling_config.py:23-32)."""
    elbow_val = 1.57
    shoulder_val = torch.tensor([0.0] * 3, dtype=torch.float32)
    x[:, :, 13:16] = shoulder_val
    x[:, :, 16] = elbow_val
    x[:, :, 17:20] = shoulder_val
    x[:, :, 20] = elbow_val
    return x


import weakref

_PROBES = {}
_VERIFIED = weakref.WeakKeyDictionary()   # conditioning function -> {(shape, device)} verified as pure overwrites


def _probes(shape, device):
    """Device-resident probe tensors (all-NaN, seeded normal), cached per (shape, device); callers clone them."""
    key = (shape, str(device))
    if key not in _PROBES:
        if len(_PROBES) >= 4:
            _PROBES.pop(next(iter(_PROBES)))
        g = torch.Generator(device="cpu").manual_seed(5)
        _PROBES[key] = (torch.full(shape, float("nan"), device=device, dtype=torch.float32),
                        torch.randn(shape, generator=g).to(device))
    return _PROBES[key]


def trace_conditioning(fn, shape, device):
    """Returns (mask uint8, value fp32) with shapes [H,T] (batch-uniform) or [B,H,T], or (None, None) for the
    identity.  Raises DmError if fn is not an x-independent overwrite."""
    if fn is None or fn is apply_conditioning or fn is apply_conditioning_in_betweening \
            or fn is apply_conditioning_motion_blending:
        return None, None
    nan_probe, keep = _probes(tuple(shape), device)
    out = fn(nan_probe.clone())
    if not torch.is_tensor(out) or tuple(out.shape) != tuple(shape):
        raise DmError("conditioning_fn must return a tensor of the input shape")
    mask = ~torch.isnan(out)
    value = torch.where(mask, out, torch.zeros_like(out))
    # x-independence is a property of the function: verified with a second probe the first time this function
    # object is seen for this shape (its mask/values are re-read on every call above)
    key = (tuple(shape), str(device))
    try:
        seen = _VERIFIED.setdefault(fn, set())
    except TypeError:          # not weak-referenceable: verify every time
        seen = set()
    if key in seen:
        pure = torch.ones((), dtype=torch.bool, device=out.device)
    else:
        out2 = fn(keep.clone())
        pure = (out2 == torch.where(mask, value, keep)).all()
    # one device->host read for every property the host needs
    flags = torch.stack([pure, mask.any(), (mask == mask[:1]).all(), (value == value[:1]).all()]).cpu().tolist()
    if flags[0]:
        seen.add(key)
    if not flags[0]:
        raise DmError("conditioning_fn is not an x-independent overwrite x = where(mask, value, x); it cannot be "
                      "fused into the sampling kernels (no fallback)")
    if not flags[1]:
        return None, None
    if flags[2]:
        mask = mask[0]
    if flags[3] and mask.dim() == 2:
        value = value[0]
    elif mask.dim() == 2:
        mask = mask.unsqueeze(0).expand(shape).contiguous()
    return mask.to(torch.uint8).contiguous(), value.contiguous()
