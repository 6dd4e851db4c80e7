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
# conditioning function -> {(shape, device): (mask is batch-uniform, value is batch-uniform)}: the STRUCTURE of a function
# verified once as a pure overwrite.  Values (and which entries are masked) are re-read on every call; only the
# host-side decisions are cached, so that a sampling call needs no device->host round trip.  A round trip per call
# serialises the host behind the previous 50 ms sampling kernel (measured: 2 % of the pass).
_STRUCTURE = weakref.WeakKeyDictionary()
_PENDING = []   # (pinned flags, event, fn name, expected): deferred re-verification of the cached structure


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


def check_pending(block=False):
    """Deferred half of the tracer: flags computed on the device during earlier calls are compared with the cached
    structure once they have arrived (never waits unless block=True).  A mismatch means an earlier sampling call
    used a stale [H,T]-vs-[B,H,T] decision: the cache entry is dropped and the error is raised now, loudly."""
    keep = []
    for flags, ev, name, expect, fn_ref, key in _PENDING:
        if not block and ev is not None and not ev.query():
            keep.append((flags, ev, name, expect, fn_ref, key))
            continue
        if ev is not None:
            ev.synchronize()
        got = tuple(bool(v) for v in flags.tolist())
        if got != expect:
            fn = fn_ref()
            if fn is not None and fn in _STRUCTURE:
                _STRUCTURE[fn].pop(key, None)
            _PENDING[:] = keep
            raise DmError(f"conditioning_fn {name} changed its structure (any / batch-uniform mask / batch-uniform "
                          f"value = {got}, was {expect}) between calls; the previous sampling call used the old one")
    _PENDING[:] = keep


def _finish(mask, value, shape, any_, mask_uniform, value_uniform):
    if not any_:
        return None, None
    if mask_uniform:
        mask = mask[0]
    if value_uniform and mask.dim() == 2:
        value = value[0]
    elif mask.dim() == 2:
        mask = mask.unsqueeze(0).expand(shape).contiguous()
    return mask.to(torch.uint8).contiguous(), value.contiguous()


def trace_conditioning(fn, shape, device):
    """Returns (mask uint8, value fp32) with shapes [H,T] (batch-uniform) or [B,H,T], or (None, None) for the
    identity.  Raises DmError if fn is not an x-independent overwrite."""
    if fn is None or fn is apply_conditioning or fn is apply_conditioning_in_betweening \
            or fn is apply_conditioning_motion_blending:
        return None, None
    check_pending()
    shape = tuple(shape)
    nan_probe, keep = _probes(shape, device)
    out = fn(nan_probe.clone())
    if not torch.is_tensor(out) or tuple(out.shape) != shape:
        raise DmError("conditioning_fn must return a tensor of the input shape")
    mask = ~torch.isnan(out)
    value = torch.where(mask, out, torch.zeros_like(out))
    key = (shape, str(device))
    try:
        known = _STRUCTURE.setdefault(fn, {})
    except TypeError:          # not weak-referenceable: verify every time
        known = {}
    flags_dev = torch.stack([mask.any(), (mask == mask[:1]).all(), (value == value[:1]).all()])
    if key in known and out.is_cuda:
        # fast path: no host round trip.  The flags travel to pinned memory behind the kernels above and are compared
        # with the cached structure by a later call (check_pending).
        host = torch.empty(3, dtype=torch.bool, pin_memory=True)
        host.copy_(flags_dev, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(out.device))
        if len(_PENDING) < 64:
            _PENDING.append((host, ev, getattr(fn, "__name__", repr(fn)), (True,) + known[key], weakref.ref(fn), key))
        return _finish(mask, value, shape, True, *known[key])
    # first sight of this function for this shape: x-independence is verified with a second probe, and every
    # property the host needs comes back in ONE device->host read
    out2 = fn(keep.clone())
    pure = (out2 == torch.where(mask, value, keep)).all()
    flags = torch.cat([pure.reshape(1), flags_dev]).cpu().tolist()
    if not flags[0]:
        raise DmError("conditioning_fn is not an x-independent overwrite x = where(mask, value, x); it cannot be "
                      "fused into the sampling kernels (no fallback)")
    if flags[1]:
        known[key] = (bool(flags[2]), bool(flags[3]))   # functions that overwrite nothing are re-verified every call
    return _finish(mask, value, shape, flags[1], flags[2], flags[3])
