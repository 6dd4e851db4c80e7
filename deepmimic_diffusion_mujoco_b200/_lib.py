"""ctypes binding of libdm_b200.so (C ABI in include/dm_b200.h).

There is no CPU or PyTorch fallback: if the library is missing or a call fails, a DmError is raised.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# DM_B200_LIB: developer override used to A/B two builds of the same library on one GPU box
LIB_PATH = os.environ.get("DM_B200_LIB") or os.path.join(_HERE, "libdm_b200.so")
CSRC = os.path.join(_HERE, "csrc")

EXPORTS = [
    "dm_last_error", "dm_plan_create", "dm_plan_free", "dm_plan_set_schedule", "dm_unet_forward", "dm_ddpm_step",
    "dm_sample_loop", "dm_fk_reward", "dm_fk_reward_noalloc", "dm_fid_moments", "dm_window_moments", "dm_mocap_to_qpos", "dm_mocap_to_qvel", "dm_motion_dataset", "dm_plan_set_sample_offset", "dm_plan_num_layers", "dm_plan_layer_info",
    "dm_plan_read_layer", "dm_plan_ctas_per_group", "dm_plan_last_launch_count", "dm_device_error_code", "dm_plan_set_trace",
]


class DmError(RuntimeError):
    pass


class DmTensor(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("data", ctypes.c_void_p), ("ndim", ctypes.c_int),
                ("shape", ctypes.c_int64 * 4)]


class DmSchedule(ctypes.Structure):
    _fields_ = [("n_timesteps", ctypes.c_int), ("sqrt_recip_alphas_cumprod", ctypes.c_void_p),
                ("sqrt_recipm1_alphas_cumprod", ctypes.c_void_p), ("posterior_mean_coef1", ctypes.c_void_p),
                ("posterior_mean_coef2", ctypes.c_void_p), ("posterior_log_variance_clipped", ctypes.c_void_p),
                ("time_values", ctypes.c_void_p)]


class DmCond(ctypes.Structure):
    _fields_ = [("mask_dev", ctypes.c_void_p), ("value_dev", ctypes.c_void_p), ("mask_batched", ctypes.c_int),
                ("value_batched", ctypes.c_int), ("flags_dev", ctypes.c_void_p)]


def build(verbose=False):
    """Compile the CUDA sources for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise DmError("building libdm_b200.so failed")
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DmError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(there is no CPU fallback)")
    L = ctypes.CDLL(LIB_PATH)
    vp, ci, u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint64
    L.dm_last_error.restype = ctypes.c_char_p
    L.dm_plan_create.argtypes = [ctypes.POINTER(DmTensor), ci, ci, ci, ci, ci, ctypes.POINTER(ci), ci,
                                 ctypes.POINTER(vp)]
    L.dm_plan_free.argtypes = [vp]
    L.dm_plan_free.restype = None
    L.dm_plan_set_schedule.argtypes = [vp, ctypes.POINTER(DmSchedule)]
    L.dm_unet_forward.argtypes = [vp, vp, vp, vp, vp]
    L.dm_ddpm_step.argtypes = [vp, vp, vp, vp, u64, ci, ci, ci, ctypes.POINTER(DmCond), vp]
    L.dm_sample_loop.argtypes = [vp, vp, ci, vp, u64, ci, ci, ci, ctypes.POINTER(DmCond), vp, ci, vp]
    L.dm_fk_reward.argtypes = [vp, vp, ci, ci, ctypes.c_float, vp, vp, vp, vp]
    L.dm_fk_reward_noalloc.argtypes = [vp, vp, ci, ci, ctypes.c_float, vp, vp, vp, vp]
    L.dm_fid_moments.argtypes = [vp, ci, ci, vp, vp, vp]
    L.dm_window_moments.argtypes = [vp, ci, ci, ci, ci, ci, vp, vp, vp]
    L.dm_mocap_to_qpos.argtypes = [vp, ci, vp, vp]
    L.dm_mocap_to_qvel.argtypes = [vp, ci, vp, vp]
    L.dm_motion_dataset.argtypes = [vp, ci, vp, vp]
    L.dm_plan_set_sample_offset.argtypes = [vp, ctypes.c_longlong]
    L.dm_plan_num_layers.argtypes = [vp]
    L.dm_plan_layer_info.argtypes = [vp, ci, ctypes.c_char_p, ci, ctypes.POINTER(ci), ctypes.POINTER(ci)]
    L.dm_plan_read_layer.argtypes = [vp, ci, vp, vp]
    L.dm_plan_last_launch_count.argtypes = [vp]
    L.dm_plan_ctas_per_group.argtypes = [vp]
    L.dm_plan_set_trace.argtypes = [vp, vp, ci, ci]
    L.dm_device_error_code.argtypes = [ctypes.POINTER(ctypes.c_uint)]
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise DmError(f"libdm_b200 error {rc}: {lib().dm_last_error().decode()}")


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
