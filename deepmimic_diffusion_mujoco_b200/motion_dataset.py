"""Drop-in for data_loaders/motion_dataset.py:9-56 (`MotionDataset`, `Batch`) and the `.npy` motion writer.

Host-side boundary code, like the reference's: DeepMimic mocap JSON ("Frames": [duration, root pos 3, root quat 4,
12 joints as quaternions / hinge angles in DeepMimic order]) -> qpos rows [F, 35] in the MuJoCo humanoid's joint
order (root pos, root quat, then Euler xyz per 3-dof joint / angle per hinge), as utils/mocap_v2.py:105-291 does.
The arithmetic is vectorised over frames (no pyquaternion / transformations dependency):
  align_rotation  (utils/mocap_util.py:31-40)  q' = qL * q * qR with qL/qR = +/-90 deg about x
  align_position  (utils/mocap_util.py:42-48)  (x, y, z) -> (x, -z, y)
  euler 'rxyz'    (utils/transformations.py:1031-1097) from the rotation matrix of the aligned quaternion
"""
import json
from collections import namedtuple

import numpy as np
import torch

Batch = namedtuple("Batch", "trajectories conditions")

_BODY_JOINTS = ["chest", "neck", "right_shoulder", "right_elbow", "left_shoulder", "left_elbow", "right_hip",
                "right_knee", "right_ankle", "left_hip", "left_knee", "left_ankle"]
_DP_ORDER = ["chest", "neck", "right_hip", "right_knee", "right_ankle", "right_shoulder", "right_elbow", "left_hip",
             "left_knee", "left_ankle", "left_shoulder", "left_elbow"]
_HINGES = {"right_elbow", "left_elbow", "right_knee", "left_knee"}
_R = np.sqrt(0.5)


def _qmul(a, b):
    aw, ax, ay, az = (a[..., i] for i in range(4))
    bw, bx, by, bz = (b[..., i] for i in range(4))
    return np.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw], axis=-1)


def _align_rotation(q):
    left = np.array([_R, _R, 0.0, 0.0])      # +90 deg about x
    right = np.array([_R, -_R, 0.0, 0.0])    # -90 deg about x
    return _qmul(_qmul(np.broadcast_to(left, q.shape), q), np.broadcast_to(right, q.shape))


def _euler_rxyz(q):
    """q: [..., 4] (w, x, y, z) -> intrinsic x-y-z Euler angles [..., 3]."""
    n = np.sum(q * q, axis=-1, keepdims=True)
    w, x, y, z = ((q * np.sqrt(2.0 / n))[..., i] for i in range(4))
    m00 = 1.0 - y * y - z * z
    m01, m02 = x * y - z * w, x * z + y * w
    m12, m22 = y * z - x * w, 1.0 - x * x - y * y
    m11 = 1.0 - x * x - z * z
    m10 = x * y + z * w
    cy = np.sqrt(m22 * m22 + m12 * m12)
    ok = cy > np.finfo(float).eps * 4.0
    ax = np.where(ok, np.arctan2(m01, m00), np.arctan2(-m10, m11))
    ay = np.arctan2(-m02, cy)
    az = np.where(ok, np.arctan2(m12, m22), 0.0)
    return np.stack([-az, -ay, -ax], axis=-1)


def mocap_to_qpos(frames):
    """[F, 44] DeepMimic frames -> [F, 35] float64 qpos rows (`MocapDM.data_config`)."""
    fr = np.asarray(frames, dtype=np.float64)
    cols = [np.stack([fr[:, 1], -fr[:, 3], fr[:, 2]], axis=-1), _align_rotation(fr[:, 4:8])]
    joint, off = {}, 8
    for name in _DP_ORDER:
        width = 1 if name in _HINGES else 4
        joint[name] = fr[:, off:off + width]
        off += width
    for name in _BODY_JOINTS:
        cols.append(joint[name] if name in _HINGES else _euler_rxyz(_align_rotation(joint[name])))
    return np.concatenate(cols, axis=-1)


def _rot_vel(q_now, q_prev, dura):
    """utils/mocap_v2.py:155-178 (calc_rot_vel(data[k], data[k-1], dura)), vectorised: q_diff = conj(q_now) * q_prev,
    velocity = angle / dura * axis with pyquaternion's .axis / .angle semantics (normalise unless already unit to
    1e-14, axis = vector / |vector| or zeros below 1e-17, angle = 2 atan2(|vector|, scalar) wrapped into (-pi, pi])."""
    conj = q_now * np.array([1.0, -1.0, -1.0, -1.0])
    qd = _qmul(conj, q_prev)
    ss = np.sum(qd * qd, axis=-1, keepdims=True)
    unit = np.abs(1.0 - ss) < 1e-14
    n = np.sqrt(ss)
    qd = np.where(unit | (n <= 0), qd, qd / np.where(n > 0, n, 1.0))
    nv = np.sqrt(qd[..., 1] ** 2 + qd[..., 2] ** 2 + qd[..., 3] ** 2)
    axis = np.where((nv < 1e-17)[..., None], 0.0, qd[..., 1:] / np.where(nv < 1e-17, 1.0, nv)[..., None])
    theta = 2.0 * np.arctan2(nv, qd[..., 0])
    ang = np.fmod(theta + np.pi, 2 * np.pi) - np.pi
    ang = np.where(ang == -np.pi, np.pi, ang)
    return (ang / dura)[..., None] * axis


def mocap_to_qvel(frames):
    """[F, 44] DeepMimic frames -> [F, 34] float64 velocity rows (`MocapDM.data_vel`, utils/mocap_v2.py:196-291): root
    linear 3, root angular 3, then per joint in MuJoCo order an angular velocity (3) or a hinge rate (1).  Row 0 is zero;
    row k differences frame k against frame k-1 over durations[k-1] (0.0167 when that is zero)."""
    fr = np.asarray(frames, dtype=np.float64)
    F = fr.shape[0]
    out = np.zeros((F, 34), dtype=np.float64)
    if F < 2:
        return out
    dura = fr[:-1, 0].copy()
    dura[dura == 0] = 0.0167
    pos = np.stack([fr[:, 1], -fr[:, 3], fr[:, 2]], axis=-1)
    cols = [(pos[1:] - pos[:-1]) * 1.0 / dura[:, None]]
    rq = _align_rotation(fr[:, 4:8])
    cols.append(_rot_vel(rq[1:], rq[:-1], dura))
    joint, off = {}, 8
    for name in _DP_ORDER:
        width = 1 if name in _HINGES else 4
        joint[name] = fr[:, off:off + width]
        off += width
    for name in _BODY_JOINTS:
        if name in _HINGES:
            cols.append((joint[name][1:] - joint[name][:-1]) * 1.0 / dura[:, None])
        else:
            q = _align_rotation(joint[name])
            cols.append(_rot_vel(q[1:], q[:-1], dura))
    out[1:] = np.concatenate(cols, axis=-1)
    return out


def load_mocap_frames(path):
    with open(path, "r") as f:
        return np.array(json.load(f)["Frames"], dtype=np.float64)


class MotionDataset(torch.utils.data.Dataset):
    """motion_dataset.py:12-56: every cyclic shift of the clip (truncated to a multiple of 8 frames)."""

    def __init__(self, motion_src_path, device=None):
        """device=None: host tensors like the reference.  device='cuda': the JSON is parsed on the host (it is text),
        the frame conversion and the [n, n, 35] tensor of cyclic shifts are built by CUDA kernels (dm_mocap_to_qpos,
        dm_motion_dataset) and stay on the device, ready to be passed as starting_motion / conditioning values."""
        frames = load_mocap_frames(motion_src_path)
        self.dt = float(frames[0, 0])
        n = frames.shape[0] - frames.shape[0] % 8      # temporal unet needs a multiple of 8 (motion_dataset.py:26-28)
        if device is not None and torch.device(device).type == "cuda":
            self.motion_data = mocap_dataset_on_device(frames[:n], torch.device(device))
            return
        cfg = torch.from_numpy(mocap_to_qpos(frames)[:n]).float()
        idx = (torch.arange(n)[:, None] + torch.arange(n)[None, :]) % n
        self.motion_data = cfg[idx]                    # [n, n, 35]: row i = cat(cfg[i:], cfg[:i])

    def __len__(self):
        return len(self.motion_data)

    def __getitem__(self, idx):
        motion = self.motion_data[idx]
        return Batch(trajectories=motion, conditions=self.get_conditions(motion))

    def get_conditions(self, data_configs):
        return {0: data_configs[0]}


BatchV2 = namedtuple("BatchV2", "trajectories motion_class")

MOTION_CLASSES = {"humanoid3d_walk": 0, "humanoid3d_run": 1, "humanoid3d_spinkick": 2, "humanoid3d_roll": 3,
                  "humanoid3d_dance_a": 4, "humanoid3d_dance_b": 5, "humanoid3d_jump": 6, "humanoid3d_cartwheel": 7,
                  "humanoid3d_backflip": 8}      # data_loaders/motion_dataset_v2.py:11-21


class MotionDatasetV2(torch.utils.data.Dataset):
    """Drop-in for data_loaders/motion_dataset_v2.py:30-136, the loader of the 69-wide (qpos | qvel) models
    (eval.py:189-197, TemporalUnet dim=128): every item is a [F, 69] float32 trajectory plus its motion class.
    shuffle=False: the clip itself, 1000 times (:104-107); shuffle=True: every cyclic shift with the root translation
    kept continuous across the wrap (:85-103).  device='cuda': rows are computed by dm_mocap_to_qpos / dm_mocap_to_qvel
    and stay on the device.  One file per dataset (a directory keeps its first two .txt files, as the reference does)."""

    def __init__(self, motion_src_path, shuffle=False, device=None):
        import os
        if not motion_src_path.endswith(".txt"):
            files = [os.path.join(motion_src_path, f) for f in os.listdir(motion_src_path) if f.endswith(".txt")][:2]
        else:
            files = [motion_src_path]
        self.motion_data, classes = [], []
        self.max_sequence_length = 0
        on_dev = device is not None and torch.device(device).type == "cuda"
        for path in files:
            cls = MOTION_CLASSES[path.split("/")[-1].split(".")[0]]
            frames = load_mocap_frames(path)
            if on_dev:
                combined = torch.cat([mocap_qpos_on_device(frames, torch.device(device)),
                                      mocap_qvel_on_device(frames, torch.device(device))], dim=1)
            else:
                combined = torch.from_numpy(np.concatenate([mocap_to_qpos(frames), mocap_to_qvel(frames)], axis=1))
            F = combined.shape[0]
            self.max_sequence_length = max(self.max_sequence_length, F)
            if shuffle:
                c = combined.double() if not on_dev else combined
                diff = c[-1] - c[0]
                for i in range(F):
                    prefix, suffix = c[i:].clone(), c[:i].clone()
                    if i != 0 and i != F - 1:
                        suffix[:, :3] += diff[:3]
                        first = prefix[0] - c[0]
                        prefix[:, :3] -= first[:3]
                        suffix[:, :3] -= first[:3]
                    self.motion_data.append(torch.cat([prefix.float(), suffix.float()], dim=0))
                    classes.append(cls)
            else:
                row = combined.float()
                self.motion_data += [row] * 1000
                classes += [cls] * 1000
        self.motion_classes = torch.tensor(classes)

    def __len__(self):
        return len(self.motion_classes)

    def __getitem__(self, idx):
        return BatchV2(trajectories=self.motion_data[idx], motion_class=self.motion_classes[idx])

    def collate_fn(self, batch):
        """motion_dataset_v2.py:116-130: a jagged nested tensor of the batch's trajectories, stacked classes."""
        traj = torch.nested.nested_tensor([b.trajectories for b in batch], layout=torch.jagged)
        return BatchV2(trajectories=traj, motion_class=torch.stack([b.motion_class for b in batch]))

    def get_conditions(self, data_configs):
        return {0: data_configs[0]}


def mocap_qvel_on_device(frames, device):
    """[F, 44] DeepMimic frames (host array) -> qvel fp32 [F, 34] on `device`, computed by dm_mocap_to_qvel."""
    from . import _lib
    fr = torch.as_tensor(np.ascontiguousarray(frames, dtype=np.float64)).to(device)
    out = torch.empty(fr.shape[0], 34, device=device, dtype=torch.float32)
    with torch.cuda.device(device):
        _lib.check(_lib.lib().dm_mocap_to_qvel(fr.data_ptr(), fr.shape[0], out.data_ptr(), _lib.stream_ptr()))
    return out


def mocap_qpos_on_device(frames, device):
    """[F, 44] DeepMimic frames (host array) -> qpos fp32 [F, 35] on `device`, computed by dm_mocap_to_qpos."""
    from . import _lib
    fr = torch.as_tensor(np.ascontiguousarray(frames, dtype=np.float64)).to(device)
    out = torch.empty(fr.shape[0], 35, device=device, dtype=torch.float32)
    with torch.cuda.device(device):
        _lib.check(_lib.lib().dm_mocap_to_qpos(fr.data_ptr(), fr.shape[0], out.data_ptr(), _lib.stream_ptr()))
    return out


def mocap_dataset_on_device(frames, device):
    """All cyclic shifts of the clip, fp32 [n, n, 35] on `device` (dm_mocap_to_qpos + dm_motion_dataset)."""
    from . import _lib
    q = mocap_qpos_on_device(frames, device)
    n = q.shape[0]
    out = torch.empty(n, n, 35, device=device, dtype=torch.float32)
    with torch.cuda.device(device):
        _lib.check(_lib.lib().dm_motion_dataset(q.data_ptr(), n, out.data_ptr(), _lib.stream_ptr()))
    return out


def tile_clip(clip, horizon):
    """[F, 35] -> [horizon, 35] by cycling the clip, the way the dataset's shifted rows continue it."""
    clip = torch.as_tensor(clip)
    reps = -(-horizon // clip.shape[0])
    return clip.repeat(reps, 1)[:horizon].contiguous()


def save_motion_npy(path, trajectory):
    """The `.npy` contract of the notebooks (float32 [H, 35]) that mocap_player.play_mocap_np_file (:51) loads."""
    t = torch.as_tensor(trajectory).detach().float().cpu().numpy()
    if t.ndim != 2 or t.shape[1] != 35:
        raise ValueError("expected one trajectory [H, 35]")
    np.save(path, t.astype(np.float32))


def save_motions(samples, output_dir, filenames=None, joint_indices=None):
    """Bulk `.npy` export, the contract of transformer_inference.py:213-276 (`save_motions`): one float `[frames, 35]`
    file per sample (first 35 features unless `joint_indices` says otherwise; padded with zeros / truncated to 35).
    The whole sample set crosses PCIe ONCE (one device-to-host copy into pinned memory) instead of once per sample.
    Returns the absolute paths."""
    import os
    os.makedirs(output_dir, exist_ok=True)
    s = torch.as_tensor(samples).detach()
    if s.dim() != 3:
        raise ValueError("expected samples [N, frames, features]")
    idx = list(range(35)) if joint_indices is None else list(joint_indices)
    if len(idx) != 35:   # reference: "MuJoCo expects exactly 35 joints"
        idx = idx[:35] if len(idx) > 35 else idx + list(range(len(idx), 35))
    idx = [i for i in idx if i < s.shape[2]]
    sel = s[:, :, idx].float()
    if s.is_cuda:
        host = torch.empty(sel.shape, dtype=torch.float32, pin_memory=True)
        host.copy_(sel, non_blocking=True)
        torch.cuda.current_stream(s.device).synchronize()
    else:
        host = sel.contiguous()
    arr = host.numpy()
    if arr.shape[2] < 35:
        arr = np.concatenate([arr, np.zeros(arr.shape[:2] + (35 - arr.shape[2],), dtype=arr.dtype)], axis=2)
    names = list(filenames or [])
    names += [f"motion_{i}.npy" for i in range(len(names), arr.shape[0])]
    out = []
    for i in range(arr.shape[0]):
        path = os.path.join(output_dir, names[i])
        np.save(path, arr[i])
        out.append(os.path.abspath(path))
    return out
