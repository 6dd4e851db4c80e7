"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy float64, plain loops) of the reference's mocap loader.

Follows, line by line:
  utils/mocap_util.py:31-48      align_rotation / align_position
  utils/mocap_v2.py:105-153      read_raw_data_from_file (frame parse, DeepMimic joint order)
  utils/mocap_v2.py:180-291      convert_raw_data (emit in BODY_JOINTS order, quaternion -> Euler 'rxyz'; data_vel:
                                 finite differences and calc_rot_vel :155-178)
  data_loaders/motion_dataset.py:13-56   MotionDataset (truncate to a multiple of 8, cyclic shifts)
(all under /root/reference/diffusion/).

Third-party arithmetic restated here because the dependency is absent from /root/reference and from this image:
  pyquaternion 0.9.9 (environment.yml:107): Quaternion(matrix=...) trace method, Hamilton product, conjugate,
  .axis (get_axis: normalise, vector / |vector|, zeros below 1e-17) and .angle (normalise, 2 atan2(|vector|, scalar)
  wrapped into (-pi, pi]);
  transformations.py (vendored in the reference, utils/transformations.py:1031-1097,1174-1193): quaternion_matrix,
  euler_from_matrix(axes='rxyz').

PARITY UNPINNED: the reference has no test or golden vector for the loader and pyquaternion cannot be imported
here.  Weak anchors checked by tests/test_loader.py: the notebook output of data_config[0] for the cartwheel clip
(archive/0_temporal_unet_diffusion.ipynb cell 3) and the reference's own transformations.euler_from_quaternion,
imported when tests/golden/make_loader_golden.py generated the fixture.
"""
import json
import math

import numpy as np

BODY_JOINTS = ["chest", "neck", "right_shoulder", "right_elbow", "left_shoulder", "left_elbow", "right_hip",
               "right_knee", "right_ankle", "left_hip", "left_knee", "left_ankle"]               # mocap_util.py:5-7
BODY_JOINTS_IN_DP_ORDER = ["chest", "neck", "right_hip", "right_knee", "right_ankle", "right_shoulder",
                           "right_elbow", "left_hip", "left_knee", "left_ankle", "left_shoulder",
                           "left_elbow"]                                                           # mocap_util.py:9-11
DOF_DEF = {"root": 3, "chest": 3, "neck": 3, "right_shoulder": 3, "right_elbow": 1, "right_wrist": 0,
           "left_shoulder": 3, "left_elbow": 1, "left_wrist": 0, "right_hip": 3, "right_knee": 1, "right_ankle": 3,
           "left_hip": 3, "left_knee": 1, "left_ankle": 3}                                         # mocap_util.py:13-16


def quat_from_matrix(m):
    """pyquaternion 0.9.9 Quaternion._from_matrix (trace method); returns (w, x, y, z)."""
    m = np.asarray(m, dtype=np.float64).T
    if m[2, 2] < 0:
        if m[0, 0] > m[1, 1]:
            t = 1 + m[0, 0] - m[1, 1] - m[2, 2]
            q = [m[1, 2] - m[2, 1], t, m[0, 1] + m[1, 0], m[2, 0] + m[0, 2]]
        else:
            t = 1 - m[0, 0] + m[1, 1] - m[2, 2]
            q = [m[2, 0] - m[0, 2], m[0, 1] + m[1, 0], t, m[1, 2] + m[2, 1]]
    else:
        if m[0, 0] < -m[1, 1]:
            t = 1 - m[0, 0] - m[1, 1] + m[2, 2]
            q = [m[0, 1] - m[1, 0], m[2, 0] + m[0, 2], m[1, 2] + m[2, 1], t]
        else:
            t = 1 + m[0, 0] + m[1, 1] + m[2, 2]
            q = [t, m[1, 2] - m[2, 1], m[2, 0] - m[0, 2], m[0, 1] - m[1, 0]]
    return np.array(q, dtype=np.float64) * (0.5 / math.sqrt(t))


def quat_mul(a, b):
    """Hamilton product, (w, x, y, z)."""
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([aw * bw - ax * bx - ay * by - az * bz,
                     aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw])


Q_ALIGN_RIGHT = quat_from_matrix([[1.0, 0.0, 0.0], [0.0, 0.0, 1.0], [0.0, -1.0, 0.0]])   # mocap_util.py:33-35
Q_ALIGN_LEFT = quat_from_matrix([[1.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])    # mocap_util.py:36-38


def align_rotation(rot):
    """mocap_util.py:31-40: q_out = q_left * q * q_right."""
    return quat_mul(quat_mul(Q_ALIGN_LEFT, np.asarray(rot, dtype=np.float64)), Q_ALIGN_RIGHT)


def align_position(pos):
    """mocap_util.py:42-48: (x, y, z) -> (x, -z, y)."""
    return np.array([pos[0], -pos[2], pos[1]], dtype=np.float64)


def quaternion_matrix_xyzw(q):
    """transformations.quaternion_matrix for [x, y, z, w] input (the order the loader passes, mocap_v2.py:277)."""
    q = np.array(q, dtype=np.float64)
    n = float(np.dot(q, q))
    if n < np.finfo(float).eps * 4.0:
        return np.identity(3)
    q = q * math.sqrt(2.0 / n)
    o = np.outer(q, q)
    return np.array([[1.0 - o[1, 1] - o[2, 2], o[0, 1] - o[2, 3], o[0, 2] + o[1, 3]],
                     [o[0, 1] + o[2, 3], 1.0 - o[0, 0] - o[2, 2], o[1, 2] - o[0, 3]],
                     [o[0, 2] - o[1, 3], o[1, 2] + o[0, 3], 1.0 - o[0, 0] - o[1, 1]]])


def euler_rxyz_from_matrix(M):
    """transformations.euler_from_matrix(M, 'rxyz'): firstaxis 2, parity 1, no repetition, rotating frame."""
    i, j, k = 2, 1, 0
    cy = math.sqrt(M[i, i] * M[i, i] + M[j, i] * M[j, i])
    if cy > np.finfo(float).eps * 4.0:
        ax = math.atan2(M[k, j], M[k, k])
        ay = math.atan2(-M[k, i], cy)
        az = math.atan2(M[j, i], M[i, i])
    else:
        ax = math.atan2(-M[j, k], M[j, j])
        ay = math.atan2(-M[k, i], cy)
        az = 0.0
    ax, ay, az = -ax, -ay, -az      # parity
    return az, ay, ax               # rotating frame: swap first and last


def data_config_from_frames(frames):
    """frames: [F, 44] DeepMimic rows (duration, root pos 3, root quat 4, joints in DP order).
    Returns data_config [F, 35] float64 (mocap_v2.py:196-283, the tmp_angle list)."""
    out = []
    for fr in np.asarray(frames, dtype=np.float64):
        state = {"root_pos": align_position(fr[1:4]), "root_rot": align_rotation(fr[4:8])}
        off = 8
        for name in BODY_JOINTS_IN_DP_ORDER:
            if DOF_DEF[name] == 1:
                state[name] = fr[off:off + 1]
                off += 1
            else:
                state[name] = align_rotation(fr[off:off + 4])
                off += 4
        row = list(state["root_pos"]) + list(state["root_rot"])
        for name in BODY_JOINTS:
            if DOF_DEF[name] == 1:
                row += list(state[name])
            else:
                q = state[name]
                row += list(euler_rxyz_from_matrix(quaternion_matrix_xyzw([q[1], q[2], q[3], q[0]])))
        out.append(row)
    return np.array(out, dtype=np.float64)


def _normalise(q):
    """pyquaternion Quaternion._normalise: divide by the norm unless |1 - sum of squares| < 1e-14 (is_unit)."""
    ss = float(np.dot(q, q))
    if abs(1.0 - ss) < 1e-14:
        return q
    n = math.sqrt(ss)
    return q / n if n > 0 else q


def _wrap_angle(theta):
    """pyquaternion Quaternion._wrap_angle: into (-pi, pi]."""
    r = ((theta + math.pi) % (2 * math.pi)) - math.pi
    return math.pi if r == -math.pi else r


def calc_rot_vel(seg_0, seg_1, dura):
    """mocap_v2.py:155-178: q_diff = conj(q_0) * q_1; angular velocity = angle / dura * axis (pyquaternion .axis, .angle)."""
    q0 = np.asarray(seg_0, dtype=np.float64)
    q1 = np.asarray(seg_1, dtype=np.float64)
    qd = _normalise(quat_mul(np.array([q0[0], -q0[1], -q0[2], -q0[3]]), q1))
    nv = math.sqrt(qd[1] * qd[1] + qd[2] * qd[2] + qd[3] * qd[3])       # np.linalg.norm(self.vector)
    axis = np.zeros(3) if nv < 1e-17 else qd[1:] / nv
    angle = _wrap_angle(2.0 * math.atan2(nv, qd[0]))
    return angle / dura * axis


def data_vel_from_frames(frames):
    """frames: [F, 44] DeepMimic rows.  Returns data_vel [F, 34] float64 (mocap_v2.py:196-291, the tmp_vel list):
    root linear velocity 3, root angular velocity 3, then per joint in BODY_JOINTS order the angular velocity (3) of a
    3-dof joint or the hinge rate (1).  Frame 0 is all zeros; frame k differences against frame k-1 over durations[k-1]
    (0.0167 if that is zero); the rotation difference is taken from frame k TO frame k-1, as the reference does
    (calc_rot_vel(data[k], data[k-1], dura), :230-234 and :266-270)."""
    fr = np.asarray(frames, dtype=np.float64)
    states = []
    for row in fr:
        st = {"root_pos": align_position(row[1:4]), "root_rot": align_rotation(row[4:8])}
        off = 8
        for name in BODY_JOINTS_IN_DP_ORDER:
            if DOF_DEF[name] == 1:
                st[name] = row[off:off + 1].copy()
                off += 1
            else:
                st[name] = align_rotation(row[off:off + 4])
                off += 4
        states.append(st)
    durations = fr[:, 0]
    out = []
    for k, st in enumerate(states):
        dura = durations[k] if k == 0 else durations[k - 1]
        if dura == 0:
            dura = 0.0167
        if k == 0:
            out.append(np.zeros(34))
            continue
        pv = states[k - 1]
        v = list((st["root_pos"] - pv["root_pos"]) * 1.0 / dura)
        v += list(calc_rot_vel(st["root_rot"], pv["root_rot"], dura))
        for name in BODY_JOINTS:
            if DOF_DEF[name] == 1:
                v += list((st[name] - pv[name]) * 1.0 / dura)
            else:
                v += list(calc_rot_vel(st[name], pv[name], dura))
        out.append(np.array(v))
    return np.array(out, dtype=np.float64)


def load_frames(path):
    with open(path, "r") as f:
        return np.array(json.load(f)["Frames"], dtype=np.float64)


def motion_dataset_v2_rows(frames):
    """data_loaders/motion_dataset_v2.py:66-83: combined = [data_config | data_vel] -> [F, 69] float32 (no truncation)."""
    return np.concatenate([data_config_from_frames(frames), data_vel_from_frames(frames)], axis=1).astype(np.float32)


def motion_dataset_v2_shuffled(frames):
    """motion_dataset_v2.py:85-103 (shuffle=True): row i = clip rolled by i frames with the root translation made
    continuous across the wrap and re-based to the clip's first frame; rows 0 and F-1 are plain rolls."""
    combined = np.concatenate([data_config_from_frames(frames), data_vel_from_frames(frames)], axis=1)
    diff = combined[-1] - combined[0]
    rows = []
    for i in range(len(combined)):
        prefix = combined[i:].copy()
        suffix = combined[:i].copy()
        if i != 0 and i != len(combined) - 1:
            suffix[:, :3] += diff[:3]
            first_diff = prefix[0] - combined[0]
            prefix[:, :3] -= first_diff[:3]
            suffix[:, :3] -= first_diff[:3]
        rows.append(np.concatenate([prefix.astype(np.float32), suffix.astype(np.float32)], axis=0))
    return np.stack(rows)


def motion_dataset(frames):
    """MotionDataset.__init__ (motion_dataset.py:13-42): [F8, F8, 35] float32, row i = clip rolled by i frames."""
    cfg = data_config_from_frames(frames)
    n = cfg.shape[0] - cfg.shape[0] % 8
    cfg = cfg[:n]
    return np.stack([np.concatenate([cfg[i:], cfg[:i]], axis=0) for i in range(n)]).astype(np.float32)
