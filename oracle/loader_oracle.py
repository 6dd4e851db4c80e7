"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy float64, plain loops) of the reference's mocap loader.

Follows, line by line:
  utils/mocap_util.py:31-48      align_rotation / align_position
  utils/mocap_v2.py:105-153      read_raw_data_from_file (frame parse, DeepMimic joint order)
  utils/mocap_v2.py:180-291      convert_raw_data (emit in BODY_JOINTS order, quaternion -> Euler 'rxyz')
  data_loaders/motion_dataset.py:13-56   MotionDataset (truncate to a multiple of 8, cyclic shifts)
(all under /root/reference/diffusion/).

Third-party arithmetic restated here because the dependency is absent from /root/reference and from this image:
  pyquaternion 0.9.9 (environment.yml:107): Quaternion(matrix=...) trace method, Hamilton product;
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


def load_frames(path):
    with open(path, "r") as f:
        return np.array(json.load(f)["Frames"], dtype=np.float64)


def motion_dataset(frames):
    """MotionDataset.__init__ (motion_dataset.py:13-42): [F8, F8, 35] float32, row i = clip rolled by i frames."""
    cfg = data_config_from_frames(frames)
    n = cfg.shape[0] - cfg.shape[0] % 8
    cfg = cfg[:n]
    return np.stack([np.concatenate([cfg[i:], cfg[:i]], axis=0) for i in range(n)]).astype(np.float32)
