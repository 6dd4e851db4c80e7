"""CPU ORACLE (test infrastructure, NOT a product path): float64 numpy restatement of the scoring half of the
hot path -- humanoid forward kinematics, DeepMimic tracking reward, motion-FID statistics.

PARITY UNPINNED for FK and reward: the reference delegates FK to MuJoCo 2.0 through mujoco-py 2.1.2.14
(mocap_player.py:67-73), which is neither vendored under /root/reference nor installed here, and it has no
reward implementation and no tests for either.  This file restates MuJoCo's published mj_kinematics semantics
on the reference's own model file (assets/dp_env_v2.xml:20-107) and fixes the reward formula of SURVEY.md
8(a) a15 (weights: utils/mocap_util.py:26-29).  FID statistics follow metrics/fid_score.py:24-81, which IS
importable; tests/golden/fid_reference.npz pins them against the unmodified reference.
"""
import numpy as np

# body order of the body_pos output == BODY_DEFS (utils/mocap_util.py:17-19)
BODY_DEFS = ["root", "chest", "neck", "right_hip", "right_knee", "right_ankle", "right_shoulder", "right_elbow",
             "right_wrist", "left_hip", "left_knee", "left_ankle", "left_shoulder", "left_elbow", "left_wrist"]


def qmul(a, b):
    w1, x1, y1, z1 = a
    w2, x2, y2, z2 = b
    return np.array([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2])


def qconj(q):
    return np.array([q[0], -q[1], -q[2], -q[3]])


def qmat(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def qaxis(axis, ang):
    axis = np.asarray(axis, dtype=np.float64)
    return np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * axis])


def qangle(q):
    return 2.0 * np.arctan2(np.linalg.norm(q[1:]), abs(q[0]))


def qlogvel(q, dt):
    """axis * angle / dt of a unit quaternion (utils/mocap_v2.py:155-178), shortest rotation."""
    vn = np.linalg.norm(q[1:])
    if vn < 1e-12:
        return np.zeros(3)
    w, s = (q[0], 1.0) if q[0] >= 0 else (-q[0], -1.0)
    return q[1:] * (2.0 * np.arctan2(vn, w) / dt * s / vn)


# ---- kinematic tree of assets/dp_env_v2.xml: (name, parent, body offset, [(qpos index, axis, anchor)]) ----
X, Y, Z, NY = (1, 0, 0), (0, 1, 0), (0, 0, 1), (0, -1, 0)


def _ball(i, y_anchor=(0, 0, 0)):
    return [(i, X, (0, 0, 0)), (i + 1, Y, y_anchor), (i + 2, Z, (0, 0, 0))]


TREE = [
    ("chest", "root", (0, 0, 0.236151), _ball(7)),
    ("neck", "chest", (0, 0, 0.223894), _ball(10)),
    ("right_shoulder", "chest", (-0.02405, -0.18311, 0.2435), _ball(13)),
    ("right_elbow", "right_shoulder", (0, 0, -0.274788), [(16, NY, (0.028, 0, 0))]),
    ("left_shoulder", "chest", (-0.02405, 0.18311, 0.2435), _ball(17)),
    ("left_elbow", "left_shoulder", (0, 0, -0.274788), [(20, NY, (0.028, 0, 0))]),
    ("right_hip", "root", (0, -0.094887, -0.05), _ball(21, (0.02, 0, 0))),
    ("right_knee", "right_hip", (0, 0, -0.421546), [(24, NY, (-0.035, 0, 0))]),
    ("right_ankle", "right_knee", (0, 0, -0.40987), _ball(25, (-0.02, 0, 0))),
    ("left_hip", "root", (0, 0.094887, -0.05), _ball(28, (0.02, 0, 0))),
    ("left_knee", "left_hip", (0, 0, -0.421546), [(31, NY, (-0.035, 0, 0))]),
    ("left_ankle", "left_knee", (0, 0, -0.40987), _ball(32, (-0.02, 0, 0))),
]
WRIST_LOCAL = (0, 0, -0.258947)  # geom right_wrist / left_wrist in the elbow body frame (dp_env_v2.xml:51,65)
# (mass, body, local centre) of every geom (dp_env_v2.xml geoms; capsule centre = midpoint of fromto)
GEOMS = [(2.5, "root", (0, 0, 0.07)), (3.0, "chest", (0, 0, 0.12)), (2.0, "neck", (0, 0, 0.175))]
for side in ("right", "left"):
    GEOMS += [(1.5, side + "_shoulder", (0, 0, -0.14)), (1.0, side + "_elbow", (0, 0, -0.12)),
              (0.5, side + "_elbow", WRIST_LOCAL), (4.5, side + "_hip", (0, 0, -0.21)),
              (3.0, side + "_knee", (0, 0, -0.2)), (1.0, side + "_ankle", (0.045, 0, -0.0425))]


def fk_frame(q):
    """mj_kinematics for one qpos[35] -> {body: (xpos[3], xquat[4])}."""
    q = np.asarray(q, dtype=np.float64)
    rq = q[3:7].copy()
    n = np.linalg.norm(rq)
    rq = np.array([1.0, 0, 0, 0]) if n < 1e-10 else rq / n
    out = {"root": (q[0:3].copy(), rq)}
    for name, parent, off, joints in TREE:
        ppos, pq = out[parent]
        pos = ppos + qmat(pq) @ np.asarray(off, dtype=np.float64)
        quat = pq.copy()
        for qi, axis, anchor in joints:
            anchor = np.asarray(anchor, dtype=np.float64)
            a_world = pos + qmat(quat) @ anchor
            quat = qmul(quat, qaxis(axis, q[qi]))
            pos = a_world - qmat(quat) @ anchor
        out[name] = (pos, quat)
    return out


def body_positions(qpos):
    """qpos [..., 35] -> [..., 15, 3] in BODY_DEFS order (wrists = elbow frame * WRIST_LOCAL)."""
    qpos = np.asarray(qpos, dtype=np.float64)
    flat = qpos.reshape(-1, 35)
    res = np.zeros((flat.shape[0], 15, 3))
    for i, q in enumerate(flat):
        fk = fk_frame(q)
        for k, name in enumerate(BODY_DEFS):
            if name.endswith("wrist"):
                p, o = fk[name.replace("wrist", "elbow")]
                res[i, k] = p + qmat(o) @ np.asarray(WRIST_LOCAL)
            else:
                res[i, k] = fk[name][0]
    return res.reshape(qpos.shape[:-1] + (15, 3))


J3 = [("chest", 7, 0.5), ("neck", 10, 0.3), ("right_shoulder", 13, 0.3), ("left_shoulder", 17, 0.3),
      ("right_hip", 21, 0.5), ("right_ankle", 25, 0.2), ("left_hip", 28, 0.5), ("left_ankle", 32, 0.2)]
J1 = [(16, 0.2), (20, 0.2), (24, 0.3), (31, 0.3)]
RATE_W = np.array([0.5] * 3 + [0.3] * 3 + [0.3] * 3 + [0.2] + [0.3] * 3 + [0.2] + [0.5] * 3 + [0.3] + [0.2] * 3 +
                  [0.5] * 3 + [0.3] + [0.2] * 3)


def frame_features(q, qp, dt):
    q = np.asarray(q, dtype=np.float64)
    qp = np.asarray(qp, dtype=np.float64)
    fk = fk_frame(q)
    rq = fk["root"][1]
    pq = qp[3:7] / np.linalg.norm(qp[3:7])
    f = {"rootq": rq, "rootw": qlogvel(qmul(qconj(pq), rq), dt)}
    f["jq"] = [qmul(qmul(qaxis(X, q[i]), qaxis(Y, q[i + 1])), qaxis(Z, q[i + 2])) for _, i, _ in J3]
    f["ja"] = [q[i] for i, _ in J1]
    f["rate"] = (q[7:35] - qp[7:35]) / dt
    rp = q[0:3]
    ee = []
    for name in ("right_ankle", "left_ankle"):
        ee.append(fk[name][0] - rp)
    for name in ("right_elbow", "left_elbow"):
        p, o = fk[name]
        ee.append(p + qmat(o) @ np.asarray(WRIST_LOCAL) - rp)
    f["ee"] = ee
    com, mass = np.zeros(3), 0.0
    for m, body, c in GEOMS:
        p, o = fk[body]
        com += m * (p + qmat(o) @ np.asarray(c, dtype=np.float64))
        mass += m
    f["com"] = com / mass
    return f


def tracking_reward(qpos, ref, dt):
    """qpos [N,H,35], ref [H,35] -> reward [N,H] and the four exp terms [N,H,4].
    r = 0.65 exp(-2 pose) + 0.1 exp(-0.1 vel) + 0.15 exp(-40 ee) + 0.1 exp(-10 com)   """
    qpos = np.asarray(qpos, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    N, H, _ = qpos.shape
    rf = [frame_features(ref[h], ref[max(h - 1, 0)], dt) for h in range(H)]
    r = np.zeros((N, H))
    terms = np.zeros((N, H, 4))
    for n in range(N):
        for h in range(H):
            f = frame_features(qpos[n, h], qpos[n, max(h - 1, 0)], dt)
            g = rf[h]
            pose = qangle(qmul(qconj(g["rootq"]), f["rootq"])) ** 2
            for k, (_, _, w) in enumerate(J3):
                pose += w * qangle(qmul(qconj(g["jq"][k]), f["jq"][k])) ** 2
            for k, (_, w) in enumerate(J1):
                pose += w * (f["ja"][k] - g["ja"][k]) ** 2
            vel = np.sum((f["rootw"] - g["rootw"]) ** 2) + np.sum(RATE_W * (f["rate"] - g["rate"]) ** 2)
            ee = sum(np.sum((a - b) ** 2) for a, b in zip(f["ee"], g["ee"]))
            com = np.sum((f["com"] - g["com"]) ** 2)
            t = np.array([np.exp(-2 * pose), np.exp(-0.1 * vel), np.exp(-40 * ee), np.exp(-10 * com)])
            terms[n, h] = t
            r[n, h] = 0.65 * t[0] + 0.1 * t[1] + 0.15 * t[2] + 0.1 * t[3]
    return r, terms


# ---- motion FID: metrics/fid_score.py:24-81 ----
def activation_statistics(data):
    """mu, sigma of flattened trajectories (fid_score.py:24-43), float64."""
    f = np.asarray(data, dtype=np.float64).reshape(len(data), -1)
    mu = f.mean(0)
    c = f - mu
    return mu, (c.T @ c) / (f.shape[0] - 1)


def fid_from_statistics(mu1, sigma1, mu2, sigma2, eps=1e-6):
    """The reference's SVD-based formula (fid_score.py:45-81) -- NOT the principal matrix square root."""
    diff = mu1 - mu2
    U, s, Vh = np.linalg.svd(sigma1 @ sigma2)
    covmean = U @ np.diag(np.sqrt(np.clip(s, eps, None))) @ Vh
    return float(diff @ diff + np.trace(sigma1) + np.trace(sigma2) - 2 * np.trace(covmean))


# ---- SiFID / diversity: eval.py:15-21, 54-89 ----
def slice_windows(sample, window_size, step_size=10):
    sample = np.asarray(sample, dtype=np.float64)
    offs = np.arange(sample.shape[0] - window_size + 1)[0::step_size]
    return np.stack([sample[o:o + window_size] for o in offs])


def sifid(gen_samples, gt_sample, window_size=10):
    gmu, gsig = activation_statistics(slice_windows(gt_sample, window_size))
    vals = []
    for x in np.asarray(gen_samples, dtype=np.float64):
        mu, sig = activation_statistics(slice_windows(x, window_size))
        vals.append(fid_from_statistics(gmu, gsig, mu, sig))
    return float(np.mean(vals))


def inter_diversity(samples):
    s = np.asarray(samples, dtype=np.float64)
    h = s.shape[0] // 2
    return float(np.linalg.norm(s[:h] - s[h:2 * h], axis=1).mean())


def intra_diversity(samples, offsets, window_size=10):
    s = np.asarray(samples, dtype=np.float64)
    return float(np.mean([np.linalg.norm(s[i, o[0]:o[0] + window_size] - s[i, o[1]:o[1] + window_size])
                          for i, o in enumerate(offsets)]))
