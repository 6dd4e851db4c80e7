"""CPU: the kinematic tables behind K5 against a fixture PARSED from the reference's MuJoCo model
(assets/dp_env_v2.xml -> tests/golden/fk_tree.json by tests/golden/make_fk_tree.py).

MuJoCo itself is absent, so FK parity stays "unpinned" against mj_kinematics; what these tests pin is
everything short of that: the hand-written tables of oracle/scoring_oracle.py equal the model file, a table-free FK
driven ONLY by the fixture agrees with the oracle and with the product arithmetic (csrc/fk_math.cuh compiled for the
host), and hand-derived poses give the positions the XML numbers imply."""
import json
import os

import numpy as np
import pytest

from oracle import scoring_oracle as SO
from test_fk_host import harness, run  # noqa: F401  (fixture + runner of the host-compiled product arithmetic)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def tree():
    return json.load(open(os.path.join(ROOT, "tests", "golden", "fk_tree.json")))


def rot(axis, ang):
    """Rodrigues rotation matrix (independent of the oracle's quaternion helpers)."""
    a = np.asarray(axis, dtype=np.float64)
    a = a / np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * (K @ K)


def quat_matrix(q):
    w, x, y, z = np.asarray(q, dtype=np.float64) / np.linalg.norm(q)
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def fixture_fk(tree, q):
    """mj_kinematics semantics from the fixture alone: child frame = parent frame * body pos; every hinge, in document
    order, rotates the body about its axis through its anchor (both in the body frame).  Returns {body: (pos, R)},
    the centre of mass and the wrist geom centres."""
    q = np.asarray(q, dtype=np.float64)
    out = {}
    for b in tree["bodies"]:
        if b["parent"] is None:
            pos, R = q[0:3].copy(), quat_matrix(q[3:7])       # free joint: qpos replaces the body's frame
        else:
            ppos, pR = out[b["parent"]]
            pos, R = ppos + pR @ np.asarray(b["pos"]), pR.copy()
            for j in b["joints"]:
                anchor = np.asarray(j["pos"])
                world_anchor = pos + R @ anchor
                R = R @ rot(j["axis"], q[j["qpos"]])
                pos = world_anchor - R @ anchor
        out[b["name"]] = (pos, R)
    mass, com, geoms = 0.0, np.zeros(3), {}
    for b in tree["bodies"]:
        pos, R = out[b["name"]]
        for g in b["geoms"]:
            c = pos + R @ np.asarray(g["centre"])
            geoms[g["name"]] = c
            com += g["mass"] * c
            mass += g["mass"]
    return out, com / mass, geoms


def fixture_body_positions(tree, qpos):
    res = np.zeros(qpos.shape[:-1] + (15, 3))
    for idx in np.ndindex(*qpos.shape[:-1]):
        fk, _, geoms = fixture_fk(tree, qpos[idx])
        for k, name in enumerate(SO.BODY_DEFS):
            res[idx + (k,)] = geoms[name] if name.endswith("wrist") else fk[name][0]
    return res


def test_oracle_tables_equal_the_model_file(tree):
    bodies = {b["name"]: b for b in tree["bodies"]}
    assert tree["nq"] == 35 and [b["name"] for b in tree["bodies"]][0] == "root"
    assert [t[0] for t in SO.TREE] == [b["name"] for b in tree["bodies"][1:]]          # document order
    for name, parent, off, joints in SO.TREE:
        b = bodies[name]
        assert b["parent"] == parent and tuple(b["pos"]) == tuple(float(v) for v in off), name
        assert len(joints) == len(b["joints"]), name
        for (qi, axis, anchor), j in zip(joints, b["joints"]):
            assert j["type"] == "hinge" and j["qpos"] == qi, (name, j["name"])
            assert tuple(j["axis"]) == tuple(float(v) for v in axis) and tuple(j["pos"]) == tuple(float(v) for v in anchor), j["name"]
    # every geom with mass, its body and its centre; the wrists are geoms of the elbow bodies
    mine = sorted((m, body, tuple(round(float(v), 12) for v in c)) for m, body, c in SO.GEOMS)
    theirs = sorted((g["mass"], b["name"], tuple(round(v, 12) for v in g["centre"])) for b in tree["bodies"] for g in b["geoms"])
    assert mine == theirs
    for side in ("right", "left"):
        g = [g for g in bodies[side + "_elbow"]["geoms"] if g["name"] == side + "_wrist"][0]
        assert tuple(g["centre"]) == SO.WRIST_LOCAL
    assert abs(sum(m for m, _, _ in SO.GEOMS) - 30.5) < 1e-12


def test_hand_derived_poses(tree, harness, tmp_path):   # noqa: F811
    """Positions that follow from the XML numbers by hand (dp_env_v2.xml lines cited), for the fixture FK, the oracle and
    the product arithmetic."""
    H = 4
    q = np.zeros((1, H, 35))
    q[..., 2] = 0.9
    q[..., 3] = 1.0
    q[0, 1, 16] = np.pi / 2          # right elbow flexed 90 degrees: axis (0,-1,0), anchor (0.028,0,0)  (xml:48-49)
    q[0, 2, 24] = -np.pi / 2         # right knee flexed 90 degrees: axis (0,-1,0), anchor (-0.035,0,0) (xml:78)
    q[0, 3, 3:7] = [np.cos(np.pi / 4), 0, 0, np.sin(np.pi / 4)]    # whole body turned 90 degrees about z
    want = {}
    z_chest = 0.9 + 0.236151                                        # xml:28
    sh = np.array([-0.02405, -0.18311, z_chest + 0.2435])           # xml:42
    el0 = sh + [0, 0, -0.274788]                                    # xml:47
    want[(0, "chest")] = [0, 0, z_chest]
    want[(0, "neck")] = [0, 0, z_chest + 0.223894]                  # xml:34
    want[(0, "right_shoulder")] = sh
    want[(0, "right_elbow")] = el0
    want[(0, "right_wrist")] = el0 + [0, 0, -0.258947]              # xml:51
    want[(0, "left_ankle")] = [0, 0.094887, 0.9 - 0.05 - 0.421546 - 0.40987]    # xml:89,95,99
    # elbow hinge by +90 deg about -y: R maps x -> +z and -z -> +x, so the forearm points forward;
    # body origin = anchor_world - R anchor = el0 + (0.028,0,0) - (0,0,0.028)
    want[(1, "right_elbow")] = el0 + [0.028, 0, -0.028]
    want[(1, "right_wrist")] = el0 + [0.028 + 0.258947, 0, -0.028]
    # knee hinge by -90 deg about -y: R maps x -> -z and -z -> -x: the shank points backward
    kn0 = np.array([0, -0.094887, 0.9 - 0.05 - 0.421546])           # xml:71,76
    want[(2, "right_knee")] = kn0 + [-0.035, 0, -0.035]
    want[(2, "right_ankle")] = kn0 + [-0.035 - 0.40987, 0, -0.035]  # xml:80
    # yaw by 90 degrees: (x, y) -> (-y, x)
    want[(3, "right_shoulder")] = [0.18311, -0.02405, z_chest + 0.2435]
    want[(3, "left_hip")] = [-0.094887, 0, 0.9 - 0.05]
    fix = fixture_body_positions(tree, q)
    ora = SO.body_positions(q)
    _, _, prod = run(harness, tmp_path, q.astype(np.float32), q[0].astype(np.float32), 1 / 60)
    for (h, name), p in want.items():
        k = SO.BODY_DEFS.index(name)
        assert np.abs(fix[0, h, k] - p).max() < 1e-12, (h, name)
        assert np.abs(ora[0, h, k] - p).max() < 1e-12, (h, name)
        assert np.abs(prod[0, h, k] - p).max() < 2e-6, (h, name)


@pytest.mark.parametrize("seed,scale", [(0, 0.4), (1, 1.5)])
def test_fixture_fk_agrees_with_oracle_and_product(tree, harness, tmp_path, seed, scale):   # noqa: F811
    g = np.random.default_rng(seed)
    N, H = 4, 9
    q = np.zeros((N, H, 35))
    q[..., 0:3] = g.standard_normal((N, H, 3))
    q[..., 3:7] = g.standard_normal((N, H, 4))
    q[..., 7:] = scale * g.standard_normal((N, H, 28))
    fix = fixture_body_positions(tree, q)
    assert np.abs(fix - SO.body_positions(q)).max() < 1e-12
    ref = q[0]
    _, terms, prod = run(harness, tmp_path, q.astype(np.float32), ref.astype(np.float32), 1 / 60)
    q32 = q.astype(np.float32).astype(np.float64)
    fix32 = fixture_body_positions(tree, q32)
    assert np.abs(prod - fix32).max() < 1e-5                        # north_star: FK positions within 1e-5 m
    # centre of mass (every geom's mass and centre enters): the product's com error term against the fixture's
    for n in range(1, N):
        for h in range(H):
            _, com, _ = fixture_fk(tree, q32[n, h])
            _, com_ref, _ = fixture_fk(tree, q32[0, h])
            d = com - com_ref                                       # world-frame centre of mass
            want = float(d @ d)
            assert abs(terms[n, h, 3] - want) < 2e-5 + 1e-3 * want, (n, h)
