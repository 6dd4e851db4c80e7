"""CPU check of the K5 arithmetic: csrc/fk_math.cuh (the source the CUDA kernel inlines) is compiled for the host by
a test-only harness and compared with the float64 oracle.  This does not make a CPU product path: the library's
dm_fk_reward only launches the CUDA kernel."""
import os
import struct
import subprocess

import numpy as np
import pytest

from oracle import scoring_oracle as SO

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    out = tmp_path_factory.mktemp("fk") / "fk_host_harness"
    subprocess.run(["g++", "-O2", "-std=c++17", os.path.join(ROOT, "tests", "fk_host_harness.cpp"), "-o", str(out)], check=True)
    return str(out)


def run(harness, tmp_path, qpos, ref, dt):
    N, H, _ = qpos.shape
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(fin, "wb") as f:
        f.write(struct.pack("iif", N, H, dt))
        f.write(qpos.astype(np.float32).tobytes())
        f.write(ref.astype(np.float32).tobytes())
    subprocess.run([harness, str(fin), str(fout)], check=True)
    raw = np.fromfile(fout, dtype=np.float32)
    n = N * H
    return raw[:n].reshape(N, H), raw[n:5 * n].reshape(N, H, 4), raw[5 * n:].reshape(N, H, 15, 3)


def make(seed, N, H, spread, angle_scale=0.3):
    g = np.random.default_rng(seed)
    ref = np.zeros((H, 35))
    ref[:, 2] = 0.9
    ref[:, 3] = 1.0
    ref[:, 7:] = angle_scale * g.standard_normal((H, 28))
    ref[:, 3:7] += 0.1 * g.standard_normal((H, 4))
    qpos = ref[None] + spread * g.standard_normal((N, H, 35))
    qpos[0] = ref
    return qpos.astype(np.float32), ref.astype(np.float32)


@pytest.mark.parametrize("seed,spread,angle_scale", [(0, 0.15, 0.3), (1, 0.02, 1.0), (2, 0.5, 2.5)])
def test_fk_math_matches_oracle(harness, tmp_path, seed, spread, angle_scale):
    dt = 1 / 60
    qpos, ref = make(seed, 5, 12, spread, angle_scale)
    reward, terms, bp = run(harness, tmp_path, qpos, ref, dt)
    q64, r64 = qpos.astype(np.float64), ref.astype(np.float64)
    bp_o = SO.body_positions(q64)
    assert np.abs(bp - bp_o).max() < 1e-5          # north_star: FK positions within 1e-5 m
    r_o, t_o = SO.tracking_reward(q64, r64, dt)
    assert np.abs(reward - r_o).max() < 1e-4
    assert abs(reward[0].mean() - 1.0) < 1e-5      # a perfect tracker scores 1


def test_large_angles_wrap(harness, tmp_path):
    """hinge angles many turns away from zero: the half-angle reduction may flip q to -q, never the rotation."""
    dt = 1 / 30
    qpos, ref = make(3, 3, 6, 0.1)
    qpos[1, :, 7:] += 2 * np.pi * 3
    qpos[2, :, 7:] -= 2 * np.pi * 5
    _, _, bp = run(harness, tmp_path, qpos, ref, dt)
    bp_o = SO.body_positions(qpos.astype(np.float64))
    assert np.abs(bp - bp_o).max() < 2e-5
