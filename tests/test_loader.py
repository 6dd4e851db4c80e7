"""CPU tests of the mocap loader boundary (SURVEY.md 8(b), 8(c) loader oracle): fixture <-> oracle <-> product."""
import json
import os

import numpy as np
import torch

from deepmimic_diffusion_mujoco_b200 import motion_dataset as MD
from oracle import loader_oracle as LO


def clips(golden_dir):
    return np.load(os.path.join(golden_dir, "mocap_clips.npz"))


def test_oracle_reproduces_fixture_and_notebook_anchor(golden_dir):
    z = clips(golden_dir)
    for name in ["walk", "cartwheel", "spinkick"]:
        cfg = LO.data_config_from_frames(z[name + "_frames"])
        assert cfg.shape == (z[name + "_frames"].shape[0], 35)
        np.testing.assert_allclose(cfg, z[name + "_config"], rtol=0, atol=1e-13)
    # archive/0_temporal_unet_diffusion.ipynb cell 3 prints data_config[0] of the cartwheel clip
    anchor = [0.0, 0.0, 0.85536, 0.996643, -0.007001, 0.08157, 0.000573, 0.0423037, -0.0560887]
    np.testing.assert_allclose(z["cartwheel_config"][0][:9], anchor, atol=5e-7)


def test_product_loader_matches_oracle(golden_dir):
    z = clips(golden_dir)
    for name in ["walk", "cartwheel", "spinkick"]:
        got = MD.mocap_to_qpos(z[name + "_frames"])
        np.testing.assert_allclose(got, z[name + "_config"], rtol=0, atol=1e-12)


def test_velocity_rows_oracle_fixture_and_product(golden_dir):
    """data_vel (utils/mocap_v2.py:196-291, calc_rot_vel :155-178): oracle == fixture, product host code == oracle, and
    properties the reference arithmetic implies: row 0 is zero, linear / hinge rates are finite differences over the
    PREVIOUS frame's duration, and angle * axis rotates frame k's root orientation onto frame k-1's."""
    z = clips(golden_dir)
    for name in ["walk", "cartwheel", "spinkick"]:
        fr = z[name + "_frames"]
        vel = LO.data_vel_from_frames(fr)
        assert vel.shape == (fr.shape[0], 34)
        np.testing.assert_allclose(vel, z[name + "_vel"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(MD.mocap_to_qvel(fr), z[name + "_vel"], rtol=0, atol=1e-11)
        assert not vel[0].any()
        cfg = z[name + "_config"]
        dura = fr[:-1, 0]
        np.testing.assert_allclose(vel[1:, :3], (cfg[1:, :3] - cfg[:-1, :3]) / dura[:, None], atol=1e-12)
        for qpos_i, qvel_i in [(16, 15), (20, 19), (24, 23), (31, 30)]:      # elbows, knees: hinge rates
            np.testing.assert_allclose(vel[1:, qvel_i], (cfg[1:, qpos_i] - cfg[:-1, qpos_i]) / dura, atol=1e-12)
        k = 5
        rv = vel[k, 3:6] * dura[k - 1]
        ang = np.linalg.norm(rv)
        half = np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * rv / ang])
        q_prev = LO.quat_mul(cfg[k, 3:7], half)                                  # q_k * q_diff == q_{k-1}
        want = cfg[k - 1, 3:7] / np.linalg.norm(cfg[k - 1, 3:7])                 # the clip's quaternions are unit to ~1e-7
        q_prev = q_prev / np.linalg.norm(q_prev)
        np.testing.assert_allclose(q_prev * np.sign(q_prev[0]), want * np.sign(want[0]), atol=1e-9)


def test_motion_dataset_v2_semantics(golden_dir, tmp_path):
    """data_loaders/motion_dataset_v2.py: 69-wide rows (qpos | qvel), motion class from the file name, 1000 copies
    (shuffle=False) or root-continuous cyclic shifts (shuffle=True)."""
    z = clips(golden_dir)
    path = tmp_path / "humanoid3d_spinkick.txt"
    path.write_text(json.dumps({"Loop": "wrap", "Frames": z["spinkick_frames"].tolist()}))
    ds = MD.MotionDatasetV2(str(path))
    ref = LO.motion_dataset_v2_rows(z["spinkick_frames"])
    assert len(ds) == 1000 and ds.max_sequence_length == 78
    b = ds[3]
    assert int(b.motion_class) == 2 and b.trajectories.shape == (78, 69)
    assert torch.equal(b.trajectories, torch.from_numpy(ref))
    assert torch.equal(ds.get_conditions(b.trajectories)[0], b.trajectories[0])
    sh = MD.MotionDatasetV2(str(path), shuffle=True)
    ref_sh = LO.motion_dataset_v2_shuffled(z["spinkick_frames"])
    assert len(sh) == 78
    for i in (0, 1, 40, 77):
        assert torch.equal(sh[i].trajectories, torch.from_numpy(ref_sh[i])), i
    batch = sh.collate_fn([sh[0], sh[5]])
    assert batch.trajectories.is_nested and batch.motion_class.tolist() == [2, 2]


def test_motion_dataset_semantics(golden_dir, tmp_path):
    z = clips(golden_dir)
    path = tmp_path / "humanoid3d_walk.txt"
    path.write_text(json.dumps({"Loop": "wrap", "Frames": z["walk_frames"].tolist()}))
    ds = MD.MotionDataset(str(path))
    ref = LO.motion_dataset(z["walk_frames"])
    assert len(ds) == 32 and ref.shape == (32, 32, 35)            # 39 frames -> 32 (multiple of 8)
    assert torch.equal(ds.motion_data, torch.from_numpy(ref))
    b = ds[5]
    assert torch.equal(b.trajectories, torch.from_numpy(ref[5])) and torch.equal(b.conditions[0], b.trajectories[0])
    assert torch.equal(ds[0].trajectories[5], ds[5].trajectories[0])   # cyclic shift
    t = MD.tile_clip(ds[0].trajectories, 64)
    assert t.shape == (64, 35) and torch.equal(t[32:], t[:32])


def test_npy_writer_contract(tmp_path):
    traj = torch.randn(64, 35, dtype=torch.float64)
    p = tmp_path / "motion.npy"
    MD.save_motion_npy(str(p), traj)
    back = np.load(str(p))
    assert back.dtype == np.float32 and back.shape == (64, 35)
    np.testing.assert_array_equal(back, traj.float().numpy())


def test_bulk_writer_contract(tmp_path):
    """save_motions (transformer_inference.py:213-276): one [frames, 35] file per sample, default names, 69 -> 35 features."""
    s = torch.randn(3, 32, 69)
    paths = MD.save_motions(s, str(tmp_path / "out"), filenames=["a.npy"])
    assert [os.path.basename(p) for p in paths] == ["a.npy", "motion_1.npy", "motion_2.npy"]
    for i, p in enumerate(paths):
        back = np.load(p)
        assert back.shape == (32, 35)
        np.testing.assert_array_equal(back, s[i, :, :35].numpy())
    small = MD.save_motions(torch.randn(1, 8, 20), str(tmp_path / "pad"))
    assert np.load(small[0]).shape == (8, 35) and not np.load(small[0])[:, 20:].any()
