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
