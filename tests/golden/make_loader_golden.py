"""Generates tests/golden/mocap_clips.npz (run HERE, where /root/reference exists; the result is committed).

Contents: the raw DeepMimic frames of the three clips the configs use (data, not code), the loader oracle's
data_config for them, and a cross-check of the oracle's quaternion->Euler step against the reference's own
utils/transformations.euler_from_quaternion (importable as-is; pyquaternion is not, so align_rotation is restated)."""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, "/root/reference/diffusion/utils")
sys.dont_write_bytecode = True
from oracle import loader_oracle as LO
import transformations as T   # the reference's vendored file

out = {}
for name in ["walk", "cartwheel", "spinkick"]:
    fr = LO.load_frames(f"/root/reference/diffusion/data/motions/humanoid3d_{name}.txt")
    cfg = LO.data_config_from_frames(fr)
    # cross-check Euler extraction against the reference function on every 3-dof joint of every frame
    off, worst = 8, 0.0
    for j in LO.BODY_JOINTS_IN_DP_ORDER:
        if LO.DOF_DEF[j] == 1:
            off += 1
            continue
        for f in range(fr.shape[0]):
            q = LO.align_rotation(fr[f, off:off + 4])
            ref = T.euler_from_quaternion(np.array([q[1], q[2], q[3], q[0]]), axes="rxyz")
            mine = LO.euler_rxyz_from_matrix(LO.quaternion_matrix_xyzw([q[1], q[2], q[3], q[0]]))
            worst = max(worst, float(np.abs(np.array(ref) - np.array(mine)).max()))
        off += 4
    print(name, fr.shape, "euler vs reference transformations.py max abs diff", worst)
    assert worst < 1e-12
    out[name + "_frames"] = fr
    out[name + "_config"] = cfg
np.savez_compressed(os.path.join(HERE, "mocap_clips.npz"), **out)
print("cartwheel data_config[0][:9] =", out["cartwheel_config"][0][:9])
