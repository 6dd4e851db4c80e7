"""Generates tests/golden/mocap_clips.npz (run HERE, where /root/reference exists; the result is committed).

Contents: the raw DeepMimic frames of the three clips the configs use (data, not code), the loader oracle's
data_config and data_vel for them, and a cross-check of the oracle's quaternion->Euler step against the reference's own
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
    # data_vel (mocap_v2.py:196-291): pyquaternion is not importable, so calc_rot_vel's q_diff, axis and angle are
    # cross-checked against the reference's own transformations.py: the Hamilton product conj(q_k) * q_{k-1} against
    # quaternion_multiply / quaternion_conjugate, and rotation_matrix(angle, axis) against quaternion_matrix(q_diff)
    vel = LO.data_vel_from_frames(fr)
    worst_mul = worst_rot = 0.0
    rq = np.array([LO.align_rotation(r[4:8]) for r in fr])
    for k in range(1, fr.shape[0]):
        q0, q1 = rq[k], rq[k - 1]
        mine = LO.quat_mul(np.array([q0[0], -q0[1], -q0[2], -q0[3]]), q1)
        ref = T.quaternion_multiply(T.quaternion_conjugate([q0[1], q0[2], q0[3], q0[0]]), [q1[1], q1[2], q1[3], q1[0]])
        worst_mul = max(worst_mul, float(np.abs(mine - np.array([ref[3], ref[0], ref[1], ref[2]])).max()))
        dura = fr[k - 1, 0] if fr[k - 1, 0] != 0 else 0.0167
        rv = vel[k, 3:6] * dura                         # angle * axis
        ang = float(np.linalg.norm(rv))
        if ang > 1e-9:
            R_mine = T.rotation_matrix(ang, rv / ang)[:3, :3]
            R_ref = T.quaternion_matrix([ref[0], ref[1], ref[2], ref[3]])[:3, :3]
            worst_rot = max(worst_rot, float(np.abs(R_mine - R_ref).max()))
    print(name, "data_vel", vel.shape, "q_diff vs transformations.py", worst_mul, "axis-angle rotation vs quaternion_matrix", worst_rot)
    assert worst_mul < 1e-12 and worst_rot < 1e-9
    out[name + "_frames"] = fr
    out[name + "_config"] = cfg
    out[name + "_vel"] = vel
np.savez_compressed(os.path.join(HERE, "mocap_clips.npz"), **out)
print("cartwheel data_config[0][:9] =", out["cartwheel_config"][0][:9])
