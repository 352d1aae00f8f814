"""CPU checks of oracle/ik_oracle.py, the restatement of the reference's action consumer (ArmManager / Pinocchio calls,
common/body/ArmManager.py:88-254). Pinocchio is not installed: its log6 / Jlog6 / LOCAL joint Jacobian are validated
against first principles (finite differences through scipy's matrix exponential), and the damped least-squares step
against its defining equation and by convergence."""
import os

import numpy as np
import pytest
from scipy.linalg import expm

from oracle import ik_oracle as iko

UR5E_URDF = "/root/reference/robo_manip_baselines/envs/assets/common/robots/ur5e/ur5e.urdf"


def ur5e_like_arm():
    """The UR5e chain with the values of the reference's URDF typed in (six z-axis revolute joints), so that the test
    does not need the reference mount."""
    H = np.pi / 2
    org = [((0, 0, np.pi), (0, 0, 0)), ]   # fixed base_link -> base_link_inertia
    joints = [((0, 0, 0), (0, 0, 0.1625)), ((H, 0, 0), (0, 0, 0)), ((0, 0, 0), (-0.425, 0, 0)),
              ((0, 0, 0), (-0.3922, 0, 0.1333)), ((H, 0, 0), (0, -0.0997, 0)), ((H, np.pi, np.pi), (0, 0.0996, 0))]
    placements = []
    acc = (iko.rpy_to_matrix(org[0][0]), np.array(org[0][1], float))
    for rpy, xyz in joints:
        placements.append(iko.se3_mul(acc, (iko.rpy_to_matrix(rpy), np.array(xyz, float))))
        acc = (np.eye(3), np.zeros(3))
    return iko.SerialArm(placements, [np.array([0.0, 0.0, 1.0])] * 6)


def exp6(xi):
    """SE3 exponential of a Motion [v; w] through the 4 x 4 matrix exponential."""
    T = np.zeros((4, 4))
    T[:3, :3] = iko.skew(xi[3:])
    T[:3, 3] = xi[:3]
    E = expm(T)
    return E[:3, :3], E[:3, 3]


def test_log6_inverts_exp6_and_Jlog6_is_its_right_jacobian():
    rng = np.random.default_rng(0)
    for angle in (1e-6, 0.3, 2.0, 3.0):               # rotation angle of the twist, below pi (the log's principal branch)
        xi = rng.normal(size=6)
        xi[3:] *= angle / np.linalg.norm(xi[3:])
        scale = angle
        M = exp6(xi)
        assert np.allclose(iko.log6(M), xi, atol=1e-9)
        J = iko.Jlog6(M)
        num = np.zeros((6, 6))
        h = 1e-6
        for i in range(6):
            d = np.zeros(6)
            d[i] = h
            num[:, i] = (iko.log6(iko.se3_mul(M, exp6(d))) - iko.log6(iko.se3_mul(M, exp6(-d)))) / (2 * h)
        assert np.allclose(J, num, atol=2e-6), (scale, np.abs(J - num).max())


def test_pose_conversions_round_trip():
    rng = np.random.default_rng(1)
    for _ in range(20):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        pose = np.concatenate([rng.normal(size=3), q])
        R, p = iko.se3_from_pose(pose)
        assert np.allclose(R @ R.T, np.eye(3), atol=1e-12) and np.isclose(np.linalg.det(R), 1.0)
        back = iko.pose_from_se3(R, p)
        assert np.allclose(back[:3], pose[:3]) and min(np.abs(back[3:] - q).max(), np.abs(back[3:] + q).max()) < 1e-12
    R = iko.rpy_to_matrix([0.3, -0.2, 1.1])
    Rz = lambda a: np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    Ry = lambda a: np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]])
    Rx = lambda a: np.array([[1, 0, 0], [0, np.cos(a), -np.sin(a)], [0, np.sin(a), np.cos(a)]])
    assert np.allclose(R, Rz(1.1) @ Ry(-0.2) @ Rx(0.3))


def test_local_joint_jacobian_matches_finite_differences_of_fk():
    arm = ur5e_like_arm()
    rng = np.random.default_rng(2)
    q = rng.uniform(-1.5, 1.5, size=6)
    J = arm.joint_jacobian_local(q, 6)
    cur = arm.joint_frames(q)[5]
    h = 1e-6
    for k in range(6):
        d = np.zeros(6)
        d[k] = h
        nxt = arm.joint_frames(q + d)[5]
        col = iko.log6(iko.se3_mul(iko.se3_inv(cur), nxt)) / h
        assert np.allclose(J[:, k], col, atol=1e-5), k


def test_ik_step_solves_its_defining_equation_and_converges():
    arm = ur5e_like_arm()
    q0 = np.array([0.2, -1.3, 1.6, -1.9, -1.5, 0.3])
    goal_q = q0 + np.array([0.15, -0.1, 0.2, 0.1, -0.15, 0.2])
    target = arm.joint_frames(goal_q)[5]
    q = q0.copy()
    errs = []
    for _ in range(25):
        cur = arm.joint_frames(q)[5]
        errs.append(np.linalg.norm(iko.log6(iko.se3_mul(iko.se3_inv(cur), target))))
        q = iko.ik_step(arm, q, target, 6)
    assert errs[-1] < 1e-9 and errs[5] < errs[0] * 0.2
    # the step is the damped least-squares solution: (J J^T + lambda I) y = e, dq = -J^T y with J = -Jlog6(err^-1) J_local
    cur = arm.joint_frames(q0)[5]
    E = iko.se3_mul(iko.se3_inv(cur), target)
    e = iko.log6(E)
    Jm = -iko.Jlog6(iko.se3_inv(E)) @ arm.joint_jacobian_local(q0, 6)
    dq = iko.ik_step(arm, q0, target, 6) - q0
    lam = e @ e + 1e-6
    assert np.allclose((Jm @ Jm.T + lam * np.eye(6)) @ np.linalg.lstsq(Jm.T, -dq, rcond=None)[0], e, atol=1e-9)


@pytest.mark.skipif(not os.path.exists(UR5E_URDF), reason="reference mount not present (build container only)")
def test_urdf_chain_of_the_reference_ur5e():
    arm = iko.SerialArm.from_urdf(UR5E_URDF)
    ref = ur5e_like_arm()
    assert arm.n == 6
    q = np.array([0.1, -1.2, 1.4, -1.7, -1.5, 0.2])
    a, b = arm.joint_frames(q)[5], ref.joint_frames(q)[5]
    assert np.allclose(a[0], b[0], atol=1e-8) and np.allclose(a[1], b[1], atol=1e-8)
    # wrist_3 at the zero configuration of a UR5e: 0.817 m reach along -x (base rotated by pi) is a known datum
    z = arm.joint_frames(np.zeros(6))[5]
    assert abs(abs(z[1][0]) - 0.8172) < 1e-3 and abs(z[1][2] - 0.0628) < 1e-3


def test_arm_manager_oracle_command_keys():
    arm = ur5e_like_arm()
    q0 = np.array([0.2, -1.3, 1.6, -1.9, -1.5, 0.3])
    m = iko.ArmManagerOracle(arm, 6, q0, np.zeros(1), np.arange(6), np.array([6]), [0.0], [255.0])
    m.set_command_data("command_joint_pos", np.concatenate([q0 + 0.1, [300.0]]))
    assert np.allclose(m.arm_joint_pos, q0 + 0.1) and m.gripper_joint_pos[0] == 255.0
    assert np.allclose(m.get_command_eef_pose()[:3], arm.joint_frames(q0 + 0.1)[5][1])
    m.set_command_data("command_joint_pos_rel", np.concatenate([np.full(6, 0.01), [-10.0]]), is_skip=True)
    assert np.allclose(m.arm_joint_pos, q0 + 0.1)
    m.set_command_data("command_joint_pos_rel", np.concatenate([np.full(6, 0.01), [-10.0]]))
    assert np.allclose(m.arm_joint_pos, q0 + 0.11) and m.gripper_joint_pos[0] == 245.0
    before = m.target
    m.set_command_data("command_eef_pose_rel", np.array([0.01, 0.0, -0.02, 0.0, 0.05, 0.0]))
    want = iko.se3_mul(before, iko.se3_from_rel_pose([0.01, 0.0, -0.02, 0.0, 0.05, 0.0]))
    assert np.allclose(m.target[0], want[0]) and np.allclose(m.target[1], want[1])
    parts = iko.split_action(np.arange(14.0), ["command_joint_pos", "command_eef_pose"], {"command_joint_pos": 7, "command_eef_pose": 7})
    assert np.array_equal(parts["command_eef_pose"], np.arange(7.0, 14.0))
    with pytest.raises(ValueError):
        m.set_command_data("command_mobile_omni_vel", np.zeros(3))
