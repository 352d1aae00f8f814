"""Batched action consumer on the device vs oracle/ik_oracle.py (the numpy restatement of
ArmManager.set_command_data / inverse_kinematics, common/body/ArmManager.py:88-254; Pinocchio calls restated -- parity
unpinned for those, validated against first principles in tests/test_ik_cpu.py). fp64 on both sides: agreement to 1e-10
(different summation order / device libm), exact for pure data movement."""
import numpy as np
import pytest
import torch

from oracle import ik_oracle as iko
from test_ik_cpu import ur5e_like_arm

pytestmark = pytest.mark.gpu


def _managers(cuda, B, rng):
    from robomanipbaselines_b200.common.action_consumer import BatchedArmManager

    arm = ur5e_like_arm()
    q0 = np.array([0.2, -1.3, 1.6, -1.9, -1.5, 0.3])
    dev = BatchedArmManager(B, arm.placements, arm.axes, 6, np.arange(6), np.array([6]), q0, np.zeros(1), [0.0], [255.0],
                            device=str(cuda))
    ors = [iko.ArmManagerOracle(arm, 6, q0, np.zeros(1), np.arange(6), np.array([6]), [0.0], [255.0]) for _ in range(B)]
    return arm, dev, ors


def _check(dev, ors, tol=1e-10):
    q = dev.arm_joint_pos.cpu().numpy()
    g = dev.gripper_joint_pos.cpu().numpy()
    t = dev.target_se3.cpu().numpy()
    for b, o in enumerate(ors):
        assert np.abs(q[b] - o.arm_joint_pos).max() <= tol, b
        assert np.abs(g[b, :1] - o.gripper_joint_pos).max() <= tol, b
        assert np.abs(t[b, :9].reshape(3, 3) - o.target[0]).max() <= tol and np.abs(t[b, 9:] - o.target[1]).max() <= tol, b


def test_arm_commands_match_oracle(cuda):
    rng = np.random.default_rng(0)
    B = 37
    arm, dev, ors = _managers(cuda, B, rng)
    _check(dev, ors)                                            # reset: target = FK(init)
    # joint position commands (+ gripper clipping), then relative ones, skipped and applied
    cmd = np.concatenate([rng.uniform(-1.5, 1.5, (B, 6)), rng.uniform(-50, 300, (B, 1))], axis=1)
    dev.set_command_data("command_joint_pos", torch.from_numpy(cmd).to(cuda))
    for b, o in enumerate(ors):
        o.set_command_data("command_joint_pos", cmd[b])
    _check(dev, ors)
    rel = np.concatenate([rng.normal(0, 0.02, (B, 6)), rng.normal(0, 5, (B, 1))], axis=1)
    for skip in (True, False):
        dev.set_command_data("command_joint_pos_rel", torch.from_numpy(rel).to(cuda), is_skip=skip)
        for b, o in enumerate(ors):
            o.set_command_data("command_joint_pos_rel", rel[b], is_skip=skip)
        _check(dev, ors)
    # end-effector pose commands: 12 IK steps towards a pose reachable from a perturbed configuration
    goal_q = np.stack([o.arm_joint_pos for o in ors]) + rng.normal(0, 0.08, (B, 6))
    poses = np.stack([iko.pose_from_se3(*arm.joint_frames(goal_q[b])[5]) for b in range(B)])
    d_pose = torch.from_numpy(poses).to(cuda)
    for it in range(12):
        dev.set_command_data("command_eef_pose", d_pose)
        for b, o in enumerate(ors):
            o.set_command_data("command_eef_pose", poses[b])
        _check(dev, ors, tol=1e-9)
    got = dev.get_command_data("command_eef_pose").cpu().numpy()
    assert np.abs(got - poses).max() <= 1e-9                    # the target SE3 round-trips through the quaternion
    reached = np.stack([iko.pose_from_se3(*arm.joint_frames(o.arm_joint_pos)[5]) for o in ors])
    assert np.abs(reached[:, :3] - poses[:, :3]).max() < 1e-6   # and the iteration has converged on it
    # relative end-effector commands (always applied: ArmManager.set_command_data does not forward is_skip)
    relp = np.concatenate([rng.normal(0, 0.01, (B, 3)), rng.normal(0, 0.03, (B, 3))], axis=1)
    for it in range(3):
        dev.set_command_data("command_eef_pose_rel", torch.from_numpy(relp).to(cuda), is_skip=bool(it & 1))
        for b, o in enumerate(ors):
            o.set_command_data("command_eef_pose_rel", relp[b], is_skip=bool(it & 1))
        _check(dev, ors, tol=1e-9)
    jp = dev.get_command_data("command_joint_pos").cpu().numpy()
    assert np.abs(jp - np.stack([o.get_command_joint_pos() for o in ors])).max() <= 1e-9
    # per-episode reset
    mask = np.zeros(B, dtype=bool)
    mask[::3] = True
    dev.reset(torch.from_numpy(mask))
    for b, o in enumerate(ors):
        if mask[b]:
            o.reset()
    _check(dev, ors, tol=1e-9)
    with pytest.raises(ValueError):
        dev.set_command_data("command_mobile_omni_vel", torch.zeros(B, 3, device=cuda, dtype=torch.float64))
    with pytest.raises(ValueError):
        dev.set_command_data("command_eef_pose", torch.zeros(B, 7, device=cuda, dtype=torch.float32))


def test_action_consumer_splits_by_key_and_tallies_results(cuda):
    from robomanipbaselines_b200.common.action_consumer import BatchedActionConsumer, ResultTally

    rng = np.random.default_rng(1)
    B = 5
    arm, dev, ors = _managers(cuda, B, rng)
    keys, dims = ["command_gripper_joint_pos", "command_eef_pose"], {"command_gripper_joint_pos": 1, "command_eef_pose": 7}
    cons = BatchedActionConsumer(keys, dims, [dev])
    goal_q = np.stack([o.arm_joint_pos for o in ors]) + rng.normal(0, 0.05, (B, 6))
    poses = np.stack([iko.pose_from_se3(*arm.joint_frames(goal_q[b])[5]) for b in range(B)])
    act = np.concatenate([rng.uniform(0, 255, (B, 1)), poses], axis=1)           # (B, Da = 8) policy actions
    cons.set_command_data(torch.from_numpy(act).to(cuda), is_skip=False)
    parts = iko.split_action(act, keys, dims)
    for b, o in enumerate(ors):
        o.set_command_data("command_gripper_joint_pos", parts["command_gripper_joint_pos"][b])
        o.set_command_data("command_eef_pose", parts["command_eef_pose"][b])
    _check(dev, ors, tol=1e-9)
    env_action = cons.get_env_action(["command_joint_pos"]).cpu().numpy()
    assert np.abs(env_action - np.stack([o.get_command_joint_pos() for o in ors])).max() <= 1e-9
    with pytest.raises(ValueError):
        BatchedActionConsumer(["command_mobile_omni_vel"], {"command_mobile_omni_vel": 3}, [dev]).set_command_data(
            torch.zeros(B, 3, device=cuda, dtype=torch.float64))
    tally = ResultTally(B, device=str(cuda))
    tally.record(torch.tensor([True, False, True, False, False]), torch.tensor([1.0, 0.0, 0.2, 0.0, 0.0]), torch.full((B,), 3.5))
    tally.record(torch.tensor([True, True, False, False, True]), torch.tensor([0.0, 1.0, 0.0, 0.0, 1.0]), torch.full((B,), 9.0))
    res = tally.result()
    assert res["success"] == [True, True, False, False, True]            # a finished episode keeps its first result
    assert res["duration"] == [3.5, 9.0, 3.5, 0.0, 9.0] and res["reward"][2] == pytest.approx(0.2)
