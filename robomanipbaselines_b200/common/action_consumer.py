"""Batched action consumer: what the reference does with `policy_action` right after infer_policy, for B
episodes at once on the device.

  RolloutBase.set_command_data      common/base/RolloutBase.py:496-509   split the action vector by action key
  MotionManager.set_command_data    common/manager/MotionManager.py:25-39 dispatch to the body managers
  ArmManager.set_command_data ...   common/body/ArmManager.py:88-160     joint-position passthrough (+ FK for the target
                                                                          pose, gripper clipped to the action-space limits),
                                                                          relative joint commands, end-effector pose commands
  ArmManager.inverse_kinematics     common/body/ArmManager.py:230-254    ONE damped least-squares step per command
  RolloutPhase result tally         common/base/RolloutBase.py:101-107   success / reward / duration lists (gathered over
                                                                          ranks by common/sharding.py::gather_results)

The kinematic model is read from the arm's URDF (body_config.arm_urdf_path) as a serial chain of revolute joints -- what
pin.buildModelFromUrdf produces for the reference's arms; Pinocchio itself is not needed. The arithmetic runs in
rmb_arm_command (csrc/ik.cuh), fp64, one thread per episode; oracle/ik_oracle.py is its numpy twin.
"""
import ctypes
import xml.etree.ElementTree as ET

import numpy as np
import torch

from .. import _lib
from .._lib import check, cur_stream, ptr

KEY_CODES = {"command_joint_pos": 0, "command_joint_pos_rel": 1, "command_gripper_joint_pos": 2, "command_eef_pose": 3,
             "command_eef_pose_rel": 4}
KEY_RESET = 5


def _rpy(rpy):
    r, p, y = rpy
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                     [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr], [-sp, cp * sr, cp * cr]])


def _quat_wxyz(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def chain_from_urdf(path, root_pose=None, exclude_joint_names=()):
    """[(R, p)] placements and [axis] of the revolute joints along the URDF's single chain; fixed (and excluded) joints
    are merged into the next revolute joint's placement, root_pose (tx, ty, tz, qw, qx, qy, qz) = ArmConfig.arm_root_pose
    multiplies the first placement (ArmManager.py:47-71)."""
    root = ET.parse(path).getroot()
    by_parent, children = {}, set()
    for j in root.findall("joint"):
        org = j.find("origin")
        xyz = np.array([float(v) for v in (org.get("xyz", "0 0 0") if org is not None else "0 0 0").split()])
        rpy = [float(v) for v in (org.get("rpy", "0 0 0") if org is not None else "0 0 0").split()]
        ax = j.find("axis")
        axis = np.array([float(v) for v in ax.get("xyz").split()]) if ax is not None else np.array([1.0, 0.0, 0.0])
        child = j.find("child").get("link")
        by_parent.setdefault(j.find("parent").get("link"), []).append((j.get("name"), j.get("type"), child, _rpy(rpy), xyz, axis))
        children.add(child)
    base = [l.get("name") for l in root.findall("link") if l.get("name") not in children][0]
    if root_pose is None:
        acc = (np.eye(3), np.zeros(3))
    else:
        root_pose = np.asarray(root_pose, np.float64)
        acc = (_quat_wxyz(root_pose[3:7]), root_pose[:3].copy())
    placements, axes, link = [], [], base
    while link in by_parent:
        name, typ, child, R, p, axis = by_parent[link][0]
        acc = (acc[0] @ R, acc[0] @ p + acc[1])
        if typ in ("revolute", "continuous") and name not in exclude_joint_names:
            placements.append(acc)
            axes.append(axis / np.linalg.norm(axis))
            acc = (np.eye(3), np.zeros(3))
        link = child
    return placements, axes


class BatchedArmManager:
    """ArmManager for B episodes: arm joint positions (B, n), gripper joint positions (B, g), target SE3 (B, 12) on the
    device; set_command_data / get_command_data with the reference's keys and semantics."""

    SUPPORTED_COMMAND_KEYS = tuple(KEY_CODES)

    def __init__(self, batch, placements, axes, ik_eef_joint_id, arm_joint_idxes, gripper_joint_idxes, init_arm_joint_pos,
                 init_gripper_joint_pos, gripper_low, gripper_high, gripper_joint_idxes_in_gripper_joint_pos=None,
                 eef_idx=0, device="cuda:0"):
        self.lib = _lib.load()
        self.B, self.device = int(batch), torch.device(device)
        n, g = len(axes), len(gripper_joint_idxes)
        if n > _lib.ARM_MAX_JOINTS or g > _lib.ARM_MAX_GRIPPER:
            raise ValueError(f"at most {_lib.ARM_MAX_JOINTS} arm joints and {_lib.ARM_MAX_GRIPPER} gripper joints")
        m = _lib.ArmModel()
        m.n_joints, m.ik_joint_id = n, int(ik_eef_joint_id)
        for k, ((R, p), a) in enumerate(zip(placements, axes)):
            for i in range(9):
                m.placement_R[k][i] = float(np.asarray(R).reshape(-1)[i])
            for i in range(3):
                m.placement_p[k][i] = float(p[i])
                m.axis[k][i] = float(a[i])
        lay = _lib.ArmLayout()
        lay.n_gripper, lay.eef_idx = g, int(eef_idx if eef_idx is not None else 0)
        in_grip = gripper_joint_idxes_in_gripper_joint_pos if gripper_joint_idxes_in_gripper_joint_pos is not None else range(g)
        for k in range(n):
            lay.arm_idx[k] = int(arm_joint_idxes[k])
            lay.init_arm[k] = float(init_arm_joint_pos[k])
        for i in range(g):
            lay.gripper_idx[i] = int(gripper_joint_idxes[i])
            lay.gripper_idx_in_gripper[i] = int(list(in_grip)[i])
            lay.init_gripper[i] = float(init_gripper_joint_pos[i])
            lay.gripper_low[i], lay.gripper_high[i] = float(gripper_low[i]), float(gripper_high[i])
        self.model, self.layout, self.n, self.g = m, lay, n, g
        self.joint_dim = n + g
        dev = self.device
        self.arm_joint_pos = torch.zeros(batch, n, device=dev, dtype=torch.float64)
        self.gripper_joint_pos = torch.zeros(batch, max(g, 1), device=dev, dtype=torch.float64)
        self.target_se3 = torch.zeros(batch, 12, device=dev, dtype=torch.float64)
        self.reset()

    @classmethod
    def from_body_config(cls, batch, body_config, action_low, action_high, device="cuda:0"):
        """From the reference's ArmConfig (ArmManager.py:264-315) and the env's action-space limits."""
        pl, ax = chain_from_urdf(body_config.arm_urdf_path, getattr(body_config, "arm_root_pose", None),
                                 getattr(body_config, "exclude_joint_names", None) or ())
        lim = getattr(body_config, "gripper_joint_idxes_for_limit", None)
        lim = body_config.gripper_joint_idxes if lim is None else lim
        return cls(batch, pl, ax, body_config.ik_eef_joint_id, body_config.arm_joint_idxes, body_config.gripper_joint_idxes,
                   body_config.init_arm_joint_pos, body_config.init_gripper_joint_pos, np.asarray(action_low)[lim],
                   np.asarray(action_high)[lim], body_config.gripper_joint_idxes_in_gripper_joint_pos, body_config.eef_idx,
                   device=device)

    def _call(self, key_code, command, is_skip, mask):
        cmd_ptr, stride = None, 0
        if command is not None:
            if command.dtype != torch.float64 or command.dim() != 2 or command.shape[0] != self.B or not command.is_cuda:
                raise ValueError(f"command must be a float64 CUDA tensor ({self.B}, dim)")
            if command.stride(1) != 1:
                command = command.contiguous()
            cmd_ptr, stride = ptr(command), command.stride(0)
        m = None
        if mask is not None:
            m = torch.as_tensor(mask, device=self.device).to(torch.uint8).contiguous()
        check(self.lib.rmb_arm_command(ctypes.byref(self.model), ctypes.byref(self.layout), key_code, cmd_ptr, stride,
                                       int(bool(is_skip)), ptr(m), ptr(self.arm_joint_pos), ptr(self.gripper_joint_pos),
                                       ptr(self.target_se3), self.B, cur_stream()), "rmb_arm_command")

    def reset(self, mask=None):
        """ArmManager.reset (:76-86) for the episodes in mask (default all)."""
        self._call(KEY_RESET, None, False, mask)

    def set_command_data(self, key, command, is_skip=False):
        """ArmManager.set_command_data (:88-123). command: (B, dim of that key) float64 on the device."""
        if key not in KEY_CODES:
            raise ValueError(f"[{self.__class__.__name__}] Invalid command data key: {key}")
        need = {0: self.joint_dim, 1: self.joint_dim, 2: self.g, 3: 7 * (self.layout.eef_idx + 1), 4: 6 * (self.layout.eef_idx + 1)}[KEY_CODES[key]]
        if command.shape[1] < need:
            raise ValueError(f"{key}: command has {command.shape[1]} columns, needs {need}")
        self._call(KEY_CODES[key], command, is_skip, None)

    def get_command_data(self, key):
        """ArmManager.get_command_data (:162-172): command_joint_pos -> (B, n + g); command_gripper_joint_pos -> (B, g);
        command_eef_pose -> (B, 7) pose of the target SE3."""
        if key == "command_gripper_joint_pos":
            return self.gripper_joint_pos[:, :self.g]
        jp = torch.zeros(self.B, self.joint_dim, device=self.device, dtype=torch.float64) if key == "command_joint_pos" else None
        pose = torch.zeros(self.B, 7, device=self.device, dtype=torch.float64) if key == "command_eef_pose" else None
        if jp is None and pose is None:
            raise ValueError(f"[{self.__class__.__name__}] Invalid command data key: {key}")
        check(self.lib.rmb_arm_get_command(ctypes.byref(self.model), ctypes.byref(self.layout), ptr(self.arm_joint_pos),
                                           ptr(self.gripper_joint_pos), ptr(self.target_se3), ptr(jp), self.joint_dim,
                                           ptr(pose), self.B, cur_stream()), "rmb_arm_get_command")
        return jp if jp is not None else pose


class BatchedActionConsumer:
    """RolloutBase.set_command_data (:496-509) + MotionManager.set_command_data (:25-39) for B episodes: slice the (B, Da)
    action matrix by action key and hand every slice to the body managers that support the key."""

    def __init__(self, action_keys, key_dims, body_managers):
        self.action_keys, self.key_dims, self.body_managers = list(action_keys), dict(key_dims), list(body_managers)
        if sum(self.key_dims[k] for k in self.action_keys) <= 0:
            raise ValueError("no action dimensions")

    def set_command_data(self, policy_action, is_skip=False, action_keys=None):
        keys = self.action_keys if action_keys is None else action_keys
        idx = 0
        for key in keys:
            dim = self.key_dims[key]
            handled = False
            for bm in self.body_managers:
                if key in bm.SUPPORTED_COMMAND_KEYS:
                    bm.set_command_data(key, policy_action[:, idx:idx + dim], is_skip)
                    handled = True
            if not handled:
                raise ValueError(f"[{self.__class__.__name__}] Command data key is not supported by any body manager: {key}")
            idx += dim

    def get_env_action(self, command_keys_for_step):
        """RolloutBase.run (:396-402): the env action = concatenation of get_command_data over command_keys_for_step."""
        return torch.cat([self.body_managers[0].get_command_data(k) for k in command_keys_for_step], dim=1)


class ResultTally:
    """The result lists RolloutPhase.check_transition appends to (RolloutBase.py:101-107) for B episodes per rank, and
    their union over ranks in global episode order (AutoEval reads `success`, misc/AutoEval.py:423-435)."""

    def __init__(self, batch, device="cuda:0"):
        self.B, self.device = batch, torch.device(device)
        self.success = torch.zeros(batch, device=self.device, dtype=torch.uint8)
        self.reward = torch.zeros(batch, device=self.device, dtype=torch.float32)
        self.duration = torch.zeros(batch, device=self.device, dtype=torch.float32)
        self.done = torch.zeros(batch, device=self.device, dtype=torch.bool)

    def record(self, finished_mask, reward, elapsed_duration):
        """finished_mask bool (B,): episodes whose rollout phase ended at this tick; reward / duration (B,)."""
        fm = torch.as_tensor(finished_mask, device=self.device) & ~self.done
        self.success = torch.where(fm, (torch.as_tensor(reward, device=self.device) >= 1.0).to(torch.uint8), self.success)
        self.reward = torch.where(fm, torch.as_tensor(reward, device=self.device, dtype=torch.float32), self.reward)
        self.duration = torch.where(fm, torch.as_tensor(elapsed_duration, device=self.device, dtype=torch.float32), self.duration)
        self.done |= fm

    def result(self, total_episodes=None):
        """{"success": [...], "reward": [...], "duration": [...]} as yaml.dump writes it (RolloutBase.py:417-422)."""
        from .sharding import gather_results

        s, r, d = self.success, self.reward, self.duration
        if total_episodes is not None:
            s, r, d = gather_results(s, r, d, total_episodes)
        return {"success": [bool(v) for v in s.tolist()], "reward": [float(v) for v in r.tolist()],
                "duration": [float(v) for v in d.tolist()]}
