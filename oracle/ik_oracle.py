"""TEST INFRASTRUCTURE ONLY (parity checker; never on the product path).

numpy-f64 restatement of the reference's action consumer for arm bodies:
  * RolloutBase.set_command_data          -- robo_manip_baselines/common/base/RolloutBase.py:496-509 (split by action key)
  * ArmManager.set_command_data & friends -- robo_manip_baselines/common/body/ArmManager.py:88-160
  * ArmManager.inverse_kinematics         -- robo_manip_baselines/common/body/ArmManager.py:230-254 (ONE damped
                                             least-squares step per command, damping |e|^2 + 1e-6)
  * get_se3_from_pose / get_pose_from_se3 / get_se3_from_rel_pose -- common/utils/MathUtils.py:27-46

PARITY UNPINNED for the kinematics: the reference calls Pinocchio (pin.forwardKinematics, computeJointJacobian, log,
Jlog6, integrate; pinocchio is not installed here and is a compiled dependency, not in the mount). Its published
algorithms are restated: URDF chain -> joint placements (fixed joints merged into the next joint's placement, as
pin.buildModelFromUrdf does), LOCAL-frame joint Jacobian, log6 / Jlog6 from pinocchio's explog.hpp. tests/test_ik_cpu.py
checks the restatement against first principles (Jlog6 against finite differences of log6 through scipy's matrix
exponential, the Jacobian against finite differences of FK, convergence of the IK iteration on the UR5e URDF of the
reference's assets, which is read in place, never copied).
"""
import xml.etree.ElementTree as ET

import numpy as np

PREC = 1e-8 ** 0.25      # pinocchio TaylorSeriesExpansion<double>::precision<3>() is of this order


def skew(v):
    return np.array([[0.0, -v[2], v[1]], [v[2], 0.0, -v[0]], [-v[1], v[0], 0.0]])


def rpy_to_matrix(rpy):
    """pin.rpy.rpyToMatrix: R = Rz(yaw) Ry(pitch) Rx(roll)."""
    r, p, y = rpy
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                     [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                     [-sp, cp * sr, cp * cr]])


def quat_wxyz_to_matrix(q):
    """pin.Quaternion(w, x, y, z).toRotationMatrix() (Eigen; the quaternion is used as given, not normalised)."""
    w, x, y, z = q
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    return np.array([[1 - (tyy + tzz), txy - twz, txz + twy],
                     [txy + twz, 1 - (txx + tzz), tyz - twx],
                     [txz - twy, tyz + twx, 1 - (txx + tyy)]])


def matrix_to_quat_wxyz(R):
    """Eigen's rotation-matrix -> quaternion (the branch structure of Eigen::Quaternion::operator=(Matrix3))."""
    t = R[0, 0] + R[1, 1] + R[2, 2]
    if t > 0:
        t = np.sqrt(t + 1.0)
        w = 0.5 * t
        t = 0.5 / t
        return np.array([w, (R[2, 1] - R[1, 2]) * t, (R[0, 2] - R[2, 0]) * t, (R[1, 0] - R[0, 1]) * t])
    i = 0
    if R[1, 1] > R[0, 0]:
        i = 1
    if R[2, 2] > R[i, i]:
        i = 2
    j, k = (i + 1) % 3, (i + 2) % 3
    t = np.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0)
    q = np.zeros(4)
    q[1 + i] = 0.5 * t
    t = 0.5 / t
    q[0] = (R[k, j] - R[j, k]) * t
    q[1 + j] = (R[j, i] + R[i, j]) * t
    q[1 + k] = (R[k, i] + R[i, k]) * t
    return q


def se3_from_pose(pose):
    """MathUtils.get_se3_from_pose (:34-36): pose = (tx, ty, tz, qw, qx, qy, qz) -> (R, p)."""
    pose = np.asarray(pose, np.float64)
    return quat_wxyz_to_matrix(pose[3:7]), pose[0:3].copy()


def pose_from_se3(R, p):
    """MathUtils.get_pose_from_se3 (:27-31)."""
    return np.concatenate([p, matrix_to_quat_wxyz(R)])


def se3_from_rel_pose(rel_pose):
    """MathUtils.get_se3_from_rel_pose (:44-46): (tx, ty, tz, roll, pitch, yaw) -> (R, p)."""
    rel_pose = np.asarray(rel_pose, np.float64)
    return rpy_to_matrix(rel_pose[3:6]), rel_pose[0:3].copy()


def se3_mul(a, b):
    return a[0] @ b[0], a[0] @ b[1] + a[1]


def se3_inv(a):
    return a[0].T, -a[0].T @ a[1]


def log3(R):
    """pinocchio log3: rotation matrix -> (axis * angle, angle)."""
    tr = R[0, 0] + R[1, 1] + R[2, 2]
    if tr >= 3.0:
        theta = 0.0
    elif tr <= -1.0:
        theta = np.pi
    else:
        theta = np.arccos((tr - 1.0) / 2.0)
    if theta >= np.pi - 1e-2:
        cphi = -(tr - 1.0) / 2.0
        beta = theta * theta / (1.0 + cphi)
        tmp = (np.diag(R) + cphi) * beta
        sgn = np.array([1.0 if R[2, 1] > R[1, 2] else -1.0, 1.0 if R[0, 2] > R[2, 0] else -1.0,
                        1.0 if R[1, 0] > R[0, 1] else -1.0])
        return sgn * np.sqrt(np.maximum(tmp, 0.0)), theta
    t = theta / np.sin(theta) if theta > PREC else 1.0 + theta * theta / 6.0
    return 0.5 * t * np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]]), theta


def log6(M):
    """pinocchio log6: SE3 -> Motion vector [v; w] (linear first)."""
    R, p = M
    w, t = log3(R)
    t2 = t * t
    if t < PREC:
        alpha = 1.0 - t2 / 12.0 - t2 * t2 / 720.0
        beta = 1.0 / 12.0 + t2 / 720.0
    else:
        st, ct = np.sin(t), np.cos(t)
        alpha = t * st / (2.0 * (1.0 - ct))
        beta = 1.0 / t2 - st / (2.0 * t * (1.0 - ct))
    v = alpha * p - 0.5 * np.cross(w, p) + beta * np.dot(w, p) * w
    return np.concatenate([v, w])


def Jlog3(theta, w):
    if theta < PREC:
        alpha = 1.0 / 12.0 + theta * theta / 720.0
        diag = 0.5 * (2.0 - theta * theta / 6.0)
    else:
        st, ct = np.sin(theta), np.cos(theta)
        st_1mct = st / (1.0 - ct)
        alpha = 1.0 / (theta * theta) - st_1mct / (2.0 * theta)
        diag = 0.5 * (theta * st_1mct)
    return alpha * np.outer(w, w) + diag * np.eye(3) + 0.5 * skew(w)


def Jlog6(M):
    """pinocchio Jlog6: d log6(M exp6(delta)) / d delta at 0 (right Jacobian of the log map), 6 x 6."""
    R, p = M
    w, t = log3(R)
    A = Jlog3(t, w)
    t2 = t * t
    if t < PREC:
        beta = 1.0 / 12.0 + t2 / 720.0
        beta_dot_over_theta = 1.0 / 360.0
    else:
        tinv = 1.0 / t
        t2inv = tinv * tinv
        st, ct = np.sin(t), np.cos(t)
        inv_2_2ct = 1.0 / (2.0 * (1.0 - ct))
        beta = t2inv - st * tinv * inv_2_2ct
        beta_dot_over_theta = -2.0 * t2inv * t2inv + (1.0 + st * tinv) * t2inv * inv_2_2ct
    wTp = np.dot(w, p)
    v3 = (beta_dot_over_theta * wTp) * w - (t2 * beta_dot_over_theta + 2.0 * beta) * p
    C = np.outer(v3, w) + beta * np.outer(w, p) + wTp * beta * np.eye(3) + 0.5 * skew(p)
    J = np.zeros((6, 6))
    J[:3, :3] = A
    J[3:, 3:] = A
    J[:3, 3:] = C @ A
    return J


class SerialArm:
    """A serial chain of revolute joints as Pinocchio models it: joint k has a placement (R_k, p_k) in its parent joint's
    frame and rotates about `axes[k]` in its own frame."""

    def __init__(self, placements, axes):
        self.placements = [(np.asarray(R, np.float64), np.asarray(p, np.float64)) for R, p in placements]
        self.axes = [np.asarray(a, np.float64) for a in axes]
        self.n = len(self.axes)

    @classmethod
    def from_urdf(cls, path, root_pose=None, exclude_joint_names=()):
        """pin.buildModelFromUrdf(path) for a single chain: walk from the root link, merge fixed (and excluded, fixed at
        0) joints into the placement of the next revolute joint; root_pose = ArmConfig.arm_root_pose is applied to the
        first joint's placement (ArmManager.py:67-71)."""
        root = ET.parse(path).getroot()
        joints = {}
        children = set()
        for j in root.findall("joint"):
            parent, child = j.find("parent").get("link"), j.find("child").get("link")
            org = j.find("origin")
            xyz = [float(v) for v in (org.get("xyz", "0 0 0") if org is not None else "0 0 0").split()]
            rpy = [float(v) for v in (org.get("rpy", "0 0 0") if org is not None else "0 0 0").split()]
            ax = j.find("axis")
            axis = [float(v) for v in ax.get("xyz").split()] if ax is not None else [1.0, 0.0, 0.0]
            joints.setdefault(parent, []).append((j.get("name"), j.get("type"), child, rpy_to_matrix(rpy), np.array(xyz), np.array(axis)))
            children.add(child)
        links = [l.get("name") for l in root.findall("link")]
        base = [l for l in links if l not in children][0]
        placements, axes = [], []
        acc = (np.eye(3), np.zeros(3)) if root_pose is None else se3_from_pose(root_pose)
        link = base
        while link in joints:
            name, typ, child, R, p, axis = joints[link][0]        # single chain: first child
            acc = se3_mul(acc, (R, p))
            if typ in ("revolute", "continuous") and name not in exclude_joint_names:
                placements.append(acc)
                axes.append(axis / np.linalg.norm(axis))
                acc = (np.eye(3), np.zeros(3))
            link = child
        return cls(placements, axes)

    def joint_frames(self, q):
        """pin.forwardKinematics: oMi for joints 1..n (returned 0-based)."""
        out = []
        cur = (np.eye(3), np.zeros(3))
        for k in range(self.n):
            a = self.axes[k]
            K = skew(a)
            Rq = np.eye(3) + np.sin(q[k]) * K + (1.0 - np.cos(q[k])) * (K @ K)
            cur = se3_mul(se3_mul(cur, self.placements[k]), (Rq, np.zeros(3)))
            out.append(cur)
        return out

    def joint_jacobian_local(self, q, joint_id):
        """pin.computeJointJacobian (LOCAL frame of joint `joint_id`, 1-based as in Pinocchio): 6 x n, rows [v; w]."""
        frames = self.joint_frames(q)
        oMj = frames[joint_id - 1]
        J = np.zeros((6, self.n))
        for k in range(joint_id):
            R, p = se3_mul(se3_inv(oMj), frames[k])
            a = R @ self.axes[k]
            J[:3, k] = np.cross(p, a)
            J[3:, k] = a
        return J


def ik_step(arm, q, target, joint_id):
    """ArmManager.inverse_kinematics (:230-254): one damped least-squares step towards `target` (R, p); returns the new
    arm joint positions."""
    cur = arm.joint_frames(q)[joint_id - 1]
    err_se3 = se3_mul(se3_inv(cur), target)                       # current_se3.actInv(target_se3)
    err = log6(err_se3)
    J = arm.joint_jacobian_local(q, joint_id)
    J = -1.0 * (Jlog6(se3_inv(err_se3)) @ J)
    damping_scale = 1e-6
    delta = -1.0 * J.T.dot(np.linalg.solve(J.dot(J.T) + (np.dot(err, err) + damping_scale) * np.identity(6), err))
    return np.asarray(q, np.float64) + delta                      # pin.integrate for revolute joints


class ArmManagerOracle:
    """One episode of ArmManager's command handling (:88-160) on a SerialArm: state = arm joint positions, gripper joint
    positions, target SE3."""

    def __init__(self, arm, joint_id, init_arm_joint_pos, init_gripper_joint_pos, arm_joint_idxes, gripper_joint_idxes,
                 gripper_low, gripper_high, eef_idx=0):
        self.arm, self.joint_id = arm, joint_id
        self.init_arm, self.init_grip = np.asarray(init_arm_joint_pos, np.float64), np.asarray(init_gripper_joint_pos, np.float64)
        self.arm_idx, self.grip_idx = np.asarray(arm_joint_idxes), np.asarray(gripper_joint_idxes)
        self.low, self.high = np.asarray(gripper_low, np.float64), np.asarray(gripper_high, np.float64)
        self.eef_idx = eef_idx
        self.reset()

    def reset(self):
        self.arm_joint_pos = self.init_arm.copy()
        self.gripper_joint_pos = self.init_grip.copy()
        self.target = self.arm.joint_frames(self.arm_joint_pos)[self.joint_id - 1]

    def set_command_data(self, key, command, is_skip=False):
        command = np.asarray(command, np.float64)
        if key == "command_joint_pos":
            self.arm_joint_pos = command[self.arm_idx].copy()
            self.target = self.arm.joint_frames(self.arm_joint_pos)[self.joint_id - 1]
            self.gripper_joint_pos = np.clip(command[self.grip_idx], self.low, self.high)
        elif key == "command_joint_pos_rel":
            arm, grip = self.arm_joint_pos.copy(), self.gripper_joint_pos.copy()
            if not is_skip:
                arm += command[self.arm_idx]
                grip += command[self.grip_idx]
            self.set_command_data("command_joint_pos", self._join(arm, grip))
        elif key == "command_gripper_joint_pos":
            self.gripper_joint_pos = np.clip(command, self.low, self.high)
        elif key == "command_eef_pose":
            self.target = se3_from_pose(command[7 * self.eef_idx:7 * (self.eef_idx + 1)])
            self.arm_joint_pos = ik_step(self.arm, self.arm_joint_pos, self.target, self.joint_id)
        elif key == "command_eef_pose_rel":
            # set_command_data does not forward is_skip to set_command_eef_pose_rel (ArmManager.py:113-117): always applied
            self.target = se3_mul(self.target, se3_from_rel_pose(command[6 * self.eef_idx:6 * (self.eef_idx + 1)]))
            self.arm_joint_pos = ik_step(self.arm, self.arm_joint_pos, self.target, self.joint_id)
        else:
            raise ValueError(f"[ArmManager] Invalid command data key: {key}")

    def _join(self, arm, grip):
        out = np.zeros(len(self.arm_idx) + len(self.grip_idx))
        out[self.arm_idx], out[self.grip_idx] = arm, grip
        return out

    def get_command_joint_pos(self):
        return self._join(self.arm_joint_pos, self.gripper_joint_pos)

    def get_command_eef_pose(self):
        return pose_from_se3(*self.target)


def split_action(policy_action, action_keys, key_dims):
    """RolloutBase.set_command_data (:496-509): consecutive slices of the action vector, one per action key."""
    out, i = {}, 0
    for key in action_keys:
        out[key] = policy_action[..., i:i + key_dims[key]]
        i += key_dims[key]
    return out
