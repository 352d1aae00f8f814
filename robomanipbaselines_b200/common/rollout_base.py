"""RolloutBase: the hot-path half of robo_manip_baselines.common.base.RolloutBase (common/base/RolloutBase.py), with the
reference's attribute and method names, on top of the batched device rollouts.

What is kept (the part `RolloutPhase.pre_update` drives, RolloutBase.py:56-62, and what a subclass must provide):
    setup_model_meta_info (:289-309)   state/action keys, camera names, dims, default skip
    setup_policy          (:311-313)   abstract: build the policy module, load the checkpoint
    load_ckpt             (:376-385)   strict load_state_dict of a reference checkpoint, policy.eval()
    reset_variables       (:364-365)   per-episode buffers
    infer_policy          (:459-461)   abstract: one policy step -> self.policy_action (denormalised float64)
    get_state / get_images (:463-490)  raw env observation -> device tensors (normalisation happens on the device)
    set_command_data      (:496-509)   split policy_action by action key and hand it to the motion manager
    draw_plot             (:492-494)   abstract; here headless: fills self.plot_data (no window system on this box)
    print_statistics / calc_model_size (:563-571,607-616)
What is NOT here (out of scope): argparse CLI, gymnasium env construction, phase manager, matplotlib /
cv2 windows, yaml result file. The driver that owns the env sets `self.obs` / `self.info` (as the reference's run loop
does, :404) and calls `policy_step()`, which is RolloutPhase.pre_update.

One object drives ONE episode like the reference (B = 1 through the same kernels); the batched classes
(BatchedRollout, BatchedDiffusionRollout) are the B > 1 surface.
"""
import os
import pickle
import time
from abc import ABC, abstractmethod
from types import SimpleNamespace

import numpy as np
import torch

ARM_JOINT_KEYS = ("measured_joint_pos", "command_joint_pos", "measured_joint_pos_rel", "command_joint_pos_rel",
                  "measured_joint_vel", "command_joint_vel", "measured_joint_torque", "command_joint_torque")
GRIPPER_KEYS = ("measured_gripper_joint_pos", "command_gripper_joint_pos")
EEF_POSE_KEYS = ("measured_eef_pose", "command_eef_pose")
EEF_6D_KEYS = ("measured_eef_pose_rel", "command_eef_pose_rel", "measured_eef_vel", "command_eef_vel",
               "measured_eef_wrench", "command_eef_wrench")
MOBILE_KEYS = ("measured_mobile_omni_vel", "command_mobile_omni_vel")


def data_key_dim(key, env):
    """DataKey.get_dim (common/data/DataKey.py:96-148) on a duck-typed env: body configs that have arm joints are arms."""
    bodies = [b for b in env.unwrapped.body_config_list if hasattr(b, "arm_joint_idxes")]
    if key == "time":
        return 1
    if key in ARM_JOINT_KEYS:
        return sum(len(b.arm_joint_idxes) + len(b.gripper_joint_idxes) for b in bodies)
    if key in GRIPPER_KEYS:
        return sum(len(b.gripper_joint_idxes) for b in bodies)
    if key in EEF_POSE_KEYS or key in EEF_6D_KEYS:
        n_eef = len([b for b in bodies if getattr(b, "eef_idx", None) is not None])
        return (7 if key in EEF_POSE_KEYS else 6) * n_eef
    if key in MOBILE_KEYS:
        return 3
    raise ValueError(f"[DataKey] Invalid data key: {key}")


def rgb_image_key(camera_name):
    """DataKey.get_rgb_image_key (common/data/DataKey.py)."""
    return camera_name.lower() + "_rgb_image"


class RolloutBase(ABC):
    require_task_desc = False
    policy_name = "Policy"

    def __init__(self, env, checkpoint=None, model_meta_info=None, motion_manager=None, device="cuda:0", **args):
        """env: the reference's gymnasium env (duck-typed: only `.unwrapped` attributes are read).
        checkpoint: path to policy_*.ckpt; model_meta_info.pkl is read from its directory (RolloutBase.py:289-295)
        unless `model_meta_info` is given as a dict. motion_manager: the reference's MotionManager (get_data /
        set_command_data); optional when the caller passes raw states itself. args: the reference's CLI options that
        matter on the path (skip, no_plot, no_temp_ensem, task_desc, seed, ...)."""
        self.env = env
        self.args = SimpleNamespace(checkpoint=checkpoint, skip=None, skip_draw=None, no_plot=True, seed=0,
                                    no_temp_ensem=False, task_desc=None)
        for k, v in args.items():
            setattr(self.args, k, v)
        if self.require_task_desc and self.args.task_desc is None:
            raise ValueError(f"[{self.__class__.__name__}] task_desc is required")
        self.device = torch.device(device)
        self.motion_manager = motion_manager
        self.setup_model_meta_info(model_meta_info)
        self.setup_policy()
        self.setup_variables()
        self.reset_variables()

    # ---- RolloutBase.setup_model_meta_info (:289-309)
    def setup_model_meta_info(self, model_meta_info=None):
        if model_meta_info is None:
            path = os.path.join(os.path.split(self.args.checkpoint)[0], "model_meta_info.pkl")
            with open(path, "rb") as f:
                model_meta_info = pickle.load(f)
            print(f"[{self.__class__.__name__}] Load model meta info: {path}")
        self.model_meta_info = model_meta_info
        self.state_keys = model_meta_info["state"]["keys"]
        self.action_keys = model_meta_info["action"]["keys"]
        self.camera_names = model_meta_info["image"]["camera_names"]
        self.state_dim = len(model_meta_info["state"]["example"])
        self.action_dim = len(model_meta_info["action"]["example"])
        if self.args.skip is None:
            self.args.skip = model_meta_info["data"].get("skip", 1)
        if self.args.skip_draw is None:
            self.args.skip_draw = self.args.skip

    @abstractmethod
    def setup_policy(self):
        pass

    def setup_variables(self):
        self.episode_idx = 0
        self.rollout_time_idx = 0
        self.result = {key: [] for key in ("success", "reward", "duration")}
        self.inference_duration_list = []
        self.plot_data = {}
        self.obs, self.info, self.reward = None, None, 0.0

    def reset_variables(self):
        """:364-365 -- subclasses extend it (and reset their device-side episode state)."""
        self.policy_action_list = np.empty((0, self.action_dim))
        self.rollout_time_idx = 0

    def print_policy_info(self):
        print(f"[{self.__class__.__name__}] Construct {self.policy_name} policy.\n"
              f"  - state dim: {self.state_dim}, action dim: {self.action_dim}, camera num: {len(self.camera_names)}\n"
              f"  - state keys: {self.state_keys}\n  - action keys: {self.action_keys}\n"
              f"  - camera names: {self.camera_names}\n  - skip: {self.args.skip}")

    def load_ckpt(self, device=None):
        """:376-385. The policy mirrors keep their parameters on the host (the engine owns the device copy), so the
        checkpoint is mapped to the CPU; strict=True as in the reference."""
        print(f"[{self.__class__.__name__}] Load {self.args.checkpoint}")
        if device is not None:
            self.device = torch.device(device)
        self.policy.load_state_dict(torch.load(self.args.checkpoint, map_location="cpu", weights_only=True))
        self.policy.eval()

    # ---- one control tick of RolloutPhase (:56-62, :70)
    def policy_step(self):
        if self.rollout_time_idx % self.args.skip == 0:
            t0 = time.time()
            self.infer_policy()
            self.inference_duration_list.append(time.time() - t0)
        if self.motion_manager is not None:
            self.set_command_data()
        if (not self.args.no_plot) and self.rollout_time_idx % self.args.skip_draw == 0:
            self.draw_plot()
        self.rollout_time_idx += 1
        return self.policy_action

    @abstractmethod
    def infer_policy(self):
        pass

    # ---- observations: the raw values; normalize_data / ToDtype(scale) run inside the device step
    def get_raw_state(self):
        """The concatenation of :464-472 (float64), before normalize_data."""
        if len(self.state_keys) == 0:
            return np.zeros(0, dtype=np.float64)
        return np.concatenate([np.asarray(self.motion_manager.get_data(k, self.obs), dtype=np.float64)
                               for k in self.state_keys])

    def get_state(self):
        """(1, Ds) float64 raw state on the device (the reference returns the normalised float32 tensor; here
        normalize_data + the float32 cast are fused into the device step, bit-exact)."""
        return torch.from_numpy(self.get_raw_state()[np.newaxis]).to(self.device)

    def get_images(self):
        """(1, Ncam, H, W, 3) uint8 on the device: the env's frames as they are (:481-484); HWC -> CHW, 1/255 scaling
        (:486-488) and the ImageNet normalisation are fused into the first kernel of the policy."""
        images = np.stack([self.info["rgb_images"][c] for c in self.camera_names], axis=0)
        return torch.from_numpy(np.ascontiguousarray(images[np.newaxis])).to(self.device)

    def store_action(self, action):
        """The tail of every infer_policy (e.g. RolloutAct.py:98-101)."""
        self.policy_action = np.asarray(action, dtype=np.float64)
        self.policy_action_list = np.concatenate([self.policy_action_list, self.policy_action[np.newaxis]])

    def set_command_data(self, action_keys=None):
        """:496-509."""
        if action_keys is None:
            action_keys = self.action_keys
        is_skip = self.rollout_time_idx % self.args.skip != 0
        action_idx = 0
        for key in action_keys:
            dim = data_key_dim(key, self.env)
            self.motion_manager.set_command_data(key, self.policy_action[action_idx:action_idx + dim], is_skip)
            action_idx += dim

    @abstractmethod
    def draw_plot(self):
        pass

    def plot_images(self):
        return {c: self.info["rgb_images"][c] for c in self.camera_names}

    def plot_action(self, history_size=100):
        return self.policy_action_list[-history_size:]

    def record_result(self, elapsed_duration):
        """RolloutPhase.check_transition's tally (:101-107)."""
        self.result["success"].append(bool(self.reward >= 1.0))
        self.result["reward"].append(float(self.reward))
        self.result["duration"].append(float(elapsed_duration))

    def calc_model_size(self):
        """:607-616 -- works because self.policy stays an nn.Module."""
        size = sum(p.nelement() * p.element_size() for p in self.policy.parameters())
        return size + sum(b.nelement() * b.element_size() for b in self.policy.buffers())

    def print_statistics(self):
        d = np.array(self.inference_duration_list) if self.inference_duration_list else np.zeros(1)
        print(f"[{self.__class__.__name__}] Statistics on policy inference\n"
              f"  - Policy model size [MB] | {self.calc_model_size() / 1024**2:.2f}\n"
              f"  - GPU memory usage [GB] | {torch.cuda.max_memory_reserved() / 1024**3:.3f}\n"
              f"  - Inference duration [s] | mean: {d.mean():.2e}, std: {d.std():.2e} min: {d.min():.2e}, max: {d.max():.2e}")
