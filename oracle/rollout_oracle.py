"""TEST INFRASTRUCTURE ONLY (parity checker; never on the product path).

numpy-f64 restatement of the host side of the reference's ACT rollout step:
  * normalize_data / denormalize_data      -- robo_manip_baselines/common/utils/DataUtils.py:9-40
  * RolloutAct.infer_policy's chunk history, temporal ensembling and FIFO mode
                                           -- robo_manip_baselines/policy/act/RolloutAct.py:62-101
  * RolloutBase.get_state / get_images     -- robo_manip_baselines/common/base/RolloutBase.py:463-490

Parity pinned (ACT): tests/golden/make_golden.py executes the reference's own RolloutAct.infer_policy /
normalize_data / denormalize_data (imported from /root/reference with stub modules for the GUI / robotics
dependencies) on seeded inputs and stores the results in tests/golden/rollout_act_golden.npz; tests/test_oracle.py
checks this restatement against those vectors bit-for-bit.
Parity pinned (Diffusion Policy / DP3 glue): tests/golden/make_dp_rollout_golden.py executes the reference's own
RolloutDiffusionPolicy(3d).infer_policy / get_state / get_images / get_pointcloud on a stand-in `self` with a recording
fake predict_action; tests/test_oracle.py replays DiffusionRolloutOracle + dp_get_image / pointcloud_oracle against
tests/golden/dp_rollout_golden.npz bit-for-bit (policy inputs at every inference, inference steps, float64 actions).
"""
import numpy as np


def normalize_data(data, stats):
    """DataUtils.py:9-23."""
    norm_type = stats["norm_config"]["type"] if "norm_config" in stats else "gaussian"
    if norm_type == "gaussian":
        return (data - stats["mean"]) / stats["std"]
    if norm_type == "limits":
        cfg = stats["norm_config"]
        scale = (cfg["out_max"] - cfg["out_min"]) / stats["range"]
        return scale * (data - stats["min"]) + cfg["out_min"]
    raise ValueError(f"invalid normalization type: {norm_type}")


def denormalize_data(data, stats):
    """DataUtils.py:26-40."""
    norm_type = stats["norm_config"]["type"] if "norm_config" in stats else "gaussian"
    if norm_type == "gaussian":
        return stats["std"] * data + stats["mean"]
    if norm_type == "limits":
        cfg = stats["norm_config"]
        scale = stats["range"] / (cfg["out_max"] - cfg["out_min"])
        return scale * (data - cfg["out_min"]) + stats["min"]
    raise ValueError(f"invalid normalization type: {norm_type}")


class ActRolloutOracle:
    """One episode of RolloutAct.infer_policy (RolloutAct.py:68-101) with the policy call factored out.

    `step(chunk)` takes the (C, Da) float32 chunk the policy returned at this call (or None in FIFO mode when the
    reference would not call the policy) and returns the denormalised float64 action.
    """

    def __init__(self, chunk_size, action_dim, action_stats, no_temp_ensem=False):
        self.chunk_size = chunk_size
        self.action_dim = action_dim
        self.action_stats = action_stats
        self.no_temp_ensem = no_temp_ensem
        self.reset()

    def reset(self):
        """RolloutAct.reset_variables (:62-66)."""
        self.policy_action_buf = []
        self.policy_action_buf_history = []

    def needs_inference(self):
        """:70"""
        return (not self.no_temp_ensem) or (len(self.policy_action_buf) == 0)

    def step(self, chunk):
        if self.needs_inference():
            self.policy_action_buf = list(np.asarray(chunk).astype(np.float64))        # :74-76
            if not self.no_temp_ensem:
                self.policy_action_buf_history.append(self.policy_action_buf)           # :77-83
                if len(self.policy_action_buf_history) > self.chunk_size:
                    self.policy_action_buf_history.pop(0)
        if self.no_temp_ensem:
            action = self.policy_action_buf.pop(0)                                      # :86-87
        else:
            k = 0.01                                                                    # :90-97
            exp_weights = np.exp(-k * np.arange(len(self.policy_action_buf_history)))
            exp_weights = exp_weights / exp_weights.sum()
            action = np.zeros(self.action_dim)
            for action_idx, buf in enumerate(reversed(self.policy_action_buf_history)):
                action += exp_weights[::-1][action_idx] * buf[action_idx]
        self.normalized_action = action
        return denormalize_data(action, self.action_stats)                              # :98

    def ring_view(self, n_calls):
        """History as the engine's ring buffer: slot (j mod C) holds chunk j. Returns (C, C, Da) f32 + filled mask."""
        C = self.chunk_size
        ring = np.zeros((C, C, self.action_dim), dtype=np.float32)
        filled = np.zeros(C, dtype=bool)
        m = len(self.policy_action_buf_history)
        for age, buf in enumerate(reversed(self.policy_action_buf_history)):
            j = n_calls - 1 - age
            ring[j % C] = np.asarray(buf, dtype=np.float32)
            filled[j % C] = True
        assert m == min(n_calls, C)
        return ring, filled


def get_state(raw_state, state_stats):
    """RolloutBase.get_state (:463-477) for one episode: f64 concat -> normalize -> float32 (1, Ds)."""
    state = normalize_data(np.asarray(raw_state, dtype=np.float64), state_stats)
    return state[np.newaxis].astype(np.float32)


def get_images(rgb_images):
    """RolloutBase.get_images (:479-490): list of (H, W, 3) uint8 -> float32 (1, Ncam, 3, H, W) in [0, 1]."""
    images = np.stack(rgb_images, axis=0)
    images = np.moveaxis(images, -1, -3)
    # v2.ToDtype(float32, scale=True) multiplies by 1/255 in float32 (RolloutBase.py:353,488)
    return (images.astype(np.float32) * np.float32(1.0 / 255.0))[np.newaxis]


def dp_get_image(rgb_image, image_size):
    """RolloutDiffusionPolicy.get_images (:113-125) for one camera frame: cv2.resize (the reference's own dependency) ->
    HWC to CHW -> uint8 to float32 with 1/255 scaling (v2.ToDtype(scale=True), RolloutBase.py:353) -> image * 2.0 - 1.0,
    all in float32. uint8 (H, W, 3) -> float32 (3, h, w)."""
    import cv2

    image = cv2.resize(rgb_image, tuple(image_size))
    image = np.moveaxis(image, -1, -3).astype(np.float32) * np.float32(1.0 / 255.0)
    return image * np.float32(2.0) - np.float32(1.0)


def dp3_resize(rgb_image, depth_image, image_size):
    """RolloutDiffusionPolicy3d.get_pointcloud's two cv2.resize calls (:134-137)."""
    import cv2

    return cv2.resize(rgb_image, tuple(image_size)), cv2.resize(depth_image, tuple(image_size))


class DiffusionRolloutOracle:
    """One episode of RolloutDiffusionPolicy(3d).infer_policy with the policy call factored out
    (policy/diffusion_policy/RolloutDiffusionPolicy.py:60-108, policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:73-123).

    Observation ring buffers are pushed ONLY when the policy is called (get_state / get_images / get_pointcloud sit
    inside the `if len(policy_action_buf) == 0` branch): the first call fills all n_obs_steps slots with the same
    observation, later calls pop the oldest and append. The action FIFO holds the n_action_steps actions of one
    inference as float64 and pops one per policy step; each popped action is denormalised.
    """

    def __init__(self, n_obs_steps, state_stats, action_stats):
        self.n_obs_steps = n_obs_steps
        self.state_stats, self.action_stats = state_stats, action_stats
        self.reset()

    def reset(self):
        self.obs_buf = None              # dict name -> list of n_obs_steps arrays
        self.policy_action_buf = None

    def needs_inference(self):
        return self.policy_action_buf is None or len(self.policy_action_buf) == 0

    def push_obs(self, raw_state, **others):
        """-> dict of stacked (n_obs_steps, ...) arrays, state normalised and cast to float32 (get_state :87-108)."""
        new = {"state": normalize_data(np.asarray(raw_state, np.float64), self.state_stats).astype(np.float32)}
        new.update({k: np.asarray(v) for k, v in others.items()})
        if self.obs_buf is None:
            self.obs_buf = {k: [v for _ in range(self.n_obs_steps)] for k, v in new.items()}
        else:
            for k, v in new.items():
                self.obs_buf[k].pop(0)
                self.obs_buf[k].append(v)
        return {k: np.stack(v, axis=0) for k, v in self.obs_buf.items()}

    def step(self, actions=None):
        """actions: (n_action_steps, Da) float32 from predict_action when needs_inference() was True."""
        if self.needs_inference():
            self.policy_action_buf = list(np.asarray(actions).astype(np.float64))
        return denormalize_data(self.policy_action_buf.pop(0), self.action_stats)
