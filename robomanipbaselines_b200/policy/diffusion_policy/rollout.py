"""RolloutDiffusionPolicy: drop-in for robo_manip_baselines.policy.diffusion_policy.RolloutDiffusionPolicy
(policy/diffusion_policy/RolloutDiffusionPolicy.py:24-160) on the B200 engine; one episode per object."""
import torch

from ...common.batched_diffusion_rollout import BatchedDiffusionRollout, DpImageFrontend
from ...common.rollout_base import RolloutBase, rgb_image_key
from ...common.schedulers import DDPMScheduler
from .policy import DiffusionUnetHybridImagePolicy


class RolloutDiffusionPolicy(RolloutBase):
    policy_name = "DiffusionPolicy"

    def setup_policy(self):
        """:25-46."""
        self.print_policy_info()
        data = self.model_meta_info["data"]
        print(f"  - use ema: {self.model_meta_info['policy'].get('use_ema')}")
        print(f"  - horizon: {data['horizon']}, obs steps: {data['n_obs_steps']}, action steps: {data['n_action_steps']}")
        print(f"  - image size: {data['image_size']}, image crop size: {data.get('image_crop_size')}")
        noise_scheduler = DDPMScheduler(**self.model_meta_info["policy"]["noise_scheduler_args"])
        self.policy = DiffusionUnetHybridImagePolicy(noise_scheduler=noise_scheduler,
                                                     **self.model_meta_info["policy"]["args"])
        if self.args.checkpoint is not None:
            self.load_ckpt()
        self.rollout = None

    def reset_variables(self):
        """:60-64 -- state_buf / images_buf / policy_action_buf are the device ring buffers and FIFO."""
        super().reset_variables()
        if getattr(self, "rollout", None) is not None:
            self.rollout.reset()

    def _make_rollout(self):
        keys = tuple(rgb_image_key(c) for c in self.camera_names)
        frontend = DpImageFrontend(self.camera_names, self.model_meta_info["data"]["image_size"])
        r = BatchedDiffusionRollout(self.policy, 1, self.model_meta_info, device=str(self.device), obs_keys=keys,
                                    frontend=frontend)
        r.reset()
        return r

    def infer_policy(self, x_init=None, step_noise=None):
        """:66-85. get_state (:87-108) and get_images (:110-143) run on the device inside the rollout: normalise,
        cv2-exact resize, `* 2 - 1`, ring buffers pushed only when the action FIFO is empty."""
        if self.rollout is None:
            self.rollout = self._make_rollout()
        action = self.rollout.infer_env(self.get_state(), x_init=x_init, step_noise=step_noise, rgb_u8=self.get_images())
        self.store_action(action[0].cpu().numpy())

    def draw_plot(self):
        """:145-160, headless."""
        self.plot_data = {"images": self.plot_images(), "action": self.plot_action()}
        return self.plot_data
