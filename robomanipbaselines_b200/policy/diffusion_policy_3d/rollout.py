"""RolloutDiffusionPolicy3d: drop-in for robo_manip_baselines.policy.diffusion_policy_3d.RolloutDiffusionPolicy3d
(policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:29-231) on the B200 engine; one episode per object."""
import numpy as np
import torch

from ...common.batched_diffusion_rollout import BatchedDiffusionRollout, Dp3Frontend
from ...common.rollout_base import RolloutBase
from ...common.schedulers import DDIMScheduler
from .policy import DP3


class RolloutDiffusionPolicy3d(RolloutBase):
    policy_name = "DiffusionPolicy3d"

    def setup_policy(self):
        """:36-59."""
        self.print_policy_info()
        data = self.model_meta_info["data"]
        print(f"  - use ema: {self.model_meta_info['policy'].get('use_ema')}")
        print(f"  - horizon: {data['horizon']}, obs steps: {data['n_obs_steps']}, action steps: {data['n_action_steps']}")
        print(f"  - with color: {data.get('use_pc_color')}, num points: {data['num_points']}, image size: {data['image_size']}, "
              f"min bound: {data.get('min_bound')}, max bound: {data.get('max_bound')}")
        noise_scheduler = DDIMScheduler(**self.model_meta_info["policy"]["noise_scheduler_args"])
        self.policy = DP3(noise_scheduler=noise_scheduler, **self.model_meta_info["policy"]["args"])
        if self.args.checkpoint is not None:
            self.load_ckpt()
        self.rollout = None
        self.pointcloud_plot = None

    def reset_variables(self):
        """:73-81."""
        super().reset_variables()
        self.pointcloud_plot = None
        if getattr(self, "rollout", None) is not None:
            self.rollout.reset()

    def _make_rollout(self):
        fovy = self.env.unwrapped.get_camera_fovy(self.camera_names[0])                     # :131
        frontend = Dp3Frontend(1, self.model_meta_info, fovy, device=str(self.device))
        r = BatchedDiffusionRollout(self.policy, 1, self.model_meta_info, device=str(self.device),
                                    obs_keys=("point_cloud",), frontend=frontend)
        r.reset()
        return r

    def get_depth_and_rgb(self):
        """The frames get_pointcloud reads (:127-130), moved to the device as they are."""
        cam = self.camera_names[0]
        rgb = torch.from_numpy(np.ascontiguousarray(self.info["rgb_images"][cam][np.newaxis])).to(self.device)
        depth = torch.from_numpy(np.ascontiguousarray(self.info["depth_images"][cam][np.newaxis], dtype=np.float32)).to(self.device)
        return rgb, depth

    def infer_policy(self, x_init=None, step_noise=None):
        """:83-101. get_state (:103-123) and get_pointcloud (:125-171: two cv2.resize, back-projection, crop, FPS,
        normalisation) run on the device inside the rollout."""
        if self.rollout is None:
            self.rollout = self._make_rollout()
        rgb, depth = self.get_depth_and_rgb()
        action = self.rollout.infer_env(self.get_state(), x_init=x_init, step_noise=step_noise, rgb_u8=rgb, depth=depth)
        self.store_action(action[0].cpu().numpy())

    def draw_plot(self):
        """:173-231, headless: the newest (normalised) cloud of the ring buffer stands in for pointcloud_plot."""
        pc = None
        if self.rollout is not None and "point_cloud" in self.rollout.obs_buf:
            pc = self.rollout.obs_buf["point_cloud"][0, -1].cpu().numpy()
        self.pointcloud_plot = pc
        self.plot_data = {"pointcloud": pc, "action": self.plot_action()}
        return self.plot_data
