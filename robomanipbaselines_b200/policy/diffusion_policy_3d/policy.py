"""Drop-in for `DP3` (3D Diffusion Policy) at inference time, as the reference builds and calls it:

    DP3(noise_scheduler=DDIMScheduler(**noise_scheduler_args), **model_meta_info["policy"]["args"])
        policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:50-56 ; args from TrainDiffusionPolicy3d.py:185-218
    policy.predict_action({"state": (B, To, Ds), "point_cloud": (B, To, Np, n_point_dim)})["action"]   :83-101

DP3's source (third_party/3D-Diffusion-Policy @ 95ad86a) is not in the reference mount; module names follow the public
upstream code ("obs_encoder.extractor.mlp.0.weight", "model.down_modules...") so a reference checkpoint loads strictly.
"""
import torch
from torch import nn

from ...engine import PointNetEngine
from ..diffusion_common import DiffusionPolicyBase, register_tree


def _get(cfg, key, default=None):
    if cfg is None:
        return default
    if isinstance(cfg, dict):
        return cfg.get(key, default)
    return getattr(cfg, key, default)


def dp3_encoder_param_shapes(in_channels, out_channels, state_dim):
    """DP3Encoder(PointNetEncoderXYZ use_layernorm=True final_norm="layernorm", state_mlp_size=(64, 64))."""
    s = {}
    dims = [in_channels, 64, 128, 256]
    for i in range(3):
        s[f"extractor.mlp.{3 * i}.weight"] = (dims[i + 1], dims[i])
        s[f"extractor.mlp.{3 * i}.bias"] = (dims[i + 1],)
        s[f"extractor.mlp.{3 * i + 1}.weight"] = (dims[i + 1],)
        s[f"extractor.mlp.{3 * i + 1}.bias"] = (dims[i + 1],)
    s["extractor.final_projection.0.weight"] = (out_channels, 256)
    s["extractor.final_projection.0.bias"] = (out_channels,)
    s["extractor.final_projection.1.weight"] = (out_channels,)
    s["extractor.final_projection.1.bias"] = (out_channels,)
    if state_dim > 0:
        s["state_mlp.0.weight"] = (64, state_dim)
        s["state_mlp.0.bias"] = (64,)
        s["state_mlp.2.weight"] = (64, 64)
        s["state_mlp.2.bias"] = (64,)
    return s


class DP3(DiffusionPolicyBase):
    def __init__(self, shape_meta, noise_scheduler, horizon, n_action_steps, n_obs_steps, num_inference_steps=None,
                 obs_as_global_cond=True, diffusion_step_embed_dim=256, down_dims=(256, 512, 1024), kernel_size=5,
                 n_groups=8, condition_type="film", use_down_condition=True, use_mid_condition=True,
                 use_up_condition=True, encoder_output_dim=256, crop_shape=None, use_pc_color=False,
                 pointnet_type="pointnet", pointcloud_encoder_cfg=None, max_batch=1, **kwargs):
        if not obs_as_global_cond:
            raise NotImplementedError("RoboManipBaselines uses obs_as_global_cond=True (TrainDiffusionPolicy3d.py:190)")
        if condition_type != "film" or pointnet_type != "pointnet":
            raise NotImplementedError("only condition_type='film' / pointnet_type='pointnet' are supported")
        action_dim = int(shape_meta["action"]["shape"][0])
        obs = shape_meta["obs"]
        self_state_dim = int(obs["state"]["shape"][0]) if "state" in obs else 0
        n_point_dim = int(obs["point_cloud"]["shape"][1])
        num_points = int(obs["point_cloud"]["shape"][0])
        out_ch = int(_get(pointcloud_encoder_cfg, "out_channels", encoder_output_dim))
        feat = out_ch + (64 if self_state_dim > 0 else 0)
        super().__init__(noise_scheduler, action_dim, feat, horizon, n_action_steps, n_obs_steps, num_inference_steps,
                         diffusion_step_embed_dim, down_dims, kernel_size, n_groups, max_batch)
        self.state_dim, self.num_points, self.n_point_dim = self_state_dim, num_points, n_point_dim
        self.use_pc_color = bool(use_pc_color)
        self.in_channels = 6 if self.use_pc_color else 3
        self.encoder_out = out_ch
        self.obs_encoder = register_tree(nn.Module(), dp3_encoder_param_shapes(self.in_channels, out_ch, self_state_dim))
        self._pn = None

    def _invalidate(self):
        super()._invalidate()
        if getattr(self, "_pn", None) is not None:
            self._pn.close()
        self._pn = None

    def pointnet_for(self, n_clouds, device):
        dev = device.index if device.index is not None else torch.cuda.current_device()
        if self._pn is None or self._pn.max_clouds < n_clouds or self._pn.device.index != dev:
            if self._pn is not None:
                self._pn.close()
            self._pn = PointNetEngine(max_clouds=max(n_clouds, self.max_batch * self.n_obs_steps),
                                      num_points=self.num_points, in_channels=self.in_channels,
                                      out_channels=self.encoder_out, state_dim=self.state_dim, device=dev)
            self._pn.load_state_dict(self.obs_encoder.state_dict())
        return self._pn

    def encode_obs(self, obs_dict):
        """{"point_cloud": (B, To, Np, n_point_dim), "state": (B, To, Ds)} -> global_cond (B, To * feat)."""
        pc = obs_dict["point_cloud"]
        if not pc.is_cuda:
            raise RuntimeError("DP3 runs on the B200 engine only: move inputs to cuda (no CPU fallback)")
        B, To = pc.shape[0], self.n_obs_steps
        pc = pc[:, :To].to(torch.float32).contiguous().view(B * To, self.num_points, pc.shape[-1])
        st = None
        if self.state_dim > 0:
            st = obs_dict["state"][:, :To].to(torch.float32).contiguous().view(B * To, self.state_dim)
        feat = self.pointnet_for(B * To, pc.device).encode(pc, st)
        return feat.view(B, To * self.obs_feature_dim)
