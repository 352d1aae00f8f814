"""Drop-in for `DiffusionUnetHybridImagePolicy` at inference time, as the reference builds and calls it:

    DiffusionUnetHybridImagePolicy(noise_scheduler=DDPMScheduler(**noise_scheduler_args), **policy["args"])
        policy/diffusion_policy/RolloutDiffusionPolicy.py:37-43 ; args from TrainDiffusionPolicy.py:100-138
    policy.predict_action({"state": (B, To, Ds), "<cam>_rgb_image": (B, To, 3, H, W) in [-1, 1]})["action"]   :66-79

The policy's source (third_party/diffusion_policy @ 7e8fea7, robomimic 0.2.0 observation encoder) is not in the
reference mount; module names follow the public upstream code ("obs_encoder.obs_nets.<key>.backbone.nets.0.weight",
"model.down_modules...") so that a reference checkpoint loads with strict=True. robomimic's VisualCore registers its
backbone and pool twice (as attributes and inside `nets`), so both spellings of those keys exist here too.
"""
import torch
from torch import nn

from ...engine import ImgEncEngine
from ..diffusion_common import DiffusionPolicyBase, register_tree

_CH = [64, 128, 256, 512]


def resnet18_gn_param_shapes(prefix):
    """torchvision resnet18 children[:-2] with GroupNorm in place of BatchNorm (weight/bias only)."""
    s = {f"{prefix}0.weight": (64, 3, 7, 7), f"{prefix}1.weight": (64,), f"{prefix}1.bias": (64,)}
    for li in range(4):
        for bi in range(2):
            p = f"{prefix}{4 + li}.{bi}."
            cin = _CH[li - 1] if (bi == 0 and li > 0) else _CH[li]
            s[p + "conv1.weight"] = (_CH[li], cin, 3, 3)
            s[p + "bn1.weight"] = (_CH[li],)
            s[p + "bn1.bias"] = (_CH[li],)
            s[p + "conv2.weight"] = (_CH[li], _CH[li], 3, 3)
            s[p + "bn2.weight"] = (_CH[li],)
            s[p + "bn2.bias"] = (_CH[li],)
            if bi == 0 and li > 0:
                s[p + "downsample.0.weight"] = (_CH[li], cin, 1, 1)
                s[p + "downsample.1.weight"] = (_CH[li],)
                s[p + "downsample.1.bias"] = (_CH[li],)
    return s


def visual_core_param_shapes(num_kp=32, feat_dim=64):
    s = {}
    for bb, pool in (("backbone.nets.", "pool.nets."), ("nets.0.nets.", "nets.1.nets.")):
        s.update(resnet18_gn_param_shapes(bb))
        s[pool + "weight"] = (num_kp, 512, 1, 1)
        s[pool + "bias"] = (num_kp,)
    s["nets.3.weight"] = (feat_dim, 2 * num_kp)
    s["nets.3.bias"] = (feat_dim,)
    return s


class DiffusionUnetHybridImagePolicy(DiffusionPolicyBase):
    def __init__(self, shape_meta, noise_scheduler, horizon, n_action_steps, n_obs_steps, num_inference_steps=None,
                 obs_as_global_cond=True, crop_shape=(76, 76), diffusion_step_embed_dim=256,
                 down_dims=(256, 512, 1024), kernel_size=5, n_groups=8, cond_predict_scale=True,
                 obs_encoder_group_norm=False, eval_fixed_crop=False, max_batch=1, **kwargs):
        if not obs_as_global_cond or not cond_predict_scale:
            raise NotImplementedError("RoboManipBaselines uses obs_as_global_cond=True, cond_predict_scale=True")
        if not obs_encoder_group_norm or not eval_fixed_crop:
            raise NotImplementedError("RoboManipBaselines uses obs_encoder_group_norm=True, eval_fixed_crop=True "
                                      "(TrainDiffusionPolicy.py:124-125)")
        action_dim = int(shape_meta["action"]["shape"][0])
        self_obs = [(k, dict(v)) for k, v in shape_meta["obs"].items()]
        rgb = [(k, tuple(int(x) for x in v["shape"])) for k, v in self_obs if v.get("type", "low_dim") == "rgb"]
        low = [(k, int(v["shape"][0])) for k, v in self_obs if v.get("type", "low_dim") == "low_dim"]
        feat = 64 * len(rgb) + sum(d for _, d in low)
        super().__init__(noise_scheduler, action_dim, feat, horizon, n_action_steps, n_obs_steps, num_inference_steps,
                         diffusion_step_embed_dim, down_dims, kernel_size, n_groups, max_batch)
        self.obs_order = [k for k, _ in self_obs]                 # robomimic concatenates in shape_meta order
        self.rgb_keys, self.low_dims = dict(rgb), dict(low)
        self.crop_shape = None if crop_shape is None else (int(crop_shape[0]), int(crop_shape[1]))
        self.obs_encoder = nn.Module()
        self.obs_encoder.add_module("obs_nets", nn.Module())
        for k, _ in rgb:
            self.obs_encoder.obs_nets.add_module(k, register_tree(nn.Module(), visual_core_param_shapes()))
        self._enc = {}

    def _invalidate(self):
        super()._invalidate()
        for e in getattr(self, "_enc", {}).values():
            e.close()
        self._enc = {}

    def encoder_for(self, key, n_images, device):
        dev = device.index if device.index is not None else torch.cuda.current_device()
        e = self._enc.get(key)
        if e is None or e.max_images < n_images or e.device.index != dev:
            if e is not None:
                e.close()
            _, h, w = self.rgb_keys[key]
            ch, cw = self.crop_shape if self.crop_shape is not None else (h, w)
            e = ImgEncEngine(max_images=max(n_images, self.max_batch * self.n_obs_steps), in_h=h, in_w=w, crop_h=ch,
                             crop_w=cw, device=dev)
            e.load_state_dict(getattr(self.obs_encoder.obs_nets, key).state_dict())
            self._enc[key] = e
        return e

    def encode_obs(self, obs_dict):
        """-> global_cond (B, To * feat): per obs step the features of every obs key in shape_meta order."""
        To = self.n_obs_steps
        first = obs_dict[self.obs_order[0]]
        if not first.is_cuda:
            raise RuntimeError("DiffusionUnetHybridImagePolicy runs on the B200 engine only: move inputs to cuda")
        B = first.shape[0]
        cond = torch.empty(B * To, self.obs_feature_dim, device=first.device, dtype=torch.float32)
        col = 0
        for k in self.obs_order:
            x = obs_dict[k][:, :To]
            if k in self.rgb_keys:
                c, h, w = self.rgb_keys[k]
                imgs = x.to(torch.float32).contiguous().view(B * To, c, h, w)
                self.encoder_for(k, B * To, imgs.device).encode(imgs, out=cond[:, col:col + 64])
                col += 64
            else:
                d = self.low_dims[k]
                cond[:, col:col + d] = x.to(torch.float32).reshape(B * To, d)
                col += d
        return cond.view(B, To * self.obs_feature_dim)
