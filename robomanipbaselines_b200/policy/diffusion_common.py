"""Shared host side of the two diffusion policies (Diffusion Policy and DP3).

The reference constructs them as
    DiffusionUnetHybridImagePolicy(noise_scheduler=..., **model_meta_info["policy"]["args"])   RolloutDiffusionPolicy.py:40-43
    DP3(noise_scheduler=..., **model_meta_info["policy"]["args"])                              RolloutDiffusionPolicy3d.py:53-56
and calls `policy.predict_action(obs_dict)["action"]` (RolloutDiffusionPolicy.py:74, RolloutDiffusionPolicy3d.py:90).
The modules' source is not in the reference mount (un-vendored submodules); parameter names below follow the public
upstream code so that a reference checkpoint loads with strict=True. The arithmetic runs in the sm_100a engine.
"""
import torch
from torch import nn

from ..engine import UnetEngine


def register_tree(root, shapes, buffers=()):
    """Create nested nn.Module containers so that root.state_dict() has exactly the dotted keys of `shapes`."""
    for key, shape in shapes.items():
        mod = root
        parts = key.split(".")
        for p in parts[:-1]:
            if not hasattr(mod, p):
                mod.add_module(p, nn.Module())
            mod = getattr(mod, p)
        t = torch.zeros(tuple(shape))
        if key in buffers:
            mod.register_buffer(parts[-1], t)
        else:
            mod.register_parameter(parts[-1], nn.Parameter(t, requires_grad=False))
    return root


def unet1d_param_shapes(input_dim, global_cond_dim, dsed=128, down_dims=(512, 1024, 2048), kernel_size=5):
    """State-dict keys/shapes of upstream ConditionalUnet1D (cond_predict_scale=True, no local_cond)."""
    s = {}
    cond_dim = dsed + global_cond_dim
    k = kernel_size

    def crb(prefix, cin, cout):
        for i, ci in ((0, cin), (1, cout)):
            s[f"{prefix}.blocks.{i}.block.0.weight"] = (cout, ci, k)
            s[f"{prefix}.blocks.{i}.block.0.bias"] = (cout,)
            s[f"{prefix}.blocks.{i}.block.1.weight"] = (cout,)
            s[f"{prefix}.blocks.{i}.block.1.bias"] = (cout,)
        s[f"{prefix}.cond_encoder.1.weight"] = (2 * cout, cond_dim)
        s[f"{prefix}.cond_encoder.1.bias"] = (2 * cout,)
        if cin != cout:
            s[f"{prefix}.residual_conv.weight"] = (cout, cin, 1)
            s[f"{prefix}.residual_conv.bias"] = (cout,)

    s["diffusion_step_encoder.1.weight"] = (4 * dsed, dsed)
    s["diffusion_step_encoder.1.bias"] = (4 * dsed,)
    s["diffusion_step_encoder.3.weight"] = (dsed, 4 * dsed)
    s["diffusion_step_encoder.3.bias"] = (dsed,)
    all_dims = [input_dim] + list(down_dims)
    in_out = list(zip(all_dims[:-1], all_dims[1:]))
    mid = all_dims[-1]
    crb("mid_modules.0", mid, mid)
    crb("mid_modules.1", mid, mid)
    for i, (di, do) in enumerate(in_out):
        crb(f"down_modules.{i}.0", di, do)
        crb(f"down_modules.{i}.1", do, do)
        if i < len(in_out) - 1:
            s[f"down_modules.{i}.2.conv.weight"] = (do, do, 3)
            s[f"down_modules.{i}.2.conv.bias"] = (do,)
    for j, (di, do) in enumerate(reversed(in_out[1:])):
        crb(f"up_modules.{j}.0", do * 2, di)
        crb(f"up_modules.{j}.1", di, di)
        s[f"up_modules.{j}.2.conv.weight"] = (di, di, 4)      # ConvTranspose1d: (in, out, k)
        s[f"up_modules.{j}.2.conv.bias"] = (di,)
    s["final_conv.0.block.0.weight"] = (down_dims[0], down_dims[0], k)
    s["final_conv.0.block.0.bias"] = (down_dims[0],)
    s["final_conv.0.block.1.weight"] = (down_dims[0],)
    s["final_conv.0.block.1.bias"] = (down_dims[0],)
    s["final_conv.1.weight"] = (input_dim, down_dims[0], 1)
    s["final_conv.1.bias"] = (input_dim,)
    return s


class DiffusionPolicyBase(nn.Module):
    """predict_action() = obs encoder -> conditional_sample (whole scheduler loop on device) -> action slice."""

    def __init__(self, noise_scheduler, action_dim, obs_feature_dim, horizon, n_action_steps, n_obs_steps,
                 num_inference_steps=None, diffusion_step_embed_dim=128, down_dims=(256, 512, 1024), kernel_size=5,
                 n_groups=8, max_batch=1):
        super().__init__()
        self.noise_scheduler = noise_scheduler
        self.action_dim = action_dim
        self.obs_feature_dim = obs_feature_dim
        self.horizon, self.n_action_steps, self.n_obs_steps = horizon, n_action_steps, n_obs_steps
        self.num_inference_steps = noise_scheduler.num_train_timesteps if num_inference_steps is None else num_inference_steps
        self.unet_args = dict(input_dim=action_dim, global_cond_dim=obs_feature_dim * n_obs_steps,
                              diffusion_step_embed_dim=diffusion_step_embed_dim, down_dims=tuple(down_dims),
                              kernel_size=kernel_size, n_groups=n_groups)
        self.model = register_tree(nn.Module(), unet1d_param_shapes(action_dim, obs_feature_dim * n_obs_steps,
                                                                    diffusion_step_embed_dim, down_dims, kernel_size))
        self.max_batch = max_batch
        self._unet = None
        self._steps = None
        # the scheduler loop is thousands of small launches: replay it as ONE CUDA graph (captured per batch size)
        self.use_cuda_graph = True
        self._graphs = {}

    # Keys the upstream classes carry but that hold no arithmetic the engine needs (public upstream layout; the sources
    # are not in the reference mount, so this list is part of the "parity unpinned" hypothesis):
    #   ModuleAttrMixin._dummy_variable on the policy and its mask generator; LinearNormalizer.params_dict.* (RMB
    #   normalises outside the policy: RolloutDiffusionPolicy.py:95,123); robomimic SpatialSoftmax's persistent buffers
    #   (temperature, pos_x, pos_y), which exist twice per camera (VisualCore registers `pool` and `nets.1`).
    INERT_KEYS = ("_dummy_variable", "mask_generator._dummy_variable")
    INERT_PREFIXES = ("normalizer.",)
    INERT_SUFFIXES = (".pool.pos_x", ".pool.pos_y", ".nets.1.pos_x", ".nets.1.pos_y")
    TEMPERATURE_SUFFIXES = (".pool.temperature", ".nets.1.temperature")

    def strip_inert_keys(self, state_dict):
        """RolloutBase.load_ckpt calls load_state_dict(strict=True) on a reference checkpoint (RolloutBase.py:376-385):
        drop the inert upstream entries first so that strictness still guards every tensor the engine uses."""
        out = {}
        for k, v in state_dict.items():
            if k in self.INERT_KEYS or k.startswith(self.INERT_PREFIXES) or k.endswith(self.INERT_SUFFIXES):
                continue
            if k.endswith(self.TEMPERATURE_SUFFIXES):
                if abs(float(torch.as_tensor(v).reshape(-1)[0]) - 1.0) > 1e-6:
                    raise NotImplementedError(f"{k} = {float(v.reshape(-1)[0])}: the engine's SpatialSoftmax assumes temperature 1")
                continue
            out[k] = v
        return out

    def load_state_dict(self, state_dict, strict=True, **kw):
        out = super().load_state_dict(self.strip_inert_keys(state_dict), strict=strict, **kw)
        self._invalidate()
        return out

    def _invalidate(self):
        self._graphs = {}
        if self._unet is not None:
            self._unet.close()
        self._unet = None

    def unet_for(self, batch, device):
        dev = device.index if device.index is not None else torch.cuda.current_device()
        if self._unet is None or self._unet.max_batch < batch or self._unet.device.index != dev:
            if self._unet is not None:
                self._unet.close()
            self.max_batch = max(self.max_batch, batch)
            a = self.unet_args
            self._unet = UnetEngine(max_batch=self.max_batch, horizon=self.horizon, input_dim=a["input_dim"],
                                    global_cond_dim=a["global_cond_dim"],
                                    diffusion_step_embed_dim=a["diffusion_step_embed_dim"], down_dims=a["down_dims"],
                                    kernel_size=a["kernel_size"], n_groups=a["n_groups"],
                                    num_train_timesteps=self.noise_scheduler.num_train_timesteps, device=dev)
            self._unet.load_state_dict(self.model.state_dict())
        return self._unet

    def step_table(self):
        if self._steps is None or len(self._steps) != self.num_inference_steps:
            self._steps = self.noise_scheduler.step_table(self.num_inference_steps)
        return self._steps

    def conditional_sample(self, global_cond, x_init=None, step_noise=None, generator=None, out=None):
        """global_cond float32 (B, G) on device. x_init (B, T, Da) and step_noise (n_steps, B, T, Da) default to
        torch.randn on the device (pass them for reproducible parity with a reference run)."""
        B, dev = global_cond.shape[0], global_cond.device
        steps = self.step_table()
        if x_init is None:
            x_init = torch.randn(B, self.horizon, self.action_dim, device=dev, dtype=torch.float32, generator=generator)
        if step_noise is None and self.noise_scheduler.kind == 0:
            step_noise = torch.randn(len(steps), B, self.horizon, self.action_dim, device=dev, dtype=torch.float32,
                                     generator=generator)
        unet = self.unet_for(B, dev)
        if not self.use_cuda_graph:
            return unet.denoise(global_cond.contiguous(), x_init.contiguous(), steps, noise=step_noise, out=out)
        key = (B, dev.index, len(steps), id(unet))
        g = self._graphs.get(key)
        if g is None:
            st = {"gc": torch.empty_like(global_cond, memory_format=torch.contiguous_format),
                  "x": torch.empty_like(x_init, memory_format=torch.contiguous_format),
                  "noise": None if step_noise is None else torch.empty_like(step_noise, memory_format=torch.contiguous_format)}
            st["out"] = torch.empty_like(st["x"])
            st["gc"].copy_(global_cond); st["x"].copy_(x_init)
            if step_noise is not None:
                st["noise"].copy_(step_noise)
            unet.denoise(st["gc"], st["x"], steps, noise=st["noise"], out=st["out"])    # warm-up: plans, func attributes
            torch.cuda.synchronize(dev)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                unet.denoise(st["gc"], st["x"], steps, noise=st["noise"], out=st["out"])
            st["graph"] = graph
            self._graphs = {key: st}                # one live graph (it pins the engine's workspace addresses)
            g = st
        g["gc"].copy_(global_cond); g["x"].copy_(x_init)
        if step_noise is not None:
            g["noise"].copy_(step_noise)
        g["graph"].replay()
        if out is not None:
            out.copy_(g["out"])
            return out
        return g["out"].clone()

    def encode_obs(self, obs_dict):
        raise NotImplementedError

    def predict_action(self, obs_dict, x_init=None, step_noise=None, generator=None):
        """obs_dict as the reference's rollout builds it -> {"action": (B, n_action_steps, Da), "action_pred": (B, T, Da)}."""
        global_cond = self.encode_obs(obs_dict)
        naction = self.conditional_sample(global_cond, x_init, step_noise, generator)
        start = self.n_obs_steps - 1
        return {"action": naction[:, start:start + self.n_action_steps], "action_pred": naction}

    def forward(self, *a, **k):
        raise NotImplementedError("training (compute_loss) is out of scope of the rollout engine")
