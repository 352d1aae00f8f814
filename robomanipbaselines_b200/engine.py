"""Python handle around rmb_act_engine_t (include/rmb_b200.h). PyTorch is used only for device memory and streams."""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import ActConfig, check, cur_stream, ptr


class ActEngine:
    """One engine per GPU. Owns the bf16 weight arena and the activation workspace (sized for max_batch)."""

    def __init__(self, *, max_batch, num_cams, img_h, img_w, state_dim, action_dim, hidden_dim=512,
                 dim_feedforward=3200, nheads=8, enc_layers=4, dec_layers=7, chunk_size=100, device=0,
                 multi_task=False, task_emb_dim=384, task_token=False):
        if not torch.cuda.is_available():
            raise _lib.RmbError("ActEngine needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.cfg = ActConfig(max_batch, num_cams, img_h, img_w, state_dim, action_dim, hidden_dim, dim_feedforward,
                             nheads, enc_layers, dec_layers, chunk_size, device, int(bool(multi_task)),
                             int(task_emb_dim) if multi_task else 0, int(bool(task_token)))
        self.device = torch.device("cuda", device)
        self._h = ctypes.c_void_p()
        check(self.lib.rmb_act_engine_create(ctypes.byref(self.cfg), ctypes.byref(self._h)), "rmb_act_engine_create")
        self.finalized = False

    # ---- weights: keys are the reference checkpoint's (RolloutBase.load_ckpt, common/base/RolloutBase.py:376-385)
    def load_state_dict(self, state_dict):
        for name, t in state_dict.items():
            a = np.ascontiguousarray(t.detach().to("cpu", torch.float32).numpy())
            check(self.lib.rmb_act_load_tensor(self._h, name.encode(), a.ctypes.data_as(ctypes.c_void_p), a.size),
                  f"rmb_act_load_tensor({name})")
        check(self.lib.rmb_act_finalize(self._h), "rmb_act_finalize")
        self.finalized = True

    @property
    def max_batch(self):
        return self.cfg.max_batch

    @property
    def num_launches(self):
        return self.lib.rmb_act_forward_num_launches(self._h)

    def flops(self, batch):
        return self.lib.rmb_act_forward_flops(self._h, batch)

    def forward(self, images, qpos, out=None):
        """images: uint8 (B, Ncam, H, W, 3) or float32 (B, Ncam, 3, H, W) in [0,1]; qpos float32 (B, Ds), on device.
        Returns a_hat float32 (B, chunk, Da), normalised action space."""
        B = qpos.shape[0]
        c = self.cfg
        if images.dtype == torch.uint8:
            fmt, want = 0, (B, c.num_cams, c.img_h, c.img_w, 3)
        elif images.dtype == torch.float32:
            fmt, want = 1, (B, c.num_cams, 3, c.img_h, c.img_w)
        else:
            raise TypeError("images must be uint8 HWC or float32 CHW")
        if tuple(images.shape) != want:
            raise ValueError(f"images shape {tuple(images.shape)} != {want}")
        if qpos.dtype != torch.float32 or tuple(qpos.shape) != (B, c.state_dim):
            raise ValueError("qpos must be float32 (B, state_dim)")
        if not (images.is_cuda and qpos.is_cuda and images.is_contiguous() and qpos.is_contiguous()):
            raise ValueError("inputs must be contiguous CUDA tensors")
        if out is None:
            out = torch.empty(B, c.chunk_size, c.action_dim, device=self.device, dtype=torch.float32)
        check(self.lib.rmb_act_forward(self._h, ptr(images), fmt, ptr(qpos), ptr(out), B, cur_stream()),
              "rmb_act_forward")
        return out

    def forward_profile(self, images, qpos, out=None):
        """forward() with CUDA events around every launch; returns (a_hat, [dict(name, kernel, ms, flops, bytes)])."""
        B = qpos.shape[0]
        c = self.cfg
        fmt = 0 if images.dtype == torch.uint8 else 1
        if out is None:
            out = torch.empty(B, c.chunk_size, c.action_dim, device=self.device, dtype=torch.float32)
        cap = 256
        recs = (_lib.LaunchProf * cap)()
        n = check(self.lib.rmb_act_forward_profile(self._h, ptr(images), fmt, ptr(qpos), ptr(out), B, cur_stream(),
                                                   recs, cap), "rmb_act_forward_profile")
        return out, [dict(name=recs[i].name.decode(), kernel=recs[i].kernel.decode(), ms=recs[i].ms,
                          flops=recs[i].flops, bytes=recs[i].bytes) for i in range(n)]

    @property
    def seq_len(self):
        return 2 + self.cfg.task_token + (self.cfg.img_h // 32) * (self.cfg.img_w // 32) * self.cfg.num_cams

    def set_task_embedding(self, task_emb):
        """MT-ACT: float32 (B, task_emb_dim) sentence embeddings, one per episode; the projected token and the FiLM
        gamma / beta stay resident in the engine until the next call (RolloutMtAct.get_task_emb caches likewise)."""
        if task_emb.dtype != torch.float32 or task_emb.dim() != 2 or task_emb.shape[1] != self.cfg.task_emb_dim:
            raise ValueError(f"task_emb must be float32 (B, {self.cfg.task_emb_dim})")
        if not task_emb.is_cuda:
            raise ValueError("task_emb must be a CUDA tensor")
        task_emb = task_emb.contiguous()
        check(self.lib.rmb_act_set_task_embedding(self._h, ptr(task_emb), task_emb.shape[0], cur_stream()),
              "rmb_act_set_task_embedding")

    def attention_debug(self, on=True, episode=0):
        """Opt-in: keep the head-averaged encoder self-attention maps of one episode (RolloutAct.draw_plot)."""
        check(self.lib.rmb_act_set_attention_debug(self._h, int(bool(on)), int(episode)), "rmb_act_set_attention_debug")

    def attention_maps(self):
        """float32 (enc_layers, L, L) of the last forward (attention_debug must be on)."""
        L = self.seq_len
        out = torch.empty(self.cfg.enc_layers, L, L, device=self.device)
        check(self.lib.rmb_act_read_attention_maps(self._h, ptr(out), cur_stream()), "rmb_act_read_attention_maps")
        return out

    def check_saturation(self, on=True):
        """Diagnostic mode: scan every 16-bit tensor a forward stores for values at the fp16 clamp (+-65504; bf16
        build: inf / nan). Read the count with saturation_count(). Off by default (costs extra launches)."""
        check(self.lib.rmb_act_set_check_saturation(self._h, int(bool(on))), "rmb_act_set_check_saturation")

    def saturation_count(self, clear=True):
        return check(self.lib.rmb_saturation_count(int(bool(clear))), "rmb_saturation_count")

    def debug_read(self, which, batch):
        L = self.seq_len
        out = torch.empty(batch, L, 512, device=self.device)
        check(self.lib.rmb_act_debug_read(self._h, which, ptr(out), batch, cur_stream()), "rmb_act_debug_read")
        return out

    def close(self, check_fault=True):
        """Destroy the engine. A kernel fault recorded while it was alive (results were garbage) is raised here at the
        latest, so that it cannot go unnoticed."""
        if getattr(self, "_h", None) is not None and self._h.value:
            fault = self.lib.rmb_device_fault(1) if check_fault else 0
            self.lib.rmb_act_engine_destroy(self._h)
            self._h = ctypes.c_void_p()
            if fault > 0:
                raise _lib.RmbError("ActEngine.close: a kernel recorded a device fault (mbarrier timeout) while this "
                                    "engine was in use; its results are invalid")

    def __del__(self):
        try:
            self.close(check_fault=False)
        except Exception:
            pass


def _load_tensors(lib, handle, fn, state_dict, what):
    for name, t in state_dict.items():
        a = np.ascontiguousarray(t.detach().to("cpu", torch.float32).numpy())
        check(fn(handle, name.encode(), a.ctypes.data_as(ctypes.c_void_p), a.size), f"{what}({name})")


class UnetEngine:
    """Handle around rmb_unet_engine_t: ConditionalUnet1D + the scheduler loop (include/rmb_b200.h)."""

    def __init__(self, *, max_batch, horizon, input_dim, global_cond_dim, diffusion_step_embed_dim=128,
                 down_dims=(512, 1024, 2048), kernel_size=5, n_groups=8, num_train_timesteps=100, device=0):
        if not torch.cuda.is_available():
            raise _lib.RmbError("UnetEngine needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        dd = (ctypes.c_int32 * 4)(*(list(down_dims) + [0] * (4 - len(down_dims))))
        self.cfg = _lib.UnetConfig(max_batch, horizon, input_dim, global_cond_dim, diffusion_step_embed_dim,
                                   len(down_dims), dd, kernel_size, n_groups, num_train_timesteps, device)
        self.device = torch.device("cuda", device)
        self._h = ctypes.c_void_p()
        check(self.lib.rmb_unet_engine_create(ctypes.byref(self.cfg), ctypes.byref(self._h)), "rmb_unet_engine_create")

    def load_state_dict(self, state_dict):
        """state_dict of the ConditionalUnet1D module (keys relative to it)."""
        _load_tensors(self.lib, self._h, self.lib.rmb_unet_load_tensor, state_dict, "rmb_unet_load_tensor")
        check(self.lib.rmb_unet_finalize(self._h), "rmb_unet_finalize")

    @property
    def max_batch(self):
        return self.cfg.max_batch

    @property
    def launches_per_step(self):
        return self.lib.rmb_unet_launches_per_step(self._h)

    def step_flops(self, batch):
        return self.lib.rmb_unet_step_flops(self._h, batch)

    def _check(self, x, global_cond):
        B = x.shape[0]
        c = self.cfg
        if tuple(x.shape) != (B, c.horizon, c.input_dim) or x.dtype != torch.float32:
            raise ValueError(f"trajectory must be float32 (B, {c.horizon}, {c.input_dim})")
        if tuple(global_cond.shape) != (B, c.global_cond_dim) or global_cond.dtype != torch.float32:
            raise ValueError(f"global_cond must be float32 (B, {c.global_cond_dim})")
        if not (x.is_cuda and global_cond.is_cuda and x.is_contiguous() and global_cond.is_contiguous()):
            raise ValueError("inputs must be contiguous CUDA tensors")
        return B

    def forward(self, sample, timestep, global_cond, out=None):
        """ConditionalUnet1D.forward(sample (B,T,Da), timestep int, global_cond (B,G)) -> (B,T,Da)."""
        B = self._check(sample, global_cond)
        if out is None:
            out = torch.empty_like(sample)
        check(self.lib.rmb_unet1d_forward(self._h, ptr(sample), int(timestep), ptr(global_cond), ptr(out), B,
                                          cur_stream()), "rmb_unet1d_forward")
        return out

    def forward_profile(self, sample, timestep, global_cond):
        B = self._check(sample, global_cond)
        out = torch.empty_like(sample)
        cap = 256
        recs = (_lib.LaunchProf * cap)()
        n = check(self.lib.rmb_unet1d_forward_profile(self._h, ptr(sample), int(timestep), ptr(global_cond), ptr(out),
                                                      B, cur_stream(), recs, cap), "rmb_unet1d_forward_profile")
        return out, [dict(name=recs[i].name.decode(), kernel=recs[i].kernel.decode(), ms=recs[i].ms,
                          flops=recs[i].flops, bytes=recs[i].bytes) for i in range(n)]

    def denoise(self, global_cond, x_init, steps, noise=None, out=None):
        """The conditional_sample loop. steps: ctypes array of _lib.SchedStep (see common/schedulers.py);
        noise float32 (n_steps, B, T, Da) or None."""
        B = self._check(x_init, global_cond)
        n = len(steps)
        if noise is not None and (tuple(noise.shape) != (n,) + tuple(x_init.shape) or noise.dtype != torch.float32
                                  or not noise.is_cuda or not noise.is_contiguous()):
            raise ValueError("noise must be a contiguous float32 CUDA tensor (n_steps, B, T, Da)")
        if out is None:
            out = torch.empty_like(x_init)
        check(self.lib.rmb_unet1d_denoise(self._h, ptr(global_cond), ptr(x_init), ptr(noise), steps, n, ptr(out), B,
                                          cur_stream()), "rmb_unet1d_denoise")
        return out

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.rmb_unet_engine_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PointNetEngine:
    """Handle around rmb_pointnet_engine_t: DP3Encoder (PointNetEncoderXYZ + state MLP)."""

    def __init__(self, *, max_clouds, num_points, in_channels=3, out_channels=64, state_dim=0, device=0):
        if not torch.cuda.is_available():
            raise _lib.RmbError("PointNetEngine needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.cfg = _lib.PointNetConfig(max_clouds, num_points, in_channels, out_channels, state_dim, device)
        self.device = torch.device("cuda", device)
        self._h = ctypes.c_void_p()
        check(self.lib.rmb_pointnet_engine_create(ctypes.byref(self.cfg), ctypes.byref(self._h)),
              "rmb_pointnet_engine_create")

    def load_state_dict(self, state_dict):
        _load_tensors(self.lib, self._h, self.lib.rmb_pointnet_load_tensor, state_dict, "rmb_pointnet_load_tensor")
        check(self.lib.rmb_pointnet_finalize(self._h), "rmb_pointnet_finalize")

    @property
    def max_clouds(self):
        return self.cfg.max_clouds

    @property
    def out_dim(self):
        return self.cfg.out_channels + (64 if self.cfg.state_dim > 0 else 0)

    def encode(self, points, state=None, out=None):
        """points float32 (n, Np, >= in_channels), state float32 (n, Ds) -> (n, out_channels [+ 64])."""
        n, c = points.shape[0], self.cfg
        if points.dtype != torch.float32 or points.dim() != 3 or points.shape[1] != c.num_points or not points.is_cuda \
                or not points.is_contiguous():
            raise ValueError(f"points must be a contiguous float32 CUDA tensor (n, {c.num_points}, >= {c.in_channels})")
        if c.state_dim > 0 and (state is None or tuple(state.shape) != (n, c.state_dim) or state.dtype != torch.float32
                                or not state.is_contiguous()):
            raise ValueError(f"state must be contiguous float32 (n, {c.state_dim})")
        if out is None:
            out = torch.empty(n, self.out_dim, device=self.device, dtype=torch.float32)
        check(self.lib.rmb_pointnet_encode(self._h, ptr(points), points.shape[2], ptr(state), ptr(out), n,
                                           cur_stream()), "rmb_pointnet_encode")
        return out

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.rmb_pointnet_engine_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ImgEncEngine:
    """Handle around rmb_imgenc_engine_t: Diffusion Policy's per-camera image encoder (ResNet-18 + GroupNorm +
    SpatialSoftmax + Linear) with the centre crop fused into the input pass."""

    def __init__(self, *, max_images, in_h, in_w, crop_h, crop_w, num_kp=32, feat_dim=64, device=0):
        if not torch.cuda.is_available():
            raise _lib.RmbError("ImgEncEngine needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.cfg = _lib.ImgEncConfig(max_images, in_h, in_w, crop_h, crop_w, num_kp, feat_dim, device)
        self.device = torch.device("cuda", device)
        self._h = ctypes.c_void_p()
        check(self.lib.rmb_imgenc_engine_create(ctypes.byref(self.cfg), ctypes.byref(self._h)), "rmb_imgenc_engine_create")

    def load_state_dict(self, state_dict):
        """keys relative to obs_encoder.obs_nets.<rgb key> (the duplicated `nets.0/1.*` aliases are ignored)."""
        sd = {k: v for k, v in state_dict.items() if not (k.startswith("nets.0.") or k.startswith("nets.1."))}
        _load_tensors(self.lib, self._h, self.lib.rmb_imgenc_load_tensor, sd, "rmb_imgenc_load_tensor")
        check(self.lib.rmb_imgenc_finalize(self._h), "rmb_imgenc_finalize")

    @property
    def max_images(self):
        return self.cfg.max_images

    def flops(self, n):
        return self.lib.rmb_imgenc_flops(self._h, n)

    def encode(self, images, out=None):
        """images float32 (n, 3, in_h, in_w) -> features (n, feat_dim); `out` may be a strided view (n, feat_dim) of a
        wider row-major buffer (the global_cond matrix)."""
        n, c = images.shape[0], self.cfg
        if tuple(images.shape) != (n, 3, c.in_h, c.in_w) or images.dtype != torch.float32 or not images.is_cuda \
                or not images.is_contiguous():
            raise ValueError(f"images must be a contiguous float32 CUDA tensor (n, 3, {c.in_h}, {c.in_w})")
        if out is None:
            out = torch.empty(n, c.feat_dim, device=self.device, dtype=torch.float32)
        if out.dtype != torch.float32 or out.shape != (n, c.feat_dim) or out.stride(1) != 1:
            raise ValueError("out must be float32 (n, feat_dim) with unit inner stride")
        check(self.lib.rmb_imgenc_encode(self._h, ptr(images), ptr(out), out.stride(0), n, cur_stream()),
              "rmb_imgenc_encode")
        return out

    def debug_layer4(self, n):
        c = self.cfg
        hp, wp = (c.crop_h + 31) // 32 * 32, (c.crop_w + 31) // 32 * 32
        out = torch.empty(n, hp // 32, wp // 32, 512, device=self.device)
        check(self.lib.rmb_imgenc_debug_read(self._h, ptr(out), n, cur_stream()), "rmb_imgenc_debug_read")
        return out

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.rmb_imgenc_engine_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
