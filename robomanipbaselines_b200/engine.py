"""Python handle around rmb_act_engine_t (include/rmb_b200.h). PyTorch is used only for device memory and streams."""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import ActConfig, check, cur_stream, ptr


class ActEngine:
    """One engine per GPU. Owns the bf16 weight arena and the activation workspace (sized for max_batch)."""

    def __init__(self, *, max_batch, num_cams, img_h, img_w, state_dim, action_dim, hidden_dim=512,
                 dim_feedforward=3200, nheads=8, enc_layers=4, dec_layers=7, chunk_size=100, device=0):
        if not torch.cuda.is_available():
            raise _lib.RmbError("ActEngine needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.cfg = ActConfig(max_batch, num_cams, img_h, img_w, state_dim, action_dim, hidden_dim, dim_feedforward,
                             nheads, enc_layers, dec_layers, chunk_size, device)
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

    def debug_read(self, which, batch):
        L = 2 + (self.cfg.img_h // 32) * (self.cfg.img_w // 32) * self.cfg.num_cams
        out = torch.empty(batch, L, 512, device=self.device)
        check(self.lib.rmb_act_debug_read(self._h, which, ptr(out), batch, cur_stream()), "rmb_act_debug_read")
        return out

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.rmb_act_engine_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
