"""BatchedRollout: RolloutAct.infer_policy (policy/act/RolloutAct.py:68-101) for B independent episodes at once.

Per-episode state lives on the device: the chunk ring buffer (B, C, C, Da), the policy-call counter (B,), and the
normalisation statistics of model_meta_info. One `infer()` =
  normalize_data(state)            (RolloutBase.get_state, common/base/RolloutBase.py:463-477)
  -> ACT forward on uint8 images   (get_images :479-490 + ACTPolicy.__call__ + DETRVAE.forward)
  -> temporal ensembling + denormalize_data (RolloutAct.py:77-98)
all enqueued on the current CUDA stream through the C ABI; with B = 1 it reproduces the reference step for step.
"""
import numpy as np
import torch

from .. import _lib
from .._lib import check, cur_stream, ptr
from .ensemble import make_weight_table


def _norm_params(stats, direction):
    """(type, p0, p1, out_min) for rmb_normalize / rmb_denormalize, computed with the reference's numpy expressions
    (common/utils/DataUtils.py:9-40) so the device multiplies bit-identical coefficients."""
    norm_type = stats["norm_config"]["type"] if "norm_config" in stats else "gaussian"
    if norm_type == "gaussian":
        mean, std = np.asarray(stats["mean"], np.float64), np.asarray(stats["std"], np.float64)
        return (0, mean, std, 0.0) if direction == "normalize" else (0, std, mean, 0.0)
    if norm_type == "limits":
        cfg = stats["norm_config"]
        rng_, mn = np.asarray(stats["range"], np.float64), np.asarray(stats["min"], np.float64)
        if direction == "normalize":
            scale = (cfg["out_max"] - cfg["out_min"]) / rng_
            return 1, mn, scale, float(cfg["out_min"])
        scale = rng_ / (cfg["out_max"] - cfg["out_min"])
        return 1, scale, mn, float(cfg["out_min"])
    raise ValueError(f"[normalize_data] Invalid normalization type: {norm_type}")


class BatchedRollout:
    def __init__(self, policy, batch, model_meta_info, device="cuda:0", no_temp_ensem=False, use_cuda_graph=None):
        """policy: robomanipbaselines_b200.policy.act.ACTPolicy with weights loaded.
        use_cuda_graph: replay one policy step (normalise -> ACT forward -> ensembling, ~50 launches) as a single CUDA
        graph with static input buffers; default on for batch <= 8, where the step is launch-latency bound."""
        self.lib = _lib.load()
        self.policy = policy
        self.B = batch
        self.device = torch.device(device)
        self.meta = model_meta_info
        self.C = int(model_meta_info["data"]["chunk_size"])
        self.Da = len(model_meta_info["action"]["example"])
        self.Ds = len(model_meta_info["state"]["example"])
        self.no_temp_ensem = no_temp_ensem
        dev = self.device
        self.ring = torch.zeros(batch, self.C, self.C, self.Da, device=dev, dtype=torch.float32)
        self.step = torch.zeros(batch, device=dev, dtype=torch.int32)
        self.reset_mask = torch.zeros(batch, device=dev, dtype=torch.uint8)
        self.wtab = torch.from_numpy(make_weight_table(self.C)).to(dev)
        t, p0, p1, om = _norm_params(model_meta_info["state"], "normalize")
        self.s_type, self.s_p0, self.s_p1, self.s_out_min = t, torch.from_numpy(p0).to(dev), torch.from_numpy(p1).to(dev), om
        t, p0, p1, om = _norm_params(model_meta_info["action"], "denormalize")
        self.a_type, self.a_p0, self.a_p1, self.a_out_min = t, torch.from_numpy(p0).to(dev), torch.from_numpy(p1).to(dev), om
        self.qpos = torch.empty(batch, self.Ds, device=dev, dtype=torch.float32)
        self.chunks = torch.empty(batch, self.C, self.Da, device=dev, dtype=torch.float32)
        self.action = torch.zeros(batch, self.Da, device=dev, dtype=torch.float64)
        self.action_norm = torch.zeros(batch, self.Da, device=dev, dtype=torch.float64)
        self._fifo_pos = 0   # no_temp_ensem: index into the current chunk (all episodes are stepped in lock-step)
        self.use_cuda_graph = (batch <= 8) if use_cuda_graph is None else bool(use_cuda_graph)
        self._graph = None   # (graph, static images, static raw state, image shape)

    def reset(self, mask=None):
        """RolloutAct.reset_variables (:62-66) for the episodes in `mask` (bool (B,), default all)."""
        if mask is None:
            self.reset_mask.fill_(1)
        else:
            self.reset_mask |= torch.as_tensor(mask, device=self.device).to(torch.uint8)
        if self.no_temp_ensem:
            self._fifo_pos = 0

    def normalize_state(self, raw_state):
        """raw_state float64 (B, Ds) on device -> self.qpos float32."""
        check(self.lib.rmb_normalize(ptr(raw_state), self.s_type, ptr(self.s_p0), ptr(self.s_p1), self.s_out_min,
                                     ptr(self.qpos), self.B, self.Ds, cur_stream()), "rmb_normalize")
        return self.qpos

    def infer(self, images_u8, raw_state):
        """images_u8 uint8 (B, Ncam, H, W, 3), raw_state float64 (B, Ds), both on device.
        Returns the denormalised float64 actions (B, Da) (device tensor, valid until the next call)."""
        if self.no_temp_ensem:
            return self._infer_fifo(images_u8, raw_state)
        if self.use_cuda_graph:
            return self._infer_graph(images_u8, raw_state)
        return self._infer_eager(images_u8, raw_state)

    def _infer_graph(self, images_u8, raw_state):
        key = (tuple(images_u8.shape), images_u8.dtype)
        if self._graph is None or self._graph[3] != key:
            s_img, s_state = torch.empty_like(images_u8), torch.empty_like(raw_state)
            s_img.copy_(images_u8); s_state.copy_(raw_state)
            # warm-up outside the capture (plans, kernel attributes) on copies of the rollout state, so that the
            # per-episode counters and the ring buffer are exactly what the caller left them
            saved = (self.ring.clone(), self.step.clone(), self.reset_mask.clone())
            self._infer_eager(s_img, s_state, clear_reset=False)
            torch.cuda.synchronize(self.device)
            self.ring.copy_(saved[0]); self.step.copy_(saved[1]); self.reset_mask.copy_(saved[2])
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._infer_eager(s_img, s_state, clear_reset=True)
            # the capture did not execute: state is still the caller's
            self._graph = (g, s_img, s_state, key)
        g, s_img, s_state, _ = self._graph
        s_img.copy_(images_u8); s_state.copy_(raw_state)
        g.replay()
        return self.action

    def _infer_eager(self, images_u8, raw_state, clear_reset=True):
        self.normalize_state(raw_state)
        img_h, img_w = images_u8.shape[2], images_u8.shape[3]
        eng = self.policy.engine_for(self.B, img_h, img_w, self.device)
        eng.forward(images_u8, self.qpos, out=self.chunks)
        check(self.lib.rmb_ensemble_step(ptr(self.ring), ptr(self.chunks), ptr(self.step), ptr(self.reset_mask),
                                         ptr(self.wtab), self.a_type, ptr(self.a_p0), ptr(self.a_p1), self.a_out_min,
                                         ptr(self.action), None, ptr(self.action_norm), self.B, self.C, self.Da,
                                         cur_stream()), "rmb_ensemble_step")
        if clear_reset:
            self.reset_mask.zero_()
        return self.action

    def _infer_fifo(self, images_u8, raw_state):
        """--no_temp_ensem (RolloutAct.py:70,86-87): one inference per C policy steps, then pop one action per call."""
        if self._fifo_pos == 0:
            self.normalize_state(raw_state)
            eng = self.policy.engine_for(self.B, images_u8.shape[2], images_u8.shape[3], self.device)
            eng.forward(images_u8, self.qpos, out=self.chunks)
        a = self.chunks[:, self._fifo_pos].to(torch.float64).contiguous()
        check(self.lib.rmb_denormalize(ptr(a), self.a_type, ptr(self.a_p0), ptr(self.a_p1), self.a_out_min,
                                       ptr(self.action), self.B, self.Da, cur_stream()), "rmb_denormalize")
        self._fifo_pos = (self._fifo_pos + 1) % self.C
        self.reset_mask.zero_()
        return self.action
