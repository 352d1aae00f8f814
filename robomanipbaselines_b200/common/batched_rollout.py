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


def _check_device_tensor(t, dtype, shape, device, name):
    """The C ABI takes raw device pointers: reject anything whose memory is not what the kernel will read."""
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (no CPU fallback)")
    if t.device.index is not None and device.index is not None and t.device.index != device.index:
        raise ValueError(f"{name} is on {t.device}, the rollout runs on {device}")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError(f"{name} must have shape {tuple(shape)}, got {tuple(t.shape)}")
    return t if t.is_contiguous() else t.contiguous()


def raise_on_device_fault(lib=None, what=""):
    """Synchronise and turn the kernels' sticky fault flag (an mbarrier wait that timed out: results are garbage) into
    an exception; the flag is cleared so that the next run starts clean. Call wherever results leave the device."""
    lib = lib or _lib.load()
    v = lib.rmb_device_fault(1)
    if v != 0:
        msg = lib.rmb_last_error()
        raise _lib.RmbError(f"{what}: device fault {v}: {msg.decode() if msg else '?'}")


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
        # --no_temp_ensem: per-episode chunk FIFO (the reference's policy_action_buf, RolloutAct.py:70,86-87) as float64
        # (B, C, Da) + read position; positions are a function of the reset history, mirrored exactly on the host
        if no_temp_ensem:
            self.fifo = torch.zeros(batch, self.C, self.Da, device=dev, dtype=torch.float64)
            self.fifo_pos = torch.full((batch,), self.C, device=dev, dtype=torch.int32)
            self._host_pos = np.full(batch, self.C, dtype=np.int64)
            self._sub_images = None
            self._sub_qpos = torch.empty(batch, self.Ds, device=dev, dtype=torch.float32)
            self._all_rows = torch.arange(batch, device=dev, dtype=torch.int32)
        self.n_inferences = 0
        self._task_emb = None    # MT-ACT: float32 (B or 1, 384) sentence embeddings (set_task_embedding)
        self.use_cuda_graph = (batch <= 8) if use_cuda_graph is None else bool(use_cuda_graph)
        self._graph = None   # (graph, static images, static raw state, image shape, engine)

    def reset(self, mask=None):
        """RolloutAct.reset_variables (:62-66) for the episodes in `mask` (bool (B,), default all)."""
        if self.no_temp_ensem:
            m = np.ones(self.B, dtype=bool) if mask is None else np.asarray(torch.as_tensor(mask).cpu().numpy(), dtype=bool)
            if m.shape != (self.B,):
                raise ValueError(f"reset mask must have shape ({self.B},)")
            self._host_pos[m] = self.C
            self.fifo_pos.masked_fill_(torch.from_numpy(m).to(self.device), self.C)
            return
        if mask is None:
            self.reset_mask.fill_(1)
        else:
            self.reset_mask |= torch.as_tensor(mask, device=self.device).to(torch.uint8)

    def set_task_embedding(self, task_emb):
        """MT-ACT (policy/mt_act/RolloutMtAct.py:84-85): the `task_emb=` every following policy call gets, one row per
        episode (or one row for all). The engine keeps the projected token / FiLM parameters resident."""
        if not hasattr(self.policy, "set_task_embedding"):
            raise ValueError("this policy takes no task embedding")
        self._task_emb = task_emb.to(self.device, torch.float32)

    def _engine(self, img_h, img_w, rows=None):
        """The policy's engine for this batch; MT-ACT: with the task embedding of the episodes in `rows` applied."""
        if self._task_emb is None:
            if hasattr(self.policy, "set_task_embedding"):
                raise ValueError("MT-ACT rollout without a task embedding: call set_task_embedding first")
            return self.policy.engine_for(self.B, img_h, img_w, self.device)
        if rows is None:
            return self.policy.set_task_embedding(self._task_emb, self.B, img_h, img_w, self.device)
        te = self._task_emb if self._task_emb.shape[0] == 1 else self._task_emb.index_select(0, rows.long()).contiguous()
        return self.policy.set_task_embedding(te, int(rows.numel()), img_h, img_w, self.device)

    def normalize_state(self, raw_state):
        """raw_state float64 (B, Ds) on device -> self.qpos float32."""
        raw_state = _check_device_tensor(raw_state, torch.float64, (self.B, self.Ds), self.device, "raw_state")
        check(self.lib.rmb_normalize(ptr(raw_state), self.s_type, ptr(self.s_p0), ptr(self.s_p1), self.s_out_min,
                                     ptr(self.qpos), self.B, self.Ds, cur_stream()), "rmb_normalize")
        return self.qpos

    def infer(self, images_u8, raw_state):
        """images_u8 uint8 (B, Ncam, H, W, 3), raw_state float64 (B, Ds), both on device.
        Returns the denormalised float64 actions (B, Da) (device tensor, valid until the next call)."""
        if images_u8.dim() != 5 or images_u8.shape[0] != self.B:
            raise ValueError(f"images must be (B={self.B}, Ncam, H, W, 3) uint8 or (B, Ncam, 3, H, W) float32")
        images_u8 = _check_device_tensor(images_u8, images_u8.dtype, None, self.device, "images")
        raw_state = _check_device_tensor(raw_state, torch.float64, (self.B, self.Ds), self.device, "raw_state")
        if self.no_temp_ensem:
            return self._infer_fifo(images_u8, raw_state)
        if self.use_cuda_graph:
            return self._infer_graph(images_u8, raw_state)
        return self._infer_eager(images_u8, raw_state)

    def _infer_graph(self, images_u8, raw_state):
        key = (tuple(images_u8.shape), images_u8.dtype)
        img_h, img_w = (images_u8.shape[2], images_u8.shape[3]) if images_u8.dtype == torch.uint8 else (images_u8.shape[3], images_u8.shape[4])
        eng = self._engine(img_h, img_w)
        # the captured graph pins the engine's workspace: recapture when the policy rebuilt its engine (another rollout
        # asked for a larger batch / other image size) instead of replaying into freed memory
        if self._graph is None or self._graph[3] != key or self._graph[4] is not eng:
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
            self._graph = (g, s_img, s_state, key, eng)
        g, s_img, s_state, _, _ = self._graph
        s_img.copy_(images_u8); s_state.copy_(raw_state)
        g.replay()
        return self.action

    def _infer_eager(self, images_u8, raw_state, clear_reset=True):
        self.normalize_state(raw_state)
        img_h, img_w = (images_u8.shape[2], images_u8.shape[3]) if images_u8.dtype == torch.uint8 else (images_u8.shape[3], images_u8.shape[4])
        eng = self._engine(img_h, img_w)
        eng.forward(images_u8, self.qpos, out=self.chunks)
        self.n_inferences += 1
        check(self.lib.rmb_ensemble_step(ptr(self.ring), ptr(self.chunks), ptr(self.step), ptr(self.reset_mask),
                                         ptr(self.wtab), self.a_type, ptr(self.a_p0), ptr(self.a_p1), self.a_out_min,
                                         ptr(self.action), None, ptr(self.action_norm), self.B, self.C, self.Da,
                                         cur_stream()), "rmb_ensemble_step")
        if clear_reset:
            self.reset_mask.zero_()
        return self.action

    def _infer_fifo(self, images_u8, raw_state):
        """--no_temp_ensem (RolloutAct.py:70,86-87), per episode: the policy runs only for the episodes whose chunk is
        used up (compacted into a dense sub-batch), every episode pops one action per call."""
        need = self._host_pos >= self.C
        sub_rows, act_new, stride = None, None, 0
        if need.any():
            n = int(need.sum())
            self.normalize_state(raw_state)
            img_h, img_w = (images_u8.shape[2], images_u8.shape[3]) if images_u8.dtype == torch.uint8 else (images_u8.shape[3], images_u8.shape[4])
            if n == self.B:
                self._engine(img_h, img_w).forward(images_u8, self.qpos, out=self.chunks)
                sub_rows = self._all_rows
            else:
                idx = np.nonzero(need)[0].astype(np.int32)
                idx_dev = torch.from_numpy(idx).to(self.device)
                eng = self._engine(img_h, img_w, rows=idx_dev)
                if self._sub_images is None or self._sub_images.shape != images_u8.shape or self._sub_images.dtype != images_u8.dtype:
                    self._sub_images = torch.empty_like(images_u8)
                row_bytes = images_u8[0].numel() * images_u8.element_size()
                check(self.lib.rmb_gather_rows(ptr(images_u8), ptr(idx_dev), ptr(self._sub_images), n, row_bytes, cur_stream()), "rmb_gather_rows")
                check(self.lib.rmb_gather_rows(ptr(self.qpos), ptr(idx_dev), ptr(self._sub_qpos), n, self.Ds * 4, cur_stream()), "rmb_gather_rows")
                eng.forward(self._sub_images[:n], self._sub_qpos[:n], out=self.chunks[:n])
                rows = np.full(self.B, -1, dtype=np.int32)
                rows[idx] = np.arange(n, dtype=np.int32)
                sub_rows = torch.from_numpy(rows).to(self.device)
            act_new, stride = self.chunks, self.C * self.Da
            self.n_inferences += 1
            self._host_pos[need] = 0
        check(self.lib.rmb_action_fifo_step(ptr(self.fifo), ptr(self.fifo_pos), ptr(act_new), stride, ptr(sub_rows),
                                            self.a_type, ptr(self.a_p0), ptr(self.a_p1), self.a_out_min, ptr(self.action),
                                            ptr(self.action_norm), self.B, self.C, self.Da, cur_stream()),
              "rmb_action_fifo_step")
        self._host_pos += 1
        return self.action

    def check_fault(self):
        """Raise RmbError if a kernel of this process recorded a fault since the last check (synchronises)."""
        raise_on_device_fault(self.lib, "BatchedRollout")
