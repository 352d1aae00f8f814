"""BatchedDiffusionRollout: RolloutDiffusionPolicy.infer_policy / RolloutDiffusionPolicy3d.infer_policy
(policy/diffusion_policy/RolloutDiffusionPolicy.py:66-143, policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:83-171)
for B independent episodes at once.

Per-episode state on the device: observation ring buffers (B, To, ...) that advance ONLY when that episode calls the
policy (the reference pushes inside the `if len(policy_action_buf) == 0` branch), the float64 action FIFO
(B, n_action_steps, Da) and its read position. All of it is moved by three kernels of the C ABI (rmb_obs_ring_push,
rmb_gather_rows, rmb_action_fifo_step); nothing on the step path is eager PyTorch arithmetic. Episodes reset
independently; predict_action runs only for the episodes whose FIFO is empty (compacted into a dense sub-batch; the
engines are batch-invariant), and is skipped entirely on the n_action_steps - 1 steps in between.

FIFO positions are a pure function of the reset history, so the host keeps an exact numpy mirror and never reads the
device to decide what to launch.
"""
import numpy as np
import torch

from .. import _lib
from .._lib import check, cur_stream, ptr
from .batched_rollout import _check_device_tensor, _norm_params


class _PinnedUploader:
    """Small per-step host -> device control arrays (modes, sub-batch rows) through rotating pinned buffers."""

    def __init__(self, n, dtype, device, depth=4):
        self.host = [torch.empty(n, dtype=dtype).pin_memory() for _ in range(depth)]
        self.dev = [torch.empty(n, dtype=dtype, device=device) for _ in range(depth)]
        self.events = [None] * depth
        self.i = 0

    def upload(self, array):
        i = self.i
        self.i = (i + 1) % len(self.host)
        if self.events[i] is not None:
            self.events[i].synchronize()           # the copy that last used this pinned buffer (long finished)
        n = len(array)
        self.host[i][:n].copy_(torch.from_numpy(np.ascontiguousarray(array)))
        self.dev[i][:n].copy_(self.host[i][:n], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self.events[i] = ev
        return self.dev[i]


class BatchedDiffusionRollout:
    def __init__(self, policy, batch, model_meta_info, device="cuda:0", obs_keys=("point_cloud",), frontend=None):
        """policy: DP3 / DiffusionUnetHybridImagePolicy mirror with weights loaded. obs_keys: the non-state entries of
        the obs dict (e.g. ("point_cloud",) or ("front_rgb_image",)), already in the policy's input format per step.
        frontend: optional callable(**env_frames) -> {obs_key: float32 (B, ...)} that turns env-layout frames into
        policy inputs on the device (DpImageFrontend / Dp3Frontend below); used by infer_env()."""
        self.lib = _lib.load()
        self.policy, self.B, self.device, self.meta = policy, batch, torch.device(device), model_meta_info
        self.To = int(model_meta_info["data"]["n_obs_steps"])
        self.Na = int(model_meta_info["data"]["n_action_steps"])
        self.Da = len(model_meta_info["action"]["example"])
        self.Ds = len(model_meta_info["state"]["example"])
        self.obs_keys = tuple(obs_keys)
        self.frontend = frontend
        dev = self.device
        t, p0, p1, om = _norm_params(model_meta_info["state"], "normalize")
        self.s_type, self.s_p0, self.s_p1, self.s_out_min = t, torch.from_numpy(p0).to(dev), torch.from_numpy(p1).to(dev), om
        t, p0, p1, om = _norm_params(model_meta_info["action"], "denormalize")
        self.a_type, self.a_p0, self.a_p1, self.a_out_min = t, torch.from_numpy(p0).to(dev), torch.from_numpy(p1).to(dev), om
        self.state_now = torch.empty(batch, self.Ds, device=dev, dtype=torch.float32)
        self.obs_buf = {}                                     # key -> (B, To, ...) float32
        self._sub_buf = {}                                    # key -> (B, To, ...) float32 gather target
        self.fifo = torch.zeros(batch, self.Na, self.Da, device=dev, dtype=torch.float64)
        self.fifo_pos = torch.full((batch,), self.Na, device=dev, dtype=torch.int32)    # Na == empty
        self.action = torch.zeros(batch, self.Da, device=dev, dtype=torch.float64)
        self.action_norm = torch.zeros(batch, self.Da, device=dev, dtype=torch.float64)
        self.n_inferences = 0
        self.n_inferred_episodes = 0
        # exact host mirrors (FIFO positions and "has an observation" are functions of the reset history only)
        self._host_pos = np.full(batch, self.Na, dtype=np.int64)
        self._host_has = np.zeros(batch, dtype=bool)
        self._modes = _PinnedUploader(batch, torch.uint8, dev)
        self._rows = _PinnedUploader(batch, torch.int32, dev)
        self._idx = _PinnedUploader(batch, torch.int32, dev)
        self._all_rows = torch.arange(batch, device=dev, dtype=torch.int32)

    def reset(self, mask=None):
        """reset_variables (RolloutDiffusionPolicy.py:60-64) for the episodes in `mask` (bool (B,), default all)."""
        m = np.ones(self.B, dtype=bool) if mask is None else np.asarray(torch.as_tensor(mask).cpu().numpy(), dtype=bool)
        if m.shape != (self.B,):
            raise ValueError(f"reset mask must have shape ({self.B},)")
        self._host_pos[m] = self.Na
        self._host_has[m] = False
        if m.all():
            self.fifo_pos.fill_(self.Na)
        else:
            self.fifo_pos.masked_fill_(torch.from_numpy(m).to(self.device), self.Na)

    def _push(self, key, new, mode_dev):
        new = _check_device_tensor(new, torch.float32, None, self.device, key)
        if new.shape[0] != self.B:
            raise ValueError(f"{key}: leading dimension must be the batch ({self.B})")
        if key not in self.obs_buf:
            shape = (self.B, self.To) + tuple(new.shape[1:])
            self.obs_buf[key] = torch.zeros(shape, device=self.device, dtype=torch.float32)
        buf = self.obs_buf[key]
        if tuple(buf.shape[2:]) != tuple(new.shape[1:]):
            raise ValueError(f"{key}: shape {tuple(new.shape)} does not match the buffer {tuple(buf.shape)}")
        row_bytes = new[0].numel() * 4
        check(self.lib.rmb_obs_ring_push(ptr(buf), ptr(new), ptr(mode_dev), self.B, self.To, row_bytes, cur_stream()),
              "rmb_obs_ring_push")

    def _gather(self, key, idx_dev, n):
        buf = self.obs_buf[key]
        if key not in self._sub_buf:
            self._sub_buf[key] = torch.empty_like(buf)
        dst = self._sub_buf[key]
        row_bytes = buf[0].numel() * 4
        check(self.lib.rmb_gather_rows(ptr(buf), ptr(idx_dev), ptr(dst), n, row_bytes, cur_stream()), "rmb_gather_rows")
        return dst[:n]

    def infer(self, raw_state, x_init=None, step_noise=None, **obs):
        """raw_state float64 (B, Ds); obs[key] float32 (B, ...) per obs_keys, all on device. Returns the denormalised
        float64 actions (B, Da) (device tensor, valid until the next call)."""
        raw_state = _check_device_tensor(raw_state, torch.float64, (self.B, self.Ds), self.device, "raw_state")
        need = self._host_pos >= self.Na
        sub_rows_dev = None
        act_new, act_stride = None, 0
        if need.any():
            n = int(need.sum())
            mode = np.where(need, np.where(self._host_has, 2, 1), 0).astype(np.uint8)
            mode_dev = self._modes.upload(mode)
            check(self.lib.rmb_normalize(ptr(raw_state), self.s_type, ptr(self.s_p0), ptr(self.s_p1), self.s_out_min,
                                         ptr(self.state_now), self.B, self.Ds, cur_stream()), "rmb_normalize")
            self._push("state", self.state_now, mode_dev)
            for k in self.obs_keys:
                self._push(k, obs[k], mode_dev)
            self._host_has |= need
            if n == self.B:
                sub_obs, sub_rows_dev = self.obs_buf, self._all_rows
                xi, sn = x_init, step_noise
            else:
                idx = np.nonzero(need)[0].astype(np.int32)
                # sub-batch sizes are rounded up to a power of two (padded with the first episode) so that the
                # policy's per-batch-size CUDA graph is reused instead of recaptured for every distinct count
                n_pad = min(self.B, 1 << (n - 1).bit_length())
                idx_pad = np.concatenate([idx, np.full(n_pad - n, idx[0], dtype=np.int32)])
                idx_dev = self._idx.upload(idx_pad)
                sub_obs = {k: self._gather(k, idx_dev, n_pad) for k in ("state",) + self.obs_keys}
                rows = np.full(self.B, -1, dtype=np.int32)
                rows[idx] = np.arange(n, dtype=np.int32)
                sub_rows_dev = self._rows.upload(rows)
                sel = torch.from_numpy(idx_pad.astype(np.int64)).to(self.device)
                xi = None if x_init is None else x_init.index_select(0, sel)
                sn = None if step_noise is None else step_noise.index_select(1, sel)
            act_new = self.policy.predict_action(sub_obs, x_init=xi, step_noise=sn)["action"]
            if act_new.dtype != torch.float32 or act_new.stride(-1) != 1 or act_new.stride(-2) != self.Da:
                act_new = act_new.to(torch.float32).contiguous()
            act_stride = act_new.stride(0)
            self.n_inferences += 1
            self.n_inferred_episodes += n
            self._host_pos[need] = 0
        check(self.lib.rmb_action_fifo_step(ptr(self.fifo), ptr(self.fifo_pos), ptr(act_new), act_stride,
                                            ptr(sub_rows_dev), self.a_type, ptr(self.a_p0), ptr(self.a_p1),
                                            self.a_out_min, ptr(self.action), ptr(self.action_norm), self.B, self.Na,
                                            self.Da, cur_stream()), "rmb_action_fifo_step")
        self._host_pos += 1
        return self.action

    def infer_env(self, raw_state, x_init=None, step_noise=None, **env_frames):
        """Same as infer() but from env-layout frames (uint8 RGB / float32 depth as the simulator renders them): the
        frontend (cv2-exact resize + scaling, or depth -> point cloud) runs on the device first."""
        if self.frontend is None:
            raise ValueError("infer_env needs a frontend (DpImageFrontend / Dp3Frontend)")
        return self.infer(raw_state, x_init=x_init, step_noise=step_noise, **self.frontend(**env_frames))


class DpImageFrontend:
    """RolloutDiffusionPolicy.get_images (:111-125) for B episodes: per camera cv2.resize to data["image_size"], HWC ->
    CHW, uint8 -> float32 / 255, `* 2.0 - 1.0`, in one kernel per camera (rmb_resize_u8_bilinear, cv2 bit-exact)."""

    def __init__(self, camera_names, image_size):
        self.camera_names, self.image_size = list(camera_names), (int(image_size[0]), int(image_size[1]))
        self._out = {}

    def __call__(self, rgb_u8):
        """rgb_u8 uint8 (B, Ncam, H, W, 3) on device -> {"<cam>_rgb_image": float32 (B, 3, h, w)}."""
        from .image_ops import resize_u8_to_policy_input

        if rgb_u8.dim() != 5 or rgb_u8.shape[1] != len(self.camera_names):
            raise ValueError("rgb_u8 must be uint8 (B, Ncam, H, W, 3)")
        out = {}
        for i, cam in enumerate(self.camera_names):
            key = f"{cam}_rgb_image"                      # DataKey.get_rgb_image_key (common/data/DataKey.py)
            frames = rgb_u8[:, i]
            if not frames.is_contiguous():
                frames = frames.contiguous()
            self._out[key] = resize_u8_to_policy_input(frames, self.image_size, out=self._out.get(key))
            out[key] = self._out[key]
        return out


class Dp3Frontend:
    """RolloutDiffusionPolicy3d.get_pointcloud (:125-153) for B episodes (Dp3ObservationPipeline)."""

    def __init__(self, batch, model_meta_info, fovy, device="cuda:0"):
        from .pointcloud import Dp3ObservationPipeline

        self.pipe = Dp3ObservationPipeline(batch, model_meta_info, fovy, device=device)

    def __call__(self, rgb_u8, depth):
        return {"point_cloud": self.pipe(rgb_u8, depth)}
