"""BatchedDiffusionRollout: RolloutDiffusionPolicy.infer_policy / RolloutDiffusionPolicy3d.infer_policy
(policy/diffusion_policy/RolloutDiffusionPolicy.py:66-143, policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:83-171)
for B independent episodes at once.

Per-episode state on the device: observation ring buffers (B, To, ...) that advance ONLY when that episode calls the
policy (the reference pushes inside the `if len(policy_action_buf) == 0` branch), the float64 action FIFO
(B, n_action_steps, Da) and its read position. Episodes reset independently; an inference is run for the whole batch
whenever at least one episode's FIFO is empty and is consumed only by those episodes.
"""
import torch

from .. import _lib
from .._lib import check, cur_stream, ptr
from .batched_rollout import _norm_params


class BatchedDiffusionRollout:
    def __init__(self, policy, batch, model_meta_info, device="cuda:0", obs_keys=("point_cloud",)):
        """policy: DP3 / DiffusionUnetHybridImagePolicy mirror with weights loaded. obs_keys: the non-state entries of
        the obs dict (e.g. ("point_cloud",) or ("front_rgb_image",)), already in the policy's input format per step."""
        self.lib = _lib.load()
        self.policy, self.B, self.device, self.meta = policy, batch, torch.device(device), model_meta_info
        self.To = int(model_meta_info["data"]["n_obs_steps"])
        self.Na = int(model_meta_info["data"]["n_action_steps"])
        self.Da = len(model_meta_info["action"]["example"])
        self.Ds = len(model_meta_info["state"]["example"])
        self.obs_keys = tuple(obs_keys)
        dev = self.device
        t, p0, p1, om = _norm_params(model_meta_info["state"], "normalize")
        self.s_type, self.s_p0, self.s_p1, self.s_out_min = t, torch.from_numpy(p0).to(dev), torch.from_numpy(p1).to(dev), om
        t, p0, p1, om = _norm_params(model_meta_info["action"], "denormalize")
        self.a_type, self.a_p0, self.a_p1, self.a_out_min = t, torch.from_numpy(p0).to(dev), torch.from_numpy(p1).to(dev), om
        self.state_now = torch.empty(batch, self.Ds, device=dev, dtype=torch.float32)
        self.obs_buf = {}                                     # key -> (B, To, ...) float32
        self.has_obs = torch.zeros(batch, device=dev, dtype=torch.bool)
        self.fifo = torch.zeros(batch, self.Na, self.Da, device=dev, dtype=torch.float64)
        self.fifo_pos = torch.full((batch,), self.Na, device=dev, dtype=torch.int64)    # Na == empty
        self.popped = torch.empty(batch, self.Da, device=dev, dtype=torch.float64)
        self.action = torch.zeros(batch, self.Da, device=dev, dtype=torch.float64)
        self.n_inferences = 0
        self._host_pos = [self.Na] * batch                     # mirror of fifo_pos: no device sync to decide

    def reset(self, mask=None):
        """reset_variables (RolloutDiffusionPolicy.py:60-64) for the episodes in `mask` (bool (B,), default all)."""
        if mask is None:
            mask_l = [True] * self.B
        else:
            mask_l = [bool(m) for m in torch.as_tensor(mask).tolist()]
        m = torch.tensor(mask_l, device=self.device)
        self.has_obs &= ~m
        self.fifo_pos[m] = self.Na
        self._host_pos = [self.Na if r else p for r, p in zip(mask_l, self._host_pos)]

    def _push(self, key, new, need):
        """Ring-buffer update for the episodes in `need`: first call fills every slot, later calls shift by one."""
        if key not in self.obs_buf:
            self.obs_buf[key] = torch.zeros((self.B, self.To) + tuple(new.shape[1:]), device=self.device, dtype=torch.float32)
        buf = self.obs_buf[key]
        shp = (self.B,) + (1,) * (buf.dim() - 1)
        first = (need & ~self.has_obs).view(shp)
        later = (need & self.has_obs).view(shp)
        shifted = torch.cat([buf[:, 1:], new.unsqueeze(1)], dim=1)
        filled = new.unsqueeze(1).expand_as(buf)
        self.obs_buf[key] = torch.where(first, filled, torch.where(later, shifted, buf))

    def infer(self, raw_state, x_init=None, step_noise=None, **obs):
        """raw_state float64 (B, Ds); obs[key] float32 (B, ...) per obs_keys, all on device. Returns the denormalised
        float64 actions (B, Da)."""
        need_l = [p >= self.Na for p in self._host_pos]
        if any(need_l):
            need = torch.tensor(need_l, device=self.device)
            check(self.lib.rmb_normalize(ptr(raw_state), self.s_type, ptr(self.s_p0), ptr(self.s_p1), self.s_out_min,
                                         ptr(self.state_now), self.B, self.Ds, cur_stream()), "rmb_normalize")
            self._push("state", self.state_now, need)
            for k in self.obs_keys:
                self._push(k, obs[k].to(torch.float32), need)
            self.has_obs |= need
            act = self.policy.predict_action(self.obs_buf, x_init=x_init, step_noise=step_noise)["action"]
            self.n_inferences += 1
            self.fifo = torch.where(need.view(-1, 1, 1), act.to(torch.float64), self.fifo)
            self.fifo_pos = torch.where(need, torch.zeros_like(self.fifo_pos), self.fifo_pos)
            self._host_pos = [0 if n else p for n, p in zip(need_l, self._host_pos)]
        idx = self.fifo_pos.view(-1, 1, 1).expand(-1, 1, self.Da)
        self.popped.copy_(self.fifo.gather(1, idx).squeeze(1))
        check(self.lib.rmb_denormalize(ptr(self.popped), self.a_type, ptr(self.a_p0), ptr(self.a_p1), self.a_out_min,
                                       ptr(self.action), self.B, self.Da, cur_stream()), "rmb_denormalize")
        self.fifo_pos += 1
        self._host_pos = [p + 1 for p in self._host_pos]
        return self.action
