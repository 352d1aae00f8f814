"""Host side of the noise schedulers the reference builds with diffusers == 0.11.1:

    DDPMScheduler(**model_meta_info["policy"]["noise_scheduler_args"])   policy/diffusion_policy/RolloutDiffusionPolicy.py:37-39
    DDIMScheduler(**model_meta_info["policy"]["noise_scheduler_args"])   policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:50-52

Same constructor keywords (TrainDiffusionPolicy.py:127-137, TrainDiffusionPolicy3d.py:208-217). diffusers is not
installed here, so the classes restate its schedule arithmetic (fp32 torch scalars, same op order) and hand the
per-step scalars to the device as an rmb_sched_step_t table: the step itself runs fused in unet_final_step_kernel.
"""
import math

import numpy as np
import torch

from .. import _lib


def _betas(num_train_timesteps, beta_start, beta_end, beta_schedule):
    if beta_schedule == "linear":
        return torch.linspace(beta_start, beta_end, num_train_timesteps, dtype=torch.float32)
    if beta_schedule == "scaled_linear":
        return torch.linspace(beta_start ** 0.5, beta_end ** 0.5, num_train_timesteps, dtype=torch.float32) ** 2
    if beta_schedule == "squaredcos_cap_v2":
        def alpha_bar(ts):
            return math.cos((ts + 0.008) / 1.008 * math.pi / 2) ** 2

        n = num_train_timesteps
        return torch.tensor([min(1 - alpha_bar((i + 1) / n) / alpha_bar(i / n), 0.999) for i in range(n)],
                            dtype=torch.float32)
    raise NotImplementedError(f"{beta_schedule} does is not implemented")


class _SchedulerBase:
    kind = -1

    def __init__(self, num_train_timesteps, beta_start, beta_end, beta_schedule, clip_sample, prediction_type):
        if prediction_type not in ("epsilon", "sample"):
            raise ValueError(f"prediction_type given as {prediction_type} must be one of `epsilon` or `sample`")
        self.num_train_timesteps = int(num_train_timesteps)
        self.betas = _betas(self.num_train_timesteps, beta_start, beta_end, beta_schedule)
        self.alphas = 1.0 - self.betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.one = torch.tensor(1.0)
        self.clip_sample = bool(clip_sample)
        self.prediction_type = prediction_type
        self.config = dict(num_train_timesteps=num_train_timesteps, beta_start=beta_start, beta_end=beta_end,
                           beta_schedule=beta_schedule, clip_sample=clip_sample, prediction_type=prediction_type)
        self.num_inference_steps = None
        self.timesteps = np.arange(0, self.num_train_timesteps)[::-1].copy()

    def step_table(self, num_inference_steps):
        """ctypes array of rmb_sched_step_t for one conditional_sample loop."""
        self.set_timesteps(num_inference_steps)
        rows = [self._step_scalars(int(t)) for t in self.timesteps]
        arr = (_lib.SchedStep * len(rows))()
        for i, r in enumerate(rows):
            arr[i] = _lib.SchedStep(r["t"], self.kind, int(self.prediction_type == "sample"), int(self.clip_sample),
                                    int(r.get("derive_eps", 0)), r["s_beta"], r["s_alpha"], r["c_x0"], r["c_xt"],
                                    r.get("sigma", 0.0))
        return arr


class DDPMScheduler(_SchedulerBase):
    """diffusers 0.11.1 DDPMScheduler, variance_type "fixed_small"."""

    kind = 0

    def __init__(self, num_train_timesteps=1000, beta_start=0.0001, beta_end=0.02, beta_schedule="linear",
                 trained_betas=None, variance_type="fixed_small", clip_sample=True, prediction_type="epsilon"):
        if trained_betas is not None:
            raise NotImplementedError("trained_betas is not used by RoboManipBaselines")
        if variance_type != "fixed_small":
            raise NotImplementedError("only variance_type='fixed_small' (TrainDiffusionPolicy.py:136) is supported")
        super().__init__(num_train_timesteps, beta_start, beta_end, beta_schedule, clip_sample, prediction_type)
        self.variance_type = variance_type

    def set_timesteps(self, num_inference_steps):
        n = min(self.num_train_timesteps, int(num_inference_steps))
        self.num_inference_steps = n
        self.timesteps = np.arange(0, self.num_train_timesteps, self.num_train_timesteps // n)[::-1].copy()

    def _step_scalars(self, t):
        a_t = self.alphas_cumprod[t]
        a_prev = self.alphas_cumprod[t - 1] if t > 0 else self.one
        b_t, b_prev = 1 - a_t, 1 - a_prev
        c_x0 = (a_prev ** 0.5 * self.betas[t]) / b_t
        c_xt = self.alphas[t] ** 0.5 * b_prev / b_t
        sigma = 0.0
        if t > 0:
            variance = torch.clamp((1 - a_prev) / (1 - a_t) * self.betas[t], min=1e-20)
            sigma = float(variance ** 0.5)
        return dict(t=t, s_beta=float(b_t ** 0.5), s_alpha=float(a_t ** 0.5), c_x0=float(c_x0), c_xt=float(c_xt),
                    sigma=sigma)


class DDIMScheduler(_SchedulerBase):
    """diffusers 0.11.1 DDIMScheduler with eta = 0 (DP3.conditional_sample passes no eta).

    With prediction_type "sample" release 0.11.1 multiplies the direction term by the raw model output;
    derive_epsilon=True selects the later-release behaviour (epsilon derived from x_t and the predicted x_0)."""

    kind = 1

    def __init__(self, num_train_timesteps=1000, beta_start=0.0001, beta_end=0.02, beta_schedule="linear",
                 trained_betas=None, clip_sample=True, set_alpha_to_one=True, steps_offset=0,
                 prediction_type="epsilon", derive_epsilon=False):
        if trained_betas is not None:
            raise NotImplementedError("trained_betas is not used by RoboManipBaselines")
        super().__init__(num_train_timesteps, beta_start, beta_end, beta_schedule, clip_sample, prediction_type)
        self.final_alpha_cumprod = torch.tensor(1.0) if set_alpha_to_one else self.alphas_cumprod[0]
        self.steps_offset = int(steps_offset)
        self.derive_epsilon = bool(derive_epsilon)

    def set_timesteps(self, num_inference_steps):
        self.num_inference_steps = int(num_inference_steps)
        step_ratio = self.num_train_timesteps // self.num_inference_steps
        self.timesteps = (np.arange(0, self.num_inference_steps) * step_ratio).round()[::-1].copy().astype(np.int64)
        self.timesteps += self.steps_offset

    def _step_scalars(self, t):
        prev_t = t - self.num_train_timesteps // self.num_inference_steps
        a_t = self.alphas_cumprod[t]
        a_prev = self.alphas_cumprod[prev_t] if prev_t >= 0 else self.final_alpha_cumprod
        b_t = 1 - a_t
        return dict(t=t, s_beta=float(b_t ** 0.5), s_alpha=float(a_t ** 0.5), c_x0=float(a_prev ** 0.5),
                    c_xt=float((1 - a_prev) ** 0.5), derive_eps=self.derive_epsilon)
