"""TEST INFRASTRUCTURE ONLY (parity checker; never on the product path).

fp32 PyTorch-CPU restatement of the Diffusion Policy / DP3 denoising path that RoboManipBaselines calls at
  robo_manip_baselines/policy/diffusion_policy/RolloutDiffusionPolicy.py:37-43,74        (DDPMScheduler + policy.predict_action)
  robo_manip_baselines/policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:50-56,90  (DDIMScheduler + DP3.predict_action)
with the hyper-parameters of TrainDiffusionPolicy.py:114-138 / TrainDiffusionPolicy3d.py:185-218.

PARITY UNPINNED. The arithmetic lives in sources that are NOT in /root/reference:
  * third_party/diffusion_policy      = github.com/mmurooka/diffusion_policy @ 7e8fea759bfcf6680b5d5b35cd48f836baf010cd
  * third_party/3D-Diffusion-Policy   = github.com/Naoki-Shibayama/3D-Diffusion-Policy @ 95ad86abde3cedd436bd1ad7cdbcb8bfb9d0806b
  * diffusers == 0.11.1 (pyproject.toml:69,79), not installed here
so this file restates their published algorithms (ConditionalUnet1D, DDPMScheduler.step, DDIMScheduler.step,
DP3Encoder / PointNetEncoderXYZ) from the upstream public code as recalled in SURVEY.md 3.5; module and parameter
names follow upstream so that a real checkpoint's state-dict keys ("model.down_modules.0.0.blocks.0.block.0.weight",
...) load. Nothing of the reference pins it; the CUDA engine is tested against THIS restatement.
"""
import math

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------------ ConditionalUnet1D
class SinusoidalPosEmb(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.dim = dim

    def forward(self, x):
        half = self.dim // 2
        emb = math.log(10000) / (half - 1)
        emb = torch.exp(torch.arange(half, dtype=torch.float32) * -emb)
        emb = x[:, None].float() * emb[None, :]
        return torch.cat((emb.sin(), emb.cos()), dim=-1)


class Downsample1d(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.conv = nn.Conv1d(dim, dim, 3, 2, 1)

    def forward(self, x):
        return self.conv(x)


class Upsample1d(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.conv = nn.ConvTranspose1d(dim, dim, 4, 2, 1)

    def forward(self, x):
        return self.conv(x)


class Conv1dBlock(nn.Module):
    """Conv1d --> GroupNorm --> Mish"""

    def __init__(self, inp_channels, out_channels, kernel_size, n_groups=8):
        super().__init__()
        self.block = nn.Sequential(
            nn.Conv1d(inp_channels, out_channels, kernel_size, padding=kernel_size // 2),
            nn.GroupNorm(n_groups, out_channels),
            nn.Mish(),
        )

    def forward(self, x):
        return self.block(x)


class ConditionalResidualBlock1D(nn.Module):
    def __init__(self, in_channels, out_channels, cond_dim, kernel_size=3, n_groups=8, cond_predict_scale=True):
        super().__init__()
        self.blocks = nn.ModuleList([
            Conv1dBlock(in_channels, out_channels, kernel_size, n_groups=n_groups),
            Conv1dBlock(out_channels, out_channels, kernel_size, n_groups=n_groups),
        ])
        cond_channels = out_channels * 2 if cond_predict_scale else out_channels
        self.cond_predict_scale = cond_predict_scale
        self.out_channels = out_channels
        self.cond_encoder = nn.Sequential(nn.Mish(), nn.Linear(cond_dim, cond_channels), nn.Unflatten(-1, (-1, 1)))
        self.residual_conv = nn.Conv1d(in_channels, out_channels, 1) if in_channels != out_channels else nn.Identity()

    def forward(self, x, cond):
        out = self.blocks[0](x)
        embed = self.cond_encoder(cond)
        if self.cond_predict_scale:
            embed = embed.reshape(embed.shape[0], 2, self.out_channels, 1)
            out = embed[:, 0] * out + embed[:, 1]
        else:
            out = out + embed
        out = self.blocks[1](out)
        return out + self.residual_conv(x)


class ConditionalUnet1D(nn.Module):
    def __init__(self, input_dim, global_cond_dim, diffusion_step_embed_dim=128, down_dims=(512, 1024, 2048),
                 kernel_size=5, n_groups=8, cond_predict_scale=True):
        super().__init__()
        all_dims = [input_dim] + list(down_dims)
        start_dim = down_dims[0]
        dsed = diffusion_step_embed_dim
        self.diffusion_step_encoder = nn.Sequential(
            SinusoidalPosEmb(dsed), nn.Linear(dsed, dsed * 4), nn.Mish(), nn.Linear(dsed * 4, dsed))
        cond_dim = dsed + global_cond_dim
        in_out = list(zip(all_dims[:-1], all_dims[1:]))
        kw = dict(cond_dim=cond_dim, kernel_size=kernel_size, n_groups=n_groups, cond_predict_scale=cond_predict_scale)
        mid_dim = all_dims[-1]
        self.mid_modules = nn.ModuleList([
            ConditionalResidualBlock1D(mid_dim, mid_dim, **kw), ConditionalResidualBlock1D(mid_dim, mid_dim, **kw)])
        down_modules = nn.ModuleList([])
        for ind, (dim_in, dim_out) in enumerate(in_out):
            is_last = ind >= (len(in_out) - 1)
            down_modules.append(nn.ModuleList([
                ConditionalResidualBlock1D(dim_in, dim_out, **kw), ConditionalResidualBlock1D(dim_out, dim_out, **kw),
                Downsample1d(dim_out) if not is_last else nn.Identity()]))
        up_modules = nn.ModuleList([])
        for ind, (dim_in, dim_out) in enumerate(reversed(in_out[1:])):
            is_last = ind >= (len(in_out) - 1)
            up_modules.append(nn.ModuleList([
                ConditionalResidualBlock1D(dim_out * 2, dim_in, **kw), ConditionalResidualBlock1D(dim_in, dim_in, **kw),
                Upsample1d(dim_in) if not is_last else nn.Identity()]))
        self.final_conv = nn.Sequential(Conv1dBlock(start_dim, start_dim, kernel_size=kernel_size),
                                        nn.Conv1d(start_dim, input_dim, 1))
        self.up_modules = up_modules
        self.down_modules = down_modules

    def forward(self, sample, timestep, global_cond):
        """sample (B, T, input_dim), timestep int or (B,), global_cond (B, global_cond_dim) -> (B, T, input_dim)."""
        sample = sample.transpose(1, 2)                       # b h t -> b t h
        timesteps = torch.as_tensor(timestep, dtype=torch.long)
        if timesteps.ndim == 0:
            timesteps = timesteps[None]
        timesteps = timesteps.expand(sample.shape[0])
        global_feature = torch.cat([self.diffusion_step_encoder(timesteps), global_cond], dim=-1)
        x = sample
        h = []
        for resnet, resnet2, downsample in self.down_modules:
            x = resnet(x, global_feature)
            x = resnet2(x, global_feature)
            h.append(x)
            x = downsample(x)
        for mid_module in self.mid_modules:
            x = mid_module(x, global_feature)
        for resnet, resnet2, upsample in self.up_modules:
            x = torch.cat((x, h.pop()), dim=1)
            x = resnet(x, global_feature)
            x = resnet2(x, global_feature)
            x = upsample(x)
        x = self.final_conv(x)
        return x.transpose(1, 2)


# ------------------------------------------------------------------------------------------------ noise schedulers
def betas_for_alpha_bar(num_diffusion_timesteps, max_beta=0.999):
    """diffusers 0.11.1 `betas_for_alpha_bar` (beta_schedule "squaredcos_cap_v2")."""
    def alpha_bar(time_step):
        return math.cos((time_step + 0.008) / 1.008 * math.pi / 2) ** 2

    betas = []
    for i in range(num_diffusion_timesteps):
        t1 = i / num_diffusion_timesteps
        t2 = (i + 1) / num_diffusion_timesteps
        betas.append(min(1 - alpha_bar(t2) / alpha_bar(t1), max_beta))
    return torch.tensor(betas, dtype=torch.float32)


def _make_betas(num_train_timesteps, beta_start, beta_end, beta_schedule):
    if beta_schedule == "linear":
        return torch.linspace(beta_start, beta_end, num_train_timesteps, dtype=torch.float32)
    if beta_schedule == "scaled_linear":
        return torch.linspace(beta_start ** 0.5, beta_end ** 0.5, num_train_timesteps, dtype=torch.float32) ** 2
    if beta_schedule == "squaredcos_cap_v2":
        return betas_for_alpha_bar(num_train_timesteps)
    raise NotImplementedError(beta_schedule)


class DDPMSchedulerOracle:
    """diffusers 0.11.1 DDPMScheduler restricted to what TrainDiffusionPolicy.py:127-137 configures
    (variance_type fixed_small, prediction_type epsilon | sample, clip_sample)."""

    def __init__(self, num_train_timesteps=100, beta_start=0.0001, beta_end=0.02, beta_schedule="squaredcos_cap_v2",
                 variance_type="fixed_small", clip_sample=True, prediction_type="epsilon"):
        assert variance_type == "fixed_small"
        self.num_train_timesteps = num_train_timesteps
        self.betas = _make_betas(num_train_timesteps, beta_start, beta_end, beta_schedule)
        self.alphas = 1.0 - self.betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.one = torch.tensor(1.0)
        self.clip_sample = clip_sample
        self.prediction_type = prediction_type
        self.timesteps = np.arange(0, num_train_timesteps)[::-1].copy()

    def set_timesteps(self, num_inference_steps):
        n = min(self.num_train_timesteps, num_inference_steps)
        self.num_inference_steps = n
        self.timesteps = np.arange(0, self.num_train_timesteps, self.num_train_timesteps // n)[::-1].copy()

    def step(self, model_output, t, sample, noise=None):
        t = int(t)
        alpha_prod_t = self.alphas_cumprod[t]
        alpha_prod_t_prev = self.alphas_cumprod[t - 1] if t > 0 else self.one
        beta_prod_t = 1 - alpha_prod_t
        beta_prod_t_prev = 1 - alpha_prod_t_prev
        if self.prediction_type == "epsilon":
            x0 = (sample - beta_prod_t ** 0.5 * model_output) / alpha_prod_t ** 0.5
        else:
            x0 = model_output
        if self.clip_sample:
            x0 = torch.clamp(x0, -1, 1)
        c_x0 = (alpha_prod_t_prev ** 0.5 * self.betas[t]) / beta_prod_t
        c_xt = self.alphas[t] ** 0.5 * beta_prod_t_prev / beta_prod_t
        prev = c_x0 * x0 + c_xt * sample
        if t > 0:
            variance = torch.clamp((1 - alpha_prod_t_prev) / (1 - alpha_prod_t) * self.betas[t], min=1e-20)
            prev = prev + (variance ** 0.5) * noise
        return prev


class DDIMSchedulerOracle:
    """diffusers 0.11.1 DDIMScheduler (eta = 0) as configured by TrainDiffusionPolicy3d.py:208-217.

    derive_epsilon=False is the 0.11.1 behaviour for prediction_type "sample": the direction term multiplies the raw
    model output (the x0 prediction); later diffusers releases first derive epsilon from (x_t, x0) -- see SURVEY 3.5.
    """

    def __init__(self, num_train_timesteps=100, beta_start=0.0001, beta_end=0.02, beta_schedule="squaredcos_cap_v2",
                 clip_sample=True, set_alpha_to_one=True, steps_offset=0, prediction_type="sample",
                 derive_epsilon=False):
        self.num_train_timesteps = num_train_timesteps
        self.betas = _make_betas(num_train_timesteps, beta_start, beta_end, beta_schedule)
        self.alphas = 1.0 - self.betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.final_alpha_cumprod = torch.tensor(1.0) if set_alpha_to_one else self.alphas_cumprod[0]
        self.clip_sample = clip_sample
        self.steps_offset = steps_offset
        self.prediction_type = prediction_type
        self.derive_epsilon = derive_epsilon

    def set_timesteps(self, num_inference_steps):
        self.num_inference_steps = num_inference_steps
        step_ratio = self.num_train_timesteps // num_inference_steps
        self.timesteps = (np.arange(0, num_inference_steps) * step_ratio).round()[::-1].copy().astype(np.int64)
        self.timesteps += self.steps_offset

    def step(self, model_output, t, sample, noise=None):
        t = int(t)
        prev_t = t - self.num_train_timesteps // self.num_inference_steps
        alpha_prod_t = self.alphas_cumprod[t]
        alpha_prod_t_prev = self.alphas_cumprod[prev_t] if prev_t >= 0 else self.final_alpha_cumprod
        beta_prod_t = 1 - alpha_prod_t
        if self.prediction_type == "epsilon":
            x0 = (sample - beta_prod_t ** 0.5 * model_output) / alpha_prod_t ** 0.5
            direction_src = model_output
        else:
            x0 = model_output
            direction_src = model_output
        if self.clip_sample:
            x0 = torch.clamp(x0, -1, 1)
        if self.derive_epsilon and self.prediction_type != "epsilon":
            direction_src = (sample - alpha_prod_t ** 0.5 * x0) / beta_prod_t ** 0.5
        direction = (1 - alpha_prod_t_prev) ** 0.5 * direction_src
        return alpha_prod_t_prev ** 0.5 * x0 + direction


def conditional_sample(model, scheduler, global_cond, x_init, step_noise, num_inference_steps):
    """DiffusionUnet*Policy.conditional_sample / DP3.conditional_sample with obs_as_global_cond: x_init (B, T, Da) is the
    initial torch.randn trajectory, step_noise[i] the DDPM noise drawn at loop iteration i (None for DDIM)."""
    x = x_init.clone()
    scheduler.set_timesteps(num_inference_steps)
    with torch.no_grad():
        for i, t in enumerate(scheduler.timesteps):
            out = model(x, int(t), global_cond)
            x = scheduler.step(out, int(t), x, None if step_noise is None else step_noise[i])
    return x


# ------------------------------------------------------------------------------------------------ DP3 encoder
class PointNetEncoderXYZ(nn.Module):
    """DP3 `PointNetEncoderXYZ` (use_layernorm=True, final_norm="layernorm"): per-point MLP 3-64-128-256 with
    LayerNorm + ReLU, max over points, Linear 256 -> out, LayerNorm."""

    def __init__(self, in_channels=3, out_channels=64):
        super().__init__()
        block_channel = [64, 128, 256]
        self.mlp = nn.Sequential(
            nn.Linear(in_channels, block_channel[0]), nn.LayerNorm(block_channel[0]), nn.ReLU(),
            nn.Linear(block_channel[0], block_channel[1]), nn.LayerNorm(block_channel[1]), nn.ReLU(),
            nn.Linear(block_channel[1], block_channel[2]), nn.LayerNorm(block_channel[2]), nn.ReLU())
        self.final_projection = nn.Sequential(nn.Linear(block_channel[-1], out_channels), nn.LayerNorm(out_channels))

    def forward(self, x):
        x = self.mlp(x)
        x = torch.max(x, 1)[0]
        return self.final_projection(x)


class DP3Encoder(nn.Module):
    """DP3 `DP3Encoder`: point cloud (B, Np, >=3) -> PointNetEncoderXYZ(xyz) ; state (B, Ds) -> MLP Ds-64-64 (ReLU between);
    output cat([pn_feat, state_feat]) of size out_channel + 64."""

    def __init__(self, state_dim, out_channel=64, use_pc_color=False):
        super().__init__()
        self.use_pc_color = use_pc_color
        self.extractor = PointNetEncoderXYZ(6 if use_pc_color else 3, out_channel)
        self.state_mlp = nn.Sequential(nn.Linear(state_dim, 64), nn.ReLU(), nn.Linear(64, 64))
        self.n_output_channels = out_channel + 64

    def forward(self, observations):
        points = observations["point_cloud"]
        if not self.use_pc_color:
            points = points[..., :3]
        pn_feat = self.extractor(points)
        state_feat = self.state_mlp(observations["agent_pos"])
        return torch.cat([pn_feat, state_feat], dim=-1)


# ------------------------------------------------------------------------------------------------ DP image encoder
class SpatialSoftmax(nn.Module):
    """robomimic 0.2.0 `SpatialSoftmax` (num_kp keypoints, temperature 1, no noise): 1x1 conv to num_kp maps, softmax
    over H*W, expected (x, y) with x, y in linspace(-1, 1)."""

    def __init__(self, in_c, in_h, in_w, num_kp=32):
        super().__init__()
        self.nets = nn.Conv2d(in_c, num_kp, kernel_size=1)
        pos_x, pos_y = np.meshgrid(np.linspace(-1.0, 1.0, in_w), np.linspace(-1.0, 1.0, in_h))
        self.register_buffer("pos_x", torch.from_numpy(pos_x.reshape(1, in_h * in_w)).float(), persistent=False)
        self.register_buffer("pos_y", torch.from_numpy(pos_y.reshape(1, in_h * in_w)).float(), persistent=False)
        self.num_kp, self.in_h, self.in_w = num_kp, in_h, in_w

    def forward(self, feature):
        feature = self.nets(feature).reshape(-1, self.in_h * self.in_w)
        attention = F.softmax(feature, dim=-1)
        expected_x = torch.sum(self.pos_x * attention, dim=1, keepdim=True)
        expected_y = torch.sum(self.pos_y * attention, dim=1, keepdim=True)
        return torch.cat([expected_x, expected_y], 1).view(-1, self.num_kp, 2)


class _ResNet18Conv(nn.Module):
    def __init__(self):
        super().__init__()
        import torchvision

        net = torchvision.models.resnet18(weights=None)
        self.nets = nn.Sequential(*(list(net.children())[:-2]))

    def forward(self, x):
        return self.nets(x)


def _bn_to_gn(module):
    """diffusion_policy `replace_submodules(BatchNorm2d -> GroupNorm(num_features // 16, num_features))`."""
    for name, child in list(module.named_children()):
        if isinstance(child, nn.BatchNorm2d):
            setattr(module, name, nn.GroupNorm(child.num_features // 16, child.num_features))
        else:
            _bn_to_gn(child)
    return module


class DpImageEncoder(nn.Module):
    """robomimic VisualCore for one rgb key as DiffusionUnetHybridImagePolicy builds it: centre crop (CropRandomizer in
    eval mode) -> ResNet18Conv with GroupNorm -> SpatialSoftmax(32) -> Flatten -> Linear(64, 64)."""

    def __init__(self, crop_h, crop_w, num_kp=32, feat_dim=64):
        super().__init__()
        self.crop_h, self.crop_w = crop_h, crop_w
        self.backbone = _bn_to_gn(_ResNet18Conv())
        fh, fw = int(math.ceil(crop_h / 32.0)), int(math.ceil(crop_w / 32.0))
        self.pool = SpatialSoftmax(512, fh, fw, num_kp)
        self.nets = nn.Sequential(self.backbone, self.pool, nn.Flatten(start_dim=1), nn.Linear(num_kp * 2, feat_dim))

    def forward(self, x):
        h, w = x.shape[-2:]
        t, l = (h - self.crop_h) // 2, (w - self.crop_w) // 2
        return self.nets(x[..., t:t + self.crop_h, l:l + self.crop_w])
