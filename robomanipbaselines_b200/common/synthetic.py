"""Deterministic random-init weights and observations for benchmarks, smoke tests and parity fixtures.

There is no network for checkpoints, so "random-init weights of that architecture" are generated here as a pure
function of (key name, shape, seed): the same state dict can be rebuilt on any machine and loaded both into the
reference modules (golden generation) and into this package's policies.

synthetic_state_dict draws from distributions chosen so that the network is conditioned like a trained one
(BatchNorm-bounded activations, unsaturated softmax, non-trivial FrozenBN statistics so that BN folding is exercised);
the golden fixtures are made with it. default_init_state_dict reproduces the distributions of a freshly constructed
reference policy ("random-init weights" in BASELINE.json's words); bench.py uses it, and
tests/test_dtype_builds_gpu.py::test_default_init_weights_engine_error_vs_reference_sensitivity measures the engine on
it next to the fp32 reference's own sensitivity to bf16-rounded inputs (B200: engine 8.7e-4 / 9.8e-4 against a
sensitivity of 7.3e-4 / 1.3e-4 on the small / c1_1cam cases, tolerance 1e-2).
"""
import zlib

import numpy as np
import torch


def _gen(key, seed):
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(key.encode()) + 1000003 * int(seed)) % (2 ** 31 - 1))
    return g


def synthetic_tensor(key, shape, seed=0):
    shape = tuple(int(s) for s in shape)
    g = _gen(key, seed)
    is_bn = (".bn" in key) or (".downsample.1." in key)
    if key.endswith("nets.1.weight") and len(shape) == 1:        # GroupNorm gain after the stem of robomimic's ResNet18Conv
        return torch.rand(shape, generator=g) + 0.5
    if key.endswith("running_var"):
        return torch.rand(shape, generator=g) + 0.5
    if key.endswith("bn2.weight"):
        # small gain on the last BN of every residual branch (cf. torchvision's zero_init_residual): keeps the
        # activations O(1) through 8 residual blocks, as trained BN statistics do
        return 0.15 + 0.2 * torch.rand(shape, generator=g)
    if is_bn and key.endswith(".weight"):
        return torch.rand(shape, generator=g) + 0.5
    if is_bn:
        return torch.randn(shape, generator=g) * 0.1
    if ".norm" in key and key.endswith(".weight"):
        return 1.0 + 0.1 * torch.randn(shape, generator=g)
    is_diffusion = key.split(".")[0] in ("model", "obs_encoder") and (".block." in key or "_modules." in key or "extractor." in key or "final_conv" in key) \
        or key.split(".")[0] in ("down_modules", "up_modules", "mid_modules", "final_conv", "extractor")
    if is_diffusion and len(shape) == 3:     # Conv1d / ConvTranspose1d of ConditionalUnet1D
        return torch.randn(shape, generator=g) * (1.0 / (shape[1] * shape[2])) ** 0.5
    if is_diffusion and len(shape) == 1 and key.endswith(".weight"):   # GroupNorm / LayerNorm gains of the diffusion policies
        return 1.0 + 0.1 * torch.randn(shape, generator=g)
    if len(shape) == 4:
        fan_in = shape[1] * shape[2] * shape[3]
        return torch.randn(shape, generator=g) * (2.0 / fan_in) ** 0.5
    if key.endswith("in_proj_weight"):
        return torch.randn(shape, generator=g) * 0.5 * (1.0 / shape[1]) ** 0.5     # moderate attention logits
    if "embed" in key and len(shape) == 2:
        return torch.randn(shape, generator=g) * 0.5
    if len(shape) == 2:
        return torch.randn(shape, generator=g) * (1.0 / shape[1]) ** 0.5
    if len(shape) == 1:
        return torch.randn(shape, generator=g) * 0.05
    return torch.randn(shape, generator=g)


def synthetic_state_dict(shapes, seed=0):
    """shapes: mapping key -> shape (e.g. {k: v.shape for k, v in module.state_dict().items()})."""
    return {k: synthetic_tensor(k, s, seed) for k, s in shapes.items()}


def synthetic_norm_stats(dim, seed=0, norm_type="gaussian"):
    """model_meta_info['state'|'action']-style statistics with non-trivial values (TrainBase.py:309-344 layout)."""
    rng = np.random.default_rng(1000 + seed)
    mn = rng.normal(size=dim)
    rg = np.clip(rng.uniform(0.5, 2.0, size=dim), 1e-3, None)
    stats = {"mean": rng.normal(size=dim), "std": np.clip(rng.uniform(0.2, 1.5, size=dim), 1e-3, None), "min": mn,
             "max": mn + rg, "range": rg, "example": np.zeros(dim)}
    if norm_type == "limits":
        stats["norm_config"] = {"type": "limits", "out_min": -1.0, "out_max": 1.0}
    elif norm_type == "gaussian":
        stats["norm_config"] = {"type": "gaussian"}
    return stats


def synthetic_act_observations(batch, num_cams, img_h, img_w, state_dim, seed=1):
    """uint8 U{0..255} images (B, Ncam, H, W, 3) HWC and N(0,1) float64 raw joint states (B, Ds)."""
    rng = np.random.default_rng(seed)
    images = rng.integers(0, 256, size=(batch, num_cams, img_h, img_w, 3), dtype=np.uint8)
    state = rng.standard_normal(size=(batch, state_dim))
    return images, state


def default_init_tensor(key, shape, seed=0):
    """The distributions PyTorch / the reference give a freshly constructed ACTPolicy (BASELINE.json: "random-init
    weights"): torchvision resnet18 convs kaiming_normal(fan_out, relu); FrozenBatchNorm2d buffers weight 1, bias 0,
    mean 0, var 1 (third_party/act/detr/models/backbone.py:30-36); every >= 2-D transformer parameter xavier_uniform
    (third_party/act/detr/models/transformer.py:44-47); nn.Linear / Conv2d elsewhere kaiming_uniform(a = sqrt 5) with
    U(+-1/sqrt(fan_in)) biases; nn.Embedding N(0, 1); LayerNorm 1 / 0. Same distributions, own RNG stream."""
    shape = tuple(int(s) for s in shape)
    g = _gen("default." + key, seed)
    if ".bn" in key or ".downsample.1." in key:
        if key.endswith("weight") or key.endswith("running_var"):
            return torch.ones(shape)
        return torch.zeros(shape)
    if "backbones" in key and len(shape) == 4:
        fan_out = shape[0] * shape[2] * shape[3]
        return torch.randn(shape, generator=g) * (2.0 / fan_out) ** 0.5
    if ".norm" in key:
        return torch.ones(shape) if key.endswith("weight") else torch.zeros(shape)
    in_transformer = ".transformer." in key or key.startswith("model.encoder.")
    if in_transformer and len(shape) >= 2:
        bound = (6.0 / (shape[0] + shape[1])) ** 0.5
        return (torch.rand(shape, generator=g) * 2 - 1) * bound
    if in_transformer and key.endswith("in_proj_bias"):
        return torch.zeros(shape)
    if in_transformer and key.endswith("out_proj.bias"):
        return torch.zeros(shape)
    if "embed" in key and len(shape) == 2:
        return torch.randn(shape, generator=g)
    if len(shape) >= 2:
        fan_in = int(np.prod(shape[1:]))
        bound = (1.0 / fan_in) ** 0.5                      # kaiming_uniform(a = sqrt(5)) == U(+-1/sqrt(fan_in))
        return (torch.rand(shape, generator=g) * 2 - 1) * bound
    if key.endswith("pos_table"):
        return torch.zeros(shape)
    # biases of Linear / Conv layers: U(+-1/sqrt(fan_in)); fan_in is not recoverable from a 1-D shape alone, and the
    # scale of a bias does not condition the network: use the hidden width
    return (torch.rand(shape, generator=g) * 2 - 1) * (1.0 / 512) ** 0.5


def default_init_state_dict(shapes, seed=0):
    return {k: default_init_tensor(k, s, seed) for k, s in shapes.items()}
