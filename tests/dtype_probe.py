"""Runs in a SUBPROCESS of tests/test_dtype_builds_gpu.py with RMB_LIB_DTYPE=fp16|bf16 (one process loads exactly one
build of the library) and prints one JSON line: errors of the golden ACT cases and two UNet cases, and the saturation
diagnostics on a network whose trunk activations are scaled out of the fp16 range.

    RMB_LIB_DTYPE=bf16 python tests/dtype_probe.py
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.join(HERE, "golden"))

from cases import ACT_CASES, act_inputs, policy_args  # noqa: E402
from robomanipbaselines_b200 import _lib  # noqa: E402
from robomanipbaselines_b200.common.synthetic import synthetic_state_dict  # noqa: E402
from robomanipbaselines_b200.policy.act import ACTPolicy  # noqa: E402

GOLD = os.path.join(HERE, "golden")
TRUNK = "model.backbones.0.0.body."


def scaled_trunk_state_dict(sd, S):
    """Same function in exact arithmetic, but every ResNet trunk activation is S times larger: conv1 weights and every
    folded BN bias of the trunk x S (ReLU / max-pool / residual adds are positively homogeneous), input_proj weights / S."""
    out = {k: v.clone() for k, v in sd.items()}
    out[TRUNK + "conv1.weight"] = sd[TRUNK + "conv1.weight"] * S
    for k in sd:                       # FrozenBatchNorm2d: y = c * scale + (b - rm * scale); S * y needs b * S, rm * S
        is_bn = k.startswith(TRUNK) and (".bn" in k or ".downsample.1." in k or k.startswith(TRUNK + "bn1."))
        if is_bn and (k.endswith(".bias") or k.endswith(".running_mean")):
            out[k] = sd[k] * S
    out["model.input_proj.weight"] = sd["model.input_proj.weight"] / S
    return out


def act_case(name, cuda, sd_transform=None, check_sat=False):
    case = ACT_CASES[name]
    gold = np.load(os.path.join(GOLD, f"act_golden_{name}.npz"))
    pol = ACTPolicy(policy_args(case), state_dim=case["state_dim"], action_dim=case["action_dim"])
    sd = synthetic_state_dict({k: tuple(v.shape) for k, v in pol.state_dict().items()}, seed=case["weight_seed"])
    if sd_transform is not None:
        sd = sd_transform(sd)
    pol.load_state_dict(sd, strict=True)
    qpos, img = act_inputs(case)
    eng = pol.engine_for(case["batch"], case["img_h"], case["img_w"], cuda)
    eng.saturation_count(clear=True)
    eng.check_saturation(check_sat)
    a = pol(torch.from_numpy(qpos).to(cuda), torch.from_numpy(img).to(cuda))
    torch.cuda.synchronize()
    sat = eng.saturation_count(clear=True) if check_sat else None
    a = a.cpu().numpy()
    err = float(np.abs(a - gold["a_hat"]).max()) if np.isfinite(a).all() else float("inf")
    return err, sat


def unet_case(cuda, down_dims, B):
    from oracle import diffusion_oracle as dor
    from robomanipbaselines_b200.engine import UnetEngine

    T, Da, G, dsed, k = 16, 7, 142, 128, 5
    ref = dor.ConditionalUnet1D(Da, G, dsed, down_dims, k).eval()
    sd = synthetic_state_dict({kk: tuple(v.shape) for kk, v in ref.state_dict().items()}, seed=0)
    ref.load_state_dict(sd)
    eng = UnetEngine(max_batch=B, horizon=T, input_dim=Da, global_cond_dim=G, diffusion_step_embed_dim=dsed,
                     down_dims=down_dims, kernel_size=k, device=0)
    eng.load_state_dict(sd)
    g = torch.Generator().manual_seed(2)
    x, gc = torch.randn(B, T, Da, generator=g), torch.randn(B, G, generator=g)
    with torch.no_grad():
        want = ref(x, torch.full((B,), 37), gc)
    got = eng.forward(x.to(cuda), 37, gc.to(cuda)).cpu()
    return float((got - want).abs().max())


def main():
    cuda = torch.device("cuda:0")
    lib = _lib.load()
    out = {"dtype": lib.rmb_act_dtype().decode(), "lib": os.path.basename(_lib.LIB_PATH)}
    for name in ("small", "c1_1cam", "c1_2cam"):
        out[f"act_{name}_err"], _ = act_case(name, cuda)
    # diagnostics: off-range trunk (x 4096: stem outputs reach ~1e5) and the normal network with the check on
    out["normal_err"], out["normal_sat"] = act_case("small", cuda, None, check_sat=True)
    out["hostile_err"], out["hostile_sat"] = act_case("small", cuda, lambda sd: scaled_trunk_state_dict(sd, 4096.0), check_sat=True)
    out["unet_mid3_err"] = unet_case(cuda, (64, 128, 256), 5)
    out["unet_full_err"] = unet_case(cuda, (512, 1024, 2048), 2)
    out["device_fault"] = int(lib.rmb_device_fault(1))
    print("DTYPE_PROBE " + json.dumps(out))


if __name__ == "__main__":
    main()
