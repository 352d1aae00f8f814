"""CPU tests of the drop-in boundary: the shared library loads without a GPU driver, exports every symbol that
include/rmb_b200.h declares, rejects bad arguments with error codes, and the Python mirror keeps the reference's
state-dict layout and weight-table arithmetic."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from robomanipbaselines_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "rmb_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rmb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in rmb_b200.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature in _lib.py"
    assert set(_lib.SIGNATURES) == set(names)
    assert lib.rmb_act_dtype() in (b"fp16", b"bf16")
    assert b"sm_100a" in lib.rmb_version()


def test_argument_errors_without_gpu():
    lib = _lib.load()
    assert lib.rmb_act_engine_create(None, None) < 0
    assert b"null" in lib.rmb_last_error()
    cfg = _lib.ActConfig(1, 1, 480, 640, 7, 7, 256, 3200, 8, 4, 7, 100, 0)     # hidden 256 unsupported
    h = ctypes.c_void_p()
    assert lib.rmb_act_engine_create(ctypes.byref(cfg), ctypes.byref(h)) < 0
    assert b"hidden_dim" in lib.rmb_last_error()
    cfg = _lib.ActConfig(1, 1, 470, 640, 7, 7, 512, 3200, 8, 4, 7, 100, 0)     # H not a multiple of 32
    assert lib.rmb_act_engine_create(ctypes.byref(cfg), ctypes.byref(h)) < 0
    assert lib.rmb_ensemble_step(None, None, None, None, None, 0, None, None, 0.0, None, None, None, 1, 1, 1, None) < 0
    assert lib.rmb_normalize(None, 0, None, None, 0.0, None, 1, 1, None) < 0
    with pytest.raises(_lib.RmbError):
        _lib.check(-1, "x")


def test_engine_requires_cuda():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from robomanipbaselines_b200.engine import ActEngine

    with pytest.raises(_lib.RmbError):
        ActEngine(max_batch=1, num_cams=1, img_h=64, img_w=96, state_dim=7, action_dim=7)


def test_policy_shell_state_dict_layout():
    from robomanipbaselines_b200.policy.act import ACTPolicy, DETRVAE

    args = {"num_queries": 100, "hidden_dim": 512, "dim_feedforward": 3200, "enc_layers": 4, "dec_layers": 7,
            "nheads": 8, "camera_names": ["front", "side"], "backbone": "resnet18", "kl_weight": 10}
    DETRVAE.set_state_dim(9)
    DETRVAE.set_action_dim(8)
    a = ACTPolicy(args)
    DETRVAE.set_state_dim(7)
    DETRVAE.set_action_dim(7)
    b = ACTPolicy(args)
    assert (a.state_dim, a.action_dim, b.state_dim, b.action_dim) == (9, 8, 7, 7)      # per instance
    sd = b.state_dict()
    assert len(sd) == 344                                                            # reference key count
    assert sd["model.pos_table"].shape == (1, 102, 512)
    assert sd["model.transformer.decoder.layers.6.multihead_attn.in_proj_weight"].shape == (1536, 512)
    assert sd["model.backbones.0.0.body.layer4.0.downsample.1.running_var"].shape == (512,)
    assert sd["model.input_proj_robot_state.weight"].shape == (512, 7)
    assert a.state_dict()["model.action_head.weight"].shape == (8, 512)
    # strict load incl. the tolerated num_batches_tracked keys of BatchNorm checkpoints (backbone.py:36-45)
    sd2 = dict(sd)
    sd2["model.backbones.0.0.body.bn1.num_batches_tracked"] = torch.tensor(0)
    b.load_state_dict(sd2, strict=True)
    with pytest.raises(RuntimeError):
        b.load_state_dict({k: v for k, v in sd.items() if "action_head" not in k}, strict=True)
    with pytest.raises(RuntimeError):
        b(torch.zeros(1, 7), torch.zeros(1, 2, 3, 480, 640))                       # CPU tensors: no fallback
    assert sum(p.numel() for p in b.parameters()) > 80e6                             # calc_model_size works


def test_weight_table_matches_reference_expression():
    from robomanipbaselines_b200.common.ensemble import make_weight_table

    C = 100
    tab = make_weight_table(C)
    for m in (1, 2, 37, 100):
        w = np.exp(-0.01 * np.arange(m))                       # RolloutAct.py:90-92
        w = w / w.sum()
        assert np.array_equal(tab[m - 1, :m], w)
        assert np.all(tab[m - 1, m:] == 0)
    assert tab[99, 0] > tab[99, 99]                            # oldest chunk gets the largest weight


def test_norm_params_follow_datautils():
    from oracle.rollout_oracle import denormalize_data, normalize_data
    from robomanipbaselines_b200.common.batched_rollout import _norm_params

    rng = np.random.default_rng(0)
    st = {"mean": rng.normal(size=5), "std": rng.uniform(0.1, 2, size=5), "min": rng.normal(size=5),
          "range": rng.uniform(0.1, 2, size=5)}
    x = rng.normal(size=(4, 5))
    t, p0, p1, om = _norm_params(st, "normalize")
    assert t == 0 and np.array_equal((x - p0) / p1, normalize_data(x, st))
    t, p0, p1, om = _norm_params(st, "denormalize")
    assert np.array_equal(p0 * x + p1, denormalize_data(x, st))
    lim = dict(st, norm_config={"type": "limits", "out_min": -1.0, "out_max": 1.0})
    t, p0, p1, om = _norm_params(lim, "normalize")
    assert t == 1 and np.array_equal(p1 * (x - p0) + om, normalize_data(x, lim))
    t, p0, p1, om = _norm_params(lim, "denormalize")
    assert np.array_equal(p0 * (x - om) + p1, denormalize_data(x, lim))
    with pytest.raises(ValueError):
        _norm_params(dict(st, norm_config={"type": "bogus"}), "normalize")
