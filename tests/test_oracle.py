"""CPU tests: the oracle restatements must reproduce the golden vectors produced by the real reference
(tests/golden/make_golden.py, run in the build container against /root/reference)."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

from cases import ACT_CASES, act_inputs, policy_args  # noqa: E402
from oracle import act_oracle, rollout_oracle  # noqa: E402
from robomanipbaselines_b200.common.synthetic import synthetic_state_dict  # noqa: E402
from robomanipbaselines_b200.policy.act.policy import ACTPolicy  # noqa: E402

GOLD = os.path.join(HERE, "golden")


def _case_state_dict(case):
    shell = ACTPolicy(policy_args(case), state_dim=case["state_dim"], action_dim=case["action_dim"])
    shapes = {k: tuple(v.shape) for k, v in shell.state_dict().items()}
    return synthetic_state_dict(shapes, seed=case["weight_seed"])


@pytest.mark.parametrize("name", ["small", "c1_1cam", "c1_2cam"])
def test_act_oracle_matches_reference_golden(name):
    case = ACT_CASES[name]
    gold = np.load(os.path.join(GOLD, f"act_golden_{name}.npz"))
    sd = _case_state_dict(case)
    assert len(sd) == int(gold["n_keys"][0])
    wsum = float(sum(v.double().abs().sum() for v in sd.values()))
    assert abs(wsum - gold["weight_checksum"][0]) <= 1e-9 * wsum      # same synthetic weights as at golden time
    qpos, img_u8 = act_inputs(case)
    assert abs(float(np.abs(qpos).sum()) - gold["input_checksum"][0]) < 1e-4
    assert int(img_u8.astype(np.int64).sum()) == int(gold["input_checksum"][1])
    image = torch.from_numpy(np.concatenate([rollout_oracle.get_images(list(img_u8[b])) for b in range(case["batch"])]))
    taps = {}
    cfg = dict(policy_args(case))
    a_full = act_oracle.act_forward(sd, cfg, torch.from_numpy(qpos), image, live_only=False, taps=taps)
    err = np.abs(a_full.numpy() - gold["a_hat"]).max()
    print(f"[oracle {name}] max-abs vs reference {err:.3e}")
    assert err <= 2e-5
    # dead-layer claim: dropping decoder layers 1.. changes nothing
    a_live = act_oracle.act_forward(sd, cfg, torch.from_numpy(qpos), image, live_only=True)
    assert torch.equal(a_live, a_full)
    if "memory" in gold:
        assert np.abs(taps["memory"].numpy() - gold["memory"]).max() <= 2e-4          # post-LayerNorm, O(1)
        scale = np.abs(gold["src_tokens"]).max()                                      # raw backbone features
        assert np.abs(taps["src_tokens"].numpy() - gold["src_tokens"]).max() <= 1e-5 * scale
    else:
        assert np.abs(taps["memory"].mean(-1).numpy() - gold["memory_token_mean"]).max() <= 2e-4


@pytest.mark.parametrize("norm", ["gaussian", "limits", "legacy"])
def test_rollout_oracle_matches_reference_golden(norm):
    g = np.load(os.path.join(GOLD, "rollout_act_golden.npz"))
    stats = {k: g[f"{norm}_stat_{k}"] for k in ("mean", "std", "min", "range")}
    if norm == "limits":
        stats["norm_config"] = {"type": "limits", "out_min": -1.0, "out_max": 1.0}
    elif norm == "gaussian":
        stats["norm_config"] = {"type": "gaussian"}
    chunks = g[f"{norm}_chunks"]
    T, C, Da = chunks.shape
    for mode in ("ensemble", "fifo"):
        o = rollout_oracle.ActRolloutOracle(C, Da, stats, no_temp_ensem=(mode == "fifo"))
        calls, acts = 0, []
        for t in range(T):
            need = o.needs_inference()
            calls += int(need)
            acts.append(o.step(chunks[t] if need else None))
        assert np.array_equal(np.array(acts), g[f"{norm}_{mode}_actions"])          # bit-exact float64
        assert calls == int(g[f"{norm}_{mode}_calls"][0])
        if mode == "ensemble":
            assert np.array_equal(np.array(o.policy_action_buf_history), g[f"{norm}_history"])
    x = g[f"{norm}_norm_in"]
    assert np.array_equal(rollout_oracle.normalize_data(x, stats), g[f"{norm}_normalize"])
    assert np.array_equal(rollout_oracle.denormalize_data(x, stats), g[f"{norm}_denormalize"])


def test_get_state_and_images_oracle():
    g = np.load(os.path.join(GOLD, "rollout_act_golden.npz"))
    stats = {"mean": g["get_state_mean"], "std": g["get_state_std"], "norm_config": {"type": "gaussian"}}
    assert np.array_equal(rollout_oracle.get_state(g["get_state_raw"], stats), g["get_state_out"])
    assert np.array_equal(rollout_oracle.get_images(list(g["get_images_in"])), g["get_images_out"])


@pytest.mark.parametrize("name", ["small", "img84"])
def test_pointcloud_oracle_matches_reference_golden(name):
    """oracle/pointcloud_oracle.py against the outputs of the reference's own VisionUtils / Vision3dUtils functions
    (tests/golden/make_pointcloud_golden.py): bit for bit, including the float64 dtype flow."""
    from cases import POINTCLOUD_CASES, pointcloud_inputs
    from oracle import pointcloud_oracle as po

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "pointcloud_golden.npz"))
    c = POINTCLOUD_CASES[name]
    depth, rgb = pointcloud_inputs(c)
    pc = np.concat(po.convert_depth_image_to_pointcloud(depth, c["fovy"], rgb), axis=1)
    assert pc.dtype == g[name + "_pointcloud"].dtype and np.array_equal(pc, g[name + "_pointcloud"])
    assert np.array_equal(po.crop_pointcloud_bb(pc, c["min_bound"], c["max_bound"]), g[name + "_cropped"])
    idx = po.farthest_point_sampling(g[name + "_cropped"], 32)
    assert idx[0] == 0 and len(set(idx.tolist())) == 32            # distinct points, start index 0
    few = po.farthest_point_sampling(g[name + "_cropped"][:5], 8)
    assert sorted(few[:5].tolist()) == [0, 1, 2, 3, 4] and (few[5:] == -1).all()


def _dp_golden_stats(g, prefix, norm):
    s = {k: g[f"{prefix}_{k}"] for k in ("mean", "std", "min", "range")}
    s["norm_config"] = {"type": "limits", "out_min": -1.0, "out_max": 1.0} if norm == "limits" else {"type": "gaussian"}
    return s


@pytest.mark.parametrize("norm", ["gaussian", "limits"])
def test_dp_rollout_oracle_matches_reference_golden(norm):
    """DiffusionRolloutOracle + dp_get_image vs the reference's own RolloutDiffusionPolicy.infer_policy / get_state /
    get_images run (tests/golden/make_dp_rollout_golden.py): every policy input and every action bit-for-bit."""
    g = np.load(os.path.join(GOLD, "dp_rollout_golden.npz"))
    p = f"dp_{norm}"
    sstats, astats = _dp_golden_stats(g, f"{p}_state", norm), _dp_golden_stats(g, f"{p}_action", norm)
    raw, imgs, fake = g[f"{p}_raw_states"], g[f"{p}_images_u8"], g[f"{p}_fake_actions"]
    h, w = g[f"{p}_in_front"].shape[-2:]
    o = rollout_oracle.DiffusionRolloutOracle(g[f"{p}_in_state"].shape[1], sstats, astats)
    steps, k = [], 0
    for t in range(raw.shape[0]):
        if o.needs_inference():
            obs = o.push_obs(raw[t], front=rollout_oracle.dp_get_image(imgs[t, 0], (w, h)),
                             side=rollout_oracle.dp_get_image(imgs[t, 1], (w, h)))
            assert np.array_equal(obs["state"], g[f"{p}_in_state"][k])
            assert np.array_equal(obs["front"], g[f"{p}_in_front"][k])
            assert np.array_equal(obs["side"], g[f"{p}_in_side"][k])
            steps.append(t)
            a = o.step(fake[k])
            k += 1
        else:
            a = o.step()
        assert np.array_equal(a, g[f"{p}_actions"][t]), t
    assert steps == list(g[f"{p}_infer_steps"])


@pytest.mark.parametrize("norm", ["gaussian", "limits"])
def test_dp3_rollout_oracle_matches_reference_golden(norm):
    """Same for RolloutDiffusionPolicy3d (get_pointcloud: cv2.resize x2 -> back-projection -> crop -> [FPS stand-in of
    the golden run: first num_points rows] -> normalize_data -> float32 ring buffer)."""
    from oracle import pointcloud_oracle as pco

    g = np.load(os.path.join(GOLD, "dp_rollout_golden.npz"))
    p = f"dp3_{norm}"
    sstats, astats = _dp_golden_stats(g, f"{p}_state", norm), _dp_golden_stats(g, f"{p}_action", norm)
    pstats = _dp_golden_stats(g, f"{p}_pointcloud", norm)
    raw, rgb, depth, fake = g[f"{p}_raw_states"], g[f"{p}_rgb"], g[f"{p}_depth"], g[f"{p}_fake_actions"]
    n_points = g[f"{p}_in_pc"].shape[2]
    size = (12, 10)                                   # make_dp_rollout_golden.PC_SIZE (W, H)
    o = rollout_oracle.DiffusionRolloutOracle(g[f"{p}_in_state"].shape[1], sstats, astats)
    steps, k = [], 0
    for t in range(raw.shape[0]):
        if o.needs_inference():
            r, d = rollout_oracle.dp3_resize(rgb[t], depth[t], size)
            pc = np.concatenate(pco.convert_depth_image_to_pointcloud(d, float(g[f"{p}_fovy"][0]), r), axis=1)
            pc = pco.crop_pointcloud_bb(pc, [-0.6, -0.6, 0.3], [0.6, 0.6, 1.7])
            pc = pc[:n_points] if len(pc) >= n_points else np.concatenate([pc, np.zeros((n_points - len(pc), 6))])
            pc = rollout_oracle.normalize_data(pc, pstats).astype(np.float32)
            obs = o.push_obs(raw[t], point_cloud=pc)
            assert np.array_equal(obs["state"], g[f"{p}_in_state"][k])
            assert np.array_equal(obs["point_cloud"], g[f"{p}_in_pc"][k])
            steps.append(t)
            a = o.step(fake[k])
            k += 1
        else:
            a = o.step()
        assert np.array_equal(a, g[f"{p}_actions"][t]), t
    assert steps == list(g[f"{p}_infer_steps"])


def test_mt_act_oracle_hypothesis_reduces_to_act():
    """The MT-ACT branch of the oracle (architecture hypothesis, parity unpinned) with zero FiLM generators and no task
    token must be the ACT network: it reproduces the reference's DETRVAE golden. And the policy mirror carries the
    hypothesised extra keys next to all of ACT's."""
    from robomanipbaselines_b200.policy.mt_act import ACTPolicy as MtPolicy

    case = ACT_CASES["small"]
    gold = np.load(os.path.join(GOLD, "act_golden_small.npz"))
    pol = MtPolicy(policy_args(case), state_dim=case["state_dim"], action_dim=case["action_dim"])
    shapes = {k: tuple(v.shape) for k, v in pol.state_dict().items()}
    act_keys = set(ACTPolicy(policy_args(case), state_dim=case["state_dim"], action_dim=case["action_dim"]).state_dict())
    extra = sorted(set(shapes) - act_keys)
    assert extra == sorted(["model.proj_text_emb.weight", "model.proj_text_emb.bias"] +
                           [f"model.backbones.0.0.film_models.{i}.{w}" for i in (1, 2, 3) for w in ("weight", "bias")])
    assert shapes["model.backbones.0.0.film_models.2.weight"] == (2 * 2 * 256, 512)
    sd = synthetic_state_dict(shapes, seed=case["weight_seed"])
    for k in sd:
        if "film_models" in k:
            sd[k] = torch.zeros_like(sd[k])
    qpos, img_u8 = act_inputs(case)
    image = torch.from_numpy(np.concatenate([rollout_oracle.get_images(list(img_u8[b])) for b in range(case["batch"])]))
    te = torch.randn(1, 384)
    a = act_oracle.act_forward(sd, policy_args(case), torch.from_numpy(qpos), image, task_emb=te)
    assert np.abs(a.numpy() - gold["a_hat"]).max() <= 2e-5
    for k in sd:
        if "film_models" in k:
            sd[k] = torch.randn_like(sd[k]) * 0.1
    b = act_oracle.act_forward(sd, policy_args(case), torch.from_numpy(qpos), image, task_emb=te)
    assert (a - b).abs().max() > 1e-3                       # the FiLM term is live
    tok = dict(policy_args(case), task_token=True)
    pol3 = MtPolicy(tok, state_dim=case["state_dim"], action_dim=case["action_dim"])
    assert tuple(pol3.state_dict()["model.additional_pos_embed.weight"].shape) == (3, 512)
