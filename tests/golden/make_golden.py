"""Generate the committed golden vectors by EXECUTING THE REFERENCE (/root/reference, build container only).

    python tests/golden/make_golden.py

Writes tests/golden/act_golden_<case>.npz (reference DETRVAE outputs on synthetic weights / inputs) and
tests/golden/rollout_act_golden.npz (reference RolloutAct.infer_policy / RolloutBase.get_state / get_images /
normalize_data / denormalize_data outputs). Inputs and weights are pure functions of seeds
(robomanipbaselines_b200/common/synthetic.py), so only outputs and a few checksums are stored.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import ref_loader  # noqa: E402
from cases import ACT_CASES, act_inputs, policy_args  # noqa: E402
from robomanipbaselines_b200.common.synthetic import synthetic_state_dict  # noqa: E402


def make_act_case(name, case):
    torch.manual_seed(0)
    args = policy_args(case)
    ref = ref_loader.ReferenceACTPolicy(args, case["state_dim"], case["action_dim"])
    shapes = {k: tuple(v.shape) for k, v in ref.state_dict().items()}

    # our nn.Module shell must expose exactly the reference's keys and shapes (strict load_state_dict contract)
    from robomanipbaselines_b200.policy.act.policy import ACTPolicy

    shell = ACTPolicy(args, state_dim=case["state_dim"], action_dim=case["action_dim"])
    shell_shapes = {k: tuple(v.shape) for k, v in shell.state_dict().items()}
    assert shell_shapes == shapes, (set(shapes) ^ set(shell_shapes))

    sd = synthetic_state_dict(shapes, seed=case["weight_seed"])
    ref.load_state_dict(sd)
    qpos, image_u8 = act_inputs(case)
    from torchvision.transforms import v2
    image = v2.ToDtype(torch.float32, scale=True)(torch.from_numpy(np.moveaxis(image_u8, -1, -3).copy()))  # get_images (:479-490)
    taps = {}
    enc = ref.model.transformer.encoder
    h1 = enc.register_forward_pre_hook(lambda m, a: taps.__setitem__("src_tokens", a[0].detach().permute(1, 0, 2).contiguous()))
    h2 = enc.register_forward_hook(lambda m, a, o: taps.__setitem__("memory", o.detach().permute(1, 0, 2).contiguous()))
    # the hook RolloutAct.setup_policy registers (policy/act/RolloutAct.py:36-43), here keeping every episode's map
    attn = []
    hooks = [layer.self_attn.register_forward_hook(lambda _l, _i, _o: attn.append(_o[1].detach().numpy().copy()))
             for layer in enc.layers]
    a_hat = ref(torch.from_numpy(qpos), image)
    h1.remove(); h2.remove()
    for h in hooks:
        h.remove()
    out = {
        "a_hat": a_hat.numpy(),
        "weight_checksum": np.array([float(sum(v.double().abs().sum() for v in sd.values()))]),
        "input_checksum": np.array([float(np.abs(qpos).sum()), float(image_u8.astype(np.int64).sum())]),
        "n_keys": np.array([len(sd)]),
    }
    if case.get("store_taps"):
        out["src_tokens"] = taps["src_tokens"].numpy()
        out["memory"] = taps["memory"].numpy()
        out["attn_maps"] = np.stack(attn)            # (enc_layers, B, L, L) head-averaged attention weights
    else:  # large case: keep the fixture small, store per-token statistics only
        out["memory_token_mean"] = taps["memory"].mean(-1).numpy()
        out["src_token_mean"] = taps["src_tokens"].mean(-1).numpy()
    path = os.path.join(HERE, f"act_golden_{name}.npz")
    np.savez_compressed(path, **out)
    print(f"[golden] {name}: a_hat {a_hat.shape} |a|max {a_hat.abs().max():.3f} -> {path} ({os.path.getsize(path)} B)")


def make_rollout_golden():
    """Drive the reference's own RolloutAct.infer_policy / RolloutBase.get_state / get_images with a stand-in `self`."""
    ref_loader.install()
    from robo_manip_baselines.common.base.RolloutBase import RolloutBase
    from robo_manip_baselines.common.utils.DataUtils import denormalize_data, normalize_data
    from robo_manip_baselines.policy.act.RolloutAct import RolloutAct
    from torchvision.transforms import v2

    rng = np.random.default_rng(7)
    C, Da, Ds, T = 7, 3, 4, 25
    out = {}
    for norm in ("gaussian", "limits", "legacy"):
        stats = {"mean": rng.normal(size=Da), "std": rng.uniform(0.5, 2.0, size=Da), "min": rng.normal(size=Da),
                 "range": rng.uniform(0.5, 2.0, size=Da)}
        if norm == "limits":
            stats["norm_config"] = {"type": "limits", "out_min": -1.0, "out_max": 1.0}
        elif norm == "gaussian":
            stats["norm_config"] = {"type": "gaussian"}
        chunks = rng.normal(size=(T, C, Da)).astype(np.float32)
        for mode in ("ensemble", "fifo"):
            state = {"i": 0, "calls": 0}

            def policy(_s, _im):
                state["calls"] += 1
                return torch.from_numpy(chunks[state["i"]])[None]

            me = types.SimpleNamespace(
                args=types.SimpleNamespace(no_temp_ensem=(mode == "fifo")),
                get_state=lambda: None, get_images=lambda: None, policy=policy,
                model_meta_info={"data": {"chunk_size": C}, "action": stats}, action_dim=Da,
                policy_action_buf=[], policy_action_buf_history=[], policy_action_list=np.empty((0, Da)))
            for t in range(T):
                state["i"] = t
                RolloutAct.infer_policy(me)
            out[f"{norm}_{mode}_actions"] = me.policy_action_list.copy()
            out[f"{norm}_{mode}_calls"] = np.array([state["calls"]])
            if mode == "ensemble":
                out[f"{norm}_history"] = np.array(me.policy_action_buf_history)
        out[f"{norm}_chunks"] = chunks
        for k in ("mean", "std", "min", "range"):
            out[f"{norm}_stat_{k}"] = stats[k]
        x = rng.normal(size=(6, Da))
        out[f"{norm}_norm_in"] = x
        out[f"{norm}_normalize"] = normalize_data(x, stats)
        out[f"{norm}_denormalize"] = denormalize_data(x, stats)

    # RolloutBase.get_state (:463-477) / get_images (:479-490) on a stand-in self
    sstats = {"mean": rng.normal(size=Ds), "std": rng.uniform(0.5, 2.0, size=Ds), "norm_config": {"type": "gaussian"}}
    raw = rng.normal(size=Ds)
    imgs = {"front": rng.integers(0, 256, size=(8, 12, 3), dtype=np.uint8),
            "side": rng.integers(0, 256, size=(8, 12, 3), dtype=np.uint8)}
    me = types.SimpleNamespace(
        state_keys=["k0", "k1"], obs=None, device=torch.device("cpu"), model_meta_info={"state": sstats},
        motion_manager=types.SimpleNamespace(get_data=lambda key, obs: raw[:2] if key == "k0" else raw[2:]),
        info={"rgb_images": imgs}, camera_names=["front", "side"],
        image_transforms=v2.ToDtype(torch.float32, scale=True))
    out["get_state_raw"] = raw
    out["get_state_mean"], out["get_state_std"] = sstats["mean"], sstats["std"]
    out["get_state_out"] = RolloutBase.get_state(me).numpy()
    out["get_images_in"] = np.stack([imgs["front"], imgs["side"]])
    out["get_images_out"] = RolloutBase.get_images(me).numpy()
    path = os.path.join(HERE, "rollout_act_golden.npz")
    np.savez_compressed(path, **out)
    print(f"[golden] rollout -> {path} ({os.path.getsize(path)} B)")


if __name__ == "__main__":
    only = sys.argv[1:]                              # e.g. `make_golden.py small` regenerates one fixture
    if not only or "rollout" in only:
        make_rollout_golden()
    for name, case in ACT_CASES.items():
        if not only or name in only:
            make_act_case(name, case)
