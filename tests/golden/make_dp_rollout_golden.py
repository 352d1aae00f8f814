"""Golden vectors for the Diffusion Policy / DP3 ROLLOUT GLUE, produced by EXECUTING THE REFERENCE's own
RolloutDiffusionPolicy.infer_policy / get_state / get_images (policy/diffusion_policy/RolloutDiffusionPolicy.py:60-143)
and RolloutDiffusionPolicy3d.infer_policy / get_state / get_pointcloud
(policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:73-171) from /root/reference (build container only).

    python tests/golden/make_dp_rollout_golden.py

The reference methods are called unmodified on a stand-in `self` (they need no env: only the attributes they read).
`policy.predict_action` is a recording fake that returns seeded actions, so what is pinned is exactly the glue: the
observation ring buffers (fill on the first call, shift afterwards, pushed ONLY when the action FIFO is empty), the
cv2 resize + [-1, 1] scaling, the state normalisation, the float64 action FIFO and the denormalised actions.
For DP3, pytorch3d is absent: `downsample_pointcloud_fps` is replaced by a documented stand-in (first num_points rows),
which the replay in the tests uses too; back-projection / crop / normalisation are the reference's own code.
Writes tests/golden/dp_rollout_golden.npz.
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

T_STEPS = 14          # policy steps per episode (n_action_steps = 4 -> 4 inferences, the last FIFO partly consumed)
N_OBS, N_ACT = 2, 4
DS, DA = 5, 3
IMG_HW, IMG_SIZE = (24, 32), (16, 12)      # env image (H, W); model_meta_info image_size is (W, H) for cv2.resize
PC_HW, PC_SIZE, NUM_POINTS = (20, 24), (12, 10), 32


def _stats(rng, dim, kind):
    s = {"mean": rng.normal(size=dim), "std": rng.uniform(0.5, 2.0, size=dim), "min": rng.normal(size=dim),
         "range": rng.uniform(0.5, 2.0, size=dim)}
    if kind == "limits":
        s["norm_config"] = {"type": "limits", "out_min": -1.0, "out_max": 1.0}
    elif kind == "gaussian":
        s["norm_config"] = {"type": "gaussian"}
    return s


def _pack_stats(out, prefix, s):
    for k in ("mean", "std", "min", "range"):
        out[f"{prefix}_{k}"] = s[k]


def make_dp(out, norm):
    from robo_manip_baselines.policy.diffusion_policy.RolloutDiffusionPolicy import RolloutDiffusionPolicy
    from torchvision.transforms import v2

    rng = np.random.default_rng({"gaussian": 101, "limits": 102}[norm])
    sstats, astats = _stats(rng, DS, norm), _stats(rng, DA, norm)
    cams = ["front", "side"]
    raw_states = rng.normal(size=(T_STEPS, DS))
    images = rng.integers(0, 256, size=(T_STEPS, len(cams)) + IMG_HW + (3,), dtype=np.uint8)
    fake_actions = rng.normal(size=(T_STEPS, N_ACT, DA)).astype(np.float32)
    rec = {"state": [], "front": [], "side": [], "step": []}
    cur = {"t": 0}

    def predict_action(input_data):
        rec["step"].append(cur["t"])
        rec["state"].append(input_data["state"].numpy().copy())
        rec["front"].append(input_data["front_rgb_image"].numpy().copy())
        rec["side"].append(input_data["side_rgb_image"].numpy().copy())
        return {"action": torch.from_numpy(fake_actions[len(rec["step"]) - 1])[None]}

    me = types.SimpleNamespace(
        state_keys=["k0", "k1"], obs=None, device=torch.device("cpu"), camera_names=cams,
        model_meta_info={"state": sstats, "action": astats,
                         "data": {"n_obs_steps": N_OBS, "image_size": IMG_SIZE}},
        image_transforms=v2.Compose([v2.ToDtype(torch.float32, scale=True)]),      # RolloutBase.py:352-354
        policy=types.SimpleNamespace(predict_action=predict_action),
        state_buf=None, images_buf=None, policy_action_buf=None, policy_action_list=np.empty((0, DA)))
    me.motion_manager = types.SimpleNamespace(
        get_data=lambda key, obs: raw_states[cur["t"], :2] if key == "k0" else raw_states[cur["t"], 2:])
    me.get_state = lambda: RolloutDiffusionPolicy.get_state(me)
    me.get_images = lambda: RolloutDiffusionPolicy.get_images(me)
    for t in range(T_STEPS):
        cur["t"] = t
        me.info = {"rgb_images": {c: images[t, i] for i, c in enumerate(cams)}}
        RolloutDiffusionPolicy.infer_policy(me)
    p = f"dp_{norm}"
    out[f"{p}_raw_states"], out[f"{p}_images_u8"], out[f"{p}_fake_actions"] = raw_states, images, fake_actions
    out[f"{p}_actions"] = me.policy_action_list.copy()
    out[f"{p}_infer_steps"] = np.array(rec["step"])
    out[f"{p}_in_state"] = np.concatenate(rec["state"])          # (n_inf, To, Ds) float32
    out[f"{p}_in_front"] = np.concatenate(rec["front"])          # (n_inf, To, 3, h, w) float32 in [-1, 1]
    out[f"{p}_in_side"] = np.concatenate(rec["side"])
    _pack_stats(out, f"{p}_state", sstats)
    _pack_stats(out, f"{p}_action", astats)
    print(f"[golden] dp {norm}: {len(rec['step'])} inferences at steps {rec['step']}")


def make_dp3(out, norm):
    import importlib

    # the package __init__ re-exports the class under the module's name, so fetch the module object itself
    mod = importlib.import_module("robo_manip_baselines.policy.diffusion_policy_3d.RolloutDiffusionPolicy3d")
    mod = sys.modules["robo_manip_baselines.policy.diffusion_policy_3d.RolloutDiffusionPolicy3d"]
    RolloutDiffusionPolicy3d = mod.RolloutDiffusionPolicy3d
    # pytorch3d is absent: stand-in FPS (documented in the header); everything else on the path is reference code
    mod.downsample_pointcloud_fps = lambda pc, n: pc[:n] if len(pc) >= n else np.concatenate([pc, np.zeros((n - len(pc), pc.shape[1]))])
    if not hasattr(np, "concat"):
        np.concat = np.concatenate

    rng = np.random.default_rng({"gaussian": 201, "limits": 202}[norm])
    sstats, astats = _stats(rng, DS, norm), _stats(rng, DA, norm)
    pstats = _stats(rng, 6, norm)
    fovy = 52.0
    raw_states = rng.normal(size=(T_STEPS, DS))
    rgb = rng.integers(0, 256, size=(T_STEPS,) + PC_HW + (3,), dtype=np.uint8)
    depth = rng.uniform(0.4, 1.6, size=(T_STEPS,) + PC_HW).astype(np.float32)
    fake_actions = rng.normal(size=(T_STEPS, N_ACT, DA)).astype(np.float32)
    rec = {"state": [], "pc": [], "step": []}
    cur = {"t": 0}

    def predict_action(input_data):
        rec["step"].append(cur["t"])
        rec["state"].append(input_data["state"].numpy().copy())
        rec["pc"].append(input_data["point_cloud"].numpy().copy())
        return {"action": torch.from_numpy(fake_actions[len(rec["step"]) - 1])[None]}

    me = types.SimpleNamespace(
        state_keys=["k0", "k1"], obs=None, device=torch.device("cpu"), camera_names=["front"],
        model_meta_info={"state": sstats, "action": astats, "pointcloud": pstats,
                         "data": {"n_obs_steps": N_OBS, "image_size": PC_SIZE, "num_points": NUM_POINTS,
                                  "min_bound": [-0.6, -0.6, 0.3], "max_bound": [0.6, 0.6, 1.7]}},
        policy=types.SimpleNamespace(predict_action=predict_action),
        env=types.SimpleNamespace(unwrapped=types.SimpleNamespace(get_camera_fovy=lambda name: fovy)),
        state_buf=None, pointcloud_buf=None, policy_action_buf=None, policy_action_list=np.empty((0, DA)))
    me.motion_manager = types.SimpleNamespace(
        get_data=lambda key, obs: raw_states[cur["t"], :2] if key == "k0" else raw_states[cur["t"], 2:])
    me.get_state = lambda: RolloutDiffusionPolicy3d.get_state(me)
    me.get_pointcloud = lambda: RolloutDiffusionPolicy3d.get_pointcloud(me)
    for t in range(T_STEPS):
        cur["t"] = t
        me.info = {"rgb_images": {"front": rgb[t]}, "depth_images": {"front": depth[t]}}
        RolloutDiffusionPolicy3d.infer_policy(me)
    p = f"dp3_{norm}"
    out[f"{p}_raw_states"], out[f"{p}_rgb"], out[f"{p}_depth"], out[f"{p}_fake_actions"] = raw_states, rgb, depth, fake_actions
    out[f"{p}_fovy"] = np.array([fovy])
    out[f"{p}_actions"] = me.policy_action_list.copy()
    out[f"{p}_infer_steps"] = np.array(rec["step"])
    out[f"{p}_in_state"] = np.concatenate(rec["state"])
    out[f"{p}_in_pc"] = np.concatenate(rec["pc"])                # (n_inf, To, Np, 6) float32, normalised
    _pack_stats(out, f"{p}_state", sstats)
    _pack_stats(out, f"{p}_action", astats)
    _pack_stats(out, f"{p}_pointcloud", pstats)
    print(f"[golden] dp3 {norm}: {len(rec['step'])} inferences at steps {rec['step']}")


if __name__ == "__main__":
    ref_loader.install()
    out = {}
    for norm in ("gaussian", "limits"):
        make_dp(out, norm)
        make_dp3(out, norm)
    path = os.path.join(HERE, "dp_rollout_golden.npz")
    np.savez_compressed(path, **out)
    print(f"[golden] -> {path} ({os.path.getsize(path)} B)")
