"""The reference-shaped rollout classes (RolloutAct, RolloutDiffusionPolicy, RolloutDiffusionPolicy3d) driven the way
RolloutPhase.pre_update drives them (common/base/RolloutBase.py:56-62) with a stand-in env object, against the golden
vectors produced by executing the reference's own infer_policy (tests/golden/make_golden.py,
tests/golden/make_dp_rollout_golden.py)."""
import os
import sys
import types

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

from cases import ACT_CASES, act_inputs, policy_args  # noqa: E402
from robomanipbaselines_b200.common.synthetic import synthetic_norm_stats, synthetic_state_dict  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(HERE, "golden")


class _Env:
    """Stand-in for the gymnasium env: only what the rollout classes read."""

    def __init__(self, n_arm=2, n_grip=1, fovy=52.0):
        body = types.SimpleNamespace(arm_joint_idxes=list(range(n_arm)), gripper_joint_idxes=list(range(n_grip)), eef_idx=0)
        self.unwrapped = types.SimpleNamespace(body_config_list=[body], get_camera_fovy=lambda name: fovy)


class _MotionManager:
    def __init__(self):
        self.raw, self.commands = None, []

    def get_data(self, key, obs):
        return self.raw[:2] if key == "k0" else self.raw[2:]

    def set_command_data(self, key, command, is_skip=False):
        self.commands.append((key, np.array(command), is_skip))


class _FakeActEngine:
    """Returns the golden run's chunk of call k (the golden faked the network the same way)."""

    def __init__(self, chunks, device):
        self.chunks, self.k, self.device = chunks, 0, device

    def forward(self, images, qpos, out=None):
        out.copy_(torch.from_numpy(self.chunks[self.k][None]).to(self.device))
        self.k += 1
        return out


class _FakeActPolicy(torch.nn.Module):
    def __init__(self, chunks, device):
        super().__init__()
        self.eng = _FakeActEngine(chunks, device)
        enc = types.SimpleNamespace(layers=[])
        self.model = types.SimpleNamespace(transformer=types.SimpleNamespace(encoder=enc))

    def engine_for(self, *a):
        return self.eng


@pytest.mark.parametrize("norm", ["gaussian", "limits", "legacy"])
@pytest.mark.parametrize("mode", ["ensemble", "fifo"])
def test_rollout_act_class_matches_reference_action_list(cuda, norm, mode):
    """RolloutAct.infer_policy for 25 policy steps == the action list of the reference's RolloutAct.infer_policy
    (rollout_act_golden.npz: chunk size 7, so the history wraps three times), bit-for-bit, incl. policy call counts."""
    from robomanipbaselines_b200.policy.act import RolloutAct

    g = np.load(os.path.join(GOLD, "rollout_act_golden.npz"))
    stats = {k: g[f"{norm}_stat_{k}"] for k in ("mean", "std", "min", "range")}
    if norm == "limits":
        stats["norm_config"] = {"type": "limits", "out_min": -1.0, "out_max": 1.0}
    elif norm == "gaussian":
        stats["norm_config"] = {"type": "gaussian"}
    chunks = g[f"{norm}_chunks"]
    T, C, Da = chunks.shape
    stats["example"], stats["keys"] = np.zeros(Da), ["command_joint_pos"]
    sstats = dict(synthetic_norm_stats(4, 1), keys=["k0", "k1"])
    meta = {"data": {"chunk_size": C, "skip": 1}, "state": sstats, "action": stats, "image": {"camera_names": ["front"]},
            "policy": {"args": {}}}

    class R(RolloutAct):
        def make_policy(self):
            return _FakeActPolicy(chunks, cuda)

    mm = _MotionManager()
    r = R(_Env(), model_meta_info=meta, motion_manager=mm, device=str(cuda), no_temp_ensem=(mode == "fifo"),
          use_cuda_graph=False)      # the fake engine copies from the host: not capturable
    rng = np.random.default_rng(0)
    for t in range(T):
        r.policy.eng.k = t                      # the golden fed chunk t at step t (in FIFO mode only when it inferred)
        mm.raw = rng.normal(size=4)
        r.info = {"rgb_images": {"front": rng.integers(0, 256, size=(64, 96, 3), dtype=np.uint8)}}
        r.policy_step()
    assert np.array_equal(r.policy_action_list, g[f"{norm}_{mode}_actions"])
    assert r.rollout.n_inferences == int(g[f"{norm}_{mode}_calls"][0])
    # set_command_data handed the whole action to the one action key at every step (RolloutBase.py:496-509)
    assert len(mm.commands) == T and all(k == "command_joint_pos" for k, _, _ in mm.commands)
    assert np.array_equal(mm.commands[-1][1], r.policy_action)


def test_rollout_act_class_end_to_end_with_attention_maps(cuda):
    """The real engine behind RolloutAct on the 'small' golden case: chunk-consistent actions and the draw_plot
    attention images (head-averaged encoder self-attention of episode 0, correlation_mat[2:, 1]) against the maps the
    reference's own forward hook recorded (act_golden_small.npz)."""
    from oracle import rollout_oracle
    from robomanipbaselines_b200.policy.act import RolloutAct

    case = ACT_CASES["small"]
    gold = np.load(os.path.join(GOLD, "act_golden_small.npz"))
    Ds, Da, C = case["state_dim"], case["action_dim"], case["chunk_size"]
    cams = [f"cam{i}" for i in range(case["num_cams"])]
    # identity state normalisation so that the golden's (already normalised) qpos is what the network sees
    sstats = {"mean": np.zeros(Ds), "std": np.ones(Ds), "norm_config": {"type": "gaussian"}, "example": np.zeros(Ds),
              "keys": ["k0", "k1"]}
    astats = dict(synthetic_norm_stats(Da, 2), keys=["command_joint_pos"])
    meta = {"data": {"chunk_size": C, "skip": 1}, "state": sstats, "action": astats, "image": {"camera_names": cams},
            "policy": {"args": policy_args(case)}}
    mm = _MotionManager()
    r = RolloutAct(_Env(n_arm=Da - 1), model_meta_info=meta, motion_manager=mm, device=str(cuda), no_plot=False)
    sd = synthetic_state_dict({k: tuple(v.shape) for k, v in r.policy.state_dict().items()}, seed=case["weight_seed"])
    r.policy.load_state_dict(sd, strict=True)
    qpos, img = act_inputs(case)
    mm.raw = qpos[0].astype(np.float64)
    r.info = {"rgb_images": {c: img[0, i] for i, c in enumerate(cams)}}
    r.policy_step()
    # first step: the ensembled action is the chunk's first row, denormalised
    want = rollout_oracle.denormalize_data(gold["a_hat"][0, 0].astype(np.float64), astats)
    assert np.abs(r.policy_action - want).max() <= 1e-2 * np.abs(astats["std"]).max()
    maps = np.stack([l.self_attn.correlation_mat for l in r.policy.model.transformer.encoder.layers])
    err = np.abs(maps - gold["attn_maps"][:, 0]).max()
    print(f"[attention debug output] max-abs err vs the reference's hooked attention weights {err:.3e}")
    assert maps.shape == gold["attn_maps"][:, 0].shape and err <= 2e-3
    assert np.abs(maps.sum(-1) - 1).max() <= 1e-4
    plot = r.draw_plot()
    assert plot["attention"][0].shape == (case["img_h"] // 32, case["img_w"] // 32 * case["num_cams"])
    assert np.array_equal(plot["attention"][1], maps[1][2:, 1].reshape(plot["attention"][1].shape))
    assert r.calc_model_size() > 0


@pytest.mark.parametrize("norm", ["gaussian", "limits"])
def test_rollout_dp_class_matches_reference_golden(cuda, norm):
    """RolloutDiffusionPolicy.infer_policy from raw env frames vs the reference's own run: every recorded policy input
    (state + both cameras' ring buffers) and the action list bit-for-bit."""
    from test_rollout_glue_gpu import _RecordingPolicy, _stats
    from robomanipbaselines_b200.policy.diffusion_policy import RolloutDiffusionPolicy

    g = np.load(os.path.join(GOLD, "dp_rollout_golden.npz"))
    p = f"dp_{norm}"
    raw, imgs, fake = g[f"{p}_raw_states"], g[f"{p}_images_u8"], g[f"{p}_fake_actions"]
    To, Ds = g[f"{p}_in_state"].shape[1:]
    Na, Da = fake.shape[1:]
    h, w = g[f"{p}_in_front"].shape[-2:]
    meta = {"data": {"n_obs_steps": To, "n_action_steps": Na, "horizon": 8, "image_size": (w, h), "skip": 1},
            "state": dict(_stats(g, f"{p}_state", norm, Ds), keys=["k0", "k1"]),
            "action": dict(_stats(g, f"{p}_action", norm, Da), keys=["command_joint_pos"]),
            "image": {"camera_names": ["front", "side"]}, "policy": {}}
    pol = _RecordingPolicy(fake)

    class R(RolloutDiffusionPolicy):
        def setup_policy(self):
            self.policy, self.rollout = pol, None

    mm = _MotionManager()
    r = R(_Env(), model_meta_info=meta, motion_manager=mm, device=str(cuda))
    for t in range(raw.shape[0]):
        mm.raw = raw[t]
        r.info = {"rgb_images": {"front": imgs[t, 0], "side": imgs[t, 1]}}
        r.policy_step()
    assert np.array_equal(r.policy_action_list, g[f"{p}_actions"])
    assert len(pol.calls) == len(g[f"{p}_infer_steps"])
    for k, c in enumerate(pol.calls):
        assert np.array_equal(c["state"][0].cpu().numpy(), g[f"{p}_in_state"][k])
        assert np.array_equal(c["front_rgb_image"][0].cpu().numpy(), g[f"{p}_in_front"][k])
        assert np.array_equal(c["side_rgb_image"][0].cpu().numpy(), g[f"{p}_in_side"][k])


def test_rollout_dp3_class_runs_from_depth_frames(cuda):
    """RolloutDiffusionPolicy3d from raw rgb + depth frames with the real DP3 engine: one inference per n_action_steps,
    finite actions, the point-cloud ring buffer equals the device pipeline's own output for the same frames."""
    from robomanipbaselines_b200.common.pointcloud import Dp3ObservationPipeline
    from robomanipbaselines_b200.policy.diffusion_policy_3d import RolloutDiffusionPolicy3d

    Ds, Da, To, Na, Np = 4, 3, 2, 4, 64
    sched = dict(beta_end=0.02, beta_schedule="squaredcos_cap_v2", beta_start=0.0001, clip_sample=True,
                 num_train_timesteps=100, set_alpha_to_one=True, prediction_type="sample", steps_offset=0)
    args = dict(shape_meta={"obs": {"state": {"shape": [Ds]}, "point_cloud": {"shape": (Np, 6)}}, "action": {"shape": [Da]}},
                horizon=8, n_action_steps=Na, n_obs_steps=To, num_inference_steps=4, diffusion_step_embed_dim=64,
                down_dims=[64, 128], kernel_size=5, n_groups=8, pointcloud_encoder_cfg={"out_channels": 64},
                encoder_output_dim=64)
    meta = {"data": {"n_obs_steps": To, "n_action_steps": Na, "horizon": 8, "image_size": (24, 20), "num_points": Np,
                     "min_bound": [-0.6, -0.6, 0.3], "max_bound": [0.6, 0.6, 1.7], "use_pc_color": False, "skip": 1},
            "state": dict(synthetic_norm_stats(Ds, 1), keys=["k0", "k1"]),
            "action": dict(synthetic_norm_stats(Da, 2), keys=["command_joint_pos"]),
            "pointcloud": synthetic_norm_stats(6, 3), "image": {"camera_names": ["front"]},
            "policy": {"args": args, "noise_scheduler_args": sched, "use_ema": False}}
    mm = _MotionManager()
    r = RolloutDiffusionPolicy3d(_Env(), model_meta_info=meta, motion_manager=mm, device=str(cuda))
    r.policy.load_state_dict(synthetic_state_dict({k: tuple(v.shape) for k, v in r.policy.state_dict().items()}, seed=7))
    pipe = Dp3ObservationPipeline(1, meta, 52.0, device=str(cuda))
    rng = np.random.default_rng(4)
    for t in range(6):
        mm.raw = rng.normal(size=Ds)
        rgb = rng.integers(0, 256, size=(40, 48, 3), dtype=np.uint8)
        depth = rng.uniform(0.4, 1.6, size=(40, 48)).astype(np.float32)
        r.info = {"rgb_images": {"front": rgb}, "depth_images": {"front": depth}}
        r.policy_step()
        if t % Na == 0:
            want = pipe(torch.from_numpy(rgb[None]).to(cuda), torch.from_numpy(depth[None]).to(cuda))
            assert torch.equal(r.rollout.obs_buf["point_cloud"][0, -1], want[0])
    assert r.rollout.n_inferences == 2 and np.isfinite(r.policy_action_list).all()
    assert r.draw_plot()["pointcloud"].shape == (Np, 6)
