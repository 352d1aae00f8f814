"""CPU-side checks of the Diffusion Policy / DP3 host logic: scheduler step tables against the oracle's scalars, policy
mirrors' state-dict layout against the oracle modules (upstream names), rollout-glue oracle behaviour, and that the
product path refuses to run without a GPU."""
import numpy as np
import pytest
import torch

from oracle import diffusion_oracle as dor
from oracle import rollout_oracle as ror
from robomanipbaselines_b200.common.schedulers import DDIMScheduler, DDPMScheduler
from robomanipbaselines_b200.common.synthetic import synthetic_norm_stats

SCHED = dict(beta_end=0.02, beta_schedule="squaredcos_cap_v2", beta_start=0.0001, clip_sample=True,
             num_train_timesteps=100)


@pytest.mark.parametrize("schedule", ["squaredcos_cap_v2", "linear", "scaled_linear"])
def test_ddpm_step_table_matches_oracle(schedule):
    args = dict(SCHED, beta_schedule=schedule)
    mine, orc = DDPMScheduler(**args, prediction_type="epsilon", variance_type="fixed_small"), dor.DDPMSchedulerOracle(**args)
    steps = mine.step_table(100)
    orc.set_timesteps(100)
    assert [s.t for s in steps] == [int(t) for t in orc.timesteps] == list(range(99, -1, -1))
    g = torch.Generator().manual_seed(0)
    x, out, nz = torch.randn(5, 16, 7, generator=g), torch.randn(5, 16, 7, generator=g), torch.randn(5, 16, 7, generator=g)
    for s in steps:
        # the device kernel's formula, evaluated op by op in fp32 like unet_final_step_kernel / sched_step()
        x0 = torch.clamp((x - torch.tensor(s.s_beta) * out) / torch.tensor(s.s_alpha), -1, 1)
        prev = torch.tensor(s.c_x0) * x0 + torch.tensor(s.c_xt) * x
        if s.sigma > 0:
            prev = prev + torch.tensor(s.sigma) * nz
        assert torch.equal(prev, orc.step(out, s.t, x, nz)), s.t
    assert steps[-1].sigma == 0.0 and all(s.sigma > 0 for s in steps[:-1])


@pytest.mark.parametrize("n_steps", [10, 16, 100])
@pytest.mark.parametrize("derive", [False, True])
def test_ddim_step_table_matches_oracle(n_steps, derive):
    args = dict(SCHED, set_alpha_to_one=True, steps_offset=0, prediction_type="sample")
    mine, orc = DDIMScheduler(**args, derive_epsilon=derive), dor.DDIMSchedulerOracle(**args, derive_epsilon=derive)
    steps = mine.step_table(n_steps)
    orc.set_timesteps(n_steps)
    assert [s.t for s in steps] == [int(t) for t in orc.timesteps]
    g = torch.Generator().manual_seed(1)
    x, out = torch.randn(3, 16, 7, generator=g), torch.randn(3, 16, 7, generator=g)
    for s in steps:
        x0 = torch.clamp(out, -1, 1)
        d = (x - torch.tensor(s.s_alpha) * x0) / torch.tensor(s.s_beta) if derive else out
        prev = torch.tensor(s.c_x0) * x0 + torch.tensor(s.c_xt) * d
        assert torch.equal(prev, orc.step(out, s.t, x)), s.t


def test_scheduler_rejects_unknown_options():
    with pytest.raises(NotImplementedError):
        DDPMScheduler(variance_type="learned")
    with pytest.raises(ValueError):
        DDIMScheduler(prediction_type="v_prediction")
    with pytest.raises(NotImplementedError):
        DDPMScheduler(beta_schedule="sigmoid")


def test_policy_mirrors_have_upstream_state_dict_layout():
    from robomanipbaselines_b200.policy.diffusion_policy import DiffusionUnetHybridImagePolicy
    from robomanipbaselines_b200.policy.diffusion_policy_3d import DP3

    dp3 = DP3(noise_scheduler=DDIMScheduler(**SCHED, prediction_type="sample"),
              shape_meta={"obs": {"state": {"shape": [7], "type": "low_dim"},
                                  "point_cloud": {"shape": (1024, 6), "type": "point_cloud"}}, "action": {"shape": [7]}},
              horizon=16, n_action_steps=8, n_obs_steps=2, num_inference_steps=10, obs_as_global_cond=True,
              diffusion_step_embed_dim=128, down_dims=[512, 1024, 2048], kernel_size=5, n_groups=8,
              pointcloud_encoder_cfg={"in_channels": 6, "out_channels": 64, "use_layernorm": True,
                                      "final_norm": "layernorm", "normal_channel": False},
              use_pc_color=False, encoder_output_dim=64)
    want = {"model." + k: tuple(v.shape) for k, v in dor.ConditionalUnet1D(7, 256).state_dict().items()}
    want.update({"obs_encoder." + k: tuple(v.shape) for k, v in dor.DP3Encoder(7, 64).state_dict().items()})
    assert {k: tuple(v.shape) for k, v in dp3.state_dict().items()} == want
    dp = DiffusionUnetHybridImagePolicy(
        noise_scheduler=DDPMScheduler(**SCHED),
        shape_meta={"obs": {"state": {"shape": [7], "type": "low_dim"},
                            "front_rgb_image": {"shape": [3, 240, 320], "type": "rgb"},
                            "side_rgb_image": {"shape": [3, 240, 320], "type": "rgb"}}, "action": {"shape": [7]}},
        horizon=16, n_action_steps=8, n_obs_steps=2, num_inference_steps=100, obs_as_global_cond=True,
        crop_shape=(216, 288), diffusion_step_embed_dim=128, down_dims=[512, 1024, 2048], kernel_size=5, n_groups=8,
        cond_predict_scale=True, obs_encoder_group_norm=True, eval_fixed_crop=True)
    assert dp.obs_feature_dim == 2 * 64 + 7
    want = {"model." + k: tuple(v.shape) for k, v in dor.ConditionalUnet1D(7, 2 * 135).state_dict().items()}
    enc = {k: tuple(v.shape) for k, v in dor.DpImageEncoder(216, 288).state_dict().items()}
    for cam in ("front_rgb_image", "side_rgb_image"):
        want.update({f"obs_encoder.obs_nets.{cam}.{k}": v for k, v in enc.items()})
    assert {k: tuple(v.shape) for k, v in dp.state_dict().items()} == want
    # A checkpoint written by the upstream classes also carries entries with no arithmetic behind them (public
    # upstream layout: ModuleAttrMixin._dummy_variable on the policy and its mask generator, LinearNormalizer params,
    # robomimic SpatialSoftmax's persistent buffers, twice per camera). RolloutBase.load_ckpt loads with strict=True
    # (common/base/RolloutBase.py:376-385): those keys must not make the load fail, a real missing key still must.
    sd = {k: torch.randn_like(v) for k, v in dp.state_dict().items()}
    sd.update({"_dummy_variable": torch.zeros(0), "mask_generator._dummy_variable": torch.zeros(0),
               "normalizer.params_dict.action.scale": torch.ones(7), "normalizer.params_dict.action.offset": torch.zeros(7)})
    for cam in ("front_rgb_image", "side_rgb_image"):
        for pool in ("pool", "nets.1"):
            sd[f"obs_encoder.obs_nets.{cam}.{pool}.temperature"] = torch.ones(1)
            sd[f"obs_encoder.obs_nets.{cam}.{pool}.pos_x"] = torch.zeros(1, 63)
            sd[f"obs_encoder.obs_nets.{cam}.{pool}.pos_y"] = torch.zeros(1, 63)
    dp.load_state_dict(sd, strict=True)
    assert torch.equal(dp.state_dict()["model.final_conv.1.bias"], sd["model.final_conv.1.bias"])
    bad = dict(sd)
    bad["obs_encoder.obs_nets.front_rgb_image.pool.temperature"] = torch.full((1,), 0.5)
    with pytest.raises(NotImplementedError):
        dp.load_state_dict(bad, strict=True)
    missing = {k: v for k, v in sd.items() if k != "model.final_conv.1.bias"}
    with pytest.raises(RuntimeError):
        dp.load_state_dict(missing, strict=True)
    sd3 = {k: torch.randn_like(v) for k, v in dp3.state_dict().items()}
    sd3.update({"_dummy_variable": torch.zeros(0), "mask_generator._dummy_variable": torch.zeros(0),
                "normalizer.params_dict.point_cloud.scale": torch.ones(6)})
    dp3.load_state_dict(sd3, strict=True)
    with pytest.raises(RuntimeError):          # no CPU fallback
        dp3.predict_action({"state": torch.zeros(1, 2, 7), "point_cloud": torch.zeros(1, 2, 1024, 6)})
    with pytest.raises(NotImplementedError):
        DP3(noise_scheduler=DDIMScheduler(**SCHED), shape_meta={"obs": {"point_cloud": {"shape": (8, 3)}}, "action": {"shape": [7]}},
            horizon=16, n_action_steps=8, n_obs_steps=2, obs_as_global_cond=False)


def test_diffusion_rollout_oracle_glue():
    """Obs buffers advance only at inference; the FIFO pops n_action_steps float64 actions, each denormalised."""
    st, ac = synthetic_norm_stats(7, 1, "limits"), synthetic_norm_stats(7, 2)
    o = ror.DiffusionRolloutOracle(2, st, ac)
    rng = np.random.default_rng(0)
    chunk0, chunk1 = rng.standard_normal((8, 7)).astype(np.float32), rng.standard_normal((8, 7)).astype(np.float32)
    raws = rng.standard_normal((17, 7))
    hist = []
    for i in range(17):
        if o.needs_inference():
            obs = o.push_obs(raws[i])
            hist.append(i)
            a = o.step(chunk0 if len(hist) == 1 else chunk1)
        else:
            a = o.step()
        src = (chunk0 if i < 8 else chunk1)[i % 8].astype(np.float64)
        assert np.array_equal(a, ror.denormalize_data(src, ac))
    assert hist == [0, 8, 16]
    want = np.stack([ror.normalize_data(raws[8], st), ror.normalize_data(raws[16], st)]).astype(np.float32)
    assert np.array_equal(obs["state"], want)
