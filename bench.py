#!/usr/bin/env python
"""bench.py -- batched ACT rollout inference throughput on B200 (BASELINE.json metric, configs[1]).

  python bench.py --gpus N --steps K --warmup W            # this engine (under torchrun for N > 1)
  python bench.py --impl reference --steps K --warmup W    # CPU arm: the reference's own code on the host cores

A "step" is one BatchedRollout.infer() over B = 256 episodes per GPU (2 RGB cameras 480x640 uint8 + 7-d joint state,
chunk 100, temporal ensembling): state normalise -> ACT forward -> ensembling ring buffer -> denormalise, i.e. 256
policy steps of RolloutAct.infer_policy. Prints ONE JSON line on rank 0.

Besides the contract's keys the line carries `other_configs` (the other BASELINE.json configs at their per-GPU share,
under torchrun too, each with its own action all-gather): Diffusion Policy C3 (DDPM-100 and DDIM-16, through
predict_action and through the batched rollout from uint8 frames), MT-ACT C4 (2048 episodes over the N GPUs), DP3 C5
(512 = 4096 / 8 per GPU, DDIM-10 / -16, through predict_action and through the batched rollout from depth frames).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = ("ACT batched rollout inference (ResNet-18, 4 enc / 7 dec, hidden 512, ff 3200, chunk 100, temporal "
            "ensembling), 256 parallel episodes per GPU, 2 RGB cameras 480x640 uint8 + 7-d joint state "
            "(BASELINE.json configs[1])")
POLICY_ARGS = {"lr": 1e-5, "num_queries": 100, "kl_weight": 10, "hidden_dim": 512, "dim_feedforward": 3200,
               "lr_backbone": 1e-5, "backbone": "resnet18", "enc_layers": 4, "dec_layers": 7, "nheads": 8,
               "camera_names": ["front", "side"]}
IMG_H, IMG_W, NCAM, DS, DA, CHUNK = 480, 640, 2, 7, 7, 100
METRIC, UNIT = "policy steps/sec (batched rollout)", "policy_steps/s"
CPU_SAMPLE = 4          # episodes per CPU-arm step (the same bounded sample in --impl reference and in cpu_baseline)
SCHED = dict(beta_end=0.02, beta_schedule="squaredcos_cap_v2", beta_start=0.0001, clip_sample=True, num_train_timesteps=100)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"bf16_tflops": p["bf16_tflops"], "bf16_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "hbm_gbs": p["hbm_gbs"], "source": "measured (MEASURED_PEAKS.json)"}
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0,
            "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed regions run."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.lines, self.proc, self.thread = [], None, None
        try:
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            ident = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
        except Exception:
            ident = str(device_index)
        self.cmd = ["nvidia-smi", "-i", ident, f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200"]

    def start(self):
        try:
            self.proc = subprocess.Popen(self.cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True)
        self.thread.start()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); smax.append(float(parts[1])); power.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v == "Active":
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def act_weights(policy):
    """Random-init weights of the architecture (BASELINE.json): the distributions of a freshly constructed reference
    ACTPolicy (robomanipbaselines_b200/common/synthetic.py::default_init_state_dict), seed 0."""
    from robomanipbaselines_b200.common.synthetic import default_init_state_dict

    return default_init_state_dict({k: tuple(v.shape) for k, v in policy.state_dict().items()}, seed=0)


def act_meta(chunk=CHUNK):
    from robomanipbaselines_b200.common.synthetic import synthetic_norm_stats

    return {"data": {"chunk_size": chunk}, "state": synthetic_norm_stats(DS, 1), "action": synthetic_norm_stats(DA, 2)}


def build_rollout(device, batch):
    from robomanipbaselines_b200.common.batched_rollout import BatchedRollout
    from robomanipbaselines_b200.policy.act import ACTPolicy

    pol = ACTPolicy(POLICY_ARGS, state_dim=DS, action_dim=DA, max_batch=batch)
    pol.load_state_dict(act_weights(pol), strict=True)
    meta = act_meta()
    return pol, meta, BatchedRollout(pol, batch, meta, device=str(device))


class Timer:
    """K timed iterations between barriers + device synchronisation, CUDA events, max over ranks."""

    def __init__(self, device, dist):
        self.device, self.dist = device, dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def run(self, fn, iters, warmup):
        for i in range(warmup):
            fn(i)
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
        self.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=self.device, dtype=torch.float64)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item()) / iters


def make_gather(world, per_gpu, da, device):
    from robomanipbaselines_b200.common.sharding import AsyncEpisodeGather

    return AsyncEpisodeGather(world * per_gpu, (per_gpu, da), torch.float64, device, depth=2)


def bench_diffusion(device, timer, world, reps=2):
    """BASELINE.json configs[2] and the per-GPU share of configs[4]: one predict_action = obs encoder + the whole
    scheduler loop with synthetic observations resident in HBM, AND the batched rollout driven from env-layout frames
    (uint8 RGB / float32 depth) for n_action_steps policy steps = one inference + the FIFO pops. Each timed call ends
    with the all-gather of the (per-GPU batch, Da) actions when world > 1. policy steps/s = 8 x inferences/s."""
    from robomanipbaselines_b200.common.batched_diffusion_rollout import (BatchedDiffusionRollout, Dp3Frontend,
                                                                          DpImageFrontend)
    from robomanipbaselines_b200.common.schedulers import DDIMScheduler, DDPMScheduler
    from robomanipbaselines_b200.common.synthetic import synthetic_norm_stats, synthetic_state_dict
    from robomanipbaselines_b200.policy.diffusion_policy import DiffusionUnetHybridImagePolicy
    from robomanipbaselines_b200.policy.diffusion_policy.policy import visual_core_param_shapes
    from robomanipbaselines_b200.policy.diffusion_policy_3d import DP3

    out = {}
    gen = torch.Generator(device=device).manual_seed(3)
    To, Na = 2, 8
    stats = {"state": synthetic_norm_stats(DS, 1), "action": synthetic_norm_stats(DA, 2)}

    # ---- configs[2]: Diffusion Policy, batch 1024 per GPU, one 240x320 camera (crop 216x288) + 7-d state
    B = 1024
    for name, sched_cls, n_steps, extra in (("ddpm100", DDPMScheduler, 100, dict(prediction_type="epsilon", variance_type="fixed_small")),
                                            ("ddim16", DDIMScheduler, 16, dict(set_alpha_to_one=True, prediction_type="epsilon", steps_offset=0))):
        pol = DiffusionUnetHybridImagePolicy(
            noise_scheduler=sched_cls(**SCHED, **extra),
            shape_meta={"obs": {"state": {"shape": [DS], "type": "low_dim"},
                                "front_rgb_image": {"shape": [3, 240, 320], "type": "rgb"}}, "action": {"shape": [DA]}},
            horizon=16, n_action_steps=Na, n_obs_steps=To, num_inference_steps=n_steps, obs_as_global_cond=True,
            crop_shape=(216, 288), diffusion_step_embed_dim=128, down_dims=[512, 1024, 2048], kernel_size=5, n_groups=8,
            cond_predict_scale=True, obs_encoder_group_norm=True, eval_fixed_crop=True, max_batch=B)
        shapes = {k: tuple(v.shape) for k, v in pol.state_dict().items()}
        sd = synthetic_state_dict({k: v for k, v in shapes.items() if k.startswith("model.")}, seed=0)
        vc_shapes = visual_core_param_shapes()
        vc = synthetic_state_dict({k: v for k, v in vc_shapes.items() if not k.startswith(("nets.0.", "nets.1."))}, seed=0)
        for k in vc_shapes:
            if k.startswith("nets.0."):
                vc[k] = vc["backbone." + k[7:]]
            elif k.startswith("nets.1."):
                vc[k] = vc["pool." + k[7:]]
        sd.update({"obs_encoder.obs_nets.front_rgb_image." + k: v for k, v in vc.items()})
        pol.load_state_dict(sd, strict=True)
        obs = {"state": torch.randn(B, To, DS, device=device, generator=gen),
               "front_rgb_image": torch.rand(B, To, 3, 240, 320, device=device, generator=gen) * 2 - 1}
        gat = make_gather(world, B, DA, device)

        def predict(i):
            a = pol.predict_action(obs)["action"]
            gat.submit(a[:, 0].to(torch.float64))
            gat.result()

        ms_enc = timer.run(lambda i: pol.encode_obs(obs), reps, 1)
        ms = timer.run(predict, reps, 1)
        unet = pol.unet_for(B, device)
        enc = pol.encoder_for("front_rgb_image", B * To, device)
        flops = unet.step_flops(B) * n_steps + enc.flops(B * To)
        entry = {
            "workload": f"DiffusionUnetHybridImagePolicy.predict_action, batch {B} per GPU, horizon 16, {name}, ResNet-18 "
                        "(GroupNorm) obs encoder on one 240x320 camera (crop 216x288) x 2 obs steps + 7-d state",
            "ms_per_inference_batch": ms, "ms_obs_encoder": ms_enc, "inferences_per_s": world * B / ms * 1e3,
            "policy_steps_per_s": world * Na * B / ms * 1e3, "model_tflops_per_gpu": flops / ms / 1e9,
            "gpu_launches": unet.launches_per_step * n_steps + 3 + 70}
        # the batched rollout from env frames: Na policy steps = 1 inference (resize + ring push + predict) + Na FIFO pops
        meta = dict(stats, data={"n_obs_steps": To, "n_action_steps": Na, "image_size": (320, 240)})
        br = BatchedDiffusionRollout(pol, B, meta, device=str(device), obs_keys=("front_rgb_image",),
                                     frontend=DpImageFrontend(["front"], (320, 240)))
        frames = torch.randint(0, 256, (B, 1, IMG_H, IMG_W, 3), device=device, dtype=torch.uint8, generator=gen)
        raw = torch.randn(B, DS, device=device, dtype=torch.float64, generator=gen)
        br.reset()

        def rollout_cycle(i):
            for _ in range(Na):
                gat.submit(br.infer_env(raw, rgb_u8=frames))
                gat.result()

        ms_cycle = timer.run(rollout_cycle, reps, 1)
        entry["rollout"] = {"what": f"BatchedDiffusionRollout.infer_env from uint8 480x640 frames, {Na} policy steps = one "
                                    "inference (cv2-exact resize, ring push, predict_action) + FIFO pops + denormalise"
                                    + (", all-gather of the actions every step" if world > 1 else ""),
                            "ms_per_cycle": ms_cycle, "policy_steps_per_s": world * Na * B / ms_cycle * 1e3}
        out[f"diffusion_policy_c3_{name}"] = entry
        del pol, obs, unet, enc, br, frames
        torch.cuda.empty_cache()

    # ---- configs[4] per-GPU share: DP3, batch 512 (= 4096 / 8), 1024-point clouds, DDIM 10 and 16
    B = 512
    for n_steps in (10, 16):
        pol = DP3(noise_scheduler=DDIMScheduler(**SCHED, set_alpha_to_one=True, prediction_type="sample", steps_offset=0),
                  shape_meta={"obs": {"state": {"shape": [DS]}, "point_cloud": {"shape": (1024, 6)}}, "action": {"shape": [DA]}},
                  horizon=16, n_action_steps=Na, n_obs_steps=To, num_inference_steps=n_steps, diffusion_step_embed_dim=128,
                  down_dims=[512, 1024, 2048], kernel_size=5, n_groups=8, pointcloud_encoder_cfg={"out_channels": 64},
                  use_pc_color=False, encoder_output_dim=64, max_batch=B)
        pol.load_state_dict(synthetic_state_dict({k: tuple(v.shape) for k, v in pol.state_dict().items()}, seed=0))
        obs = {"state": torch.randn(B, To, DS, device=device, generator=gen),
               "point_cloud": torch.rand(B, To, 1024, 6, device=device, generator=gen) * 2 - 1}
        gat = make_gather(world, B, DA, device)

        def predict3(i):
            a = pol.predict_action(obs)["action"]
            gat.submit(a[:, 0].to(torch.float64))
            gat.result()

        ms_enc = timer.run(lambda i: pol.encode_obs(obs), reps, 1)
        ms = timer.run(predict3, reps, 1)
        unet = pol.unet_for(B, device)
        entry = {
            "workload": f"DP3.predict_action, batch 512 per GPU (4096 / 8), 2 x 1024-point clouds + 7-d state, DDIM {n_steps} steps",
            "ms_per_inference_batch": ms, "ms_obs_encoder": ms_enc, "inferences_per_s": world * B / ms * 1e3,
            "policy_steps_per_s": world * Na * B / ms * 1e3, "model_tflops_per_gpu": unet.step_flops(B) * n_steps / ms / 1e9,
            "gpu_launches": unet.launches_per_step * n_steps + 3 + 7}
        if n_steps == 10:
            meta = dict(stats, pointcloud=synthetic_norm_stats(6, 3),
                        data={"n_obs_steps": To, "n_action_steps": Na, "image_size": (84, 84), "num_points": 1024,
                              "min_bound": [-1.5, -1.5, 0.2], "max_bound": [1.5, 1.5, 2.5]})
            br = BatchedDiffusionRollout(pol, B, meta, device=str(device), obs_keys=("point_cloud",),
                                         frontend=Dp3Frontend(B, meta, 45.0, device=str(device)))
            rgb = torch.randint(0, 256, (B, IMG_H, IMG_W, 3), device=device, dtype=torch.uint8, generator=gen)
            depth = torch.rand(B, IMG_H, IMG_W, device=device, generator=gen) * 1.5 + 0.5
            raw = torch.randn(B, DS, device=device, dtype=torch.float64, generator=gen)
            br.reset()

            def rollout_cycle3(i):
                for _ in range(Na):
                    gat.submit(br.infer_env(raw, rgb_u8=rgb, depth=depth))
                    gat.result()

            ms_cycle = timer.run(rollout_cycle3, reps, 1)
            entry["rollout"] = {"what": f"BatchedDiffusionRollout.infer_env from 480x640 rgb + depth frames, {Na} policy steps "
                                        "= one inference (resize, back-projection, crop, FPS to 1024 points, ring push, "
                                        "predict_action) + FIFO pops + denormalise",
                                "ms_per_cycle": ms_cycle, "policy_steps_per_s": world * Na * B / ms_cycle * 1e3}
            del br, rgb, depth
        out[f"dp3_c5_ddim{n_steps}"] = entry
        del pol, obs, unet
        torch.cuda.empty_cache()
    return out


def bench_mt_act(device, timer, world, iters=3):
    """BASELINE.json configs[3]: MT-ACT (architecture hypothesis, parity unpinned), 2048 episodes over the N GPUs, chunk
    size 20 (policy/mt_act/TrainMtAct.py:36-38), one 384-d task embedding per episode resident in the engine."""
    from robomanipbaselines_b200.common.batched_rollout import BatchedRollout
    from robomanipbaselines_b200.policy.mt_act import ACTPolicy as MtActPolicy

    total = 2048
    B = total // world
    args = dict(POLICY_ARGS, num_queries=20)
    pol = MtActPolicy(args, state_dim=DS, action_dim=DA, max_batch=B)
    pol.load_state_dict(act_weights(pol), strict=True)
    br = BatchedRollout(pol, B, act_meta(20), device=str(device))
    gen = torch.Generator(device=device).manual_seed(5)
    te = torch.randn(B, 384, device=device, generator=gen)
    br.set_task_embedding(te / te.norm(dim=1, keepdim=True))
    imgs = torch.randint(0, 256, (B, NCAM, IMG_H, IMG_W, 3), generator=gen, device=device, dtype=torch.uint8)
    state = torch.randn(B, DS, generator=gen, device=device, dtype=torch.float64)
    gat = make_gather(world, B, DA, device)
    br.reset()

    def step(i):
        gat.submit(br.infer(imgs, state))
        gat.result()

    ms = timer.run(step, iters, 2)
    eng = pol.engine_for(B, IMG_H, IMG_W, device)
    res = {"workload": f"MT-ACT batched rollout inference, {total} episodes over {world} GPU(s) = {B} per GPU, 2 cameras "
                       "480x640 + 7-d state + 384-d task embedding, chunk 20, temporal ensembling",
           "ms_per_step": ms, "policy_steps_per_s": total / ms * 1e3, "model_tflops_per_gpu": eng.flops(B) / ms / 1e9,
           "scaling": "strong", "parity": "unpinned (architecture hypothesis, robomanipbaselines_b200/policy/mt_act/policy.py)"}
    del br, imgs
    pol._invalidate()
    torch.cuda.empty_cache()
    return res


def bench_micro(device):
    """Stand-alone HBM micro-benchmarks of the two elementwise kernels that are too small to time inside a step
    (SURVEY 8d): the scheduler step over 2^26 elements and the ensembling ring buffer over 65 536 episodes."""
    import ctypes

    from robomanipbaselines_b200 import _lib
    from robomanipbaselines_b200._lib import check, cur_stream, ptr
    from robomanipbaselines_b200.common.ensemble import make_weight_table
    from robomanipbaselines_b200.common.schedulers import DDPMScheduler

    lib = _lib.load()
    peak = measured_peaks()["hbm_gbs"]
    out = {}

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    n = 1 << 26
    x, o, z, y = (torch.randn(n, device=device) for _ in range(4))
    st = DDPMScheduler(num_train_timesteps=100, beta_schedule="squaredcos_cap_v2").step_table(100)[40]
    ms = timed(lambda: check(lib.rmb_scheduler_step(ptr(x), ptr(o), ptr(z), ctypes.byref(st), ptr(y), n, cur_stream()), "rmb_scheduler_step"), 10)
    gbs = 16.0 * n / ms / 1e6
    out["scheduler_step"] = {"kernel": "sched_step_kernel", "elements": n, "algorithmic_bytes": 16 * n, "ms": ms,
                             "achieved_gbs": gbs, "peak_gbs": peak, "frac": gbs / peak,
                             "what": "DDPM step t=59 (x, model output, noise in; x' out), 4 x 256 MB arrays >> L2"}
    del x, o, z, y
    B, C, Da = 65536, CHUNK, DA
    ring = torch.zeros(B, C, C, Da, device=device)
    chunks = torch.randn(B, C, Da, device=device)
    step = torch.zeros(B, device=device, dtype=torch.int32)
    wtab = torch.from_numpy(make_weight_table(C)).to(device)
    p0, p1 = torch.ones(Da, device=device, dtype=torch.float64), torch.zeros(Da, device=device, dtype=torch.float64)
    act = torch.empty(B, Da, device=device, dtype=torch.float64)

    def ens():
        check(lib.rmb_ensemble_step(ptr(ring), ptr(chunks), ptr(step), None, ptr(wtab), 0, ptr(p0), ptr(p1), 0.0,
                                    ptr(act), None, None, B, C, Da, cur_stream()), "rmb_ensemble_step")

    for _ in range(C):                       # fill the ring: every later call gathers the full 100-entry diagonal
        ens()
    ms = timed(ens, 10)
    per_ep = 3 * C * Da * 4 + Da * 8         # chunk in + ring slot write + diagonal gather + action out
    gbs = per_ep * B / ms / 1e6
    out["ensemble_step"] = {"kernel": "ensemble_step_kernel", "episodes": B, "algorithmic_bytes": per_ep * B, "ms": ms,
                            "achieved_gbs": gbs, "peak_gbs": peak, "frac": gbs / peak,
                            "what": "temporal ensembling with a full 100-chunk history per episode, 18 GB ring buffer"}
    del ring, chunks
    torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------------------------------------------- CPU arm
def cpu_arm(sample, iters, warmup):
    """The reference path on the host cores for `sample` episodes per step: the reference's OWN ACTPolicy / DETRVAE and
    RolloutAct.infer_policy (oracle/_ref, made by oracle/make_ref.py from the reference mount; kind "reference") or, when
    those sources are not there, the oracle port (kind "port"). Returns (seconds per step list, cores, kind, description)."""
    from robomanipbaselines_b200.common.synthetic import synthetic_act_observations
    from robomanipbaselines_b200.policy.act import ACTPolicy

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    shell = ACTPolicy(POLICY_ARGS, state_dim=DS, action_dim=DA)
    sd = act_weights(shell)
    meta = act_meta()
    images, state = synthetic_act_observations(sample, NCAM, IMG_H, IMG_W, DS, seed=1)
    import contextlib

    from oracle import reference_runner

    if reference_runner.reference_root() is not None:
        with contextlib.redirect_stdout(sys.stderr):      # the reference prints its parameter count while it is built
            runner = reference_runner.ReferenceActRollout(POLICY_ARGS, DS, DA, sd, meta, POLICY_ARGS["camera_names"], sample)
        kind = "reference"
        what = ("the reference's own ACTPolicy/DETRVAE (third_party/act) + RolloutAct.infer_policy, unmodified sources from "
                f"{os.path.relpath(runner.root, ROOT) if runner.root.startswith(ROOT) else runner.root}, fp32 eager PyTorch")
        step = lambda: runner.step(images, state)
    else:
        from oracle import act_oracle, rollout_oracle

        oracles = [rollout_oracle.ActRolloutOracle(CHUNK, DA, meta["action"]) for _ in range(sample)]
        kind = "port"
        what = "oracle port (oracle/act_oracle.py + oracle/rollout_oracle.py), fp32 eager PyTorch, all 7 decoder layers"

        def step():
            qpos = np.concatenate([rollout_oracle.get_state(state[b], meta["state"]) for b in range(sample)])
            image = np.concatenate([rollout_oracle.get_images(list(images[b])) for b in range(sample)])
            chunks = act_oracle.act_forward(sd, POLICY_ARGS, torch.from_numpy(qpos), torch.from_numpy(image), live_only=False).numpy()
            return np.stack([oracles[b].step(chunks[b]) for b in range(sample)])

    times = []
    for i in range(warmup + iters):
        t0 = time.perf_counter()
        step()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return times, cores, kind, what


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the box's host cores, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    times, cores, kind, what = cpu_arm(CPU_SAMPLE, args.steps, max(min(args.warmup, 2), 1))
    ms = 1e3 * float(np.mean(times))
    value = CPU_SAMPLE / (ms / 1e3)
    sample = (f"{args.steps} steps x {CPU_SAMPLE} episodes (a bounded sample of the 256-episode batch; the reference steps "
              f"episodes one at a time at batch 1), {what}, {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": f"{CPU_SAMPLE} episodes per step",
                   "weights": "random-init (default_init_state_dict, seed 0)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------- roofline
def roofline_block(recs_runs, ms_per_step, peaks, batch):
    """recs_runs: per-launch CUDA-event records of several back-to-back profiled forwards taken right after the timed
    loop (GPU at its sustained clocks). Figures per kernel: `isolated` = algorithmic work / the launches' own event
    times; `in_step` = the same scaled by (sum of all launches) / ms_per_step, i.e. what the kernel delivers inside the
    timed step, where clocks sag under the power cap and launches queue back to back."""
    n = len(recs_runs)
    by_kernel = {}
    for recs in recs_runs:
        for r in recs:
            k = by_kernel.setdefault(r["kernel"], {"ms": 0.0, "flops": 0.0, "bytes": 0.0, "launches": 0})
            k["ms"] += r["ms"] / n; k["flops"] += r["flops"] / n; k["bytes"] += r["bytes"] / n; k["launches"] += 1
    for k in by_kernel.values():
        k["launches"] //= n
    prof_ms = sum(k["ms"] for k in by_kernel.values())
    stretch = ms_per_step / prof_ms                      # > 1: the timed step is slower than the sum of its launches
    tpk, hpk = peaks["bf16_tflops_sustained"], peaks["hbm_gbs"]

    def entry(name, v):
        tf = v["flops"] / (v["ms"] * 1e-3) / 1e12 if v["ms"] > 0 else 0.0
        gb = v["bytes"] / (v["ms"] * 1e-3) / 1e9 if v["ms"] > 0 else 0.0
        bound = "tensor" if tf / tpk >= gb / hpk else "hbm"
        ach, peak, unit = (tf, tpk, "TFLOP/s") if bound == "tensor" else (gb, hpk, "GB/s")
        return {"bound": bound, "unit": unit, "peak": peak, "achieved_isolated": ach, "frac_isolated": ach / peak,
                "achieved": ach / stretch, "frac": ach / stretch / peak, "ms_per_step": v["ms"], "launches_per_step": v["launches"],
                "share_of_step": v["ms"] * stretch / ms_per_step, "tflops_isolated": tf, "hbm_gbs_isolated": gb}

    dom = max(by_kernel, key=lambda k: by_kernel[k]["ms"])
    d = entry(dom, by_kernel[dom])
    roof = {"kernel": dom, "bound": d["bound"], "achieved": d["achieved"], "peak": d["peak"], "unit": d["unit"], "frac": d["frac"],
            "traffic": None, "achieved_isolated": d["achieved_isolated"], "frac_isolated": d["frac_isolated"],
            "what": "achieved / frac are IN-STEP figures (per-launch event times stretched to the timed ms_per_step); "
                    "*_isolated are the launches' own CUDA-event times in profiled forwards run right after the timed loop",
            "peak_source": peaks["source"] + ", sustained figure (kernel timed inside a long step)",
            "launches_per_step": d["launches_per_step"], "share_of_step": d["share_of_step"],
            "algorithmic_flops_per_step": by_kernel[dom]["flops"], "algorithmic_bytes_per_launch": by_kernel[dom]["bytes"] / by_kernel[dom]["launches"],
            "avg_launch_ms": by_kernel[dom]["ms"] / by_kernel[dom]["launches"],
            "profile_sum_ms": prof_ms, "step_stretch": stretch,
            "kernels": {k: entry(k, v) for k, v in by_kernel.items()}}
    # DRAM bytes per launch: from the committed ncu capture of this same command at this commit's kernels (bench.py
    # never runs a profiler itself): the newest profiles/r02*_dram_traffic.json, written by scripts/ncu_traffic.py
    import glob
    tpaths = sorted(glob.glob(os.path.join(ROOT, "profiles", "r02*_dram_traffic.json")))
    tpath = tpaths[-1] if tpaths else ""
    if batch == 256 and tpath:
        with open(tpath) as f:
            tj = json.load(f)
        for k, v in tj.get("kernels", {}).items():
            if k in roof["kernels"] and v.get("launches") == roof["kernels"][k]["launches_per_step"]:
                roof["kernels"][k]["traffic"] = v["dram_bytes_per_launch_mean"]
        if dom in tj.get("kernels", {}) and "traffic" in roof["kernels"][dom]:
            roof["traffic"] = roof["kernels"][dom]["traffic"]
            roof["traffic_unit"] = "bytes/launch (dram__bytes_read.sum + dram__bytes_write.sum)"
            roof["traffic_source"] = "profiles/" + os.path.basename(tpath) + " (" + tj.get("commit", "?") + ")"
    return roof


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="episodes per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-other-configs", "--no-diffusion", dest="no_other", action="store_true",
                    help="skip the Diffusion Policy / DP3 / MT-ACT side measurements")
    ap.add_argument("--sync-gather", action="store_true", help="all-gather on the compute stream every step (A/B)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=device)
    from robomanipbaselines_b200 import _lib
    from robomanipbaselines_b200.common.batched_rollout import raise_on_device_fault
    from robomanipbaselines_b200.common.sharding import gather_episodes

    B, K, W = args.batch, args.steps, args.warmup
    pol, meta, br = build_rollout(device, B)
    timer = Timer(device, dist)

    # synthetic observations: 2 alternating input sets; one set (B*2*480*640*3 = 472 MB of uint8) is already 3.7x the L2
    gen = torch.Generator(device=device).manual_seed(1 + rank)
    imgs = [torch.randint(0, 256, (B, NCAM, IMG_H, IMG_W, 3), generator=gen, device=device, dtype=torch.uint8) for _ in range(2)]
    states = [torch.randn(B, DS, generator=gen, device=device, dtype=torch.float64) for _ in range(2)]
    # north_star: every rank sees all episodes' actions (KB-sized NCCL all-gather over NVLink). It runs on a side stream,
    # double-buffered: step i's gather overlaps step i + 1, whose result is consumed one step later.
    gat = make_gather(world, B, DA, device)

    def step(i):
        act = br.infer(imgs[i & 1], states[i & 1])
        if world > 1:
            if args.sync_gather:
                return gather_episodes(act, world * B)
            if gat.n_submit - gat.n_result >= gat.depth:
                gat.result()
            gat.submit(act)
        return act

    def barrier():
        gat.drain()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    br.reset()
    for i in range(W):
        step(i)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    # ---------------- device-resident throughput
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for i in range(K):
        step(i)
    gat.drain()                                   # the last steps' gathers belong to the timed region
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)

    # ---------------- per-launch CUDA-event profile, back to back with the timed loop (sustained clocks)
    eng = pol.engine_for(B, IMG_H, IMG_W, device)
    br.normalize_state(states[0])
    recs_runs = []
    for r in range(4):
        _, recs = eng.forward_profile(imgs[r & 1], br.qpos, out=br.chunks)
        if r > 0:
            recs_runs.append(recs)
    torch.cuda.synchronize()

    # ---------------- end to end: pinned host observations in, host actions out, every step (copies double-buffered
    # against compute on a second stream)
    h_imgs = [t.cpu().pin_memory() for t in imgs]
    h_states = [t.cpu().pin_memory() for t in states]
    h_act = torch.empty(B, DA, dtype=torch.float64).pin_memory()
    d_imgs = [torch.empty_like(imgs[0]) for _ in range(2)]
    d_states = [torch.empty_like(states[0]) for _ in range(2)]
    copy_stream = torch.cuda.Stream(device=device)
    main_stream = torch.cuda.current_stream(device)
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]

    def e2e_run(n):
        for i in range(n):
            j = i & 1
            with torch.cuda.stream(copy_stream):
                if i >= 2:
                    copy_stream.wait_event(consumed[j])
                d_imgs[j].copy_(h_imgs[j], non_blocking=True)
                d_states[j].copy_(h_states[j], non_blocking=True)
                copied[j].record(copy_stream)
            main_stream.wait_event(copied[j])
            act = br.infer(d_imgs[j], d_states[j])
            consumed[j].record(main_stream)
            if world > 1:
                if gat.n_submit - gat.n_result >= gat.depth:
                    gat.result()
                gat.submit(act)
            h_act.copy_(act, non_blocking=True)

    e2e_run(2)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_run(K)
    gat.drain()
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    clocks = sampler.stop()

    t = torch.tensor([ms_total, ms_e2e], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = t.tolist()

    latency = None
    if not args.no_latency and rank == 0:
        from robomanipbaselines_b200.common.batched_rollout import BatchedRollout

        br1 = BatchedRollout(pol, 1, meta, device=str(device))
        i1, s1 = imgs[0][:1].contiguous(), states[0][:1].contiguous()
        lat = []
        for i in range(60):
            t0 = time.perf_counter()
            a = br1.infer(i1, s1)
            a.cpu()
            if i >= 10:
                lat.append(1e3 * (time.perf_counter() - t0))
        latency = {"p50_batch1_ms": float(np.median(lat)), "what": "BatchedRollout(B=1).infer + D2H of the action, host wall clock, 50 iterations"}

    line = None
    if rank == 0:
        peaks = measured_peaks()
        n_launch = eng.num_launches + 2      # + normalize_state + ensemble_step
        line = {
            "metric": METRIC, "value": world * B * K / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": _lib.load().rmb_act_dtype().decode(), "data": "synthetic",
            "config": {"workload": WORKLOAD, "episodes_per_gpu": B, "policy_steps_per_step": world * B,
                       "weights": "random-init: the distributions of a freshly constructed reference ACTPolicy "
                                  "(robomanipbaselines_b200/common/synthetic.py::default_init_state_dict, seed 0)",
                       "l2": "inputs larger than L2: 472 MB of uint8 images per step, two alternating input sets",
                       "parallelism": f"{world} independent replicas, episodes sharded by rank"
                                      + (", NCCL all-gather of the actions every step on a side stream (double-buffered)" if world > 1 else ""),
                       "flops_per_policy_step": eng.flops(B) / B},
            "clocks": clocks,
            "e2e": {"value": world * B * K / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e / K,
                    "h2d_bytes_per_step": int(imgs[0].numel() + states[0].numel() * 8),
                    "d2h_bytes_per_step": int(B * DA * 8),
                    "what": "BatchedRollout.infer on pinned host uint8 images + float64 states, float64 actions back to host every step"},
            "gpu_launches": n_launch * K,
            "roofline": roofline_block(recs_runs, ms_total / K, peaks, B),
            "model_tflops": eng.flops(B) * K / (ms_total * 1e-3) / 1e12,
        }
        if latency:
            line["latency"] = latency
    raise_on_device_fault(what="bench.py (ACT)")
    if not args.no_other:
        del imgs, d_imgs, h_imgs, br
        pol._invalidate()
        torch.cuda.empty_cache()
        other = bench_diffusion(device, timer, world)
        other["mt_act_c4"] = bench_mt_act(device, timer, world)
        raise_on_device_fault(what="bench.py (other configs)")
        if rank == 0:
            line["other_configs"] = other
            if world == 1:
                line["microbench"] = bench_micro(device)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            times, cores, kind, what = cpu_arm(CPU_SAMPLE, 3, 1)
            line["cpu_baseline"] = {"value": CPU_SAMPLE / float(np.mean(times)), "unit": UNIT, "cores": cores, "kind": kind,
                                    "sample": f"3 steps x {CPU_SAMPLE} episodes of the same workload, {what}, {cores} threads"}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
