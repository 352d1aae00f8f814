#!/usr/bin/env python
"""bench.py -- batched ACT rollout inference throughput on B200 (BASELINE.json metric, configs[1]).

  python bench.py --gpus N --steps K --warmup W            # this engine (under torchrun for N > 1)
  python bench.py --impl reference --steps K --warmup W    # CPU arm: the oracle port of the reference path on host cores

A "step" is one BatchedRollout.infer() over B = 256 episodes per GPU (2 RGB cameras 480x640 uint8 + 7-d joint state,
chunk 100, temporal ensembling): state normalise -> ACT forward -> ensembling ring buffer -> denormalise, i.e. 256
policy steps of RolloutAct.infer_policy. Prints ONE JSON line on rank 0.
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


def build_rollout(device, batch):
    from robomanipbaselines_b200.common.batched_rollout import BatchedRollout
    from robomanipbaselines_b200.common.synthetic import synthetic_norm_stats, synthetic_state_dict
    from robomanipbaselines_b200.policy.act import ACTPolicy

    pol = ACTPolicy(POLICY_ARGS, state_dim=DS, action_dim=DA, max_batch=batch)
    shapes = {k: tuple(v.shape) for k, v in pol.state_dict().items()}
    pol.load_state_dict(synthetic_state_dict(shapes, seed=0), strict=True)
    meta = {"data": {"chunk_size": CHUNK}, "state": synthetic_norm_stats(DS, 1), "action": synthetic_norm_stats(DA, 2)}
    return pol, meta, BatchedRollout(pol, batch, meta, device=str(device))


def bench_diffusion(device, reps=2):
    """Secondary workloads of BASELINE.json (configs[2] and the per-GPU share of configs[4]): one predict_action =
    obs encoder + the whole scheduler loop, synthetic observations resident in HBM, CUDA events. Reported as
    inferences/s and policy steps/s (one inference is amortised over n_action_steps = 8 policy steps)."""
    from robomanipbaselines_b200.common.schedulers import DDIMScheduler, DDPMScheduler
    from robomanipbaselines_b200.common.synthetic import synthetic_state_dict
    from robomanipbaselines_b200.policy.diffusion_policy import DiffusionUnetHybridImagePolicy
    from robomanipbaselines_b200.policy.diffusion_policy.policy import visual_core_param_shapes
    from robomanipbaselines_b200.policy.diffusion_policy_3d import DP3

    sched = dict(beta_end=0.02, beta_schedule="squaredcos_cap_v2", beta_start=0.0001, clip_sample=True,
                 num_train_timesteps=100)
    out = {}
    gen = torch.Generator(device=device).manual_seed(3)

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    # ---- configs[2]: Diffusion Policy, batch 1024, DDPM 100 and DDIM-like 16 steps, one 240x320 camera + 7-d state
    B, To = 1024, 2
    pol = DiffusionUnetHybridImagePolicy(
        noise_scheduler=DDPMScheduler(**sched, prediction_type="epsilon", variance_type="fixed_small"),
        shape_meta={"obs": {"state": {"shape": [DS], "type": "low_dim"},
                            "front_rgb_image": {"shape": [3, 240, 320], "type": "rgb"}}, "action": {"shape": [DA]}},
        horizon=16, n_action_steps=8, n_obs_steps=To, num_inference_steps=100, obs_as_global_cond=True,
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
    ms_enc = timed(lambda: pol.encode_obs(obs))
    ms = timed(lambda: pol.predict_action(obs))
    unet = pol.unet_for(B, device)
    enc = pol.encoder_for("front_rgb_image", B * To, device)
    flops = unet.step_flops(B) * 100 + enc.flops(B * To)
    out["diffusion_policy_c3"] = {
        "workload": "DiffusionUnetHybridImagePolicy.predict_action, batch 1024, horizon 16, DDPM 100 steps, ResNet-18 "
                    "(GroupNorm) obs encoder on one 240x320 camera (crop 216x288) x 2 obs steps + 7-d state",
        "ms_per_inference_batch": ms, "ms_obs_encoder": ms_enc, "inferences_per_s": B / ms * 1e3,
        "policy_steps_per_s": 8 * B / ms * 1e3, "model_tflops": flops / ms / 1e9,
        "gpu_launches": unet.launches_per_step * 100 + 3 + 70}
    del pol, obs, unet, enc
    torch.cuda.empty_cache()

    # ---- configs[4] per-GPU share: DP3, batch 512 (= 4096 / 8), 1024-point clouds, DDIM 10 and 16
    B = 512
    for n_steps in (10, 16):
        pol = DP3(noise_scheduler=DDIMScheduler(**sched, set_alpha_to_one=True, prediction_type="sample", steps_offset=0),
                  shape_meta={"obs": {"state": {"shape": [DS]}, "point_cloud": {"shape": (1024, 6)}}, "action": {"shape": [DA]}},
                  horizon=16, n_action_steps=8, n_obs_steps=To, num_inference_steps=n_steps, diffusion_step_embed_dim=128,
                  down_dims=[512, 1024, 2048], kernel_size=5, n_groups=8, pointcloud_encoder_cfg={"out_channels": 64},
                  use_pc_color=False, encoder_output_dim=64, max_batch=B)
        pol.load_state_dict(synthetic_state_dict({k: tuple(v.shape) for k, v in pol.state_dict().items()}, seed=0))
        obs = {"state": torch.randn(B, To, DS, device=device, generator=gen),
               "point_cloud": torch.rand(B, To, 1024, 6, device=device, generator=gen) * 2 - 1}
        ms_enc = timed(lambda: pol.encode_obs(obs))
        ms = timed(lambda: pol.predict_action(obs))
        unet = pol.unet_for(B, device)
        out[f"dp3_c5_ddim{n_steps}"] = {
            "workload": f"DP3.predict_action, batch 512 per GPU (4096 / 8), 2 x 1024-point clouds + 7-d state, DDIM {n_steps} steps",
            "ms_per_inference_batch": ms, "ms_obs_encoder": ms_enc, "inferences_per_s": B / ms * 1e3,
            "policy_steps_per_s": 8 * B / ms * 1e3, "model_tflops": unet.step_flops(B) * n_steps / ms / 1e9,
            "gpu_launches": unet.launches_per_step * n_steps + 3 + 7}
        del pol, obs, unet
        torch.cuda.empty_cache()
    return out


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
                            "what": "temporal ensembling with a full 100-chunk history per episode, 18 GB ring buffer; "
                                    "the diagonal gather touches one 28-byte row per 2.8 KB slot (sector-granular reads)"}
    del ring, chunks
    torch.cuda.empty_cache()
    return out


def cpu_reference_step(sd, meta, oracles, images_u8, raw_state, live_only):
    """One policy step of the reference path for len(oracles) episodes on the CPU (oracle port, fp32 PyTorch + numpy)."""
    from oracle import act_oracle, rollout_oracle

    B = len(oracles)
    qpos = np.concatenate([rollout_oracle.get_state(raw_state[b], meta["state"]) for b in range(B)])
    image = np.concatenate([rollout_oracle.get_images(list(images_u8[b])) for b in range(B)])
    chunks = act_oracle.act_forward(sd, POLICY_ARGS, torch.from_numpy(qpos), torch.from_numpy(image), live_only=live_only).numpy()
    return np.stack([oracles[b].step(chunks[b]) for b in range(B)])


def time_cpu_baseline(sample_batch, iters, warmup, live_only):
    from oracle import rollout_oracle
    from robomanipbaselines_b200.common.synthetic import (synthetic_act_observations, synthetic_norm_stats,
                                                          synthetic_state_dict)
    from robomanipbaselines_b200.policy.act import ACTPolicy

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    pol = ACTPolicy(POLICY_ARGS, state_dim=DS, action_dim=DA)
    sd = synthetic_state_dict({k: tuple(v.shape) for k, v in pol.state_dict().items()}, seed=0)
    meta = {"data": {"chunk_size": CHUNK}, "state": synthetic_norm_stats(DS, 1), "action": synthetic_norm_stats(DA, 2)}
    oracles = [rollout_oracle.ActRolloutOracle(CHUNK, DA, meta["action"]) for _ in range(sample_batch)]
    images, state = synthetic_act_observations(sample_batch, NCAM, IMG_H, IMG_W, DS, seed=1)
    times = []
    for i in range(warmup + iters):
        t0 = time.perf_counter()
        cpu_reference_step(sd, meta, oracles, images, state, live_only)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return times, cores


def run_reference(args):
    """--impl reference: the reference's CPU path (fp32 eager PyTorch, as executed: all 7 decoder layers) timed on the
    host cores through the oracle port -- the reference's Python sources cannot travel to the GPU box."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample = 8
    times, cores = time_cpu_baseline(sample, args.steps, max(args.warmup, 1), live_only=False)
    ms = 1e3 * float(np.mean(times))
    value = sample / (ms / 1e3)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": f"{sample} episodes per step (bounded sample of the 256-episode batch)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps x {sample} episodes, oracle/act_oracle.py + oracle/rollout_oracle.py, "
                                   f"fp32 torch CPU, {cores} threads, all 7 decoder layers as the reference executes them"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="episodes per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-diffusion", action="store_true", help="skip the Diffusion Policy / DP3 side measurements")
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
    B, K, W = args.batch, args.steps, args.warmup
    pol, meta, br = build_rollout(device, B)

    # synthetic observations: 2 alternating input sets; one set (B*2*480*640*3 = 472 MB of uint8) is already 3.7x the L2
    gen = torch.Generator(device=device).manual_seed(1 + rank)
    imgs = [torch.randint(0, 256, (B, NCAM, IMG_H, IMG_W, 3), generator=gen, device=device, dtype=torch.uint8) for _ in range(2)]
    states = [torch.randn(B, DS, generator=gen, device=device, dtype=torch.float64) for _ in range(2)]
    from robomanipbaselines_b200.common.sharding import gather_episodes

    def step(i):
        act = br.infer(imgs[i & 1], states[i & 1])
        if world > 1:   # north_star: every rank sees all episodes' actions (KB-sized NCCL all-gather over NVLink)
            act = gather_episodes(act, world * B)
        return act

    def barrier():
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
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)

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
                gather_episodes(act, world * B)
            h_act.copy_(act, non_blocking=True)

    e2e_run(2)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_run(K)
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)
    clocks = sampler.stop()

    # ---------------- per-launch CUDA-event profile of one step (roofline of the dominant kernel)
    eng = pol.engine_for(B, IMG_H, IMG_W, device)
    br.normalize_state(states[0])
    for _ in range(2):
        _, recs = eng.forward_profile(imgs[0], br.qpos, out=br.chunks)
    torch.cuda.synchronize()

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

    if rank == 0:
        peaks = measured_peaks()
        by_kernel = {}
        for r in recs:
            k = by_kernel.setdefault(r["kernel"], {"ms": 0.0, "flops": 0.0, "bytes": 0.0, "launches": 0})
            k["ms"] += r["ms"]; k["flops"] += r["flops"]; k["bytes"] += r["bytes"]; k["launches"] += 1
        step_ms_prof = sum(r["ms"] for r in recs)
        dom = max(by_kernel, key=lambda k: by_kernel[k]["ms"])
        d = by_kernel[dom]
        achieved = d["flops"] / (d["ms"] * 1e-3) / 1e12
        roofline = {
            "kernel": dom, "bound": "tensor", "achieved": achieved, "peak": peaks["bf16_tflops_sustained"],
            "unit": "TFLOP/s", "frac": achieved / peaks["bf16_tflops_sustained"], "traffic": None,
            "peak_source": peaks["source"] + ", sustained figure (kernel timed inside a long step)",
            "launches_per_step": d["launches"], "share_of_step": d["ms"] / step_ms_prof,
            "algorithmic_flops_per_step": d["flops"], "avg_launch_ms": d["ms"] / d["launches"],
            "other_kernels": {k: {"ms": v["ms"], "launches": v["launches"],
                                  "hbm_gbs": (v["bytes"] / (v["ms"] * 1e-3) / 1e9) if v["ms"] > 0 else None,
                                  "tflops": (v["flops"] / (v["ms"] * 1e-3) / 1e12) if v["ms"] > 0 else None}
                              for k, v in by_kernel.items()},
        }
        # DRAM bytes per launch of the dominant kernel: from the committed ncu capture of this same command (bench.py never
        # runs a profiler itself); mean over the launches of one step, like `achieved`
        tpath = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01o_gemm_pair_dram_traffic.json")
        if dom == "gemm_tc_pair_kernel" and B == 256 and os.path.exists(tpath):
            with open(tpath) as f:
                tj = json.load(f)
            if tj.get("launches") == d["launches"]:
                roofline["traffic"] = tj["dram_bytes_per_launch_mean"]
                roofline["traffic_unit"] = "bytes/launch (dram__bytes_read.sum + dram__bytes_write.sum)"
                roofline["traffic_source"] = "profiles/r01o_gemm_pair_dram_traffic.json"
                roofline["algorithmic_bytes_per_launch"] = d["bytes"] / d["launches"]
        n_launch = eng.num_launches + 2      # + normalize_state + ensemble_step
        line = {
            "metric": METRIC, "value": world * B * K / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "fp16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "episodes_per_gpu": B, "policy_steps_per_step": world * B,
                       "weights": "random-init (robomanipbaselines_b200/common/synthetic.py, seed 0)",
                       "l2": "inputs larger than L2: 472 MB of uint8 images per step, two alternating input sets",
                       "parallelism": f"{world} independent replicas, episodes sharded by rank"
                                      + (", NCCL all-gather of actions per step" if world > 1 else ""),
                       "flops_per_policy_step": eng.flops(B) / B},
            "clocks": clocks,
            "e2e": {"value": world * B * K / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e / K,
                    "h2d_bytes_per_step": int(imgs[0].numel() + states[0].numel() * 8),
                    "d2h_bytes_per_step": int(B * DA * 8),
                    "what": "BatchedRollout.infer on pinned host uint8 images + float64 states, float64 actions back to host every step"},
            "gpu_launches": n_launch * K,
            "roofline": roofline,
            "model_tflops": eng.flops(B) * K / (ms_total * 1e-3) / 1e12,
        }
        if latency:
            line["latency"] = latency
        if world == 1 and not args.no_diffusion:
            del imgs, d_imgs, h_imgs
            pol._invalidate()
            torch.cuda.empty_cache()
            line["other_configs"] = bench_diffusion(device)
            line["microbench"] = bench_micro(device)
        if world == 1 and not args.no_cpu_baseline:
            sample = 4
            times, cores = time_cpu_baseline(sample, 4, 1, live_only=False)
            line["cpu_baseline"] = {"value": sample / float(np.mean(times)), "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"4 steps x {sample} episodes of the same workload, oracle port (fp32 torch CPU, "
                                              f"{cores} threads, all 7 decoder layers as the reference executes them)"}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
