#!/usr/bin/env python
"""Per-launch CUDA-event table of one ConditionalUnet1D evaluation + timed conditional_sample loops.

    python scripts/profile_unet.py [--batch 1024] [--cond 142] [--kind ddpm --steps 100] [--once]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robomanipbaselines_b200.common.schedulers import DDIMScheduler, DDPMScheduler  # noqa: E402
from robomanipbaselines_b200.common.synthetic import synthetic_state_dict  # noqa: E402
from robomanipbaselines_b200.engine import UnetEngine  # noqa: E402
from robomanipbaselines_b200.policy.diffusion_common import unet1d_param_shapes  # noqa: E402

SCHED = dict(beta_end=0.02, beta_schedule="squaredcos_cap_v2", beta_start=0.0001, clip_sample=True,
             num_train_timesteps=100)


def build_unet(dev, B, G, Da=7, T=16, down_dims=(512, 1024, 2048)):
    eng = UnetEngine(max_batch=B, horizon=T, input_dim=Da, global_cond_dim=G, down_dims=down_dims, device=dev.index or 0)
    eng.load_state_dict(synthetic_state_dict(unet1d_param_shapes(Da, G, 128, down_dims, 5), seed=0))
    return eng


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--cond", type=int, default=142)
    ap.add_argument("--kind", default="ddpm")
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--out", default=None)
    ap.add_argument("--once", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    B, G, Da, T = args.batch, args.cond, 7, 16
    eng = build_unet(dev, B, G)
    gen = torch.Generator(device=dev).manual_seed(1)
    x = torch.randn(B, T, Da, device=dev, generator=gen)
    gc = torch.randn(B, G, device=dev, generator=gen)
    for _ in range(3):
        eng.forward(x, 50, gc)
    torch.cuda.synchronize()
    if args.once:
        eng.forward(x, 50, gc)
        torch.cuda.synchronize()
        print("ok")
        return
    recs = None
    for _ in range(5):
        _, rr = eng.forward_profile(x, 50, gc)
        if recs is None:
            recs = rr
        else:
            for a, b in zip(recs, rr):
                a["ms"] = min(a["ms"], b["ms"])
    total = sum(r["ms"] for r in recs)
    print(f"{'launch':44s} {'kernel':24s} {'ms':>8s} {'%':>6s} {'TFLOP/s':>9s} {'GB/s':>8s}")
    for r in recs:
        tf = r["flops"] / (r["ms"] * 1e-3) / 1e12 if r["ms"] > 0 else 0
        gb = r["bytes"] / (r["ms"] * 1e-3) / 1e9 if r["ms"] > 0 else 0
        print(f"{r['name']:44s} {r['kernel']:24s} {r['ms']:8.4f} {100 * r['ms'] / total:6.2f} {tf:9.1f} {gb:8.1f}")
    by = {}
    for r in recs:
        k = by.setdefault(r["kernel"], [0.0, 0.0, 0])
        k[0] += r["ms"]; k[1] += r["flops"]; k[2] += 1
    for k, (ms, fl, n) in by.items():
        print(f"  {k}: {n} launches, {ms:.3f} ms ({100 * ms / total:.1f} %), {fl / ms / 1e9:.1f} TFLOP/s")
    print(f"sum of launches {total:.3f} ms per UNet evaluation at B={B}; {eng.step_flops(B) / total / 1e9:.1f} TFLOP/s "
          f"algorithmic ({eng.step_flops(1) / 1e9:.3f} GFLOP/sample/step)")
    # whole loop, CUDA events, back to back
    sched = DDPMScheduler(**SCHED, prediction_type="epsilon") if args.kind == "ddpm" else \
        DDIMScheduler(**SCHED, prediction_type="sample")
    steps = sched.step_table(args.steps)
    noise = torch.randn(len(steps), B, T, Da, device=dev, generator=gen) if args.kind == "ddpm" else None
    out = torch.empty_like(x)
    eng.denoise(gc, x, steps, noise=noise, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record()
    for _ in range(reps):
        eng.denoise(gc, x, steps, noise=noise, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{args.kind}-{len(steps)} loop at B={B}: {ms:.2f} ms per inference batch -> {B / ms * 1e3:.0f} inferences/s, "
          f"{8 * B / ms * 1e3:.0f} policy steps/s (n_action_steps 8); {eng.step_flops(B) * len(steps) / ms / 1e9:.1f} TFLOP/s")
    # the same loop replayed as one CUDA graph (what DiffusionPolicyBase.conditional_sample does)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        eng.denoise(gc, x, steps, noise=noise, out=out)
    graph.replay()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    msg = e0.elapsed_time(e1) / reps
    print(f"  as one CUDA graph: {msg:.2f} ms -> {B / msg * 1e3:.0f} inferences/s, {8 * B / msg * 1e3:.0f} policy steps/s; "
          f"{eng.step_flops(B) * len(steps) / msg / 1e9:.1f} TFLOP/s")
    if args.out:
        with open(args.out, "w") as f:
            json.dump({"batch": B, "total_ms": total, "loop_ms": ms, "launches": recs}, f, indent=1)


if __name__ == "__main__":
    main()
