#!/usr/bin/env python
"""Per-launch CUDA-event table of one ACT forward at the bench workload (B episodes, 2 cams 480x640).

    python scripts/profile_step.py [--batch 256] [--out gpurun_out/launches.json] [--once]

--once runs a single un-profiled forward after two warm-ups (the command line to put under ncu).
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--out", default=None)
    ap.add_argument("--once", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    B = args.batch
    pol, meta, br = bench.build_rollout(dev, B)
    gen = torch.Generator(device=dev).manual_seed(1)
    img = torch.randint(0, 256, (B, bench.NCAM, bench.IMG_H, bench.IMG_W, 3), generator=gen, device=dev, dtype=torch.uint8)
    st = torch.randn(B, bench.DS, generator=gen, device=dev, dtype=torch.float64)
    br.reset()
    for _ in range(2):
        br.infer(img, st)
    torch.cuda.synchronize()
    if args.once:
        br.infer(img, st)
        torch.cuda.synchronize()
        print("ok")
        return
    eng = pol.engine_for(B, bench.IMG_H, bench.IMG_W, dev)
    recs = None
    for _ in range(5):            # per-launch minimum over 5 profiled forwards (boxes differ by ~10 % run to run)
        _, rr = eng.forward_profile(img, br.qpos, out=br.chunks)
        if recs is None:
            recs = rr
        else:
            for a, b in zip(recs, rr):
                a["ms"] = min(a["ms"], b["ms"])
    total = sum(r["ms"] for r in recs)
    print(f"{'launch':34s} {'kernel':26s} {'ms':>8s} {'%':>6s} {'TFLOP/s':>9s} {'GB/s':>8s}")
    for r in recs:
        tf = r["flops"] / (r["ms"] * 1e-3) / 1e12 if r["ms"] > 0 else 0
        gb = r["bytes"] / (r["ms"] * 1e-3) / 1e9 if r["ms"] > 0 else 0
        print(f"{r['name']:34s} {r['kernel']:26s} {r['ms']:8.4f} {100 * r['ms'] / total:6.2f} {tf:9.1f} {gb:8.1f}")
    print(f"total {total:.3f} ms for {B} episodes -> {B / total * 1e3:.0f} policy steps/s; "
          f"{eng.flops(B) / total / 1e9:.1f} TFLOP/s overall")
    if args.out:
        with open(args.out, "w") as f:
            json.dump({"batch": B, "total_ms": total, "launches": recs}, f, indent=1)


if __name__ == "__main__":
    main()
