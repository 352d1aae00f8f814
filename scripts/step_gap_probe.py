"""Whole-step time (CUDA events around 10 back-to-back steps) for eager launches vs one CUDA graph per step at B=256,
next to the sum of per-launch times: where do the ~2 ms between "sum of kernels" and "step" go?"""
import os
import sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from robomanipbaselines_b200.common.batched_rollout import BatchedRollout

dev = torch.device("cuda:0")
B = 256
pol, meta, br = bench.build_rollout(dev, B)
gen = torch.Generator(device=dev).manual_seed(1)
imgs = [torch.randint(0, 256, (B, bench.NCAM, bench.IMG_H, bench.IMG_W, 3), generator=gen, device=dev, dtype=torch.uint8) for _ in range(2)]
states = [torch.randn(B, bench.DS, generator=gen, device=dev, dtype=torch.float64) for _ in range(2)]


def run(r, n):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(3):
        r.infer(imgs[i & 1], states[i & 1])
    torch.cuda.synchronize()
    e0.record()
    for i in range(n):
        r.infer(imgs[i & 1], states[i & 1])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


br.reset()
print(f"eager: {run(br, 10):.3f} ms/step", flush=True)
brg = BatchedRollout(pol, B, meta, device=str(dev), use_cuda_graph=True)
brg.reset()
print(f"graph: {run(brg, 10):.3f} ms/step", flush=True)
print(f"eager again: {run(br, 10):.3f} ms/step", flush=True)
