"""p50 batch-1 action latency: BatchedRollout(B).infer on device-resident observations + D2H of the action, host wall
clock (the `latency` entry of bench.py), for a few small batch sizes."""
import os
import sys
import time
import numpy as np
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
from robomanipbaselines_b200.common.batched_rollout import BatchedRollout

dev = torch.device("cuda:0")
for B in (1, 2, 8):
    pol, meta, br = bench.build_rollout(dev, B)
    gen = torch.Generator(device=dev).manual_seed(1)
    img = torch.randint(0, 256, (B, bench.NCAM, bench.IMG_H, bench.IMG_W, 3), generator=gen, device=dev, dtype=torch.uint8)
    st = torch.randn(B, bench.DS, generator=gen, device=dev, dtype=torch.float64)
    br.reset()
    for _ in range(5):
        br.infer(img, st)
    torch.cuda.synchronize()
    lat = []
    for _ in range(100):
        t0 = time.perf_counter()
        a = br.infer(img, st).cpu()
        lat.append(1e3 * (time.perf_counter() - t0))
    print(f"B={B}: p50 {np.median(lat):.4f} ms  p90 {np.percentile(lat, 90):.4f} ms  (cuda graph: {br.use_cuda_graph})", flush=True)
