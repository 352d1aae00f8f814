"""Times rmb_pointcloud_from_depth (DP3 observation pipeline) on B frames of 84 x 84 pixels: python scripts/pointcloud_probe.py [B]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from robomanipbaselines_b200.common.pointcloud import DepthToPointcloud  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
rng = np.random.default_rng(0)
depth = torch.from_numpy(rng.uniform(0.3, 2.0, (B, 84, 84)).astype(np.float32)).cuda()
rgb = torch.from_numpy(rng.integers(0, 256, (B, 84, 84, 3), dtype=np.uint8)).cuda()
pipe = DepthToPointcloud(B, (84, 84), 45.0, 1024, [-1.5, -1.5, 0.2], [1.5, 1.5, 2.5], None, device="cuda:0")
for _ in range(2):
    pipe(depth, rgb)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    pipe(depth, rgb)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
print(f"pointcloud_from_depth: {B} frames 84x84 -> 1024 points: {ms:.3f} ms ({B / ms * 1e3:.0f} clouds/s), valid points {pipe.count[:4].tolist()}")
