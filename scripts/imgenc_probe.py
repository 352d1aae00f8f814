"""One DP image-encoder pass at the bench shape (2048 images 240x320, crop 216x288): the command line to put under
`ncu --metrics gpu__time_duration.sum` for a per-kernel launch list; prints the CUDA-event time of a pass."""
import os
import sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from robomanipbaselines_b200.engine import ImgEncEngine
from robomanipbaselines_b200.common.synthetic import synthetic_state_dict
from robomanipbaselines_b200.policy.diffusion_policy.policy import visual_core_param_shapes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
dev = torch.device("cuda:0")
eng = ImgEncEngine(max_images=n, in_h=240, in_w=320, crop_h=216, crop_w=288, device=0)
shapes = {k: v for k, v in visual_core_param_shapes().items()}
eng.load_state_dict(synthetic_state_dict(shapes, seed=5))
img = torch.rand(n, 3, 240, 320, device=dev) * 2 - 1
for _ in range(2):
    eng.encode(img)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
eng.encode(img)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print(f"imgenc n={n}: {ms:.3f} ms  {eng.flops(n) / ms / 1e9:.1f} TFLOP/s", flush=True)
