#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
PT="python -m pytest -q -s -m gpu -p no:cacheprovider"
timeout 300 $PT tests/test_kernels_gpu.py -k "ensemble" 2>&1 | tail -2
timeout 600 $PT tests/test_act_engine_gpu.py tests/test_rollout_classes_gpu.py 2>&1 | tail -2
timeout 900 $PT tests/test_diffusion_gpu.py -k "bimanual or unet_forward or denoise" 2>&1 | tail -3
timeout 300 python - <<'PY' 2>&1 | tail -3
import torch, json, sys
sys.path.insert(0, ".")
import bench
print(json.dumps(bench.bench_micro(torch.device("cuda:0"))["ensemble_step"]))
PY
