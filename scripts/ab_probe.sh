#!/bin/bash
# A/B of the kernel switches on the per-launch table of one ACT forward (B = 256), same box
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 120 python -m pytest -q -m gpu -p no:cacheprovider tests/test_kernels_gpu.py -k "residual_layernorm" 2>&1 | tail -2
RMB_LN_FUSE=always timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/ab_ln_always.txt 2>&1
RMB_LN_FUSE=never timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/ab_ln_never.txt 2>&1
tail -n 1 gpurun_out/ab_ln_always.txt gpurun_out/ab_ln_never.txt
grep -E "enc1 (out|norm|linear2)" gpurun_out/ab_ln_always.txt gpurun_out/ab_ln_never.txt
