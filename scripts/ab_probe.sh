#!/bin/bash
# A/B of the kernel switches on the per-launch table of one ACT forward (B = 256)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/ab_default.txt 2>&1
RMB_LN_FUSE=never timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/ab_ln_never.txt 2>&1
RMB_ATTN_V1=1 timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/ab_attn_v1.txt 2>&1
tail -n 1 gpurun_out/ab_*.txt
