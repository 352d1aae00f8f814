#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/ab_base.txt 2>&1
RMB_HALO_OUT_BUFS=21 timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/ab_halo21.txt 2>&1
tail -n 1 gpurun_out/ab_base.txt gpurun_out/ab_halo21.txt
grep -E "layer1" gpurun_out/ab_base.txt gpurun_out/ab_halo21.txt
