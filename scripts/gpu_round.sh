#!/bin/bash
# One gpurun call: per-group GPU tests (separate processes so a trapped kernel cannot poison later groups),
# then smoke + a short bench. Everything is bounded by `timeout`; logs land in gpurun_out/.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { # name, timeout, cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout "$t" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "rc=$rc" | tee -a gpurun_out/summary.txt
  tail -n 6 "gpurun_out/$name.log" | tee -a gpurun_out/summary.txt
  return $rc
}
: > gpurun_out/summary.txt
PT="python -m pytest -q -s -m gpu -p no:cacheprovider"
run t_elem 300 $PT tests/test_kernels_gpu.py -k "maxpool or layernorm or ensemble or normalize"
run t_gemm 300 $PT tests/test_kernels_gpu.py -k "test_gemm"
run t_conv 300 $PT tests/test_kernels_gpu.py -k "test_conv"
run t_stem 300 $PT tests/test_kernels_gpu.py -k "preprocess_and_stem or stem_pool"
run t_attn 300 $PT tests/test_kernels_gpu.py -k "test_attention"
RMB_ATTN_V2=1 run t_attn_v2 300 $PT tests/test_kernels_gpu.py -k "test_attention"
run t_engine 600 $PT tests/test_act_engine_gpu.py
ENGINE_RC=$?
run t_diffusion 900 $PT tests/test_diffusion_gpu.py
run t_pointcloud 300 $PT tests/test_pointcloud_gpu.py
run t_resize 300 $PT tests/test_resize_gpu.py
run t_glue 300 $PT tests/test_rollout_glue_gpu.py
run t_mtact 600 $PT tests/test_mt_act_gpu.py
run t_consumer 300 $PT tests/test_action_consumer_gpu.py
run t_classes 600 $PT tests/test_rollout_classes_gpu.py
run t_dtype 900 $PT tests/test_dtype_builds_gpu.py
if [ "$1" != "--no-bench" ] && [ $ENGINE_RC -eq 0 ]; then
  run smoke 300 python -c "import __graft_entry__ as g; g.smoke()"
  run bench 900 python bench.py --steps 10 --warmup 3
  timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/table.txt 2>&1
  RMB_LN_FUSE=never timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/table_ln_never.txt 2>&1
  tail -n 1 gpurun_out/table.txt gpurun_out/table_ln_never.txt
fi
exit 0
