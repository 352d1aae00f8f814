#!/bin/bash
# kernel tests for the kernels under work + A/B launch table + ncu full of selected kernels
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
PT="python -m pytest -q -s -m gpu -p no:cacheprovider"
timeout 300 $PT tests/test_kernels_gpu.py -k "attention or layernorm or stem_pool" > gpurun_out/r2_kernels.log 2>&1; echo "kernels rc=$?"; tail -3 gpurun_out/r2_kernels.log
timeout 600 $PT tests/test_act_engine_gpu.py tests/test_mt_act_gpu.py > gpurun_out/r2_engine.log 2>&1; echo "engine rc=$?"; tail -3 gpurun_out/r2_engine.log
timeout 300 python scripts/profile_step.py --batch 256 > gpurun_out/r2_table.txt 2>&1; tail -1 gpurun_out/r2_table.txt
for SPEC in "$@"; do
  REGEX=${SPEC%%:*}; SKIP=${SPEC##*:}
  NAME="r2_$(echo "$REGEX" | tr -c 'a-zA-Z0-9' '_')"
  timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$REGEX" -s "$SKIP" -c 1 -f -o "gpurun_out/$NAME" \
      python scripts/profile_step.py --batch 64 --once > "gpurun_out/$NAME.log" 2>&1
  echo "ncu $REGEX rc=$?"
  ncu -i "gpurun_out/$NAME.ncu-rep" --page details > "gpurun_out/$NAME.details.txt" 2>/dev/null
  ncu -i "gpurun_out/$NAME.ncu-rep" --page source --csv > "gpurun_out/$NAME.source.csv" 2>/dev/null
  rm -f "gpurun_out/$NAME.ncu-rep"
done
