#!/bin/bash
# One `ncu --set full` capture of the tensor-core kernels of one ACT forward (B200_PROFILING.md recipe):
# the plain run must exit 0 first; output lands in gpurun_out/<name>.ncu-rep.
#   bash scripts/ncu_full.sh <name> <batch> <kernel-regex> <skip> <count>
cd "$(dirname "$0")/.." || exit 1
NAME=${1:-full}; BATCH=${2:-64}; REGEX=${3:-gemm_tc|attention_tc}; SKIP=${4:-102}; COUNT=${5:-51}
mkdir -p gpurun_out
timeout 300 python scripts/profile_step.py --batch "$BATCH" --once > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$REGEX" -s "$SKIP" -c "$COUNT" \
    -f -o "gpurun_out/$NAME" python scripts/profile_step.py --batch "$BATCH" --once > gpurun_out/ncu.log 2>&1
echo "rc=$?"
tail -n 3 gpurun_out/plain.log gpurun_out/ncu.log
# the report must come back inside gpurun's 64 MiB limit: always export the raw page as CSV, drop an oversized report
ncu -i "gpurun_out/$NAME.ncu-rep" --page raw --csv > "gpurun_out/$NAME.raw.csv" 2>/dev/null
SZ=$(stat -c %s "gpurun_out/$NAME.ncu-rep" 2>/dev/null || echo 0)
echo "report bytes: $SZ"
if [ "$SZ" -gt 50000000 ]; then
  ncu -i "gpurun_out/$NAME.ncu-rep" --page source --csv > "gpurun_out/$NAME.source.csv" 2>/dev/null
  rm -f "gpurun_out/$NAME.ncu-rep"
fi
