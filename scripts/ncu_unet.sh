#!/bin/bash
# One `ncu --set full` capture of a few GEMM launches of one ConditionalUnet1D evaluation (B200_PROFILING.md recipe):
# the plain run must exit 0 first; CSV exports land in gpurun_out/ (the .ncu-rep is dropped when large).
#   bash scripts/ncu_unet.sh <name> <batch> <skip> <count>
cd "$(dirname "$0")/.." || exit 1
NAME=${1:-unet}; BATCH=${2:-1024}; SKIP=${3:-115}; COUNT=${4:-2}
mkdir -p gpurun_out
timeout 300 python scripts/profile_unet.py --batch "$BATCH" --once > gpurun_out/plain_unet.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm_tc" -s "$SKIP" -c "$COUNT" \
    -f -o "gpurun_out/$NAME" python scripts/profile_unet.py --batch "$BATCH" --once > gpurun_out/ncu_unet.log 2>&1
echo "rc=$?"
tail -n 2 gpurun_out/plain_unet.log gpurun_out/ncu_unet.log
ncu -i "gpurun_out/$NAME.ncu-rep" --page raw --csv > "gpurun_out/$NAME.raw.csv" 2>/dev/null
ncu -i "gpurun_out/$NAME.ncu-rep" --page source --csv --print-source sass > "gpurun_out/$NAME.source.csv" 2>/dev/null
SZ=$(stat -c %s "gpurun_out/$NAME.ncu-rep" 2>/dev/null || echo 0)
if [ "$SZ" -gt 20000000 ]; then rm -f "gpurun_out/$NAME.ncu-rep"; fi
ls -la gpurun_out/$NAME.*
