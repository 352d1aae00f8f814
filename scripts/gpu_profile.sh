#!/bin/bash
# One gpurun call for the measured evidence of a round: the bench line, the per-launch CUDA-event table, the ncu launch
# list (times + DRAM bytes) of one forward, and `ncu --set full` summaries of the named kernels. Everything is bounded
# by `timeout`; results land in gpurun_out/.  bash scripts/gpu_profile.sh <tag> [full-kernel-regex ...]
cd "$(dirname "$0")/.." || exit 1
TAG=${1:-r02}; shift
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 > "gpurun_out/${TAG}_bench.json" 2> "gpurun_out/${TAG}_bench.err"
echo "bench rc=$?"; tail -c 600 "gpurun_out/${TAG}_bench.err"
timeout 300 python scripts/profile_step.py --batch 256 --out "gpurun_out/${TAG}_launches.json" > "gpurun_out/${TAG}_act_c2_launches_cuda_events.txt" 2>&1
echo "profile_step rc=$?"
timeout 300 python scripts/profile_step.py --batch 256 --once > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file "gpurun_out/${TAG}_ncu_launches.csv" python scripts/profile_step.py --batch 256 --once > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
for SPEC in "$@"; do
  REGEX=${SPEC%%:*}; SKIP=${SPEC##*:}
  NAME="${TAG}_$(echo "$REGEX" | tr -c 'a-zA-Z0-9' '_')"
  timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$REGEX" -s "$SKIP" -c 1 -f -o "gpurun_out/$NAME" \
      python scripts/profile_step.py --batch 64 --once > "gpurun_out/$NAME.log" 2>&1
  echo "ncu full $REGEX rc=$?"
  ncu -i "gpurun_out/$NAME.ncu-rep" --page raw --csv > "gpurun_out/$NAME.raw.csv" 2>/dev/null
  ncu -i "gpurun_out/$NAME.ncu-rep" --page details > "gpurun_out/$NAME.details.txt" 2>/dev/null
  ncu -i "gpurun_out/$NAME.ncu-rep" --page source --csv > "gpurun_out/$NAME.source.csv" 2>/dev/null
  rm -f "gpurun_out/$NAME.ncu-rep"
done
ls -la gpurun_out | head -40
