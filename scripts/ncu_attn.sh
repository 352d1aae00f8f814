#!/bin/bash
# ncu --set full of ONE attention launch (64 episodes, 602 keys) of the given kernel variant:
#   bash scripts/ncu_attn.sh <variant 1|2|3> <name>
cd "$(dirname "$0")/.." || exit 1
V=${1:-3}; NAME=${2:-attn_v$V}
mkdir -p gpurun_out
timeout 120 python scripts/attn_probe.py "$V" --once > "gpurun_out/$NAME.plain.log" 2>&1 || { echo "plain run failed"; tail -5 "gpurun_out/$NAME.plain.log"; exit 1; }
cat "gpurun_out/$NAME.plain.log"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:attention" -s 3 -c 1 -f \
    -o "gpurun_out/$NAME" python scripts/attn_probe.py "$V" --once > "gpurun_out/$NAME.ncu.log" 2>&1
echo "ncu rc=$?"
ncu -i "gpurun_out/$NAME.ncu-rep" --page details > "gpurun_out/$NAME.details.txt" 2>/dev/null
ncu -i "gpurun_out/$NAME.ncu-rep" --page source --csv --print-source sass > "gpurun_out/$NAME.source.csv" 2>/dev/null
ncu -i "gpurun_out/$NAME.ncu-rep" --page raw --csv > "gpurun_out/$NAME.raw.csv" 2>/dev/null
python scripts/ncu_top_stalls.py "gpurun_out/$NAME.source.csv" 40 > "gpurun_out/$NAME.stalls.txt" 2>&1
rm -f "gpurun_out/$NAME.ncu-rep"
tail -45 "gpurun_out/$NAME.stalls.txt"
