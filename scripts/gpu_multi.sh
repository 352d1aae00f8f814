#!/bin/bash
# N-GPU check of bench.py under torchrun exactly as the driver launches it:  bash scripts/gpu_multi.sh N
cd "$(dirname "$0")/.." || exit 1
N=${1:-2}
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus "$N" --steps 10 --warmup 3 > "gpurun_out/scale_n${N}.json" 2> "gpurun_out/scale_n${N}.err"
echo "rc=$?"; tail -c 1500 "gpurun_out/scale_n${N}.err"; head -c 400 "gpurun_out/scale_n${N}.json"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus "$N" --steps 10 --warmup 3 --no-other-configs --sync-gather > "gpurun_out/scale_n${N}_syncgather.json" 2> "gpurun_out/scale_n${N}_syncgather.err"
echo "rc=$?"; head -c 300 "gpurun_out/scale_n${N}_syncgather.json"
