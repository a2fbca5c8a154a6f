#!/bin/bash
# Weak-scaling lines of the headline workloads under torchrun (one rank per GPU): bash profiles/run_scale.sh <N> <tag>
set -u
N=${1:-2}; TAG=${2:-rX}; OUT=gpurun_out
mkdir -p $OUT
: > $OUT/${TAG}_scale${N}_bench_lines.jsonl
for w in poker_hu poker_6max tfe_4x4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --workload $w --steps 512 --warmup 16 --cpu-seconds 0 --no-extras \
    >> $OUT/${TAG}_scale${N}_bench_lines.jsonl 2>> $OUT/${TAG}_scale${N}.err
done
exit 0
