#!/bin/bash
# 2048 / Tetris: prefetched scalars carried raw (no consumer right behind the load)
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp9}
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_envs.py tests/test_gpu_edges.py tests/test_gpu_million.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_lines.jsonl
for i in 1 2; do
for w in tfe_4x4 tetris; do
  timeout 300 python bench.py --workload $w --steps 512 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
done
exit 0
