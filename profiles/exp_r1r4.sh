#!/bin/bash
# 2048 step under programmatic dependent launch vs plain launches; train-loop / scripted with and without the L2 look-ahead
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp4}
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_envs.py tests/test_gpu_edges.py tests/test_gpu_million.py -m gpu -x -q -k "tfe or TFE or 2048" > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_lines.jsonl
for pdl in 0 1 0 1; do
  echo "{\"pdl\": $pdl}" >> $OUT/${TAG}_lines.jsonl
  PULSE_PDL=$pdl timeout 300 python bench.py --workload tfe_4x4 --steps 512 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
for ahead in 0 296 0 296; do
  for w in poker_hu_train_loop; do
    echo "{\"ahead\": $ahead, \"workload\": \"$w\"}" >> $OUT/${TAG}_lines.jsonl
    PULSE_POKER_L2_AHEAD=$ahead timeout 300 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
  done
done
exit 0
