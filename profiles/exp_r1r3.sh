#!/bin/bash
# 2048 step: adaptive (compute rows until the table copy lands) vs waiting; poker look-ahead with / without ev words
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp3}
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_envs.py tests/test_gpu_edges.py tests/test_gpu_million.py -m gpu -x -q -k "tfe or TFE or 2048" > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
PULSE_TFE_ADAPT=0 timeout 600 python -m pytest tests/test_gpu_envs.py tests/test_gpu_edges.py -m gpu -x -q -k "tfe or TFE or 2048" >> $OUT/${TAG}_pytest.log 2>&1
echo "pytest (no adapt) exit $?" >> $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_lines.jsonl
for ad in 0 1 0 1; do
  echo "{\"adapt\": $ad}" >> $OUT/${TAG}_lines.jsonl
  PULSE_TFE_ADAPT=$ad timeout 300 python bench.py --workload tfe_4x4 --steps 512 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
for ev in 0 1; do
  for w in poker_hu poker_6max; do
    echo "{\"ev\": $ev, \"workload\": \"$w\"}" >> $OUT/${TAG}_lines.jsonl
    PULSE_POKER_L2_EV=$ev timeout 300 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
  done
done
for w in poker_hu_scripted poker_hu_train_loop; do
  echo "{\"workload\": \"$w\"}" >> $OUT/${TAG}_lines.jsonl
  timeout 300 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
exit 0
