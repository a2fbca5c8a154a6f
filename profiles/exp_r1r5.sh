#!/bin/bash
# resident-CTA choice for 6-max / 10-max now that the look-ahead hides most of the DRAM latency
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp5}
mkdir -p $OUT
: > $OUT/${TAG}_lines.jsonl
for minb in 8 10 12; do
  echo "{\"minb\": $minb, \"workload\": \"poker_6max\"}" >> $OUT/${TAG}_lines.jsonl
  PULSE_POKER_MINB=$minb timeout 300 python bench.py --workload poker_6max --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
for minb in 5 8 10; do
  for ahead in 148 74; do
    echo "{\"minb\": $minb, \"ahead\": $ahead, \"workload\": \"poker_10max\"}" >> $OUT/${TAG}_lines.jsonl
    PULSE_POKER_L2_AHEAD=$ahead PULSE_POKER_MINB=$minb timeout 300 python bench.py --workload poker_10max --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
  done
done
echo "{\"minb\": 12, \"workload\": \"poker_hu\"}" >> $OUT/${TAG}_lines.jsonl
PULSE_POKER_MINB=12 timeout 300 python bench.py --workload poker_hu --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
timeout 300 python -m pytest tests/test_gpu_poker_packed.py tests/test_gpu_edges.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1
exit 0
