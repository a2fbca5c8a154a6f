#!/bin/bash
# fewer, fatter CTAs for the 6- and 10-seat step (register spill vs occupancy), with the look-ahead scaled to the resident set
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp6}
mkdir -p $OUT
: > $OUT/${TAG}_lines.jsonl
for minb in 5 6 8; do
  for ahead in 74 148; do
    echo "{\"minb\": $minb, \"ahead\": $ahead, \"workload\": \"poker_6max\"}" >> $OUT/${TAG}_lines.jsonl
    PULSE_POKER_L2_AHEAD=$ahead PULSE_POKER_MINB=$minb timeout 300 python bench.py --workload poker_6max --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
  done
done
for minb in 3 4 5; do
  for ahead in 37 74; do
    echo "{\"minb\": $minb, \"ahead\": $ahead, \"workload\": \"poker_10max\"}" >> $OUT/${TAG}_lines.jsonl
    PULSE_POKER_L2_AHEAD=$ahead PULSE_POKER_MINB=$minb timeout 300 python bench.py --workload poker_10max --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
  done
done
for minb in 5 6; do
PULSE_POKER_MINB=$minb timeout 300 python -m pytest tests/test_gpu_poker_packed.py tests/test_gpu_edges.py -m gpu -x -q >> $OUT/${TAG}_pytest.log 2>&1
done
exit 0
