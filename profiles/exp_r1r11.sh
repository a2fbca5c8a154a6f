#!/bin/bash
# ten seats with the ordered loads (42 values live at once): resident CTAs 3 / 4 / 5
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp11}
mkdir -p $OUT
: > $OUT/${TAG}_lines.jsonl
for minb in 3 4 5; do
  echo "{\"minb\": $minb, \"workload\": \"poker_10max\"}" >> $OUT/${TAG}_lines.jsonl
  PULSE_POKER_MINB=$minb timeout 300 python bench.py --workload poker_10max --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
for minb in 5 6 8; do
  echo "{\"minb\": $minb, \"workload\": \"poker_6max\"}" >> $OUT/${TAG}_lines.jsonl
  PULSE_POKER_MINB=$minb timeout 300 python bench.py --workload poker_6max --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
exit 0
