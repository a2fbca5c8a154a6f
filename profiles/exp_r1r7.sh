#!/bin/bash
# prefetch issued by the first warp only; new resident-CTA defaults (6 seats: 6, 10 seats: 4); heads-up at 8 vs 10
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp7}
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_poker_packed.py tests/test_gpu_edges.py tests/test_gpu_policy.py tests/test_gpu_million.py tests/test_gpu_handstats.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_lines.jsonl
for w in poker_hu poker_6max poker_10max poker_hu_scripted poker_hu_train_loop; do
  echo "{\"workload\": \"$w\"}" >> $OUT/${TAG}_lines.jsonl
  timeout 300 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
for minb in 8 10; do
  for ahead in 148 296; do
    echo "{\"minb\": $minb, \"ahead\": $ahead, \"workload\": \"poker_hu\"}" >> $OUT/${TAG}_lines.jsonl
    PULSE_POKER_L2_AHEAD=$ahead PULSE_POKER_MINB=$minb timeout 300 python bench.py --workload poker_hu --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
  done
done
for ahead in 37 74 148; do
  echo "{\"ahead\": $ahead, \"workload\": \"poker_6max\"}" >> $OUT/${TAG}_lines.jsonl
  PULSE_POKER_L2_AHEAD=$ahead timeout 300 python bench.py --workload poker_6max --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
exit 0
