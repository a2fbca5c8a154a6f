#!/bin/bash
# redeal pass with batched counter loads + warp-aggregated append
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp8}
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_poker_packed.py tests/test_gpu_edges.py tests/test_gpu_policy.py tests/test_gpu_handstats.py tests/test_gpu_replay.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_lines.jsonl
for w in poker_hu_scripted poker_hu_train_loop; do
  timeout 300 python bench.py --workload $w --steps 512 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv \
  --log-file $OUT/${TAG}_launches_train_loop_steps16.csv \
  python bench.py --workload poker_hu_train_loop --steps 16 --warmup 3 --cpu-seconds 0 > $OUT/${TAG}_ncu_launches_loop.log 2>&1
exit 0
