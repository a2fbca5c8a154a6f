#!/bin/bash
# poker step: ordered (volatile) first-level loads, table word last
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp10}
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_poker_packed.py tests/test_gpu_edges.py tests/test_gpu_policy.py tests/test_gpu_handstats.py tests/test_gpu_million.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_lines.jsonl
for w in poker_hu poker_6max poker_10max poker_hu_scripted poker_hu_train_loop; do
  timeout 300 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
exit 0
