#!/bin/bash
# A/B of the round-1r kernel changes on one B200: parity subset first, then bench lines with the poker step's
# L2 look-ahead off / at several distances, and the 2048 step (hashed table slots, fused reset path).
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp}
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_envs.py tests/test_gpu_poker_packed.py tests/test_gpu_edges.py tests/test_gpu_policy.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_lines.jsonl
for ahead in 0 -1 592 2368 4736; do
  for w in poker_hu poker_6max poker_hu_scripted; do
    if [ $ahead -ge 0 ]; then export PULSE_POKER_L2_AHEAD=$ahead; else unset PULSE_POKER_L2_AHEAD; fi
    echo "{\"ahead\": $ahead, \"workload\": \"$w\"}" >> $OUT/${TAG}_lines.jsonl
    timeout 300 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
  done
done
unset PULSE_POKER_L2_AHEAD
for i in 1 2; do
  timeout 300 python bench.py --workload tfe_4x4 --steps 512 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
done
exit 0
