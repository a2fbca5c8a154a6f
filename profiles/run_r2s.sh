#!/bin/bash
# round 2, GPU call S: on-device random-policy rollout at the headline size (SURVEY 8(b) pulse_poker_rollout): the
# registers-resident path against stepping through global memory, the rollout parity tests, and the headline line (the
# step body was factored out of packed_step_core for it: no regression allowed)
set -u
OUT=gpurun_out; TAG=r2s
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_policy.py tests/test_gpu_poker.py tests/test_gpu_handstats.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -4 $OUT/${TAG}_pytest.log | cut -c1-300
: > $OUT/${TAG}_bench_lines.jsonl
for r in 1 0; do
PULSE_POKER_ROLLOUT_REGS=$r timeout 400 python bench.py --workload poker_hu_rollout --steps 256 --warmup 32 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench.err
done
tail -5 $OUT/${TAG}_bench.err
timeout 400 python bench.py --steps 256 --warmup 32 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench.err
timeout 400 python bench.py --workload poker_hu_autoreset --steps 256 --warmup 32 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench.err
python - <<'PY'
import json
for l in open('gpurun_out/r2s_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d.get('e2e') or {}
    print(d['config']['workload'][:24], '%.3e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'kern', d['roofline'].get('kernel_ms'), 'frac %.3f'%d['roofline']['frac'], 'e2e %.3e'%e.get('value',0))
PY
exit 0
