#!/bin/bash
# round 2, GPU call H: records-only host step, poker rollout, wide-row direct stores A/B (6 / 10 seats)
set -u
OUT=gpurun_out; TAG=r2h
mkdir -p $OUT
timeout 1200 python -m pytest tests/test_gpu_host_env.py tests/test_gpu_policy.py tests/test_gpu_envs.py tests/test_gpu_multirank.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -6 $OUT/${TAG}_pytest.log | cut -c1-300
: > $OUT/${TAG}_bench_lines.jsonl
timeout 900 python bench.py --steps 64 --warmup 16 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_default.err; tail -3 $OUT/${TAG}_bench_default.err
for w in poker_6max poker_10max; do
  timeout 400 python bench.py --workload $w --steps 48 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_$w.err
  PULSE_POKER_OBS_DIRECT_WIDE=1 timeout 400 python bench.py --workload $w --steps 48 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_wide.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r2h_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d['e2e']
    print(d['config']['workload'][:30], '%.3e'%d['value'], 'ms/step %.4f'%d.get('ms_per_step'), 'kern %.4f'%d.get('roofline',{}).get('kernel_ms',0), 'frac %.3f'%d.get('roofline',{}).get('frac',0), 'e2e %.3e'%e['value'], 'rec %.3e'%d.get('e2e_records',{}).get('value',0), e.get('host_threads'), e.get('host_ms_per_step'), e.get('d2h_bytes_per_step'))
PY
exit 0
