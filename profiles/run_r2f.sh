#!/bin/bash
# round 2, GPU call F: full GPU suite (fixed expander), default bench, rollouts
set -u
OUT=gpurun_out; TAG=r2f
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -6 $OUT/${TAG}_pytest.log | cut -c1-300
: > $OUT/${TAG}_bench_lines.jsonl
timeout 900 python bench.py --steps 64 --warmup 16 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_default.err; tail -3 $OUT/${TAG}_bench_default.err
for w in blackjack blackjack_per_step tfe_4x4_rollout; do
  timeout 400 python bench.py --workload $w --steps 4096 --warmup 128 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_$w.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r2f_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d['e2e']
    print(d['config']['workload'][:30], '%.3e'%d['value'], 'ms/step %.4f'%d.get('ms_per_step'), 'kern %.4f'%d.get('roofline',{}).get('kernel_ms',0), 'frac %.3f'%d.get('roofline',{}).get('frac',0), 'e2e %.3e'%e['value'], e.get('host_threads'), e.get('chunks'), e.get('host_ms_per_step'), e.get('d2h_bytes_per_step'), d.get('live_step_fraction'))
    for k,v in (d.get('workloads') or {}).items(): print('   ', k, v.get('value'), (v.get('roofline') or {}).get('frac'), v.get('live_step_fraction'), v.get('error'), (v.get('e2e') or {}).get('value'))
PY
exit 0
