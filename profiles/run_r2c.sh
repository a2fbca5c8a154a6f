#!/bin/bash
# round 2, GPU call C: full GPU suite on the reworked step (direct 256-bit rows, closed-form heads-up showdown, branch-light
# execute), host-env pipeline timing, A/B of the direct row stores, Blackjack rollout
set -u
OUT=gpurun_out; TAG=r2c
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -6 $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_bench_lines.jsonl
timeout 900 python bench.py --steps 64 --warmup 16 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_default.err; tail -3 $OUT/${TAG}_bench_default.err
for w in poker_hu_autoreset poker_hu_dropin blackjack blackjack_per_step; do
  timeout 400 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_$w.err
done
PULSE_POKER_OBS_DIRECT=0 timeout 400 python bench.py --workload poker_hu_autoreset --steps 256 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_ab.err
PULSE_POKER_OBS_DIRECT=0 timeout 400 python bench.py --steps 64 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_ab.err
PULSE_POKER_MINB=10 timeout 400 python bench.py --steps 64 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_ab.err
for th in 4 8; do
  timeout 400 python bench.py --steps 32 --warmup 8 --no-extras --cpu-seconds 0 --e2e-threads $th >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_threads.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r2c_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d['e2e']
    print(d['config']['workload'][:30], '%.3e'%d['value'], 'ms/step %.4f'%d.get('ms_per_step'), 'kern %.4f'%d.get('roofline',{}).get('kernel_ms',0), 'frac %.3f'%d.get('roofline',{}).get('frac',0), 'e2e %.3e'%e['value'], e.get('host_threads'), e.get('host_ms_per_step'), d.get('live_step_fraction'))
    for k,v in (d.get('workloads') or {}).items(): print('   ', k, v.get('value'), (v.get('roofline') or {}).get('frac'), v.get('live_step_fraction'), v.get('error'), (v.get('e2e') or {}).get('value'))
PY
exit 0
