#!/bin/bash
# round 2, GPU call AH: kernel events around every 5th launch (default) -- the default line and two eager workloads
set -u
OUT=gpurun_out; TAG=r2ah
mkdir -p $OUT
: > $OUT/${TAG}_bench_lines.jsonl
timeout 900 python bench.py >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench.err; tail -2 $OUT/${TAG}_bench.err
for w in poker_hu_autoreset poker_6max tfe_4x4; do
timeout 400 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r2ah_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d.get('e2e') or {}; r=d['roofline']
    print(d['config']['workload'][:22], '%.4e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'kern %.4f'%r['kernel_ms'], 'frac %.3f'%r['frac'], r.get('kernel_time','')[:60], 'e2e %.3e'%e.get('value',0))
PY
exit 0
