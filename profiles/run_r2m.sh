#!/bin/bash
# round 2, GPU call M: the reference's own benchmark size (256 tables) through the drop-in class and the packed
# engine (per-call host cost), and the launch list of the default bench command
set -u
OUT=gpurun_out; TAG=r2m
mkdir -p $OUT
: > $OUT/${TAG}_bench_lines.jsonl
for w in poker_hu_dropin poker_hu; do
  timeout 300 python bench.py --workload $w --tables 256 --steps 512 --warmup 64 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_${w}_256.err; tail -2 $OUT/${TAG}_bench_${w}_256.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r2m_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d.get('e2e') or {}
    print(d['config']['workload'][:40], d['config'].get('tables_per_gpu'), '%.3e'%d['value'], 'ms/step %.4f'%d.get('ms_per_step'), 'e2e %.3e'%e.get('value',0))
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches_default_bench.csv python bench.py --steps 32 --warmup 8 --cpu-seconds 0 --no-extras > $OUT/${TAG}_ncu.log 2>&1
grep -c "" $OUT/${TAG}_launches_default_bench.csv
exit 0
