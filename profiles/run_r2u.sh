#!/bin/bash
# round 2, GPU call U: table-driven hand evaluator -- the whole GPU suite, then every workload that deals
set -u
OUT=gpurun_out; TAG=r2u
mkdir -p $OUT
timeout 2400 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log | cut -c1-300; grep -h "showdown_equity:" $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_bench_lines.jsonl
timeout 600 python bench.py --steps 512 --warmup 32 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench.err; tail -2 $OUT/${TAG}_bench.err
for w in poker_hu_train_loop poker_6max poker_10max; do
timeout 400 python bench.py --workload $w --steps 128 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r2u_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d.get('e2e') or {}
    print(d['config']['workload'][:24], '%.3e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'kern %.4f'%d['roofline'].get('kernel_ms',0), 'frac %.3f'%d['roofline']['frac'], 'e2e %.3e'%e.get('value',0))
    for k,v in (d.get('workloads') or {}).items(): print('   ', k, '%.3e'%v.get('value',0) if 'value' in v else v, 'ms %.4f'%v.get('ms_per_step',0), (v.get('roofline') or {}).get('frac'))
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $OUT/${TAG}_launches.csv python bench.py --workload poker_hu_autoreset --steps 16 --warmup 3 --cpu-seconds 0 > $OUT/${TAG}_ncu.log 2>&1
grep "reset_kernel\|redeal_kernel" $OUT/${TAG}_launches.csv | awk -F'","' '{print $5, $NF}' | tail -4 | cut -c1-200
exit 0
