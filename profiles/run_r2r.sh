#!/bin/bash
# round 2, GPU call R: final build -- the whole GPU suite (-x), smoke(), the default bench line and the reference arm
set -u
OUT=gpurun_out; TAG=r2r
mkdir -p $OUT
timeout 2400 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -6 $OUT/${TAG}_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/${TAG}_smoke.log 2>&1; tail -2 $OUT/${TAG}_smoke.log | cut -c1-300
: > $OUT/${TAG}_bench_lines.jsonl
( time timeout 900 python bench.py >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_default.err ) 2>&1 | grep real; tail -3 $OUT/${TAG}_bench_default.err
( time timeout 600 python bench.py --impl reference >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_ref.err ) 2>&1 | grep real; tail -3 $OUT/${TAG}_bench_ref.err
python - <<'PY'
import json
for l in open('gpurun_out/r2r_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d.get('e2e') or {}
    r=d.get('roofline') or {}
    print(d.get('impl','ours'), d['config']['workload'][:30], '%.3e'%d['value'], 'ms/step %.4f'%d.get('ms_per_step'), 'frac %.3f'%r.get('frac',0), 'by traffic', r.get('frac_by_traffic'), 'e2e %.3e'%e.get('value',0), 'rec %.3e'%(d.get('e2e_records') or {}).get('value',0), d.get('clocks'))
    for k,v in (d.get('workloads') or {}).items(): print('   ', k, '%.3e'%v.get('value',0) if 'value' in v else v, (v.get('roofline') or {}).get('frac'), (v.get('e2e') or {}).get('value'))
PY
exit 0
