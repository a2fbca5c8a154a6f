#!/bin/bash
# round 2, GPU call J: the static-scratch append test alone (fresh process), the whole GPU suite (-x, the new file last),
# smoke(), the default bench line and the reference arm
set -u
OUT=gpurun_out; TAG=r2j
mkdir -p $OUT
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/${TAG}_noalloc_alone.log 2>&1; echo "rc=$?" >> $OUT/${TAG}_noalloc_alone.log
tail -5 $OUT/${TAG}_noalloc_alone.log | cut -c1-300
timeout 600 python -m pytest tests/test_deck_repeat.py -m gpu -x -q > $OUT/${TAG}_deck.log 2>&1; echo "rc=$?" >> $OUT/${TAG}_deck.log
tail -5 $OUT/${TAG}_deck.log | cut -c1-300
timeout 2400 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -8 $OUT/${TAG}_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/${TAG}_smoke.log 2>&1; tail -2 $OUT/${TAG}_smoke.log | cut -c1-300
: > $OUT/${TAG}_bench_lines.jsonl
timeout 900 python bench.py >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_default.err; tail -3 $OUT/${TAG}_bench_default.err
timeout 600 python bench.py --impl reference >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_ref.err; tail -3 $OUT/${TAG}_bench_ref.err
python - <<'PY'
import json
for l in open('gpurun_out/r2j_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d.get('e2e') or {}
    print(d.get('impl','ours'), d['config']['workload'][:30], '%.3e'%d['value'], 'ms/step %.4f'%d.get('ms_per_step'), 'frac %.3f'%(d.get('roofline') or {}).get('frac',0), 'e2e %.3e'%e.get('value',0), 'rec %.3e'%(d.get('e2e_records') or {}).get('value',0), d.get('clocks'))
PY
exit 0
