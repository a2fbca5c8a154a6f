#!/bin/bash
# round 2, GPU call Y: host expander writes whole reward / done groups (no read-for-ownership) -- parity, then e2e A/B by thread count
set -u
OUT=gpurun_out; TAG=r2y
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_host_env.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -4 $OUT/${TAG}_pytest.log | cut -c1-300
: > $OUT/${TAG}_bench_lines.jsonl
for i in 1 2 3; do
timeout 400 python bench.py --steps 64 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench.err
done
timeout 400 python bench.py --steps 64 --warmup 16 --cpu-seconds 0 --no-extras --e2e-threads 8 >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench.err
python - <<'PY'
import json
for l in open('gpurun_out/r2y_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d.get('e2e') or {}
    print(d['config']['workload'][:12], '%.3e'%d['value'], 'e2e %.3e'%e.get('value',0), 'rec %.3e'%(d.get('e2e_records') or {}).get('value',0), e.get('host_threads'), e.get('host_ms_per_step'))
PY
exit 0
