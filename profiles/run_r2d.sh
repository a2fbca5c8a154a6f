#!/bin/bash
# round 2, GPU call D: new wire format + vectorised expander, rollouts (2048, Blackjack), e2e thread sweep
set -u
OUT=gpurun_out; TAG=r2d
mkdir -p $OUT
timeout 1200 python -m pytest tests/test_gpu_host_env.py tests/test_gpu_envs.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -6 $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_bench_lines.jsonl
timeout 900 python bench.py --steps 64 --warmup 16 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_default.err; tail -3 $OUT/${TAG}_bench_default.err
for w in tfe_4x4 tfe_4x4_rollout blackjack blackjack_per_step; do
  timeout 400 python bench.py --workload $w --steps 512 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_$w.err
done
for th in 2 4 8 12; do
  timeout 400 python bench.py --steps 32 --warmup 8 --no-extras --cpu-seconds 0 --e2e-threads $th >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_threads.err
done
for ch in 4 16 32; do
  timeout 400 python bench.py --steps 32 --warmup 8 --no-extras --cpu-seconds 0 --e2e-chunks $ch >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_chunks.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r2d_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d['e2e']
    print(d['config']['workload'][:30], '%.3e'%d['value'], 'ms/step %.4f'%d.get('ms_per_step'), 'kern %.4f'%d.get('roofline',{}).get('kernel_ms',0), 'frac %.3f'%d.get('roofline',{}).get('frac',0), d.get('roofline',{}).get('frac_by_moved_bytes'), 'e2e %.3e'%e['value'], e.get('host_threads'), e.get('chunks'), e.get('host_ms_per_step'), e.get('d2h_bytes_per_step'))
    for k,v in (d.get('workloads') or {}).items(): print('   ', k, v.get('value'), (v.get('roofline') or {}).get('frac'), v.get('live_step_fraction'), v.get('error'), (v.get('e2e') or {}).get('value'))
PY
exit 0
