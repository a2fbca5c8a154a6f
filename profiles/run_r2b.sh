#!/bin/bash
# round 2, GPU call B: host-env parity tests, full default bench line (e2e through the host-buffer env), the all-live
# workload, ncu full captures of the auto-reset step / redeal pass / wire step
set -u
OUT=gpurun_out; TAG=r2b
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_host_env.py tests/test_gpu_envs.py tests/test_gpu_poker_packed.py -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -15 $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_bench_lines.jsonl
timeout 900 python bench.py --steps 64 --warmup 16 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_default.err; tail -3 $OUT/${TAG}_bench_default.err
for w in poker_hu_autoreset poker_hu_scripted; do
  timeout 400 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_$w.err
done
for ch in 4 16; do
  timeout 400 python bench.py --steps 32 --warmup 8 --no-extras --cpu-seconds 0 --e2e-chunks $ch >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_chunks.err
done
timeout 400 python bench.py --impl reference --steps 4 --warmup 3 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_ref.err
python - <<'PY'
import json
for l in open('gpurun_out/r2b_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    print(d['config']['workload'][:28], '%.3e'%d['value'], 'ms/step', d.get('ms_per_step'), 'frac', d.get('roofline',{}).get('frac'), 'e2e %.3e'%d['e2e']['value'], d['e2e'].get('h2d_bytes_per_step'), d['e2e'].get('d2h_bytes_per_step'), d['e2e'].get('chunks'), d.get('live_step_fraction'))
    for k,v in (d.get('workloads') or {}).items(): print('   ', k, v.get('value'), (v.get('roofline') or {}).get('frac'), v.get('live_step_fraction'), v.get('error'))
PY
cap() {  # name, kernel regex, skip, bench args...
  local name=$1 regex=$2 skip=$3; shift 3
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 \
    -o $OUT/${TAG}_${name} -f python bench.py "$@" --cpu-seconds 0 --no-extras > $OUT/${TAG}_ncu_${name}.log 2>&1
  ncu -i $OUT/${TAG}_${name}.ncu-rep --page raw --csv > $OUT/${TAG}_${name}_raw.csv 2>/dev/null
  ncu -i $OUT/${TAG}_${name}.ncu-rep --page source --csv --print-source cuda,sass > $OUT/${TAG}_${name}_source.csv 2>/dev/null
  gzip -f $OUT/${TAG}_${name}_source.csv
  rm -f $OUT/${TAG}_${name}.ncu-rep
}
cap autoreset_step poker_packed_step_kernel 40 --workload poker_hu_autoreset --steps 16 --warmup 3
cap autoreset_redeal poker_packed_redeal_kernel 40 --workload poker_hu_autoreset --steps 16 --warmup 3
cap host_step poker_host_step_kernel 30 --steps 16 --warmup 3
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv \
  --log-file $OUT/${TAG}_launches_autoreset_steps16.csv \
  python bench.py --workload poker_hu_autoreset --steps 16 --warmup 3 --cpu-seconds 0 > $OUT/${TAG}_ncu_launches.log 2>&1
ls -la $OUT | tail -30
exit 0
