#!/bin/bash
# round 2, GPU call E: tile-local redeal inside the step kernel (A/B against the separate pass), host env with pinned
# epoch flags + overlapped submission, ncu of the fused auto-reset step
set -u
OUT=gpurun_out; TAG=r2e
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
tail -6 $OUT/${TAG}_pytest.log
: > $OUT/${TAG}_bench_lines.jsonl
timeout 900 python bench.py --steps 64 --warmup 16 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_default.err; tail -3 $OUT/${TAG}_bench_default.err
for w in poker_hu_autoreset poker_hu_scripted poker_hu_train_loop; do
  timeout 400 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_$w.err
  PULSE_POKER_REDEAL_PASS=1 timeout 400 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_ab.err
done
for th in 4 8; do
  timeout 400 python bench.py --steps 32 --warmup 8 --no-extras --cpu-seconds 0 --e2e-threads $th >> $OUT/${TAG}_bench_lines.jsonl 2>> $OUT/${TAG}_bench_threads.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r2e_bench_lines.jsonl'):
    try: d=json.loads(l)
    except Exception: continue
    e=d['e2e']
    print(d['config']['workload'][:30], '%.3e'%d['value'], 'ms/step %.4f'%d.get('ms_per_step'), 'kern %.4f'%d.get('roofline',{}).get('kernel_ms',0), 'frac %.3f'%d.get('roofline',{}).get('frac',0), 'e2e %.3e'%e['value'], e.get('host_threads'), e.get('chunks'), e.get('host_ms_per_step'), e.get('d2h_bytes_per_step'), d.get('live_step_fraction'))
    for k,v in (d.get('workloads') or {}).items(): print('   ', k, v.get('value'), (v.get('roofline') or {}).get('frac'), v.get('live_step_fraction'), v.get('error'), (v.get('e2e') or {}).get('value'))
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
cap poker_hu_step poker_packed_step_kernel 20 --steps 16 --warmup 3
exit 0
