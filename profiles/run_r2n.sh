#!/bin/bash
# round 2, GPU call N: source-level ncu capture of the step kernel (all-live workload and headline workload)
set -u
OUT=gpurun_out; TAG=r2n
mkdir -p $OUT
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
ls -la $OUT | tail -8
exit 0
