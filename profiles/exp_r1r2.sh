#!/bin/bash
# finer scan of the poker step's L2 look-ahead + one full ncu capture of the 2048 step with the source page
set -u
OUT=gpurun_out; TAG=${1:-r1r_exp2}
mkdir -p $OUT
: > $OUT/${TAG}_lines.jsonl
for ahead in 74 148 296 444 592 888; do
  for w in poker_hu poker_6max; do
    export PULSE_POKER_L2_AHEAD=$ahead
    echo "{\"ahead\": $ahead, \"workload\": \"$w\"}" >> $OUT/${TAG}_lines.jsonl
    timeout 300 python bench.py --workload $w --steps 256 --warmup 16 --cpu-seconds 0 --no-extras >> $OUT/${TAG}_lines.jsonl 2>> $OUT/${TAG}_err.log
  done
done
unset PULSE_POKER_L2_AHEAD
name=tfe_step
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tfe_step_kernel -s 40 -c 1 \
  -o $OUT/${TAG}_${name} -f python bench.py --workload tfe_4x4 --steps 64 --warmup 8 --cpu-seconds 0 --no-extras > $OUT/${TAG}_ncu_${name}.log 2>&1
ncu -i $OUT/${TAG}_${name}.ncu-rep --page raw --csv > $OUT/${TAG}_${name}_raw.csv 2>/dev/null
ncu -i $OUT/${TAG}_${name}.ncu-rep --page source --csv --print-source cuda,sass > $OUT/${TAG}_${name}_source.csv 2>/dev/null
gzip -f $OUT/${TAG}_${name}_source.csv
rm -f $OUT/${TAG}_${name}.ncu-rep
exit 0
