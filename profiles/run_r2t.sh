#!/bin/bash
# round 2, GPU call T: source-level ncu capture of the whole-batch reset (the deal)
set -u
OUT=gpurun_out; TAG=r2t
mkdir -p $OUT
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poker_packed_reset_kernel -s 3 -c 1 \
  -o $OUT/${TAG}_reset -f python bench.py --steps 32 --warmup 3 --cpu-seconds 0 --no-extras > $OUT/${TAG}_ncu_reset.log 2>&1
ncu -i $OUT/${TAG}_reset.ncu-rep --page raw --csv > $OUT/${TAG}_reset_raw.csv 2>/dev/null
ncu -i $OUT/${TAG}_reset.ncu-rep --page source --csv --print-source cuda,sass > $OUT/${TAG}_reset_source.csv 2>/dev/null
gzip -f $OUT/${TAG}_reset_source.csv
rm -f $OUT/${TAG}_reset.ncu-rep
ls -la $OUT | tail -4
exit 0
