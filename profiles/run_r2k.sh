#!/bin/bash
# round 2, GPU call K: showdown equity (N4) parity + timing
set -u
OUT=gpurun_out; TAG=r2k
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_strength.py -m gpu -x -q -s > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/${TAG}_pytest.log
grep -n "showdown_equity:" $OUT/${TAG}_pytest.log; tail -6 $OUT/${TAG}_pytest.log | cut -c1-300
exit 0
