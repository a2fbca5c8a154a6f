#!/bin/bash
# round 2, GPU call A: full GPU test suite, smoke, default bench + drop-in workload, host link probe
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; tail -2 gpurun_out/r2a_smoke.log
python bench.py --steps 64 --warmup 16 --no-extras > gpurun_out/r2a_bench_hu.jsonl 2> gpurun_out/r2a_bench_hu.err; tail -c 1500 gpurun_out/r2a_bench_hu.jsonl
python bench.py --steps 64 --warmup 16 --workload poker_hu_dropin --no-extras --cpu-seconds 0 > gpurun_out/r2a_bench_dropin.jsonl 2> gpurun_out/r2a_bench_dropin.err; tail -c 1500 gpurun_out/r2a_bench_dropin.jsonl
tools/_build/hostlink_probe > gpurun_out/r2a_hostlink.jsonl 2>&1; tail -30 gpurun_out/r2a_hostlink.jsonl
nproc; lscpu | head -20 > gpurun_out/r2a_lscpu.txt; free -g | head -2 >> gpurun_out/r2a_lscpu.txt
