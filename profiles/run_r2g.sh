#!/bin/bash
# round 2, GPU call G (2 GPUs): the driver's launch line at N=2 -- weak scaling, the stats all-reduce, e2e with host threads split over ranks
set -u
OUT=gpurun_out; TAG=r2g
mkdir -p $OUT
nvidia-smi -L > $OUT/${TAG}_gpus.txt; nproc >> $OUT/${TAG}_gpus.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 64 --warmup 16 > $OUT/${TAG}_bench_n2.jsonl 2> $OUT/${TAG}_bench_n2.err
tail -c 2500 $OUT/${TAG}_bench_n2.jsonl; tail -5 $OUT/${TAG}_bench_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --steps 8 --warmup 3 --impl reference > $OUT/${TAG}_bench_ref_n2.jsonl 2> $OUT/${TAG}_bench_ref_n2.err
tail -c 600 $OUT/${TAG}_bench_ref_n2.jsonl
python bench.py --gpus 1 --steps 64 --warmup 16 --no-extras --cpu-seconds 0 > $OUT/${TAG}_bench_n1.jsonl 2> $OUT/${TAG}_bench_n1.err
tail -c 1500 $OUT/${TAG}_bench_n1.jsonl
tools/_build/hostlink_probe > $OUT/${TAG}_hostlink_2gpu.jsonl 2>&1; grep -E "d2h|h2d|host_expand|info" $OUT/${TAG}_hostlink_2gpu.jsonl | tail -24
exit 0
