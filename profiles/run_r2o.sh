#!/bin/bash
# round 2, GPU call O (N GPUs): the driver's launch line at N ranks on the final build -- default bench flags (weak
# scaling, stats all-reduce, e2e with host threads split over ranks) and the reference arm
set -u
N=${1:-4}
OUT=gpurun_out; TAG=r2o
mkdir -p $OUT
nvidia-smi -L > $OUT/${TAG}_gpus_n$N.txt; nproc >> $OUT/${TAG}_gpus_n$N.txt; free -g | head -2 >> $OUT/${TAG}_gpus_n$N.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 256 --warmup 16 > $OUT/${TAG}_bench_n$N.jsonl 2> $OUT/${TAG}_bench_n$N.err
tail -c 3000 $OUT/${TAG}_bench_n$N.jsonl; tail -5 $OUT/${TAG}_bench_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 8 --warmup 3 --impl reference > $OUT/${TAG}_bench_ref_n$N.jsonl 2> $OUT/${TAG}_bench_ref_n$N.err
tail -c 500 $OUT/${TAG}_bench_ref_n$N.jsonl
exit 0
