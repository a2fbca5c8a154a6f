#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench lines of every workload, ncu launch list, ncu full captures.
# Usage (from the repo root, under gpurun):  bash profiles/run_gpu_round.sh <tag>
# Everything lands in gpurun_out/<tag>_*; the summaries worth judging are copied to profiles/ by hand.
set -u
TAG=${1:-rX}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
nproc > $OUT/${TAG}_nproc.txt

timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/${TAG}_smoke.log 2>&1
echo "smoke exit $?" >> $OUT/${TAG}_smoke.log

: > $OUT/${TAG}_bench_lines.jsonl
timeout 600 python bench.py >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_default.err
for w in poker_6max poker_10max tfe_4x4 tetris particle2d blackjack poker_hu_scripted poker_hu_train_loop; do
  timeout 400 python bench.py --workload $w --steps 512 --warmup 16 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_$w.err
done
timeout 400 python bench.py --impl reference --steps 8 --warmup 3 >> $OUT/${TAG}_bench_lines.jsonl 2> $OUT/${TAG}_bench_ref.err

# launch list of the bench command (cold-cache, serialised: only the kernel's share of the step is comparable)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file $OUT/${TAG}_launches_bench_steps16.csv \
  python bench.py --steps 16 --warmup 3 --cpu-seconds 0 > $OUT/${TAG}_ncu_launches.log 2>&1

# full captures of the step kernels (one launch each, after warm-up launches)
cap() {  # name, kernel regex, skip, bench args...
  local name=$1 regex=$2 skip=$3; shift 3
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 \
    -o $OUT/${TAG}_${name} -f python bench.py "$@" --cpu-seconds 0 --no-extras > $OUT/${TAG}_ncu_${name}.log 2>&1
  ncu -i $OUT/${TAG}_${name}.ncu-rep --page raw --csv > $OUT/${TAG}_${name}_raw.csv 2>/dev/null
  ncu -i $OUT/${TAG}_${name}.ncu-rep --page source --csv --print-source cuda,sass > $OUT/${TAG}_${name}_source.csv 2>/dev/null
  gzip -f $OUT/${TAG}_${name}_source.csv
  rm -f $OUT/${TAG}_${name}.ncu-rep   # gpurun_out/ is capped at 64 MiB: keep the CSV pages, not the report
}
cap poker_hu_step poker_packed_step_kernel 20 --steps 16 --warmup 3
cap poker_hu_reset poker_packed_reset 1 --steps 16 --warmup 3
cap poker_6max_step poker_packed_step_kernel 20 --workload poker_6max --steps 16 --warmup 3
cap tfe_step tfe_step_kernel 40 --workload tfe_4x4 --steps 64 --warmup 8
cap tetris_step tetris_step_kernel 40 --workload tetris --steps 64 --warmup 8
# the callers either side of the step (N1-N3): fused policy step, redeal pass, hand statistics, ring append
cap loop_policy_step poker_packed_step_kernel 12 --workload poker_hu_train_loop --steps 16 --warmup 3
cap loop_redeal poker_packed_redeal_kernel 12 --workload poker_hu_train_loop --steps 16 --warmup 3
cap loop_hand_stats hand_stats_kernel 12 --workload poker_hu_train_loop --steps 16 --warmup 3
cap loop_replay_write replay_write_kernel 12 --workload poker_hu_train_loop --steps 16 --warmup 3
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv \
  --log-file $OUT/${TAG}_launches_train_loop_steps16.csv \
  python bench.py --workload poker_hu_train_loop --steps 16 --warmup 3 --cpu-seconds 0 > $OUT/${TAG}_ncu_launches_loop.log 2>&1
timeout 300 python profiles/time_strength.py > $OUT/${TAG}_time_strength.json 2> $OUT/${TAG}_time_strength.err
ls -la $OUT > $OUT/${TAG}_ls.txt
exit 0
