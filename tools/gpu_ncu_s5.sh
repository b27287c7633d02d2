#!/bin/bash
# ncu --set full of the table-mode accumulate / reduce kernels out of the bench command (final geometry of the round).
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ntt --no-prove"
$CMD > gpurun_out/s16_bench_short.json 2> gpurun_out/s16_bench_short.err || { echo "bench failed"; exit 1; }
ncu --set full --clock-control none --import-source on --kernel-name regex:'k_accumulate|k_reduce_chunks|k_window_sum' \
    --launch-skip 9 --launch-count 6 -o gpurun_out/prof_r1s5_tables2 -f $CMD > gpurun_out/s16_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep
