#!/bin/bash
# ncu --set full capture of the MSM kernels (and one NTT pass) out of the bench command.
set -u
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pre_ncu.json 2>&1 || { echo "bench failed"; exit 1; }
ncu --set full --clock-control none --import-source on \
    --kernel-name regex:'k_accumulate|k_reduce_chunks|k_scatter|k_histogram|k_ntt_pass' \
    --launch-skip 24 --launch-count 10 \
    -o gpurun_out/prof_r1_full -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/*.ncu-rep
