#!/bin/bash
# Round 2, call M: full gpu suite, then both bench arms on one GPU (the numbers of this tree).
mkdir -p gpurun_out
(time timeout 2400 python -m pytest tests -q -m gpu -x) 2>&1 | tail -6
timeout 1200 python bench.py > gpurun_out/r2b_bench_n1.json 2> gpurun_out/r2b_bench_n1.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r2b_bench_n1.err
timeout 900 python bench.py --impl reference > gpurun_out/r2b_bench_reference_n1.json 2> gpurun_out/r2b_bench_reference_n1.err; echo "ref rc=$?"
head -c 600 gpurun_out/r2b_bench_n1.json
