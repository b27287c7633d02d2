#!/bin/bash
# Round 2, call A: the allocation-free entry point alone in a fresh process, the whole gpu suite, then both bench arms.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
echo "== noalloc alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | tail -5
echo "== gpu suite"; (time timeout 2400 python -m pytest tests -q -m gpu -x --durations=15) 2>&1 | tail -40
echo "== bench"; timeout 900 python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 600 gpurun_out/r2_bench_n1.err; head -c 1500 gpurun_out/r2_bench_n1.json
