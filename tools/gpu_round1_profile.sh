#!/bin/bash
# Round-1 evidence run: GPU tests, both bench arms, then the ncu launch list of the same bench command.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r1.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu_r1.log
python bench.py > gpurun_out/bench_r1.json 2> gpurun_out/bench_r1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_r1.json 2> gpurun_out/bench_ref_r1.err; echo "ref rc=$?"
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r1_short.json 2>&1; rc=$?; echo "short rc=$rc"
if [ $rc -eq 0 ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv \
      --log-file gpurun_out/launches_r1_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline \
      > gpurun_out/ncu_launches.log 2>&1; echo "ncu rc=$?"
fi
