#!/bin/bash
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -8) 2>&1 | tee gpurun_out/s7_pytest.log
timeout 900 python bench.py > gpurun_out/s7_bench.json 2> gpurun_out/s7_bench.err
tail -c 1200 gpurun_out/s7_bench.json; tail -5 gpurun_out/s7_bench.err
