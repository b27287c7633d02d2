#!/bin/bash
# Round-1 session-5 check on one GPU: the whole gpu-marked suite, then the default bench line.
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -15) 2>&1 | tee gpurun_out/s5_pytest.log
timeout 900 python bench.py > gpurun_out/s5_bench.json 2> gpurun_out/s5_bench.err
tail -c 3000 gpurun_out/s5_bench.json
tail -5 gpurun_out/s5_bench.err
