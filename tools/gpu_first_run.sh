#!/bin/bash
# First GPU contact: MSM parity tests + a timing probe. Run via gpurun from the repo root.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests/test_msm_gpu.py -x -q -m gpu 2>&1 | tail -30
python tools/msm_probe.py 2>&1 | tee gpurun_out/msm_probe.log
