#!/bin/bash
# GPU contact: parity tests + a timing probe. Run via gpurun from the repo root.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
timeout 900 python -m pytest tests/test_msm_gpu.py -x -q -m gpu 2>&1 | tail -40
timeout 900 python -m pytest tests/test_ntt_gpu.py -x -q -m gpu 2>&1 | tail -40
timeout 600 python tools/msm_probe.py 2>&1 | tee gpurun_out/msm_probe.log
