#!/bin/bash
# Standard GPU check: all gpu-marked tests, then the MSM timing probe.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -15
timeout 600 python tools/msm_probe.py 2>&1 | tee gpurun_out/msm_probe.log
