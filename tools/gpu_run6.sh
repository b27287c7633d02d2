#!/bin/bash
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests/test_msm_gpu.py tests/test_groth16_gpu.py -q -m gpu -x 2>&1 | tail -15) 2>&1 | tee gpurun_out/s6_pytest.log
timeout 900 python tools/msm_probe.py 2>&1 | tee gpurun_out/s6_msm_probe.log
