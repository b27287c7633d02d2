#!/bin/bash
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests/test_msm_gpu.py tests/test_groth16_gpu.py -q -m gpu -x 2>&1 | grep -v "^\*\|^  \*\|^(\|^$" | tail -4) | tee gpurun_out/s15_pytest.log
timeout 600 python tools/msm_probe.py alt_bn128 bls12_377:1 2>&1 | grep -E "^(alt|bls)|iter 3|==|parity" | tee gpurun_out/s15_msm_probe.log
timeout 600 python tools/msm_sweep.py --curves alt_bn128 --min-log 14 --max-log 18 --ref-max-log 0 --tables 2>&1 | cut -c1-420 | tee gpurun_out/s15_small_sweep.log
