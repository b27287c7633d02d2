#!/bin/bash
# Round 2, call Z11: overlay after the sampled B-index fingerprint: cache tests, entry-point timing.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_groth16_gpu.py -q -m gpu -x -k "overlay or caches or device_resident or proof_bit_exact" 2>&1 | grep -v "^\*\|^$" | tail -3
B200_OVERLAY_TIMES=1 timeout 900 python tools/overlay_stock_prove.py --log-m 20 --skip-cpu 2>&1 | grep "fused prover\|curve" | tail -3 | cut -c1-330
