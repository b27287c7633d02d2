#!/bin/bash
# Round 2, call W: the reference's prover entry point specialised in the overlay (zero-source-change path).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_groth16_gpu.py -q -m gpu -x -k "proof_bit_exact or device_resident or caches or interposed" 2>&1 | grep -v "^\* \|^$" | tail -5
B200_OVERLAY_TIMES=1 timeout 900 python tools/overlay_stock_prove.py --log-m 20 --skip-cpu 2>&1 | grep "fused prover" | tail -2 | cut -c1-260
timeout 600 python tools/overlay_stock_prove.py --log-m 16 --real-keys 2>/dev/null | tail -1 | tee gpurun_out/r2_overlay_stock_prove.jsonl
timeout 900 python tools/overlay_stock_prove.py --log-m 20 2>/dev/null | tail -1 | tee -a gpurun_out/r2_overlay_stock_prove.jsonl
