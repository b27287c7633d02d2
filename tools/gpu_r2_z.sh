#!/bin/bash
# Round 2, call Z14: chain-by-chain witness map on step and extended domains.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_groth16_gpu.py -q -m gpu -x -k "chain_parallel or sharded or step or extended" 2>&1 | grep -v "^\*\|^$" | tail -6
