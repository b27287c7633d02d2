#!/bin/bash
# Round 2, call T: extended_radix2_domain (new), the step functions of the sharded prover on one GPU, then the
# extended domain at its real size (alt_bn128, 2^29 elements).
mkdir -p gpurun_out
free -g | head -2
echo "== new tests"
timeout 1500 python -m pytest tests/test_ntt_gpu.py tests/test_groth16_gpu.py -q -m gpu -x -k "extended or chain_parallel or step" --durations=8 2>&1 | tail -25
echo "== extended domain, 2^22 (two-adicity lowered to 21)"
timeout 300 python tools/ext_fullsize.py 22 2>&1 | tail -2 | tee gpurun_out/r2_ext_2p22.json
echo "== extended domain, 2^29 (the field's own parameters)"
timeout 900 python tools/ext_fullsize.py 29 2>&1 | tail -3 | tee gpurun_out/r2_ext_2p29.json
nvidia-smi --query-gpu=memory.used,memory.total --format=csv,noheader
