#!/bin/bash
# Round 2, call Z16: extended-domain kernels after the body refactor: NTT and prover tests on step / extended domains.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ntt_gpu.py tests/test_groth16_gpu.py -q -m gpu -x -k "extended or step or zk_patch or chain_parallel or fft_ops" 2>&1 | tail -2
