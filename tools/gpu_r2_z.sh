#!/bin/bash
# Round 2, call Z17: R1CS evaluation after the body / plan refactor: witness-map and prover tests.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ntt_gpu.py tests/test_groth16_gpu.py -q -m gpu -x -k "r1cs or witness or golden or c_abi_prover or zk_patch" 2>&1 | tail -2
