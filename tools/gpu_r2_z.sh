#!/bin/bash
# Round 2, call Z2: L2 reuse between the two gathers of an affine round (tools/microbench/gather_l2.cu).
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/gather_l2 tools/microbench/gather_l2.cu && /tmp/gather_l2 | tee gpurun_out/r2_gather_l2.json
