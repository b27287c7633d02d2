#!/bin/bash
# Sharded prove only, at N GPUs (BN254 2^20 with the parity check, then bls12_377 2^22).
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_groth16_gpu.py -q -m gpu -x -k "chain_parallel or sharded" 2>&1 | grep -v "^\*\|^  \*\|^(\|^$" | tail -3
timeout 900 $TR --master-port 29561 tools/prove_multi.py --log-m 20 --check > gpurun_out/multi_n${N}_prove20.json 2> gpurun_out/multi_n${N}_prove20.err
tail -c 500 gpurun_out/multi_n${N}_prove20.json; tail -2 gpurun_out/multi_n${N}_prove20.err
