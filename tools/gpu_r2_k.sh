#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
for r in 0 1 2; do
B200_AFFINE_ROUNDS=$r python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2957$r tools/prove_multi.py --log-m 20 --steps 8 2>&1 | grep '^{\|rror\|Traceback' | tee -a gpurun_out/r2_prove_multi_rounds_n$N.jsonl
done
