#!/bin/bash
# Round 2, final tree at N GPUs: the driver's own launch line for bench.py (both arms) and the NCCL prover test.
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_groth16_gpu.py -q -m gpu -x -k "nccl" 2>&1 | grep -v "^\*\|^  \*\|^(\|^$" | tail -2
(time timeout 900 $TR --master-port 29571 bench.py --gpus $N > gpurun_out/r2s3_bench_n${N}.json 2> gpurun_out/r2s3_bench_n${N}.err) 2>&1 | grep real
tail -c 300 gpurun_out/r2s3_bench_n${N}.err; head -c 600 gpurun_out/r2s3_bench_n${N}.json
