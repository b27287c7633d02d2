#!/bin/bash
# Multi-GPU check: NCCL test of the sharded prover, the sharded prove at 2^20 (BN254) and the bench line at N GPUs.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_groth16_gpu.py -q -m gpu -x -k "distributed_prover_nccl" 2>&1 | tail -5 | tee gpurun_out/multi_n${N}_pytest.log
timeout 900 $TR --master-port 29561 tools/prove_multi.py --log-m 20 --check > gpurun_out/multi_n${N}_prove20.json 2> gpurun_out/multi_n${N}_prove20.err
tail -c 1500 gpurun_out/multi_n${N}_prove20.json; tail -3 gpurun_out/multi_n${N}_prove20.err
timeout 900 $TR --master-port 29562 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/multi_n${N}_bench.json 2> gpurun_out/multi_n${N}_bench.err
tail -c 2500 gpurun_out/multi_n${N}_bench.json; tail -3 gpurun_out/multi_n${N}_bench.err
