#!/bin/bash
# Multi-GPU measurements at N GPUs: bench line, sharded prove (BN254 2^20 with parity check vs one GPU; bls12_377 2^22 =
# BASELINE.json configs[3]), MSM sweep (configs[4]).
N=${1:-8}
MAXLOG=${2:-24}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" -gt 1 ]; then
timeout 600 python -m pytest tests/test_groth16_gpu.py -q -m gpu -x -k "distributed_prover_nccl" 2>&1 | tail -3 | tee gpurun_out/multi_n${N}_pytest.log
fi
timeout 900 $TR --master-port 29561 tools/prove_multi.py --log-m 20 --check > gpurun_out/multi_n${N}_prove20.json 2> gpurun_out/multi_n${N}_prove20.err
tail -c 600 gpurun_out/multi_n${N}_prove20.json; tail -2 gpurun_out/multi_n${N}_prove20.err
timeout 900 $TR --master-port 29562 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/multi_n${N}_bench.json 2> gpurun_out/multi_n${N}_bench.err
head -c 600 gpurun_out/multi_n${N}_bench.json; tail -2 gpurun_out/multi_n${N}_bench.err
timeout 1200 $TR --master-port 29563 tools/prove_multi.py --curve bls12_377 --log-m 22 --steps 3 --warmup 2 > gpurun_out/multi_n${N}_prove22_bls12_377.json 2> gpurun_out/multi_n${N}_prove22.err
tail -c 600 gpurun_out/multi_n${N}_prove22_bls12_377.json; tail -2 gpurun_out/multi_n${N}_prove22.err
[ "$MAXLOG" -gt 0 ] && timeout 1500 $TR --master-port 29564 tools/msm_sweep.py --curves bls12_377 bw6_761 --min-log 16 --max-log $MAXLOG --ref-max-log 0 --tables > gpurun_out/multi_n${N}_sweep.jsonl 2> gpurun_out/multi_n${N}_sweep.err
[ "$MAXLOG" -gt 0 ] && (cat gpurun_out/multi_n${N}_sweep.jsonl | cut -c1-330; tail -2 gpurun_out/multi_n${N}_sweep.err)
exit 0
