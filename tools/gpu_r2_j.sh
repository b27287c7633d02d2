#!/bin/bash
# Round 2, call J (N GPUs): sharded prover with background witness-only MSMs and parallel folds: parity and stage split.
mkdir -p gpurun_out
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 tools/prove_multi.py --log-m 20 --check 2>&1 | grep '^{' | tee gpurun_out/r2_prove_multi_n$N.jsonl
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29562 tools/prove_multi.py --log-m 16 --check 2>&1 | grep '^{' | tee -a gpurun_out/r2_prove_multi_n$N.jsonl
timeout 600 python -m pytest tests/test_groth16_gpu.py -q -m gpu -x -k "nccl or sharded" 2>&1 | tail -3
if [ "$N" = "4" ]; then
( python tools/prove_quick.py
  B200_MSM_SLOTS=2,1,4,0,3 python tools/prove_quick.py
  B200_MSM_SLOTS=2,1,4,5,3 python tools/prove_quick.py ) 2>/dev/null | tee gpurun_out/r2_prove_knobs4.jsonl
fi
