#!/bin/bash
# Round 2, call I: shared witness sort + parallel host fold + background shard MSMs: parity, then prover timing.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_groth16_gpu.py tests/test_msm_gpu.py -q -m gpu -x 2>&1 | tail -5
( python tools/prove_quick.py
  B200_SHARED_SORT=0 python tools/prove_quick.py
  B200_MSM_SLOTS=2,1,0,0,3 python tools/prove_quick.py
  B200_MSM_SLOTS=2,1,0,0,3 B200_SHARED_SORT=0 python tools/prove_quick.py
  B200_MSM_SLOTS=0,1,2,0,3 python tools/prove_quick.py
  B200_MSM_SLOTS=3,1,2,0,3 python tools/prove_quick.py
  B200_WM_FIRST=0 python tools/prove_quick.py ) 2>/dev/null | tee gpurun_out/r2_prove_knobs3.jsonl
