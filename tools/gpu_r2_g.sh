#!/bin/bash
# Round 2, call G: prover under tuning knobs (affine rounds, slot plans) with the per-field register caps built in.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_msm_gpu.py -q -m gpu -x 2>&1 | tail -3
( python tools/prove_quick.py
  B200_AFFINE_ROUNDS=4 python tools/prove_quick.py
  B200_AFFINE_ROUNDS=2 python tools/prove_quick.py
  B200_WM_FIRST=0 python tools/prove_quick.py
  B200_MSM_SLOTS=2,1,3,0,3 python tools/prove_quick.py
  B200_MSM_SLOTS=2,1,2,3,0 python tools/prove_quick.py
  B200_MSM_SLOTS=0,1,2,0,3 python tools/prove_quick.py
  B200_MSM_SLOTS=2,1,0,0,3 python tools/prove_quick.py ) 2>/dev/null | tee gpurun_out/r2_prove_knobs.jsonl
