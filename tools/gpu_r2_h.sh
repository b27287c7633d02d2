#!/bin/bash
# Round 2, call H: short kernels on a high-priority stream per slot: parity, standalone MSM, prover under slot plans.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_msm_gpu.py tests/test_groth16_gpu.py -q -m gpu -x 2>&1 | tail -3
python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
( python tools/prove_quick.py
  B200_HI_STREAM=0 python tools/prove_quick.py
  B200_FLAT_PRIO=1 python tools/prove_quick.py
  B200_MSM_SLOTS=2,1,0,0,3 python tools/prove_quick.py
  B200_MSM_SLOTS=2,1,0,0,3 B200_FLAT_PRIO=1 python tools/prove_quick.py
  B200_WM_FIRST=0 python tools/prove_quick.py
  B200_WM_FIRST=0 B200_MSM_SLOTS=2,1,0,0,3 python tools/prove_quick.py
  B200_WM_FIRST=0 B200_FLAT_PRIO=1 python tools/prove_quick.py
  B200_MSM_SLOTS=0,1,2,0,3 python tools/prove_quick.py
  B200_MSM_SLOTS=0,1,2,3,3 B200_WM_FIRST=0 python tools/prove_quick.py ) 2>/dev/null | tee gpurun_out/r2_prove_knobs2.jsonl
