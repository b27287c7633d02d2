#!/bin/bash
# Round 2, call R: Fp::mul2 (two products side by side) in the NTT butterflies and the affine round: parity + timing.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ntt_gpu.py tests/test_msm_gpu.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | tail -3
( python tools/ntt_probe.py 24; python tools/ntt_probe.py 20
  python tools/affine_probe.py --check-log 12 | tail -1
  python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 1 --log-n 20 | tail -1
  python tools/prove_quick.py | tail -1 ) 2>/dev/null | tee gpurun_out/r2_mul2.log
