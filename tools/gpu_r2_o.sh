#!/bin/bash
# Round 2, call O: affine rounds with block ownership (element-major arrays), fused or not: parity and timing per field.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_msm_gpu.py tests/test_kat_gpu.py -q -m gpu -x 2>&1 | tail -3
for fuse in 0 1; do
( B200_AFF_FUSE=$fuse python tools/affine_probe.py --check-log 12
  B200_AFF_FUSE=$fuse python tools/affine_probe.py --check-log 12 --group 2
  B200_AFF_FUSE=$fuse python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 1 --log-n 20
  B200_AFF_FUSE=$fuse python tools/affine_probe.py --check-log 12 --curve bw6_761 --log-n 18
  B200_AFF_FUSE=$fuse python tools/prove_quick.py ) 2>/dev/null | grep '^{' | tee -a gpurun_out/r2_fused_rounds2.jsonl
done
python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 2 --log-n 18 2>/dev/null | grep '^{' | tee -a gpurun_out/r2_fused_rounds2.jsonl
