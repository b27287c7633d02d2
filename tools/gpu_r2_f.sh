#!/bin/bash
# Round 2, call F: register cap (resident CTAs/SM) of the level-0 affine kernels, per field.
mkdir -p gpurun_out
for mb in 1 2 3 4 5; do
  echo "== AFF_MB $mb"
  export B200_AFF_MB=$mb
  python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --group 2 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 1 --log-n 20 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 2 --log-n 18 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --curve bw6_761 --log-n 18 2>/dev/null | tail -1
done 2>&1 | tee gpurun_out/r2_aff_mb.log
