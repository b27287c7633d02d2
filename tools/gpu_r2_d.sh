#!/bin/bash
# Round 2, call D: two-level reduce with side-by-side products: parity, then stage times per buckets-per-thread setting.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_kat_gpu.py tests/test_msm_gpu.py -q -m gpu -x 2>&1 | tail -5
for cpt in 0 1 2 3; do
  echo "== reduce cpt $cpt"
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 --no-tables 2>/dev/null | tail -1
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 --group 2 2>/dev/null | tail -1
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 2 --log-n 18 2>/dev/null | tail -1
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 --curve bw6_761 --log-n 18 2>/dev/null | tail -1
done 2>&1 | tee gpurun_out/r2_reduce_cpt2.log
