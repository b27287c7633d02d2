#!/bin/bash
# Round 2, call C: new two-level reduce + L2 fetch granularity knob: parity tests, then timing probes.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_kat_gpu.py tests/test_msm_gpu.py tests/test_groth16_gpu.py -q -m gpu -x 2>&1 | tail -15
for fg in 0 32 64 128; do
  echo "== L2 fetch $fg"
  if [ $fg = 0 ]; then unset B200_L2_FETCH; else export B200_L2_FETCH=$fg; fi
  python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
  B200_AFFINE_ROUNDS=0 python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
done 2>&1 | tee gpurun_out/r2_l2fetch.log
unset B200_L2_FETCH
for cpt in 0 1 2 3; do
  echo "== reduce cpt $cpt"
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 --no-tables 2>/dev/null | tail -1
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 --group 2 2>/dev/null | tail -1
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 2 --log-n 18 2>/dev/null | tail -1
  B200_REDUCE_CPT=$cpt python tools/affine_probe.py --check-log 12 --curve bw6_761 --log-n 18 2>/dev/null | tail -1
done 2>&1 | tee gpurun_out/r2_reduce_cpt.log
