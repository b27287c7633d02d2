#!/bin/bash
# Round 2, call E: chunked vs tree reduce on every field, and ncu --set full of the G2 affine kernels and both reduces.
mkdir -p gpurun_out
for mode in chunks tree; do
  echo "== reduce $mode"
  export B200_REDUCE=$mode
  python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --no-tables 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --group 2 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 1 --log-n 20 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 2 --log-n 18 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --curve bw6_761 --log-n 18 2>/dev/null | tail -1
  python tools/affine_probe.py --check-log 12 --curve bw6_761 --log-n 16 2>/dev/null | tail -1
done 2>&1 | tee gpurun_out/r2_reduce_ab.log
export B200_REDUCE=tree
CMD="python tools/affine_probe.py --check-log 10 --group 2"
ncu --set full --clock-control none --import-source on --kernel-name regex:'k_aff_up0|k_aff_down0|k_accumulate_pts|k_reduce_rows|k_reduce_windows' \
    --launch-skip 40 --launch-count 9 -o gpurun_out/r2_prof_g2 -f $CMD > gpurun_out/ncu_full_g2.log 2>&1; echo "ncu g2 rc=$?"
CMD="python tools/affine_probe.py --check-log 10"
ncu --set full --clock-control none --import-source on --kernel-name regex:'k_reduce_rows|k_reduce_windows' \
    --launch-skip 10 --launch-count 2 -o gpurun_out/r2_prof_reduce_tree -f $CMD > gpurun_out/ncu_full_red.log 2>&1; echo "ncu reduce rc=$?"
ls -la gpurun_out
