#!/bin/bash
# Round 2, call B: where the standalone MSM's time goes with the affine rounds: stage times per round count, the
# per-kernel launch list, and ncu --set full of the affine / accumulate / reduce kernels (alt_bn128 G1 2^20, tables).
mkdir -p gpurun_out
for r in 0 1 2 3; do B200_AFFINE_ROUNDS=$r python tools/affine_probe.py --check-log 14 2>/dev/null | tail -1; done | tee gpurun_out/r2_affine_rounds_g1.jsonl
for r in 0 2 3; do B200_AFFINE_ROUNDS=$r python tools/affine_probe.py --group 2 --check-log 12 2>/dev/null | tail -1; done | tee gpurun_out/r2_affine_rounds_g2.jsonl
CMD="python tools/affine_probe.py --check-log 10"
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_msm_affine.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on --kernel-name regex:'k_aff_up0|k_aff_down0|k_accumulate_pts|k_reduce_chunks|k_scatter|k_histogram|k_aff_top|k_aff_up|k_aff_down' \
    --launch-skip 120 --launch-count 40 -o gpurun_out/r2_prof_affine -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/
