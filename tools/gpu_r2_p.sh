#!/bin/bash
# Round 2, call P: the ncu evidence of the headline command: launch list of a short prove-only bench run and --set full
# of the kernels that dominate it (raw CSV written on the box; the .ncu-rep itself is too large to bring back).
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ntt --no-msm --no-config4 --no-sweep"
$CMD > gpurun_out/r2_bench_short.json 2> gpurun_out/r2_bench_short.err || { echo "bench failed"; tail -5 gpurun_out/r2_bench_short.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_prove.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --kernel-name regex:'k_aff_round|k_aff_up_round|k_accumulate_pts|k_ntt_pass|k_reduce_chunks' \
    --launch-skip 420 --launch-count 36 -o /tmp/r2_prof_prove -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/r2_prof_prove.ncu-rep --page raw --csv > gpurun_out/r2_prof_prove_raw.csv 2>/dev/null
python tools/ncu_summary.py /tmp/r2_prof_prove.ncu-rep > gpurun_out/r2_ncu_prove.md 2>/dev/null
ls -la gpurun_out | tail -8
