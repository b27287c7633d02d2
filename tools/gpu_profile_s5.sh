#!/bin/bash
# Round-1 (session 5) evidence on one GPU: launch list of the bench's MSM step over the precomputed tables, one
# ncu --set full capture of its accumulate / reduce kernels, config 4 (bls12_377 2^22 prove) and the MSM sweep at N = 1.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ntt --no-prove"
$CMD > gpurun_out/s9_bench_short.json 2> gpurun_out/s9_bench_short.err; rc=$?; echo "short rc=$rc"
if [ $rc -eq 0 ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv \
      --log-file gpurun_out/launches_r1s5_bench.csv $CMD > gpurun_out/s9_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
  ncu --set full --clock-control none --import-source on --kernel-name regex:'k_accumulate|k_reduce_chunks' \
      --launch-skip 6 --launch-count 4 -o gpurun_out/prof_r1s5_tables -f $CMD > gpurun_out/s9_ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
timeout 900 python tools/prove_multi.py --curve bls12_377 --log-m 22 --steps 3 --warmup 2 > gpurun_out/multi_n1_prove22_bls12_377.json 2> gpurun_out/multi_n1_prove22.err
tail -c 700 gpurun_out/multi_n1_prove22_bls12_377.json
timeout 900 python tools/prove_multi.py --log-m 20 --check > gpurun_out/multi_n1_prove20.json 2> gpurun_out/multi_n1_prove20.err
tail -c 500 gpurun_out/multi_n1_prove20.json
timeout 1200 python tools/msm_sweep.py --curves bls12_377 bw6_761 --min-log 16 --max-log 24 --ref-max-log 0 --tables > gpurun_out/multi_n1_sweep.jsonl 2> gpurun_out/multi_n1_sweep.err
cut -c1-200 gpurun_out/multi_n1_sweep.jsonl; tail -2 gpurun_out/multi_n1_sweep.err
ls -la gpurun_out/*.ncu-rep
