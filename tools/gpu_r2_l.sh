#!/bin/bash
# Round 2, call L: affine rounds vs MSM size (standalone, tables): where do the rounds start to pay?
mkdir -p gpurun_out
for ln in 16 17 18 19; do for r in 0 1 2 3; do
  B200_AFFINE_ROUNDS=$r python tools/affine_probe.py --log-n $ln --check-log 10 2>/dev/null | tail -1
done; done | tee gpurun_out/r2_rounds_vs_size.jsonl
for ln in 17 18; do for r in 0 2 3; do
  B200_AFFINE_ROUNDS=$r python tools/affine_probe.py --group 2 --log-n $ln --check-log 10 2>/dev/null | tail -1
done; done | tee -a gpurun_out/r2_rounds_vs_size.jsonl
