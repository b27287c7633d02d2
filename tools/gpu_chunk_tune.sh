#!/bin/bash
mkdir -p gpurun_out
for c in 2 4 8; do
  echo "== B200_TABLE_CHUNK=$c"
  B200_TABLE_CHUNK=$c timeout 600 python tools/msm_probe.py alt_bn128 bw6_761 2>&1 | grep -E "^(alt|bw6|  tables iter 3|  tables ==|  parity)"
done | tee gpurun_out/s10_chunk_tune.log
