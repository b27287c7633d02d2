#!/bin/bash
# Round 2, call S: NTT passes planned for whole waves: parity + timing; prover knobs on the final tree.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ntt_gpu.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | tail -3
( for l in 18 19 20 21 22 24; do python tools/ntt_probe.py $l | head -2; B200_NTT_WAVES=0 python tools/ntt_probe.py $l | head -2 | sed 's/^/   waves off: /'; done
  python tools/prove_quick.py | tail -1
  B200_NTT_WAVES=0 python tools/prove_quick.py | tail -1
  B200_WM_FIRST=0 python tools/prove_quick.py | tail -1 ) 2>/dev/null | tee gpurun_out/r2_ntt_waves.log
