#!/bin/bash
# Round 2, call Z15: NTT launcher after the plan_transform refactor: the NTT and witness-map tests, one prove.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ntt_gpu.py tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | tail -2
timeout 300 python tools/prove_quick.py --steps 6 2>/dev/null | tail -1 | cut -c1-300
