#!/bin/bash
# Round 2, call Z13: L2 prefetch of the next block's points in round 1 of the affine rounds.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_msm_gpu.py tests/test_kat_gpu.py -q -m gpu -x 2>&1 | tail -2
{
python tools/affine_probe.py --check-log 12 2>/dev/null | tail -1
python tools/affine_probe.py --check-log 12 --group 2 2>/dev/null | tail -1
python tools/affine_probe.py --check-log 12 --curve bls12_377 2>/dev/null | tail -1
python tools/affine_probe.py --check-log 12 --curve bls12_377 --group 2 --log-n 18 2>/dev/null | tail -1
python tools/affine_probe.py --check-log 12 --curve bw6_761 --log-n 18 2>/dev/null | tail -1
python tools/prove_quick.py --steps 10 2>/dev/null | tail -1
} | tee gpurun_out/r2_aff_prefetch.jsonl
