#!/bin/bash
# Round 2, call V: the sibling provers (PGHR13, SE) through the overlay build, and the <G1,G1> kc MSM.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_groth16_gpu.py -q -m gpu -k "sibling or g1_pairs or overlay_operators or extended" --durations=8 > gpurun_out/r2_sibling_tests.log 2>&1
grep -v "^\* \|^$" gpurun_out/r2_sibling_tests.log | tail -80
