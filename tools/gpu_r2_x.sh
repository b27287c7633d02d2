#!/bin/bash
# Round 2, call X: the sibling provers, reference build vs overlay build.
mkdir -p gpurun_out
timeout 1500 python tools/sibling_prove.py --log-nc 16 2>/dev/null | grep "^{" | tee gpurun_out/r2_sibling_prove.jsonl
