#!/bin/bash
# Round 2, call X: SAP witness map through the overlay; the sibling provers, reference build vs overlay build.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_groth16_gpu.py -q -m gpu -x -k "sap_witness or sibling" 2>&1 | grep -v "^\* \|^$" | tail -8
timeout 1500 python tools/sibling_prove.py --log-nc 16 2>/dev/null | grep "^{" | tee gpurun_out/r2_sibling_prove.jsonl
