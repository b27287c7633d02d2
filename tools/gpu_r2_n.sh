#!/bin/bash
mkdir -p gpurun_out
python tools/affine_probe.py --group 2 --check-log 12 2>/dev/null | tail -1
python tools/prove_quick.py 2>/dev/null | tail -1
