#!/bin/bash
# Round 2, call Y: the whole gpu suite on the current tree, then smoke().
mkdir -p gpurun_out
(time timeout 2400 python -m pytest tests -q -m gpu -x --durations=12 2>&1 | grep -v "^\* \|^$" | tail -30) 2>&1 | tee gpurun_out/r2_gpu_suite.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
