#!/bin/bash
# Round 2, closing run on the final tree: smoke() and the whole gpu suite.
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
(time timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | grep -v "^\*\|^  \*\|^(\|^$" | tail -3) 2>&1 | tee gpurun_out/r2s6_pytest.log
