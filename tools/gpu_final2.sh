#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
(timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | grep -v "^\*\|^  \*\|^(\|^$" | tail -4) 2>&1 | tee gpurun_out/s14_pytest.log
timeout 900 python bench.py > gpurun_out/s14_bench.json 2> gpurun_out/s14_bench.err
tail -c 200 gpurun_out/s14_bench.json; tail -3 gpurun_out/s14_bench.err
