#!/bin/bash
# Round 2, final bench on the final tree (N = 1): the whole gpu suite once more, then bench.py (the reference arm of this
# tree's inputs was run by tools/gpu_r2_final.sh: profiles/r2s3_bench_reference_n1.json).
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | grep -v "^\*\|^  \*\|^(\|^$" | tail -3) 2>&1 | tee gpurun_out/r2s3_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
(time timeout 900 python bench.py > gpurun_out/r2s3_bench_n1.json 2> gpurun_out/r2s3_bench_n1.err) 2>&1 | grep real
tail -c 300 gpurun_out/r2s3_bench_n1.err; head -c 300 gpurun_out/r2s3_bench_n1.json
