#!/bin/bash
# Round 2, final validation on the final tree: smoke(), the whole gpu suite, both bench arms (N = 1).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
(time timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | grep -v "^\*\|^  \*\|^(\|^$" | tail -4) 2>&1 | tee gpurun_out/r2s3_pytest.log
(time timeout 900 python bench.py --impl reference > gpurun_out/r2s3_bench_reference_n1.json 2> gpurun_out/r2s3_bench_reference_n1.err) 2>&1 | grep real
(time timeout 900 python bench.py > gpurun_out/r2s3_bench_n1.json 2> gpurun_out/r2s3_bench_n1.err) 2>&1 | grep real
tail -c 300 gpurun_out/r2s3_bench_n1.err; head -c 900 gpurun_out/r2s3_bench_n1.json; echo; head -c 400 gpurun_out/r2s3_bench_reference_n1.json
