#!/bin/bash
mkdir -p gpurun_out
(time timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -5) 2>&1 | tee gpurun_out/s11_pytest.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s11_bench_reference.json 2> gpurun_out/s11_bench_reference.err
timeout 900 python bench.py > gpurun_out/s11_bench.json 2> gpurun_out/s11_bench.err
tail -c 300 gpurun_out/s11_bench.json; tail -3 gpurun_out/s11_bench.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
