#!/bin/bash
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_keygen_gpu.py -q -m gpu -x 2>&1 | tail -15) | tee gpurun_out/s13_keygen_pytest.log
timeout 600 python tools/keygen_probe.py 16 20 2>&1 | tee gpurun_out/s13_keygen_probe.jsonl
