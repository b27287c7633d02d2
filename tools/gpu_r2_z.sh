#!/bin/bash
# Round 2, last call: the ZK-patch golden test against the regenerated tests/golden/zk.npz.
timeout 200 python -m pytest tests/test_ntt_gpu.py -q -m gpu -x -k "zk_patch" 2>&1 | tail -2
