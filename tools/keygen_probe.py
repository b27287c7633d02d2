"""Key-generation operator timing: b200_batch_exp vs the reference's batch_exp + batch_to_special on the
host cores, same base and scalars, bit-exact check.   python tools/keygen_probe.py [log_n ...]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libsnark_b200 as b  # noqa: E402
from oracle import refo  # noqa: E402

b.init(0)
ref = refo.RefOracle()
ref.set_num_threads(len(os.sched_getaffinity(0)))
logs = [int(x) for x in sys.argv[1:]] or [16, 20]
for curve, group in [("alt_bn128", refo.G1), ("alt_bn128", refo.G2), ("bls12_377", refo.G1), ("bw6_761", refo.G1)]:
    g = ref.make_bases(curve, group, 1, 31)[0]
    for log_n in logs:
        if curve != "alt_bn128" and log_n > 18:
            continue
        n = 1 << log_n
        sc = b.random_scalars(curve, n, seed=log_n)
        b.batch_exp(curve, group, g, sc[:16])                      # table build + warm-up
        best = 1e30
        for _ in range(3):
            t0 = time.perf_counter()
            out = b.batch_exp(curve, group, g, sc)
            best = min(best, (time.perf_counter() - t0) * 1e3)
        t0 = time.perf_counter()
        want = ref.batch_exp(curve, group, g, sc)
        ref_ms = (time.perf_counter() - t0) * 1e3
        print(json.dumps({"op": "batch_exp + batch_to_special", "curve": curve, "group": "G%d" % group, "log_n": log_n,
                          "b200_ms_host_to_host": round(best, 3), "reference_ms": round(ref_ms, 1),
                          "reference_threads": ref.num_threads(), "speedup": round(ref_ms / best, 1),
                          "bit_exact": bool(np.array_equal(out, want))}), flush=True)
