"""Timing probe: alt_bn128 G1 MSM at 2^20 with stage split, plus the IMAD peak microbenchmark."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libsnark_b200 as b  # noqa: E402
from oracle import refo  # noqa: E402

b.init(0)
print("imad peak (wide, lo) mads/s:", b.imad_peak())
ref = refo.RefOracle()
CASES = [("alt_bn128", refo.G1, 20), ("alt_bn128", refo.G2, 18), ("alt_bn128", refo.G2, 20), ("bls12_377", refo.G1, 18),
         ("bls12_377", refo.G1, 20), ("bls12_377", refo.G2, 18), ("bw6_761", refo.G1, 16), ("bw6_761", refo.G1, 18)]
if len(sys.argv) > 1:
    CASES = [c for c in CASES if c[0] in sys.argv[1:] or f"{c[0]}:{c[1]}" in sys.argv[1:]]
for curve, group, logn in CASES:
    n = 1 << logn
    t0 = time.time()
    bases = ref.make_bases(curve, group, n, 1)
    scalars = ref.field_random(curve, refo.FR, 0, n)
    print(curve, group, logn, "inputs", round(time.time() - t0, 2), "s")
    rb = b.ResidentBases(curve, group, bases)
    for it in range(4):
        t0 = time.time()
        out = rb.multi_exp(scalars)
        dt = time.time() - t0
        print("  iter", it, "wall ms", round(dt * 1e3, 3), {k: round(v, 3) for k, v in b.last_msm_timing().items()})
    t0 = time.time()
    rows = rb.precompute()
    print("  precompute:", rows, "rows,", round(time.time() - t0, 3), "s")
    for it in range(4):
        t0 = time.time()
        out_t = rb.multi_exp(scalars)
        dt = time.time() - t0
        print("  tables iter", it, "wall ms", round(dt * 1e3, 3), {k: round(v, 3) for k, v in b.last_msm_timing().items()})
    print("  tables == plain:", np.array_equal(out, out_t))
    if logn <= 18:
        want = ref.msm(curve, group, bases, scalars, chunks=ref.num_threads())
        print("  parity:", np.array_equal(out, want))
    rb.free()
