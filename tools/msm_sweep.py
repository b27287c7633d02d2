"""BASELINE.json configs[4]: MSM size sweep on bls12_377 and bw6_761 G1 (plus alt_bn128), resident bases,
with the reference's multi_exp timed on the host cores up to --ref-max-log (parity checked where both run).

    python tools/msm_sweep.py [--curves bls12_377 bw6_761] [--min-log 16] [--max-log 24] [--ref-max-log 20]
One JSON object per (curve, size) on stdout."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libsnark_b200 as b  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--curves", nargs="+", default=["bls12_377", "bw6_761"])
    ap.add_argument("--min-log", type=int, default=16)
    ap.add_argument("--max-log", type=int, default=24)
    ap.add_argument("--ref-max-log", type=int, default=20)
    args = ap.parse_args()
    b.init(0)
    ref = None
    try:
        from oracle import refo
        ref = refo.RefOracle()
        ref.set_num_threads(len(os.sched_getaffinity(0)))
    except (OSError, FileNotFoundError):
        pass
    for curve in args.curves:
        for log_n in range(args.min_log, args.max_log + 1, 2):
            n = 1 << log_n
            bases = b.make_bases(curve, b.G1, n, k0=7)
            scalars = b.random_scalars(curve, n, seed=log_n)
            rb = b.ResidentBases(curve, b.G1, bases)
            out = rb.multi_exp(scalars)
            best, stage = 1e30, None
            for _ in range(3):
                t0 = time.perf_counter()
                out = rb.multi_exp(scalars)
                dt = (time.perf_counter() - t0) * 1e3
                if dt < best:
                    best, stage = dt, b.last_msm_timing()
            line = {"curve": curve, "group": "G1", "log_n": log_n, "e2e_ms": best, "device_ms": stage["total"],
                    "stage_ms": stage, "mpts_per_s_e2e": n / best / 1e3, "mpts_per_s_resident": n / stage["total"] / 1e3,
                    "window_bits": int(b.lib().b200_msm_window_bits(b._cid(curve), n))}
            if ref is not None and log_n <= args.ref_max_log:
                t0 = time.perf_counter()
                want = ref.msm(curve, 1, bases, scalars, chunks=ref.num_threads())
                line["reference_ms"] = (time.perf_counter() - t0) * 1e3
                line["reference_threads"] = ref.num_threads()
                line["bit_exact"] = bool(np.array_equal(out, want))
                line["speedup_e2e"] = line["reference_ms"] / best
            print(json.dumps(line), flush=True)
            rb.free()
            del bases, scalars


if __name__ == "__main__":
    main()
