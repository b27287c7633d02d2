#!/usr/bin/env python
"""Probe of the batched-affine rounds (msm_affine.cuh): one resident MSM per (curve, group, size) with the number of
rounds taken from B200_AFFINE_ROUNDS (read once per process), stage times and a parity check against the reference.

    B200_AFFINE_ROUNDS=3 python tools/affine_probe.py [--curve alt_bn128] [--group 1] [--log-n 20] [--no-tables]"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libsnark_b200 as b  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--curve", default="alt_bn128")
    ap.add_argument("--group", type=int, default=1)
    ap.add_argument("--log-n", type=int, default=20)
    ap.add_argument("--no-tables", action="store_true")
    ap.add_argument("--check-log", type=int, default=16, help="parity on a prefix of this size (0 = whole MSM)")
    args = ap.parse_args()
    b.init(0)
    n = 1 << args.log_n
    bases = b.make_bases(args.curve, args.group, n, k0=3)
    sc = b.random_scalars(args.curve, n, seed=5)
    rb = b.ResidentBases(args.curve, args.group, bases)
    rows = 0 if args.no_tables else rb.precompute()
    d = b.DeviceBuffer(sc.nbytes)
    d.upload(sc)
    for _ in range(3):
        out = rb.multi_exp(None, device_ptr=d.ptr)
    st = []
    for _ in range(5):
        rb.multi_exp(None, device_ptr=d.ptr)
        st.append(b.last_msm_timing())
    best = min(st, key=lambda s: s["total"])
    line = {"curve": args.curve, "group": args.group, "log_n": args.log_n, "rounds_env": os.environ.get("B200_AFFINE_ROUNDS"),
            "table_rows": rows, "stage_ms": {k: round(v, 4) for k, v in best.items()}}
    try:
        from oracle import refo
        ref = refo.RefOracle()
        ref.set_num_threads(len(os.sched_getaffinity(0)))
        k = n if not args.check_log else min(n, 1 << args.check_log)
        z = sc.copy()
        z[k:] = 0
        got = rb.multi_exp(z)
        want = ref.msm(args.curve, args.group, bases[:k], sc[:k], chunks=ref.num_threads())
        line["bit_exact_prefix"] = bool(np.array_equal(got, want))
        line["prefix"] = k
    except (OSError, FileNotFoundError):
        pass
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
