#!/usr/bin/env python
"""The sibling provers (r1cs_ppzksnark = PGHR13, r1cs_se_ppzksnark) instantiated with BDLO12_signed / special: the
reference's own prover entry point called from the harness built as is and built with the overlay, same key and witness;
every proof checked by the reference's strong-IC verifier (plain build). One JSON line per scheme.

    python tools/sibling_prove.py [--curve alt_bn128] [--log-nc 16] [--steps 3]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--curve", default="alt_bn128")
    ap.add_argument("--log-nc", type=int, default=16)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    import libsnark_b200 as b200
    from oracle import refo
    b200.init(0)
    ref, ovl = refo.RefOracle(), refo.RefOracle(overlay=True)
    for scheme, name in [(0, "r1cs_ppzksnark"), (1, "r1cs_se_ppzksnark")]:
        t0 = time.perf_counter()
        fx = ref.sibling_setup(scheme, args.curve, (1 << args.log_nc) - 11, 10)
        out = {"prover": name, "curve": args.curve, "num_constraints": (1 << args.log_nc) - 11,
               "generator_s": round(time.perf_counter() - t0, 1)}
        proof, ms = fx.prove()
        assert fx.verify(proof) == 1
        out["reference_build_ms"] = round(ms, 1)
        calls = []
        for _ in range(args.steps + 2):
            proof, ms = fx.prove(via=ovl)
            assert fx.verify(proof) == 1
            calls.append(ms)
        out["overlay_build_first_call_ms"] = round(calls[0], 1)
        out["overlay_build_second_call_ms"] = round(calls[1], 1)
        out["overlay_build_ms"] = round(float(np.mean(calls[2:])), 2)
        out["verified"] = True
        print(json.dumps(out), flush=True)
        fx.close()


if __name__ == "__main__":
    main()
