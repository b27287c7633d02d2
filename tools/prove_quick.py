#!/usr/bin/env python
"""Quick prover timing under the tuning knobs of the environment (B200_AFFINE_ROUNDS, B200_MSM_SLOTS, ...): synthetic
key (device-made points, no reference needed), resident witness, a few proofs; one JSON line.

    B200_AFFINE_ROUNDS=4 python tools/prove_quick.py [--curve alt_bn128] [--log-m 20] [--steps 8]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench  # noqa: E402
import libsnark_b200 as b  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--curve", default="alt_bn128")
    ap.add_argument("--log-m", type=int, default=20)
    ap.add_argument("--steps", type=int, default=8)
    args = ap.parse_args()
    b.init(0)
    d = bench.prove_inputs_synthetic(b, args.curve, args.log_m)
    cs = b.R1CS(args.curve, d["nc"], d["ni"], d["nv"], d["row_ptr"], d["col"], d["val"])
    pts = bench.split_points(b, args.curve, d["points"])
    pk = b.Groth16ProvingKey(args.curve, *pts, d["A"], d["Bg2"], d["Bg1"], d["Bidx"], d["H"], d["L"], cs, d["m"], precompute=True)
    rs = d["rs"]
    pk.upload_witness(d["primary"], d["auxiliary"])
    proof = None
    for _ in range(3):
        proof, _st = pk.prove_resident(rs[0], rs[1])
    ms, stages = [], []
    for _ in range(args.steps):
        b.lib().b200_sync()
        t0 = time.perf_counter()
        p, st = pk.prove_resident(rs[0], rs[1])
        ms.append((time.perf_counter() - t0) * 1e3)
        stages.append(st)
        assert np.array_equal(p, proof)
    knobs = {k: v for k, v in os.environ.items() if k.startswith("B200_")}
    print(json.dumps({"curve": args.curve, "log_m": args.log_m, "knobs": knobs, "ms": round(float(np.mean(ms)), 3),
                      "min_ms": round(float(np.min(ms)), 3),
                      "stage_ms": {k: round(float(np.mean([s[k] for s in stages])), 2) for k in stages[0]},
                      "proof_crc": int(np.bitwise_xor.reduce(proof.view(np.uint64)))}), flush=True)
    pk.free()
    cs.free()


if __name__ == "__main__":
    main()
