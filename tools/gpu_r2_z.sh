#!/bin/bash
# Round 2, call Z10: staged host-to-device copies of pageable bases / scalars (b200_msm, b200_bases_upload).
mkdir -p gpurun_out
python - <<'PY' | tee gpurun_out/r2_cold_msm.json
import json, time, numpy as np, libsnark_b200 as b
b.init(0)
out = {}
for curve, group, logn in [("alt_bn128", b.G1, 20), ("alt_bn128", b.G2, 20), ("bls12_377", b.G1, 20), ("alt_bn128", b.G1, 16)]:
    n = 1 << logn
    bases = b.make_bases(curve, group, n, k0=1)
    rng = np.random.default_rng(3)
    limbs = b.sizeof_fr(curve) // 8
    sc = rng.integers(0, 1 << 62, size=(n, limbs), dtype=np.uint64)
    r0 = b.multi_exp(curve, group, bases, sc)
    t = []
    for _ in range(8):
        b.lib().b200_sync()
        t0 = time.perf_counter()
        r = b.multi_exp(curve, group, bases, sc)
        t.append((time.perf_counter() - t0) * 1e3)
    assert np.array_equal(r, r0)
    rb = b.ResidentBases(curve, group, bases)
    want = rb.multi_exp(sc)
    rb.free()
    assert np.array_equal(want, r0)
    out[f"{curve}.g{group}.2^{logn}"] = {"cold_ms": round(float(np.mean(t)), 2), "min_ms": round(float(np.min(t)), 2),
                                          "h2d_MiB": round((bases.nbytes + sc.nbytes) / 2**20, 1), "all_ms": [round(x, 1) for x in t]}
print(json.dumps(out))
PY
