"""extended_radix2_domain at the one size the fields' own parameters reach on a GPU: alt_bn128 Fr, m = 2^29 (s = 28),
a 16 GiB vector. FFT in place on the device, two outputs (one per half of the domain) checked against the unmodified
reference's evaluate_polynomial at get_evaluation_domain(2^29)->get_domain_element(i) (reference parameters untouched:
the factory hands out extended_radix2_domain by itself), then iFFT and a bit-exact compare with the input.
Needs ~20 GiB of host memory and ~40 GiB of HBM; prints one JSON line.  python tools/ext_fullsize.py [log2_m]"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    lg = int(sys.argv[1]) if len(sys.argv) > 1 else 29
    avail = 0
    for line in open("/proc/meminfo"):
        if line.startswith("MemAvailable"):
            avail = int(line.split()[1]) >> 20
    need = (32 << lg) * 2.5 / (1 << 30) + 8     # the vector, the harness's std::vector copy of it, slack
    if avail < need:
        print(json.dumps({"skipped": f"host memory: {avail} GiB available, {need:.0f} GiB wanted"}))
        return
    import libsnark_b200 as b200
    from oracle import refo
    b200.init(0)
    ref = refo.RefOracle()
    curve = "alt_bn128"
    if lg != 29:
        ref.set_two_adicity(curve, lg - 1)
        b200.set_two_adicity(curve, lg - 1)
    m, tile = 1 << lg, 1 << min(24, lg)
    rng = np.random.default_rng(29)
    base = rng.integers(0, 1 << 63, size=(tile, 4), dtype=np.uint64)
    base[:, 3] >>= np.uint64(3)                      # < 2^252 < r: canonical residues
    a = np.empty((m, 4), dtype=np.uint64)
    for t in range(m // tile):
        a[t * tile:(t + 1) * tile] = base
        a[t * tile:(t + 1) * tile, 1] ^= np.uint64(0x9E3779B97F4A7C15 * (t + 1) & (2**64 - 1))
    L = b200.lib()
    dev = b200.DeviceBuffer(a.nbytes)
    for t in range(m // tile):
        part = a[t * tile:(t + 1) * tile]
        assert L.b200_memcpy_h2d(C.c_void_p(dev.ptr + t * tile * 32), part.ctypes.data_as(C.c_void_p), part.nbytes) == 0
    cid = b200.CURVES[curve]
    out = {"curve": curve, "log2_m": lg, "domain": "extended_radix2_domain"}
    for rep in range(2):                              # first call builds the tables of the 2^(lg-1) sub-domain
        t0 = time.perf_counter()
        rc = L.b200_ntt_dev(cid, b200.FFT, C.c_void_p(dev.ptr), m, None)
        assert rc == 0, L.b200_last_error()
        out["fft_ms_call%d" % rep] = round((time.perf_counter() - t0) * 1e3, 2)
        if rep == 0:
            idx = np.array([12345, (m >> 1) + 54321], dtype=np.uint64)
            got = np.zeros((2, 4), dtype=np.uint64)
            for k, i in enumerate(idx):
                assert L.b200_memcpy_d2h(got[k].ctypes.data_as(C.c_void_p), C.c_void_p(dev.ptr + int(i) * 32), 32) == 0
            t0 = time.perf_counter()
            want = ref.poly_eval_at_factory_domain(curve, a, idx)
            out["reference_horner_s"] = round(time.perf_counter() - t0, 1)
            out["spot_check_indices"] = [int(i) for i in idx]
            out["spot_check_equal"] = bool(np.array_equal(got, want))
            t0 = time.perf_counter()
            assert L.b200_ntt_dev(cid, b200.IFFT, C.c_void_p(dev.ptr), m, None) == 0, L.b200_last_error()
            out["ifft_ms"] = round((time.perf_counter() - t0) * 1e3, 2)
            back = np.empty((tile, 4), dtype=np.uint64)
            same = True
            for t in range(m // tile):
                assert L.b200_memcpy_d2h(back.ctypes.data_as(C.c_void_p), C.c_void_p(dev.ptr + t * tile * 32), back.nbytes) == 0
                same = same and np.array_equal(back, a[t * tile:(t + 1) * tile])
            out["round_trip_equal"] = bool(same)
    dev.free()
    print(json.dumps(out))
    if not (out["spot_check_equal"] and out["round_trip_equal"]):
        sys.exit(1)


if __name__ == "__main__":
    main()
