"""Timing probe: b200_ntt_dev on alt_bn128 Fr at 2^log_m (device-resident data)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libsnark_b200 as b  # noqa: E402

log_m = int(sys.argv[1]) if len(sys.argv) > 1 else 24
b.init(0)
m = 1 << log_m
a = b.random_scalars("alt_bn128", m, seed=3)
buf = b.DeviceBuffer(a.nbytes)
buf.upload(a)
L = b.lib()
import ctypes as C
# Fr::multiplicative_generator = 5 for alt_bn128 (alt_bn128_init.cpp:72), Montgomery form via the library's own tables
g = np.array([0x1b0d0ef99fffffe6, 0xeaba68a3a32a913f, 0x47d8eb76d8dd0689, 0x15d0085520f5bbc3], dtype=np.uint64)
for op in (0, 1, 2, 3):
    best = 1e9
    for it in range(4):
        b.timer_start()
        rc = L.b200_ntt_dev(0, op, C.c_void_p(buf.ptr), m, g.ctypes.data_as(C.c_void_p))
        assert rc == 0
        best = min(best, b.timer_stop())
    print(f"2^{log_m} op {op}: {best:.3f} ms  {m / best / 1e6:.2f} Gelem/s")
buf.free()
