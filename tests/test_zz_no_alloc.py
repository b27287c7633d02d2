"""b200_ntt_dev_noalloc (the transforms over a static device scratch, no cudaMalloc) against b200_ntt_dev, through
the C-ABI. Collected last on purpose. Passes alone in a fresh process and after the
other tests: the entry point rebuilds its tables whenever curve, size, generator or coset direction change."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

OPS = ("FFT", "IFFT", "COSET_FFT", "ICOSET_FFT")


def _gen(b200, curve, seed):
    return b200.random_scalars(curve, 1, seed=seed)[0]


def _run(b200, entry, curve, op, a, g):
    """One in-place transform of the device copy of `a` through `entry` (a C-ABI function), result back on the host."""
    buf = b200.DeviceBuffer(a.nbytes)
    try:
        buf.upload(a)
        g_ = None if g is None else np.ascontiguousarray(g, dtype=np.uint64)
        rc = entry(b200.CURVES[curve], op, C.c_void_p(buf.ptr), a.shape[0], None if g_ is None else g_.ctypes.data_as(C.c_void_p))
        assert rc == b200.OK, b200.lib().b200_last_error().decode()
        return buf.download(a.shape)
    finally:
        buf.free()


@pytest.mark.parametrize("curve", ["alt_bn128", "bls12_377", "bw6_761"])
def test_noalloc_matches_ntt_dev_two_calls_in_a_row(b200, curve):
    L = b200.lib()
    for lg in (1, 2, 5, 8, 9, 13, 16):
        m = 1 << lg
        a = b200.random_scalars(curve, m, seed=100 + lg)
        for k, name in enumerate(OPS):
            op = getattr(b200, name)
            g = _gen(b200, curve, 7 + k) if "COSET" in name else None
            want = _run(b200, L.b200_ntt_dev, curve, op, a, g)
            first = _run(b200, L.b200_ntt_dev_noalloc, curve, op, a, g)
            second = _run(b200, L.b200_ntt_dev_noalloc, curve, op, a, g)  # tables now come from the scratch as left
            assert np.array_equal(first, want), (curve, lg, name, "first call")
            assert np.array_equal(second, want), (curve, lg, name, "second call")


def test_noalloc_key_changes_between_calls(b200):
    """The scratch keeps the last call's tables: a change of generator, size, direction or curve must rebuild them."""
    L = b200.lib()
    m = 1 << 10
    a = b200.random_scalars("alt_bn128", m, seed=5)
    g1, g2 = _gen(b200, "alt_bn128", 1), _gen(b200, "alt_bn128", 2)
    seq = [("alt_bn128", b200.COSET_FFT, a, g1), ("alt_bn128", b200.COSET_FFT, a, g2),
           ("alt_bn128", b200.ICOSET_FFT, a, g2), ("alt_bn128", b200.FFT, a, None),
           ("alt_bn128", b200.ICOSET_FFT, a, g1), ("alt_bn128", b200.IFFT, a[:256], None),
           ("bls12_377", b200.COSET_FFT, b200.random_scalars("bls12_377", m, seed=6), _gen(b200, "bls12_377", 3)),
           ("alt_bn128", b200.COSET_FFT, a, g1)]
    for curve, op, x, g in seq:
        got = _run(b200, L.b200_ntt_dev_noalloc, curve, op, x, g)
        assert np.array_equal(got, _run(b200, L.b200_ntt_dev, curve, op, x, g)), (curve, op, x.shape[0])
    # round trip inside the scratch path alone
    back = _run(b200, L.b200_ntt_dev_noalloc, "alt_bn128", b200.IFFT,
                _run(b200, L.b200_ntt_dev_noalloc, "alt_bn128", b200.FFT, a, None), None)
    assert np.array_equal(back, a)


def test_noalloc_rejects_sizes_beyond_the_scratch(b200):
    L = b200.lib()
    m = 1 << 17
    buf = b200.DeviceBuffer(m * 32)
    try:
        rc = L.b200_ntt_dev_noalloc(b200.ALT_BN128, b200.FFT, C.c_void_p(buf.ptr), m, None)
        assert rc == b200.ERR_ARG and b"exceeds the static scratch" in L.b200_last_error()
        rc = L.b200_ntt_dev_noalloc(b200.ALT_BN128, b200.FFT, C.c_void_p(buf.ptr), 3 << 10, None)
        assert rc == b200.ERR_DOMAIN
        rc = L.b200_ntt_dev_noalloc(b200.ALT_BN128, b200.COSET_FFT, C.c_void_p(buf.ptr), 1 << 10, None)
        assert rc == b200.ERR_ARG
        assert L.b200_ntt_dev_noalloc(b200.ALT_BN128, b200.FFT, None, 1 << 10, None) == b200.ERR_ARG
    finally:
        buf.free()
