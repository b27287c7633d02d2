"""The reference's own known-answer inputs replayed on the CUDA path:
 - the signed-digit goldens of libff's field tests (test_fields.cpp:282-395) against the digits the device's
   k_prepare + window_digit produce;
 - TestEdgeCase1 and the closed-form family of test_multiexp.cpp (:204-255, :337-383) through b200_msm with and
   without precomputed tables;
 - scalars whose every lower window digit is -2^(c-1) (the carry chain of field_get_signed_digit at its extreme);
 - knowledge-commitment MSMs on all three curves against tests/golden/msm.npz.
The goldens were written by the unmodified reference (tools/gen_golden.py); no reference code runs here except
where a test asks `ref` for a fresh answer."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CURVES = ["alt_bn128", "bls12_377", "bw6_761"]
BITS = {"alt_bn128": 254, "bls12_377": 253, "bw6_761": 377}


@pytest.mark.parametrize("curve", CURVES)
def test_device_signed_digits_equal_libff_goldens(b200, curve):
    g = np.load(os.path.join(GOLD, "digits.npz"))
    sc = np.ascontiguousarray(g[f"{curve}.scalars"])
    for c in [2, 5, 12, 13, 16, 21]:
        want = g[f"{curve}.c{c}"]
        got = b200.test_signed_digits(curve, sc, c)
        W = (BITS[curve] + 1 + c - 1) // c
        assert got.shape[1] == W
        # every window below the top one: libff's digit exactly
        assert np.array_equal(got[:, :W - 1], want[:, :W - 1]), (curve, c)
        # the top window keeps its unsigned value (msm.cuh header: it cannot exceed 2^(c-1), the bucket count), i.e.
        # libff's digit W-1 with its carry digit W folded in: d[W-1] + 2^c d[W]
        top = want[:, W - 1].copy()
        if want.shape[1] > W:
            top += want[:, W] << c
            assert not want[:, W + 1:].any()
        assert np.array_equal(got[:, W - 1], top), (curve, c)
        assert (got[:, W - 1] >= 0).all() and (got[:, W - 1] <= 1 << (c - 1)).all()


def _msm_both_ways(b200, curve, group, bases, scalars):
    """b200_msm (cold entry point) and a resident set before / after b200_bases_precompute."""
    outs = [b200.multi_exp(curve, group, bases, scalars)]
    rb = b200.ResidentBases(curve, group, bases)
    try:
        outs.append(rb.multi_exp(scalars))
        rb.precompute()
        outs.append(rb.multi_exp(scalars))
    finally:
        rb.free()
    return outs


def test_reference_msm_kats_on_the_device(b200):
    g = np.load(os.path.join(GOLD, "msm.npz"))
    for k in [1, 2, 3, 4, 5, 256, 257]:
        for got in _msm_both_ways(b200, "alt_bn128", b200.G1, g["family.bases"][:k], g[f"family.scalars{k}"]):
            assert np.array_equal(got, g[f"family.out{k}"]), ("family", k)
    # TestEdgeCase1: a negative top digit with carry (needs num_bits + 2 digits in the reference)
    for got in _msm_both_ways(b200, "alt_bn128", b200.G1, g["edge.bases"], g["edge.scalars"]):
        assert np.array_equal(got, g["edge.out"])
    # the edge scalars padded into larger MSMs, so that other window widths (c = 5 ... 12) see the same carries
    for n in [40, 700, 5000]:
        bases = np.concatenate([g["edge.bases"], np.repeat(g["family.bases"], 1 + n // 257, axis=0)[: n - 2]])
        scalars = np.concatenate([g["edge.scalars"], np.zeros((n - 2, 4), dtype=np.uint64)])
        for got in _msm_both_ways(b200, "alt_bn128", b200.G1, bases, scalars):
            assert np.array_equal(got, g["edge.out"]), n


@pytest.mark.parametrize("curve", CURVES)
def test_reference_random_kats_on_the_device(b200, curve):
    g = np.load(os.path.join(GOLD, "msm.npz"))
    for gname, gid in [("g1", 1), ("g2", 2)]:
        k = f"rand.{curve}.{gname}"
        for got in _msm_both_ways(b200, curve, gid, g[k + ".bases"], g[k + ".scalars"]):
            assert np.array_equal(got, g[k + ".out"]), k


def _all_min_digit_scalar(bits, c, top):
    """The scalar whose signed digits are -2^(c-1) in every window below the last and `top` in the last:
    t = s + K has all-zero lower base-2^c digits (msm.cuh header)."""
    W = (bits + 1 + c - 1) // c
    K = sum(1 << (c - 1 + c * w) for w in range(W - 1))
    return (top << (c * (W - 1))) - K


@pytest.mark.parametrize("curve,group,n", [("alt_bn128", 1, 300), ("alt_bn128", 1, 1 << 14), ("alt_bn128", 2, 2000),
                                           ("bls12_377", 1, 3000), ("bw6_761", 1, 600)])
def test_all_digits_at_the_negative_extreme(ref, b200, curve, group, n):
    """SURVEY.md §8(d) config 2's adversarial family: every point lands in bucket 2^(c-1) of every window, with the
    sign bit set — for the window width of the plain geometry, of the table geometry, and their neighbours."""
    L = b200.lib()
    bases = ref.make_bases(curve, group, n, 31)
    rb = b200.ResidentBases(curve, group, bases)
    try:
        c_plain = L.b200_msm_window_bits(b200.CURVES[curve], n)
        rb.precompute()
        c_table = L.b200_bases_table_window_bits(rb.handle)
        r = b200.FR_MODULUS[b200.CURVES[curve]]
        vals = []
        for c in sorted({c_plain, c_table, c_plain + 1, max(2, c_plain - 1)}):
            for top in (1, 2):
                s = _all_min_digit_scalar(BITS[curve], c, top)
                if 0 < s < r:
                    vals.append(s)
        assert len(vals) >= 4
        for s in vals:
            sc = b200.scalars_from_ints(curve, [s] * n)
            want = ref.msm(curve, group, bases, sc)
            assert np.array_equal(rb.multi_exp(sc), want), hex(s)
            assert np.array_equal(b200.multi_exp(curve, group, bases, sc), want), hex(s)
        # mixed: the extreme scalars interleaved with random ones
        sc = ref.field_random(curve, 0, 77, n)
        sc[::3] = b200.scalars_from_ints(curve, [vals[0]])[0]
        sc[1::7] = b200.scalars_from_ints(curve, [vals[-1]])[0]
        want = ref.msm(curve, group, bases, sc)
        assert np.array_equal(rb.multi_exp(sc), want)
    finally:
        rb.free()


@pytest.mark.parametrize("curve", CURVES)
def test_kc_msm_goldens_on_the_device(b200, curve):
    """kc_multi_exp_with_mixed_addition (kc_multiexp.tcc:31-113) over the committed sparse vectors."""
    g = np.load(os.path.join(GOLD, "msm.npz"))
    idx, vals, sc = g[f"kc.{curve}.indices"], g[f"kc.{curve}.values"], g[f"kc.{curve}.scalars"]
    for lo, hi in [(0, 64), (5, 50)]:
        g2, g1 = b200.kc_multi_exp_with_mixed_addition(curve, idx, vals, lo, hi, sc[: hi - lo])
        assert np.array_equal(g2, g[f"kc.{curve}.out_g2.{lo}_{hi}"]), (curve, lo, hi)
        assert np.array_equal(g1, g[f"kc.{curve}.out_g1.{lo}_{hi}"]), (curve, lo, hi)


@pytest.mark.parametrize("curve", ["bls12_377", "bw6_761"])
def test_kc_msm_other_curves_against_reference(ref, b200, curve):
    from oracle import refo
    nnz, dom = 300, 450
    values = np.concatenate([ref.make_bases(curve, refo.G2, nnz, 21), ref.make_bases(curve, refo.G1, nnz, 22)], axis=1)
    rng = np.random.default_rng(8)
    indices = np.sort(rng.choice(dom, nnz, replace=False)).astype(np.uint64)
    scalars = ref.field_random(curve, refo.FR, 123, dom)
    scalars[2] = 0
    scalars[9] = ref.field_consts(curve, refo.FR)["one"]
    for lo, hi in [(0, dom), (7, 401)]:
        sc = scalars[: hi - lo]
        w2, w1 = ref.kc_msm(curve, indices, values, dom, lo, hi, sc)
        g2o, g1o = b200.kc_multi_exp_with_mixed_addition(curve, indices, values, lo, hi, sc)
        assert np.array_equal(g2o, w2) and np.array_equal(g1o, w1), (curve, lo, hi)
