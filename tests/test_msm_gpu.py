"""GPU parity: b200_msm (through the C-ABI) vs the reference's libff::multi_exp on the same inputs.
Bit-exact comparison of the normalised result."""
import numpy as np
import pytest

from oracle import refo

pytestmark = pytest.mark.gpu

CURVE_GROUPS = [("alt_bn128", refo.G1), ("alt_bn128", refo.G2), ("bls12_377", refo.G1), ("bls12_377", refo.G2),
                ("bw6_761", refo.G1), ("bw6_761", refo.G2)]


def _inputs(ref, curve, group, n, seed=1):
    bases = ref.make_bases(curve, group, n, seed)
    scalars = ref.field_random(curve, refo.FR, 1000 * seed, n)
    return bases, scalars


@pytest.mark.parametrize("curve,group", CURVE_GROUPS)
@pytest.mark.parametrize("n", [1, 2, 3, 5, 257, 4096])
def test_msm_matches_reference(ref, b200, curve, group, n):
    bases, scalars = _inputs(ref, curve, group, n)
    want = ref.msm(curve, group, bases, scalars)
    got = b200.multi_exp(curve, group, bases, scalars)
    assert np.array_equal(got, want)


def test_msm_empty(b200, ref):
    i = ref.info("alt_bn128")
    got = b200.multi_exp("alt_bn128", refo.G1, np.zeros((0, i["g1_limbs"]), np.uint64), np.zeros((0, 4), np.uint64),
                         base_stride=96)
    zero = ref.group_op("alt_bn128", refo.G1, refo.GOP_TO_AFFINE, np.zeros((1, 12), np.uint64))  # Z = 0
    assert got[8:].tolist() == [0, 0, 0, 0]  # Z == 0  <=> infinity
    assert np.array_equal(got, zero[0])


@pytest.mark.parametrize("log_n", [16, 18, 20])
def test_msm_alt_bn128_g1_large(ref, b200, log_n):
    n = 1 << log_n
    bases, scalars = _inputs(ref, "alt_bn128", refo.G1, n, seed=3)
    want = ref.msm("alt_bn128", refo.G1, bases, scalars, chunks=ref.num_threads())
    got = b200.multi_exp("alt_bn128", refo.G1, bases, scalars)
    assert np.array_equal(got, want)


def test_msm_closed_form_family(ref, b200):
    """libff test_multiexp.cpp:204-255: bases [1]G..[n]G, scalars n..1; expected = reference result."""
    for n in [1, 2, 3, 4, 5, 256, 257]:
        k = np.zeros((n, 4), dtype=np.uint64)
        k[:, 0] = np.arange(1, n + 1, dtype=np.uint64)
        kk = ref.to_mont("alt_bn128", refo.FR, k)
        bases = ref.scalar_mul_gen("alt_bn128", refo.G1, kk)
        s = np.zeros((n, 4), dtype=np.uint64)
        s[:, 0] = np.arange(n, 0, -1, dtype=np.uint64)
        ss = ref.to_mont("alt_bn128", refo.FR, s)
        want = ref.msm("alt_bn128", refo.G1, bases, ss, method=refo.MSM_NAIVE_PLAIN)
        got = b200.multi_exp("alt_bn128", refo.G1, bases, ss)
        assert np.array_equal(got, want), n


def test_msm_adversarial_scalars(ref, b200):
    """all-equal scalars, zeros/ones mix, repeated bases (P+P and P-P collisions)."""
    n = 2048
    curve, group = "alt_bn128", refo.G1
    bases, scalars = _inputs(ref, curve, group, n, seed=5)
    eq = np.repeat(scalars[:1], n, axis=0)
    assert np.array_equal(b200.multi_exp(curve, group, bases, eq), ref.msm(curve, group, bases, eq))
    one = ref.field_consts(curve, refo.FR)["one"]
    mix = scalars.copy()
    mix[::2] = 0
    mix[1::4] = one
    want = ref.msm(curve, group, bases, mix, method=refo.MSM_FILTER_ONE_ZERO)
    got, stats = b200.multi_exp_filter_one_zero(curve, group, bases, mix, return_stats=True)
    assert np.array_equal(got, want)
    assert stats.tolist() == [n // 2, n // 4, n // 4]
    # repeated bases with equal and opposite scalars
    rep = np.repeat(bases[:4], n // 4, axis=0)
    assert np.array_equal(b200.multi_exp(curve, group, rep, scalars), ref.msm(curve, group, rep, scalars))
    neg = ref.field_op(curve, refo.FR, refo.OP_NEG, scalars[: n // 2])
    pm = np.concatenate([scalars[: n // 2], neg])
    rep2 = np.concatenate([bases[: n // 2], bases[: n // 2]])
    got = b200.multi_exp(curve, group, rep2, pm)
    assert got[8:].tolist() == [0, 0, 0, 0]
    assert np.array_equal(got, ref.msm(curve, group, rep2, pm))


def test_msm_infinity_bases(ref, b200):
    curve, group = "alt_bn128", refo.G1
    n = 300
    bases, scalars = _inputs(ref, curve, group, n, seed=7)
    bases = bases.copy()
    zero = ref.group_op(curve, group, refo.GOP_TO_AFFINE, np.zeros((1, 12), np.uint64))[0]
    bases[5] = zero
    bases[77] = zero
    assert np.array_equal(b200.multi_exp(curve, group, bases, scalars), ref.msm(curve, group, bases, scalars))


def test_kc_msm(ref, b200):
    curve = "alt_bn128"
    i = ref.info(curve)
    nnz, dom = 500, 700
    g2 = ref.make_bases(curve, refo.G2, nnz, 11)
    g1 = ref.make_bases(curve, refo.G1, nnz, 12)
    values = np.concatenate([g2, g1], axis=1)
    rng = np.random.default_rng(5)
    indices = np.sort(rng.choice(dom, nnz, replace=False)).astype(np.uint64)
    scalars = ref.field_random(curve, refo.FR, 99, dom)
    scalars[3] = 0
    scalars[10] = ref.field_consts(curve, refo.FR)["one"]
    for lo, hi in [(0, dom), (13, 555)]:
        sc = scalars[: hi - lo]
        w2, w1 = ref.kc_msm(curve, indices, values, dom, lo, hi, sc)
        g2o, g1o = b200.kc_multi_exp_with_mixed_addition(curve, indices, values, lo, hi, sc)
        assert np.array_equal(g2o, w2) and np.array_equal(g1o, w1)


def test_resident_and_sharded(ref, b200):
    curve, group = "alt_bn128", refo.G1
    n = 1 << 14
    bases, scalars = _inputs(ref, curve, group, n, seed=9)
    want = ref.msm(curve, group, bases, scalars, chunks=4)
    rb = b200.ResidentBases(curve, group, bases)
    assert np.array_equal(rb.multi_exp(scalars), want)
    # 4 shards by base range, window sums gathered and folded
    q = n // 4
    ws = np.concatenate([rb.window_sums(scalars[k * q:(k + 1) * q], first=k * q, n=q) for k in range(4)])
    assert np.array_equal(b200.fold_windows(curve, group, q, ws, 4), want)
    rb.free()


@pytest.mark.parametrize("curve,group,log_n", [("bls12_377", refo.G1, 18), ("bw6_761", refo.G1, 16),
                                               ("alt_bn128", refo.G2, 18)])
def test_msm_full_size_linearity(ref, b200, curve, group, log_n):
    """BASELINE config 5 sizes through a size-independent property: one device MSM over
    the whole range equals the group sum of reference MSMs over disjoint quarters, and
    MSM(bases, s1 + s2) == MSM(bases, s1) + MSM(bases, s2) on the device."""
    n = 1 << log_n
    bases, scalars = _inputs(ref, curve, group, n, seed=11)
    got = b200.multi_exp(curve, group, bases, scalars)
    q = n // 4
    parts = [ref.msm(curve, group, bases[i * q:(i + 1) * q], scalars[i * q:(i + 1) * q], chunks=ref.num_threads())
             for i in range(4)]
    acc = parts[0]
    for p_ in parts[1:]:
        acc = ref.group_op(curve, group, refo.GOP_ADD, acc[None, :], p_[None, :])[0]
    acc = ref.group_op(curve, group, refo.GOP_TO_AFFINE, acc[None, :])[0]
    assert np.array_equal(got, acc)
    s2 = ref.field_random(curve, refo.FR, 991, n)
    ssum = ref.field_op(curve, refo.FR, 0, scalars, s2)   # op 0 = add
    lhs = b200.multi_exp(curve, group, bases, ssum)
    r2 = b200.multi_exp(curve, group, bases, s2)
    rhs = ref.group_op(curve, group, refo.GOP_ADD, got[None, :], r2[None, :])
    rhs = ref.group_op(curve, group, refo.GOP_TO_AFFINE, rhs)[0]
    assert np.array_equal(lhs, rhs)


@pytest.mark.parametrize("curve,group", CURVE_GROUPS)
def test_msm_over_precomputed_tables(ref, b200, curve, group):
    """b200_bases_precompute: rows [2^(c w)] P built on the device, all windows into one bucket set. Same normalised
    point as the reference for whole sets, sub-ranges, zero/one/equal scalars, an infinity base and window sums."""
    n = 3000 if curve != "bw6_761" else 700
    bases, scalars = _inputs(ref, curve, group, n, seed=11)
    bases = bases.copy()
    bases[7] = ref.group_op(curve, group, refo.GOP_TO_AFFINE, np.zeros((1, bases.shape[1]), np.uint64))[0]  # infinity
    one = ref.field_consts(curve, refo.FR)["one"]
    scalars[3] = 0
    scalars[4] = one
    rb = b200.ResidentBases(curve, group, bases)
    try:
        plain = rb.multi_exp(scalars)
        rows = rb.precompute()
        assert rows >= 8 and rb.precompute() == rows            # idempotent
        want = ref.msm(curve, group, bases, scalars)
        assert np.array_equal(plain, want)
        assert np.array_equal(rb.multi_exp(scalars), want)
        for first, cnt in [(0, 1), (5, n // 3), (n - 17, 17), (1, n - 1)]:
            w = ref.msm(curve, group, bases[first:first + cnt], scalars[first:first + cnt])
            assert np.array_equal(rb.multi_exp(scalars[first:first + cnt], first=first, n=cnt), w), (first, cnt)
        eq = np.repeat(scalars[:1], n, axis=0)
        assert np.array_equal(rb.multi_exp(eq), ref.msm(curve, group, bases, eq))
        # window sums of two sub-ranges folded like two shards
        half = n // 2
        wins = np.concatenate([rb.window_sums(scalars[:half], first=0, n=half),
                               rb.window_sums(scalars[half:2 * half], first=half, n=half)])
        got = b200.fold_windows(curve, group, half, wins, 2)
        assert np.array_equal(got, ref.msm(curve, group, bases[:2 * half], scalars[:2 * half]))
    finally:
        rb.free()


@pytest.mark.parametrize("log_n", [16, 18, 20])
def test_msm_tables_large(ref, b200, log_n):
    n = 1 << log_n
    bases, scalars = _inputs(ref, "alt_bn128", refo.G1, n, seed=4)
    rb = b200.ResidentBases("alt_bn128", refo.G1, bases)
    try:
        rb.precompute()
        want = ref.msm("alt_bn128", refo.G1, bases, scalars, chunks=ref.num_threads())
        assert np.array_equal(rb.multi_exp(scalars), want)
    finally:
        rb.free()
