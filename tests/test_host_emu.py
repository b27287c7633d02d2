"""CPU tier: the DEVICE formulation executed on the host. tests/host_emu/emu.cpp compiles the product's
field / curve / MSM-index templates (libsnark_b200/csrc/{fp,curve,msm}.cuh) with
-DB200_HOST_EMULATE_PRIMS, i.e. the 32-bit mad.cc carry chains run through a bit-exact emulation of the
PTX carry flag and the kernel bodies run as plain loops. Checked against the golden vectors produced by
the unmodified reference. This pins, without a GPU: the Montgomery carry chains, the XYZZ group law
incl. its exceptional cases, the K-offset signed-digit extraction (== libff field_get_signed_digit) and
the bucket/task/reduction index logic."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
EMU_DIR = os.path.join(HERE, "host_emu")
CSRC = os.path.join(os.path.dirname(HERE), "libsnark_b200", "csrc")
CURVES = {"alt_bn128": 0, "bls12_377": 1, "bw6_761": 2}
BITS = {"alt_bn128": 254, "bls12_377": 253, "bw6_761": 377}


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(EMU_DIR, "libemu.so")
    src = os.path.join(EMU_DIR, "emu.cpp")
    if os.environ.get("B200_EMU_SO"):   # an instrumented build (AddressSanitizer / UBSan), see test_emulation_under_sanitizers
        so = os.environ["B200_EMU_SO"]
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".hpp"))]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(f) for f in hdrs + [src]):
        subprocess.run(["g++", "-std=c++17", "-O1", "-shared", "-fPIC", "-DB200_HOST_EMULATE_PRIMS", "-I", CSRC,
                        "-o", so, src], check=True)
    L = C.CDLL(so)
    vp, sz, i = C.c_void_p, C.c_size_t, C.c_int
    L.emu_field_op.argtypes = [i, i, i, vp, vp, vp, sz]
    L.emu_fp_convert.argtypes = [i, i, i, vp, vp, sz]
    L.emu_group_op.argtypes = [i, i, i, vp, vp, vp, sz]
    L.emu_signed_digits.argtypes = [i, vp, sz, i, vp, C.POINTER(C.c_int)]
    L.emu_msm.argtypes = [i, i, vp, sz, vp, i, vp]
    L.emu_msm_table.argtypes = [i, i, vp, sz, vp, i, vp]
    L.emu_batch_exp.argtypes = [i, i, vp, vp, vp, sz, vp]
    L.emu_geometry.argtypes = [C.c_uint64, i, i, vp]
    L.emu_msm_affine.argtypes = [i, i, vp, sz, vp, i, i, i, vp]
    L.emu_inverse_fast.argtypes = [i, i, vp, vp, sz]
    L.emu_msm_shared.argtypes = [i, i, vp, sz, vp, i, i, C.c_long, vp]
    L.emu_ntt.argtypes = [i, i, vp, sz, vp]
    L.emu_ext_ntt.argtypes = [i, i, vp, sz, vp]
    L.emu_step_ntt.argtypes = [i, i, vp, sz, vp]
    L.emu_r1cs_evaluate.argtypes = [i, sz, sz, sz, vp, vp, vp, vp, vp, vp, vp]
    return L


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_field_ops(emu, curve):
    g = np.load(os.path.join(GOLD, "fields.npz"))
    for fname, fid in [("fr", 0), ("fq", 1), ("fq2", 2)]:
        k = f"{curve}.{fname}"
        if k + ".a" not in g:
            continue
        a, b = np.ascontiguousarray(g[k + ".a"]), np.ascontiguousarray(g[k + ".b"])
        for oname, op in [("add", 0), ("sub", 1), ("mul", 2), ("sqr", 3), ("neg", 4), ("inv", 5)]:
            out = np.zeros_like(a)
            assert emu.emu_field_op(CURVES[curve], fid, op, _p(a), _p(b), _p(out), a.shape[0]) == 0
            assert np.array_equal(out, g[f"{k}.{oname}"]), (k, oname)


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_group_law(emu, curve):
    g = np.load(os.path.join(GOLD, "groups.npz"))
    for gname, gid in [("g1", 1), ("g2", 2)]:
        k = f"{curve}.{gname}"
        a, b = np.ascontiguousarray(g[k + ".a"]), np.ascontiguousarray(g[k + ".b"])
        for op, want in [(0, "add"), (1, "add"), (2, "dbl")]:
            out = np.zeros_like(a)
            assert emu.emu_group_op(CURVES[curve], gid, op, _p(a), _p(b), _p(out), a.shape[0]) == 0
            assert np.array_equal(out, g[f"{k}.{want}"]), (k, op)


@pytest.mark.parametrize("curve", list(CURVES))
def test_device_digit_extraction_equals_libff(emu, curve):
    """msm.cuh window_digit over t = s + K  ==  libff field_get_signed_digit (field_utils.tcc:82-115)."""
    g = np.load(os.path.join(GOLD, "digits.npz"))
    sc = np.ascontiguousarray(g[f"{curve}.scalars"])
    for c in [5, 12, 13, 16]:
        want = g[f"{curve}.c{c}"]
        W = C.c_int(0)
        out = np.zeros((sc.shape[0], 512), dtype=np.int64)
        flat = np.zeros(sc.shape[0] * 512, dtype=np.int64)
        assert emu.emu_signed_digits(CURVES[curve], _p(sc), sc.shape[0], c, _p(flat), C.byref(W)) == 0
        out = flat[: sc.shape[0] * W.value].reshape(sc.shape[0], W.value)
        assert W.value == (BITS[curve] + 1 + c - 1) // c
        k = min(W.value, want.shape[1])
        assert np.array_equal(out[:, :k], want[:, :k]), c
        assert not want[:, k:].any() and not out[:, k:].any()


def _emu_msm(emu, curve, gid, bases, scalars, force_c=0):
    bases, scalars = np.ascontiguousarray(bases), np.ascontiguousarray(scalars)
    out = np.zeros(bases.shape[1], dtype=np.uint64)
    assert emu.emu_msm(CURVES[curve], gid, _p(bases), scalars.shape[0], _p(scalars), force_c, _p(out)) == 0
    return out


def test_emulated_msm_kats(emu):
    g = np.load(os.path.join(GOLD, "msm.npz"))
    for k in [1, 2, 3, 4, 5, 256, 257]:
        got = _emu_msm(emu, "alt_bn128", 1, g["family.bases"][:k], g[f"family.scalars{k}"])
        assert np.array_equal(got, g[f"family.out{k}"]), k
    for c in [0, 5, 9, 13]:
        assert np.array_equal(_emu_msm(emu, "alt_bn128", 1, g["edge.bases"], g["edge.scalars"], c), g["edge.out"]), c


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_msm_pipeline(emu, curve):
    """zeros, ones, an infinity base and a repeated base through the sort / task / reduce index logic."""
    g = np.load(os.path.join(GOLD, "msm.npz"))
    for gname, gid in [("g1", 1), ("g2", 2)]:
        k = f"rand.{curve}.{gname}"
        for c in ([0, 6] if gid == 1 else [0]):
            assert np.array_equal(_emu_msm(emu, curve, gid, g[k + ".bases"], g[k + ".scalars"], c), g[k + ".out"]), (k, c)


def test_emulated_msm_heavy_bucket(emu):
    """All-equal scalars: every point lands in the same bucket of each window -> many tasks per bucket
    (the k_collapse path) and the narrow top window."""
    g = np.load(os.path.join(GOLD, "msm.npz"))
    bases = np.repeat(g["family.bases"], 3, axis=0)[:700]   # repeated bases as well: P + P collisions
    sc = np.repeat(g["rand.alt_bn128.g1.scalars"][:1], 700, axis=0)
    from oracle import port
    want = port.OraclePort().msm("alt_bn128", 1, bases, sc)
    assert np.array_equal(_emu_msm(emu, "alt_bn128", 1, bases, sc, 5), want)


def _emu_msm_table(emu, curve, gid, bases, scalars, c=0):
    bases, scalars = np.ascontiguousarray(bases), np.ascontiguousarray(scalars)
    out = np.zeros(bases.shape[1], dtype=np.uint64)
    assert emu.emu_msm_table(CURVES[curve], gid, _p(bases), scalars.shape[0], _p(scalars), c, _p(out)) == 0
    return out


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_msm_over_precomputed_tables(emu, curve):
    """b200_bases_precompute's formulation: rows [2^(c w)] P built by build_table_body, every window scattered into
    ONE bucket set, a single-window reduction — the same group element as the reference's multi_exp. (Each table
    entry costs an emulated field inversion, so the wide fields get few points and wide windows.)"""
    from oracle import port
    P = port.OraclePort()
    g = np.load(os.path.join(GOLD, "msm.npz"))
    npts = {"alt_bn128": 96, "bls12_377": 40, "bw6_761": 12}[curve]
    for gname, gid in [("g1", 1), ("g2", 2)]:
        k = f"rand.{curve}.{gname}"
        bases, sc = g[k + ".bases"][:npts], g[k + ".scalars"][:npts]
        want = g[k + ".out"] if npts == 96 else P.msm(curve, gid, bases, sc)
        for c in ([0, 13] if gid == 1 and curve == "alt_bn128" else [16]):
            assert np.array_equal(_emu_msm_table(emu, curve, gid, bases, sc, c), want), (k, c)
    if curve == "alt_bn128":
        for kk in [1, 2, 5, 257]:
            got = _emu_msm_table(emu, curve, 1, g["family.bases"][:kk], g[f"family.scalars{kk}"], 12)
            assert np.array_equal(got, g[f"family.out{kk}"]), kk
        assert np.array_equal(_emu_msm_table(emu, curve, 1, g["edge.bases"], g["edge.scalars"], 9), g["edge.out"])


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_batch_exp(emu, curve):
    """batch_exp.cuh (fixed-base table, one mixed addition per 8-bit digit, simultaneous inversion) on the host ==
    (coeff * v[i]) * g normalised, i.e. libff::batch_exp(_with_coeff) + batch_to_special, via the C oracle."""
    from oracle import port
    P = port.OraclePort()
    g = np.load(os.path.join(GOLD, "msm.npz"))
    for gname, gid in [("g1", 1), ("g2", 2)]:
        k = f"rand.{curve}.{gname}"
        base = np.ascontiguousarray(g[k + ".bases"][5])
        sc = np.ascontiguousarray(g[k + ".scalars"][:11]).copy()
        sc[3] = 0
        one = np.load(os.path.join(GOLD, "fields.npz"))
        coeff = np.ascontiguousarray(g[k + ".scalars"][20])
        for c in (None, coeff):
            out = np.zeros((sc.shape[0], base.shape[0]), dtype=np.uint64)
            assert emu.emu_batch_exp(CURVES[curve], gid, _p(base), None if c is None else _p(c), _p(sc), sc.shape[0], _p(out)) == 0
            eff = sc if c is None else P.field_op(curve, 0, 2, np.repeat(c[None, :], sc.shape[0], axis=0), sc)
            for i in range(sc.shape[0]):
                want = P.msm(curve, gid, base[None, :], eff[i:i + 1])
                assert np.array_equal(out[i], want), (k, i, c is not None)
        assert not out[3][2 * (base.shape[0] // 3):].any()       # zero scalar -> Z = 0
    # and against the vectors the unmodified reference produced (tests/golden/keygen.npz, tools/gen_golden.py)
    kg = np.load(os.path.join(GOLD, "keygen.npz"))
    sc, coeff = np.ascontiguousarray(kg[f"{curve}.scalars"]), np.ascontiguousarray(kg[f"{curve}.coeff"][0])
    for gname, gid in [("g1", 1), ("g2", 2)]:
        base = np.ascontiguousarray(kg[f"{curve}.{gname}.base"])
        for c, want in ((None, kg[f"{curve}.{gname}.out"]), (coeff, kg[f"{curve}.{gname}.out_coeff"])):
            out = np.zeros_like(want)
            assert emu.emu_batch_exp(CURVES[curve], gid, _p(base), None if c is None else _p(c), _p(sc), sc.shape[0], _p(out)) == 0
            assert np.array_equal(out, want), (gname, c is not None)


def test_geometry_invariants(emu):
    """Window / bucket / reduce geometry (msm_geometry.hpp) over the sizes and scalar widths in scope: the windows cover
    the scalar plus its sign carry, the reduce chunks tile the buckets exactly, CTA sizes are powers of two, and the
    chunk follows the reduce-thread budget."""
    for bits in (253, 254, 377):
        for log_n in list(range(0, 27, 2)) + [17, 21]:
            n = max(1, (1 << log_n) - (log_n % 3))
            for table in (0, 1):
                if table and 30 * n >= (1 << 31):
                    continue
                out = np.zeros(11, dtype=np.uint32)
                assert emu.emu_geometry(n, bits, table, _p(out)) == 0
                c, W, B, NB, cpw, rb, opw, collapse, bstride, estride, chunk = [int(x) for x in out]
                assert 5 <= c <= 20 and W * c >= bits + 1 and (W - 1) * c < bits + 1
                assert B == 1 << (c - 1) and NB == (B if table else W * B)
                assert chunk in (2, 4, 8) and cpw * chunk == B and opw * rb == cpw
                assert rb & (rb - 1) == 0 and 1 <= rb <= 128
                assert (bstride, estride) == ((0, n) if table else (B, 0))
                threads = (1 if table else W) * cpw
                assert chunk == 8 or threads <= 40000
                assert chunk == 2 or 2 * threads > 40000       # the next smaller chunk would exceed the budget
                assert collapse >= 4


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_inverse_fast(emu, curve):
    """Fp::inverse_fast / Fp2::inverse_fast (binary Euclid, the top of the affine rounds' inversion trees) == the
    reference's invert() on the golden field vectors; zero maps to zero."""
    g = np.load(os.path.join(GOLD, "fields.npz"))
    for fname, fid in [("fr", 0), ("fq", 1), ("fq2", 2)]:
        k = f"{curve}.{fname}"
        if k + ".a" not in g:
            continue
        a = np.ascontiguousarray(g[k + ".a"]).copy()
        want = np.ascontiguousarray(g[k + ".inv"]).copy()
        a[0] = 0
        want[0] = 0
        out = np.zeros_like(a)
        assert emu.emu_inverse_fast(CURVES[curve], fid, _p(a), _p(out), a.shape[0]) == 0
        assert np.array_equal(out, want), k


def _emu_msm_affine(emu, curve, gid, bases, scalars, c, table, rounds):
    bases, scalars = np.ascontiguousarray(bases), np.ascontiguousarray(scalars)
    out = np.zeros(bases.shape[1], dtype=np.uint64)
    assert emu.emu_msm_affine(CURVES[curve], gid, _p(bases), scalars.shape[0], _p(scalars), c, table, rounds, _p(out)) == 0
    return out


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_msm_affine_rounds(emu, curve):
    """The batched-affine rounds of msm_affine.cuh (padding of the counting sort, pair denominators, three-level
    inversion tree, affine additions, XYZZ accumulation over the survivors) give the reference's group element, for
    1-3 rounds, plain and over tables, with narrow windows so that buckets hold many entries."""
    from oracle import port
    P = port.OraclePort()
    g = np.load(os.path.join(GOLD, "msm.npz"))
    npts = {"alt_bn128": 96, "bls12_377": 40, "bw6_761": 16}[curve]
    for gname, gid in [("g1", 1), ("g2", 2)]:
        k = f"rand.{curve}.{gname}"
        bases, sc = g[k + ".bases"][:npts], g[k + ".scalars"][:npts]
        want = g[k + ".out"] if npts == 96 else P.msm(curve, gid, bases, sc)
        for rounds in ([1, 2, 3] if gid == 1 else [2]):
            assert np.array_equal(_emu_msm_affine(emu, curve, gid, bases, sc, 5, 0, rounds), want), (k, rounds)
    if curve == "alt_bn128":
        bases, sc = g["rand.alt_bn128.g1.bases"][:48], g["rand.alt_bn128.g1.scalars"][:48]
        want = P.msm(curve, 1, bases, sc)
        for rounds in (1, 3):
            assert np.array_equal(_emu_msm_affine(emu, curve, 1, bases, sc, 13, 1, rounds), want), ("table", rounds)


def test_emulated_msm_over_a_shared_sort(emu):
    """rt::SharedSort (msm_host.cuh): one sort of ALL the scalars with entries point | window << k, read by a base set with
    its own table row length that lacks the first `skip` points (the prover's A / B queries: skip 0; its L query: skip =
    num_inputs + 1). The result must be the MSM of that set with the scalars from `skip` on."""
    from oracle import port
    P = port.OraclePort()
    g = np.load(os.path.join(GOLD, "msm.npz"))
    for curve, gname, gid, npts in [("alt_bn128", "g1", 1, 96), ("alt_bn128", "g2", 2, 48), ("bls12_377", "g1", 1, 40)]:
        k = f"rand.{curve}.{gname}"
        bases, sc = np.ascontiguousarray(g[k + ".bases"][:npts]).copy(), np.ascontiguousarray(g[k + ".scalars"][:npts])
        third = bases.shape[1] // 3
        for i in np.nonzero(~bases[:, 2 * third:].any(axis=1))[0]:   # a shared sort takes no base at infinity (Z == 0)
            bases[i] = bases[(i + 1) % npts]
        for skip in (0, 11):
            sub = np.ascontiguousarray(bases[skip:])
            want = P.msm(curve, gid, sub, sc[skip:])
            for rounds in (1, 3):
                out = np.zeros(bases.shape[1], dtype=np.uint64)
                assert emu.emu_msm_shared(CURVES[curve], gid, _p(sub), npts, _p(sc), 6, rounds, skip, _p(out)) == 0
                assert np.array_equal(out, want), (k, skip, rounds)


def test_emulated_msm_affine_exceptional_pairs(emu):
    """Pairs the affine formula cannot add: P + P (repeated bases with equal scalars -> doubling), P + (-P) (a base with
    scalars s and r - s -> the neutral pair (0, 0)), pads, zero and one scalars, an infinity base, the reference KATs."""
    from oracle import port
    P = port.OraclePort()
    g = np.load(os.path.join(GOLD, "msm.npz"))
    # all-equal scalars over repeated bases: every bucket holds runs of identical points
    bases = np.repeat(g["family.bases"], 3, axis=0)[:300]
    sc = np.repeat(g["rand.alt_bn128.g1.scalars"][:1], 300, axis=0)
    want = P.msm("alt_bn128", 1, bases, sc)
    for rounds in (1, 2, 3):
        assert np.array_equal(_emu_msm_affine(emu, "alt_bn128", 1, bases, sc, 5, 0, rounds), want), rounds
    # s and -s on the same base, interleaved with ordinary entries, zeros and ones
    b2 = np.ascontiguousarray(g["rand.alt_bn128.g1.bases"][:64]).copy()
    s2 = np.ascontiguousarray(g["rand.alt_bn128.g1.scalars"][:64]).copy()
    neg = P.field_op("alt_bn128", 0, 4, s2[:16], s2[:16])      # op 4 = neg
    b2[16:32] = b2[:16]
    s2[16:32] = neg
    s2[40] = 0
    r_mod = 21888242871839275222246405745257275088548364400416034343698204186575808495617
    s2[41] = np.frombuffer(((1 << 256) % r_mod).to_bytes(32, "little"), dtype=np.uint64)   # Montgomery form of 1
    b2[50] = 0
    b2[50, 4] = b2[0, 8]    # (0, 1, 0): y = one copied from an affine point's z
    want = P.msm("alt_bn128", 1, b2, s2)
    for rounds in (1, 2, 3):
        for c in (5, 8):
            assert np.array_equal(_emu_msm_affine(emu, "alt_bn128", 1, b2, s2, c, 0, rounds), want), (rounds, c)
    for kk in [1, 2, 5, 257]:
        got = _emu_msm_affine(emu, "alt_bn128", 1, g["family.bases"][:kk], g[f"family.scalars{kk}"], 6, 0, 2)
        assert np.array_equal(got, g[f"family.out{kk}"]), kk
    assert np.array_equal(_emu_msm_affine(emu, "alt_bn128", 1, g["edge.bases"], g["edge.scalars"], 9, 0, 3), g["edge.out"])


def _emu_ntt(emu, curve, op, a, g=None):
    a = np.array(a, dtype=np.uint64, order="C", copy=True)
    g_ = None if g is None else np.ascontiguousarray(g, dtype=np.uint64)
    assert emu.emu_ntt(CURVES[curve], op, _p(a), a.shape[0], None if g_ is None else _p(g_)) == 0
    return a


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_ntt_passes(emu, curve):
    """k_ntt_pass (ntt.cuh: Stockham passes of up to 8 stage bits, rows gathered bit-reversed, radix-8 register steps, global
    twiddles merged into the butterflies, coset / 1/m tables fused into the first loads and last stores) run on the host
    with the launcher's own pass plan: the four transforms == the vectors of the unmodified reference for every committed
    size (one pass), and == the pinned C oracle at 2^9 ... 2^13 (two passes, uneven splits, a coset generator that is
    not the field's) — the index arithmetic of the device formulation, without a GPU."""
    from oracle import port
    P = port.OraclePort()
    g = np.load(os.path.join(GOLD, "fft.npz"))
    names = {0: "fft", 1: "ifft", 2: "cosetfft", 3: "icosetfft"}
    gen = P.fft  # noqa: F841  (keeps the import visibly used)
    for m in [2, 4, 8, 32, 256]:
        a = g[f"{curve}.m{m}.in"]
        for op, name in names.items():
            assert np.array_equal(_emu_ntt(emu, curve, op, a), g[f"{curve}.m{m}.{name}"]), (m, name)
    rng = np.random.default_rng(41)
    limbs = g[f"{curve}.m2.in"].shape[1]
    for lg in ([9, 10, 13] if curve == "alt_bn128" else [9, 11]):
        m = 1 << lg
        a = rng.integers(0, 1 << 62, size=(m, limbs), dtype=np.uint64)
        a[:, -1] >>= np.uint64(10)                         # canonical residues (below the modulus)
        other_g = a[5].copy()
        for op in range(4):
            assert np.array_equal(_emu_ntt(emu, curve, op, a), P.fft(curve, op, a)), (lg, op)
        for op in (2, 3):
            assert np.array_equal(_emu_ntt(emu, curve, op, a, other_g), P.fft(curve, op, a, other_g)), (lg, op, "g")
        assert np.array_equal(_emu_ntt(emu, curve, 1, _emu_ntt(emu, curve, 0, a)), a)


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_extended_domain(emu, curve):
    """The device formulation of extended_radix2_domain (ntt.cuh: ext_fft_pre_body / ext_ifft_post_body / mul_pow_body /
    ext_divz_body over the two-level power tables, constants from ext_domain_consts_host, two emulated radix-2 transforms per
    call) == the vectors the unmodified reference produced with FieldT::s lowered (tests/golden/extended.npz): all five
    operations at 2^2 ... 2^9 — the same functions ntt.cu launches, without a GPU."""
    g = np.load(os.path.join(GOLD, "extended.npz"))
    coset = g[f"{curve}.generator"]
    for s in [1, 2, 5, 8]:
        a = g[f"{curve}.s{s}.in"]
        for op, name in [(0, "fft"), (1, "ifft"), (2, "coset_fft"), (3, "icoset_fft"), (4, "div_z")]:
            out = np.array(a, dtype=np.uint64, order="C", copy=True)
            assert emu.emu_ext_ntt(CURVES[curve], op, _p(out), out.shape[0], _p(np.ascontiguousarray(coset))) == 0
            assert np.array_equal(out, g[f"{curve}.s{s}.{name}"]), (s, name)


@pytest.mark.parametrize("curve", list(CURVES))
def test_emulated_step_domain(emu, curve):
    """The device formulation of step_radix2_domain (ntt.cuh: step_fft_pre_body / fold_halve_body / step_ifft_tmp_body /
    step_ifft_post_body / step_zinv_body / step_divz_body, constants from step_domain_consts_host, the emulated radix-2
    transforms of sizes big and small) == the vectors the unmodified reference produced through get_evaluation_domain
    (tests/golden/step.npz): all five operations at m = 3 ... 288, incl. small = 1 and small = big / 2."""
    g = np.load(os.path.join(GOLD, "step.npz"))
    coset = np.ascontiguousarray(g[f"{curve}.generator"])
    for m in [3, 6, 24, 80, 288]:
        a = g[f"{curve}.m{m}.in"]
        for op, name in [(0, "fft"), (1, "ifft"), (2, "coset_fft"), (3, "icoset_fft"), (4, "div_z")]:
            out = np.array(a, dtype=np.uint64, order="C", copy=True)
            assert emu.emu_step_ntt(CURVES[curve], op, _p(out), m, _p(coset)) == 0
            assert np.array_equal(out, g[f"{curve}.m{m}.{name}"]), (m, name)


def test_emulated_r1cs_evaluation(emu):
    """The device's sparse R1CS evaluation (ntt.cuh r1cs_eval_body for every output element, the long rows cut by
    r1cs_long_plan and summed per segment and per row as the two long-row kernels do) == the aA, aB, aC the unmodified
    reference built (r1cs_to_qap.tcc:236-248, :293-296) for the committed circuits — each of which closes with a constraint
    over all variables, i.e. a long row — on all three fields, incl. the zero padding and the input-consistency rows."""
    cases = [("step.npz", "alt_bn128", "witness")]
    z = np.load(os.path.join(GOLD, "zk.npz"))
    for ci, nc, _s, _m in z["cases"]:
        cases.append(("zk.npz", list(CURVES)[int(ci)], f"{list(CURVES)[int(ci)]}.nc{int(nc)}"))
    for fname, curve, k in cases:
        g = np.load(os.path.join(GOLD, fname))
        n_c, ni, nv, m = [int(x) for x in g[k + ".sizes"]]
        row_ptr = np.ascontiguousarray(g[k + ".row_ptr"], dtype=np.uint32)
        col = np.ascontiguousarray(g[k + ".col"], dtype=np.uint32)
        val = np.ascontiguousarray(g[k + ".val"], dtype=np.uint64)
        w = np.ascontiguousarray(g[k + ".w"], dtype=np.uint64)
        assert w.shape[0] == nv + 1 and row_ptr.shape[0] == 3 * n_c + 1
        out = [np.full((m, w.shape[1]), 0xAB, dtype=np.uint64) for _ in range(3)]   # garbage: every element must be written
        nlong = emu.emu_r1cs_evaluate(CURVES[curve], n_c, ni, m, _p(row_ptr), _p(col), _p(val), _p(w), *[_p(o) for o in out])
        assert nlong >= (1 if nv > 70 else 0), (k, nlong)   # the closing constraint sums every variable
        for got, name in zip(out, ["aA", "aB", "aC"]):
            assert np.array_equal(got, g[f"{k}.{name}"]), (k, name)
