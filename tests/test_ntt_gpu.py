"""GPU parity: b200_ntt / b200_witness_map_h (through the C-ABI) vs libfqfft's basic_radix2_domain
and libsnark's r1cs_to_qap_witness_map run by the unmodified reference. Bit-exact."""
import numpy as np
import pytest

from oracle import refo

pytestmark = pytest.mark.gpu

CURVES = ["alt_bn128", "bls12_377", "bw6_761"]


def _gen(ref, curve):
    return ref.field_consts(curve, refo.FR)["generator"]


@pytest.mark.parametrize("curve", CURVES)
@pytest.mark.parametrize("log_m", [1, 2, 3, 4, 5, 8, 9, 10, 13, 16])
def test_fft_ops_match_reference(ref, b200, curve, log_m):
    if curve != "alt_bn128" and log_m in (5, 9, 13):
        pytest.skip("covered on alt_bn128")
    m = 1 << log_m
    a = ref.field_random(curve, refo.FR, 7 * m, m)
    g = _gen(ref, curve)
    dom = b200.get_evaluation_domain(curve, m)
    assert np.array_equal(dom.FFT(a), ref.fft(curve, refo.FFT, a))
    assert np.array_equal(dom.iFFT(a), ref.fft(curve, refo.IFFT, a))
    assert np.array_equal(dom.cosetFFT(a, g), ref.fft(curve, refo.COSET_FFT, a, g))
    assert np.array_equal(dom.icosetFFT(a, g), ref.fft(curve, refo.ICOSET_FFT, a, g))
    assert np.array_equal(dom.divide_by_Z_on_coset(a), ref.fft(curve, refo.DIV_Z_COSET, a))


def test_fft_other_coset_generator(ref, b200):
    curve, m = "alt_bn128", 1 << 10
    a = ref.field_random(curve, refo.FR, 1, m)
    g = ref.field_random(curve, refo.FR, 424242, 1)[0]
    dom = b200.get_evaluation_domain(curve, m)
    assert np.array_equal(dom.cosetFFT(a, g), ref.fft(curve, refo.COSET_FFT, a, g))
    assert np.array_equal(dom.icosetFFT(a, g), ref.fft(curve, refo.ICOSET_FFT, a, g))
    g2 = _gen(ref, curve)  # back to the default: tables are rebuilt
    assert np.array_equal(dom.cosetFFT(a, g2), ref.fft(curve, refo.COSET_FFT, a, g2))


def test_fft_small_known_vector(ref, b200):
    """libfqfft evaluation_domain_test.cpp:40-80: f = {2, 5, 3, 8}, FFT(f)[i] == f(omega^i)."""
    curve = "alt_bn128"
    f = np.zeros((4, 4), dtype=np.uint64)
    f[:, 0] = [2, 5, 3, 8]
    fm = ref.to_mont(curve, refo.FR, f)
    got = b200.get_evaluation_domain(curve, 4).FFT(fm)
    want = ref.poly_eval_at_domain(curve, fm, np.arange(4, dtype=np.uint64))
    assert np.array_equal(got, want)


def test_domain_size_exception(b200):
    """basic_radix2_domain.tcc:57,65: a.size() != m throws DomainSizeException; a size that is neither 2^k nor
    2^k + 2^r has no radix-2 domain (the factory never asks for one: it rounds, get_evaluation_domain.tcc:41-58)."""
    a = np.zeros((6, 4), dtype=np.uint64)
    with pytest.raises(b200.DomainSizeException):
        b200.evaluation_domain("alt_bn128", 7).FFT(np.zeros((7, 4), dtype=np.uint64))
    with pytest.raises(b200.DomainSizeException):
        b200.get_evaluation_domain("alt_bn128", 8).FFT(a)


@pytest.mark.parametrize("log_m", [20, 24])
def test_fft_large_roundtrip_and_spot_check(ref, b200, log_m):
    """Size-independent properties at scale: iFFT(FFT(a)) == a, icosetFFT(cosetFFT(a)) == a, and
    naive Horner evaluation at a few domain points."""
    curve, m = "alt_bn128", 1 << log_m
    a = ref.field_random(curve, refo.FR, 5, m)
    g = _gen(ref, curve)
    dom = b200.get_evaluation_domain(curve, m)
    fa = dom.FFT(a)
    assert np.array_equal(dom.iFFT(fa), a)
    assert np.array_equal(dom.icosetFFT(dom.cosetFFT(a, g), g), a)
    idx = np.array([0, 1, 2, m // 2, m - 1, 12345, 777777 % m], dtype=np.uint64)
    assert np.array_equal(fa[idx.astype(np.int64)], ref.poly_eval_at_domain(curve, a, idx))
    if log_m <= 20:     # the reference's own FFT at 2^24 takes minutes on the host; properties cover it
        assert np.array_equal(fa, ref.fft(curve, refo.FFT, a))


@pytest.mark.parametrize("curve,log_m", [("alt_bn128", 4), ("alt_bn128", 10), ("alt_bn128", 16), ("bls12_377", 12),
                                         ("bw6_761", 11)])
def test_witness_map_h(ref, b200, curve, log_m):
    """coefficients_for_H of r1cs_to_qap_witness_map: the reference's own H vs the device chain fed with
    the reference's aA, aB, aC evaluations (recovered from its H? no: recomputed here from the R1CS)."""
    m = 1 << log_m
    ni = 2 if log_m < 6 else 10
    fx = ref.groth16_setup(curve, m - ni - 1, ni, real_keys=False)
    assert fx.sizes["m"] == m
    want, _ = fx.witness_map()
    aA, aB, aC = fx.abc_evaluations()
    got = b200.r1cs_to_qap_witness_map_h(curve, aA, aB, aC)
    assert np.array_equal(got, want)
    fx.close()


@pytest.mark.parametrize("curve", CURVES)
@pytest.mark.parametrize("m", [3, 5, 6, 12, 20, 48, 1024 + 64, (1 << 14) + (1 << 13), (1 << 16) + 1, (1 << 16) + 256])
def test_step_radix2_domain_matches_reference(ref, b200, curve, m):
    """step_radix2_domain (sizes 2^k + 2^r: what get_evaluation_domain returns for every non-power-of-two size,
    get_evaluation_domain.tcc:41-58): all four transforms and divide_by_Z_on_coset, bit-exact
    (step_radix2_domain.tcc:65-276)."""
    a = ref.field_random(curve, refo.FR, 100 + m, m)
    g = _gen(ref, curve)
    dom = b200.get_evaluation_domain(curve, m)
    assert dom.m == m
    for name, op in [("FFT", refo.FFT), ("iFFT", refo.IFFT), ("cosetFFT", refo.COSET_FFT), ("icosetFFT", refo.ICOSET_FFT)]:
        got = getattr(dom, name)(a, g) if "coset" in name else getattr(dom, name)(a)
        assert np.array_equal(got, ref.fft(curve, op, a, g, via_factory=True)), (name, m)
    assert np.array_equal(dom.divide_by_Z_on_coset(a), ref.fft(curve, refo.DIV_Z_COSET, a, g, via_factory=True))
    assert np.array_equal(dom.iFFT(dom.FFT(a)), a)


@pytest.mark.parametrize("curve,s", [("alt_bn128", 1), ("alt_bn128", 3), ("alt_bn128", 10), ("alt_bn128", 15), ("alt_bn128", 17),
                                     ("bls12_377", 9), ("bls12_377", 16), ("bw6_761", 7), ("bw6_761", 12)])
def test_extended_radix2_domain_matches_reference(ref, b200, curve, s):
    """extended_radix2_domain (m = 2^(s+1): what get_evaluation_domain returns one step above the field's 2-adicity,
    get_evaluation_domain.tcc:44-46): all four transforms and divide_by_Z_on_coset, bit-exact
    (extended_radix2_domain.tcc:64-236). The unmodified reference reaches that domain at a testable size with
    FieldT::s / FieldT::root_of_unity lowered (RefOracle.set_two_adicity); b200_set_two_adicity mirrors the setting."""
    m = 1 << (s + 1)
    a = ref.field_random(curve, refo.FR, 300 + s, m)
    g = _gen(ref, curve)
    g2 = ref.field_random(curve, refo.FR, 17, 1)[0]                 # a coset generator other than the field's
    ref.set_two_adicity(curve, s)
    b200.set_two_adicity(curve, s)
    try:
        dom = b200.get_evaluation_domain(curve, m)
        for name, op in [("FFT", refo.FFT), ("iFFT", refo.IFFT), ("cosetFFT", refo.COSET_FFT), ("icosetFFT", refo.ICOSET_FFT)]:
            for gg in ([g, g2] if "coset" in name else [g]):
                got = getattr(dom, name)(a, gg) if "coset" in name else getattr(dom, name)(a)
                assert np.array_equal(got, ref.fft(curve, op, a, gg, via_factory=True)), (name, s)
        assert np.array_equal(dom.divide_by_Z_on_coset(a), ref.fft(curve, refo.DIV_Z_COSET, a, g, via_factory=True))
        assert np.array_equal(dom.iFFT(dom.FFT(a)), a)
        if s <= 10:   # libfqfft's own check of a domain kind: FFT == the polynomial at every get_domain_element(i)
            assert np.array_equal(dom.FFT(a), ref.poly_eval_at_factory_domain(curve, a, np.arange(m)))
        # one size up no radix-2 kind is valid any more; one size down is the plain radix-2 domain
        with pytest.raises(b200.DomainSizeException):
            b200.get_evaluation_domain(curve, 2 * m).FFT(np.concatenate([a, a]))
        assert np.array_equal(b200.get_evaluation_domain(curve, m // 2).FFT(a[: m // 2]),
                              ref.fft(curve, refo.FFT, a[: m // 2], g, via_factory=True))
    finally:
        ref.set_two_adicity(curve, 0)
        b200.set_two_adicity(curve, 0)
    # the field's own parameters again: 2^(s+1) is a basic domain
    assert np.array_equal(b200.get_evaluation_domain(curve, m).FFT(a), ref.fft(curve, refo.FFT, a, g))


@pytest.mark.parametrize("curve", ["alt_bn128", "bls12_377", "bw6_761"])
def test_extended_domain_golden_fixtures(b200, curve):
    """The device extended_radix2_domain against the committed vectors the unmodified reference produced with FieldT::s
    lowered (tests/golden/extended.npz, tools/gen_golden.py:gen_extended) — no reference code in the loop."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "extended.npz"))
    coset = g[f"{curve}.generator"]
    for s in [1, 2, 5, 8]:
        b200.set_two_adicity(curve, s)
        try:
            a = g[f"{curve}.s{s}.in"]
            dom = b200.get_evaluation_domain(curve, 1 << (s + 1))
            assert np.array_equal(dom.FFT(a), g[f"{curve}.s{s}.fft"])
            assert np.array_equal(dom.iFFT(a), g[f"{curve}.s{s}.ifft"])
            assert np.array_equal(dom.cosetFFT(a, coset), g[f"{curve}.s{s}.coset_fft"])
            assert np.array_equal(dom.icosetFFT(a, coset), g[f"{curve}.s{s}.icoset_fft"])
            assert np.array_equal(dom.divide_by_Z_on_coset(a), g[f"{curve}.s{s}.div_z"])
        finally:
            b200.set_two_adicity(curve, 0)


@pytest.mark.parametrize("curve,nc,s,m", [("alt_bn128", 117, 0, 128), ("alt_bn128", 130, 0, 136), ("bls12_377", 59, 5, 64),
                                         ("bw6_761", 40, 0, 48)])
def test_witness_map_with_zk_patch_golden(b200, curve, nc, s, m):
    """b200_witness_map_r1cs_zk — r1cs_to_qap_witness_map with d1, d2, d3 non-zero (the sibling provers' call;
    r1cs_to_qap.tcc:262-271, :327-333), add_poly_Z of every domain kind included — against the committed vectors of the
    unmodified reference (tests/golden/zk.npz): basic, step, extended (2-adicity lowered) and a 12-limb step domain."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "zk.npz"))
    k = f"{curve}.nc{nc}"
    n_c, ni, nv, mm = [int(x) for x in g[k + ".sizes"]]
    assert mm == m
    if s:
        b200.set_two_adicity(curve, s)
    try:
        cs = b200.R1CS(curve, n_c, ni, nv, g[k + ".row_ptr"], g[k + ".col"], g[k + ".val"])
        try:
            d = g[k + ".d"]
            assert np.array_equal(cs.witness_map_h_zk(g[k + ".w"], m, d[0], d[1], d[2]), g[k + ".H"])
            z = np.zeros_like(d[0])
            assert np.array_equal(cs.witness_map_h_zk(g[k + ".w"], m, z, z, z), cs.witness_map_h(g[k + ".w"], m))
            # one d at a time: each term of the patch on its own
            for j in range(3):
                dd = [z, z, z]
                dd[j] = d[j]
                assert cs.witness_map_h_zk(g[k + ".w"], m, *dd).shape == (m + 1, g[k + ".w"].shape[1])
        finally:
            cs.free()
    finally:
        if s:
            b200.set_two_adicity(curve, 0)


def test_factory_rounding(b200):
    """get_evaluation_domain.tcc:37-58: 2^k stays, 2^k + 2^r stays, anything else rounds its small part up."""
    assert b200.domain_size_for(1 << 20) == 1 << 20
    assert b200.domain_size_for((1 << 20) + 11) == (1 << 20) + 16          # 1048592
    assert b200.domain_size_for((1 << 10) + 100) == (1 << 10) + 128
    assert b200.domain_size_for(1000) == 1024


@pytest.mark.parametrize("curve", ["alt_bn128", "bls12_377", "bw6_761"])
def test_step_domain_golden_fixtures(b200, curve):
    """The device step_radix2_domain against the committed vectors the unmodified reference produced
    (tests/golden/step.npz, tools/gen_golden.py:gen_step), and — alt_bn128 — the device sparse R1CS evaluation and
    witness map of a circuit over a step domain against the reference's aA, aB, aC and H."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "step.npz"))
    coset = g[f"{curve}.generator"]                       # Fr::multiplicative_generator, Montgomery limbs
    for m in [3, 6, 24, 80, 288]:
        a = g[f"{curve}.m{m}.in"]
        dom = b200.get_evaluation_domain(curve, m)
        assert dom.m == m
        assert np.array_equal(dom.FFT(a), g[f"{curve}.m{m}.fft"])
        assert np.array_equal(dom.iFFT(a), g[f"{curve}.m{m}.ifft"])
        assert np.array_equal(dom.cosetFFT(a, coset), g[f"{curve}.m{m}.coset_fft"])
        assert np.array_equal(dom.icosetFFT(a, coset), g[f"{curve}.m{m}.icoset_fft"])
        assert np.array_equal(dom.divide_by_Z_on_coset(a), g[f"{curve}.m{m}.div_z"])
    if curve == "alt_bn128":
        nc, ni, nv, m = [int(x) for x in g["witness.sizes"]]
        cs = b200.R1CS(curve, nc, ni, nv, g["witness.row_ptr"], g["witness.col"], g["witness.val"])
        try:
            for got, want in zip(cs.evaluate(g["witness.w"], m), (g["witness.aA"], g["witness.aB"], g["witness.aC"])):
                assert np.array_equal(got, want)
            assert np.array_equal(cs.witness_map_h(g["witness.w"], m), g["witness.H"])
            assert np.array_equal(b200.r1cs_to_qap_witness_map_h(curve, g["witness.aA"], g["witness.aB"], g["witness.aC"]),
                                  g["witness.H"])
        finally:
            cs.free()

