"""CPU tier: pins the oracle. The C restatement (oracle/oracle_port.c) is checked against the golden
vectors in tests/golden/, which tools/gen_golden.py produced from the UNMODIFIED reference
(oracle/_ref); where that reference build is present (this container) it is re-checked against the same
vectors, so the fixtures cannot drift. KATs: libff test_multiexp.cpp:204-255, :337-383,
test_fields.cpp:282-395; libfqfft evaluation_domain_test.cpp:40-80."""
import os

import numpy as np
import pytest

from oracle import port as port_mod

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CURVES = ["alt_bn128", "bls12_377", "bw6_761"]
FIELDS = {"fr": 0, "fq": 1, "fq2": 2}
OPS = {"add": 0, "sub": 1, "mul": 2, "sqr": 3, "neg": 4, "inv": 5}


@pytest.fixture(scope="module")
def P():
    return port_mod.OraclePort()


def impls(P):
    """(name, object) of every oracle implementation available: the C port always, the reference if built."""
    out = [("port", P)]
    try:
        from oracle import refo
        out.append(("ref", refo.RefOracle()))
    except (OSError, FileNotFoundError):
        pass
    return out


@pytest.mark.parametrize("curve", CURVES)
def test_field_ops_golden(P, curve):
    g = np.load(os.path.join(GOLD, "fields.npz"))
    for name, impl in impls(P):
        for fname, fid in FIELDS.items():
            k = f"{curve}.{fname}"
            if k + ".a" not in g:
                continue
            a, b = g[k + ".a"], g[k + ".b"]
            for oname, op in OPS.items():
                assert np.array_equal(impl.field_op(curve, fid, op, a, b), g[f"{k}.{oname}"]), (name, k, oname)


@pytest.mark.parametrize("curve", CURVES)
def test_signed_digits_golden(P, curve):
    g = np.load(os.path.join(GOLD, "digits.npz"))
    sc = g[f"{curve}.scalars"]
    for name, impl in impls(P):
        for c in [2, 5, 12, 13, 16, 21]:
            want = g[f"{curve}.c{c}"]
            assert np.array_equal(impl.signed_digits(curve, sc, c, want.shape[1]), want), (name, c)
        for n, c in g["optimal_c"]:
            assert impl.optimal_c(int(n)) == int(c)


def test_signed_digit_properties(P):
    """test_fields.cpp:348-395: digits lie in [-2^(c-1), 2^(c-1)) and recompose to the value."""
    g = np.load(os.path.join(GOLD, "digits.npz"))
    r = 0x30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000001
    for c in [2, 5, 12, 13, 16, 21]:
        d = g[f"alt_bn128.c{c}"]
        assert d.min() >= -(1 << (c - 1)) and d.max() < (1 << (c - 1))
        v = sum(int(x) << (c * i) for i, x in enumerate(d[0]))
        assert v == r - 1  # Fr(-1)


def test_field_get_digit_kat():
    """test_fields.cpp:282-344: 16-bit digits of Fr(-1) of alt_bn128 (unsigned digit = signed digit + carry handling off)."""
    v = 0x30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000000
    want = [0x0000, 0xf000, 0xf593, 0x43e1, 0x7091, 0x79b9, 0xe848, 0x2833, 0x585d, 0x8181, 0x45b6, 0xb850, 0xa029,
            0xe131, 0x4e72, 0x3064]
    assert [(v >> (16 * i)) & 0xffff for i in range(16)] == want


@pytest.mark.parametrize("curve", CURVES)
def test_group_law_golden(P, curve):
    g = np.load(os.path.join(GOLD, "groups.npz"))
    for name, impl in impls(P):
        for gname, gid in [("g1", 1), ("g2", 2)]:
            k = f"{curve}.{gname}"
            a, b = g[k + ".a"], g[k + ".b"]
            norm = (lambda x: x) if name == "port" else (lambda x: impl.group_op(curve, gid, 4, x))
            assert np.array_equal(norm(impl.group_op(curve, gid, 0, a, b)), g[k + ".add"]), (name, k)
            assert np.array_equal(norm(impl.group_op(curve, gid, 2, a)), g[k + ".dbl"]), (name, k)
            assert np.array_equal(norm(impl.group_op(curve, gid, 3, a)), g[k + ".neg"]), (name, k)


def test_msm_kats(P):
    g = np.load(os.path.join(GOLD, "msm.npz"))
    for name, impl in impls(P):
        for k in [1, 2, 3, 4, 5, 256, 257]:
            got = impl.msm("alt_bn128", 1, g["family.bases"][:k], g[f"family.scalars{k}"])
            assert np.array_equal(got, g[f"family.out{k}"]), (name, k)
        assert np.array_equal(impl.msm("alt_bn128", 1, g["edge.bases"], g["edge.scalars"]), g["edge.out"]), name


@pytest.mark.parametrize("curve", CURVES)
def test_msm_random_golden(P, curve):
    g = np.load(os.path.join(GOLD, "msm.npz"))
    for name, impl in impls(P):
        for gname, gid in [("g1", 1), ("g2", 2)]:
            k = f"rand.{curve}.{gname}"
            for method in (0, 2):
                got = impl.msm(curve, gid, g[k + ".bases"], g[k + ".scalars"], method=method)
                assert np.array_equal(got, g[k + ".out"]), (name, k, method)
        k = f"kc.{curve}"
        for lo, hi in [(0, 64), (5, 50)]:
            sc = g[k + ".scalars"][: hi - lo]
            if name == "port":
                o2, o1 = impl.kc_msm(curve, g[k + ".indices"], g[k + ".values"], lo, hi, sc)
            else:
                o2, o1 = impl.kc_msm(curve, g[k + ".indices"], g[k + ".values"], 64, lo, hi, sc)
            assert np.array_equal(o2, g[f"{k}.out_g2.{lo}_{hi}"]) and np.array_equal(o1, g[f"{k}.out_g1.{lo}_{hi}"])


@pytest.mark.parametrize("curve", CURVES)
def test_fft_golden(P, curve):
    g = np.load(os.path.join(GOLD, "fft.npz"))
    for name, impl in impls(P):
        for m in [2, 4, 8, 32, 256]:
            k = f"{curve}.m{m}"
            a = g[k + ".in"]
            for op, oname in enumerate(["fft", "ifft", "cosetfft", "icosetfft", "divz"]):
                assert np.array_equal(impl.fft(curve, op, a), g[f"{k}.{oname}"]), (name, k, oname)
    assert np.array_equal(P.fft("alt_bn128", 0, g["f2538.in"]), g["f2538.fft"])


def test_fft_roundtrip_and_domain_error(P):
    g = np.load(os.path.join(GOLD, "fft.npz"))
    a = g["alt_bn128.m256.in"]
    assert np.array_equal(P.fft("alt_bn128", 1, P.fft("alt_bn128", 0, a)), a)
    assert np.array_equal(P.fft("alt_bn128", 3, P.fft("alt_bn128", 2, a)), a)
    with pytest.raises(RuntimeError):
        P.fft("alt_bn128", 0, a[:7])  # neither 2^k nor 2^k + 2^r: DomainSizeException analogue
    assert P.fft("alt_bn128", 0, a[:6]).shape[0] == 6  # 4 + 2 is a step_radix2_domain


@pytest.mark.parametrize("curve", CURVES)
def test_witness_map_golden(P, curve):
    g = np.load(os.path.join(GOLD, "fft.npz"))
    k = f"wm.{curve}"
    assert np.array_equal(P.witness_map_h(curve, g[k + ".aA"], g[k + ".aB"], g[k + ".aC"]), g[k + ".H"])
    # degree of H is m - 2: H[m-1] == H[m] == 0, H[m-2] != 0 (r1cs_gg_ppzksnark.tcc:488-490)
    H = g[k + ".H"]
    assert not H[-1].any() and not H[-2].any() and H[-3].any()


@pytest.mark.parametrize("curve", CURVES)
def test_batch_exp_golden(P, curve):
    """Key generation: get_window_table + batch_exp / batch_exp_with_coeff + batch_to_special
    (multiexp.tcc:808-974). The window width is a tuning parameter of the reference and must not change the result."""
    g = np.load(os.path.join(GOLD, "keygen.npz"))
    sc, coeff = g[f"{curve}.scalars"], g[f"{curve}.coeff"]
    for gname, gid in [("g1", 1), ("g2", 2)]:
        base = g[f"{curve}.{gname}.base"]
        for window in (1, 4, 9):
            assert np.array_equal(P.batch_exp(curve, gid, base, sc, window=window), g[f"{curve}.{gname}.out"]), (gname, window)
        assert np.array_equal(P.batch_exp(curve, gid, base, sc, coeff[0], window=6), g[f"{curve}.{gname}.out_coeff"]), gname
        assert not g[f"{curve}.{gname}.out"][3][2 * (base.shape[0] // 3):].any()          # scalar 0 -> zero(): Z = 0
        assert np.array_equal(g[f"{curve}.{gname}.out"][4], base)                          # scalar 1 -> the base itself
    # kc_batch_exp (kc_multiexp.tcc:147-239): zeros skipped, (coeff0 v) g2 | (coeff1 v) g1 per remaining entry
    idx, vals = g[f"{curve}.kc.indices"], g[f"{curve}.kc.values"]
    assert idx.tolist() == [i for i in range(sc.shape[0]) if i != 3]
    g2l = g[f"{curve}.g2.base"].shape[0]
    nz = sc[idx.astype(np.int64)]
    assert np.array_equal(vals[:, :g2l], P.batch_exp(curve, 2, g[f"{curve}.g2.base"], nz, coeff[0]))
    assert np.array_equal(vals[:, g2l:], P.batch_exp(curve, 1, g[f"{curve}.g1.base"], nz, coeff[1]))
    try:
        from oracle import refo
        ref = refo.RefOracle()
    except (OSError, FileNotFoundError):
        return
    for gname, gid in [("g1", 1), ("g2", 2)]:
        assert np.array_equal(ref.batch_exp(curve, gid, g[f"{curve}.{gname}.base"], sc, coeff[0]), g[f"{curve}.{gname}.out_coeff"])


@pytest.mark.parametrize("curve", CURVES)
def test_step_radix2_domain_golden(P, curve):
    """step_radix2_domain (sizes 2^k + 2^r; step_radix2_domain.tcc:65-279) restated in the C oracle == the vectors the
    reference produced through get_evaluation_domain; plus the whole witness map over a step domain."""
    g = np.load(os.path.join(GOLD, "step.npz"))
    ops = {"fft": 0, "ifft": 1, "coset_fft": 2, "icoset_fft": 3, "div_z": 4}
    for m in [3, 6, 24, 80, 288]:
        a = g[f"{curve}.m{m}.in"]
        for name, op in ops.items():
            assert np.array_equal(P.fft(curve, op, a, g[f"{curve}.generator"]), g[f"{curve}.m{m}.{name}"]), (m, name)
        # round trips are the identity
        assert np.array_equal(P.fft(curve, 1, P.fft(curve, 0, a)), a)
        assert np.array_equal(P.fft(curve, 3, P.fft(curve, 2, a)), a)
    if curve == "alt_bn128":
        got = P.witness_map_h(curve, g["witness.aA"], g["witness.aB"], g["witness.aC"])
        assert got.shape[0] == 1152 + 1 and np.array_equal(got, g["witness.H"])


EXTENDED_CASES = [("alt_bn128", 6, 120), ("bls12_377", 5, 59), ("bw6_761", 4, 27)]   # curve, two-adicity, constraints


@pytest.mark.parametrize("curve", CURVES)
def test_extended_radix2_domain_golden(P, curve):
    """extended_radix2_domain (m = 2^(s+1); extended_radix2_domain.tcc:21-236) restated in the C oracle == the vectors the
    unmodified reference produced through get_evaluation_domain with FieldT::s lowered (tools/gen_golden.py:gen_extended,
    where the FFT was also checked against evaluate_polynomial at every domain element)."""
    g = np.load(os.path.join(GOLD, "extended.npz"))
    ops = {"fft": 0, "ifft": 1, "coset_fft": 2, "icoset_fft": 3, "div_z": 4}
    for s in [1, 2, 5, 8]:
        P.set_two_adicity(curve, s)
        try:
            a = g[f"{curve}.s{s}.in"]
            assert a.shape[0] == 1 << (s + 1)
            for name, op in ops.items():
                assert np.array_equal(P.fft(curve, op, a, g[f"{curve}.generator"]), g[f"{curve}.s{s}.{name}"]), (s, name)
            assert np.array_equal(P.fft(curve, 1, P.fft(curve, 0, a)), a)
            assert np.array_equal(P.fft(curve, 3, P.fft(curve, 2, a)), a)
            with pytest.raises(RuntimeError):   # 2^(s+2): no radix-2 kind accepts it (DomainSizeException)
                P.fft(curve, 0, np.concatenate([a, a]))
        finally:
            P.set_two_adicity(curve, 0)


@pytest.mark.parametrize("curve,s,nc", EXTENDED_CASES)
def test_whole_proof_on_extended_domain_golden(P, curve, s, nc):
    """The restated prover over an extended domain == the reference prover's proof (reference generator, prover and
    verifier run with FieldT::s = s; the proof verified there)."""
    key, r1cs, w, r, s_, proof = groth16_fixture(curve, nc, "extended.npz")
    assert r1cs["m"] == 1 << (s + 1)
    P.set_two_adicity(curve, s)
    try:
        assert np.array_equal(P.groth16_prove(curve, key, r1cs, w, r, s_), proof)
    finally:
        P.set_two_adicity(curve, 0)


ZK_CASES = [("alt_bn128", 117, 0, 128), ("alt_bn128", 130, 0, 136), ("bls12_377", 59, 5, 64), ("bw6_761", 40, 0, 48)]


@pytest.mark.parametrize("curve,nc,s,m", ZK_CASES)
def test_witness_map_with_zk_patch_golden(P, curve, nc, s, m):
    """r1cs_to_qap_witness_map with d1, d2, d3 non-zero (the sibling provers' call; r1cs_to_qap.tcc:262-271, :327-333) restated
    in the C oracle, add_poly_Z of each domain kind included == every coefficient the reference computed (basic, step,
    extended with FieldT::s lowered, step on the 12-limb field); the sparse evaluation feeding it is pinned as well."""
    g = np.load(os.path.join(GOLD, "zk.npz"))
    k = f"{curve}.nc{nc}"
    n_c, ni, nv, mm = [int(x) for x in g[k + ".sizes"]]
    assert mm == m
    if s:
        P.set_two_adicity(curve, s)
    try:
        aA, aB, aC = P.r1cs_evaluate(curve, n_c, ni, g[k + ".row_ptr"], g[k + ".col"], g[k + ".val"], g[k + ".w"], m)
        assert np.array_equal(aA, g[k + ".aA"]) and np.array_equal(aB, g[k + ".aB"]) and np.array_equal(aC, g[k + ".aC"])
        d = g[k + ".d"]
        assert np.array_equal(P.witness_map_h_zk(curve, aA, aB, aC, d[0], d[1], d[2]), g[k + ".H"])
        z = np.zeros_like(d[0])
        assert np.array_equal(P.witness_map_h_zk(curve, aA, aB, aC, z, z, z), P.witness_map_h(curve, aA, aB, aC))
    finally:
        if s:
            P.set_two_adicity(curve, 0)


@pytest.mark.parametrize("curve,nc,s,m", [c for c in ZK_CASES if c[2] == 0])
def test_sap_witness_map_golden(P, curve, nc, s, m):
    """r1cs_to_sap_witness_map (r1cs_to_sap.tcc:297-481; the witness map of r1cs_se_ppzksnark, which the overlay specialises)
    restated in the C oracle == every coefficient of H and every extra variable the reference computed, d1 and d2 random."""
    g = np.load(os.path.join(GOLD, "zk.npz"))
    k = f"{curve}.nc{nc}"
    n_c, ni, nv, _ = [int(x) for x in g[k + ".sizes"]]
    msap = g[k + ".sap_H"].shape[0] - 1
    assert msap >= 2 * n_c + 2 * ni + 1
    ea, eb, ec = P.r1cs_evaluate(curve, n_c, ni, g[k + ".row_ptr"], g[k + ".col"], g[k + ".val"], g[k + ".w"], n_c + ni + 1)
    d = g[k + ".d"]
    H, extra = P.sap_witness_map(curve, n_c, ni, nv, ea[:n_c], eb[:n_c], ec[:n_c], g[k + ".w"], msap, d[0], d[1])
    assert np.array_equal(extra, g[k + ".sap_extra"])
    assert np.array_equal(H, g[k + ".sap_H"])


def test_r1cs_evaluation_golden(P):
    """The sparse evaluation <a_i, w>, <b_i, w>, <c_i, w> with the input-consistency rows (r1cs_to_qap.tcc:236-248,
    :293-296; linear_combination::evaluate, variable.tcc:288-299) restated over the CSR the device path uploads ==
    the aA, aB, aC the reference built for the same circuit; and from there every coefficient of H."""
    g = np.load(os.path.join(GOLD, "step.npz"))
    nc, ni, nv, m = [int(x) for x in g["witness.sizes"]]
    assert g["witness.w"].shape[0] == nv + 1 and g["witness.row_ptr"].shape[0] == 3 * nc + 1
    aA, aB, aC = P.r1cs_evaluate("alt_bn128", nc, ni, g["witness.row_ptr"], g["witness.col"], g["witness.val"], g["witness.w"], m)
    assert np.array_equal(aA, g["witness.aA"]) and np.array_equal(aB, g["witness.aB"]) and np.array_equal(aC, g["witness.aC"])
    assert np.array_equal(P.witness_map_h("alt_bn128", aA, aB, aC), g["witness.H"])


GROTH16_CASES = [("alt_bn128", 117), ("alt_bn128", 90), ("bls12_377", 59), ("bw6_761", 27)]


def groth16_fixture(curve, nc, file="groth16.npz"):
    """(key, r1cs, w, r, s, proof) of one case of tests/golden/groth16.npz (keys from the reference's generator, proof
    from the reference's prover with r, s fixed; tools/gen_golden.py:gen_groth16) or of extended.npz (gen_extended)."""
    g = np.load(os.path.join(GOLD, file))
    k = f"{curve}.nc{nc}"
    n_c, ni, nv, m = [int(x) for x in g[k + ".sizes"]]
    key = {name: g[f"{k}.{name}"] for name in ["alpha_g1", "beta_g1", "beta_g2", "delta_g1", "delta_g2", "A", "B_values",
                                                "B_indices", "H", "L"]}
    r1cs = {"nc": n_c, "ni": ni, "nv": nv, "m": m, "row_ptr": g[k + ".row_ptr"], "col": g[k + ".col"], "val": g[k + ".val"]}
    return key, r1cs, g[k + ".w"], g[k + ".r"], g[k + ".s"], g[k + ".proof"]


@pytest.mark.parametrize("curve,nc", GROTH16_CASES)
def test_whole_proof_golden(P, curve, nc):
    """r1cs_gg_ppzksnark_prover restated end to end (sparse evaluation, witness map — two of the cases sit on a step
    domain —, the four multi-exponentiations, the assembly) == the reference prover's proof for the same r, s."""
    key, r1cs, w, r, s, proof = groth16_fixture(curve, nc)
    assert np.array_equal(P.groth16_prove(curve, key, r1cs, w, r, s), proof)


@pytest.mark.parametrize("scheme", [0, 1])
def test_sibling_provers_reference_roundtrip(ref, scheme):
    """The harness around the sibling provers (r1cs_ppzksnark.tcc:570-628, r1cs_se_ppzksnark.tcc:500-554)
    instantiated with BDLO12_signed / special: generator -> prover -> strong-IC verifier, all unmodified reference code;
    a proof made for other keys is rejected. (The GPU tier runs the same prover from the overlay build.)"""
    from oracle import refo
    fx = ref.sibling_setup(scheme, "alt_bn128", 60, 3)
    other = ref.sibling_setup(scheme, "alt_bn128", 60, 3)
    try:
        proof, _ = fx.prove()
        assert fx.verify(proof) == 1
        assert other.verify(proof) == 0
        again, _ = fx.prove()
        assert not np.array_equal(again, proof) and fx.verify(again) == 1   # d1, d2, d3 / r are drawn inside the prover
    finally:
        fx.close()
        other.close()
