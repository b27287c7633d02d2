"""GPU parity of the key-generation operators: b200_batch_exp vs the reference's
get_window_table + batch_exp / batch_exp_with_coeff + batch_to_special (multiexp.tcc:808-974), the overlay's
specialisations of batch_exp / batch_exp_with_coeff / kc_batch_exp called through the harness, and the reference's
own r1cs_gg_ppzksnark_generator running over them (its keys must yield proofs its verifier accepts)."""
import ctypes

import numpy as np
import pytest

from oracle import refo

pytestmark = pytest.mark.gpu

CURVE_GROUPS = [("alt_bn128", refo.G1), ("alt_bn128", refo.G2), ("bls12_377", refo.G1), ("bls12_377", refo.G2),
                ("bw6_761", refo.G1), ("bw6_761", refo.G2)]


@pytest.fixture(scope="module")
def overlay(b200):
    try:
        return refo.RefOracle(overlay=True)
    except (FileNotFoundError, OSError) as e:
        pytest.skip(f"overlay build of the reference missing: {e}")


def _scalars(ref, curve, n, seed):
    sc = ref.field_random(curve, refo.FR, seed, n)
    if n > 3:
        sc[1] = 0
        sc[2] = ref.field_consts(curve, refo.FR)["one"]
    return sc


@pytest.mark.parametrize("curve,group", CURVE_GROUPS)
def test_batch_exp_matches_reference(ref, b200, curve, group):
    g = ref.make_bases(curve, group, 1, 31)[0]
    coeff = ref.field_random(curve, refo.FR, 77, 1)[0]
    for n in (1, 7, 1500 if curve != "bw6_761" else 300):
        sc = _scalars(ref, curve, n, 100 + n)
        assert np.array_equal(b200.batch_exp(curve, group, g, sc), ref.batch_exp(curve, group, g, sc)), n
        assert np.array_equal(b200.batch_exp(curve, group, g, sc, coeff), ref.batch_exp(curve, group, g, sc, coeff)), n
    # the group generator as the base (what the generator uses), and a second base right after (table rebuilt)
    one = np.zeros((1, sc.shape[1]), dtype=np.uint64)
    one[0] = ref.field_consts(curve, refo.FR)["one"]
    gen = ref.scalar_mul_gen(curve, group, one)[0]
    sc = _scalars(ref, curve, 64, 9)
    assert np.array_equal(b200.batch_exp(curve, group, gen, sc), ref.scalar_mul_gen(curve, group, sc))
    assert np.array_equal(b200.batch_exp(curve, group, g, sc), ref.batch_exp(curve, group, g, sc))


def test_batch_exp_large_and_errors(ref, b200):
    curve, group = "alt_bn128", refo.G1
    g = ref.make_bases(curve, group, 1, 5)[0]
    n = (1 << 16) + 3
    sc = _scalars(ref, curve, n, 4)
    assert np.array_equal(b200.batch_exp(curve, group, g, sc), ref.batch_exp(curve, group, g, sc))
    assert b200.batch_exp(curve, group, g, sc[:0]).shape[0] == 0
    bad = g.copy()
    bad[8] ^= 1                                             # Z != one: not in special form
    with pytest.raises(b200.B200Error):
        b200.batch_exp(curve, group, bad, sc[:4])
    inf = ref.group_op(curve, group, refo.GOP_TO_AFFINE, np.zeros((1, 12), np.uint64))[0]
    out = b200.batch_exp(curve, group, inf, sc[:4])
    assert np.array_equal(out, np.repeat(inf[None, :], 4, axis=0))


@pytest.mark.parametrize("curve", ["alt_bn128", "bls12_377", "bw6_761"])
def test_overlay_batch_exp_operators(ref, overlay, curve):
    """The same harness entry in the two builds of the reference: stock templates vs the overlay's specialisations."""
    n = 400 if curve != "bw6_761" else 120
    sc = _scalars(ref, curve, n, 12)
    coeff = ref.field_random(curve, refo.FR, 13, 2)
    for group in (refo.G1, refo.G2):
        g = ref.make_bases(curve, group, 1, 41)[0]
        assert np.array_equal(overlay.batch_exp(curve, group, g, sc), ref.batch_exp(curve, group, g, sc))
        assert np.array_equal(overlay.batch_exp(curve, group, g, sc, coeff[0]), ref.batch_exp(curve, group, g, sc, coeff[0]))
    g2, g1 = ref.make_bases(curve, refo.G2, 1, 42)[0], ref.make_bases(curve, refo.G1, 1, 43)[0]
    sc[5:40] = 0
    want_idx, want_vals = ref.kc_batch_exp(curve, g2, g1, coeff[0], coeff[1], sc, chunks=3)
    got_idx, got_vals = overlay.kc_batch_exp(curve, g2, g1, coeff[0], coeff[1], sc, chunks=3)
    assert np.array_equal(got_idx, want_idx) and np.array_equal(got_vals, want_vals)


@pytest.mark.parametrize("curve,nc", [("alt_bn128", 1000), ("bls12_377", 300)])
def test_overlay_generator_keys_verify(ref, overlay, curve, nc):
    """r1cs_gg_ppzksnark_generator compiled with the overlay (its batch_exp / kc_batch_exp calls run on the GPU):
    the proving key it produces gives proofs — from the reference prover and from the GPU prover — that the
    reference verifier accepts with the matching verification key."""
    fx = overlay.groth16_setup(curve, nc, 5, real_keys=True)

    def verify(p):
        # the STOCK build's verifier on the overlay-held key pair (the C++ objects are layout-identical). The overlay
        # build's own copy of the verifier is not used: its bos_coster multi_exp is compiled with libff's asm
        # ordered_exponent::operator<, which gcc 13 miscompiles — the stock harness includes
        # multiexp.hpp with USE_ASM off for that reason, and the -include'd overlay header gets there first.
        return ref.lib.ref_groth16_verify(fx.h, np.ascontiguousarray(p, dtype=np.uint64).ctypes.data_as(ctypes.c_void_p))
    try:
        rs = ref.field_random(curve, refo.FR, 0xD00D, 2)
        proof, _ = fx.prove(rs[0], rs[1], via=ref)              # the stock CPU prover on the GPU-generated key
        assert verify(proof) == 1
        again, _ = fx.prove(rs[0], rs[1])                       # the overlay's operators on the same key
        assert np.array_equal(again, proof)
        fused, _ = fx.prove_fused(rs[0], rs[1], via=overlay)
        assert np.array_equal(fused, proof)
        bad = proof.copy()
        bad[1] ^= 1
        assert verify(bad) != 1
    finally:
        fx.close()


@pytest.mark.parametrize("curve", ["alt_bn128", "bls12_377", "bw6_761"])
def test_batch_exp_golden_fixtures(b200, curve):
    """b200_batch_exp against the committed vectors the unmodified reference produced (tests/golden/keygen.npz)."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "keygen.npz"))
    sc, coeff = g[f"{curve}.scalars"], g[f"{curve}.coeff"]
    for gname, gid in [("g1", refo.G1), ("g2", refo.G2)]:
        base = g[f"{curve}.{gname}.base"]
        assert np.array_equal(b200.batch_exp(curve, gid, base, sc), g[f"{curve}.{gname}.out"])
        assert np.array_equal(b200.batch_exp(curve, gid, base, sc, coeff[0]), g[f"{curve}.{gname}.out_coeff"])
