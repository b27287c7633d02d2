"""ctypes binding of oracle/_ref/libref_oracle.so — TEST INFRASTRUCTURE ONLY.

The shared object is the unmodified reference (clearmatics/libsnark + libff +
libfqfft) compiled by oracle/Makefile from /root/reference; see ref_harness.cpp.
Only tests/, bench.py's cpu_baseline / `--impl reference` leg and
__graft_entry__.smoke() may import this module. The product package
(libsnark_b200/) never does.

All arrays are numpy uint64 views of the reference's own in-memory layouts:
  Fr / Fq element  : (n_limbs,) u64 little-endian Montgomery residue
  G1 (alt_bn128)   : X|Y|Z            -> 3*nq limbs
  G2 (alt_bn128)   : X.c0|X.c1|Y.c0|Y.c1|Z.c0|Z.c1 -> 6*nq limbs
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

CURVES = {"alt_bn128": 0, "bls12_377": 1, "bw6_761": 2}
FR, FQ, FQ2 = 0, 1, 2
G1, G2 = 1, 2

OP_ADD, OP_SUB, OP_MUL, OP_SQR, OP_NEG, OP_INV = range(6)
GOP_ADD, GOP_MIXED_ADD, GOP_DBL, GOP_NEG, GOP_TO_AFFINE, GOP_ADD_FN = range(6)
MSM_BDLO12_SIGNED_SPECIAL, MSM_NAIVE_PLAIN, MSM_FILTER_ONE_ZERO, MSM_BDLO12_SIGNED_NORMAL = range(4)
FFT, IFFT, COSET_FFT, ICOSET_FFT, DIV_Z_COSET = range(5)


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class RefOracle:
    def __init__(self, path=None, overlay=False):
        if path is None:
            name = "libref_overlay.so" if overlay else "libref_oracle.so"
            path = os.path.join(_HERE, "_ref", name)
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} missing: run `make -C oracle ref` where /root/reference is present")
        self.lib = C.CDLL(path, mode=C.RTLD_LOCAL)
        L = self.lib
        L.ref_groth16_setup.restype = C.c_void_p
        L.ref_groth16_setup.argtypes = [C.c_int, C.c_size_t, C.c_size_t, C.c_int]
        L.ref_groth16_ptr.restype = C.c_void_p
        L.ref_groth16_ptr.argtypes = [C.c_void_p, C.c_int]
        L.ref_groth16_free.argtypes = [C.c_void_p]
        L.ref_groth16_sizes.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_groth16_witness_map.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_groth16_witness_map_zk.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_groth16_sap_witness_map.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t),
                                                  C.c_void_p, C.POINTER(C.c_double)]
        L.ref_groth16_prove.argtypes = [C.c_void_p] * 5
        L.ref_groth16_abc_evaluations.argtypes = [C.c_void_p] * 4
        L.ref_groth16_prove_stock.argtypes = [C.c_void_p] * 3
        L.ref_groth16_prove_fused.argtypes = [C.c_void_p] * 5
        L.ref_groth16_r1cs_csr.argtypes = [C.c_void_p] * 5
        L.ref_groth16_verify.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_groth16_rekey_in_place.argtypes = [C.c_void_p, C.c_long]
        L.ref_bdlo12_signed_optimal_c.restype = C.c_uint64
        L.ref_bdlo12_signed_optimal_c.argtypes = [C.c_uint64]
        L.ref_field_sha512.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_size_t, C.c_void_p]
        L.ref_field_op.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.ref_fp_convert.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        L.ref_signed_digits.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p]
        L.ref_group_op.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.ref_make_bases.argtypes = [C.c_int, C.c_int, C.c_size_t, C.c_uint64, C.c_void_p]
        L.ref_scalar_mul_gen.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_batch_exp.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_kc_batch_exp.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t,
                                       C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_msm.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_kc_msm.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t,
                                 C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p]
        L.ref_kc_msm_g1.argtypes = L.ref_kc_msm.argtypes
        L.ref_fft.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
        L.ref_poly_eval_at_domain.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_poly_eval_at_factory_domain.argtypes = L.ref_poly_eval_at_domain.argtypes
        L.ref_set_two_adicity.argtypes = [C.c_int, C.c_int]
        L.ref_sibling_setup.argtypes = [C.c_int, C.c_int, C.c_size_t, C.c_size_t]
        L.ref_sibling_setup.restype = C.c_void_p
        L.ref_sibling_prove.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t),
                                        C.POINTER(C.c_double)]
        L.ref_sibling_verify.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        L.ref_sibling_free.argtypes = [C.c_void_p]
        L.ref_field_consts.argtypes = [C.c_int, C.c_int, C.c_void_p]
        L.ref_curve_consts.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_info.argtypes = [C.c_int, C.c_void_p]
        assert L.ref_init() == 0
        self._info = {}

    # ---- metadata -------------------------------------------------------
    def info(self, curve):
        c = CURVES[curve] if isinstance(curve, str) else curve
        if c not in self._info:
            out = np.zeros(16, dtype=np.uint64)
            assert self.lib.ref_info(c, _p(out)) == 0
            keys = ["nr", "nq", "sizeof_fr", "sizeof_fq", "sizeof_g1", "sizeof_g2", "fr_s", "fr_bits", "fq_bits",
                    "sizeof_g2_base"]
            d = {k: int(v) for k, v in zip(keys, out)}
            d["id"] = c
            d["g1_limbs"] = d["sizeof_g1"] // 8
            d["g2_limbs"] = d["sizeof_g2"] // 8
            self._info[c] = d
        return self._info[c]

    def num_threads(self):
        return self.lib.ref_num_threads()

    def set_num_threads(self, n):
        self.lib.ref_set_num_threads(int(n))

    def field_consts(self, curve, field):
        i = self.info(curve)
        n = i["nr"] if field == FR else i["nq"]
        out = np.zeros(6 * n + 1, dtype=np.uint64)
        assert self.lib.ref_field_consts(i["id"], field, _p(out)) == 0
        return dict(modulus=out[:n].copy(), R2=out[n:2 * n].copy(), R3=out[2 * n:3 * n].copy(), inv=int(out[3 * n]),
                    one=out[3 * n + 1:4 * n + 1].copy(), generator=out[4 * n + 1:5 * n + 1].copy(),
                    root_of_unity=out[5 * n + 1:6 * n + 1].copy())

    def curve_consts(self, curve):
        i = self.info(curve)
        nqr = np.zeros(i["nq"], dtype=np.uint64)
        g1 = np.zeros(i["g1_limbs"], dtype=np.uint64)
        g2 = np.zeros(i["g2_limbs"], dtype=np.uint64)
        assert self.lib.ref_curve_consts(i["id"], _p(nqr), _p(g1), _p(g2)) == 0
        return dict(fq2_non_residue=nqr, g1_one=g1, g2_one=g2)

    def _flimbs(self, curve, field):
        i = self.info(curve)
        if field == FR:
            return i["nr"]
        if field == FQ:
            return i["nq"]
        return i["sizeof_g2_base"] // 8

    def _glimbs(self, curve, group):
        i = self.info(curve)
        return i["g1_limbs"] if group == G1 else i["g2_limbs"]

    # ---- fields ------------------------------------------------------------
    def field_random(self, curve, field, idx0, n):
        """SHA512_rng<FieldT>(idx0 + i), i < n (libff/common/rng.tcc:26-69); field in {FR, FQ}."""
        L = self._flimbs(curve, field)
        out = np.zeros((n, L), dtype=np.uint64)
        assert self.lib.ref_field_sha512(self.info(curve)["id"], field, idx0, n, _p(out)) == 0
        return out

    def field_op(self, curve, field, op, a, b=None):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        b = a if b is None else np.ascontiguousarray(b, dtype=np.uint64)
        out = np.zeros_like(a)
        assert self.lib.ref_field_op(self.info(curve)["id"], field, op, _p(a), _p(b), _p(out), a.shape[0]) == 0
        return out

    def from_mont(self, curve, field, a):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        out = np.zeros_like(a)
        assert self.lib.ref_fp_convert(self.info(curve)["id"], field, 0, _p(a), _p(out), a.shape[0]) == 0
        return out

    def to_mont(self, curve, field, a):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        out = np.zeros_like(a)
        assert self.lib.ref_fp_convert(self.info(curve)["id"], field, 1, _p(a), _p(out), a.shape[0]) == 0
        return out

    def signed_digits(self, curve, scalars, c, num_digits):
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        out = np.zeros((scalars.shape[0], num_digits), dtype=np.int64)
        assert self.lib.ref_signed_digits(self.info(curve)["id"], _p(scalars), scalars.shape[0], c, num_digits,
                                          _p(out)) == 0
        return out

    def optimal_c(self, n):
        return int(self.lib.ref_bdlo12_signed_optimal_c(n))

    # ---- groups ------------------------------------------------------------
    def group_op(self, curve, group, op, a, b=None):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        b = a if b is None else np.ascontiguousarray(b, dtype=np.uint64)
        out = np.zeros_like(a)
        assert self.lib.ref_group_op(self.info(curve)["id"], group, op, _p(a), _p(b), _p(out), a.shape[0]) == 0
        return out

    def make_bases(self, curve, group, n, seed=1):
        out = np.zeros((n, self._glimbs(curve, group)), dtype=np.uint64)
        assert self.lib.ref_make_bases(self.info(curve)["id"], group, n, seed, _p(out)) == 0
        return out

    def batch_exp(self, curve, group, g, scalars, coeff=None):
        """libff get_window_table + batch_exp (coeff None) / batch_exp_with_coeff + batch_to_special (multiexp.tcc:808-974)."""
        g = np.ascontiguousarray(g, dtype=np.uint64)
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        coeff = None if coeff is None else np.ascontiguousarray(coeff, dtype=np.uint64)
        out = np.zeros((scalars.shape[0], self._glimbs(curve, group)), dtype=np.uint64)
        assert self.lib.ref_batch_exp(self.info(curve)["id"], group, _p(g), None if coeff is None else _p(coeff), _p(scalars),
                                      scalars.shape[0], _p(out)) == 0
        return out

    def kc_batch_exp(self, curve, g2, g1, c2, c1, scalars, chunks=1):
        """libsnark::kc_batch_exp<G2, G1, Fr>(..., output_special = true) -> (indices, values[nnz, G2 | G1])."""
        g2, g1, c2, c1 = [np.ascontiguousarray(x, dtype=np.uint64) for x in (g2, g1, c2, c1)]
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        n = scalars.shape[0]
        vals = np.zeros((n, self._glimbs(curve, G2) + self._glimbs(curve, G1)), dtype=np.uint64)
        idx = np.zeros(n, dtype=np.uint64)
        nnz = np.zeros(1, dtype=np.uint64)
        assert self.lib.ref_kc_batch_exp(self.info(curve)["id"], _p(g2), _p(g1), _p(c2), _p(c1), _p(scalars), n, chunks, _p(vals),
                                         _p(idx), _p(nnz)) == 0
        k = int(nnz[0])
        return idx[:k].copy(), vals[:k].copy()

    def scalar_mul_gen(self, curve, group, scalars):
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        out = np.zeros((scalars.shape[0], self._glimbs(curve, group)), dtype=np.uint64)
        assert self.lib.ref_scalar_mul_gen(self.info(curve)["id"], group, _p(scalars), scalars.shape[0], _p(out)) == 0
        return out

    def msm(self, curve, group, bases, scalars, method=MSM_BDLO12_SIGNED_SPECIAL, chunks=1):
        bases = np.ascontiguousarray(bases, dtype=np.uint64)
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        n = scalars.shape[0]
        assert bases.shape[0] == n
        out = np.zeros(self._glimbs(curve, group), dtype=np.uint64)
        rc = self.lib.ref_msm(self.info(curve)["id"], group, method, _p(bases), n, _p(scalars), chunks, _p(out))
        assert rc == 0, rc
        return out

    def kc_msm(self, curve, indices, values, domain_size, min_idx, max_idx, scalars, chunks=1, g1_pair=False):
        """kc_multi_exp_with_mixed_addition over knowledge_commitment<G2, G1> (or <G1, G1> with g1_pair)."""
        indices = np.ascontiguousarray(indices, dtype=np.uint64)
        values = np.ascontiguousarray(values, dtype=np.uint64)
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        g2 = np.zeros(self._glimbs(curve, G1 if g1_pair else G2), dtype=np.uint64)
        g1 = np.zeros(self._glimbs(curve, G1), dtype=np.uint64)
        fn = self.lib.ref_kc_msm_g1 if g1_pair else self.lib.ref_kc_msm
        rc = fn(self.info(curve)["id"], _p(indices), _p(values), indices.shape[0], domain_size,
                min_idx, max_idx, _p(scalars), scalars.shape[0], chunks, _p(g2), _p(g1))
        assert rc == 0, rc
        return g2, g1

    # ---- FFT ----------------------------------------------------------------
    def fft(self, curve, op, data, g=None, via_factory=False):
        """Returns a transformed COPY. rc -2 <=> the reference threw (DomainSizeException ...)."""
        a = np.array(data, dtype=np.uint64, order="C", copy=True)
        g_ = None if g is None else np.ascontiguousarray(g, dtype=np.uint64)
        rc = self.lib.ref_fft(self.info(curve)["id"], op, _p(a), a.shape[0], _p(g_), int(via_factory))
        if rc != 0:
            raise RuntimeError(f"ref_fft rc={rc}")
        return a

    def poly_eval_at_domain(self, curve, coeffs, idx):
        coeffs = np.ascontiguousarray(coeffs, dtype=np.uint64)
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        out = np.zeros((idx.shape[0], coeffs.shape[1]), dtype=np.uint64)
        assert self.lib.ref_poly_eval_at_domain(self.info(curve)["id"], _p(coeffs), coeffs.shape[0], _p(idx),
                                                idx.shape[0], _p(out)) == 0
        return out

    def poly_eval_at_factory_domain(self, curve, coeffs, idx):
        """evaluate_polynomial at get_evaluation_domain(m)->get_domain_element(idx[i]) (any domain kind)."""
        coeffs = np.ascontiguousarray(coeffs, dtype=np.uint64)
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        out = np.zeros((idx.shape[0], coeffs.shape[1]), dtype=np.uint64)
        assert self.lib.ref_poly_eval_at_factory_domain(self.info(curve)["id"], _p(coeffs), coeffs.shape[0], _p(idx),
                                                        idx.shape[0], _p(out)) == 0
        return out

    def set_two_adicity(self, curve, s):
        """Lowers FieldT::s / FieldT::root_of_unity of the reference consistently (0 restores them): the unmodified
        get_evaluation_domain then returns extended_radix2_domain at 2^(s+1)."""
        assert self.lib.ref_set_two_adicity(self.info(curve)["id"], int(s)) == 0

    def sibling_setup(self, scheme, curve, num_constraints, num_inputs):
        return SiblingFixture(self, scheme, curve, num_constraints, num_inputs)

    # ---- Groth16 -------------------------------------------------------------
    def groth16_setup(self, curve, num_constraints, num_inputs, real_keys=True):
        return Groth16Fixture(self, curve, num_constraints, num_inputs, real_keys)


PGHR13, SE_PPZKSNARK = 0, 1


class SiblingFixture:
    """r1cs_ppzksnark (PGHR13, scheme 0) or r1cs_se_ppzksnark (scheme 1) on the reference's synthetic R1CS: keys from
    the reference generator (BaseForm = special), the reference's own prover entry point instantiated with
    BDLO12_signed / special (proofs are randomised inside), the strong-IC verifier. `via` = the overlay build runs
    ITS copy of the same unmodified prover — with the operators resolved to the CUDA library — on this fixture."""

    def __init__(self, ref, scheme, curve, num_constraints, num_inputs):
        self.ref, self.scheme, self.cid = ref, int(scheme), ref.info(curve)["id"]
        self.h = ref.lib.ref_sibling_setup(self.scheme, self.cid, num_constraints, num_inputs)
        assert self.h, "ref_sibling_setup failed"

    def prove(self, via=None):
        buf = np.zeros(1 << 14, dtype=np.uint8)
        n, ms = C.c_size_t(0), C.c_double(0)
        lib = (via or self.ref).lib
        rc = lib.ref_sibling_prove(self.scheme, self.cid, self.h, _p(buf), buf.nbytes, C.byref(n), C.byref(ms))
        assert rc == 0, rc
        return buf[: n.value].copy(), ms.value

    def verify(self, proof, via=None):
        proof = np.ascontiguousarray(proof, dtype=np.uint8)
        return (via or self.ref).lib.ref_sibling_verify(self.scheme, self.cid, self.h, _p(proof), proof.nbytes)

    def close(self):
        if self.h:
            self.ref.lib.ref_sibling_free(self.h)
            self.h = None


class Groth16Fixture:
    """R1CS example (generate_r1cs_example_with_field_input) + keypair held inside the reference .so."""

    def __init__(self, ref, curve, num_constraints, num_inputs, real_keys):
        self.ref = ref
        self.curve = curve
        self.i = ref.info(curve)
        self.h = ref.lib.ref_groth16_setup(self.i["id"], num_constraints, num_inputs, int(real_keys))
        assert self.h
        out = np.zeros(16, dtype=np.uint64)
        ref.lib.ref_groth16_sizes(self.h, _p(out))
        keys = ["num_constraints", "num_inputs", "num_variables", "m", "A_size", "B_nnz", "B_domain", "H_size",
                "L_size"]
        self.sizes = {k: int(v) for k, v in zip(keys, out)}

    def _view(self, which, n, limbs, dtype=np.uint64):
        p = self.ref.lib.ref_groth16_ptr(self.h, which)
        if n == 0:
            return np.zeros((0, limbs), dtype=dtype)
        buf = (C.c_uint64 * (n * limbs)).from_address(p)
        return np.frombuffer(buf, dtype=dtype).reshape(n, limbs)

    @property
    def A_query(self):
        return self._view(0, self.sizes["A_size"], self.i["g1_limbs"])

    @property
    def B_values(self):
        return self._view(1, self.sizes["B_nnz"], self.i["g2_limbs"] + self.i["g1_limbs"])

    @property
    def B_indices(self):
        return self._view(2, self.sizes["B_nnz"], 1).reshape(-1)

    @property
    def H_query(self):
        return self._view(3, self.sizes["H_size"], self.i["g1_limbs"])

    @property
    def L_query(self):
        return self._view(4, self.sizes["L_size"], self.i["g1_limbs"])

    @property
    def primary_input(self):
        return self._view(5, self.sizes["num_inputs"], self.i["nr"])

    @property
    def auxiliary_input(self):
        return self._view(6, self.sizes["num_variables"] - self.sizes["num_inputs"], self.i["nr"])

    def pk_points(self):
        """alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2 normalised to the special (affine) form."""
        out = []
        for which, grp, key in [(7, G1, "g1_limbs"), (8, G1, "g1_limbs"), (9, G2, "g2_limbs"), (10, G1, "g1_limbs"),
                                (11, G2, "g2_limbs")]:
            v = self._view(which, 1, self.i[key]).copy()
            out.append(self.ref.group_op(self.curve, grp, GOP_TO_AFFINE, v)[0])
        return out

    def witness_map(self):
        out = np.zeros((self.sizes["m"] + 1, self.i["nr"]), dtype=np.uint64)
        ms = C.c_double(0)
        assert self.ref.lib.ref_groth16_witness_map(self.h, _p(out), C.byref(ms)) == 0
        return out, ms.value

    def witness_map_zk(self, d1, d2, d3, via=None):
        """coefficients_for_H of r1cs_to_qap_witness_map(cs, primary, auxiliary, d1, d2, d3) (m + 1 rows) and its wall
        time; via = the overlay build runs ITS r1cs_to_qap_witness_map (specialised to the device) on this fixture."""
        d = np.ascontiguousarray(np.stack([np.asarray(x, dtype=np.uint64).reshape(-1) for x in (d1, d2, d3)]))
        out = np.zeros((self.sizes["m"] + 1, self.i["nr"]), dtype=np.uint64)
        ms = C.c_double(0)
        assert (via or self.ref).lib.ref_groth16_witness_map_zk(self.h, _p(d), _p(out), C.byref(ms)) == 0
        return out, ms.value

    def sap_witness_map(self, d1, d2, via=None):
        """(coefficients_for_H, extra variables, ms) of r1cs_to_sap_witness_map(cs, primary, auxiliary, d1, d2) on this
        fixture's constraint system; via = the overlay build's specialisation."""
        d = np.ascontiguousarray(np.stack([np.asarray(x, dtype=np.uint64).reshape(-1) for x in (d1, d2)]))
        sz = self.sizes
        cap = 4 * (sz["num_constraints"] + sz["num_inputs"]) + 8
        out = np.zeros((cap, self.i["nr"]), dtype=np.uint64)
        extra = np.zeros((sz["num_constraints"] + sz["num_inputs"], self.i["nr"]), dtype=np.uint64)
        deg, ms = C.c_size_t(0), C.c_double(0)
        rc = (via or self.ref).lib.ref_groth16_sap_witness_map(self.h, _p(d), _p(out), cap, C.byref(deg), _p(extra), C.byref(ms))
        assert rc == 0, rc
        return out[: deg.value + 1].copy(), extra, ms.value

    def abc_evaluations(self):
        """aA, aB, aC of r1cs_to_qap.tcc:236-248/:293-296 (the inputs of the NTT chain)."""
        shape = (self.sizes["m"], self.i["nr"])
        a, b, c = (np.zeros(shape, dtype=np.uint64) for _ in range(3))
        assert self.ref.lib.ref_groth16_abc_evaluations(self.h, _p(a), _p(b), _p(c)) == 0
        return a, b, c

    def proof_limbs(self):
        return 2 * self.i["g1_limbs"] + self.i["g2_limbs"]

    def prove(self, r, s, via=None):
        """Proof (g_A | g_B | g_C, normalised) with caller-supplied r, s. via = another RefOracle (the
        overlay build) runs ITS prover on this fixture: the C++ objects are layout-identical."""
        r = np.ascontiguousarray(r, dtype=np.uint64)
        s = np.ascontiguousarray(s, dtype=np.uint64)
        out = np.zeros(self.proof_limbs(), dtype=np.uint64)
        ms = np.zeros(6, dtype=np.float64)
        lib = (via or self.ref).lib
        assert lib.ref_groth16_prove(self.h, _p(r), _p(s), _p(out), _p(ms)) == 0
        stages = dict(zip(["witness_map", "A", "B", "H", "L", "total"], ms.tolist()))
        return out, stages

    def prove_fused(self, r, s, via):
        """The overlay's device-resident prover (libsnark_b200::r1cs_gg_ppzksnark_prover_fused) on this fixture."""
        r = np.ascontiguousarray(r, dtype=np.uint64)
        s = np.ascontiguousarray(s, dtype=np.uint64)
        out = np.zeros(self.proof_limbs(), dtype=np.uint64)
        ms = C.c_double(0)
        rc = via.lib.ref_groth16_prove_fused(self.h, _p(r), _p(s), _p(out), C.byref(ms))
        assert rc == 0, rc
        return out, ms.value

    def r1cs_csr(self):
        """(row_ptr, col, val): the constraint system as one CSR over Fr, rows = a | b | c combinations."""
        nnz = C.c_uint64(0)
        assert self.ref.lib.ref_groth16_r1cs_csr(self.h, None, None, None, C.byref(nnz)) == 0
        row_ptr = np.zeros(3 * self.sizes["num_constraints"] + 1, dtype=np.uint32)
        col = np.zeros(max(1, nnz.value), dtype=np.uint32)
        val = np.zeros((max(1, nnz.value), self.i["nr"]), dtype=np.uint64)
        assert self.ref.lib.ref_groth16_r1cs_csr(self.h, _p(row_ptr), _p(col), _p(val), C.byref(nnz)) == 0
        return row_ptr, col[:nnz.value], val[:nnz.value]

    def full_assignment(self):
        """w = (1 | primary | auxiliary), Montgomery Fr."""
        one = self.ref.field_consts(self.curve, FR)["one"]
        return np.concatenate([np.asarray(one, dtype=np.uint64)[None, :], self.primary_input, self.auxiliary_input])

    def prove_stock(self, via=None):
        out = np.zeros(self.proof_limbs(), dtype=np.uint64)
        ms = C.c_double(0)
        assert (via or self.ref).lib.ref_groth16_prove_stock(self.h, _p(out), C.byref(ms)) == 0
        return out, ms.value

    def verify(self, proof):
        proof = np.ascontiguousarray(proof, dtype=np.uint64)
        return self.ref.lib.ref_groth16_verify(self.h, _p(proof))

    def rekey_in_place(self, mutate_row=-1):
        """A different key of identical shape written over this fixture's key at the same host addresses (test hook
        for the overlay's caches); mutate_row >= 0 also changes that constraint's coefficients in place."""
        rc = self.ref.lib.ref_groth16_rekey_in_place(self.h, int(mutate_row))
        assert rc == 0, rc

    def close(self):
        if self.h:
            self.ref.lib.ref_groth16_free(self.h)
            self.h = None
