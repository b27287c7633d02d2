"""ctypes binding of oracle/liboracle_port.so (oracle_port.c: the plain-C restatement of the
reference's algorithms) — TEST INFRASTRUCTURE ONLY. Same array conventions as oracle/refo.py."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CURVES = {"alt_bn128": 0, "bls12_377": 1, "bw6_761": 2}
NR = {0: 4, 1: 4, 2: 6}
NQ = {0: 4, 1: 6, 2: 12}
G2_FQ2 = {0: True, 1: True, 2: False}


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _cid(curve):
    return CURVES[curve] if isinstance(curve, str) else int(curve)


class OraclePort:
    def __init__(self, path=None):
        path = path or os.path.join(_HERE, "liboracle_port.so")
        if not os.path.exists(path):
            subprocess.run(["make", "-C", _HERE, "port"], check=True, stdout=subprocess.DEVNULL)
        self.lib = L = C.CDLL(path)
        vp, sz, i = C.c_void_p, C.c_size_t, C.c_int
        L.port_msm.argtypes = [i, i, i, vp, sz, sz, vp, vp]
        L.port_kc_msm.argtypes = [i, vp, vp, sz, sz, sz, vp, vp, vp]
        L.port_signed_digits.argtypes = [i, vp, sz, sz, sz, vp]
        L.port_bdlo12_signed_optimal_c.restype = C.c_uint64
        L.port_bdlo12_signed_optimal_c.argtypes = [C.c_uint64]
        L.port_fft.argtypes = [i, i, vp, sz, vp]
        L.port_witness_map_h.argtypes = [i, vp, vp, vp, sz, vp]
        L.port_witness_map_h_zk.argtypes = [i, vp, vp, vp, sz, vp, vp]
        L.port_sap_witness_map.argtypes = [i, sz, sz, sz, vp, vp, vp, vp, sz, vp, vp, vp]
        L.port_field_op.argtypes = [i, i, i, vp, vp, vp, sz]
        L.port_group_op.argtypes = [i, i, i, vp, vp, vp, sz]
        L.port_batch_exp.argtypes = [i, i, vp, vp, vp, sz, sz, vp]
        L.port_r1cs_evaluate.argtypes = [i, sz, sz, vp, vp, vp, vp, sz, vp, vp, vp]

    def group_limbs(self, curve, group):
        c = _cid(curve)
        return 3 * NQ[c] * (2 if (group == 2 and G2_FQ2[c]) else 1)

    def msm(self, curve, group, bases, scalars, method=0, stride=None):
        c = _cid(curve)
        bases = np.ascontiguousarray(bases, dtype=np.uint64)
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        n = scalars.shape[0]
        out = np.zeros(self.group_limbs(c, group), dtype=np.uint64)
        stride = stride if stride is not None else (bases.strides[0] if bases.ndim == 2 else 0)
        assert self.lib.port_msm(c, group, method, _p(bases), stride, n, _p(scalars), _p(out)) == 0
        return out

    def r1cs_evaluate(self, curve, nc, ni, row_ptr, col, val, w, m):
        """(aA, aB, aC) as r1cs_to_qap_witness_map builds them (r1cs_to_qap.tcc:236-248, :293-296) from a CSR."""
        row_ptr = np.ascontiguousarray(row_ptr, dtype=np.uint32)
        col = np.ascontiguousarray(col, dtype=np.uint32)
        val = np.ascontiguousarray(val, dtype=np.uint64)
        w = np.ascontiguousarray(w, dtype=np.uint64)
        out = [np.zeros((m, w.shape[1]), dtype=np.uint64) for _ in range(3)]
        assert self.lib.port_r1cs_evaluate(_cid(curve), nc, ni, _p(row_ptr), _p(col), _p(val), _p(w), m, _p(out[0]), _p(out[1]),
                                           _p(out[2])) == 0
        return out

    def groth16_prove(self, curve, key, r1cs, w, r, s):
        """r1cs_gg_ppzksnark_prover (r1cs_gg_ppzksnark.tcc:460-603) with caller-supplied r, s, composed from the
        restated pieces: sparse evaluation + witness map (r1cs_to_qap.tcc:209-344, d1 = d2 = d3 = 0), the four
        multi-exponentiations (:531-580; A and L through multi_exp_filter_one_zero, B through
        kc_multi_exp_with_mixed_addition, H through multi_exp) and the assembly (:583-595).
        key: dict alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2, A, B_indices, B_values (G2 | G1), H, L.
        r1cs: dict nc, ni, nv, m, row_ptr, col, val. Returns g_A | g_B | g_C normalised."""
        nc, ni, nv, m = r1cs["nc"], r1cs["ni"], r1cs["nv"], r1cs["m"]
        w = np.ascontiguousarray(w, dtype=np.uint64)
        aA, aB, aC = self.r1cs_evaluate(curve, nc, ni, r1cs["row_ptr"], r1cs["col"], r1cs["val"], w, m)
        Hc = self.witness_map_h(curve, aA, aB, aC)
        At = self.msm(curve, 1, key["A"], w, method=2)
        Bg2, Bg1 = self.kc_msm(curve, key["B_indices"], key["B_values"], 0, nv + 1, w)
        Ht = self.msm(curve, 1, key["H"][:m - 1], Hc[:m - 1])
        Lt = self.msm(curve, 1, key["L"], w[ni + 1:], method=2)

        def add(grp, x, y):
            return self.group_op(curve, grp, 0, np.asarray(x).reshape(1, -1), np.asarray(y).reshape(1, -1))[0]

        def mul(grp, k, pt):
            return self.msm(curve, grp, np.asarray(pt).reshape(1, -1), np.asarray(k).reshape(1, -1))
        r, s = np.asarray(r, dtype=np.uint64), np.asarray(s, dtype=np.uint64)
        gA = add(1, add(1, key["alpha_g1"], At), mul(1, r, key["delta_g1"]))
        gB1 = add(1, add(1, key["beta_g1"], Bg1), mul(1, s, key["delta_g1"]))
        gB2 = add(2, add(2, key["beta_g2"], Bg2), mul(2, s, key["delta_g2"]))
        rs = self.field_op(curve, 0, 2, r.reshape(1, -1), s.reshape(1, -1))[0]
        neg = self.group_op(curve, 1, 3, mul(1, rs, key["delta_g1"]).reshape(1, -1))[0]
        gC = add(1, add(1, add(1, add(1, Ht, Lt), mul(1, s, gA)), mul(1, r, gB1)), neg)
        return np.concatenate([gA, gB2, gC])

    def batch_exp(self, curve, group, g, scalars, coeff=None, window=5):
        """get_window_table + batch_exp / batch_exp_with_coeff + batch_to_special (multiexp.tcc:808-974)."""
        c = _cid(curve)
        g = np.ascontiguousarray(g, dtype=np.uint64)
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        coeff = None if coeff is None else np.ascontiguousarray(coeff, dtype=np.uint64)
        out = np.zeros((scalars.shape[0], self.group_limbs(c, group)), dtype=np.uint64)
        assert self.lib.port_batch_exp(c, group, _p(g), None if coeff is None else _p(coeff), _p(scalars), scalars.shape[0],
                                       window, _p(out)) == 0
        return out

    def kc_msm(self, curve, indices, values, min_idx, max_idx, scalars):
        c = _cid(curve)
        indices = np.ascontiguousarray(indices, dtype=np.uint64)
        values = np.ascontiguousarray(values, dtype=np.uint64)
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        g2 = np.zeros(self.group_limbs(c, 2), dtype=np.uint64)
        g1 = np.zeros(self.group_limbs(c, 1), dtype=np.uint64)
        assert self.lib.port_kc_msm(c, _p(indices), _p(values), indices.shape[0], min_idx, max_idx, _p(scalars),
                                    _p(g2), _p(g1)) == 0
        return g2, g1

    def signed_digits(self, curve, scalars, c, num_digits):
        scalars = np.ascontiguousarray(scalars, dtype=np.uint64)
        out = np.zeros((scalars.shape[0], num_digits), dtype=np.int64)
        assert self.lib.port_signed_digits(_cid(curve), _p(scalars), scalars.shape[0], c, num_digits, _p(out)) == 0
        return out

    def optimal_c(self, n):
        return int(self.lib.port_bdlo12_signed_optimal_c(n))

    def fft(self, curve, op, data, g=None):
        a = np.array(data, dtype=np.uint64, order="C", copy=True)
        g_ = None if g is None else np.ascontiguousarray(g, dtype=np.uint64)
        rc = self.lib.port_fft(_cid(curve), op, _p(a), a.shape[0], _p(g_))
        if rc:
            raise RuntimeError(f"port_fft rc={rc}")
        return a

    def set_two_adicity(self, curve, s):
        """FieldT::s as the domain factory sees it (0 = the field's own): 2^(s+1) becomes the extended domain."""
        assert self.lib.port_set_two_adicity(_cid(curve), int(s)) == 0

    def witness_map_h(self, curve, aA, aB, aC):
        aA, aB, aC = (np.ascontiguousarray(x, dtype=np.uint64) for x in (aA, aB, aC))
        m = aA.shape[0]
        out = np.zeros((m + 1, aA.shape[1]), dtype=np.uint64)
        assert self.lib.port_witness_map_h(_cid(curve), _p(aA), _p(aB), _p(aC), m, _p(out)) == 0
        return out

    def witness_map_h_zk(self, curve, aA, aB, aC, d1, d2, d3):
        """coefficients_for_H with the ZK patch (r1cs_to_qap.tcc:262-271, :327-333)."""
        aA, aB, aC = (np.ascontiguousarray(x, dtype=np.uint64) for x in (aA, aB, aC))
        m = aA.shape[0]
        d = np.ascontiguousarray(np.stack([np.asarray(x, dtype=np.uint64).reshape(-1) for x in (d1, d2, d3)]))
        out = np.zeros((m + 1, aA.shape[1]), dtype=np.uint64)
        assert self.lib.port_witness_map_h_zk(_cid(curve), _p(aA), _p(aB), _p(aC), m, _p(d), _p(out)) == 0
        return out

    def sap_witness_map(self, curve, nc, ni, nv, ea, eb, ec, w, m, d1, d2):
        """(coefficients_for_H (m + 1), the nc + ni extra variables) of r1cs_to_sap_witness_map (r1cs_to_sap.tcc:297-481);
        ea, eb, ec = the first nc rows of r1cs_evaluate, m = the SAP domain size."""
        ea, eb, ec, w = (np.ascontiguousarray(x, dtype=np.uint64) for x in (ea, eb, ec, w))
        d = np.ascontiguousarray(np.stack([np.asarray(x, dtype=np.uint64).reshape(-1) for x in (d1, d2)]))
        H = np.zeros((m + 1, w.shape[1]), dtype=np.uint64)
        extra = np.zeros((nc + ni, w.shape[1]), dtype=np.uint64)
        rc = self.lib.port_sap_witness_map(_cid(curve), nc, ni, nv, _p(ea), _p(eb), _p(ec), _p(w), m, _p(d), _p(H), _p(extra))
        assert rc == 0, rc
        return H, extra

    def field_op(self, curve, field, op, a, b=None):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        b = a if b is None else np.ascontiguousarray(b, dtype=np.uint64)
        out = np.zeros_like(a)
        assert self.lib.port_field_op(_cid(curve), field, op, _p(a), _p(b), _p(out), a.shape[0]) == 0
        return out

    def group_op(self, curve, group, op, a, b=None):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        b = a if b is None else np.ascontiguousarray(b, dtype=np.uint64)
        out = np.zeros_like(a)
        assert self.lib.port_group_op(_cid(curve), group, op, _p(a), _p(b), _p(out), a.shape[0]) == 0
        return out
