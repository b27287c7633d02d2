"""libsnark_b200 — B200-native drop-in for the Groth16 proving hot path of clearmatics/libsnark.

This package is a thin ctypes binding of the C-ABI in include/libsnark_b200.h (implemented by
libsnark_b200/libsnark_b200.so: hand-written CUDA for sm_100a). The function and class names
mirror the reference operators they replace:

  multi_exp / multi_exp_filter_one_zero   libff/algebra/scalar_multiplication/multiexp.hpp:63-88
  kc_multi_exp_with_mixed_addition        libsnark/knowledge_commitment/kc_multiexp.hpp:32-44
  evaluation_domain (basic radix-2)       libfqfft/evaluation_domain/evaluation_domain.hpp:37-105
  r1cs_to_qap_witness_map_h               libsnark/reductions/r1cs_to_qap/r1cs_to_qap.tcc:252-330

Arrays are numpy uint64 views of the reference's in-memory layouts (Montgomery limbs). There
is no CPU fallback: a missing shared library or CUDA device raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libsnark_b200.so")

ALT_BN128, BLS12_377, BW6_761 = 0, 1, 2
CURVES = {"alt_bn128": ALT_BN128, "bls12_377": BLS12_377, "bw6_761": BW6_761}
G1, G2 = 1, 2
FFT, IFFT, COSET_FFT, ICOSET_FFT = 0, 1, 2, 3
OK, ERR_ARG, ERR_CUDA, ERR_DOMAIN, ERR_HANDLE = 0, -1, -2, -3, -4

# every symbol include/libsnark_b200.h declares (tests check the .so exports all of them)
EXPORTS = [
    "b200_version", "b200_init", "b200_device_count", "b200_last_error", "b200_sizeof_fr", "b200_sizeof_group",
    "b200_dev_alloc", "b200_dev_free", "b200_memcpy_h2d", "b200_memcpy_d2h", "b200_sync",
    "b200_msm", "b200_bases_upload", "b200_bases_free", "b200_bases_size", "b200_msm_resident",
    "b200_msm_num_windows", "b200_msm_window_bits", "b200_msm_window_bytes", "b200_msm_resident_windows", "b200_msm_fold_windows",
    "b200_msm_filter_one_zero", "b200_kc_msm",
    "b200_ntt", "b200_ntt_dev", "b200_divide_by_Z_on_coset", "b200_set_two_adicity", "b200_domain_kind", "b200_witness_map_r1cs_zk", "b200_witness_map_h", "b200_witness_map_h_dev",
    "b200_msm_resident_windows_dev", "b200_timer_start", "b200_timer_stop", "b200_make_bases",
    "b200_imad_carry_peak", "b200_imad_peak", "b200_last_msm_timing", "b200_launch_count",
    "b200_r1cs_upload", "b200_r1cs_free", "b200_r1cs_evaluate", "b200_witness_map_r1cs_dev", "b200_gather_dev",
    "b200_groth16_pk_create", "b200_groth16_pk_free", "b200_groth16_prove",
    "b200_witness_chain_r1cs_dev", "b200_witness_finish_dev",
    "b200_shard_range", "b200_groth16_shard_bytes", "b200_groth16_pk_create_shard", "b200_groth16_prove_shard",
    "b200_groth16_shard_begin", "b200_groth16_shard_chain", "b200_groth16_shard_h", "b200_groth16_shard_msm",
    "b200_groth16_assemble", "b200_bases_precompute", "b200_bases_table_rows", "b200_groth16_shard_witness_copy", "b200_batch_exp", "b200_groth16_shard_finish_msm",
    "b200_groth16_witness_upload", "b200_groth16_prove_resident", "b200_bases_table_window_bits",
    "b200_ntt_dev_noalloc", "b200_groth16_shard_msm_begin", "b200_groth16_shard_msm_end",
    "b200_groth16_assemble_prepare", "b200_groth16_shard_h_copy",
]


class B200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libsnark_b200 status {code}: {msg}")
        self.code = code


class DomainSizeException(B200Error):
    """libfqfft::DomainSizeException (libfqfft/tools/exceptions.hpp:21-31)."""


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise FileNotFoundError(
                f"{_LIB_PATH} is missing: build it with `make -C libsnark_b200/csrc` "
                "(or __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(_LIB_PATH)
        vp, sz, i, u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64
        L.b200_last_error.restype = C.c_char_p
        L.b200_sizeof_fr.restype = sz
        L.b200_sizeof_fr.argtypes = [i]
        L.b200_sizeof_group.restype = sz
        L.b200_sizeof_group.argtypes = [i, i]
        L.b200_init.argtypes = [i]
        L.b200_dev_alloc.argtypes = [sz, C.POINTER(vp)]
        L.b200_dev_free.argtypes = [vp]
        L.b200_memcpy_h2d.argtypes = [vp, vp, sz]
        L.b200_memcpy_d2h.argtypes = [vp, vp, sz]
        L.b200_msm.argtypes = [i, i, vp, sz, sz, vp, vp]
        L.b200_bases_upload.argtypes = [i, i, vp, sz, sz, C.POINTER(u64)]
        L.b200_bases_free.argtypes = [u64]
        L.b200_bases_precompute.argtypes = [u64]
        L.b200_bases_table_rows.argtypes = [u64]
        L.b200_bases_table_window_bits.argtypes = [u64]
        L.b200_bases_size.restype = sz
        L.b200_bases_size.argtypes = [u64]
        L.b200_msm_resident.argtypes = [u64, sz, sz, vp, i, vp]
        L.b200_msm_num_windows.argtypes = [i, sz]
        L.b200_msm_window_bits.argtypes = [i, sz]
        L.b200_msm_window_bytes.restype = sz
        L.b200_msm_window_bytes.argtypes = [i, i]
        L.b200_msm_resident_windows.argtypes = [u64, sz, sz, vp, i, vp]
        L.b200_msm_fold_windows.argtypes = [i, i, sz, vp, sz, vp]
        L.b200_msm_filter_one_zero.argtypes = [i, i, vp, sz, sz, vp, vp, vp]
        L.b200_kc_msm.argtypes = [i, vp, vp, sz, sz, sz, vp, sz, vp, vp, vp]
        L.b200_ntt.argtypes = [i, i, vp, sz, vp]
        L.b200_ntt_dev.argtypes = [i, i, vp, sz, vp]
        L.b200_ntt_dev_noalloc.argtypes = [i, i, vp, sz, vp]
        L.b200_divide_by_Z_on_coset.argtypes = [i, vp, sz]
        L.b200_set_two_adicity.argtypes = [i, i]
        L.b200_domain_kind.argtypes = [i, sz]
        L.b200_witness_map_h.argtypes = [i, vp, vp, vp, sz, vp]
        L.b200_witness_map_h_dev.argtypes = [i, vp, vp, vp, sz, vp]
        L.b200_imad_peak.argtypes = [i, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.b200_r1cs_upload.argtypes = [i, sz, sz, sz, vp, vp, vp, C.POINTER(u64)]
        L.b200_r1cs_free.argtypes = [u64]
        L.b200_r1cs_evaluate.argtypes = [u64, vp, sz, vp, vp, vp]
        L.b200_witness_map_r1cs_dev.argtypes = [u64, vp, sz, vp]
        L.b200_witness_map_r1cs_zk.argtypes = [u64, vp, sz, vp, vp]
        L.b200_gather_dev.argtypes = [i, vp, vp, sz, vp]
        L.b200_groth16_pk_create.argtypes = [i, vp, vp, vp, vp, vp, u64, u64, u64, vp, u64, u64, u64, sz, sz, sz, sz,
                                             C.POINTER(u64)]
        L.b200_groth16_pk_free.argtypes = [u64]
        L.b200_groth16_prove.argtypes = [u64, vp, vp, vp, vp, vp, C.POINTER(C.c_double)]
        L.b200_groth16_witness_upload.argtypes = [u64, vp, vp]
        L.b200_groth16_prove_resident.argtypes = [u64, vp, vp, vp, C.POINTER(C.c_double)]
        L.b200_witness_chain_r1cs_dev.argtypes = [u64, vp, sz, i, vp]
        L.b200_witness_finish_dev.argtypes = [i, vp, vp, vp, sz, vp]
        L.b200_shard_range.restype = None
        L.b200_shard_range.argtypes = [sz, i, i, C.POINTER(sz), C.POINTER(sz)]
        L.b200_groth16_shard_bytes.restype = sz
        L.b200_groth16_shard_bytes.argtypes = [i, C.POINTER(sz), i]
        L.b200_groth16_pk_create_shard.argtypes = [i, vp, vp, vp, vp, vp, u64, u64, u64, vp, sz, u64, u64, u64, sz, sz, sz,
                                                   sz, i, i, C.POINTER(u64)]
        L.b200_groth16_prove_shard.argtypes = [u64, vp, vp, vp, C.POINTER(C.c_double)]
        L.b200_groth16_shard_begin.argtypes = [u64, vp, vp]
        L.b200_groth16_shard_chain.argtypes = [u64, i, vp]
        L.b200_groth16_shard_h_copy.argtypes = [u64, vp, sz, sz, i]
        L.b200_groth16_shard_witness_copy.argtypes = [u64, vp, i]
        L.b200_groth16_shard_h.argtypes = [u64, vp, vp, vp]
        L.b200_groth16_shard_msm.argtypes = [u64, C.c_uint, vp]
        L.b200_groth16_shard_finish_msm.argtypes = [u64, vp, vp, vp, C.c_uint, vp]
        L.b200_groth16_shard_msm_begin.argtypes = [u64, C.c_uint, vp]
        L.b200_groth16_shard_msm_end.argtypes = [u64]
        L.b200_groth16_assemble_prepare.argtypes = [i, vp, vp, vp, vp]
        L.b200_groth16_assemble.argtypes = [i, vp, vp, vp, vp, vp, C.POINTER(sz), i, vp, vp, vp, vp]
        L.b200_batch_exp.argtypes = [i, i, vp, vp, vp, sz, vp]
        L.b200_last_msm_timing.argtypes = [C.POINTER(C.c_double)]
        L.b200_launch_count.restype = u64
        L.b200_msm_resident_windows_dev.argtypes = [u64, sz, sz, vp, i, vp]
        L.b200_timer_stop.argtypes = [C.POINTER(C.c_double)]
        L.b200_make_bases.argtypes = [i, i, u64, sz, vp]
        L.b200_imad_carry_peak.restype = C.c_double
        L.b200_test_field_op.argtypes = [i, i, i, vp, vp, vp, sz]
        L.b200_test_fp_convert.argtypes = [i, i, i, vp, vp, sz]
        L.b200_test_group_op.argtypes = [i, i, i, vp, vp, vp, sz]
        L.b200_test_signed_digits.argtypes = [i, vp, sz, i, vp, C.POINTER(i)]
        _lib = L
    return _lib


def _check(rc):
    if rc != OK:
        msg = lib().b200_last_error().decode(errors="replace")
        if rc == ERR_DOMAIN:
            raise DomainSizeException(rc, msg)
        raise B200Error(rc, msg)


def _cid(curve):
    return CURVES[curve] if isinstance(curve, str) else int(curve)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _u64(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


def init(device=0):
    _check(lib().b200_init(int(device)))


def device_count():
    return lib().b200_device_count()


def launch_count():
    return int(lib().b200_launch_count())


def sizeof_fr(curve):
    return lib().b200_sizeof_fr(_cid(curve))


def sizeof_group(curve, group):
    return lib().b200_sizeof_group(_cid(curve), group)


# ---- MSM -------------------------------------------------------------------------------------

def multi_exp(curve, group, bases, scalars, base_stride=None):
    """libff::multi_exp<GroupT, Fr, BDLO12_signed, special> -> one normalised point (u64 limbs)."""
    c = _cid(curve)
    bases, scalars = _u64(bases), _u64(scalars)
    n = scalars.shape[0]
    stride = base_stride if base_stride is not None else (bases.strides[0] if bases.ndim == 2 else 0)
    out = np.zeros(sizeof_group(c, group) // 8, dtype=np.uint64)
    _check(lib().b200_msm(c, group, _p(bases), stride, n, _p(scalars), _p(out)))
    return out


def multi_exp_filter_one_zero(curve, group, bases, scalars, base_stride=None, return_stats=False):
    """libff::multi_exp_filter_one_zero (upstream: multi_exp_with_mixed_addition)."""
    c = _cid(curve)
    bases, scalars = _u64(bases), _u64(scalars)
    n = scalars.shape[0]
    stride = base_stride if base_stride is not None else (bases.strides[0] if bases.ndim == 2 else 0)
    out = np.zeros(sizeof_group(c, group) // 8, dtype=np.uint64)
    stats = np.zeros(3, dtype=np.uint64)
    _check(lib().b200_msm_filter_one_zero(c, group, _p(bases), stride, n, _p(scalars), _p(out), _p(stats)))
    return (out, stats) if return_stats else out


def kc_multi_exp_with_mixed_addition(curve, indices, values, min_idx, max_idx, scalars):
    """libsnark::kc_multi_exp_with_mixed_addition over a sparse vector of (G2, G1) pairs -> (g2, g1)."""
    c = _cid(curve)
    indices, values, scalars = _u64(indices), _u64(values), _u64(scalars)
    g2 = np.zeros(sizeof_group(c, G2) // 8, dtype=np.uint64)
    g1 = np.zeros(sizeof_group(c, G1) // 8, dtype=np.uint64)
    _check(lib().b200_kc_msm(c, _p(indices), _p(values), indices.shape[0], min_idx, max_idx, _p(scalars),
                             scalars.shape[0], _p(g2), _p(g1), None))
    return g2, g1


def batch_exp(curve, group, g, scalars, coeff=None):
    """libff::batch_exp (coeff None) / batch_exp_with_coeff followed by batch_to_special: (coeff * scalars[i]) * g for
    every i, in special form (multiexp.tcc:879-974). g: one point in special form."""
    c = _cid(curve)
    g, scalars = _u64(g), _u64(scalars)
    coeff = None if coeff is None else _u64(coeff)
    out = np.zeros((scalars.shape[0], sizeof_group(c, group) // 8), dtype=np.uint64)
    _check(lib().b200_batch_exp(c, group, _p(g), _p(coeff), _p(scalars), scalars.shape[0], _p(out)))
    return out


class ResidentBases:
    """A proving-key query kept on the device (b200_bases_upload)."""

    def __init__(self, curve, group, bases, base_stride=None):
        self.curve, self.group = _cid(curve), group
        bases = _u64(bases)
        n = bases.shape[0]
        stride = base_stride if base_stride is not None else bases.strides[0]
        h = C.c_uint64(0)
        _check(lib().b200_bases_upload(self.curve, group, _p(bases), stride, n, C.byref(h)))
        self.handle, self.n = h.value, n

    def precompute(self):
        """b200_bases_precompute: rows [2^(c w)] P on the device; returns the number of rows."""
        _check(lib().b200_bases_precompute(self.handle))
        return lib().b200_bases_table_rows(self.handle)

    def multi_exp(self, scalars, first=0, n=None, device_ptr=None):
        n = self.n - first if n is None else n
        out = np.zeros(sizeof_group(self.curve, self.group) // 8, dtype=np.uint64)
        if device_ptr is not None:
            _check(lib().b200_msm_resident(self.handle, first, n, C.c_void_p(device_ptr), 1, _p(out)))
        else:
            scalars = _u64(scalars)
            _check(lib().b200_msm_resident(self.handle, first, n, _p(scalars), 0, _p(out)))
        return out

    def window_sums(self, scalars, first=0, n=None, device_ptr=None):
        n = self.n - first if n is None else n
        W = lib().b200_msm_num_windows(self.curve, n)
        wb = lib().b200_msm_window_bytes(self.curve, self.group)
        out = np.zeros(W * wb // 8, dtype=np.uint64)
        if device_ptr is not None:
            _check(lib().b200_msm_resident_windows(self.handle, first, n, C.c_void_p(device_ptr), 1, _p(out)))
        else:
            scalars = _u64(scalars)
            _check(lib().b200_msm_resident_windows(self.handle, first, n, _p(scalars), 0, _p(out)))
        return out

    def window_sums_dev(self, dev_out, scalars=None, first=0, n=None, device_ptr=None):
        """Window sums left in device memory at address dev_out (e.g. a torch CUDA tensor)."""
        n = self.n - first if n is None else n
        if device_ptr is not None:
            _check(lib().b200_msm_resident_windows_dev(self.handle, first, n, C.c_void_p(device_ptr), 1,
                                                       C.c_void_p(dev_out)))
        else:
            scalars = _u64(scalars)
            _check(lib().b200_msm_resident_windows_dev(self.handle, first, n, _p(scalars), 0, C.c_void_p(dev_out)))

    def free(self):
        if self.handle:
            _check(lib().b200_bases_free(self.handle))
            self.handle = 0


def fold_windows(curve, group, n_per_shard, windows, num_shards):
    c = _cid(curve)
    windows = _u64(windows)
    out = np.zeros(sizeof_group(c, group) // 8, dtype=np.uint64)
    _check(lib().b200_msm_fold_windows(c, group, n_per_shard, _p(windows), num_shards, _p(out)))
    return out


def last_msm_timing():
    ms = (C.c_double * 5)()
    _check(lib().b200_last_msm_timing(ms))
    return dict(zip(["sort", "accumulate", "reduce", "fold", "total"], list(ms)))


def timer_start():
    _check(lib().b200_timer_start())


def timer_stop():
    ms = C.c_double(0)
    _check(lib().b200_timer_stop(C.byref(ms)))
    return ms.value


def make_bases(curve, group, n, k0=1):
    """n distinct affine points [k0 + i] * generator in the reference layout, generated on the device."""
    c = _cid(curve)
    out = np.zeros((n, sizeof_group(c, group) // 8), dtype=np.uint64)
    _check(lib().b200_make_bases(c, group, k0, n, _p(out)))
    return out


# the scalar-field moduli (alt_bn128_init.cpp, bls12_377_init.cpp, bw6_761_init.cpp: Fr::modulus)
FR_MODULUS = {ALT_BN128: 0x30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000001,
              BLS12_377: 0x12ab655e9a2ca55660b44d1e5c37b00159aa76fed00000010a11800000000001,
              BW6_761: 0x1ae3a4617c510eac63b05c06ca1493b1a22d9f300f5138f1ef3622fba094800170b5d44300000008508c00000000001}


def scalars_from_ints(curve, values):
    """Python integers (standard form) -> Montgomery Fr limbs, the layout every entry point takes."""
    c = _cid(curve)
    r, L = FR_MODULUS[c], sizeof_fr(c) // 8
    out = np.zeros((len(values), L), dtype=np.uint64)
    for k, v in enumerate(values):
        m = (int(v) % r) * (1 << (64 * L)) % r
        for j in range(L):
            out[k, j] = (m >> (64 * j)) & 0xFFFFFFFFFFFFFFFF
    return out


def random_scalars(curve, n, seed=0):
    """n uniformly random Fr residues (any value < r is a valid Montgomery representative)."""
    c = _cid(curve)
    r = FR_MODULUS[c]
    L = sizeof_fr(c) // 8
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 64, size=(n, L), dtype=np.uint64)
    top = r >> (64 * (L - 1))
    a[:, L - 1] %= np.uint64(top)  # top limb strictly below r's top limb -> value < r
    return a


def imad_peak(iters=4096):
    """(IMAD.WIDE no carry, IMAD lo, IMAD.WIDE.X carry chains) multiply-adds per second."""
    w, lo = C.c_double(0), C.c_double(0)
    _check(lib().b200_imad_peak(iters, C.byref(w), C.byref(lo)))
    return w.value, lo.value, lib().b200_imad_carry_peak()


# ---- NTT -------------------------------------------------------------------------------------

def set_two_adicity(curve, s):
    """Mirrors libff's mutable FieldT::s (b200_set_two_adicity): with s' in force, 2^(s'+1) is the extended domain and
    larger radix-2 sizes are refused. s = 0 restores the field's own 2-adicity."""
    _check(lib().b200_set_two_adicity(_cid(curve), int(s)))


DOMAIN_NONE, DOMAIN_BASIC, DOMAIN_EXTENDED, DOMAIN_STEP = 0, 1, 2, 3


def domain_kind(curve, m):
    """Which radix-2 domain class serves exactly m elements (b200_domain_kind); DOMAIN_NONE when none does."""
    return lib().b200_domain_kind(_cid(curve), int(m))


class evaluation_domain:
    """libfqfft::basic_radix2_domain<Fr> (m = 2^k, k <= s), extended_radix2_domain<Fr> (m = 2^(s+1)) or
    step_radix2_domain<Fr> (m = 2^k + 2^r): the domain kinds get_evaluation_domain hands out for these fields; the
    library picks the kind from m."""

    def __init__(self, curve, m):
        self.curve, self.m = _cid(curve), int(m)

    def _run(self, op, a, g=None):
        a = np.array(a, dtype=np.uint64, order="C", copy=True)
        if a.shape[0] != self.m:
            raise DomainSizeException(ERR_DOMAIN, "expected a.size() == m")  # basic_radix2_domain.tcc:57
        g_ = None if g is None else _u64(g)
        _check(lib().b200_ntt(self.curve, op, _p(a), self.m, _p(g_)))
        return a

    def FFT(self, a):
        return self._run(FFT, a)

    def iFFT(self, a):
        return self._run(IFFT, a)

    def cosetFFT(self, a, g):
        return self._run(COSET_FFT, a, g)

    def icosetFFT(self, a, g):
        return self._run(ICOSET_FFT, a, g)

    def divide_by_Z_on_coset(self, a):
        a = np.array(a, dtype=np.uint64, order="C", copy=True)
        _check(lib().b200_divide_by_Z_on_coset(self.curve, _p(a), self.m))
        return a


def domain_size_for(min_size):
    """The size get_evaluation_domain settles on (get_evaluation_domain.tcc:37-58): min_size when it is 2^k or
    2^k + 2^r, else big + the small part rounded up to a power of two."""
    min_size = int(min_size)
    if min_size < 2 or min_size & (min_size - 1) == 0:
        return min_size
    big = 1 << (min_size.bit_length() - 1)
    small = min_size - big
    rounded = 1 << (small - 1).bit_length() if small > 1 else 1
    return big + rounded


def get_evaluation_domain(curve, min_size):
    """libfqfft::get_evaluation_domain (get_evaluation_domain.tcc:32-70): basic_radix2_domain for powers of two up to
    2^s, extended_radix2_domain for 2^(s+1), step_radix2_domain otherwise (the geometric / arithmetic sequence
    domains are never reached for these fields: one of the radix-2 kinds accepts every size a GPU can hold)."""
    return evaluation_domain(curve, domain_size_for(min_size))


def r1cs_to_qap_witness_map_h(curve, aA, aB, aC):
    """The NTT chain of r1cs_to_qap_witness_map: evaluations aA, aB, aC -> coefficients_for_H (m + 1)."""
    c = _cid(curve)
    aA, aB, aC = _u64(aA), _u64(aB), _u64(aC)
    m = aA.shape[0]
    out = np.zeros((m + 1, aA.shape[1]), dtype=np.uint64)
    _check(lib().b200_witness_map_h(c, _p(aA), _p(aB), _p(aC), m, _p(out)))
    return out


class R1CS:
    """A constraint system resident on the device as one CSR (rows: a-, b-, c-combinations of the constraints;
    column 0 = the constant ONE, column k = variable k). Mirrors what libsnark::r1cs_to_qap_witness_map reads from
    r1cs_constraint_system (r1cs_to_qap.tcc:236-248)."""

    def __init__(self, curve, num_constraints, num_inputs, num_variables, row_ptr, col, val):
        self.curve, self.nc, self.ni, self.nv = curve, num_constraints, num_inputs, num_variables
        row_ptr = np.ascontiguousarray(row_ptr, dtype=np.uint32)
        col = np.ascontiguousarray(col, dtype=np.uint32)
        val = _u64(val)
        assert row_ptr.shape[0] == 3 * num_constraints + 1
        h = C.c_uint64(0)
        _check(lib().b200_r1cs_upload(_cid(curve), num_constraints, num_inputs, num_variables, _p(row_ptr), _p(col),
                                      _p(val), C.byref(h)))
        self.handle = h.value
        self._limbs = sizeof_fr(curve) // 8

    def evaluate(self, w, m):
        """(aA, aB, aC), m rows each, from w = (1 | primary | auxiliary)."""
        w = _u64(w)
        assert w.shape[0] == self.nv + 1
        out = [np.zeros((m, self._limbs), dtype=np.uint64) for _ in range(3)]
        _check(lib().b200_r1cs_evaluate(self.handle, _p(w), m, _p(out[0]), _p(out[1]), _p(out[2])))
        return out

    def witness_map_h(self, w, m):
        """coefficients_for_H (m + 1 rows) straight from the witness; everything stays on the device in between."""
        w = _u64(w)
        assert w.shape[0] == self.nv + 1
        dw, dh = DeviceBuffer(w.nbytes), DeviceBuffer((m + 1) * self._limbs * 8)
        try:
            dw.upload(w)
            _check(lib().b200_witness_map_r1cs_dev(self.handle, C.c_void_p(dw.ptr), m, C.c_void_p(dh.ptr)))
            return dh.download((m + 1, self._limbs))
        finally:
            dw.free()
            dh.free()

    def witness_map_h_zk(self, w, m, d1, d2, d3):
        """coefficients_for_H of r1cs_to_qap_witness_map(cs, primary, auxiliary, d1, d2, d3) (m + 1 rows), ZK patch included."""
        w = _u64(w)
        assert w.shape[0] == self.nv + 1
        d = np.ascontiguousarray(np.stack([_u64(d1).reshape(-1), _u64(d2).reshape(-1), _u64(d3).reshape(-1)]))
        out = np.zeros((m + 1, self._limbs), dtype=np.uint64)
        _check(lib().b200_witness_map_r1cs_zk(self.handle, _p(w), m, _p(d), _p(out)))
        return out

    def free(self):
        if self.handle:
            _check(lib().b200_r1cs_free(self.handle))
            self.handle = 0


class Groth16ProvingKey:
    """r1cs_gg_ppzksnark_proving_key resident on the device; prove() mirrors r1cs_gg_ppzksnark_prover
    (r1cs_gg_ppzksnark.tcc:460-603) with caller-supplied r, s."""

    def __init__(self, curve, alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2, A_query, B_query_g2, B_query_g1,
                 B_indices, H_query, L_query, r1cs, domain_size, precompute=False):
        self.curve, self.r1cs = curve, r1cs
        self.sets = [ResidentBases(curve, G1, A_query), ResidentBases(curve, G2, B_query_g2),
                     ResidentBases(curve, G1, B_query_g1), ResidentBases(curve, G1, H_query),
                     ResidentBases(curve, G1, L_query)]
        if precompute:
            for rb in self.sets:
                rb.precompute()
        pts = [_u64(x) for x in (alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2)]
        idx = None if B_indices is None else np.ascontiguousarray(B_indices, dtype=np.uint32)
        h = C.c_uint64(0)
        _check(lib().b200_groth16_pk_create(
            _cid(curve), *[_p(x) for x in pts], self.sets[0].handle, self.sets[1].handle, self.sets[2].handle,
            None if idx is None else _p(idx), self.sets[3].handle, self.sets[4].handle, r1cs.handle, r1cs.nc, r1cs.ni,
            r1cs.nv, domain_size, C.byref(h)))
        self.handle = h.value
        self._proof_limbs = (2 * sizeof_group(curve, G1) + sizeof_group(curve, G2)) // 8

    def prove(self, primary_input, auxiliary_input, r, s):
        """(proof, stages): proof = g_A | g_B | g_C normalised (reference layouts), stages in ms."""
        pi, ai, r, s = _u64(primary_input), _u64(auxiliary_input), _u64(r), _u64(s)
        out = np.zeros(self._proof_limbs, dtype=np.uint64)
        ms = (C.c_double * 7)()
        _check(lib().b200_groth16_prove(self.handle, _p(pi), _p(ai), _p(r), _p(s), _p(out), ms))
        return out, dict(zip(["upload", "witness_map", "A", "B", "H", "L", "total"], list(ms)))

    def upload_witness(self, primary_input, auxiliary_input):
        """(1 | primary | auxiliary) into the key's device buffer (b200_groth16_witness_upload)."""
        pi, ai = _u64(primary_input), _u64(auxiliary_input)
        _check(lib().b200_groth16_witness_upload(self.handle, _p(pi), _p(ai)))

    def prove_resident(self, r, s):
        """prove() over the witness last uploaded (b200_groth16_prove_resident): no host-to-device copy."""
        r, s = _u64(r), _u64(s)
        out = np.zeros(self._proof_limbs, dtype=np.uint64)
        ms = (C.c_double * 7)()
        _check(lib().b200_groth16_prove_resident(self.handle, _p(r), _p(s), _p(out), ms))
        return out, dict(zip(["upload", "witness_map", "A", "B", "H", "L", "total"], list(ms)))

    def free(self):
        if self.handle:
            _check(lib().b200_groth16_pk_free(self.handle))
            self.handle = 0
            for rb in self.sets:
                rb.free()


# ---- the prover sharded over several GPUs ------------------------------------------

MSM_A, MSM_B_G2, MSM_B_G1, MSM_H, MSM_L = 1, 2, 4, 8, 16


def shard_range(total, world, rank):
    """(first, count) of rank's contiguous slice of a `total`-entry query (b200_shard_range)."""
    f, c = C.c_size_t(0), C.c_size_t(0)
    lib().b200_shard_range(total, world, rank, C.byref(f), C.byref(c))
    return f.value, c.value


def groth16_shard_bytes(curve, query_sizes, world):
    qs = (C.c_size_t * 4)(*[int(x) for x in query_sizes])
    return lib().b200_groth16_shard_bytes(_cid(curve), qs, world)


def groth16_assemble(curve, alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2, query_sizes, world, blobs, r, s):
    """The gathered window sums of all ranks -> the proof (host arithmetic only; b200_groth16_assemble)."""
    c = _cid(curve)
    pts = [_u64(x) for x in (alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2)]
    blobs = np.ascontiguousarray(blobs).view(np.uint8).reshape(-1)
    qs = (C.c_size_t * 4)(*[int(x) for x in query_sizes])
    if blobs.shape[0] != world * lib().b200_groth16_shard_bytes(c, qs, world):
        raise B200Error(ERR_ARG, "blobs must hold world * b200_groth16_shard_bytes bytes")
    r, s = _u64(r), _u64(s)
    out = np.zeros((2 * sizeof_group(c, G1) + sizeof_group(c, G2)) // 8, dtype=np.uint64)
    _check(lib().b200_groth16_assemble(c, *[_p(x) for x in pts], qs, world, _p(blobs), _p(r), _p(s), _p(out)))
    return out


class Groth16ProvingKeyShard:
    """Rank `rank` of `world`'s share of a proving key: the slice b200_shard_range gives it of each of A_query,
    B_query (G2 and G1 halves, with the matching slice of the sparse vector's indices), H_query and L_query, plus
    the whole constraint system. With sliced=False the arguments are the whole queries and the slices are cut here;
    with sliced=True they already are this rank's slices and B_total = the whole B query's entry count."""

    def __init__(self, curve, alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2, A_query, B_query_g2, B_query_g1,
                 B_indices, H_query, L_query, r1cs, domain_size, rank, world, sliced=False, B_total=None,
                 precompute=False):
        self.curve, self.r1cs, self.rank, self.world, self.m = curve, r1cs, rank, world, int(domain_size)
        nv, ni = r1cs.nv, r1cs.ni
        if B_total is None:
            B_total = B_query_g2.shape[0] if not sliced else nv + 1
        self.query_sizes = [nv + 1, int(B_total), self.m - 1, nv - ni]
        rng = [shard_range(t, world, rank) for t in self.query_sizes]
        self.ranges = rng

        def cut(a, q):
            if a is None:
                return None
            return a if sliced else a[rng[q][0]:rng[q][0] + rng[q][1]]
        qs = [cut(A_query, 0), cut(B_query_g2, 1), cut(B_query_g1, 1), cut(H_query[:self.m - 1] if not sliced else H_query, 2),
              cut(L_query, 3)]
        groups = [G1, G2, G1, G1, G1]
        self.sets = [ResidentBases(curve, g, q) if q.shape[0] else None for g, q in zip(groups, qs)]
        if precompute:
            for rb in self.sets:
                if rb is not None:
                    rb.precompute()
        idx = cut(None if B_indices is None else np.ascontiguousarray(B_indices, dtype=np.uint32), 1)
        if idx is not None:
            idx = np.ascontiguousarray(idx)
        self.points = [_u64(x).copy() for x in (alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2)]
        hs = [0 if x is None else x.handle for x in self.sets]
        h = C.c_uint64(0)
        _check(lib().b200_groth16_pk_create_shard(
            _cid(curve), *[_p(x) for x in self.points], hs[0], hs[1], hs[2], None if idx is None else _p(idx),
            self.query_sizes[1], hs[3], hs[4], r1cs.handle, r1cs.nc, ni, nv, self.m, rank, world, C.byref(h)))
        self.handle = h.value
        self.blob_bytes = groth16_shard_bytes(curve, self.query_sizes, world)
        self.fr_bytes = sizeof_fr(curve)

    def prove_shard(self, primary_input, auxiliary_input, dev_blob):
        """This rank's five partial MSMs (window sums) into the device blob at address dev_blob; the witness map is
        computed locally. Returns the stage times in ms."""
        pi, ai = _u64(primary_input), _u64(auxiliary_input)
        ms = (C.c_double * 7)()
        _check(lib().b200_groth16_prove_shard(self.handle, _p(pi), _p(ai), C.c_void_p(dev_blob), ms))
        return dict(zip(["upload", "witness_map", "A", "B", "H", "L", "total"], list(ms)))

    def begin(self, primary_input, auxiliary_input):
        pi, ai = _u64(primary_input), _u64(auxiliary_input)
        _check(lib().b200_groth16_shard_begin(self.handle, _p(pi), _p(ai)))

    def witness_copy(self, dev_buf, to_key):
        _check(lib().b200_groth16_shard_witness_copy(self.handle, C.c_void_p(dev_buf), int(to_key)))

    def chain(self, which, dev_out):
        _check(lib().b200_groth16_shard_chain(self.handle, which, C.c_void_p(dev_out)))

    def h_copy(self, dev_buf, first, count, to_key):
        _check(lib().b200_groth16_shard_h_copy(self.handle, C.c_void_p(dev_buf), first, count, int(to_key)))

    def h(self, dA=None, dB=None, dC=None):
        _check(lib().b200_groth16_shard_h(self.handle, *[None if x is None else C.c_void_p(x) for x in (dA, dB, dC)]))

    def msm(self, mask, dev_blob):
        _check(lib().b200_groth16_shard_msm(self.handle, mask, C.c_void_p(dev_blob)))

    def msm_begin(self, mask, dev_blob):
        """The witness-only MSMs (A, B.g2, B.g1, L) started in the background; msm_end waits for them."""
        _check(lib().b200_groth16_shard_msm_begin(self.handle, mask, C.c_void_p(dev_blob)))

    def msm_end(self):
        _check(lib().b200_groth16_shard_msm_end(self.handle))

    def finish_msm(self, dA, dB, dC, mask, dev_blob):
        _check(lib().b200_groth16_shard_finish_msm(self.handle, C.c_void_p(dA), C.c_void_p(dB), C.c_void_p(dC), mask,
                                                   C.c_void_p(dev_blob)))

    def assemble_prepare(self, r, s):
        """Starts the r, s-only scalar multiplications of the assembly on host threads (b200_groth16_assemble_prepare)."""
        r, s = _u64(r), _u64(s)
        _check(lib().b200_groth16_assemble_prepare(_cid(self.curve), _p(self.points[3]), _p(self.points[4]), _p(r), _p(s)))

    def assemble(self, r, s, blobs):
        return groth16_assemble(self.curve, *self.points, self.query_sizes, self.world, blobs, r, s)

    def free(self):
        if self.handle:
            _check(lib().b200_groth16_pk_free(self.handle))
            self.handle = 0
            for rb in self.sets:
                if rb is not None:
                    rb.free()


class DistributedProver:
    """r1cs_gg_ppzksnark_prover over the ranks of an initialised torch.distributed (NCCL) process group, one
    Groth16ProvingKeyShard per rank. Per proof: every rank uploads the witness; with chain_parallel (>= 3 ranks) ranks 0, 1, 2 each run one of the A, B, C iFFT/cosetFFT chains and broadcast the result
    while all ranks already run their A, B and L shard MSMs, then every rank finishes H — otherwise each rank
    computes the whole witness map; the ranks' window sums are all-gathered (a few KB per rank) and folded on the
    host. torch is used for the NCCL plumbing and the device buffers the collectives work on, nothing else."""

    def __init__(self, shard, chain_parallel=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.pk = torch, dist, shard
        self.dev = torch.device("cuda", torch.cuda.current_device())
        served = domain_kind(shard.curve, shard.m) != DOMAIN_NONE   # basic, extended or step radix-2 domain
        self.chain_parallel = (shard.world >= 3 and served) if chain_parallel is None else bool(chain_parallel)
        if self.chain_parallel and not (shard.world >= 3 and served):
            raise B200Error(ERR_ARG, "chain-parallel witness map needs >= 3 ranks and a radix-2 domain size")
        self.blob = torch.zeros(shard.blob_bytes, dtype=torch.uint8, device=self.dev)
        self.gathered = torch.zeros(shard.world * shard.blob_bytes, dtype=torch.uint8, device=self.dev)
        # rank 0 holds all three chain results, ranks 1 and 2 their own; hbuf: H coefficients (whole on rank 0, a slice elsewhere)
        self.chains = ([torch.empty(shard.m * shard.fr_bytes, dtype=torch.uint8, device=self.dev)
                        if (shard.rank == 0 or shard.rank == i) else None for i in range(3)] if self.chain_parallel else None)
        self.hbuf = (torch.empty((shard.m if shard.rank == 0 else shard.ranges[2][1] + 1) * shard.fr_bytes, dtype=torch.uint8,
                                 device=self.dev) if self.chain_parallel else None)
        self._copy_stream, self._whead = None, None
        self.wbuf = (torch.empty((shard.r1cs.nv + 1) * shard.fr_bytes, dtype=torch.uint8, device=self.dev)
                     if shard.world > 1 else None)

    def _distribute_witness(self, primary_input, auxiliary_input):
        """Rank 0 uploads the witness once; the other ranks get it over NVLink (NCCL broadcast) — the witness lives in one
        process, and eight host uploads of it would share one host memory system. The upload and the broadcast are cut
        into slices: slice k travels over NVLink while slice k + 1 crosses PCIe (one copy stream; the collective of a
        slice is ordered behind its copy and runs on NCCL's own stream)."""
        pk, torch, dist = self.pk, self.torch, self.dist
        if pk.world == 1:
            pk.begin(primary_input, auxiliary_input)
            return
        fr = pk.fr_bytes
        total = self.wbuf.numel()
        head = (1 + pk.r1cs.ni) * fr
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(priority=-1)
            self._whead = torch.empty(head, dtype=torch.uint8, pin_memory=True)
        nslices = 4 if total >= (8 << 20) else 1
        cuts = [head + ((total - head) * k // nslices) // 256 * 256 for k in range(nslices)] + [total]
        works = []
        with torch.cuda.stream(self._copy_stream):
            if pk.rank == 0:
                hv = self._whead.numpy()
                hv[:fr] = np.ascontiguousarray(scalars_from_ints(pk.curve, [1])).view(np.uint8).reshape(-1)
                hv[fr:] = _u64(primary_input).view(np.uint8).reshape(-1)
                aux = torch.from_numpy(_u64(auxiliary_input).view(np.uint8).reshape(-1))
            for k in range(nslices):
                a, b_ = (0 if k == 0 else cuts[k]), cuts[k + 1]
                if pk.rank == 0:
                    if k == 0:
                        self.wbuf[:head].copy_(self._whead, non_blocking=True)
                    lo = max(a, head)
                    self.wbuf[lo:b_].copy_(aux[lo - head:b_ - head], non_blocking=True)
                works.append(dist.broadcast(self.wbuf[a:b_], src=0, async_op=True))
        for w in works:
            w.wait()
        torch.cuda.current_stream().synchronize()
        pk.witness_copy(self.wbuf.data_ptr(), 1)

    def prove(self, primary_input, auxiliary_input, r, s, fold_on_all_ranks=False):
        """The proof on rank 0 (on every rank with fold_on_all_ranks), None elsewhere. Only rank 0 reads primary_input and
        auxiliary_input. self.last_stages holds this rank's host wall clock per phase in ms."""
        import time
        torch, dist, pk = self.torch, self.dist, self.pk
        t = [time.perf_counter()]
        names = []

        def lap(name):
            t.append(time.perf_counter())
            names.append(name)
        if pk.rank == 0 or fold_on_all_ranks:
            pk.assemble_prepare(r, s)   # host threads, beside everything below
        self._distribute_witness(primary_input, auxiliary_input)
        lap("upload_broadcast")
        # the A, B and L shard MSMs need nothing but the witness: they run in the background from here on, beside the
        # chains, their broadcast and the H computation (whose kernels go to higher-priority streams)
        pk.msm_begin(MSM_A | MSM_B_G2 | MSM_B_G1 | MSM_L, self.blob.data_ptr())
        try:
            if self.chain_parallel:
                # ranks 0, 1, 2 each run one of the A, B, C chains; ranks 1 and 2 send theirs to rank 0, which finishes
                # coefficients_for_H (pointwise step + inverse coset transform) and hands every rank the slice its H
                # shard MSM reads: 2 x 32 MiB into rank 0 and 32 MiB out of it at 2^20, where broadcasting the three
                # chains moved 96 MiB to every rank (3.2 ms of an 8.6 ms proof on 8 GPUs)
                fr = pk.fr_bytes
                if pk.rank < 3:
                    pk.chain(pk.rank, self.chains[pk.rank].data_ptr())
                lap("chain")
                if pk.rank == 0:
                    ops = [dist.P2POp(dist.irecv, self.chains[i], i) for i in (1, 2)]
                    for w in dist.batch_isend_irecv(ops):
                        w.wait()
                    torch.cuda.current_stream().synchronize()
                    pk.h(self.chains[0].data_ptr(), self.chains[1].data_ptr(), self.chains[2].data_ptr())
                    pk.h_copy(self.hbuf.data_ptr(), 0, pk.m - 1, 0)
                    ops = []
                    for g in range(1, pk.world):
                        f, c = shard_range(pk.m - 1, pk.world, g)
                        if c:
                            ops.append(dist.P2POp(dist.isend, self.hbuf[f * fr:(f + c) * fr], g))
                    for w in dist.batch_isend_irecv(ops):
                        w.wait()
                else:
                    if pk.rank in (1, 2):
                        for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, self.chains[pk.rank], 0)]):
                            w.wait()
                    f, c = pk.ranges[2]
                    if c:
                        for w in dist.batch_isend_irecv([dist.P2POp(dist.irecv, self.hbuf[:c * fr], 0)]):
                            w.wait()
                        torch.cuda.current_stream().synchronize()
                        pk.h_copy(self.hbuf.data_ptr(), f, c, 1)
                lap("h_exchange")
                pk.msm(MSM_H, self.blob.data_ptr())
                lap("h_msm")
            else:
                pk.h()
                lap("witness_map")
                pk.msm(MSM_H, self.blob.data_ptr())
                lap("h_msm")
        finally:
            pk.msm_end()
        lap("other_msms_tail")
        if pk.world > 1:
            dist.all_gather_into_tensor(self.gathered, self.blob)
            src = self.gathered
        else:
            src = self.blob
        out = None
        if pk.rank == 0 or fold_on_all_ranks:
            host = src.cpu().numpy()
            lap("all_gather_d2h")
            out = pk.assemble(r, s, host)
            lap("assemble")
        else:
            torch.cuda.current_stream().synchronize()
            lap("all_gather")
        self.last_stages = {n: (t[i + 1] - t[i]) * 1e3 for i, n in enumerate(names)}
        return out


class DeviceBuffer:
    def __init__(self, nbytes):
        p = C.c_void_p(0)
        _check(lib().b200_dev_alloc(nbytes, C.byref(p)))
        self.ptr, self.nbytes = p.value, nbytes

    def upload(self, a):
        a = np.ascontiguousarray(a)
        _check(lib().b200_memcpy_h2d(C.c_void_p(self.ptr), _p(a), a.nbytes))

    def download(self, shape, dtype=np.uint64):
        out = np.zeros(shape, dtype=dtype)
        _check(lib().b200_memcpy_d2h(_p(out), C.c_void_p(self.ptr), out.nbytes))
        return out

    def free(self):
        if self.ptr:
            _check(lib().b200_dev_free(C.c_void_p(self.ptr)))
            self.ptr = 0


# ---- device unit-test hooks (selftest.cu) ------------------------------------------------------

def test_field_op(curve, field, op, a, b=None):
    a = _u64(a)
    b = a if b is None else _u64(b)
    out = np.zeros_like(a)
    _check(lib().b200_test_field_op(_cid(curve), field, op, _p(a), _p(b), _p(out), a.shape[0]))
    return out


def test_fp_convert(curve, field, to_mont, a):
    a = _u64(a)
    out = np.zeros_like(a)
    _check(lib().b200_test_fp_convert(_cid(curve), field, int(to_mont), _p(a), _p(out), a.shape[0]))
    return out


def test_group_op(curve, group, op, a, b):
    a, b = _u64(a), _u64(b)
    out = np.zeros_like(a)
    _check(lib().b200_test_group_op(_cid(curve), group, op, _p(a), _p(b), _p(out), a.shape[0]))
    return out


def test_signed_digits(curve, scalars, c):
    """The device's digit extraction (k_prepare + window_digit) for window width c: (n, W) int64."""
    scalars = _u64(scalars)
    n = scalars.shape[0]
    bits = {ALT_BN128: 254, BLS12_377: 253, BW6_761: 377}[_cid(curve)]
    W = C.c_int(0)
    out = np.zeros(n * ((bits + c) // c + 1), dtype=np.int64)
    _check(lib().b200_test_signed_digits(_cid(curve), _p(scalars), n, c, _p(out), C.byref(W)))
    return out[: n * W.value].reshape(n, W.value)


test_field_op.__test__ = test_fp_convert.__test__ = test_group_op.__test__ = test_signed_digits.__test__ = False
