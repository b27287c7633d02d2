// oracle/ref_harness.cpp — TEST INFRASTRUCTURE, not product code.
//
// A C-ABI window onto the UNMODIFIED reference (clearmatics/libsnark with its
// libff / libfqfft submodules), compiled from the sources where they lie under
// /root/reference by oracle/Makefile into oracle/_ref/libref_oracle.so.
// Everything computed here is computed by the reference's own code; this file
// only marshals raw limb buffers (the reference's in-memory layouts) in and out.
//
// The same file is compiled a second time with -DB200_OVERLAY and
// `-include libsnark_b200/overlay/libsnark_b200_overlay.hpp` into
// oracle/_ref/libref_overlay.so: there the reference's prover / witness map run
// unchanged, but libff::multi_exp*, kc_multi_exp_with_mixed_addition and
// libfqfft::get_evaluation_domain resolve to the CUDA drop-in. That build is how
// "whole proofs with r,s fixed" are compared.
//
// Only tests/, bench.py's cpu_baseline / --impl reference legs and
// __graft_entry__.smoke() may load these libraries.

#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>
#ifdef MULTICORE
#include <omp.h>
#endif

#include <libff/algebra/curves/alt_bn128/alt_bn128_pp.hpp>
#include <libff/algebra/curves/bls12_377/bls12_377_pp.hpp>
#include <libff/algebra/curves/bw6_761/bw6_761_pp.hpp>
#include <libff/algebra/fields/field_utils.hpp>
// multiexp.tcc's ordered_exponent::operator< (bos_coster, used only by the verifier's input
// accumulation) has an x86 asm compare with no "memory" clobber; gcc 13 -O2 moves the heap
// swaps across it and the verifier then rejects valid proofs for some thread counts. Including
// the header with USE_ASM off selects its own mpn_cmp branch; the field arithmetic headers were
// already included above with USE_ASM on, so nothing on the timed path changes.
#ifdef USE_ASM
#undef USE_ASM
#define REF_HARNESS_RESTORE_USE_ASM
#endif
#include <libff/algebra/scalar_multiplication/multiexp.hpp>
#ifdef REF_HARNESS_RESTORE_USE_ASM
#define USE_ASM
#undef REF_HARNESS_RESTORE_USE_ASM
#endif
#include <libff/common/profiling.hpp>
#include <libff/common/rng.hpp>
#include <libfqfft/evaluation_domain/domains/basic_radix2_domain.hpp>
#include <libfqfft/evaluation_domain/get_evaluation_domain.hpp>
#include <libfqfft/polynomial_arithmetic/naive_evaluate.hpp>
#include <libsnark/knowledge_commitment/kc_multiexp.hpp>
#include <libsnark/reductions/r1cs_to_qap/r1cs_to_qap.hpp>
#include <libsnark/reductions/r1cs_to_sap/r1cs_to_sap.hpp>
#include <libsnark/relations/constraint_satisfaction_problems/r1cs/examples/r1cs_examples.hpp>
#include <libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.hpp>
#include <libsnark/zk_proof_systems/ppzksnark/r1cs_ppzksnark/r1cs_ppzksnark.hpp>
#include <libsnark/zk_proof_systems/ppzksnark/r1cs_se_ppzksnark/r1cs_se_ppzksnark.hpp>
#include <sstream>

#define API extern "C" __attribute__((visibility("default")))

namespace
{

using namespace libff;
using namespace libsnark;

constexpr multi_exp_method kMethod = multi_exp_method_BDLO12_signed;
constexpr multi_exp_base_form kForm = multi_exp_base_form_special;

template<typename FieldT> struct is_fp2 : std::false_type {
};
template<mp_size_t n, const bigint<n> &m>
struct is_fp2<Fp2_model<n, m>> : std::true_type {
};

// Groth16 fixture: an R1CS example, keys (real or synthetic) and scratch.
template<typename ppT> struct groth16_fixture {
    r1cs_example<Fr<ppT>> example;
    r1cs_gg_ppzksnark_keypair<ppT> keypair;
    bool real_keys;
    groth16_fixture(r1cs_example<Fr<ppT>> &&e, r1cs_gg_ppzksnark_keypair<ppT> &&k, bool real)
        : example(std::move(e)), keypair(std::move(k)), real_keys(real)
    {
    }
};

// SHA512_rng<FieldT> (rng.tcc:26-69) asserts size_in_bits <= 512; bw6_761's Fq is 761 bits.
// For such fields derive a full-width element from three 64-byte draws with the
// reference's own field arithmetic: lo + hi * 2^384-ish mix (x*y + z).
template<typename FieldT, bool BIG = (FieldT::num_limbs > 8)> struct det_rng {
    static FieldT get(uint64_t idx) { return SHA512_rng<FieldT>(idx); }
};
template<typename FieldT> struct det_rng<FieldT, true> {
    static FieldT get(uint64_t idx)
    {
        typedef Fp_model<6, bls12_377_modulus_q> SmallT; // 377-bit draws
        bigint<FieldT::num_limbs> b[3];
        for (int k = 0; k < 3; ++k) {
            const auto v = SHA512_rng<SmallT>(idx * 3 + k).as_bigint();
            b[k].clear();
            for (int i = 0; i < 6; ++i) {
                b[k].data[i] = v.data[i];
            }
        }
        return FieldT(b[0]) * FieldT(b[1]) + FieldT(b[2]);
    }
};

template<typename GroupT> void make_distinct_special(std::vector<GroupT> &v, size_t n, uint64_t seed)
{
    // n distinct affine points: P_0 = [h]G, P_i = P_{i-1} + [h']G, then one
    // batched normalisation.
    typedef typename GroupT::scalar_field FrT;
    const FrT h = SHA512_rng<FrT>(seed * 2 + 0x1000000ull);
    const FrT h2 = SHA512_rng<FrT>(seed * 2 + 0x1000001ull);
    GroupT cur = h * GroupT::one();
    const GroupT step = h2 * GroupT::one();
    v.resize(n);
    for (size_t i = 0; i < n; ++i) {
        v[i] = cur;
        cur = cur + step;
    }
    batch_to_special<GroupT>(v);
}

template<typename ppT> struct H {
    typedef Fr<ppT> FrT;
    typedef Fq<ppT> FqT;
    typedef G1<ppT> G1T;
    typedef G2<ppT> G2T;

    static void info(uint64_t *out)
    {
        out[0] = FrT::num_limbs;
        out[1] = FqT::num_limbs;
        out[2] = sizeof(FrT);
        out[3] = sizeof(FqT);
        out[4] = sizeof(G1T);
        out[5] = sizeof(G2T);
        out[6] = FrT::s;
        out[7] = FrT::num_bits;
        out[8] = FqT::num_bits;
        out[9] = sizeof(typename G2T::base_field);
    }

    template<typename FieldT> static void field_consts(uint64_t *out)
    {
        // modulus | R^2 | R^3 | inv | one(mont) | multiplicative_generator(mont) | root_of_unity(mont)
        const size_t n = FieldT::num_limbs;
        memcpy(out, FieldT::mod.data, 8 * n);
        memcpy(out + n, FieldT::Rsquared.data, 8 * n);
        memcpy(out + 2 * n, FieldT::Rcubed.data, 8 * n);
        out[3 * n] = FieldT::inv;
        const FieldT one = FieldT::one();
        memcpy(out + 3 * n + 1, one.mont_repr.data, 8 * n);
        memcpy(out + 4 * n + 1, FieldT::multiplicative_generator.mont_repr.data, 8 * n);
        memcpy(out + 5 * n + 1, FieldT::root_of_unity.mont_repr.data, 8 * n);
    }

    template<typename FieldT> static int sha512(uint64_t idx0, size_t n, void *out)
    {
        FieldT *o = static_cast<FieldT *>(out);
#ifdef MULTICORE
#pragma omp parallel for
#endif
        for (size_t i = 0; i < n; ++i) {
            o[i] = det_rng<FieldT>::get(idx0 + i);
        }
        return 0;
    }

    // op: 0 add, 1 sub, 2 mul, 3 sqr, 4 neg, 5 inverse, 6 as_bigint (from Montgomery),
    //     7 from bigint (to Montgomery)
    template<typename FieldT>
    static int field_op(int op, const void *a_, const void *b_, void *out_, size_t n)
    {
        const FieldT *a = static_cast<const FieldT *>(a_);
        const FieldT *b = static_cast<const FieldT *>(b_);
        FieldT *o = static_cast<FieldT *>(out_);
        for (size_t i = 0; i < n; ++i) {
            switch (op) {
            case 0: o[i] = a[i] + b[i]; break;
            case 1: o[i] = a[i] - b[i]; break;
            case 2: o[i] = a[i] * b[i]; break;
            case 3: o[i] = a[i].squared(); break;
            case 4: o[i] = -a[i]; break;
            case 5: o[i] = a[i].is_zero() ? a[i] : a[i].inverse(); break;
            default: return -1;
            }
        }
        return 0;
    }

    template<typename FieldT>
    static int fp_convert(int to_mont, const void *a_, void *out_, size_t n)
    {
        const size_t L = FieldT::num_limbs;
        for (size_t i = 0; i < n; ++i) {
            if (to_mont) {
                bigint<FieldT::num_limbs> b;
                memcpy(b.data, static_cast<const uint64_t *>(a_) + i * L, 8 * L);
                FieldT f(b);
                memcpy(static_cast<uint64_t *>(out_) + i * L, f.mont_repr.data, 8 * L);
            } else {
                const FieldT &f = static_cast<const FieldT *>(a_)[i];
                const auto b = f.as_bigint();
                memcpy(static_cast<uint64_t *>(out_) + i * L, b.data, 8 * L);
            }
        }
        return 0;
    }

    static int signed_digits(const void *scalars, size_t n, size_t c, size_t num_digits, int64_t *out)
    {
        const FrT *s = static_cast<const FrT *>(scalars);
        for (size_t i = 0; i < n; ++i) {
            const auto b = s[i].as_bigint();
            for (size_t d = 0; d < num_digits; ++d) {
                out[i * num_digits + d] = field_get_signed_digit(b, c, d);
            }
        }
        return 0;
    }

    // op: 0 add (operator+), 1 mixed_add, 2 dbl, 3 neg, 4 to_affine_coordinates, 5 add()
    template<typename GroupT>
    static int group_op(int op, const void *a_, const void *b_, void *out_, size_t n)
    {
        const GroupT *a = static_cast<const GroupT *>(a_);
        const GroupT *b = static_cast<const GroupT *>(b_);
        GroupT *o = static_cast<GroupT *>(out_);
        for (size_t i = 0; i < n; ++i) {
            switch (op) {
            case 0: o[i] = a[i] + b[i]; break;
            case 1: o[i] = a[i].mixed_add(b[i]); break;
            case 2: o[i] = a[i].dbl(); break;
            case 3: o[i] = -a[i]; break;
            case 4:
                o[i] = a[i];
                o[i].to_affine_coordinates();
                break;
            case 5: o[i] = a[i].add(b[i]); break;
            default: return -1;
            }
        }
        return 0;
    }

    template<typename GroupT> static int make_bases(size_t n, uint64_t seed, void *out)
    {
        std::vector<GroupT> v;
        make_distinct_special<GroupT>(v, n, seed);
        memcpy(out, v.data(), n * sizeof(GroupT));
        return 0;
    }

    // The reference's key-generation operators as r1cs_gg_ppzksnark_generator uses them (.tcc:321-398):
    // get_exp_window_size + get_window_table + batch_exp / batch_exp_with_coeff + batch_to_special.
    template<typename GroupT>
    static int batch_exp_special(const void *g_, const void *coeff_, const void *scalars, size_t n, void *out)
    {
        GroupT g;
        memcpy(&g, g_, sizeof(GroupT));
        const FrT *sp = static_cast<const FrT *>(scalars);
        const std::vector<FrT> v(sp, sp + n);
        const size_t scalar_size = FrT::size_in_bits();
        const size_t window = get_exp_window_size<GroupT>(n ? n : 1);
        const window_table<GroupT> table = get_window_table(scalar_size, window, g);
        std::vector<GroupT> res;
        if (coeff_) {
            FrT coeff;
            memcpy(&coeff, coeff_, sizeof(FrT));
            res = batch_exp_with_coeff(scalar_size, window, table, coeff, v);
        } else {
            res = batch_exp(scalar_size, window, table, v);
        }
        batch_to_special<GroupT>(res);
        memcpy(out, res.data(), n * sizeof(GroupT));
        return 0;
    }

    // libsnark::kc_batch_exp<G2, G1, Fr>(..., output_special = true) over v (zeros are skipped: sparse result).
    // out_values: room for n (G2 | G1) pairs; out_indices: room for n; *nnz = entries written.
    static int kc_batch_exp_special(const void *g2_, const void *g1_, const void *c2_, const void *c1_, const void *scalars,
                                    size_t n, size_t chunks, void *out_values, uint64_t *out_indices, uint64_t *nnz)
    {
        G2T g2;
        G1T g1;
        FrT c2, c1;
        memcpy(&g2, g2_, sizeof(G2T));
        memcpy(&g1, g1_, sizeof(G1T));
        memcpy(&c2, c2_, sizeof(FrT));
        memcpy(&c1, c1_, sizeof(FrT));
        const FrT *sp = static_cast<const FrT *>(scalars);
        const std::vector<FrT> v(sp, sp + n);
        const size_t scalar_size = FrT::size_in_bits();
        const size_t w2 = get_exp_window_size<G2T>(n ? n : 1), w1 = get_exp_window_size<G1T>(n ? n : 1);
        const window_table<G2T> t2 = get_window_table(scalar_size, w2, g2);
        const window_table<G1T> t1 = get_window_table(scalar_size, w1, g1);
        const knowledge_commitment_vector<G2T, G1T> r =
            kc_batch_exp<G2T, G1T, FrT>(scalar_size, w2, w1, t2, t1, c2, c1, v, chunks, true);
        *nnz = r.values.size();
        memcpy(out_values, r.values.data(), r.values.size() * sizeof(knowledge_commitment<G2T, G1T>));
        for (size_t i = 0; i < r.indices.size(); ++i) {
            out_indices[i] = r.indices[i];
        }
        return 0;
    }

    // out[i] = [s_i] * generator, normalised to special (affine) form.
    template<typename GroupT> static int scalar_mul_gen(const void *scalars, size_t n, void *out)
    {
        const FrT *s = static_cast<const FrT *>(scalars);
        std::vector<GroupT> v(n);
#ifdef MULTICORE
#pragma omp parallel for
#endif
        for (size_t i = 0; i < n; ++i) {
            v[i] = s[i] * GroupT::one();
        }
        batch_to_special<GroupT>(v);
        memcpy(out, v.data(), n * sizeof(GroupT));
        return 0;
    }

    // method: 0 = multi_exp<BDLO12_signed, special>, 1 = multi_exp<naive_plain>,
    //         2 = multi_exp_filter_one_zero<BDLO12_signed, special>,
    //         3 = multi_exp<BDLO12_signed, normal>
    template<typename GroupT>
    static int msm(int method, const void *bases, size_t n, const void *scalars, size_t chunks, void *out)
    {
        const GroupT *b = static_cast<const GroupT *>(bases);
        const FrT *s = static_cast<const FrT *>(scalars);
        std::vector<GroupT> bv(b, b + n);
        std::vector<FrT> sv(s, s + n);
        GroupT r;
        if (n == 0) {
            r = GroupT::zero();
        } else
            switch (method) {
            case 0:
                r = multi_exp<GroupT, FrT, kMethod, kForm>(bv.begin(), bv.end(), sv.begin(), sv.end(), chunks);
                break;
            case 1:
                r = multi_exp<GroupT, FrT, multi_exp_method_naive_plain>(
                    bv.begin(), bv.end(), sv.begin(), sv.end(), chunks);
                break;
            case 2:
                r = multi_exp_filter_one_zero<GroupT, FrT, kMethod, kForm>(
                    bv.begin(), bv.end(), sv.begin(), sv.end(), chunks);
                break;
            case 3:
                r = multi_exp<GroupT, FrT, kMethod, multi_exp_base_form_normal>(
                    bv.begin(), bv.end(), sv.begin(), sv.end(), chunks);
                break;
            default:
                return -1;
            }
        r.to_affine_coordinates();
        memcpy(out, &r, sizeof(GroupT));
        return 0;
    }

    // kc_multi_exp_with_mixed_addition over a sparse vector of (G2, G1) pairs.
    // values: nnz * (sizeof(G2)+sizeof(G1)) in the reference's knowledge_commitment layout.
    // T1 = G2T: Groth16's B query; T1 = G1T: r1cs_ppzksnark's A and C queries (r1cs_ppzksnark.tcc:570-612)
    template<typename T1>
    static int kc_msm_t(
        const uint64_t *indices,
        const void *values,
        size_t nnz,
        size_t domain_size,
        size_t min_idx,
        size_t max_idx,
        const void *scalars,
        size_t nscalars,
        size_t chunks,
        void *out_g,
        void *out_h)
    {
        typedef knowledge_commitment<T1, G1T> KC;
        knowledge_commitment_vector<T1, G1T> vec;
        vec.domain_size_ = domain_size;
        vec.indices.assign(indices, indices + nnz);
        const KC *v = static_cast<const KC *>(values);
        vec.values.assign(v, v + nnz);
        const FrT *s = static_cast<const FrT *>(scalars);
        std::vector<FrT> sv(s, s + nscalars);
        KC r = kc_multi_exp_with_mixed_addition<T1, G1T, FrT, kMethod, kForm>(
            vec, min_idx, max_idx, sv.begin(), sv.end(), chunks);
        r.g.to_affine_coordinates();
        r.h.to_affine_coordinates();
        memcpy(out_g, &r.g, sizeof(T1));
        memcpy(out_h, &r.h, sizeof(G1T));
        return 0;
    }
    static int kc_msm(
        const uint64_t *indices,
        const void *values,
        size_t nnz,
        size_t domain_size,
        size_t min_idx,
        size_t max_idx,
        const void *scalars,
        size_t nscalars,
        size_t chunks,
        void *out_g2,
        void *out_g1)
    {
        return kc_msm_t<G2T>(indices, values, nnz, domain_size, min_idx, max_idx, scalars, nscalars, chunks, out_g2, out_g1);
    }
    static int kc_msm_g1(
        const uint64_t *indices,
        const void *values,
        size_t nnz,
        size_t domain_size,
        size_t min_idx,
        size_t max_idx,
        const void *scalars,
        size_t nscalars,
        size_t chunks,
        void *out_g,
        void *out_h)
    {
        return kc_msm_t<G1T>(indices, values, nnz, domain_size, min_idx, max_idx, scalars, nscalars, chunks, out_g, out_h);
    }

    // op: 0 FFT, 1 iFFT, 2 cosetFFT(g), 3 icosetFFT(g), 4 divide_by_Z_on_coset
    // via_factory != 0 obtains the domain from get_evaluation_domain (the
    // interposable entry point), else constructs basic_radix2_domain directly.
    static int fft(int op, void *data, size_t m, const void *g_, int via_factory)
    {
        FrT *d = static_cast<FrT *>(data);
        std::vector<FrT> a(d, d + m);
        std::shared_ptr<libfqfft::evaluation_domain<FrT>> dom;
        try {
            if (via_factory) {
                dom = libfqfft::get_evaluation_domain<FrT>(m);
            } else {
                dom.reset(new libfqfft::basic_radix2_domain<FrT>(m));
            }
            const FrT g = g_ ? *static_cast<const FrT *>(g_) : FrT::multiplicative_generator;
            switch (op) {
            case 0: dom->FFT(a); break;
            case 1: dom->iFFT(a); break;
            case 2: dom->cosetFFT(a, g); break;
            case 3: dom->icosetFFT(a, g); break;
            case 4: dom->divide_by_Z_on_coset(a); break;
            default: return -1;
            }
        } catch (const libfqfft::DomainSizeException &) { // not derived from std::exception
            return -2;
        } catch (const libfqfft::InvalidSizeException &) {
            return -2;
        } catch (const std::exception &) {
            return -3;
        }
        memcpy(d, a.data(), m * sizeof(FrT));
        return 0;
    }

    // Horner evaluation of the polynomial `coeffs` at omega^idx[i] (naive
    // check for large FFTs).
    static int poly_eval_at_domain(const void *coeffs, size_t m, const uint64_t *idx, size_t k, void *out)
    {
        const FrT *c = static_cast<const FrT *>(coeffs);
        std::vector<FrT> cv(c, c + m);
        libfqfft::basic_radix2_domain<FrT> dom(m);
        FrT *o = static_cast<FrT *>(out);
        for (size_t i = 0; i < k; ++i) {
            const FrT x = dom.get_domain_element(idx[i]);
            o[i] = libfqfft::evaluate_polynomial(m, cv, x);
        }
        return 0;
    }

    // The same through the factory's domain (extended_radix2_domain / step_radix2_domain elements), the
    // way libfqfft's own evaluation_domain_test.cpp checks every domain kind's FFT.
    static int poly_eval_at_factory_domain(const void *coeffs, size_t m, const uint64_t *idx, size_t k, void *out)
    {
        const FrT *c = static_cast<const FrT *>(coeffs);
        std::vector<FrT> cv(c, c + m);
        FrT *o = static_cast<FrT *>(out);
        try {
            std::shared_ptr<libfqfft::evaluation_domain<FrT>> dom = libfqfft::get_evaluation_domain<FrT>(m);
            for (size_t i = 0; i < k; ++i) {
                o[i] = libfqfft::evaluate_polynomial(m, cv, dom->get_domain_element(idx[i]));
            }
        } catch (const libfqfft::DomainSizeException &) {
            return -2;
        }
        return 0;
    }

    // libff keeps FieldT::s and FieldT::root_of_unity as mutable static parameters (set by init_*_params).
    // Lowering them consistently (root -> root^(2^(s0 - s))) makes the UNMODIFIED get_evaluation_domain hand
    // out extended_radix2_domain at 2^(s+1), a size a test can hold; s = 0 restores the field's own pair.
    static int set_two_adicity(int s)
    {
        static const size_t s0 = FrT::s;
        static const FrT root0 = FrT::root_of_unity;
        if (s < 0 || (size_t)s > s0) {
            return -1;
        }
        const size_t k = s ? (size_t)s : s0;
        FrT r = root0;
        for (size_t i = s0; i > k; --i) {
            r = r.squared();
        }
        FrT::s = k;
        FrT::root_of_unity = r;
        return 0;
    }

    // ---------------- Groth16 -------------------------------------------

    static groth16_fixture<ppT> *setup(size_t num_constraints, size_t num_inputs, int real_keys)
    {
        r1cs_example<FrT> ex = generate_r1cs_example_with_field_input<FrT>(num_constraints, num_inputs);
        if (real_keys) {
            r1cs_gg_ppzksnark_keypair<ppT> kp = r1cs_gg_ppzksnark_generator<ppT>(ex.constraint_system);
            return new groth16_fixture<ppT>(std::move(ex), std::move(kp), true);
        }
        auto *f = new groth16_fixture<ppT>(std::move(ex), r1cs_gg_ppzksnark_keypair<ppT>(), false);
        {
            // Synthetic proving key: queries of the right sizes filled with
            // distinct affine points (proofs will not verify; the prover's
            // work and the CPU/GPU parity are unchanged).
            auto &pk = f->keypair.pk;
            const auto &cs = f->example.constraint_system;
            const size_t nv = cs.num_variables();
            const size_t m = libfqfft::get_evaluation_domain<FrT>(cs.num_constraints() + cs.num_inputs() + 1)->m;
            pk.alpha_g1 = SHA512_rng<FrT>(11) * G1T::one();
            pk.beta_g1 = SHA512_rng<FrT>(12) * G1T::one();
            pk.beta_g2 = SHA512_rng<FrT>(12) * G2T::one();
            pk.delta_g1 = SHA512_rng<FrT>(13) * G1T::one();
            pk.delta_g2 = SHA512_rng<FrT>(13) * G2T::one();
            make_distinct_special<G1T>(pk.A_query, nv + 1, 21);
            make_distinct_special<G1T>(pk.H_query, m - 1, 22);
            make_distinct_special<G1T>(pk.L_query, nv - cs.num_inputs(), 23);
            std::vector<G2T> bg;
            std::vector<G1T> bh;
            make_distinct_special<G2T>(bg, nv + 1, 24);
            make_distinct_special<G1T>(bh, nv + 1, 25);
            std::vector<knowledge_commitment<G2T, G1T>> kv(nv + 1);
            for (size_t i = 0; i < nv + 1; ++i) {
                kv[i] = knowledge_commitment<G2T, G1T>(bg[i], bh[i]);
            }
            pk.B_query = knowledge_commitment_vector<G2T, G1T>(std::move(kv));
            pk.constraint_system = cs;
        }
        return f;
    }

    static void sizes(const groth16_fixture<ppT> *f, uint64_t *out)
    {
        const auto &pk = f->keypair.pk;
        const auto &cs = f->example.constraint_system;
        out[0] = cs.num_constraints();
        out[1] = cs.num_inputs();
        out[2] = cs.num_variables();
        out[3] = libfqfft::get_evaluation_domain<FrT>(cs.num_constraints() + cs.num_inputs() + 1)->m;
        out[4] = pk.A_query.size();
        out[5] = pk.B_query.indices.size();
        out[6] = pk.B_query.domain_size();
        out[7] = pk.H_query.size();
        out[8] = pk.L_query.size();
    }

    // which: 0 A_query, 1 B_query.values, 2 B_query.indices, 3 H_query, 4 L_query,
    //        5 primary_input, 6 auxiliary_input, 7 alpha_g1, 8 beta_g1, 9 beta_g2, 10 delta_g1, 11 delta_g2
    //        (single points, any projective representative)
    static const void *ptr(const groth16_fixture<ppT> *f, int which)
    {
        const auto &pk = f->keypair.pk;
        switch (which) {
        case 0: return pk.A_query.data();
        case 1: return pk.B_query.values.data();
        case 2: return pk.B_query.indices.data();
        case 3: return pk.H_query.data();
        case 4: return pk.L_query.data();
        case 5: return f->example.primary_input.data();
        case 6: return f->example.auxiliary_input.data();
        case 7: return &pk.alpha_g1;
        case 8: return &pk.beta_g1;
        case 9: return &pk.beta_g2;
        case 10: return &pk.delta_g1;
        case 11: return &pk.delta_g2;
        }
        return nullptr;
    }

    static int witness_map(const groth16_fixture<ppT> *f, void *out_H, double *ms)
    {
        const long long t0 = get_nsec_time();
        const qap_witness<FrT> w = r1cs_to_qap_witness_map(
            f->keypair.pk.constraint_system,
            f->example.primary_input,
            f->example.auxiliary_input,
            FrT::zero(),
            FrT::zero(),
            FrT::zero(),
            false);
        if (ms) {
            *ms = (get_nsec_time() - t0) * 1e-6;
        }
        memcpy(out_H, w.coefficients_for_H.data(), w.coefficients_for_H.size() * sizeof(FrT));
        return 0;
    }

    // the same with the ZK patch: d123 = d1, d2, d3 (r1cs_to_qap.tcc:262-271, :327-333; what r1cs_ppzksnark's prover passes)
    static int witness_map_zk(const groth16_fixture<ppT> *f, const void *d123, void *out_H, double *ms)
    {
        const FrT *d = static_cast<const FrT *>(d123);
        const long long t0 = get_nsec_time();
        const qap_witness<FrT> w = r1cs_to_qap_witness_map(
            f->keypair.pk.constraint_system, f->example.primary_input, f->example.auxiliary_input, d[0], d[1], d[2], false);
        if (ms) {
            *ms = (get_nsec_time() - t0) * 1e-6;
        }
        memcpy(out_H, w.coefficients_for_H.data(), w.coefficients_for_H.size() * sizeof(FrT));
        return 0;
    }

    // r1cs_to_sap_witness_map (r1cs_to_sap.tcc:297-481; r1cs_se_ppzksnark's witness map) on the fixture's constraint
    // system: d12 = d1, d2; out_H gets degree + 1 coefficients (cap elements available), *degree the SAP domain size,
    // out_extra (optional, nc + ni elements) the extra variables appended to the assignment
    static int sap_witness_map(const groth16_fixture<ppT> *f, const void *d12, void *out_H, size_t cap, size_t *degree,
                               void *out_extra, double *ms)
    {
        const FrT *d = static_cast<const FrT *>(d12);
        const long long t0 = get_nsec_time();
        const sap_witness<FrT> w = r1cs_to_sap_witness_map(
            f->keypair.pk.constraint_system, f->example.primary_input, f->example.auxiliary_input, d[0], d[1]);
        if (ms) {
            *ms = (get_nsec_time() - t0) * 1e-6;
        }
        *degree = w.degree();
        if (w.coefficients_for_H.size() > cap) {
            return -5;
        }
        memcpy(out_H, w.coefficients_for_H.data(), w.coefficients_for_H.size() * sizeof(FrT));
        if (out_extra) {
            const size_t nv = f->keypair.pk.constraint_system.num_variables();
            memcpy(out_extra, w.coefficients_for_ACs.data() + nv, (w.coefficients_for_ACs.size() - nv) * sizeof(FrT));
        }
        return 0;
    }

    // The evaluations aA, aB, aC that r1cs_to_qap_witness_map feeds its first iFFTs
    // (r1cs_to_qap.tcc:236-248 and :293-296), computed with the reference's own
    // linear_combination::evaluate. m elements each.
    static int abc_evaluations(const groth16_fixture<ppT> *f, void *oA, void *oB, void *oC)
    {
        const auto &cs = f->keypair.pk.constraint_system;
        const size_t m = libfqfft::get_evaluation_domain<FrT>(cs.num_constraints() + cs.num_inputs() + 1)->m;
        r1cs_variable_assignment<FrT> full = f->example.primary_input;
        full.insert(full.end(), f->example.auxiliary_input.begin(), f->example.auxiliary_input.end());
        std::vector<FrT> aA(m, FrT::zero()), aB(m, FrT::zero()), aC(m, FrT::zero());
        for (size_t i = 0; i <= cs.num_inputs(); ++i) {
            aA[i + cs.num_constraints()] = (i > 0 ? full[i - 1] : FrT::one());
        }
#ifdef MULTICORE
#pragma omp parallel for
#endif
        for (size_t i = 0; i < cs.num_constraints(); ++i) {
            aA[i] += cs.constraints[i].a.evaluate(full);
            aB[i] += cs.constraints[i].b.evaluate(full);
            aC[i] += cs.constraints[i].c.evaluate(full);
        }
        memcpy(oA, aA.data(), m * sizeof(FrT));
        memcpy(oB, aB.data(), m * sizeof(FrT));
        memcpy(oC, aC.data(), m * sizeof(FrT));
        return 0;
    }

    // Restates r1cs_gg_ppzksnark.tcc:477-598 with caller-supplied r, s (the
    // reference draws them from std::random_device). Every
    // operator called is the reference's (or, in the overlay build, whatever
    // the overlay header substitutes for it).
    // out: g_A (G1) | g_B (G2) | g_C (G1), each normalised, reference layouts.
    // stage_ms[6]: witness map, A, B, H, L, total.
    static int prove(const groth16_fixture<ppT> *f, const void *r_, const void *s_, void *out, double *stage_ms)
    {
        const auto &pk = f->keypair.pk;
        const FrT r = *static_cast<const FrT *>(r_);
        const FrT s = *static_cast<const FrT *>(s_);
        long long t0 = get_nsec_time(), t = t0, t1;
        double dummy[6];
        if (!stage_ms) {
            stage_ms = dummy;
        }

#ifdef B200_OVERLAY
        // overlay build: the seven transforms run as one device pipeline (b200_witness_map_h)
        const qap_witness<FrT> qap_wit = libsnark_b200::r1cs_to_qap_witness_map_fused(
            pk.constraint_system, f->example.primary_input, f->example.auxiliary_input);
#else
        const qap_witness<FrT> qap_wit = r1cs_to_qap_witness_map(
            pk.constraint_system,
            f->example.primary_input,
            f->example.auxiliary_input,
            FrT::zero(),
            FrT::zero(),
            FrT::zero(),
            false);
#endif
        t1 = get_nsec_time();
        stage_ms[0] = (t1 - t) * 1e-6;
        t = t1;

#ifdef MULTICORE
        const size_t chunks = omp_get_max_threads();
#else
        const size_t chunks = 1;
#endif
        Fr_vector<ppT> const_padded_assignment(1, FrT::one());
        const_padded_assignment.insert(
            const_padded_assignment.end(),
            qap_wit.coefficients_for_ABCs.begin(),
            qap_wit.coefficients_for_ABCs.end());

        G1T evaluation_At = multi_exp_filter_one_zero<G1T, FrT, kMethod, kForm>(
            pk.A_query.begin(),
            pk.A_query.begin() + qap_wit.num_variables() + 1,
            const_padded_assignment.begin(),
            const_padded_assignment.begin() + qap_wit.num_variables() + 1,
            chunks);
        t1 = get_nsec_time();
        stage_ms[1] = (t1 - t) * 1e-6;
        t = t1;

        knowledge_commitment<G2T, G1T> evaluation_Bt = kc_multi_exp_with_mixed_addition<G2T, G1T, FrT, kMethod, kForm>(
            pk.B_query,
            0,
            qap_wit.num_variables() + 1,
            const_padded_assignment.begin(),
            const_padded_assignment.begin() + qap_wit.num_variables() + 1,
            chunks);
        t1 = get_nsec_time();
        stage_ms[2] = (t1 - t) * 1e-6;
        t = t1;

        G1T evaluation_Ht = multi_exp<G1T, FrT, kMethod, kForm>(
            pk.H_query.begin(),
            pk.H_query.begin() + (qap_wit.degree() - 1),
            qap_wit.coefficients_for_H.begin(),
            qap_wit.coefficients_for_H.begin() + (qap_wit.degree() - 1),
            chunks);
        t1 = get_nsec_time();
        stage_ms[3] = (t1 - t) * 1e-6;
        t = t1;

        G1T evaluation_Lt = multi_exp_filter_one_zero<G1T, FrT, kMethod, kForm>(
            pk.L_query.begin(),
            pk.L_query.end(),
            const_padded_assignment.begin() + qap_wit.num_inputs() + 1,
            const_padded_assignment.begin() + qap_wit.num_variables() + 1,
            chunks);
        t1 = get_nsec_time();
        stage_ms[4] = (t1 - t) * 1e-6;
        t = t1;

        G1T g1_A = pk.alpha_g1 + evaluation_At + r * pk.delta_g1;
        G1T g1_B = pk.beta_g1 + evaluation_Bt.h + s * pk.delta_g1;
        G2T g2_B = pk.beta_g2 + evaluation_Bt.g + s * pk.delta_g2;
        G1T g1_C = evaluation_Ht + evaluation_Lt + s * g1_A + r * g1_B - (r * s) * pk.delta_g1;
        stage_ms[5] = (get_nsec_time() - t0) * 1e-6;

        g1_A.to_affine_coordinates();
        g2_B.to_affine_coordinates();
        g1_C.to_affine_coordinates();
        char *o = static_cast<char *>(out);
        memcpy(o, &g1_A, sizeof(G1T));
        memcpy(o + sizeof(G1T), &g2_B, sizeof(G2T));
        memcpy(o + sizeof(G1T) + sizeof(G2T), &g1_C, sizeof(G1T));
        return 0;
    }

    // Overlay build only: the device-resident prover of libsnark_b200_overlay.hpp, same output layout.
    static int prove_fused(const groth16_fixture<ppT> *f, const void *r_, const void *s_, void *out, double *ms)
    {
#ifdef B200_OVERLAY
        const FrT r = *static_cast<const FrT *>(r_);
        const FrT s = *static_cast<const FrT *>(s_);
        const long long t0 = get_nsec_time();
        r1cs_gg_ppzksnark_proof<ppT> p = libsnark_b200::r1cs_gg_ppzksnark_prover_fused<ppT>(
            f->keypair.pk, f->example.primary_input, f->example.auxiliary_input, r, s);
        if (ms) {
            *ms = (get_nsec_time() - t0) * 1e-6;
        }
        G1T a = p.g_A, c = p.g_C;
        G2T b = p.g_B;
        a.to_affine_coordinates();
        b.to_affine_coordinates();
        c.to_affine_coordinates();
        char *o = static_cast<char *>(out);
        memcpy(o, &a, sizeof(G1T));
        memcpy(o + sizeof(G1T), &b, sizeof(G2T));
        memcpy(o + sizeof(G1T) + sizeof(G2T), &c, sizeof(G1T));
        return 0;
#else
        (void)f; (void)r_; (void)s_; (void)out; (void)ms;
        return -7; // the stock build has no device prover
#endif
    }

    // The constraint system as one CSR (rows: a-, b-, c-combinations; col 0 = ONE), for feeding
    // b200_r1cs_upload in tests. Call with null arrays to get nnz.
    static int r1cs_csr(const groth16_fixture<ppT> *f, uint32_t *row_ptr, uint32_t *col, void *val, uint64_t *nnz_out)
    {
        const auto &cs = f->keypair.pk.constraint_system;
        const size_t nc = cs.num_constraints();
        FrT *v = static_cast<FrT *>(val);
        size_t e = 0;
        for (int mat = 0; mat < 3; ++mat) {
            for (size_t i = 0; i < nc; ++i) {
                const linear_combination<FrT> &lc =
                    mat == 0 ? cs.constraints[i].a : (mat == 1 ? cs.constraints[i].b : cs.constraints[i].c);
                if (row_ptr) {
                    row_ptr[mat * nc + i] = (uint32_t)e;
                }
                for (const auto &t : lc.terms) {
                    if (col) {
                        col[e] = (uint32_t)t.index;
                        v[e] = t.coeff;
                    }
                    ++e;
                }
            }
        }
        if (row_ptr) {
            row_ptr[3 * nc] = (uint32_t)e;
        }
        *nnz_out = e;
        return 0;
    }

    // The reference's own prover entry point (random r, s); returns the proof
    // the same way. Used to show the overlay is a drop-in for the unmodified
    // r1cs_gg_ppzksnark_prover.
    static int prove_stock(const groth16_fixture<ppT> *f, void *out, double *ms)
    {
        const long long t0 = get_nsec_time();
        r1cs_gg_ppzksnark_proof<ppT> p = r1cs_gg_ppzksnark_prover<ppT>(
            f->keypair.pk, f->example.primary_input, f->example.auxiliary_input);
        if (ms) {
            *ms = (get_nsec_time() - t0) * 1e-6;
        }
        G1T a = p.g_A, c = p.g_C;
        G2T b = p.g_B;
        a.to_affine_coordinates();
        b.to_affine_coordinates();
        c.to_affine_coordinates();
        char *o = static_cast<char *>(out);
        memcpy(o, &a, sizeof(G1T));
        memcpy(o + sizeof(G1T), &b, sizeof(G2T));
        memcpy(o + sizeof(G1T) + sizeof(G2T), &c, sizeof(G1T));
        return 0;
    }

    // Test hook for the overlay's caches: a DIFFERENT key of identical shape written over the fixture's key at the
    // same host addresses (new generator randomness: every query point and single point changes). mutate_row >= 0
    // also doubles the a- and c-coefficients of that constraint (still satisfied by the same witness) before the
    // key is generated, so that the constraint system changes in place too.
    static int rekey_in_place(const groth16_fixture<ppT> *cf, long mutate_row)
    {
        groth16_fixture<ppT> *f = const_cast<groth16_fixture<ppT> *>(cf);
        if (!f->real_keys) {
            return -1;
        }
        auto &pk = f->keypair.pk;
        if (mutate_row >= 0) {
            if ((size_t)mutate_row >= pk.constraint_system.num_constraints()) {
                return -2;
            }
            for (auto *cs : {&pk.constraint_system, &f->example.constraint_system}) {
                for (auto &t : cs->constraints[mutate_row].a.terms) {
                    t.coeff = t.coeff + t.coeff;
                }
                for (auto &t : cs->constraints[mutate_row].c.terms) {
                    t.coeff = t.coeff + t.coeff;
                }
            }
        }
        if (!pk.constraint_system.is_satisfied(f->example.primary_input, f->example.auxiliary_input)) {
            return -3;
        }
        r1cs_gg_ppzksnark_keypair<ppT> kp = r1cs_gg_ppzksnark_generator<ppT>(pk.constraint_system);
        if (kp.pk.A_query.size() != pk.A_query.size() || kp.pk.H_query.size() != pk.H_query.size() ||
            kp.pk.L_query.size() != pk.L_query.size() || kp.pk.B_query.values.size() != pk.B_query.values.size() ||
            kp.pk.B_query.indices != pk.B_query.indices) {
            return -8;
        }
        std::copy(kp.pk.A_query.begin(), kp.pk.A_query.end(), pk.A_query.begin());
        std::copy(kp.pk.H_query.begin(), kp.pk.H_query.end(), pk.H_query.begin());
        std::copy(kp.pk.L_query.begin(), kp.pk.L_query.end(), pk.L_query.begin());
        std::copy(kp.pk.B_query.values.begin(), kp.pk.B_query.values.end(), pk.B_query.values.begin());
        pk.alpha_g1 = kp.pk.alpha_g1;
        pk.beta_g1 = kp.pk.beta_g1;
        pk.beta_g2 = kp.pk.beta_g2;
        pk.delta_g1 = kp.pk.delta_g1;
        pk.delta_g2 = kp.pk.delta_g2;
        f->keypair.vk = kp.vk;
        return 0;
    }

    static int verify(const groth16_fixture<ppT> *f, const void *proof)
    {
        if (!f->real_keys) {
            return -1;
        }
        const char *p = static_cast<const char *>(proof);
        G1T a, c;
        G2T b;
        memcpy(&a, p, sizeof(G1T));
        memcpy(&b, p + sizeof(G1T), sizeof(G2T));
        memcpy(&c, p + sizeof(G1T) + sizeof(G2T), sizeof(G1T));
        r1cs_gg_ppzksnark_proof<ppT> proof_obj(std::move(a), std::move(b), std::move(c));
        return r1cs_gg_ppzksnark_verifier_strong_IC<ppT>(f->keypair.vk, f->example.primary_input, proof_obj) ? 1 : 0;
    }
};

bool g_inited = false;

} // namespace

#define DISPATCH(curve, expr)                                                                                          \
    switch (curve) {                                                                                                   \
    case 0: {                                                                                                          \
        typedef H<libff::alt_bn128_pp> HH;                                                                             \
        return expr;                                                                                                   \
    }                                                                                                                  \
    case 1: {                                                                                                          \
        typedef H<libff::bls12_377_pp> HH;                                                                             \
        return expr;                                                                                                   \
    }                                                                                                                  \
    case 2: {                                                                                                          \
        typedef H<libff::bw6_761_pp> HH;                                                                               \
        return expr;                                                                                                   \
    }                                                                                                                  \
    default: return -100;                                                                                              \
    }

// field: 0 = Fr, 1 = Fq, 2 = Fq2 (G2's base field; for bw6_761 that is Fq again)
#define DISPATCH_FIELD(curve, field, fn, ...)                                                                          \
    switch (curve * 4 + field) {                                                                                       \
    case 0: return H<libff::alt_bn128_pp>::fn<libff::alt_bn128_Fr>(__VA_ARGS__);                                       \
    case 1: return H<libff::alt_bn128_pp>::fn<libff::alt_bn128_Fq>(__VA_ARGS__);                                       \
    case 2: return H<libff::alt_bn128_pp>::fn<libff::alt_bn128_Fq2>(__VA_ARGS__);                                      \
    case 4: return H<libff::bls12_377_pp>::fn<libff::bls12_377_Fr>(__VA_ARGS__);                                       \
    case 5: return H<libff::bls12_377_pp>::fn<libff::bls12_377_Fq>(__VA_ARGS__);                                       \
    case 6: return H<libff::bls12_377_pp>::fn<libff::bls12_377_Fq2>(__VA_ARGS__);                                      \
    case 8: return H<libff::bw6_761_pp>::fn<libff::bw6_761_Fr>(__VA_ARGS__);                                           \
    case 9: return H<libff::bw6_761_pp>::fn<libff::bw6_761_Fq>(__VA_ARGS__);                                           \
    case 10: return H<libff::bw6_761_pp>::fn<libff::bw6_761_Fq>(__VA_ARGS__);                                          \
    default: return -100;                                                                                              \
    }

#define DISPATCH_FP(curve, field, fn, ...)                                                                             \
    switch (curve * 4 + field) {                                                                                       \
    case 0: return H<libff::alt_bn128_pp>::fn<libff::alt_bn128_Fr>(__VA_ARGS__);                                       \
    case 1: return H<libff::alt_bn128_pp>::fn<libff::alt_bn128_Fq>(__VA_ARGS__);                                       \
    case 4: return H<libff::bls12_377_pp>::fn<libff::bls12_377_Fr>(__VA_ARGS__);                                       \
    case 5: return H<libff::bls12_377_pp>::fn<libff::bls12_377_Fq>(__VA_ARGS__);                                       \
    case 8: return H<libff::bw6_761_pp>::fn<libff::bw6_761_Fr>(__VA_ARGS__);                                           \
    case 9: return H<libff::bw6_761_pp>::fn<libff::bw6_761_Fq>(__VA_ARGS__);                                           \
    default: return -100;                                                                                              \
    }

// group: 1 = G1, 2 = G2
#define DISPATCH_GROUP(curve, group, fn, ...)                                                                          \
    switch (curve * 4 + group) {                                                                                       \
    case 1: return H<libff::alt_bn128_pp>::fn<libff::alt_bn128_G1>(__VA_ARGS__);                                       \
    case 2: return H<libff::alt_bn128_pp>::fn<libff::alt_bn128_G2>(__VA_ARGS__);                                       \
    case 5: return H<libff::bls12_377_pp>::fn<libff::bls12_377_G1>(__VA_ARGS__);                                       \
    case 6: return H<libff::bls12_377_pp>::fn<libff::bls12_377_G2>(__VA_ARGS__);                                       \
    case 9: return H<libff::bw6_761_pp>::fn<libff::bw6_761_G1>(__VA_ARGS__);                                           \
    case 10: return H<libff::bw6_761_pp>::fn<libff::bw6_761_G2>(__VA_ARGS__);                                          \
    default: return -100;                                                                                              \
    }

API int ref_init(void)
{
    if (!g_inited) {
        libff::inhibit_profiling_info = true;
        libff::inhibit_profiling_counters = true;
        libff::alt_bn128_pp::init_public_params();
        libff::bls12_377_pp::init_public_params();
        libff::bw6_761_pp::init_public_params();
        g_inited = true;
    }
    return 0;
}

API int ref_is_overlay(void)
{
#ifdef B200_OVERLAY
    return 1;
#else
    return 0;
#endif
}

API int ref_num_threads(void)
{
#ifdef MULTICORE
    return omp_get_max_threads();
#else
    return 1;
#endif
}

API int ref_set_num_threads(int n)
{
#ifdef MULTICORE
    omp_set_num_threads(n);
#endif
    (void)n;
    return 0;
}

API int ref_set_verbose(int v)
{
    libff::inhibit_profiling_info = !v;
    return 0;
}

static int info_(int curve, uint64_t *out) { DISPATCH(curve, (HH::info(out), 0)); }
API int ref_info(int curve, uint64_t *out) { return info_(curve, out); }

static int field_consts_(int curve, int field, uint64_t *out)
{
    switch (curve * 4 + field) {
    case 0: H<libff::alt_bn128_pp>::field_consts<libff::alt_bn128_Fr>(out); return 0;
    case 1: H<libff::alt_bn128_pp>::field_consts<libff::alt_bn128_Fq>(out); return 0;
    case 4: H<libff::bls12_377_pp>::field_consts<libff::bls12_377_Fr>(out); return 0;
    case 5: H<libff::bls12_377_pp>::field_consts<libff::bls12_377_Fq>(out); return 0;
    case 8: H<libff::bw6_761_pp>::field_consts<libff::bw6_761_Fr>(out); return 0;
    case 9: H<libff::bw6_761_pp>::field_consts<libff::bw6_761_Fq>(out); return 0;
    }
    return -100;
}
API int ref_field_consts(int curve, int field, uint64_t *out) { return field_consts_(curve, field, out); }

// Fq2 non-residue (Montgomery limbs of the Fq element), curve coefficient b for
// G1 (Fq) and the twist coefficient for G2, and the generators (reference layouts).
API int ref_curve_consts(int curve, void *fq2_non_residue, void *g1_one, void *g2_one)
{
    switch (curve) {
    case 0:
        memcpy(fq2_non_residue, &libff::alt_bn128_Fq2::non_residue, sizeof(libff::alt_bn128_Fq));
        memcpy(g1_one, &libff::alt_bn128_G1::G1_one, sizeof(libff::alt_bn128_G1));
        memcpy(g2_one, &libff::alt_bn128_G2::G2_one, sizeof(libff::alt_bn128_G2));
        return 0;
    case 1:
        memcpy(fq2_non_residue, &libff::bls12_377_Fq2::non_residue, sizeof(libff::bls12_377_Fq));
        memcpy(g1_one, &libff::bls12_377_G1::G1_one, sizeof(libff::bls12_377_G1));
        memcpy(g2_one, &libff::bls12_377_G2::G2_one, sizeof(libff::bls12_377_G2));
        return 0;
    case 2:
        memset(fq2_non_residue, 0, sizeof(libff::bw6_761_Fq));
        memcpy(g1_one, &libff::bw6_761_G1::G1_one, sizeof(libff::bw6_761_G1));
        memcpy(g2_one, &libff::bw6_761_G2::G2_one, sizeof(libff::bw6_761_G2));
        return 0;
    }
    return -100;
}

static int sha512_(int curve, int field, uint64_t idx0, size_t n, void *out)
{
    DISPATCH_FP(curve, field, sha512, idx0, n, out);
}
API int ref_field_sha512(int curve, int field, uint64_t idx0, size_t n, void *out)
{
    return sha512_(curve, field, idx0, n, out);
}

static int field_op_(int curve, int field, int op, const void *a, const void *b, void *out, size_t n)
{
    DISPATCH_FIELD(curve, field, field_op, op, a, b, out, n);
}
API int ref_field_op(int curve, int field, int op, const void *a, const void *b, void *out, size_t n)
{
    return field_op_(curve, field, op, a, b, out, n);
}

static int fp_convert_(int curve, int field, int to_mont, const void *a, void *out, size_t n)
{
    DISPATCH_FP(curve, field, fp_convert, to_mont, a, out, n);
}
API int ref_fp_convert(int curve, int field, int to_mont, const void *a, void *out, size_t n)
{
    return fp_convert_(curve, field, to_mont, a, out, n);
}

static int signed_digits_(int curve, const void *s, size_t n, size_t c, size_t nd, int64_t *out)
{
    DISPATCH(curve, HH::signed_digits(s, n, c, nd, out));
}
API int ref_signed_digits(int curve, const void *s, size_t n, size_t c, size_t nd, int64_t *out)
{
    return signed_digits_(curve, s, n, c, nd, out);
}

API uint64_t ref_bdlo12_signed_optimal_c(uint64_t n) { return libff::bdlo12_signed_optimal_c(n); }

static int group_op_(int curve, int group, int op, const void *a, const void *b, void *out, size_t n)
{
    DISPATCH_GROUP(curve, group, group_op, op, a, b, out, n);
}
API int ref_group_op(int curve, int group, int op, const void *a, const void *b, void *out, size_t n)
{
    return group_op_(curve, group, op, a, b, out, n);
}

static int make_bases_(int curve, int group, size_t n, uint64_t seed, void *out)
{
    DISPATCH_GROUP(curve, group, make_bases, n, seed, out);
}
API int ref_make_bases(int curve, int group, size_t n, uint64_t seed, void *out)
{
    return make_bases_(curve, group, n, seed, out);
}

static int batch_exp_(int curve, int group, const void *g, const void *coeff, const void *s, size_t n, void *out)
{
    DISPATCH_GROUP(curve, group, batch_exp_special, g, coeff, s, n, out);
}
API int ref_batch_exp(int curve, int group, const void *g, const void *coeff, const void *s, size_t n, void *out)
{
    return batch_exp_(curve, group, g, coeff, s, n, out);
}
static int kc_batch_exp_(int curve, const void *g2, const void *g1, const void *c2, const void *c1, const void *s, size_t n,
                         size_t chunks, void *out_values, uint64_t *out_indices, uint64_t *nnz)
{
    DISPATCH(curve, HH::kc_batch_exp_special(g2, g1, c2, c1, s, n, chunks, out_values, out_indices, nnz));
}
API int ref_kc_batch_exp(int curve, const void *g2, const void *g1, const void *c2, const void *c1, const void *s, size_t n,
                         size_t chunks, void *out_values, uint64_t *out_indices, uint64_t *nnz)
{
    return kc_batch_exp_(curve, g2, g1, c2, c1, s, n, chunks, out_values, out_indices, nnz);
}

static int scalar_mul_gen_(int curve, int group, const void *s, size_t n, void *out)
{
    DISPATCH_GROUP(curve, group, scalar_mul_gen, s, n, out);
}
API int ref_scalar_mul_gen(int curve, int group, const void *s, size_t n, void *out)
{
    return scalar_mul_gen_(curve, group, s, n, out);
}

static int msm_(int curve, int group, int method, const void *b, size_t n, const void *s, size_t chunks, void *out)
{
    DISPATCH_GROUP(curve, group, msm, method, b, n, s, chunks, out);
}
API int ref_msm(int curve, int group, int method, const void *b, size_t n, const void *s, size_t chunks, void *out)
{
    return msm_(curve, group, method, b, n, s, chunks, out);
}

static int kc_msm_(
    int curve,
    const uint64_t *indices,
    const void *values,
    size_t nnz,
    size_t domain_size,
    size_t min_idx,
    size_t max_idx,
    const void *scalars,
    size_t nscalars,
    size_t chunks,
    void *out_g2,
    void *out_g1)
{
    DISPATCH(
        curve,
        HH::kc_msm(indices, values, nnz, domain_size, min_idx, max_idx, scalars, nscalars, chunks, out_g2, out_g1));
}
static int kc_msm_g1_(
    int curve,
    const uint64_t *indices,
    const void *values,
    size_t nnz,
    size_t domain_size,
    size_t min_idx,
    size_t max_idx,
    const void *scalars,
    size_t nscalars,
    size_t chunks,
    void *out_g,
    void *out_h)
{
    DISPATCH(
        curve,
        HH::kc_msm_g1(indices, values, nnz, domain_size, min_idx, max_idx, scalars, nscalars, chunks, out_g, out_h));
}
API int ref_kc_msm_g1(
    int curve,
    const uint64_t *indices,
    const void *values,
    size_t nnz,
    size_t domain_size,
    size_t min_idx,
    size_t max_idx,
    const void *scalars,
    size_t nscalars,
    size_t chunks,
    void *out_g,
    void *out_h)
{
    return kc_msm_g1_(curve, indices, values, nnz, domain_size, min_idx, max_idx, scalars, nscalars, chunks, out_g, out_h);
}
API int ref_kc_msm(
    int curve,
    const uint64_t *indices,
    const void *values,
    size_t nnz,
    size_t domain_size,
    size_t min_idx,
    size_t max_idx,
    const void *scalars,
    size_t nscalars,
    size_t chunks,
    void *out_g2,
    void *out_g1)
{
    return kc_msm_(curve, indices, values, nnz, domain_size, min_idx, max_idx, scalars, nscalars, chunks, out_g2, out_g1);
}

static int fft_(int curve, int op, void *data, size_t m, const void *g, int via_factory)
{
    DISPATCH(curve, HH::fft(op, data, m, g, via_factory));
}
API int ref_fft(int curve, int op, void *data, size_t m, const void *g, int via_factory)
{
    return fft_(curve, op, data, m, g, via_factory);
}

static int poly_eval_(int curve, const void *coeffs, size_t m, const uint64_t *idx, size_t k, void *out)
{
    DISPATCH(curve, HH::poly_eval_at_domain(coeffs, m, idx, k, out));
}
API int ref_poly_eval_at_domain(int curve, const void *coeffs, size_t m, const uint64_t *idx, size_t k, void *out)
{
    return poly_eval_(curve, coeffs, m, idx, k, out);
}

static int poly_eval_factory_(int curve, const void *coeffs, size_t m, const uint64_t *idx, size_t k, void *out)
{
    DISPATCH(curve, HH::poly_eval_at_factory_domain(coeffs, m, idx, k, out));
}
API int ref_poly_eval_at_factory_domain(int curve, const void *coeffs, size_t m, const uint64_t *idx, size_t k, void *out)
{
    return poly_eval_factory_(curve, coeffs, m, idx, k, out);
}
static int set_two_adicity_(int curve, int s)
{
    DISPATCH(curve, HH::set_two_adicity(s));
}
API int ref_set_two_adicity(int curve, int s)
{
    return set_two_adicity_(curve, s);
}

struct fixture_handle {
    int curve;
    void *p;
};

API void *ref_groth16_setup(int curve, size_t num_constraints, size_t num_inputs, int real_keys)
{
    void *p = nullptr;
    switch (curve) {
    case 0: p = H<libff::alt_bn128_pp>::setup(num_constraints, num_inputs, real_keys); break;
    case 1: p = H<libff::bls12_377_pp>::setup(num_constraints, num_inputs, real_keys); break;
    case 2: p = H<libff::bw6_761_pp>::setup(num_constraints, num_inputs, real_keys); break;
    default: return nullptr;
    }
    return new fixture_handle{curve, p};
}

API void ref_groth16_free(void *h_)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    if (!h) {
        return;
    }
    switch (h->curve) {
    case 0: delete static_cast<groth16_fixture<libff::alt_bn128_pp> *>(h->p); break;
    case 1: delete static_cast<groth16_fixture<libff::bls12_377_pp> *>(h->p); break;
    case 2: delete static_cast<groth16_fixture<libff::bw6_761_pp> *>(h->p); break;
    }
    delete h;
}

#define FIX(HH) static_cast<const groth16_fixture<typename HH::ppT_> *>(h->p)

#define DISPATCH_FIXTURE(h, fn, ...)                                                                                   \
    switch (h->curve) {                                                                                                \
    case 0:                                                                                                            \
        return H<libff::alt_bn128_pp>::fn(                                                                             \
            static_cast<const groth16_fixture<libff::alt_bn128_pp> *>(h->p), ##__VA_ARGS__);                           \
    case 1:                                                                                                            \
        return H<libff::bls12_377_pp>::fn(                                                                             \
            static_cast<const groth16_fixture<libff::bls12_377_pp> *>(h->p), ##__VA_ARGS__);                           \
    case 2:                                                                                                            \
        return H<libff::bw6_761_pp>::fn(static_cast<const groth16_fixture<libff::bw6_761_pp> *>(h->p), ##__VA_ARGS__); \
    }

API int ref_groth16_sizes(void *h_, uint64_t *out)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    switch (h->curve) {
    case 0: H<libff::alt_bn128_pp>::sizes(static_cast<const groth16_fixture<libff::alt_bn128_pp> *>(h->p), out); return 0;
    case 1: H<libff::bls12_377_pp>::sizes(static_cast<const groth16_fixture<libff::bls12_377_pp> *>(h->p), out); return 0;
    case 2: H<libff::bw6_761_pp>::sizes(static_cast<const groth16_fixture<libff::bw6_761_pp> *>(h->p), out); return 0;
    }
    return -100;
}

API const void *ref_groth16_ptr(void *h_, int which)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, ptr, which);
    return nullptr;
}

API int ref_groth16_witness_map_zk(void *h_, const void *d123, void *out_H, double *ms)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, witness_map_zk, d123, out_H, ms);
    return -100;
}

API int ref_groth16_sap_witness_map(void *h_, const void *d12, void *out_H, size_t cap, size_t *degree, void *out_extra,
                                    double *ms)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, sap_witness_map, d12, out_H, cap, degree, out_extra, ms);
    return -100;
}

API int ref_groth16_witness_map(void *h_, void *out_H, double *ms)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, witness_map, out_H, ms);
    return -100;
}

API int ref_groth16_abc_evaluations(void *h_, void *oA, void *oB, void *oC)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, abc_evaluations, oA, oB, oC);
    return -100;
}

API int ref_groth16_prove(void *h_, const void *r, const void *s, void *out, double *stage_ms)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, prove, r, s, out, stage_ms);
    return -100;
}

API int ref_groth16_prove_fused(void *h_, const void *r, const void *s, void *out, double *ms)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, prove_fused, r, s, out, ms);
    return -100;
}

API int ref_groth16_r1cs_csr(void *h_, uint32_t *row_ptr, uint32_t *col, void *val, uint64_t *nnz)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, r1cs_csr, row_ptr, col, val, nnz);
    return -100;
}

API int ref_groth16_prove_stock(void *h_, void *out, double *ms)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, prove_stock, out, ms);
    return -100;
}

API int ref_groth16_rekey_in_place(void *h_, long mutate_row)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, rekey_in_place, mutate_row);
    return -100;
}

// overlay build: every cache lookup hashes the whole vector / constraint system (libsnark_b200::set_cache_fingerprint_full)
API int ref_overlay_set_fingerprint_full(int full)
{
#ifdef B200_OVERLAY
    libsnark_b200::set_cache_fingerprint_full(full != 0);
    return 0;
#else
    (void)full;
    return -7;
#endif
}

API int ref_groth16_verify(void *h_, const void *proof)
{
    fixture_handle *h = static_cast<fixture_handle *>(h_);
    DISPATCH_FIXTURE(h, verify, proof);
    return -100;
}

// ---- the sibling provers: r1cs_ppzksnark (PGHR13) and r1cs_se_ppzksnark ------------------
// They call the same operators as the Groth16 prover (r1cs_ppzksnark.tcc:570-628: kc_multi_exp_with_mixed_addition
// over <G1,G1> and <G2,G1>, multi_exp, multi_exp_filter_one_zero; r1cs_se_ppzksnark.tcc:500-554: multi_exp over G1
// and G2; both witness maps go through evaluation_domain), instantiated here with Method = BDLO12_signed and
// BaseForm = special — the instantiation the overlay specialises. Generator, prover (the reference's own entry
// point: its d1, d2, d3 / r are drawn inside, so proofs are not reproducible) and verifier, unmodified; a proof
// travels between the plain and the overlay build as the bytes of the reference's own operator<<.
template<typename ppT> struct pghr13_scheme {
    typedef r1cs_ppzksnark_keypair<ppT> keypair;
    typedef r1cs_ppzksnark_proof<ppT> proof;
    static keypair generate(const r1cs_constraint_system<Fr<ppT>> &cs)
    {
        return r1cs_ppzksnark_generator<ppT, libff::multi_exp_base_form_special>(cs);
    }
    static proof prove(const keypair &kp, const r1cs_example<Fr<ppT>> &ex)
    {
        return r1cs_ppzksnark_prover<ppT, libff::multi_exp_method_BDLO12_signed, libff::multi_exp_base_form_special>(
            kp.pk, ex.primary_input, ex.auxiliary_input);
    }
    static bool verify(const keypair &kp, const r1cs_example<Fr<ppT>> &ex, const proof &pi)
    {
        return r1cs_ppzksnark_verifier_strong_IC<ppT>(kp.vk, ex.primary_input, pi);
    }
};
template<typename ppT> struct se_scheme {
    typedef r1cs_se_ppzksnark_keypair<ppT> keypair;
    typedef r1cs_se_ppzksnark_proof<ppT> proof;
    static keypair generate(const r1cs_constraint_system<Fr<ppT>> &cs)
    {
        return r1cs_se_ppzksnark_generator<ppT, libff::multi_exp_base_form_special>(cs);
    }
    static proof prove(const keypair &kp, const r1cs_example<Fr<ppT>> &ex)
    {
        return r1cs_se_ppzksnark_prover<ppT, libff::multi_exp_method_BDLO12_signed, libff::multi_exp_base_form_special>(
            kp.pk, ex.primary_input, ex.auxiliary_input);
    }
    static bool verify(const keypair &kp, const r1cs_example<Fr<ppT>> &ex, const proof &pi)
    {
        return r1cs_se_ppzksnark_verifier_strong_IC<ppT>(kp.vk, ex.primary_input, pi);
    }
};

struct sibling_base {
    virtual ~sibling_base() {}
    virtual int prove(void *out, size_t cap, size_t *len, double *ms) const = 0;
    virtual int verify(const void *proof, size_t len) const = 0;
};
template<typename ppT, typename Scheme> struct sibling_fixture : sibling_base {
    r1cs_example<Fr<ppT>> example;
    typename Scheme::keypair keypair;
    sibling_fixture(r1cs_example<Fr<ppT>> &&e) : example(std::move(e)), keypair(Scheme::generate(example.constraint_system)) {}
    int prove(void *out, size_t cap, size_t *len, double *ms) const override
    {
        const double t0 = omp_get_wtime();
        const typename Scheme::proof pi = Scheme::prove(keypair, example);
        if (ms) {
            *ms = (omp_get_wtime() - t0) * 1e3;
        }
        std::ostringstream ss;
        ss << pi;
        const std::string b = ss.str();
        *len = b.size();
        if (b.size() > cap) {
            return -5;
        }
        memcpy(out, b.data(), b.size());
        return 0;
    }
    int verify(const void *proof, size_t len) const override
    {
        std::istringstream ss(std::string(static_cast<const char *>(proof), len));
        typename Scheme::proof pi;
        ss >> pi;
        return Scheme::verify(keypair, example, pi) ? 1 : 0;
    }
};
template<typename ppT> static sibling_base *sibling_make(int scheme, size_t nc, size_t ni)
{
    ppT::init_public_params();
    r1cs_example<Fr<ppT>> ex = generate_r1cs_example_with_field_input<Fr<ppT>>(nc, ni);
    if (scheme == 0) {
        return new sibling_fixture<ppT, pghr13_scheme<ppT>>(std::move(ex));
    }
    return new sibling_fixture<ppT, se_scheme<ppT>>(std::move(ex));
}

// scheme: 0 = r1cs_ppzksnark (PGHR13), 1 = r1cs_se_ppzksnark; curve: 0 alt_bn128, 1 bls12_377
API void *ref_sibling_setup(int scheme, int curve, size_t num_constraints, size_t num_inputs)
{
    if (scheme < 0 || scheme > 1) {
        return nullptr;
    }
    libff::inhibit_profiling_info = true;
    libff::inhibit_profiling_counters = true;
    switch (curve) {
    case 0: return sibling_make<libff::alt_bn128_pp>(scheme, num_constraints, num_inputs);
    case 1: return sibling_make<libff::bls12_377_pp>(scheme, num_constraints, num_inputs);
    }
    return nullptr;
}
// the fixture may have been built by the other build of this harness (plain / overlay): the objects are layout-identical
// and the virtual call below would run the OTHER library's code, so both entry points re-dispatch on (scheme, curve)
template<typename ppT, typename Scheme> static const sibling_fixture<ppT, Scheme> *sib(const void *h)
{
    return static_cast<const sibling_fixture<ppT, Scheme> *>(static_cast<const sibling_base *>(h));
}
#define SIBLING_DISPATCH(call)                                                                                         \
    switch (scheme * 2 + curve) {                                                                                      \
    case 0: return sib<libff::alt_bn128_pp, pghr13_scheme<libff::alt_bn128_pp>>(h)->sibling_fixture::call;             \
    case 1: return sib<libff::bls12_377_pp, pghr13_scheme<libff::bls12_377_pp>>(h)->sibling_fixture::call;             \
    case 2: return sib<libff::alt_bn128_pp, se_scheme<libff::alt_bn128_pp>>(h)->sibling_fixture::call;                 \
    case 3: return sib<libff::bls12_377_pp, se_scheme<libff::bls12_377_pp>>(h)->sibling_fixture::call;                 \
    }                                                                                                                  \
    return -100;
API int ref_sibling_prove(int scheme, int curve, const void *h, void *out, size_t cap, size_t *len, double *ms)
{
    SIBLING_DISPATCH(prove(out, cap, len, ms))
}
API int ref_sibling_verify(int scheme, int curve, const void *h, const void *proof, size_t len)
{
    SIBLING_DISPATCH(verify(proof, len))
}
API void ref_sibling_free(void *h)
{
    delete static_cast<sibling_base *>(h);
}
