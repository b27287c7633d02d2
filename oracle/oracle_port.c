/* oracle/oracle_port.c — TEST INFRASTRUCTURE, not product code.
 *
 * A plain-C restatement of the reference's CPU algorithms for the Groth16 proving hot path
 * (clearmatics/libsnark + libff + libfqfft), used as the parity checker where the unmodified
 * reference build (oracle/_ref) is not available, and pinned against it by tests/test_oracle_cpu.py
 * through the golden vectors in tests/golden/ (generated FROM the reference by tools/gen_golden.py).
 * Only tests/, bench.py's cpu_baseline leg and __graft_entry__.smoke() may load this library.
 *
 * What is restated (file:line under /root/reference/depends/...):
 *   Fp_model  : Montgomery mul_reduce  libff/algebra/fields/fp.tcc:51-228 (generic CIOS branch :204-227),
 *               add :350-465, sub :467-547, negation :616-630, as_bigint :270-281, inverse :679-727
 *               (value only: a^(p-2) instead of mpn_gcdext)
 *   Fp2_model : libff/algebra/fields/fp2.tcc:102-151 (Karatsuba mul, complex squaring), inverse :153-167
 *   G1/G2     : Jacobian add / mixed_add / dbl for a = 0  libff/algebra/curves/alt_bn128/alt_bn128_g1.cpp:149-326
 *               (bw6_761 uses projective formulas in the reference; results are compared normalised)
 *   digits    : field_get_digit / field_get_signed_digit  libff/algebra/fields/field_utils.tcc:50-115
 *   MSM       : multi_exp_inner<BDLO12_signed>, signed_digits_round, multiexp_accumulate_buckets
 *               libff/algebra/scalar_multiplication/multiexp.tcc:45-125, :507-641; filter_one_zero :690-757
 *   kc MSM    : libsnark/knowledge_commitment/kc_multiexp.tcc:31-113
 *   FFT       : _basic_serial_radix2_FFT, _multiply_by_coset
 *               libfqfft/evaluation_domain/domains/basic_radix2_domain_aux.tcc:44-79, :174-183
 *               FFT / iFFT / cosetFFT / icosetFFT / divide_by_Z_on_coset  basic_radix2_domain.tcc:54-126
 *   step FFT  : step_radix2_domain FFT / iFFT / cosetFFT / icosetFFT / divide_by_Z_on_coset
 *               libfqfft/evaluation_domain/domains/step_radix2_domain.tcc:20-279 (sizes 2^k + 2^r)
 *   ext FFT   : extended_radix2_domain FFT / iFFT / cosetFFT / icosetFFT / divide_by_Z_on_coset
 *               libfqfft/evaluation_domain/domains/extended_radix2_domain.tcc:21-236 (size 2^(s+1); s as the factory sees it
 *               is settable, port_set_two_adicity, the way libff's mutable FieldT::s is: pinned to vectors the reference
 *               produced with s lowered, tests/golden/extended.npz)
 *   R1CS eval : aA, aB, aC incl. the input-consistency rows  libsnark/reductions/r1cs_to_qap/r1cs_to_qap.tcc:236-248,
 *               :293-296; linear_combination::evaluate  libsnark/relations/variable.tcc:288-299
 *   H chain   : libsnark/reductions/r1cs_to_qap/r1cs_to_qap.tcc:252-330 (d1 = d2 = d3 = 0), over the three domain kinds;
 *               with the ZK patch (:262-271, :327-333) and add_poly_Z of each kind: port_witness_map_h_zk (tests/golden/zk.npz)
 *   SAP map   : r1cs_to_sap_witness_map  libsnark/reductions/r1cs_to_sap/r1cs_to_sap.tcc:297-481 (port_sap_witness_map;
 *               the witness map of r1cs_se_ppzksnark; tests/golden/zk.npz)
 *   key gen   : get_window_table / windowed_exp / batch_exp / batch_exp_with_coeff / batch_to_special
 *               libff/algebra/scalar_multiplication/multiexp.tcc:808-974
 *   prover    : the orchestration r1cs_gg_ppzksnark.tcc:460-603 is composed from these in oracle/port.py::groth16_prove
 *               and pinned to whole proofs of the reference (tests/golden/groth16.npz)
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "oracle_params.h" /* generated from the reference's constants by tools/gen_golden.py */

#define API __attribute__((visibility("default")))
#define MAXL 12 /* u64 limbs: bw6_761 Fq */

typedef unsigned __int128 u128;

/* ------------------------------------------------------------------ Fp ------------------------ */

typedef struct {
    uint64_t v[MAXL];
} fp_t;

static const field_params *g_f; /* current field (set per call; the port is single-threaded) */

static int fp_is_zero(const fp_t *a)
{
    uint64_t acc = 0;
    for (int i = 0; i < g_f->n; ++i) acc |= a->v[i];
    return acc == 0;
}
static int fp_eq(const fp_t *a, const fp_t *b) { return memcmp(a->v, b->v, 8 * g_f->n) == 0; }
static void fp_zero(fp_t *r) { memset(r, 0, sizeof(*r)); }
static void fp_one(fp_t *r)
{
    fp_zero(r);
    memcpy(r->v, g_f->one, 8 * g_f->n);
}
static int ge_mod(const uint64_t *x)
{
    for (int i = g_f->n - 1; i >= 0; --i) {
        if (x[i] != g_f->mod[i]) return x[i] > g_f->mod[i];
    }
    return 1;
}
static void sub_mod(uint64_t *x)
{
    u128 borrow = 0;
    for (int i = 0; i < g_f->n; ++i) {
        u128 d = (u128)x[i] - g_f->mod[i] - (uint64_t)borrow;
        x[i] = (uint64_t)d;
        borrow = (d >> 64) & 1;
    }
}
/* fp.tcc:350-465: r = a + b, subtract the modulus if the sum is >= it */
static void fp_add(fp_t *r, const fp_t *a, const fp_t *b)
{
    u128 c = 0;
    fp_t t;
    fp_zero(&t);
    for (int i = 0; i < g_f->n; ++i) {
        c += (u128)a->v[i] + b->v[i];
        t.v[i] = (uint64_t)c;
        c >>= 64;
    }
    if (c || ge_mod(t.v)) sub_mod(t.v);
    *r = t;
}
/* fp.tcc:467-547: r = a - b, add the modulus back on borrow */
static void fp_sub(fp_t *r, const fp_t *a, const fp_t *b)
{
    u128 borrow = 0;
    fp_t t;
    fp_zero(&t);
    for (int i = 0; i < g_f->n; ++i) {
        u128 d = (u128)a->v[i] - b->v[i] - (uint64_t)borrow;
        t.v[i] = (uint64_t)d;
        borrow = (d >> 64) & 1;
    }
    if (borrow) {
        u128 c = 0;
        for (int i = 0; i < g_f->n; ++i) {
            c += (u128)t.v[i] + g_f->mod[i];
            t.v[i] = (uint64_t)c;
            c >>= 64;
        }
    }
    *r = t;
}
/* fp.tcc:616-630 */
static void fp_neg(fp_t *r, const fp_t *a)
{
    if (fp_is_zero(a)) {
        *r = *a;
        return;
    }
    fp_t m;
    fp_zero(&m);
    memcpy(m.v, g_f->mod, 8 * g_f->n);
    fp_sub(r, &m, a);
}
/* fp.tcc:204-227 (generic branch): full product, then n rounds of k = inv * res[i]; res += k * mod << 64 i */
static void fp_mul(fp_t *r, const fp_t *a, const fp_t *b)
{
    const int n = g_f->n;
    uint64_t res[2 * MAXL + 1];
    memset(res, 0, sizeof(res));
    for (int i = 0; i < n; ++i) {
        u128 carry = 0;
        for (int j = 0; j < n; ++j) {
            u128 cur = (u128)a->v[j] * b->v[i] + res[i + j] + (uint64_t)carry;
            res[i + j] = (uint64_t)cur;
            carry = cur >> 64;
        }
        res[i + n] = (uint64_t)carry;
    }
    for (int i = 0; i < n; ++i) {
        const uint64_t k = g_f->inv * res[i];
        u128 carry = 0;
        for (int j = 0; j < n; ++j) {
            u128 cur = (u128)k * g_f->mod[j] + res[i + j] + (uint64_t)carry;
            res[i + j] = (uint64_t)cur;
            carry = cur >> 64;
        }
        for (int j = i + n; carry && j <= 2 * n; ++j) { /* mpn_add_1 */
            u128 cur = (u128)res[j] + (uint64_t)carry;
            res[j] = (uint64_t)cur;
            carry = cur >> 64;
        }
    }
    fp_t t;
    fp_zero(&t);
    memcpy(t.v, res + n, 8 * n);
    if (res[2 * n] || ge_mod(t.v)) sub_mod(t.v);
    *r = t;
}
static void fp_sqr(fp_t *r, const fp_t *a) { fp_mul(r, a, a); }
/* fp.tcc:270-281: Montgomery -> standard form = mul_reduce by 1 */
static void fp_from_mont(fp_t *r, const fp_t *a)
{
    fp_t one;
    fp_zero(&one);
    one.v[0] = 1;
    fp_mul(r, a, &one);
}
/* fp.tcc:230-235: standard -> Montgomery = mul_reduce by R^2 */
static void fp_to_mont(fp_t *r, const fp_t *a)
{
    fp_t r2;
    fp_zero(&r2);
    memcpy(r2.v, g_f->r2, 8 * g_f->n);
    fp_mul(r, a, &r2);
}
/* value of fp.tcc:679-727 (the reference uses mpn_gcdext): a^(p-2) */
static void fp_inv(fp_t *r, const fp_t *a)
{
    uint64_t e[MAXL];
    memcpy(e, g_f->mod, 8 * g_f->n);
    e[0] -= 2; /* p is odd and > 2: no borrow */
    fp_t acc;
    fp_one(&acc);
    for (int i = 64 * g_f->n - 1; i >= 0; --i) {
        fp_sqr(&acc, &acc);
        if ((e[i >> 6] >> (i & 63)) & 1) fp_mul(&acc, &acc, a);
    }
    *r = acc;
}
static void fp_pow_u64(fp_t *r, const fp_t *a, uint64_t e)
{
    fp_t acc, base = *a;
    fp_one(&acc);
    while (e) {
        if (e & 1) fp_mul(&acc, &acc, &base);
        fp_sqr(&base, &base);
        e >>= 1;
    }
    *r = acc;
}

/* ------------------------------------------------------------------ coordinate field: Fq or Fq2 */

typedef struct {
    fp_t c[2];
} el_t;

static int g_deg;      /* 1: Fq, 2: Fq2 */
static int g_nr_small; /* Fq2 non-residue = -g_nr_small (alt_bn128: 1, bls12_377: 5) */

static void fp_mul_by_nr(fp_t *r, const fp_t *a)
{
    fp_t acc = *a, t;
    for (int i = 1; i < g_nr_small; ++i) fp_add(&acc, &acc, a);
    fp_neg(&t, &acc);
    *r = t;
}
static int el_is_zero(const el_t *a) { return fp_is_zero(&a->c[0]) && (g_deg == 1 || fp_is_zero(&a->c[1])); }
static int el_eq(const el_t *a, const el_t *b) { return fp_eq(&a->c[0], &b->c[0]) && (g_deg == 1 || fp_eq(&a->c[1], &b->c[1])); }
static void el_zero(el_t *r) { memset(r, 0, sizeof(*r)); }
static void el_one(el_t *r)
{
    el_zero(r);
    fp_one(&r->c[0]);
}
static void el_add(el_t *r, const el_t *a, const el_t *b)
{
    for (int i = 0; i < g_deg; ++i) fp_add(&r->c[i], &a->c[i], &b->c[i]);
}
static void el_sub(el_t *r, const el_t *a, const el_t *b)
{
    for (int i = 0; i < g_deg; ++i) fp_sub(&r->c[i], &a->c[i], &b->c[i]);
}
static void el_neg(el_t *r, const el_t *a)
{
    for (int i = 0; i < g_deg; ++i) fp_neg(&r->c[i], &a->c[i]);
}
/* fp2.tcc:102-114 (Karatsuba) */
static void el_mul(el_t *r, const el_t *a, const el_t *b)
{
    if (g_deg == 1) {
        fp_mul(&r->c[0], &a->c[0], &b->c[0]);
        return;
    }
    fp_t aA, bB, sa, sb, s, t;
    fp_mul(&aA, &a->c[0], &b->c[0]);
    fp_mul(&bB, &a->c[1], &b->c[1]);
    fp_add(&sa, &a->c[0], &a->c[1]);
    fp_add(&sb, &b->c[0], &b->c[1]);
    fp_mul(&s, &sa, &sb);
    fp_mul_by_nr(&t, &bB);
    el_t o;
    fp_add(&o.c[0], &aA, &t);
    fp_sub(&s, &s, &aA);
    fp_sub(&o.c[1], &s, &bB);
    *r = o;
}
static void el_sqr(el_t *r, const el_t *a) { el_mul(r, a, a); } /* same value as fp2.tcc:142-151 */
/* fp2.tcc:153-167 */
static void el_inv(el_t *r, const el_t *a)
{
    if (g_deg == 1) {
        fp_inv(&r->c[0], &a->c[0]);
        return;
    }
    fp_t t0, t1, t2, t3;
    fp_sqr(&t0, &a->c[0]);
    fp_sqr(&t1, &a->c[1]);
    fp_mul_by_nr(&t2, &t1);
    fp_sub(&t2, &t0, &t2);
    fp_inv(&t3, &t2);
    el_t o;
    fp_mul(&o.c[0], &a->c[0], &t3);
    fp_mul(&t0, &a->c[1], &t3);
    fp_neg(&o.c[1], &t0);
    *r = o;
}

/* ------------------------------------------------------------------ group (Jacobian, a = 0) ---- */

typedef struct {
    el_t X, Y, Z;
} pt_t;

static int pt_is_zero(const pt_t *p) { return el_is_zero(&p->Z); }
static void pt_zero(pt_t *p) /* alt_bn128_g1.cpp:86-91: (0, 1, 0) */
{
    el_zero(&p->X);
    el_one(&p->Y);
    el_zero(&p->Z);
}
/* dbl-2009-l, alt_bn128_g1.cpp:285-326 */
static void pt_dbl(pt_t *r, const pt_t *p)
{
    if (pt_is_zero(p)) {
        *r = *p;
        return;
    }
    el_t A, B, C, D, E, F, t, X3, Y3, Z3, eightC;
    el_sqr(&A, &p->X);
    el_sqr(&B, &p->Y);
    el_sqr(&C, &B);
    el_add(&t, &p->X, &B);
    el_sqr(&t, &t);
    el_sub(&t, &t, &A);
    el_sub(&t, &t, &C);
    el_add(&D, &t, &t);
    el_add(&E, &A, &A);
    el_add(&E, &E, &A);
    el_sqr(&F, &E);
    el_sub(&X3, &F, &D);
    el_sub(&X3, &X3, &D);
    el_add(&eightC, &C, &C);
    el_add(&eightC, &eightC, &eightC);
    el_add(&eightC, &eightC, &eightC);
    el_sub(&t, &D, &X3);
    el_mul(&Y3, &E, &t);
    el_sub(&Y3, &Y3, &eightC);
    el_mul(&Z3, &p->Y, &p->Z);
    el_add(&Z3, &Z3, &Z3);
    r->X = X3;
    r->Y = Y3;
    r->Z = Z3;
}
/* add-2007-bl with the reference's special cases, alt_bn128_g1.cpp:97-206 */
static void pt_add(pt_t *r, const pt_t *a, const pt_t *b)
{
    if (pt_is_zero(a)) {
        *r = *b;
        return;
    }
    if (pt_is_zero(b)) {
        *r = *a;
        return;
    }
    el_t Z1Z1, Z2Z2, U1, U2, S1, S2, H, I, J, rr, V, t, X3, Y3, Z3;
    el_sqr(&Z1Z1, &a->Z);
    el_sqr(&Z2Z2, &b->Z);
    el_mul(&U1, &a->X, &Z2Z2);
    el_mul(&U2, &b->X, &Z1Z1);
    el_mul(&t, &b->Z, &Z2Z2);
    el_mul(&S1, &a->Y, &t);
    el_mul(&t, &a->Z, &Z1Z1);
    el_mul(&S2, &b->Y, &t);
    if (el_eq(&U1, &U2) && el_eq(&S1, &S2)) {
        pt_dbl(r, a);
        return;
    }
    el_sub(&H, &U2, &U1);
    el_add(&t, &H, &H);
    el_sqr(&I, &t);
    el_mul(&J, &H, &I);
    el_sub(&rr, &S2, &S1);
    el_add(&rr, &rr, &rr);
    el_mul(&V, &U1, &I);
    el_sqr(&X3, &rr);
    el_sub(&X3, &X3, &J);
    el_sub(&X3, &X3, &V);
    el_sub(&X3, &X3, &V);
    el_sub(&t, &V, &X3);
    el_mul(&Y3, &rr, &t);
    el_mul(&t, &S1, &J);
    el_add(&t, &t, &t);
    el_sub(&Y3, &Y3, &t);
    el_add(&t, &a->Z, &b->Z);
    el_sqr(&t, &t);
    el_sub(&t, &t, &Z1Z1);
    el_sub(&t, &t, &Z2Z2);
    el_mul(&Z3, &t, &H);
    r->X = X3;
    r->Y = Y3;
    r->Z = Z3;
}
/* mixed_add (b in special form), alt_bn128_g1.cpp:208-283: same value as add */
static void pt_mixed_add(pt_t *r, const pt_t *a, const pt_t *b) { pt_add(r, a, b); }
static void pt_neg(pt_t *r, const pt_t *a)
{
    *r = *a;
    el_neg(&r->Y, &a->Y);
}
/* to_affine_coordinates, alt_bn128_g1.cpp:68-82 */
static void pt_to_affine(pt_t *p)
{
    if (pt_is_zero(p)) {
        pt_zero(p);
        return;
    }
    el_t zi, z2, z3;
    el_inv(&zi, &p->Z);
    el_sqr(&z2, &zi);
    el_mul(&z3, &z2, &zi);
    el_mul(&p->X, &p->X, &z2);
    el_mul(&p->Y, &p->Y, &z3);
    el_one(&p->Z);
}

/* reference layout <-> pt_t: X | Y | Z, each g_deg * n limbs */
static void pt_load(pt_t *p, const uint64_t *src)
{
    const int n = g_f->n;
    memset(p, 0, sizeof(*p));
    el_t *co[3] = {&p->X, &p->Y, &p->Z};
    for (int k = 0; k < 3; ++k)
        for (int d = 0; d < g_deg; ++d) memcpy(co[k]->c[d].v, src + (size_t)(k * g_deg + d) * n, 8 * n);
}
static void pt_store(uint64_t *dst, const pt_t *p)
{
    const int n = g_f->n;
    const el_t *co[3] = {&p->X, &p->Y, &p->Z};
    for (int k = 0; k < 3; ++k)
        for (int d = 0; d < g_deg; ++d) memcpy(dst + (size_t)(k * g_deg + d) * n, co[k]->c[d].v, 8 * n);
}

/* ------------------------------------------------------------------ digits --------------------- */

/* libff/common/utils.cpp:32-44: ceil(log2(n)) */
static size_t ff_log2(size_t n)
{
    size_t r = ((n & (n - 1)) == 0 ? 0 : 1);
    while (n > 1) {
        n >>= 1;
        r++;
    }
    return r;
}
/* multiexp.tcc:35-40 and :637-641 */
static size_t bdlo12_signed_optimal_c(size_t n)
{
    const size_t l = ff_log2(n);
    return l - (l / 3 - 2) + 1;
}
/* field_utils.tcc:50-79 (v: nl u64 limbs, standard form) */
static size_t field_get_digit(const uint64_t *v, int nl, size_t digit_size, size_t digit_idx)
{
    const size_t start_bit = digit_size * digit_idx, end_bit = start_bit + digit_size;
    const size_t low_limb = start_bit / 64, high_limb = end_bit / 64;
    const int use_high = (high_limb < (size_t)nl) && (high_limb != low_limb);
    const size_t low_shift = start_bit - low_limb * 64;
    const size_t high_bits = use_high ? (end_bit - high_limb * 64) : 0;
    const size_t high_shift = digit_size - high_bits;
    const uint64_t mask = (1ull << digit_size) - 1;
    const uint64_t lo = low_limb < (size_t)nl ? v[low_limb] >> low_shift : 0;
    const uint64_t hi = use_high ? (v[high_limb] << high_shift) : 0;
    return (size_t)(mask & (hi | lo));
}
/* field_utils.tcc:82-115 */
static int64_t field_get_signed_digit(const uint64_t *v, int nl, size_t digit_size, size_t digit_index)
{
    const uint64_t carry_mask = 1ull << (digit_size - 1);
    const int64_t overflow_mask = 1ll << digit_size;
    uint64_t carry = 0, digit = 0;
    int64_t overflow = 0;
    size_t i = 0;
    do {
        carry = (uint64_t)overflow | carry;
        digit = field_get_digit(v, nl, digit_size, i) + carry;
        overflow = (int64_t)((digit & (uint64_t)overflow_mask) >> digit_size);
        carry = (digit & carry_mask) >> (digit_size - 1);
        ++i;
    } while (i <= digit_index);
    return (1 - overflow) * ((int64_t)digit - (int64_t)carry * overflow_mask);
}
static size_t bigint_num_bits(const uint64_t *v, int nl)
{
    for (int i = nl - 1; i >= 0; --i) {
        if (v[i]) return (size_t)(64 * i + 64 - __builtin_clzll(v[i]));
    }
    return 0;
}

/* ------------------------------------------------------------------ MSM ------------------------ */

/* multiexp.tcc:507-633 with BaseForm = special (bases have Z = one or are zero()). */
static void multi_exp_inner(pt_t *result, const pt_t *bases, const uint64_t *scalars_std, int nr, size_t n)
{
    const size_t c = bdlo12_signed_optimal_c(n);
    size_t num_bits = 0;
    for (size_t i = 0; i < n; ++i) {
        const size_t b = bigint_num_bits(scalars_std + i * nr, nr);
        if (b > num_bits) num_bits = b;
    }
    const size_t num_rounds = (num_bits + 2 + c - 1) / c;
    const size_t num_buckets = (size_t)1 << (c - 1);
    pt_t *buckets = (pt_t *)malloc(num_buckets * sizeof(pt_t));
    unsigned char *hit = (unsigned char *)malloc(num_buckets);
    pt_zero(result);
    for (size_t round = 0; round < num_rounds; ++round) {
        const size_t digit_idx = num_rounds - 1 - round;
        if (round > 0)
            for (size_t i = 0; i < c; ++i) pt_dbl(result, result);
        /* signed_digits_round, :519-561 */
        memset(hit, 0, num_buckets);
        size_t non_zero = 0;
        for (size_t i = 0; i < n; ++i) {
            const int64_t d = field_get_signed_digit(scalars_std + i * nr, nr, c, digit_idx);
            if (d == 0) continue;
            /* multi_exp_add_element_to_bucket_with_signed_digit, :45-81 */
            pt_t b = bases[i];
            size_t idx;
            if (d < 0) {
                pt_neg(&b, &bases[i]);
                idx = (size_t)(-d - 1);
            } else {
                idx = (size_t)(d - 1);
            }
            if (hit[idx]) {
                pt_mixed_add(&buckets[idx], &buckets[idx], &b);
            } else {
                buckets[idx] = b;
                hit[idx] = 1;
            }
            ++non_zero;
        }
        if (non_zero == 0) continue; /* round result is zero() */
        /* multiexp_accumulate_buckets, :91-125 */
        pt_t sum, acc;
        pt_zero(&sum);
        pt_zero(&acc);
        for (size_t k = num_buckets; k-- > 0;) {
            if (hit[k]) pt_add(&acc, &acc, &buckets[k]);
            pt_add(&sum, &sum, &acc);
        }
        pt_add(result, result, &sum);
    }
    free(buckets);
    free(hit);
}

static int set_context(int curve, int group)
{
    if (curve < 0 || curve > 2 || (group != 1 && group != 2)) return -1;
    g_f = &curve_table[curve].fq;
    g_deg = (group == 2 && curve_table[curve].g2_over_fq2) ? 2 : 1;
    g_nr_small = curve_table[curve].nr_small;
    return 0;
}

/* scalars: Montgomery Fr -> standard bigints (multiexp.tcc:577-582) */
static uint64_t *scalars_to_std(int curve, const uint64_t *scalars, size_t n)
{
    const field_params *fq = g_f;
    g_f = &curve_table[curve].fr;
    const int nr = g_f->n;
    uint64_t *out = (uint64_t *)malloc((n ? n : 1) * nr * 8);
    for (size_t i = 0; i < n; ++i) {
        fp_t a, s;
        fp_zero(&a);
        memcpy(a.v, scalars + i * nr, 8 * nr);
        fp_from_mont(&s, &a);
        memcpy(out + i * nr, s.v, 8 * nr);
    }
    g_f = fq;
    return out;
}

/* method 0: multi_exp (chunks = 1); method 2: multi_exp_filter_one_zero (multiexp.tcc:690-757).
 * bases: n points, `stride` bytes apart, reference layout, special form. out: normalised point. */
API int port_msm(int curve, int group, int method, const void *bases, size_t stride, size_t n, const void *scalars,
                 void *out)
{
    if (set_context(curve, group)) return -1;
    const int nr = curve_table[curve].fr.n;
    pt_t *pts = (pt_t *)malloc((n ? n : 1) * sizeof(pt_t));
    uint64_t *std = scalars_to_std(curve, (const uint64_t *)scalars, n);
    pt_t acc, result;
    pt_zero(&acc);
    size_t m = 0;
    uint64_t one_std[MAXL] = {1};
    for (size_t i = 0; i < n; ++i) {
        pt_t p;
        pt_load(&p, (const uint64_t *)((const char *)bases + i * stride));
        const uint64_t *s = std + i * nr;
        if (method == 2) {
            if (bigint_num_bits(s, nr) == 0) continue; /* scalar 0: skipped */
            if (memcmp(s, one_std, 8 * nr) == 0) {      /* scalar 1: acc = acc.mixed_add(base) */
                pt_mixed_add(&acc, &acc, &p);
                continue;
            }
        }
        pts[m] = p;
        memmove(std + m * nr, s, 8 * nr);
        ++m;
    }
    if (m)
        multi_exp_inner(&result, pts, std, nr, m);
    else
        pt_zero(&result);
    pt_add(&result, &acc, &result);
    pt_to_affine(&result);
    pt_store((uint64_t *)out, &result);
    free(pts);
    free(std);
    return 0;
}

/* ---- key generation: fixed-base batch exponentiation ------------------------------------------------
 * get_window_table (multiexp.tcc:808-844): powers_of_g[outer][inner] = inner * 2^(window * outer) * g;
 * windowed_exp (:846-876): res = sum_outer powers_of_g[outer][bits of pow in window outer];
 * batch_exp (:879-914) / batch_exp_with_coeff (:916-946): res[i] = windowed_exp(coeff * v[i]);
 * batch_to_special (:949-974): every result normalised, zero -> zero().to_special() = (0, 1, 0).
 * g: one point (reference layout); coeff: Fr (Montgomery) or NULL; window: 1..16 (the reference picks it from a
 * per-curve table, get_exp_window_size :771-806 — it does not change the result). out: n points, special form. */
API int port_batch_exp(int curve, int group, const void *g_, const void *coeff, const void *scalars, size_t n,
                       size_t window, void *out)
{
    if (set_context(curve, group) || window < 1 || window > 16) return -1;
    const int nr = curve_table[curve].fr.n;
    const size_t scalar_size = bigint_num_bits(curve_table[curve].fr.mod, nr); /* Fr::size_in_bits() */
    const size_t in_window = (size_t)1 << window;
    const size_t outerc = (scalar_size + window - 1) / window;
    const size_t last_in_window = (size_t)1 << (scalar_size - (outerc - 1) * window);
    pt_t *table = (pt_t *)malloc(outerc * in_window * sizeof(pt_t));
    pt_t gouter;
    pt_load(&gouter, (const uint64_t *)g_);
    for (size_t outer = 0; outer < outerc; ++outer) { /* :830-841 */
        pt_t ginner;
        pt_zero(&ginner);
        const size_t cur = outer == outerc - 1 ? last_in_window : in_window;
        for (size_t inner = 0; inner < in_window; ++inner) {
            if (inner < cur) {
                table[outer * in_window + inner] = ginner;
                pt_add(&ginner, &ginner, &gouter);
            } else {
                pt_zero(&table[outer * in_window + inner]);
            }
        }
        for (size_t i = 0; i < window; ++i) pt_dbl(&gouter, &gouter);
    }
    /* coeff * v[i] in Fr (Montgomery product, fp.tcc:51-228), then as_bigint */
    uint64_t *eff = (uint64_t *)malloc((n ? n : 1) * nr * 8);
    {
        const field_params *fq = g_f;
        g_f = &curve_table[curve].fr;
        for (size_t i = 0; i < n; ++i) {
            fp_t a, c, p;
            fp_zero(&a);
            memcpy(a.v, (const uint64_t *)scalars + i * nr, 8 * nr);
            if (coeff) {
                fp_zero(&c);
                memcpy(c.v, coeff, 8 * nr);
                fp_mul(&p, &c, &a);
                a = p;
            }
            memcpy(eff + i * nr, a.v, 8 * nr);
        }
        g_f = fq;
    }
    uint64_t *std = scalars_to_std(curve, eff, n);
    const size_t glimbs = (size_t)3 * g_deg * g_f->n;
    for (size_t i = 0; i < n; ++i) {
        const uint64_t *pw = std + i * nr;
        pt_t res = table[0]; /* powers_of_g[0][0] = zero() */
        for (size_t outer = 0; outer < outerc; ++outer) { /* :859-871 */
            size_t inner = 0;
            for (size_t k = 0; k < window; ++k) {
                const size_t bit = outer * window + k;
                if (bit < (size_t)64 * nr && ((pw[bit >> 6] >> (bit & 63)) & 1)) inner |= (size_t)1 << k;
            }
            pt_add(&res, &res, &table[outer * in_window + inner]);
        }
        if (pt_is_zero(&res)) {
            pt_zero(&res);
        } else {
            pt_to_affine(&res);
        }
        pt_store((uint64_t *)out + i * glimbs, &res);
    }
    free(table);
    free(eff);
    free(std);
    return 0;
}

/* kc_multiexp.tcc:31-113: sparse vector of (G2, G1) pairs; entries with min_idx <= index < max_idx. */
API int port_kc_msm(int curve, const uint64_t *indices, const void *values, size_t nnz, size_t min_idx, size_t max_idx,
                    const void *scalars, void *out_g2, void *out_g1)
{
    if (curve < 0 || curve > 2) return -1;
    const int nr = curve_table[curve].fr.n, nq = curve_table[curve].fq.n;
    const size_t g2b = (size_t)3 * nq * 8 * (curve_table[curve].g2_over_fq2 ? 2 : 1), g1b = (size_t)3 * nq * 8;
    const size_t stride = g2b + g1b;
    size_t first = 0;
    while (first < nnz && indices[first] < min_idx) ++first;
    size_t cnt = 0;
    while (first + cnt < nnz && indices[first + cnt] < max_idx) ++cnt;
    uint64_t *sc = (uint64_t *)malloc((cnt ? cnt : 1) * nr * 8);
    for (size_t k = 0; k < cnt; ++k)
        memcpy(sc + k * nr, (const uint64_t *)scalars + (indices[first + k] - min_idx) * nr, 8 * nr);
    const char *v = (const char *)values + first * stride;
    int rc = port_msm(curve, 2, 2, v, stride, cnt, sc, out_g2);
    if (!rc) rc = port_msm(curve, 1, 2, v + g2b, stride, cnt, sc, out_g1);
    free(sc);
    return rc;
}

API int port_signed_digits(int curve, const void *scalars, size_t n, size_t c, size_t num_digits, int64_t *out)
{
    if (curve < 0 || curve > 2) return -1;
    g_f = &curve_table[curve].fr;
    const int nr = g_f->n;
    uint64_t *std = scalars_to_std(curve, (const uint64_t *)scalars, n);
    for (size_t i = 0; i < n; ++i)
        for (size_t d = 0; d < num_digits; ++d) out[i * num_digits + d] = field_get_signed_digit(std + i * nr, nr, c, d);
    free(std);
    return 0;
}
API uint64_t port_bdlo12_signed_optimal_c(uint64_t n) { return bdlo12_signed_optimal_c(n); }

/* ------------------------------------------------------------------ FFT ------------------------ */

static size_t bitreverse(size_t n, size_t l) /* libff/common/utils.cpp:59-67 */
{
    size_t r = 0;
    for (size_t k = 0; k < l; ++k) {
        r = (r << 1) | (n & 1);
        n >>= 1;
    }
    return r;
}
/* FieldT::s as the domain factory sees it: libff keeps s and root_of_unity as mutable static parameters; a host that
 * lowers them consistently (port_set_two_adicity) gets extended_radix2_domain at 2^(s+1). The root of a given order is
 * the same either way, so it is always derived from the field's own pair. */
static int g_two_adicity[3]; /* 0 = the field's own */
static size_t two_adicity(void)
{
    for (int c = 0; c < 3; ++c)
        if (g_f == &curve_table[c].fr && g_two_adicity[c]) return (size_t)g_two_adicity[c];
    return (size_t)g_f->s;
}
API int port_set_two_adicity(int curve, int s)
{
    if (curve < 0 || curve > 2 || s < 0 || s > curve_table[curve].fr.s) return -1;
    g_two_adicity[curve] = s;
    return 0;
}
/* field_utils.tcc:190-206: root_of_unity^(2^(s - log n)) */
static int get_root_of_unity(fp_t *w, size_t n)
{
    const size_t logn = ff_log2(n);
    if (n != ((size_t)1 << logn) || logn > two_adicity()) return -1;
    fp_zero(w);
    memcpy(w->v, g_f->root_of_unity, 8 * g_f->n);
    for (size_t i = g_f->s; i > logn; --i) fp_sqr(w, w);
    return 0;
}
/* basic_radix2_domain_aux.tcc:44-79 */
static void serial_radix2_fft(fp_t *a, size_t n, const fp_t *omega)
{
    const size_t logn = ff_log2(n);
    for (size_t k = 0; k < n; ++k) {
        const size_t rk = bitreverse(k, logn);
        if (k < rk) {
            fp_t t = a[k];
            a[k] = a[rk];
            a[rk] = t;
        }
    }
    size_t m = 1;
    for (size_t s = 1; s <= logn; ++s) {
        fp_t w_m;
        fp_pow_u64(&w_m, omega, n / (2 * m));
        for (size_t k = 0; k < n; k += 2 * m) {
            fp_t w;
            fp_one(&w);
            for (size_t j = 0; j < m; ++j) {
                fp_t t;
                fp_mul(&t, &w, &a[k + j + m]);
                fp_sub(&a[k + j + m], &a[k + j], &t);
                fp_add(&a[k + j], &a[k + j], &t);
                fp_mul(&w, &w, &w_m);
            }
        }
        m *= 2;
    }
}
/* basic_radix2_domain_aux.tcc:174-183 */
static void multiply_by_coset(fp_t *a, size_t n, const fp_t *g)
{
    fp_t u = *g;
    for (size_t i = 1; i < n; ++i) {
        fp_mul(&a[i], &a[i], &u);
        fp_mul(&u, &u, g);
    }
}
static int fft_inplace(fp_t *a, size_t m, int op, const fp_t *g)
{
    fp_t omega, t;
    if (get_root_of_unity(&omega, m)) return -2; /* DomainSizeException */
    switch (op) {
    case 0: serial_radix2_fft(a, m, &omega); return 0;
    case 1: { /* basic_radix2_domain.tcc:63-74 */
        fp_inv(&t, &omega);
        serial_radix2_fft(a, m, &t);
        fp_t mm, sconst;
        fp_zero(&mm);
        mm.v[0] = m;
        fp_to_mont(&mm, &mm);
        fp_inv(&sconst, &mm);
        for (size_t i = 0; i < m; ++i) fp_mul(&a[i], &a[i], &sconst);
        return 0;
    }
    case 2: /* :77-81 */
        multiply_by_coset(a, m, g);
        return fft_inplace(a, m, 0, g);
    case 3: /* :84-88 */
        fft_inplace(a, m, 1, g);
        fp_inv(&t, g);
        multiply_by_coset(a, m, &t);
        return 0;
    case 4: { /* divide_by_Z_on_coset :117-126, coset = multiplicative_generator */
        fp_t gen, z, one;
        fp_zero(&gen);
        memcpy(gen.v, g_f->generator, 8 * g_f->n);
        fp_pow_u64(&z, &gen, m);
        fp_one(&one);
        fp_sub(&z, &z, &one);
        fp_inv(&z, &z);
        for (size_t i = 0; i < m; ++i) fp_mul(&a[i], &a[i], &z);
        return 0;
    }
    }
    return -1;
}
/* ---- step_radix2_domain (libfqfft/evaluation_domain/domains/step_radix2_domain.tcc): m = big + small, big = 2^k >
 * small = 2^r — what get_evaluation_domain returns for sizes that are not a power of two
 * (get_evaluation_domain.tcc:41-58). Same op codes as fft_inplace. */
static void fp_from_u64(fp_t *r, uint64_t v)
{
    fp_zero(r);
    r->v[0] = v;
    fp_to_mont(r, r);
}
static int step_shape(size_t m, size_t *big, size_t *small) /* valid_for_size, :20-45 */
{
    if (m <= 1 || !(m & (m - 1))) return -2;
    size_t l = 0;
    while (((size_t)1 << l) < m) ++l; /* libff::log2 = ceil */
    *big = (size_t)1 << (l - 1);
    *small = m - *big;
    if (*small & (*small - 1)) return -2;
    if (l > two_adicity()) return -2;
    return 0;
}
static int step_inplace(fp_t *a, size_t m, int op, const fp_t *g)
{
    size_t big, small;
    if (step_shape(m, &big, &small)) return -2; /* DomainSizeException */
    fp_t omega, big_omega, small_omega, t;
    if (get_root_of_unity(&omega, 2 * big) || get_root_of_unity(&small_omega, small)) return -2; /* :57-61 */
    fp_sqr(&big_omega, &omega);
    const size_t compr = big / small;
    switch (op) {
    case 0: { /* FFT, :65-104 */
        fp_t *c = (fp_t *)calloc(big, sizeof(fp_t)), *d = (fp_t *)calloc(big, sizeof(fp_t)),
             *e = (fp_t *)calloc(small, sizeof(fp_t));
        fp_t omega_i;
        fp_one(&omega_i);
        for (size_t i = 0; i < big; ++i) {
            if (i < small) {
                fp_add(&c[i], &a[i], &a[i + big]);
                fp_sub(&t, &a[i], &a[i + big]);
            } else {
                c[i] = a[i];
                t = a[i];
            }
            fp_mul(&d[i], &omega_i, &t);
            fp_mul(&omega_i, &omega_i, &omega);
        }
        for (size_t i = 0; i < small; ++i)
            for (size_t j = 0; j < compr; ++j) fp_add(&e[i], &e[i], &d[i + j * small]);
        serial_radix2_fft(c, big, &big_omega);
        serial_radix2_fft(e, small, &small_omega);
        memcpy(a, c, big * sizeof(fp_t));
        memcpy(a + big, e, small * sizeof(fp_t));
        free(c);
        free(d);
        free(e);
        return 0;
    }
    case 1: { /* iFFT, :106-171 */
        fp_t *U0 = (fp_t *)malloc(big * sizeof(fp_t)), *U1 = (fp_t *)malloc(small * sizeof(fp_t)),
             *tmp = (fp_t *)malloc(big * sizeof(fp_t));
        memcpy(U0, a, big * sizeof(fp_t));
        memcpy(U1, a + big, small * sizeof(fp_t));
        fp_inv(&t, &big_omega);
        serial_radix2_fft(U0, big, &t);
        fp_inv(&t, &small_omega);
        serial_radix2_fft(U1, small, &t);
        fp_t inv;
        fp_from_u64(&t, big);
        fp_inv(&inv, &t);
        for (size_t i = 0; i < big; ++i) fp_mul(&U0[i], &U0[i], &inv);
        fp_from_u64(&t, small);
        fp_inv(&inv, &t);
        for (size_t i = 0; i < small; ++i) fp_mul(&U1[i], &U1[i], &inv);
        fp_t omega_i;
        fp_one(&omega_i);
        for (size_t i = 0; i < big; ++i) {
            fp_mul(&tmp[i], &U0[i], &omega_i);
            fp_mul(&omega_i, &omega_i, &omega);
        }
        for (size_t i = small; i < big; ++i) a[i] = U0[i]; /* A_suffix */
        for (size_t i = 0; i < small; ++i)
            for (size_t j = 1; j < compr; ++j) fp_sub(&U1[i], &U1[i], &tmp[i + j * small]);
        fp_t omega_inv, omega_inv_i, over_two;
        fp_inv(&omega_inv, &omega);
        fp_one(&omega_inv_i);
        for (size_t i = 0; i < small; ++i) {
            fp_mul(&U1[i], &U1[i], &omega_inv_i);
            fp_mul(&omega_inv_i, &omega_inv_i, &omega_inv);
        }
        fp_from_u64(&t, 2);
        fp_inv(&over_two, &t);
        for (size_t i = 0; i < small; ++i) {
            fp_add(&t, &U0[i], &U1[i]);
            fp_mul(&a[i], &t, &over_two);
            fp_sub(&t, &U0[i], &U1[i]);
            fp_mul(&a[big + i], &t, &over_two);
        }
        free(U0);
        free(U1);
        free(tmp);
        return 0;
    }
    case 2: /* cosetFFT, :173-178 */
        multiply_by_coset(a, m, g);
        return step_inplace(a, m, 0, g);
    case 3: /* icosetFFT, :180-185 */
        step_inplace(a, m, 1, g);
        fp_inv(&t, g);
        multiply_by_coset(a, m, &t);
        return 0;
    case 4: { /* divide_by_Z_on_coset, :249-279, coset = multiplicative_generator */
        fp_t coset, one, Z0, cs_z0, os_z0, o2s, elt, u, v, w;
        fp_zero(&coset);
        memcpy(coset.v, g_f->generator, 8 * g_f->n);
        fp_one(&one);
        fp_pow_u64(&Z0, &coset, big);
        fp_sub(&Z0, &Z0, &one);
        fp_pow_u64(&u, &coset, small);
        fp_mul(&cs_z0, &u, &Z0);
        fp_pow_u64(&u, &omega, small);
        fp_mul(&os_z0, &u, &Z0);
        fp_pow_u64(&o2s, &omega, 2 * small);
        fp_one(&elt);
        for (size_t i = 0; i < big; ++i) {
            fp_mul(&u, &cs_z0, &elt);
            fp_sub(&u, &u, &os_z0);
            fp_inv(&v, &u);
            fp_mul(&a[i], &a[i], &v);
            fp_mul(&elt, &elt, &o2s);
        }
        fp_mul(&w, &coset, &omega);
        fp_pow_u64(&u, &w, big);
        fp_sub(&u, &u, &one);
        fp_pow_u64(&v, &w, small);
        fp_pow_u64(&t, &omega, small);
        fp_sub(&v, &v, &t);
        fp_mul(&u, &u, &v);
        fp_inv(&v, &u);
        for (size_t i = 0; i < small; ++i) fp_mul(&a[big + i], &a[big + i], &v);
        return 0;
    }
    }
    return -1;
}
/* ---- extended_radix2_domain (libfqfft/evaluation_domain/domains/extended_radix2_domain.tcc): m = 2 small,
 * small = 2^s — the order-2^s subgroup and its coset by shift = coset_shift() = multiplicative_generator^2
 * (field_utils.tcc:152-155). Same op codes as fft_inplace. */
static int ext_inplace(fp_t *a, size_t m, int op, const fp_t *g)
{
    const size_t logm = ff_log2(m);
    if (m <= 1 || m != ((size_t)1 << logm) || logm != two_adicity() + 1) return -2; /* ctor, :47-62 */
    const size_t small = m / 2;
    fp_t omega, shift, gen, ssm, one, t, u;
    if (get_root_of_unity(&omega, small)) return -2;
    fp_zero(&gen);
    memcpy(gen.v, g_f->generator, 8 * g_f->n);
    fp_sqr(&shift, &gen);
    fp_pow_u64(&ssm, &shift, small);
    fp_one(&one);
    switch (op) {
    case 0: { /* FFT, :64-91 */
        fp_t *a0 = (fp_t *)malloc(small * sizeof(fp_t)), *a1 = (fp_t *)malloc(small * sizeof(fp_t));
        fp_t shift_i;
        fp_one(&shift_i);
        for (size_t i = 0; i < small; ++i) {
            fp_add(&a0[i], &a[i], &a[small + i]);
            fp_mul(&t, &ssm, &a[small + i]);
            fp_add(&t, &a[i], &t);
            fp_mul(&a1[i], &shift_i, &t);
            fp_mul(&shift_i, &shift_i, &shift);
        }
        serial_radix2_fft(a0, small, &omega);
        serial_radix2_fft(a1, small, &omega);
        memcpy(a, a0, small * sizeof(fp_t));
        memcpy(a + small, a1, small * sizeof(fp_t));
        free(a0);
        free(a1);
        return 0;
    }
    case 1: { /* iFFT, :93-120 */
        fp_t *a0 = (fp_t *)malloc(small * sizeof(fp_t)), *a1 = (fp_t *)malloc(small * sizeof(fp_t));
        memcpy(a0, a, small * sizeof(fp_t));
        memcpy(a1, a + small, small * sizeof(fp_t));
        fp_t omega_inv, sconst, shift_inv, shift_inv_i;
        fp_inv(&omega_inv, &omega);
        serial_radix2_fft(a0, small, &omega_inv);
        serial_radix2_fft(a1, small, &omega_inv);
        fp_from_u64(&t, small);
        fp_sub(&u, &one, &ssm);
        fp_mul(&t, &t, &u);
        fp_inv(&sconst, &t);
        fp_inv(&shift_inv, &shift);
        fp_one(&shift_inv_i);
        for (size_t i = 0; i < small; ++i) {
            fp_mul(&u, &shift_inv_i, &a1[i]);  /* shift^-i a1[i] */
            fp_mul(&t, &ssm, &a0[i]);
            fp_sub(&t, &u, &t);
            fp_mul(&a[i], &sconst, &t);
            fp_sub(&t, &a0[i], &u);
            fp_mul(&a[i + small], &sconst, &t);
            fp_mul(&shift_inv_i, &shift_inv_i, &shift_inv);
        }
        free(a0);
        free(a1);
        return 0;
    }
    case 2: /* cosetFFT, :122-127 */
        multiply_by_coset(a, m, g);
        return ext_inplace(a, m, 0, g);
    case 3: /* icosetFFT, :129-134 */
        ext_inplace(a, m, 1, g);
        fp_inv(&t, g);
        multiply_by_coset(a, m, &t);
        return 0;
    case 4: { /* divide_by_Z_on_coset, :216-236, coset = multiplicative_generator */
        fp_t csm, z0, z1, v;
        fp_pow_u64(&csm, &gen, small);
        fp_sub(&t, &csm, &one);
        fp_sub(&u, &csm, &ssm);
        fp_mul(&z0, &t, &u);
        fp_mul(&v, &csm, &ssm);
        fp_sub(&t, &v, &one);
        fp_sub(&u, &v, &ssm);
        fp_mul(&z1, &t, &u);
        fp_inv(&z0, &z0);
        fp_inv(&z1, &z1);
        for (size_t i = 0; i < small; ++i) {
            fp_mul(&a[i], &a[i], &z0);
            fp_mul(&a[i + small], &a[i + small], &z1);
        }
        return 0;
    }
    }
    return -1;
}
/* get_evaluation_domain (get_evaluation_domain.tcc:32-70) for the sizes these fields meet: basic radix-2 for powers
 * of two up to 2^s, extended radix-2 for 2^(s+1), step radix-2 for 2^k + 2^r. */
static int domain_inplace(fp_t *a, size_t m, int op, const fp_t *g)
{
    if (m & (m - 1)) return step_inplace(a, m, op, g);
    if (m > 1 && ff_log2(m) == two_adicity() + 1) return ext_inplace(a, m, op, g);
    return fft_inplace(a, m, op, g);
}

static fp_t *load_frs(const uint64_t *src, size_t m)
{
    fp_t *a = (fp_t *)calloc(m ? m : 1, sizeof(fp_t));
    for (size_t i = 0; i < m; ++i) memcpy(a[i].v, src + i * g_f->n, 8 * g_f->n);
    return a;
}
static void store_frs(uint64_t *dst, const fp_t *a, size_t m)
{
    for (size_t i = 0; i < m; ++i) memcpy(dst + i * g_f->n, a[i].v, 8 * g_f->n);
}

/* op: 0 FFT, 1 iFFT, 2 cosetFFT(g), 3 icosetFFT(g), 4 divide_by_Z_on_coset; in place on m Fr elements. */
API int port_fft(int curve, int op, void *data, size_t m, const void *g)
{
    if (curve < 0 || curve > 2) return -1;
    g_f = &curve_table[curve].fr;
    if (m < 1) return -2;
    fp_t gg;
    fp_zero(&gg);
    memcpy(gg.v, g ? g : (const void *)g_f->generator, 8 * g_f->n);
    fp_t *a = load_frs((const uint64_t *)data, m);
    const int rc = domain_inplace(a, m, op, &gg);
    if (!rc) store_frs((uint64_t *)data, a, m);
    free(a);
    return rc;
}

/* The evaluations r1cs_to_qap_witness_map builds before its first transform (r1cs_to_qap.tcc:236-248 for aA, aB;
 * :293-296 for aC; linear_combination::evaluate, variable.tcc:288-299). The constraint system comes as one CSR:
 * rows [0, nc) = the a-combinations, [nc, 2nc) = b, [2nc, 3nc) = c; col = linear_term::index (0 = the constant ONE);
 * w = (1 | primary | auxiliary), so that full_variable_assignment[index - 1] = w[index]. */
API int port_r1cs_evaluate(int curve, size_t nc, size_t ni, const uint32_t *row_ptr, const uint32_t *col,
                           const void *val, const void *w_, size_t m, void *aA, void *aB, void *aC)
{
    if (curve < 0 || curve > 2 || m < nc + ni + 1) return -1;
    g_f = &curve_table[curve].fr;
    const int nr = g_f->n;
    const uint64_t *w = (const uint64_t *)w_, *v = (const uint64_t *)val;
    uint64_t *out[3] = {(uint64_t *)aA, (uint64_t *)aB, (uint64_t *)aC};
    for (int k = 0; k < 3; ++k) memset(out[k], 0, m * nr * 8);
    /* the additional constraints input_i * 0 = 0 (:240-243): aA[nc + i] = (i > 0 ? full[i - 1] : one) = w[i] */
    for (size_t i = 0; i <= ni; ++i) memcpy(out[0] + (nc + i) * nr, w + i * nr, 8 * nr);
    for (int k = 0; k < 3; ++k) {
        for (size_t i = 0; i < nc; ++i) {
            fp_t acc;
            fp_zero(&acc);
            for (uint32_t e = row_ptr[k * nc + i]; e < row_ptr[k * nc + i + 1]; ++e) {
                fp_t x, c, p;
                fp_zero(&x);
                fp_zero(&c);
                memcpy(x.v, w + (size_t)col[e] * nr, 8 * nr); /* index 0 -> w[0] = one */
                memcpy(c.v, v + (size_t)e * nr, 8 * nr);
                fp_mul(&p, &x, &c);
                fp_add(&acc, &acc, &p);
            }
            memcpy(out[k] + i * nr, acc.v, 8 * nr);
        }
    }
    return 0;
}

/* r1cs_to_qap.tcc:252-330 with d1 = d2 = d3 = 0: aA, aB, aC (m each) -> coefficients_for_H (m + 1). */
/* evaluation_domain::add_poly_Z of the kind that serves m: basic_radix2_domain.tcc:109-115,
 * extended_radix2_domain.tcc:203-214, step_radix2_domain.tcc:235-245. H has m + 1 elements. */
static int add_poly_z(fp_t *H, size_t m, const fp_t *coeff)
{
    fp_t t, u, one;
    fp_one(&one);
    if (!(m & (m - 1))) {
        if (m > 1 && ff_log2(m) == two_adicity() + 1) {
            fp_t gen, shift, ssm;
            const size_t small = m / 2;
            fp_zero(&gen);
            memcpy(gen.v, g_f->generator, 8 * g_f->n);
            fp_sqr(&shift, &gen);
            fp_pow_u64(&ssm, &shift, small);
            fp_add(&H[m], &H[m], coeff);
            fp_add(&t, &ssm, &one);
            fp_mul(&u, coeff, &t);
            fp_sub(&H[small], &H[small], &u);
            fp_mul(&u, coeff, &ssm);
            fp_add(&H[0], &H[0], &u);
            return 0;
        }
        fp_add(&H[m], &H[m], coeff);
        fp_sub(&H[0], &H[0], coeff);
        return 0;
    }
    size_t big, small;
    if (step_shape(m, &big, &small)) return -2;
    fp_t omega, wsm;
    if (get_root_of_unity(&omega, 2 * big)) return -2;
    fp_pow_u64(&wsm, &omega, small);
    fp_mul(&u, coeff, &wsm);
    fp_add(&H[m], &H[m], coeff);
    fp_sub(&H[big], &H[big], &u);
    fp_sub(&H[small], &H[small], coeff);
    fp_add(&H[0], &H[0], &u);
    return 0;
}

/* r1cs_to_sap_witness_map (libsnark/reductions/r1cs_to_sap/r1cs_to_sap.tcc:297-481; the witness map of
 * r1cs_se_ppzksnark), in the reference's order. ea, eb, ec: the values <a_i, w>, <b_i, w>, <c_i, w> of the nc constraints
 * (port_r1cs_evaluate rows [0, nc)); w = (1 | primary | auxiliary); d12 = d1, d2. out_H: m + 1 coefficients with
 * m = the SAP domain size (get_evaluation_domain(2 nc + 2 ni + 1), passed in); out_extra: the nc + ni extra variables. */
API int port_sap_witness_map(int curve, size_t nc, size_t ni, size_t nv, const void *ea_, const void *eb_, const void *ec_,
                             const void *w_, size_t m, const void *d12, void *out_H, void *out_extra)
{
    if (curve < 0 || curve > 2 || m < 2 * nc + 2 * ni + 1) return -1;
    g_f = &curve_table[curve].fr;
    const int nr = g_f->n;
    fp_t gen, d[2], one, t, u;
    fp_zero(&gen);
    memcpy(gen.v, g_f->generator, 8 * nr);
    fp_one(&one);
    for (int k = 0; k < 2; ++k) {
        fp_zero(&d[k]);
        memcpy(d[k].v, (const uint64_t *)d12 + k * nr, 8 * nr);
    }
    fp_t *ea = load_frs((const uint64_t *)ea_, nc), *eb = load_frs((const uint64_t *)eb_, nc), *ec = load_frs((const uint64_t *)ec_, nc);
    fp_t *w = load_frs((const uint64_t *)w_, nv + 1); /* w[i] = full_variable_assignment[i - 1] */
    fp_t *extra = (fp_t *)calloc(nc + ni + 1, sizeof(fp_t));
    for (size_t i = 0; i < nc; ++i) { /* :330-341: (a - b)^2 */
        fp_sub(&t, &ea[i], &eb[i]);
        fp_sqr(&extra[i], &t);
    }
    for (size_t i = 1; i <= ni; ++i) { /* :342-351: (x_i - 1)^2 */
        fp_sub(&t, &w[i], &one);
        fp_sqr(&extra[nc + i - 1], &t);
    }
    fp_t *A = (fp_t *)calloc(m, sizeof(fp_t)), *Cc = (fp_t *)calloc(m, sizeof(fp_t)), *H = (fp_t *)calloc(m + 1, sizeof(fp_t));
    const size_t off = 2 * nc;
    for (size_t i = 0; i < nc; ++i) { /* :357-363, :414-420 */
        fp_add(&A[2 * i], &ea[i], &eb[i]);
        fp_sub(&A[2 * i + 1], &ea[i], &eb[i]);
        fp_add(&t, &ec[i], &ec[i]);
        fp_add(&t, &t, &t); /* times_four */
        fp_add(&Cc[2 * i], &t, &extra[i]);
        Cc[2 * i + 1] = extra[i];
    }
    A[off] = one; /* :367, :423 */
    Cc[off] = one;
    for (size_t i = 1; i <= ni; ++i) { /* :369-375, :425-433 */
        fp_add(&A[off + 2 * i - 1], &w[i], &one);
        fp_sub(&A[off + 2 * i], &w[i], &one);
        fp_add(&t, &w[i], &w[i]);
        fp_add(&t, &t, &t);
        fp_add(&Cc[off + 2 * i - 1], &t, &extra[nc + i - 1]);
        Cc[off + 2 * i] = extra[nc + i - 1];
    }
    int rc = domain_inplace(A, m, 1, &gen); /* :380 */
    if (!rc) {
        for (size_t i = 0; i < m; ++i) { /* :388-391: 2 d1 A */
            fp_mul(&t, &d[0], &A[i]);
            fp_add(&H[i], &t, &t);
        }
        fp_sub(&H[0], &H[0], &d[1]);
        fp_mul(&t, &d[0], &d[0]);
        rc = add_poly_z(H, m, &t);
    }
    if (!rc) {
        domain_inplace(A, m, 2, &gen);
        for (size_t i = 0; i < m; ++i) fp_sqr(&A[i], &A[i]);
        domain_inplace(Cc, m, 1, &gen);
        domain_inplace(Cc, m, 2, &gen);
        for (size_t i = 0; i < m; ++i) fp_sub(&A[i], &A[i], &Cc[i]);
        domain_inplace(A, m, 4, &gen);
        domain_inplace(A, m, 3, &gen);
        for (size_t i = 0; i < m; ++i) fp_add(&H[i], &H[i], &A[i]);
        store_frs((uint64_t *)out_H, H, m + 1);
        store_frs((uint64_t *)out_extra, extra, nc + ni);
    }
    (void)u;
    free(ea);
    free(eb);
    free(ec);
    free(w);
    free(extra);
    free(A);
    free(Cc);
    free(H);
    return rc;
}

/* r1cs_to_qap_witness_map with its ZK patch, in the reference's order (r1cs_to_qap.tcc:252-333): d123 = d1, d2, d3. */
API int port_witness_map_h_zk(int curve, const void *aA, const void *aB, const void *aC, size_t m, const void *d123,
                              void *out_H)
{
    if (curve < 0 || curve > 2) return -1;
    g_f = &curve_table[curve].fr;
    fp_t gen, d[3], t, u;
    fp_zero(&gen);
    memcpy(gen.v, g_f->generator, 8 * g_f->n);
    for (int k = 0; k < 3; ++k) {
        fp_zero(&d[k]);
        memcpy(d[k].v, (const uint64_t *)d123 + k * g_f->n, 8 * g_f->n);
    }
    fp_t *A = load_frs((const uint64_t *)aA, m), *B = load_frs((const uint64_t *)aB, m), *Cc = load_frs((const uint64_t *)aC, m);
    fp_t *H = (fp_t *)calloc(m + 1, sizeof(fp_t));
    int rc = domain_inplace(A, m, 1, &gen);
    if (!rc) rc = domain_inplace(B, m, 1, &gen);
    if (!rc) {
        for (size_t i = 0; i < m; ++i) { /* :266-268 */
            fp_mul(&t, &d[1], &A[i]);
            fp_mul(&u, &d[0], &B[i]);
            fp_add(&H[i], &t, &u);
        }
        fp_sub(&H[0], &H[0], &d[2]);
        fp_mul(&t, &d[0], &d[1]);
        rc = add_poly_z(H, m, &t);
    }
    if (!rc) {
        domain_inplace(A, m, 2, &gen);
        domain_inplace(B, m, 2, &gen);
        for (size_t i = 0; i < m; ++i) fp_mul(&A[i], &A[i], &B[i]);
        domain_inplace(Cc, m, 1, &gen);
        domain_inplace(Cc, m, 2, &gen);
        for (size_t i = 0; i < m; ++i) fp_sub(&A[i], &A[i], &Cc[i]);
        domain_inplace(A, m, 4, &gen);
        domain_inplace(A, m, 3, &gen);
        for (size_t i = 0; i < m; ++i) fp_add(&H[i], &H[i], &A[i]); /* :327-333 */
        store_frs((uint64_t *)out_H, H, m + 1);
    }
    free(A);
    free(B);
    free(Cc);
    free(H);
    return rc;
}

API int port_witness_map_h(int curve, const void *aA, const void *aB, const void *aC, size_t m, void *out_H)
{
    if (curve < 0 || curve > 2) return -1;
    g_f = &curve_table[curve].fr;
    fp_t gen;
    fp_zero(&gen);
    memcpy(gen.v, g_f->generator, 8 * g_f->n);
    fp_t *A = load_frs((const uint64_t *)aA, m), *B = load_frs((const uint64_t *)aB, m), *Cc = load_frs((const uint64_t *)aC, m);
    int rc = domain_inplace(A, m, 1, &gen);
    if (!rc) rc = domain_inplace(B, m, 1, &gen);
    if (!rc) {
        domain_inplace(A, m, 2, &gen);
        domain_inplace(B, m, 2, &gen);
        for (size_t i = 0; i < m; ++i) fp_mul(&A[i], &A[i], &B[i]);
        domain_inplace(Cc, m, 1, &gen);
        domain_inplace(Cc, m, 2, &gen);
        for (size_t i = 0; i < m; ++i) fp_sub(&A[i], &A[i], &Cc[i]);
        domain_inplace(A, m, 4, &gen);
        domain_inplace(A, m, 3, &gen);
        store_frs((uint64_t *)out_H, A, m);
        memset((uint64_t *)out_H + m * g_f->n, 0, 8 * g_f->n);
    }
    free(A);
    free(B);
    free(Cc);
    return rc;
}

/* ------------------------------------------------------------------ primitive hooks ------------ */

/* field: 0 Fr, 1 Fq, 2 Fq2. op: 0 add, 1 sub, 2 mul, 3 sqr, 4 neg, 5 inverse (0 -> 0) */
API int port_field_op(int curve, int field, int op, const void *a_, const void *b_, void *o_, size_t n)
{
    if (curve < 0 || curve > 2) return -1;
    g_f = field == 0 ? &curve_table[curve].fr : &curve_table[curve].fq;
    g_deg = (field == 2 && curve_table[curve].g2_over_fq2) ? 2 : 1;
    g_nr_small = curve_table[curve].nr_small;
    const int nl = g_f->n;
    for (size_t i = 0; i < n; ++i) {
        el_t a, b, o;
        el_zero(&a);
        el_zero(&b);
        for (int d = 0; d < g_deg; ++d) {
            memcpy(a.c[d].v, (const uint64_t *)a_ + (i * g_deg + d) * nl, 8 * nl);
            memcpy(b.c[d].v, (const uint64_t *)b_ + (i * g_deg + d) * nl, 8 * nl);
        }
        el_zero(&o);
        switch (op) {
        case 0: el_add(&o, &a, &b); break;
        case 1: el_sub(&o, &a, &b); break;
        case 2: el_mul(&o, &a, &b); break;
        case 3: el_sqr(&o, &a); break;
        case 4: el_neg(&o, &a); break;
        case 5:
            if (el_is_zero(&a))
                o = a;
            else
                el_inv(&o, &a);
            break;
        default: return -1;
        }
        for (int d = 0; d < g_deg; ++d) memcpy((uint64_t *)o_ + (i * g_deg + d) * nl, o.c[d].v, 8 * nl);
    }
    return 0;
}

/* op: 0 add, 1 mixed_add, 2 dbl, 3 neg, 4 to_affine; points in the reference layout */
API int port_group_op(int curve, int group, int op, const void *a_, const void *b_, void *o_, size_t n)
{
    if (set_context(curve, group)) return -1;
    const size_t limbs = (size_t)3 * g_deg * g_f->n;
    for (size_t i = 0; i < n; ++i) {
        pt_t a, b, o;
        pt_load(&a, (const uint64_t *)a_ + i * limbs);
        pt_load(&b, (const uint64_t *)b_ + i * limbs);
        switch (op) {
        case 0: pt_add(&o, &a, &b); break;
        case 1: pt_mixed_add(&o, &a, &b); break;
        case 2: pt_dbl(&o, &a); break;
        case 3: pt_neg(&o, &a); break;
        case 4:
            o = a;
            break;
        default: return -1;
        }
        pt_to_affine(&o); /* results are compared normalised */
        pt_store((uint64_t *)o_ + i * limbs, &o);
    }
    return 0;
}
