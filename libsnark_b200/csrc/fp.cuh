// fp.cuh — Montgomery-form prime-field arithmetic on 32-bit limbs.
//
// Replaces libff's Fp_model<n, modulus> (depends/libff/libff/algebra/fields/fp.hpp:38-160,
// mul_reduce fp.tcc:51-228, operator+= :350-465, operator-= :467-547, invert :679-727)
// on the device. Memory layout is bit-identical to libff: bigint<n>'s n little-endian
// u64 limbs (bigint.hpp:33) read as 2n little-endian u32 limbs; values are Montgomery
// residues a*R mod p with R = 2^(64n) = 2^(32N), always fully reduced (< p), exactly as
// Fp_model keeps them. So a device result can be memcmp'ed with the reference's mont_repr.
//
// Multiplication is an operand-scanning Montgomery product on two interleaved
// accumulators ("even"/"odd" 64-bit columns): every 32x32->64 product lands on a
// 64-bit aligned column of one of the two accumulators, so each row is one carry
// chain of mad.lo.cc/madc.hi.cc pairs (-> IMAD.WIDE.U32 + carry in SASS) and no
// product ever needs its own carry fix-up. Cost: 2*N^2 wide multiply-adds + N low
// multiplies (the quotient digits) — the MM(n32) of SURVEY.md §8(d).
#pragma once
#include "prims.cuh"

namespace b200
{

// Host code (the MSM's final O(scalar bits) window fold, host_fold.hpp) uses a plain 64-bit-limb
// CIOS instead of the device carry chains; tests/host_emu defines B200_HOST_EMULATE_PRIMS to run
// the device formulation itself on the CPU through the prims.cuh emulation.
#if !defined(__CUDA_ARCH__) && !defined(B200_HOST_EMULATE_PRIMS)
#define B200_HOST_FAST 1
#else
#define B200_HOST_FAST 0
#endif

// Translation units whose kernels are latency-bound chains of group operations run by few warps
// (bucket reduction, window sums, test hooks) are compiled with -DB200_OUTLINE_MUL: every field
// product then calls one shared out-of-line copy, so the code a warp walks through stays inside
// the instruction cache. (Measured: with inlined products those kernels were instruction-fetch
// bound at ~45 cycles per instruction; profiles/r1_reduce_icache.md.)
#if defined(B200_OUTLINE_MUL)
#define B200_OUTLINE_MUL_ALWAYS 1
#else
#define B200_OUTLINE_MUL_ALWAYS 0
#endif

template<class P> struct alignas(16) Fp {
    static constexpr int N = P::N;
    typedef P params;
    uint32_t v[N];

    HD static Fp zero()
    {
        Fp r;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            r.v[i] = 0;
        }
        return r;
    }
    // Montgomery form of 1 (= R mod p), libff's Fp_model::one().
    HD static Fp one()
    {
        Fp r;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            r.v[i] = P::one(i);
        }
        return r;
    }
    HD static Fp r2()
    {
        Fp r;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            r.v[i] = P::r2(i);
        }
        return r;
    }

    HD bool is_zero() const
    {
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            acc |= v[i];
        }
        return acc == 0;
    }
    HD bool operator==(const Fp &o) const
    {
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            acc |= v[i] ^ o.v[i];
        }
        return acc == 0;
    }
    HD bool operator!=(const Fp &o) const { return !(*this == o); }

    // r = (borrow of x - p) ? x : x - p, for x < 2p.
    HD static void final_sub(uint32_t *x)
    {
        uint32_t t[N];
        t[0] = prim::sub_cc(x[0], P::mod(0));
#pragma unroll
        for (int i = 1; i < N; ++i) {
            t[i] = prim::subc_cc(x[i], P::mod(i));
        }
        const uint32_t borrow = prim::subc(0, 0); // 0 or 0xffffffff
#pragma unroll
        for (int i = 0; i < N; ++i) {
            x[i] = borrow ? x[i] : t[i];
        }
    }

#if B200_HOST_FAST
    static inline uint64_t limb64(const uint32_t *v, int i) { return (uint64_t)v[2 * i] | ((uint64_t)v[2 * i + 1] << 32); }
    static inline uint64_t mod64(int i) { return (uint64_t)P::mod(2 * i) | ((uint64_t)P::mod(2 * i + 1) << 32); }
    static inline void put64(uint32_t *v, int i, uint64_t x)
    {
        v[2 * i] = (uint32_t)x;
        v[2 * i + 1] = (uint32_t)(x >> 32);
    }
    // r = x - p if x >= p (x < 2p given as M limbs)
    static inline void host_final_sub(uint64_t *x)
    {
        constexpr int M = N / 2;
        uint64_t t[M];
        unsigned __int128 borrow = 0;
        for (int i = 0; i < M; ++i) {
            const unsigned __int128 d = (unsigned __int128)x[i] - mod64(i) - (uint64_t)borrow;
            t[i] = (uint64_t)d;
            borrow = (d >> 64) & 1;
        }
        if (!borrow) {
            for (int i = 0; i < M; ++i) {
                x[i] = t[i];
            }
        }
    }
    static inline Fp host_add(const Fp &a, const Fp &b)
    {
        constexpr int M = N / 2;
        uint64_t x[M];
        unsigned __int128 c = 0;
        for (int i = 0; i < M; ++i) {
            c += (unsigned __int128)limb64(a.v, i) + limb64(b.v, i);
            x[i] = (uint64_t)c;
            c >>= 64;
        }
        host_final_sub(x);
        Fp r;
        for (int i = 0; i < M; ++i) {
            put64(r.v, i, x[i]);
        }
        return r;
    }
    static inline Fp host_sub(const Fp &a, const Fp &b)
    {
        constexpr int M = N / 2;
        uint64_t x[M];
        unsigned __int128 borrow = 0;
        for (int i = 0; i < M; ++i) {
            const unsigned __int128 d = (unsigned __int128)limb64(a.v, i) - limb64(b.v, i) - (uint64_t)borrow;
            x[i] = (uint64_t)d;
            borrow = (d >> 64) & 1;
        }
        if (borrow) {
            unsigned __int128 c = 0;
            for (int i = 0; i < M; ++i) {
                c += (unsigned __int128)x[i] + mod64(i);
                x[i] = (uint64_t)c;
                c >>= 64;
            }
        }
        Fp r;
        for (int i = 0; i < M; ++i) {
            put64(r.v, i, x[i]);
        }
        return r;
    }
    static inline Fp host_mul(const Fp &a, const Fp &b)
    {
        constexpr int M = N / 2;
        uint64_t A[M], B[M], Q[M], t[M + 2];
        for (int i = 0; i < M; ++i) {
            A[i] = limb64(a.v, i);
            B[i] = limb64(b.v, i);
            Q[i] = mod64(i);
            t[i] = 0;
        }
        t[M] = t[M + 1] = 0;
        uint64_t y = 1; // p^-1 mod 2^64 by Newton iteration, then negate
        for (int k = 0; k < 6; ++k) {
            y *= 2 - Q[0] * y;
        }
        const uint64_t inv64 = 0 - y;
        for (int i = 0; i < M; ++i) {
            unsigned __int128 cur, carry = 0;
            for (int j = 0; j < M; ++j) {
                cur = (unsigned __int128)A[j] * B[i] + t[j] + (uint64_t)carry;
                t[j] = (uint64_t)cur;
                carry = cur >> 64;
            }
            cur = (unsigned __int128)t[M] + (uint64_t)carry;
            t[M] = (uint64_t)cur;
            t[M + 1] = (uint64_t)(cur >> 64);
            const uint64_t m = t[0] * inv64;
            cur = (unsigned __int128)m * Q[0] + t[0];
            carry = cur >> 64;
            for (int j = 1; j < M; ++j) {
                cur = (unsigned __int128)m * Q[j] + t[j] + (uint64_t)carry;
                t[j - 1] = (uint64_t)cur;
                carry = cur >> 64;
            }
            cur = (unsigned __int128)t[M] + (uint64_t)carry;
            t[M - 1] = (uint64_t)cur;
            t[M] = t[M + 1] + (uint64_t)(cur >> 64);
        }
        host_final_sub(t);
        Fp r;
        for (int i = 0; i < M; ++i) {
            put64(r.v, i, t[i]);
        }
        return r;
    }
#endif

    HD friend Fp operator+(const Fp &a, const Fp &b)
    {
#if B200_HOST_FAST
        return host_add(a, b);
#endif
        Fp r;
        r.v[0] = prim::add_cc(a.v[0], b.v[0]);
#pragma unroll
        for (int i = 1; i < N - 1; ++i) {
            r.v[i] = prim::addc_cc(a.v[i], b.v[i]);
        }
        r.v[N - 1] = prim::addc(a.v[N - 1], b.v[N - 1]); // p has a spare top bit: a+b < 2p < 2^(32N)
        final_sub(r.v);
        return r;
    }

    HD friend Fp operator-(const Fp &a, const Fp &b)
    {
#if B200_HOST_FAST
        return host_sub(a, b);
#endif
        Fp r;
        r.v[0] = prim::sub_cc(a.v[0], b.v[0]);
#pragma unroll
        for (int i = 1; i < N; ++i) {
            r.v[i] = prim::subc_cc(a.v[i], b.v[i]);
        }
        const uint32_t borrow = prim::subc(0, 0); // all-ones iff a < b
        r.v[0] = prim::add_cc(r.v[0], P::mod(0) & borrow);
#pragma unroll
        for (int i = 1; i < N - 1; ++i) {
            r.v[i] = prim::addc_cc(r.v[i], P::mod(i) & borrow);
        }
        r.v[N - 1] = prim::addc(r.v[N - 1], P::mod(N - 1) & borrow);
        return r;
    }

    HD Fp neg() const
    {
        // libff Fp_model::operator- (fp.tcc:616-630): 0 stays 0, else p - a.
        Fp r;
        r.v[0] = prim::sub_cc(P::mod(0), v[0]);
#pragma unroll
        for (int i = 1; i < N - 1; ++i) {
            r.v[i] = prim::subc_cc(P::mod(i), v[i]);
        }
        r.v[N - 1] = prim::subc(P::mod(N - 1), v[N - 1]);
        const bool z = is_zero();
#pragma unroll
        for (int i = 0; i < N; ++i) {
            r.v[i] = z ? 0u : r.v[i];
        }
        return r;
    }
    HD Fp cneg(bool flag) const
    {
        const Fp n = neg();
        Fp r;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            r.v[i] = flag ? n.v[i] : v[i];
        }
        return r;
    }

    HD Fp dbl() const { return *this + *this; }

    // One Montgomery round on the two interleaved accumulators (see header).
    //   ev has weight 2^0 per limb index, od has weight 2^32: V = ev + 2^32 * od.
    // first: ev/od are written; else the previous round's (ev, od) arrive swapped
    // and shifted: new ev = old od (+ old ev[1]), new od = old ev >> 64.
    template<bool FIRST> HD static void mont_round(uint32_t *ev, uint32_t *od, const uint32_t *a, uint32_t bi)
    {
        if (FIRST) {
#pragma unroll
            for (int j = 0; j < N; j += 2) {
                ev[j] = prim::mul_lo(a[j], bi);
                ev[j + 1] = prim::mul_hi(a[j], bi);
                od[j] = prim::mul_lo(a[j + 1], bi);
                od[j + 1] = prim::mul_hi(a[j + 1], bi);
            }
        } else {
            // ev[0] (old od[0]) += old ev[1]; its carry has weight 2^32 = od[0]'s.
            ev[0] = prim::add_cc(ev[0], od[1]);
            // od[k] = old_ev[k+2] + (odd-limb products of a*bi), shifted down by 64 bits.
#pragma unroll
            for (int j = 1; j < N; j += 2) {
                const int k = j - 1;
                const uint32_t in0 = (k + 2 < N) ? od[k + 2] : 0u;
                const uint32_t in1 = (k + 3 < N) ? od[k + 3] : 0u;
                od[k] = prim::madc_lo_cc(a[j], bi, in0);
                if (k + 1 < N - 1) {
                    od[k + 1] = prim::madc_hi_cc(a[j], bi, in1);
                } else {
                    od[k + 1] = prim::madc_hi(a[j], bi, in1);
                }
            }
            ev[0] = prim::mad_lo_cc(a[0], bi, ev[0]);
            ev[1] = prim::madc_hi_cc(a[0], bi, ev[1]);
#pragma unroll
            for (int j = 2; j < N; j += 2) {
                ev[j] = prim::madc_lo_cc(a[j], bi, ev[j]);
                ev[j + 1] = prim::madc_hi_cc(a[j], bi, ev[j + 1]);
            }
            od[N - 1] = prim::addc(od[N - 1], 0);
        }
        const uint32_t m = prim::mul_lo(ev[0], P::INV);
        od[0] = prim::mad_lo_cc(P::mod(1), m, od[0]);
        if (N > 2) {
            od[1] = prim::madc_hi_cc(P::mod(1), m, od[1]);
        } else {
            od[1] = prim::madc_hi(P::mod(1), m, od[1]);
        }
#pragma unroll
        for (int j = 3; j < N; j += 2) {
            od[j - 1] = prim::madc_lo_cc(P::mod(j), m, od[j - 1]);
            if (j < N - 1) {
                od[j] = prim::madc_hi_cc(P::mod(j), m, od[j]);
            } else {
                od[j] = prim::madc_hi(P::mod(j), m, od[j]);
            }
        }
        ev[0] = prim::mad_lo_cc(P::mod(0), m, ev[0]);
        ev[1] = prim::madc_hi_cc(P::mod(0), m, ev[1]);
#pragma unroll
        for (int j = 2; j < N; j += 2) {
            ev[j] = prim::madc_lo_cc(P::mod(j), m, ev[j]);
            ev[j + 1] = prim::madc_hi_cc(P::mod(j), m, ev[j + 1]);
        }
        od[N - 1] = prim::addc(od[N - 1], 0);
    }

    // a * b * R^-1 mod p, fully reduced. (libff: Fp_model::mul_reduce, fp.tcc:51-228.)
    // Fields of >= 16 limbs (bw6_761 Fq: 24) call one shared out-of-line copy: a fully
    // inlined 24-limb product is ~1200 instructions per site.
    HD friend Fp operator*(const Fp &a, const Fp &b)
    {
#if defined(__CUDA_ARCH__)
        if (N >= 16 || B200_OUTLINE_MUL_ALWAYS) {
            return mul_outlined(a, b);
        }
#endif
        return mul_inline(a, b);
    }
#if defined(__CUDACC__)
    // by value: operands and result travel in registers (pointers would force them through local memory)
    static __device__ __noinline__ Fp mul_outlined(Fp a, Fp b) { return mul_inline(a, b); }
#endif
    HD static Fp mul_inline(const Fp &a, const Fp &b)
    {
        static_assert(N % 2 == 0, "even limb count");
#if B200_HOST_FAST
        return host_mul(a, b);
#endif
        uint32_t acc[2][N];
        mont_round<true>(acc[0], acc[1], a.v, b.v[0]);
#pragma unroll
        for (int i = 1; i < N; ++i) {
            mont_round<false>(acc[i & 1], acc[(i & 1) ^ 1], a.v, b.v[i]);
        }
        // after the last round (i = N-1, odd): ev = acc[1], od = acc[0], ev[0] == 0.
        const uint32_t *ev = acc[1];
        const uint32_t *od = acc[0];
        Fp r;
        r.v[0] = prim::add_cc(ev[1], od[0]);
#pragma unroll
        for (int k = 1; k < N - 1; ++k) {
            r.v[k] = prim::addc_cc(ev[k + 1], od[k]);
        }
        r.v[N - 1] = prim::addc(od[N - 1], 0);
        final_sub(r.v);
        return r;
    }

    HD Fp sqr() const { return (*this) * (*this); }

    // Montgomery -> standard representation (libff as_bigint, fp.tcc:270-281): multiply by 1.
    HD Fp from_mont() const
    {
        Fp o = zero();
        o.v[0] = 1;
        return (*this) * o;
    }
    // standard -> Montgomery (libff Fp_model(bigint) ctor, fp.tcc:230-235): multiply by R^2.
    HD Fp to_mont() const { return (*this) * r2(); }

    // a^(p-2) (libff's invert uses mpn_gcdext, fp.tcc:679-727; the value is the same).
    HD Fp inverse() const
    {
        Fp result = one();
        bool started = false;
        for (int i = N - 1; i >= 0; --i) {
            const uint32_t w = P::pm2(i);
            for (int b = 31; b >= 0; --b) {
                if (started) {
                    result = result.sqr();
                }
                if ((w >> b) & 1u) {
                    result = started ? result * (*this) : *this;
                    started = true;
                }
            }
        }
        return result;
    }

    // The same value as inverse() by the binary extended Euclidean algorithm on the residue: subtractions, shifts and
    // selects only, no multiplications but the closing fix-up. It exists for the top of the batched inversion trees of
    // the MSM's affine rounds (msm_affine.cuh), where a handful of warps invert a few thousand values and what counts
    // is the LATENCY of one inversion: the Fermat chain above is bits(p) dependent Montgomery products (~0.2 ms alone
    // on an SM for 254 bits, milliseconds for bw6_761's 761); one step here is a few short carry chains.
    //   u = a, v = p, x1 = 1, x2 = 0;  invariants x1 a == u, x2 a == v (mod p), v odd.
    //   step: u even -> u /= 2, x1 /= 2 (mod p);  u odd -> order so that u >= v, then u = (u - v) / 2, x1 = (x1 - x2) / 2.
    //   every step removes a bit of u or v, so u reaches 0 within 2 bits(p) steps, with v = gcd = 1 and x2 = a^-1.
    // Each step is written branch-free (the lanes of a warp run different inputs). this = aR (Montgomery) -> returns
    // a^-1 R = (aR)^-1 R^2 = montmul((aR)^-1, R^3). Zero maps to zero.
    HD Fp inverse_fast() const
    {
        uint32_t u[N], w[N], x1[N], x2[N];
#pragma unroll
        for (int i = 0; i < N; ++i) {
            u[i] = v[i];
            w[i] = P::mod(i);
            x1[i] = i == 0 ? 1u : 0u;
            x2[i] = 0u;
        }
        for (int step = 0; step < 2 * 32 * N; ++step) {
            uint32_t nz = 0;
#pragma unroll
            for (int i = 0; i < N; ++i) {
                nz |= u[i];
            }
            if (nz == 0) {
                break;
            }
            const uint32_t odd = 0u - (u[0] & 1u);              // all-ones when u is odd
            // borrow of u - w: u < w
            prim::sub_cc(u[0], w[0]);
#pragma unroll
            for (int i = 1; i < N; ++i) {
                prim::subc_cc(u[i], w[i]);
            }
            const uint32_t lt = prim::subc(0, 0);                // all-ones when u < w
            const uint32_t sw = odd & lt;
#pragma unroll
            for (int i = 0; i < N; ++i) {                         // conditional swap (u, x1) <-> (w, x2)
                const uint32_t du = (u[i] ^ w[i]) & sw, dx = (x1[i] ^ x2[i]) & sw;
                u[i] ^= du;
                w[i] ^= du;
                x1[i] ^= dx;
                x2[i] ^= dx;
            }
            // u -= w when odd (then u is even)
            u[0] = prim::sub_cc(u[0], w[0] & odd);
#pragma unroll
            for (int i = 1; i < N; ++i) {
                u[i] = prim::subc_cc(u[i], w[i] & odd);
            }
            // x1 = x1 - x2 mod p when odd
            x1[0] = prim::sub_cc(x1[0], x2[0] & odd);
#pragma unroll
            for (int i = 1; i < N; ++i) {
                x1[i] = prim::subc_cc(x1[i], x2[i] & odd);
            }
            const uint32_t neg = prim::subc(0, 0);
            x1[0] = prim::add_cc(x1[0], P::mod(0) & neg);
#pragma unroll
            for (int i = 1; i < N - 1; ++i) {
                x1[i] = prim::addc_cc(x1[i], P::mod(i) & neg);
            }
            x1[N - 1] = prim::addc(x1[N - 1], P::mod(N - 1) & neg);
            // u /= 2
#pragma unroll
            for (int i = 0; i < N - 1; ++i) {
                u[i] = (u[i] >> 1) | (u[i + 1] << 31);
            }
            u[N - 1] >>= 1;
            // x1 = x1 / 2 mod p: (x1 + p when odd) >> 1; x1 + p < 2p < 2^(32N)
            const uint32_t xo = 0u - (x1[0] & 1u);
            x1[0] = prim::add_cc(x1[0], P::mod(0) & xo);
#pragma unroll
            for (int i = 1; i < N - 1; ++i) {
                x1[i] = prim::addc_cc(x1[i], P::mod(i) & xo);
            }
            x1[N - 1] = prim::addc(x1[N - 1], P::mod(N - 1) & xo);
#pragma unroll
            for (int i = 0; i < N - 1; ++i) {
                x1[i] = (x1[i] >> 1) | (x1[i + 1] << 31);
            }
            x1[N - 1] >>= 1;
        }
        Fp r;
        const bool z = is_zero();
#pragma unroll
        for (int i = 0; i < N; ++i) {
            r.v[i] = z ? 0u : x2[i];
        }
        const Fp r3 = r2() * r2(); // R^2 * R^2 * R^-1 = R^3
        return r * r3;
    }
};

// Fp2 = Fp[u] / (u^2 - nr).  libff Fp2_model (fp2.hpp; mul fp2.tcc:102-114 Karatsuba,
// squared fp2.tcc:142-151 complex squaring). In-memory layout c0 | c1 as in libff.
// P2 supplies: typedef Fp<...> base;  static base mul_by_nr(const base&).
template<class P2> struct Fp2 {
    typedef typename P2::base F;
    typedef P2 params;
    F c0, c1;

    HD static Fp2 zero() { return Fp2{F::zero(), F::zero()}; }
    HD static Fp2 one() { return Fp2{F::one(), F::zero()}; }
    HD bool is_zero() const { return c0.is_zero() && c1.is_zero(); }
    HD bool operator==(const Fp2 &o) const { return c0 == o.c0 && c1 == o.c1; }
    HD bool operator!=(const Fp2 &o) const { return !(*this == o); }
    HD friend Fp2 operator+(const Fp2 &a, const Fp2 &b) { return Fp2{a.c0 + b.c0, a.c1 + b.c1}; }
    HD friend Fp2 operator-(const Fp2 &a, const Fp2 &b) { return Fp2{a.c0 - b.c0, a.c1 - b.c1}; }
    HD Fp2 neg() const { return Fp2{c0.neg(), c1.neg()}; }
    HD Fp2 cneg(bool f) const { return Fp2{c0.cneg(f), c1.cneg(f)}; }
    HD Fp2 dbl() const { return Fp2{c0.dbl(), c1.dbl()}; }
    // Karatsuba product / complex squaring (fp2.tcc:102-151). With B200_OUTLINE_FP2 the whole Fp2
    // operation is ONE out-of-line body whose base-field products are inlined side by side: the
    // three (two) independent carry chains give the scheduler latency cover that a kernel running
    // one or two warps per scheduler (255 registers) does not get from other warps.
    HD static Fp2 mul_body(const Fp2 &a, const Fp2 &b)
    {
#if defined(__CUDA_ARCH__) && defined(B200_OUTLINE_FP2)
        const F aA = F::mul_inline(a.c0, b.c0);
        const F bB = F::mul_inline(a.c1, b.c1);
        const F s = F::mul_inline(a.c0 + a.c1, b.c0 + b.c1);
#else
        const F aA = a.c0 * b.c0;
        const F bB = a.c1 * b.c1;
        const F s = (a.c0 + a.c1) * (b.c0 + b.c1);
#endif
        return Fp2{aA + P2::mul_by_nr(bB), s - aA - bB};
    }
    HD static Fp2 sqr_body(const Fp2 &a)
    {
#if defined(__CUDA_ARCH__) && defined(B200_OUTLINE_FP2)
        const F ab = F::mul_inline(a.c0, a.c1);
        const F t = F::mul_inline(a.c0 + a.c1, a.c0 + P2::mul_by_nr(a.c1));
#else
        const F ab = a.c0 * a.c1;
        const F t = (a.c0 + a.c1) * (a.c0 + P2::mul_by_nr(a.c1));
#endif
        return Fp2{t - ab - P2::mul_by_nr(ab), ab + ab};
    }
#if defined(__CUDACC__) && defined(B200_OUTLINE_FP2)
    static __device__ __noinline__ Fp2 mul_outlined(Fp2 a, Fp2 b) { return mul_body(a, b); }
    static __device__ __noinline__ Fp2 sqr_outlined(Fp2 a) { return sqr_body(a); }
#endif
    HD friend Fp2 operator*(const Fp2 &a, const Fp2 &b)
    {
#if defined(__CUDA_ARCH__) && defined(B200_OUTLINE_FP2)
        return mul_outlined(a, b);
#else
        return mul_body(a, b);
#endif
    }
    HD Fp2 sqr() const
    {
#if defined(__CUDA_ARCH__) && defined(B200_OUTLINE_FP2)
        return sqr_outlined(*this);
#else
        return sqr_body(*this);
#endif
    }
    HD Fp2 inverse() const
    {
        // fp2.tcc:153-167: (c0 - c1 u) / (c0^2 - nr c1^2)
        const F t0 = c0.sqr();
        const F t1 = c1.sqr();
        const F t2 = t0 - P2::mul_by_nr(t1);
        const F t3 = t2.inverse();
        return Fp2{c0 * t3, (c1 * t3).neg()};
    }
    HD Fp2 inverse_fast() const
    {
        const F t3 = (c0.sqr() - P2::mul_by_nr(c1.sqr())).inverse_fast();
        return Fp2{c0 * t3, (c1 * t3).neg()};
    }
};

} // namespace b200
