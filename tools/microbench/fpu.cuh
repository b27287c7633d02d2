// fpu.cuh — prime-field arithmetic on UNSATURATED signed limbs.
//
// EXPERIMENT (round 1), NOT part of the library: kept with imad_bench.cu as the evidence behind
// DESIGN.md section 7. Idea: L limbs of B bits in int32, limb products accumulated into int64 columns
// by multiply-adds that do not chain carries, carries moved by shifts on the ALU pipe. Outcome on
// B200: SLOWER than the 32-bit-limb carry chains of fp.cuh (4.5e10 vs 6.7e10 modmul/s, bit-exact in
// both cases). IMAD.WIDE issues at half the 32-bit IMAD rate whether or not it consumes a carry, and
// ptxas splits every 64-bit accumulate into IMAD.WIDE + IADD3 + IADD3.X, so the extra limbs and
// the carry shuffling are pure cost. Limbs are SIGNED: a subtraction is a limb-wise subtraction
// with no "+kp" offset, and values live in a small interval around 0 instead of [0, p).
//
// Representation: value(a) = sum_i v[i] 2^(B i). Montgomery radix R' = 2^(L B) (not the
// reference's R = 2^(32 N)); x is stored as x R' mod p up to a multiple of p. Conversion to and
// from the reference's form happens at the kernel's edges only (from_mont32 / to_mont32), so
// what reaches memory is the canonical 32-bit-limb Montgomery element every other kernel and
// the reference use: results stay bit-exact.
//
// Bounds (checked by static_assert where compile-time, by tests/host_emu otherwise):
//   "normalised": limbs 0..L-2 in [0, 2^B), top limb signed and small.
//   mul(a, b) needs L * (|a_i| |b_j|) + L * 2^(2B) < 2^63 for all i, j; with one operand
//   normalised or a difference of two normalised values (|limb| < 2^B) the other may be up to
//   MUL_HEADROOM times larger. mul's result is normalised and lies in (-|a b|/R', |a b|/R' + p).
//   The code is plain C++ (no PTX), so the same source runs on the host for testing.
#pragma once
#include "curve.cuh"

namespace b200
{

// c + a * b on 64-bit columns. On the device this is pinned to ONE signed wide multiply-add
// (IMAD.WIDE): left to the compiler, range knowledge about masked limbs makes it pick unsigned
// widening multiplies plus sign-correction instructions, which doubles the instruction count.
HD int64_t mad_wide(int32_t a, int32_t b, int64_t c)
{
#if defined(__CUDA_ARCH__)
    int64_t r;
    asm("mad.wide.s32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c));
    return r;
#else
    return c + (int64_t)a * (int64_t)b;
#endif
}

template<class P, int L_, int B_> struct FpU {
    static constexpr int L = L_;
    static constexpr int B = B_;
    static constexpr int N = P::N;
    static constexpr uint32_t MASK = (1u << B) - 1u;
    static constexpr int SHIFT = L * B - 32 * N; // R' = R 2^SHIFT
    static_assert(SHIFT >= 0 && SHIFT < B, "limb geometry");
    static_assert(L * B >= P::BITS + 6, "need slack above the modulus for lazy values");
    // L products of two B-bit limbs plus L products of the reduction must fit a signed 64-bit column
    static_assert(2 * L < (1 << (63 - 2 * B)), "column accumulators overflow");

    int32_t v[L];

    // bits [pos, pos + B) of the little-endian 32-bit-limb integer w[0..N) (compile-time pos)
    template<class W> HD static constexpr uint32_t bits_at(W w, int pos)
    {
        if (pos < 0) {
            return (w(0) << (-pos)) & MASK;
        }
        const int k = pos >> 5, off = pos & 31;
        uint32_t lo = k < N ? w(k) >> off : 0u;
        if (off != 0 && k + 1 < N) {
            lo |= w(k + 1) << (32 - off);
        }
        return lo & MASK;
    }
    struct ModW {
        HD constexpr uint32_t operator()(int i) const { return P::mod(i); }
    };
    struct OneW {
        HD constexpr uint32_t operator()(int i) const { return P::one(i); }
    };
    // modulus limb j in radix 2^B
    HD static constexpr uint32_t modl(int j) { return bits_at(ModW(), j * B); }
    // (R mod p) as an integer, limb j in radix 2^B: the factor that takes x R' to x R under mul()
    HD static constexpr uint32_t onel(int j) { return bits_at(OneW(), j * B); }
    // -p^-1 mod 2^B
    HD static constexpr uint32_t pinv()
    {
        uint32_t p0 = P::mod(0), x = 1;
        for (int i = 0; i < 5; ++i) {
            x *= 2u - p0 * x; // Newton: x = p0^-1 mod 2^32
        }
        return (0u - x) & MASK;
    }

    HD static FpU zero()
    {
        FpU r;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            r.v[i] = 0;
        }
        return r;
    }
    // R' mod p: the field's 1 in this Montgomery radix (canonical, < p)
    HD static FpU one()
    {
        FpU r;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            r.v[i] = (int32_t)P::uone(i);
        }
        return r;
    }
    // x R (reference form) -> x R' reduced to (-p/2, 1.3 p): one multiplication by one()
    HD static FpU from_mont32_reduced(const Fp<P> &x) { return from_mont32(x) * one(); }

    // x (canonical, reference Montgomery form x R) -> x R 2^SHIFT = x R' as an integer < 2^SHIFT p.
    // Normalised and non-negative. |value| < 2^SHIFT p: fine as ONE operand of a mul().
    HD static FpU from_mont32(const Fp<P> &x)
    {
        FpU r;
#pragma unroll
        for (int j = 0; j < L; ++j) {
            const int pos = j * B - SHIFT;
            uint32_t w;
            if (pos < 0) {
                w = (x.v[0] << (-pos)) & MASK;
            } else {
                const int k = pos >> 5, off = pos & 31;
                w = k < N ? x.v[k] >> off : 0u;
                if (off != 0 && k + 1 < N) {
                    w |= x.v[k + 1] << (32 - off);
                }
                if (j < L - 1) {
                    w &= MASK;
                }
            }
            r.v[j] = (int32_t)w;
        }
        return r;
    }

    HD friend FpU operator+(const FpU &a, const FpU &b)
    {
        FpU r;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            r.v[i] = a.v[i] + b.v[i];
        }
        return r;
    }
    HD friend FpU operator-(const FpU &a, const FpU &b)
    {
        FpU r;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            r.v[i] = a.v[i] - b.v[i];
        }
        return r;
    }
    // a - b - 2c in one pass
    HD static FpU sub_sub_dbl(const FpU &a, const FpU &b, const FpU &c)
    {
        FpU r;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            r.v[i] = a.v[i] - b.v[i] - 2 * c.v[i];
        }
        return r;
    }
    HD FpU cneg(bool flag) const
    {
        FpU r;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            r.v[i] = flag ? -v[i] : v[i];
        }
        return r;
    }

    // carry propagation: same value, limbs 0..L-2 in [0, 2^B)
    HD FpU norm() const
    {
        FpU r;
        int32_t c = 0;
#pragma unroll
        for (int i = 0; i < L - 1; ++i) {
            const int32_t t = v[i] + c;
            r.v[i] = (int32_t)((uint32_t)t & MASK);
            c = t >> B; // arithmetic
        }
        r.v[L - 1] = v[L - 1] + c;
        return r;
    }

    // Montgomery product a b / R' (mod p). See the bounds at the top of the file.
    HD friend FpU operator*(const FpU &a, const FpU &b)
    {
        int64_t t[L + 1];
#pragma unroll
        for (int i = 0; i <= L; ++i) {
            t[i] = 0;
        }
#pragma unroll
        for (int i = 0; i < L; ++i) {
            // row i of the product into columns i .. i+L-1 (held in t[0..L-1] after the shift below)
#pragma unroll
            for (int j = 0; j < L; ++j) {
                t[j] = mad_wide(a.v[i], b.v[j], t[j]);
            }
            // make column i divisible by 2^B
            const uint32_t m = ((uint32_t)t[0] * pinv()) & MASK;
#pragma unroll
            for (int j = 0; j < L; ++j) {
                t[j] = mad_wide((int32_t)m, (int32_t)modl(j), t[j]);
            }
            // drop it, carrying into the next column; shift the window
            const int64_t carry = t[0] >> B;
#pragma unroll
            for (int j = 0; j < L; ++j) {
                t[j] = t[j + 1];
            }
            t[0] += carry;
            t[L] = 0;
        }
        FpU r;
        int64_t c = 0;
#pragma unroll
        for (int i = 0; i < L - 1; ++i) {
            const int64_t s = t[i] + c;
            r.v[i] = (int32_t)((uint32_t)s & MASK);
            c = s >> B;
        }
        r.v[L - 1] = (int32_t)(t[L - 1] + c);
        return r;
    }
    HD FpU sqr() const { return (*this) * (*this); }

    // cheap necessary condition for value == 0 (mod p) when |value| < 2p: the low limb of a
    // multiple of p in {-p, 0, p}. False positives with probability ~ 3 / 2^B.
    HD bool maybe_zero_mod_p() const
    {
        const uint32_t l0 = (uint32_t)v[0] & MASK;
        return l0 == 0u || l0 == modl(0) || l0 == ((0u - modl(0)) & MASK);
    }

    // Canonical representative in [0, p) as normalised non-negative limbs. |value| < 2^(SLACK) p.
    HD FpU canonical() const
    {
        FpU r = norm();
        // bring into [0, p): add or subtract p while needed (a handful of iterations at most)
        for (int it = 0; it < 64; ++it) {
            if (r.v[L - 1] < 0) {
                r = r.add_p().norm();
                continue;
            }
            const FpU d = r.sub_p().norm();
            if (d.v[L - 1] < 0) {
                break;
            }
            r = d;
        }
        return r;
    }
    HD bool is_zero_mod_p() const
    {
        const FpU c = canonical();
        int32_t acc = 0;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            acc |= c.v[i];
        }
        return acc == 0;
    }

    // x R' (any representative within the bounds) -> canonical reference form x R, 32-bit limbs.
    HD Fp<P> to_mont32() const
    {
        FpU k;
#pragma unroll
        for (int j = 0; j < L; ++j) {
            k.v[j] = (int32_t)onel(j);
        }
        const FpU c = ((*this) * k).canonical();
        Fp<P> o;
#pragma unroll
        for (int w = 0; w < N; ++w) {
            // bits [32 w, 32 w + 32) of sum_j c.v[j] 2^(B j)
            const int j0 = (32 * w) / B, off = (32 * w) % B;
            uint64_t acc = (uint64_t)(uint32_t)c.v[j0] >> off;
            int have = B - off;
            for (int j = j0 + 1; have < 32 && j < L; ++j) {
                acc |= (uint64_t)(uint32_t)c.v[j] << have;
                have += B;
            }
            o.v[w] = (uint32_t)acc;
        }
        return o;
    }

  private:
    HD FpU add_p() const
    {
        FpU r;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            r.v[i] = v[i] + (int32_t)modl(i);
        }
        return r;
    }
    HD FpU sub_p() const
    {
        FpU r;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            r.v[i] = v[i] - (int32_t)modl(i);
        }
        return r;
    }
};

// The limb geometry per field comes from the generated parameters (254 bits: 9 x 29 = 261;
// 377: 14 x 28 = 392; 761: 28 x 28 = 784).
template<class P> using FpUOf = FpU<P, P::UL, P::UB>;

// ---- XYZZ bucket accumulator on unsaturated limbs ------------------------------------------------
// Same formulas as curve.cuh madd (EFD madd-2008-s); the lazy-reduction schedule and its bounds
// are in the comments. Invariant between calls: x, y, zz, zzz normalised, |x| < 4.5p, |y| < 1.3p,
// zz, zzz in (-0.1p, 1.3p); `inf` is tracked by the caller-visible flag, not by zz == 0.
template<class P> struct XYZZU {
    typedef FpUOf<P> U;
    U x, y, zz, zzz;
    bool inf;

    HD static XYZZU infinity()
    {
        XYZZU r;
        r.x = r.y = r.zz = r.zzz = U::zero();
        r.inf = true;
        return r;
    }
    HD void set_affine(const Affine<Fp<P>> &p, bool negate)
    {
        x = U::from_mont32_reduced(p.x);
        y = U::from_mont32_reduced(p.y).cneg(negate);
        zz = U::one();
        zzz = U::one();
        inf = false;
    }
    HD XYZZ<Fp<P>> to_xyzz32() const
    {
        XYZZ<Fp<P>> r;
        if (inf) {
            return XYZZ<Fp<P>>::inf();
        }
        r.x = x.to_mont32();
        r.y = y.to_mont32();
        r.zz = zz.to_mont32();
        r.zzz = zzz.to_mont32();
        return r;
    }
    HD void set_xyzz32(const XYZZ<Fp<P>> &s)
    {
        inf = s.is_inf();
        x = U::from_mont32_reduced(s.x);
        y = U::from_mont32_reduced(s.y);
        zz = U::from_mont32_reduced(s.zz);
        zzz = U::from_mont32_reduced(s.zzz);
    }
};

// rare cases of madd_u, kept out of line: pd == 0 (mod p) means P + P (double) or P + (-P) (infinity)
template<class P>
#if defined(__CUDACC__)
__device__ __host__ __noinline__
#endif
    void
    madd_u_rare(XYZZU<P> &a, const Affine<Fp<P>> &p, bool negate, const FpUOf<P> &rd)
{
    if (rd.is_zero_mod_p()) {
        Affine<Fp<P>> q = p;
        q.y = q.y.cneg(negate);
        a.set_xyzz32(dbl_affine(q));
    } else {
        a.inf = true;
    }
}

// a += (negate ? -p : p), p affine and not infinity.
template<class P> HD void madd_u(XYZZU<P> &a, const Affine<Fp<P>> &p, bool negate)
{
    typedef FpUOf<P> U;
    if (a.inf) {
        a.set_affine(p, negate);
        return;
    }
    const U x2 = U::from_mont32(p.x);               // [0, 2^SHIFT p), normalised
    const U y2 = U::from_mont32(p.y).cneg(negate);  // |.| < 2^SHIFT p, |limb| < 2^B
    const U u2 = x2 * a.zz;                         // |.| < 1.4p
    const U s2 = y2 * a.zzz;
    const U pd = u2 - a.x;                          // |limb| < 2^B, |.| < 6p
    const U rd = s2 - a.y;                          // |limb| < 2^B, |.| < 2.7p
    const U pp = pd.sqr();                          // [0, 2p): a square, reduced
    // pd == 0 (mod p) <=> pp in {0, p}; its low limb is exact, so two compares screen it
    const uint32_t l0 = (uint32_t)pp.v[0];
    if ((l0 == 0u || l0 == U::modl(0)) && pp.is_zero_mod_p()) {
        madd_u_rare<P>(a, p, negate, rd);
        return;
    }
    const U ppp = pd * pp;                          // |.| < 1.1p
    const U q = a.x * pp;                           // |.| < 1.1p
    const U x3 = U::sub_sub_dbl(rd.sqr(), ppp, q).norm(); // (-3.2p, 1.3p), normalised
    const U y3 = (rd * (q - x3) - a.y * ppp).norm();      // |.| < 1.2p, normalised
    a.zz = a.zz * pp;
    a.zzz = a.zzz * ppp;
    a.x = x3;
    a.y = y3;
}

} // namespace b200
