// curve.cuh — short-Weierstrass (a = 0) group law in XYZZ coordinates.
//
// Replaces, on the device, libff's GroupT::mixed_add / add / dbl / operator- for
//   alt_bn128 G1/G2 (alt_bn128_g1.cpp:149-326, alt_bn128_g2.cpp:166-349),
//   bls12_377 G1/G2 (bls12_377_g1.cpp:190-352, bls12_377_g2.cpp:209-368),
//   bw6_761  G1/G2 (bw6_761_g1.cpp:206-358,  bw6_761_g2.cpp:211-...).
// All six curves' groups have a = 0, so one formula set serves them. The reference
// uses Jacobian (or, for bw6_761, homogeneous projective) coordinates; group
// elements are only canonical after normalisation, so the device is
// free to use XYZZ (x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2): mixed add 8M+2S, add 12M+2S,
// no inversions. Results are normalised to affine before they leave the device and
// are then bit-identical to the reference's to_affine_coordinates().
//
// Exceptional cases follow libff's test list (test_groups.cpp:24-57): O+P, P+O,
// P+P (-> dbl), P+(-P) (-> O).
#pragma once
#include "fp.cuh"

namespace b200
{

template<class F> struct Affine {
    F x, y;
};

template<class F> struct XYZZ {
    F x, y, zz, zzz;

    HD static XYZZ inf()
    {
        XYZZ r;
        r.x = F::zero();
        r.y = F::zero();
        r.zz = F::zero();
        r.zzz = F::zero();
        return r;
    }
    HD bool is_inf() const { return zz.is_zero(); }
    HD static XYZZ from_affine(const Affine<F> &p)
    {
        XYZZ r;
        r.x = p.x;
        r.y = p.y;
        r.zz = F::one();
        r.zzz = F::one();
        return r;
    }
};

// 2 * (affine p)   — EFD dbl-2008-s-1 with ZZ = ZZZ = 1, a = 0.
template<class F> HD XYZZ<F> dbl_affine(const Affine<F> &p)
{
    const F u = p.y.dbl();
    const F v = u.sqr();
    const F w = u * v;
    const F s = p.x * v;
    const F x2 = p.x.sqr();
    const F m = x2.dbl() + x2;
    XYZZ<F> r;
    r.x = m.sqr() - s.dbl();
    r.y = m * (s - r.x) - w * p.y;
    r.zz = v;
    r.zzz = w;
    return r;
}

// acc = 2 * acc   — EFD dbl-2008-s-1, a = 0. (y == 0 cannot occur on these prime-order groups,
// but is handled: U = 0 gives ZZ = 0 = infinity, which is the right answer for a 2-torsion point.)
template<class F> HD void dbl(XYZZ<F> &a)
{
    const F u = a.y.dbl();
    const F v = u.sqr();
    const F w = u * v;
    const F s = a.x * v;
    const F x2 = a.x.sqr();
    const F m = x2.dbl() + x2;
    const F x3 = m.sqr() - s.dbl();
    a.y = m * (s - x3) - w * a.y;
    a.x = x3;
    a.zz = v * a.zz;
    a.zzz = w * a.zzz;
}

#if defined(__CUDACC__)
template<class F> __device__ __noinline__ void dbl_affine_outlined(XYZZ<F> &r, const Affine<F> &p) { r = dbl_affine(p); }
#endif
template<class F> HD XYZZ<F> dbl_affine_cold(const Affine<F> &p)
{
#if defined(__CUDA_ARCH__)
    XYZZ<F> r;
    dbl_affine_outlined(r, p);
    return r;
#else
    return dbl_affine(p);
#endif
}

// Branch-free selection between two points (per-limb selects; keeps the warp converged).
template<class F> HD XYZZ<F> select(bool take_b, const XYZZ<F> &a, const XYZZ<F> &b)
{
    XYZZ<F> r;
    const uint32_t *pa = reinterpret_cast<const uint32_t *>(&a);
    const uint32_t *pb = reinterpret_cast<const uint32_t *>(&b);
    uint32_t *pr = reinterpret_cast<uint32_t *>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(XYZZ<F>) / 4); ++i) {
        pr[i] = take_b ? pb[i] : pa[i];
    }
    return r;
}

// acc += p (p affine, never infinity)   — EFD madd-2008-s.
// Counterpart of libff mixed_add (alt_bn128_g1.cpp:208-283). Written without early exits:
// the general formula is always evaluated (P + (-P) falls out of it as ZZ = 0), an infinite
// accumulator is patched by a select, and only P + P takes a (rare, out-of-line) branch.
template<class F> HD void madd(XYZZ<F> &a, const Affine<F> &p)
{
    const bool a_inf = a.is_inf();
    const F u2 = p.x * a.zz;
    const F s2 = p.y * a.zzz;
    const F pd = u2 - a.x;
    const F rd = s2 - a.y;
    const F pp = pd.sqr();
    const F ppp = pd * pp;
    const F q = a.x * pp;
    XYZZ<F> r;
    r.x = rd.sqr() - ppp - q.dbl();
    r.y = rd * (q - r.x) - a.y * ppp;
    r.zz = a.zz * pp;
    r.zzz = a.zzz * ppp;
    if (!a_inf && pd.is_zero() && rd.is_zero()) {
        r = dbl_affine_cold(p); // P + P
    }
    a = select(a_inf, r, XYZZ<F>::from_affine(p));
}

// acc += b   — EFD add-2008-s. Counterpart of libff add / operator+ (alt_bn128_g1.cpp:97-206).
// Same structure as madd: no early exits, selects for the infinity cases.
template<class F> HD void add(XYZZ<F> &a, const XYZZ<F> &b)
{
    const bool a_inf = a.is_inf(), b_inf = b.is_inf();
    const F u1 = a.x * b.zz;
    const F u2 = b.x * a.zz;
    const F s1 = a.y * b.zzz;
    const F s2 = b.y * a.zzz;
    const F pd = u2 - u1;
    const F rd = s2 - s1;
    const F pp = pd.sqr();
    const F ppp = pd * pp;
    const F q = u1 * pp;
    XYZZ<F> r;
    r.x = rd.sqr() - ppp - q.dbl();
    r.y = rd * (q - r.x) - s1 * ppp;
    r.zz = a.zz * b.zz * pp;
    r.zzz = a.zzz * b.zzz * ppp;
    if (!a_inf && !b_inf && pd.is_zero() && rd.is_zero()) {
        r = a;
        dbl(r); // P + P
    }
    r = select(a_inf, r, b);
    a = select(b_inf, r, a);
}

// Out-of-line copies for kernels where the group law is not the inner loop (bucket
// reduction, window sums, fold): one instance per translation unit instead of one per call site.
#if defined(__CUDACC__)
template<class F> __device__ __noinline__ void add_outlined(XYZZ<F> &a, const XYZZ<F> &b) { add(a, b); }
template<class F> __device__ __noinline__ void dbl_outlined(XYZZ<F> &a) { dbl(a); }
#endif
template<class F> HD void add_cold(XYZZ<F> &a, const XYZZ<F> &b)
{
#if defined(__CUDA_ARCH__)
    add_outlined(a, b);
#else
    add(a, b);
#endif
}
template<class F> HD void dbl_cold(XYZZ<F> &a)
{
#if defined(__CUDA_ARCH__)
    dbl_outlined(a);
#else
    dbl(a);
#endif
}

template<class F> HD XYZZ<F> neg(const XYZZ<F> &a)
{
    XYZZ<F> r = a;
    r.y = a.y.neg();
    return r;
}

// Normalise: returns false for infinity, else writes affine (x, y).
// x = X/ZZ, y = Y/ZZZ: one inversion of ZZ*ZZZ... kept simple: inv(ZZZ), then
// 1/ZZ = ZZZ^-1 * (ZZZ/ZZ) is not available without Z, so two products off one inverse:
//   t = (ZZ*ZZZ)^-1;  1/ZZ = t*ZZZ;  1/ZZZ = t*ZZ.
template<class F> HD bool to_affine(const XYZZ<F> &a, Affine<F> &out)
{
    if (a.is_inf()) {
        out.x = F::zero();
        out.y = F::zero();
        return false;
    }
    const F t = (a.zz * a.zzz).inverse();
    out.x = a.x * (t * a.zzz);
    out.y = a.y * (t * a.zz);
    return true;
}

} // namespace b200
