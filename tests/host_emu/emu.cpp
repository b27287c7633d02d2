// tests/host_emu/emu.cpp — TEST INFRASTRUCTURE.
// Compiles the device field / curve templates (libsnark_b200/csrc/{fp,curve}.cuh) for the
// HOST through the carry-flag emulation in prims.cuh, so their arithmetic can be checked
// against the reference on a machine without a GPU. Not linked into the product library.
#include <cstdint>
#include <cstring>

#include "curves.cuh"

using namespace b200;

template<class F> static int field_op(int op, const void *a_, const void *b_, void *o_, size_t n)
{
    const F *a = (const F *)a_;
    const F *b = (const F *)b_;
    F *o = (F *)o_;
    for (size_t i = 0; i < n; ++i) {
        switch (op) {
        case 0: o[i] = a[i] + b[i]; break;
        case 1: o[i] = a[i] - b[i]; break;
        case 2: o[i] = a[i] * b[i]; break;
        case 3: o[i] = a[i].sqr(); break;
        case 4: o[i] = a[i].neg(); break;
        case 5: o[i] = a[i].is_zero() ? a[i] : a[i].inverse(); break;
        default: return -1;
        }
    }
    return 0;
}

template<class F> static int fp_convert(int to_mont, const void *a_, void *o_, size_t n)
{
    const F *a = (const F *)a_;
    F *o = (F *)o_;
    for (size_t i = 0; i < n; ++i) {
        o[i] = to_mont ? a[i].to_mont() : a[i].from_mont();
    }
    return 0;
}

// Points travel in the reference's Jacobian (X,Y,Z) layout; Z in {0 -> infinity, one -> affine} on input.
// op: 0 = add (a, b affine or inf) via XYZZ add, 1 = madd (a += b), 2 = dbl, 3 = add via madd-chain check.
// out: affine X|Y|Z(one) or the all-zero-X, Y=one, Z=0 encoding is NOT produced here: out is (x, y, flag).
template<class F> static int group_op(int op, const void *a_, const void *b_, void *o_, size_t n)
{
    struct J {
        F x, y, z;
    };
    const J *a = (const J *)a_;
    const J *b = (const J *)b_;
    J *o = (J *)o_;
    for (size_t i = 0; i < n; ++i) {
        XYZZ<F> A = a[i].z.is_zero() ? XYZZ<F>::inf() : XYZZ<F>::from_affine(Affine<F>{a[i].x, a[i].y});
        XYZZ<F> B = b[i].z.is_zero() ? XYZZ<F>::inf() : XYZZ<F>::from_affine(Affine<F>{b[i].x, b[i].y});
        // de-normalise A so the general formulas are exercised: A <- 2A - A would need neg; use dbl+add paths.
        switch (op) {
        case 0: add(A, B); break;
        case 1:
            if (!b[i].z.is_zero()) {
                madd(A, Affine<F>{b[i].x, b[i].y});
            }
            break;
        case 2: dbl(A); break;
        case 3: {
            // (2A) + B through the general add with a non-trivial ZZ on the left
            dbl(A);
            add(A, B);
            break;
        }
        case 4: {
            // (2A) + B through madd with a non-trivial ZZ on the left
            dbl(A);
            if (!b[i].z.is_zero()) {
                madd(A, Affine<F>{b[i].x, b[i].y});
            }
            break;
        }
        case 5: {
            // A + (2B): non-trivial ZZ on the right
            dbl(B);
            add(A, B);
            break;
        }
        default: return -1;
        }
        Affine<F> r;
        if (to_affine(A, r)) {
            o[i].x = r.x;
            o[i].y = r.y;
            o[i].z = F::one();
        } else {
            o[i].x = F::zero();
            o[i].y = F::one();
            o[i].z = F::zero();
        }
    }
    return 0;
}

#define FIELD_DISPATCH(fn, ...)                                         \
    switch (curve * 4 + field) {                                        \
    case 0: return fn<alt_bn128_fr>(__VA_ARGS__);                       \
    case 1: return fn<alt_bn128_fq>(__VA_ARGS__);                       \
    case 4: return fn<bls12_377_fr>(__VA_ARGS__);                       \
    case 5: return fn<bls12_377_fq>(__VA_ARGS__);                       \
    case 8: return fn<bw6_761_fr>(__VA_ARGS__);                         \
    case 9: return fn<bw6_761_fq>(__VA_ARGS__);                         \
    }

extern "C" int emu_field_op(int curve, int field, int op, const void *a, const void *b, void *o, size_t n)
{
    FIELD_DISPATCH(field_op, op, a, b, o, n)
    if (curve == 0 && field == 2) return field_op<alt_bn128_fq2>(op, a, b, o, n);
    if (curve == 1 && field == 2) return field_op<bls12_377_fq2>(op, a, b, o, n);
    if (curve == 2 && field == 2) return field_op<bw6_761_fq>(op, a, b, o, n);
    return -100;
}

extern "C" int emu_fp_convert(int curve, int field, int to_mont, const void *a, void *o, size_t n)
{
    FIELD_DISPATCH(fp_convert, to_mont, a, o, n)
    return -100;
}

extern "C" int emu_group_op(int curve, int group, int op, const void *a, const void *b, void *o, size_t n)
{
    switch (curve * 4 + group) {
    case 1: return group_op<alt_bn128_fq>(op, a, b, o, n);
    case 2: return group_op<alt_bn128_fq2>(op, a, b, o, n);
    case 5: return group_op<bls12_377_fq>(op, a, b, o, n);
    case 6: return group_op<bls12_377_fq2>(op, a, b, o, n);
    case 9: return group_op<bw6_761_fq>(op, a, b, o, n);
    case 10: return group_op<bw6_761_fq>(op, a, b, o, n);
    }
    return -100;
}
