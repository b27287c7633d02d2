// selftest.cu — device-side unit-test hooks for the field and group arithmetic (fp.cuh,
// curve.cuh). Not part of the reference surface: tests call these through the C-ABI to
// compare every primitive against libff's Fp_model / group law on the same inputs
// (SURVEY.md §7.1 steps 1-2).
#include "curves.cuh"
#include "msm.cuh"
#include "runtime.hpp"

using namespace b200;

namespace
{

template<class F> __global__ void k_field_op(int op, const F *a, const F *b, F *o, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) {
        return;
    }
    const F x = msm::load_vec(a + i), y = msm::load_vec(b + i);
    F r;
    switch (op) {
    case 0: r = x + y; break;
    case 1: r = x - y; break;
    case 2: r = x * y; break;
    case 3: r = x.sqr(); break;
    case 4: r = x.neg(); break;
    case 5: r = x.is_zero() ? x : x.inverse(); break;
    default: r = x; break;
    }
    msm::store_vec(o + i, r);
}

template<class F> __global__ void k_fp_convert(int to_mont, const F *a, F *o, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const F x = msm::load_vec(a + i);
        msm::store_vec(o + i, to_mont ? x.to_mont() : x.from_mont());
    }
}

// Points in the reference layout with Z in {0, one}. op: 0 add, 1 madd, 2 dbl, 3 (2A)+B via add,
// 4 (2A)+B via madd, 5 A+(2B) via add. Output normalised.
template<class F> __global__ void k_group_op(int op, const msm::RefPoint<F> *a, const msm::RefPoint<F> *b, msm::RefPoint<F> *o, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) {
        return;
    }
    const msm::RefPoint<F> pa = msm::load_vec(a + i), pb = msm::load_vec(b + i);
    XYZZ<F> A = pa.z.is_zero() ? XYZZ<F>::inf() : XYZZ<F>::from_affine(Affine<F>{pa.x, pa.y});
    XYZZ<F> B = pb.z.is_zero() ? XYZZ<F>::inf() : XYZZ<F>::from_affine(Affine<F>{pb.x, pb.y});
    const Affine<F> bq{pb.x, pb.y};
    switch (op) {
    case 0: add_cold(A, B); break;
    case 1:
        if (!pb.z.is_zero()) {
            madd(A, bq);
        }
        break;
    case 2: dbl_cold(A); break;
    case 3:
        dbl_cold(A);
        add_cold(A, B);
        break;
    case 4:
        dbl_cold(A);
        if (!pb.z.is_zero()) {
            madd(A, bq);
        }
        break;
    case 5:
        dbl_cold(B);
        add_cold(A, B);
        break;
    default: break;
    }
    msm::store_vec(o + i, msm::normalise(A));
}

rt::DevBuf g_a, g_b, g_o;

template<class T, class Launch> int run3(const void *a, const void *b, void *o, size_t n, Launch launch)
{
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    const size_t bytes = n * sizeof(T);
    if (g_a.reserve(bytes) || g_b.reserve(bytes) || g_o.reserve(bytes)) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(g_a.p, a, bytes, cudaMemcpyHostToDevice, s));
    B200_CUDA_CHECK(cudaMemcpyAsync(g_b.p, b ? b : a, bytes, cudaMemcpyHostToDevice, s));
    launch(g_a.as<T>(), g_b.as<T>(), g_o.as<T>(), s);
    B200_LAUNCH_CHECK();
    B200_CUDA_CHECK(cudaMemcpyAsync(o, g_o.p, bytes, cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

template<class F> int field_op(int op, const void *a, const void *b, void *o, size_t n)
{
    return run3<F>(a, b, o, n, [&](F *da, F *db, F *dо, cudaStream_t s) {
        k_field_op<F><<<(unsigned)((n + 63) / 64), 64, 0, s>>>(op, da, db, dо, n);
    });
}
template<class F> int fp_convert(int to_mont, const void *a, void *o, size_t n)
{
    return run3<F>(a, nullptr, o, n, [&](F *da, F *, F *dо, cudaStream_t s) {
        k_fp_convert<F><<<(unsigned)((n + 63) / 64), 64, 0, s>>>(to_mont, da, dо, n);
    });
}
template<class F> int group_op(int op, const void *a, const void *b, void *o, size_t n)
{
    typedef msm::RefPoint<F> P;
    return run3<P>(a, b, o, n, [&](P *da, P *db, P *dо, cudaStream_t s) {
        k_group_op<F><<<(unsigned)((n + 63) / 64), 64, 0, s>>>(op, da, db, dо, n);
    });
}

// The digits the MSM's sort sees: k_prepare (scalar -> t = s + K) on the device, then window_digit for every window,
// for a forced window width c (field_get_signed_digit, field_utils.tcc:82-115, is what they must equal).
static __global__ void k_dump_digits(const uint32_t *t, int NL, msm::Geometry g, int64_t *out)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < g.n) {
        for (int w = 0; w < g.W; ++w) {
            out[i * g.W + w] = msm::window_digit(t + i * NL, NL, g, w);
        }
    }
}

rt::DevBuf g_dt, g_dd;

template<class Fr> int signed_digits(const void *scalars, size_t n, int c, int64_t *out, int *W_out)
{
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    if (c < 2 || c > 24 || n == 0) {
        return rt::set_error(B200_ERR_ARG, "window width %d outside [2, 24] or no scalars", c);
    }
    const msm::Geometry g = msm::make_geometry(n, Fr::params::BITS, c);
    const size_t bytes = n * sizeof(Fr);
    if (g_a.reserve(bytes) || g_dt.reserve(bytes + 16) || g_dd.reserve(n * g.W * sizeof(int64_t))) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(g_a.p, scalars, bytes, cudaMemcpyHostToDevice, s));
    msm::k_prepare<Fr><<<(unsigned)((n + 63) / 64), 64, 0, s>>>(g_a.as<Fr>(), g_dt.as<uint32_t>(), n,
                                                                msm::make_offset_constant<Fr::N>(g), nullptr);
    B200_LAUNCH_CHECK();
    k_dump_digits<<<(unsigned)((n + 63) / 64), 64, 0, s>>>(g_dt.as<uint32_t>(), Fr::N, g, g_dd.as<int64_t>());
    B200_LAUNCH_CHECK();
    B200_CUDA_CHECK(cudaMemcpyAsync(out, g_dd.p, n * g.W * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    *W_out = g.W;
    return B200_OK;
}

} // namespace

extern "C" {

// field: 0 = Fr, 1 = Fq, 2 = Fq2 (G2's coordinate field; bw6_761: Fq)
__attribute__((visibility("default"))) int b200_test_field_op(int curve, int field, int op, const void *a,
                                                              const void *b, void *o, size_t n)
{
    switch (curve * 4 + field) {
    case 0: return field_op<alt_bn128_fr>(op, a, b, o, n);
    case 1: return field_op<alt_bn128_fq>(op, a, b, o, n);
    case 2: return field_op<alt_bn128_fq2>(op, a, b, o, n);
    case 4: return field_op<bls12_377_fr>(op, a, b, o, n);
    case 5: return field_op<bls12_377_fq>(op, a, b, o, n);
    case 6: return field_op<bls12_377_fq2>(op, a, b, o, n);
    case 8: return field_op<bw6_761_fr>(op, a, b, o, n);
    case 9:
    case 10: return field_op<bw6_761_fq>(op, a, b, o, n);
    }
    return rt::set_error(B200_ERR_ARG, "bad curve/field");
}

__attribute__((visibility("default"))) int b200_test_fp_convert(int curve, int field, int to_mont, const void *a,
                                                                void *o, size_t n)
{
    switch (curve * 4 + field) {
    case 0: return fp_convert<alt_bn128_fr>(to_mont, a, o, n);
    case 1: return fp_convert<alt_bn128_fq>(to_mont, a, o, n);
    case 4: return fp_convert<bls12_377_fr>(to_mont, a, o, n);
    case 5: return fp_convert<bls12_377_fq>(to_mont, a, o, n);
    case 8: return fp_convert<bw6_761_fr>(to_mont, a, o, n);
    case 9: return fp_convert<bw6_761_fq>(to_mont, a, o, n);
    }
    return rt::set_error(B200_ERR_ARG, "bad curve/field");
}

__attribute__((visibility("default"))) int b200_test_group_op(int curve, int group, int op, const void *a,
                                                              const void *b, void *o, size_t n)
{
    switch (curve * 4 + group) {
    case 1: return group_op<alt_bn128_fq>(op, a, b, o, n);
    case 2: return group_op<alt_bn128_fq2>(op, a, b, o, n);
    case 5: return group_op<bls12_377_fq>(op, a, b, o, n);
    case 6: return group_op<bls12_377_fq2>(op, a, b, o, n);
    case 9:
    case 10: return group_op<bw6_761_fq>(op, a, b, o, n);
    }
    return rt::set_error(B200_ERR_ARG, "bad curve/group");
}

// out: n * W signed digits (row-major, W = ceil((bits + 1) / c) returned in *W_out) as the device's digit extraction
// produces them for window width c.
__attribute__((visibility("default"))) int b200_test_signed_digits(int curve, const void *scalars, size_t n, int c,
                                                                   int64_t *out, int *W_out)
{
    switch (curve) {
    case 0: return signed_digits<alt_bn128_fr>(scalars, n, c, out, W_out);
    case 1: return signed_digits<bls12_377_fr>(scalars, n, c, out, W_out);
    case 2: return signed_digits<bw6_761_fr>(scalars, n, c, out, W_out);
    }
    return rt::set_error(B200_ERR_ARG, "bad curve");
}
}
