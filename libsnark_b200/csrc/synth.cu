// synth.cu — synthetic input generator for bench.py / smoke(): n distinct affine points
// [k0 + i] * G in the reference layout, produced on the device (the GPU box has no
// /root/reference, and the product may not call the oracle). Compiled with -DB200_OUTLINE_MUL.
#include "msm.cuh"
#include "runtime.hpp"

using namespace b200;

namespace
{

constexpr int PER_THREAD = 64;

template<class F, class Gen> __global__ void __launch_bounds__(64) k_make_bases(uint64_t k0, size_t n, msm::RefPoint<F> *out)
{
    const size_t start = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * PER_THREAD;
    if (start >= n) {
        return;
    }
    Affine<F> G;
    uint32_t *gw = reinterpret_cast<uint32_t *>(&G);
#pragma unroll 1
    for (int i = 0; i < (int)(sizeof(Affine<F>) / 4); ++i) {
        gw[i] = Gen::xy(i);
    }
    const XYZZ<F> Gx = XYZZ<F>::from_affine(G);
    const uint64_t k = k0 + start; // >= 1
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int bit = 63; bit >= 0; --bit) {
        dbl_cold(acc);
        if ((k >> bit) & 1ull) {
            add_cold(acc, Gx);
        }
    }
    for (int j = 0; j < PER_THREAD && start + j < n; ++j) {
        msm::store_vec(out + start + j, msm::normalise(acc));
        add_cold(acc, Gx);
    }
}

rt::DevBuf g_pts;

template<class F, class Gen> int make_bases(uint64_t k0, size_t n, void *out_host)
{
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    const size_t bytes = n * sizeof(msm::RefPoint<F>);
    if (g_pts.reserve(bytes)) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    const size_t threads = (n + PER_THREAD - 1) / PER_THREAD;
    k_make_bases<F, Gen><<<(unsigned)((threads + 63) / 64), 64, 0, s>>>(k0, n, g_pts.as<msm::RefPoint<F>>());
    B200_LAUNCH_CHECK();
    B200_CUDA_CHECK(cudaMemcpyAsync(out_host, g_pts.p, bytes, cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

} // namespace

// out_host[i] = [k0 + i] * generator, i < n, affine (Z = one), reference layout. k0 >= 1.
extern "C" __attribute__((visibility("default"))) int b200_make_bases(int curve, int group, uint64_t k0, size_t n,
                                                                      void *out_host)
{
    if (!out_host || k0 == 0) {
        return rt::set_error(B200_ERR_ARG, "null output or k0 == 0");
    }
    if (n == 0) {
        return B200_OK;
    }
    switch (curve * 4 + group) {
    case 1: return make_bases<alt_bn128_fq, alt_bn128_g1_generator>(k0, n, out_host);
    case 2: return make_bases<alt_bn128_fq2, alt_bn128_g2_generator>(k0, n, out_host);
    case 5: return make_bases<bls12_377_fq, bls12_377_g1_generator>(k0, n, out_host);
    case 6: return make_bases<bls12_377_fq2, bls12_377_g2_generator>(k0, n, out_host);
    case 9: return make_bases<bw6_761_fq, bw6_761_g1_generator>(k0, n, out_host);
    case 10: return make_bases<bw6_761_fq, bw6_761_g2_generator>(k0, n, out_host);
    }
    return rt::set_error(B200_ERR_ARG, "bad curve/group");
}
