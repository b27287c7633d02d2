// batch_exp.cu — kernels and C-ABI of the fixed-base batch exponentiation (batch_exp.cuh). Compiled with
// -DB200_OUTLINE_MUL like the other cold kernels.
#include <cstring>
#include <map>
#include <mutex>
#include <string>

#include "batch_exp.cuh"
#include "runtime.hpp"

using namespace b200;

namespace
{

template<class Fr, class F>
__global__ void __launch_bounds__(128) k_batch_exp(size_t n, const Fr *scalars, const Fr *coeff, const Affine<F> *table,
                                                   XYZZ<F> *out)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        bexp::exp_body<Fr, F>(i, scalars, coeff, table, out);
    }
}

template<class F> __global__ void __launch_bounds__(128) k_batch_normalise(size_t n, const XYZZ<F> *in, msm::RefPoint<F> *out)
{
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t * bexp::norm_group<F>() < n) {
        bexp::normalise_body<F>(t, n, in, out);
    }
}

std::mutex g_mu;
rt::DevBuf g_table, g_scalars, g_coeff, g_acc, g_out;
std::string g_table_key; // curve/group/g bytes of the table now in g_table

constexpr size_t CHUNK_POINTS = size_t(1) << 20; // scratch is sized for this many results at a time

template<class Fr, class F>
int batch_exp(int curve, int group, const void *g_host, const void *coeff, const void *scalars, size_t n, void *out)
{
    msm::RefPoint<F> g;
    memcpy(&g, g_host, sizeof(g));
    msm::RefPoint<F> *o = static_cast<msm::RefPoint<F> *>(out);
    if (g.z.is_zero()) { // zero * anything: every result is zero()
        msm::RefPoint<F> z;
        z.x = F::zero();
        z.y = F::one();
        z.z = F::zero();
        for (size_t i = 0; i < n; ++i) {
            o[i] = z;
        }
        return B200_OK;
    }
    if (!(g.z == F::one())) {
        return rt::set_error(B200_ERR_ARG, "the base of b200_batch_exp must be in special form (Z = one)");
    }
    cudaStream_t s = rt::stream();
    std::string key(1, (char)(curve * 4 + group));
    key.append(reinterpret_cast<const char *>(&g), sizeof(g));
    const size_t tbytes = (size_t)bexp::num_windows<Fr>() * bexp::PER_WINDOW * sizeof(Affine<F>);
    if (key != g_table_key) {
        std::vector<Affine<F>> table;
        if (!bexp::build_table<Fr, F>(g, table)) {
            return rt::set_error(B200_ERR_ARG, "a multiple of the base is the point at infinity: no window table");
        }
        if (g_table.reserve(tbytes)) {
            return B200_ERR_CUDA;
        }
        B200_CUDA_CHECK(cudaMemcpyAsync(g_table.p, table.data(), tbytes, cudaMemcpyHostToDevice, s));
        B200_CUDA_CHECK(cudaStreamSynchronize(s)); // `table` goes out of scope
        g_table_key = key;
    }
    const size_t cap = n < CHUNK_POINTS ? n : CHUNK_POINTS;
    if (g_scalars.reserve(cap * sizeof(Fr)) || g_coeff.reserve(sizeof(Fr)) || g_acc.reserve(cap * sizeof(XYZZ<F>)) ||
        g_out.reserve(cap * sizeof(msm::RefPoint<F>))) {
        return B200_ERR_CUDA;
    }
    const Fr *d_coeff = nullptr;
    if (coeff) {
        B200_CUDA_CHECK(cudaMemcpyAsync(g_coeff.p, coeff, sizeof(Fr), cudaMemcpyHostToDevice, s));
        d_coeff = g_coeff.as<Fr>();
    }
    constexpr int K = bexp::norm_group<F>();
    for (size_t first = 0; first < n; first += cap) {
        const size_t cnt = n - first < cap ? n - first : cap;
        B200_CUDA_CHECK(cudaMemcpyAsync(g_scalars.p, static_cast<const Fr *>(scalars) + first, cnt * sizeof(Fr),
                                        cudaMemcpyHostToDevice, s));
        k_batch_exp<Fr, F><<<(unsigned)((cnt + 127) / 128), 128, 0, s>>>(cnt, g_scalars.as<Fr>(), d_coeff,
                                                                         g_table.as<Affine<F>>(), g_acc.as<XYZZ<F>>());
        B200_LAUNCH_CHECK();
        const size_t threads = (cnt + K - 1) / K;
        k_batch_normalise<F><<<(unsigned)((threads + 127) / 128), 128, 0, s>>>(cnt, g_acc.as<XYZZ<F>>(),
                                                                               g_out.as<msm::RefPoint<F>>());
        B200_LAUNCH_CHECK();
        B200_CUDA_CHECK(cudaMemcpyAsync(o + first, g_out.p, cnt * sizeof(msm::RefPoint<F>), cudaMemcpyDeviceToHost, s));
        B200_CUDA_CHECK(cudaStreamSynchronize(s));
    }
    return B200_OK;
}

} // namespace

extern "C" __attribute__((visibility("default"))) int b200_batch_exp(int curve, int group, const void *g, const void *coeff,
                                                                     const void *scalars, size_t n, void *out)
{
    if (!g || (n && (!scalars || !out))) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    if (n == 0) {
        return B200_OK;
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    switch (curve * 4 + group) {
    case 1: return batch_exp<alt_bn128_fr, alt_bn128_fq>(curve, group, g, coeff, scalars, n, out);
    case 2: return batch_exp<alt_bn128_fr, alt_bn128_fq2>(curve, group, g, coeff, scalars, n, out);
    case 5: return batch_exp<bls12_377_fr, bls12_377_fq>(curve, group, g, coeff, scalars, n, out);
    case 6: return batch_exp<bls12_377_fr, bls12_377_fq2>(curve, group, g, coeff, scalars, n, out);
    case 9:
    case 10: return batch_exp<bw6_761_fr, bw6_761_fq>(curve, group, g, coeff, scalars, n, out);
    }
    return rt::set_error(B200_ERR_ARG, "bad curve/group");
}
