// bench_kernels.cu — integer-pipe peak microbenchmarks (the MSM roofline denominator,
// "peak_imad measured on the box by a dependent-free IMAD microbenchmark").
// Two rates: 32-bit IMAD (the pipe's issue rate, 64 per clock per SM) and the 32x32+64 wide
// multiply-add IMAD.WIDE.U32(.X), which holds the pipe for two issue slots (32 per clock per SM)
// and is what a multi-limb Montgomery product consists of. tools/microbench/imad_bench.cu has
// the longer set of variants.
#include "runtime.hpp"

namespace
{

// 32-bit IMAD (mad.lo.u32), 8 chains per thread. The multiplier operands come from the other
// accumulators, i.e. they are loop-VARIANT: with invariant operands ptxas hoists the product out
// of the loop and what remains (and what an earlier version of this file measured) is IADD3.
template<int UNROLL> __global__ void __launch_bounds__(256) k_imad_lo(uint32_t *out, uint32_t a, uint32_t b, int iters)
{
    uint32_t acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        acc[k] = (threadIdx.x + k) * 0x9e3779b9u + a + b * blockIdx.x;
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(acc[k]) : "r"(acc[(k + 1) % 8]), "r"(acc[(k + 3) % 8]));
            }
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        r ^= acc[k];
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// The pattern the Montgomery product is made of: one carry chain of mad.lo.cc / madc.hi.cc
// pairs across 8 limbs (ptxas fuses each pair into IMAD.WIDE.U32.X with carry in and out).
// Two independent rows per iteration, like the even/odd accumulators of fp.cuh.
template<int UNROLL> __global__ void __launch_bounds__(256) k_imad_carry(uint32_t *out, uint32_t a, uint32_t b, int iters)
{
    uint32_t e[8], o[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        e[k] = threadIdx.x + k;
        o[k] = blockIdx.x + k;
    }
    uint32_t x = a + threadIdx.x, y = b + blockIdx.x;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            asm volatile(
                "mad.lo.cc.u32 %0, %8, %9, %0;\n\t"
                "madc.hi.cc.u32 %1, %8, %9, %1;\n\t"
                "madc.lo.cc.u32 %2, %8, %10, %2;\n\t"
                "madc.hi.cc.u32 %3, %8, %10, %3;\n\t"
                "madc.lo.cc.u32 %4, %9, %10, %4;\n\t"
                "madc.hi.cc.u32 %5, %9, %10, %5;\n\t"
                "madc.lo.cc.u32 %6, %10, %10, %6;\n\t"
                "madc.hi.u32 %7, %10, %10, %7;\n\t"
                : "+r"(e[0]), "+r"(e[1]), "+r"(e[2]), "+r"(e[3]), "+r"(e[4]), "+r"(e[5]), "+r"(e[6]), "+r"(e[7])
                : "r"(x), "r"(y), "r"(a));
            asm volatile(
                "mad.lo.cc.u32 %0, %8, %9, %0;\n\t"
                "madc.hi.cc.u32 %1, %8, %9, %1;\n\t"
                "madc.lo.cc.u32 %2, %8, %10, %2;\n\t"
                "madc.hi.cc.u32 %3, %8, %10, %3;\n\t"
                "madc.lo.cc.u32 %4, %9, %10, %4;\n\t"
                "madc.hi.cc.u32 %5, %9, %10, %5;\n\t"
                "madc.lo.cc.u32 %6, %10, %10, %6;\n\t"
                "madc.hi.u32 %7, %10, %10, %7;\n\t"
                : "+r"(o[0]), "+r"(o[1]), "+r"(o[2]), "+r"(o[3]), "+r"(o[4]), "+r"(o[5]), "+r"(o[6]), "+r"(o[7])
                : "r"(y), "r"(x), "r"(b));
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        r ^= e[k] ^ o[k];
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r;
}

b200::rt::DevBuf g_buf;
double g_carry_rate = 0;

} // namespace

extern "C" int b200_imad_peak(int iters, double *wide_mads_per_s, double *lo_mads_per_s)
{
    using namespace b200;
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    if (iters <= 0) {
        iters = 4096;
    }
    cudaDeviceProp prop;
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    B200_CUDA_CHECK(cudaGetDeviceProperties(&prop, dev));
    const int blocks = prop.multiProcessorCount * 8, threads = 256;
    constexpr int UNROLL = 4;
    if (g_buf.reserve((size_t)blocks * threads * 8)) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    cudaEvent_t e0, e1;
    B200_CUDA_CHECK(cudaEventCreate(&e0));
    B200_CUDA_CHECK(cudaEventCreate(&e1));
    const double ops = (double)blocks * threads * (double)iters * UNROLL * 8;
    for (int which = 1; which < 3; ++which) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            B200_CUDA_CHECK(cudaEventRecord(e0, s));
            if (which == 1) {
                k_imad_lo<UNROLL><<<blocks, threads, 0, s>>>(g_buf.as<uint32_t>(), 12345u, 6789u, iters);
            } else {
                k_imad_carry<UNROLL><<<blocks, threads, 0, s>>>(g_buf.as<uint32_t>(), 12345u, 6789u, iters);
            }
            B200_LAUNCH_CHECK();
            B200_CUDA_CHECK(cudaEventRecord(e1, s));
            B200_CUDA_CHECK(cudaEventSynchronize(e1));
            float ms = 0;
            B200_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep > 0 && ms < best) {
                best = ms;
            }
        }
        const double rate = ops / (best * 1e-3);
        if (which == 2 && wide_mads_per_s) {
            *wide_mads_per_s = rate;
        }
        if (which == 1 && lo_mads_per_s) {
            *lo_mads_per_s = rate;
        }
        if (which == 2) {
            g_carry_rate = rate; // 8 fused wide multiply-adds per iteration as well (2 rows x 4 pairs)
        }
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return B200_OK;
}

// Wide multiply-adds per second when every one carries in and out (IMAD.WIDE.U32.X chains).
extern "C" __attribute__((visibility("default"))) double b200_imad_carry_peak(void) { return g_carry_rate; }
