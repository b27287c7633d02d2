// imad_bench.cu — integer-pipe microbenchmarks behind DESIGN.md §7 (stand-alone; not part of the library).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I libsnark_b200/csrc -I tools/microbench \
//        -o gpurun_out/imad_bench tools/microbench/imad_bench.cu && gpurun_out/imad_bench
// Every kernel keeps its multiplier operands loop-VARIANT (fed from the accumulators), otherwise ptxas
// hoists the product and the loop degenerates into IADD3s.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "curves.cuh"
#include "fpu.cuh" // tools/microbench/fpu.cuh (experiment, not in the library)

using namespace b200;

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int K = 8;

// A: IMAD.WIDE.U32 with 64-bit accumulate, operands varying
__global__ void __launch_bounds__(256) k_wide_acc(uint64_t *out, uint32_t seed, int iters)
{
    uint64_t acc[K];
    for (int k = 0; k < K; ++k) acc[k] = (uint64_t)(threadIdx.x + k) * 0x9e3779b97f4a7c15ull + seed;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const uint32_t b = (uint32_t)acc[(k + 1) % K];
                const uint32_t a = (uint32_t)(acc[(k + 3) % K] >> 32);
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"(b));
            }
        }
    }
    uint64_t r = 0;
    for (int k = 0; k < K; ++k) r ^= acc[k];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// B: signed IMAD.WIDE with 64-bit accumulate
__global__ void __launch_bounds__(256) k_wide_acc_s(uint64_t *out, uint32_t seed, int iters)
{
    int64_t acc[K];
    for (int k = 0; k < K; ++k) acc[k] = (int64_t)((uint64_t)(threadIdx.x + k) * 0x9e3779b97f4a7c15ull + seed);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const int32_t b = (int32_t)acc[(k + 1) % K];
                const int32_t a = (int32_t)(acc[(k + 3) % K] >> 32);
                asm volatile("mad.wide.s32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"(b));
            }
        }
    }
    uint64_t r = 0;
    for (int k = 0; k < K; ++k) r ^= (uint64_t)acc[k];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// C: 32-bit IMAD (mad.lo) with varying operands
__global__ void __launch_bounds__(256) k_lo(uint64_t *out, uint32_t seed, int iters)
{
    uint32_t acc[K];
    for (int k = 0; k < K; ++k) acc[k] = (threadIdx.x + k) * 0x9e3779b9u + seed;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(acc[k]) : "r"(acc[(k + 1) % K]), "r"(acc[(k + 3) % K]));
            }
        }
    }
    uint32_t r = 0;
    for (int k = 0; k < K; ++k) r ^= acc[k];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// D: carry chain (mad.lo.cc / madc.hi.cc pairs -> IMAD.WIDE.U32.X), operands varying
__global__ void __launch_bounds__(256) k_carry(uint64_t *out, uint32_t seed, int iters)
{
    uint32_t e[8], o[8];
    for (int k = 0; k < 8; ++k) { e[k] = (threadIdx.x + k) * 0x9e3779b9u + seed; o[k] = (blockIdx.x + k) * 0x85ebca6bu + seed; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            asm volatile(
                "mad.lo.cc.u32 %0, %8, %9, %0;\n\t madc.hi.cc.u32 %1, %8, %9, %1;\n\t"
                "madc.lo.cc.u32 %2, %8, %10, %2;\n\t madc.hi.cc.u32 %3, %8, %10, %3;\n\t"
                "madc.lo.cc.u32 %4, %9, %10, %4;\n\t madc.hi.cc.u32 %5, %9, %10, %5;\n\t"
                "madc.lo.cc.u32 %6, %10, %11, %6;\n\t madc.hi.u32 %7, %10, %11, %7;\n\t"
                : "+r"(e[0]), "+r"(e[1]), "+r"(e[2]), "+r"(e[3]), "+r"(e[4]), "+r"(e[5]), "+r"(e[6]), "+r"(e[7])
                : "r"(o[0]), "r"(o[2]), "r"(o[4]), "r"(o[6]));
            asm volatile(
                "mad.lo.cc.u32 %0, %8, %9, %0;\n\t madc.hi.cc.u32 %1, %8, %9, %1;\n\t"
                "madc.lo.cc.u32 %2, %8, %10, %2;\n\t madc.hi.cc.u32 %3, %8, %10, %3;\n\t"
                "madc.lo.cc.u32 %4, %9, %10, %4;\n\t madc.hi.cc.u32 %5, %9, %10, %5;\n\t"
                "madc.lo.cc.u32 %6, %10, %11, %6;\n\t madc.hi.u32 %7, %10, %11, %7;\n\t"
                : "+r"(o[0]), "+r"(o[1]), "+r"(o[2]), "+r"(o[3]), "+r"(o[4]), "+r"(o[5]), "+r"(o[6]), "+r"(o[7])
                : "r"(e[1]), "r"(e[3]), "r"(e[5]), "r"(e[7]));
        }
    }
    uint32_t r = 0;
    for (int k = 0; k < 8; ++k) r ^= e[k] ^ o[k];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// G: FP64 fused multiply-add, 8 independent chains, operands varying
__global__ void __launch_bounds__(256) k_dfma(uint64_t *out, uint32_t seed, int iters)
{
    double acc[K];
    for (int k = 0; k < K; ++k) acc[k] = 1.0 + 1e-9 * (threadIdx.x + k + seed);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(acc[k]) : "d"(acc[(k + 1) % K]), "d"(acc[(k + 3) % K]));
            }
        }
    }
    double r = 0;
    for (int k = 0; k < K; ++k) r += acc[k];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = (uint64_t)__double_as_longlong(r);
}

// H: IMAD.WIDE.U32 with a zero addend (independent products, folded by XOR on the ALU pipe)
__global__ void __launch_bounds__(256) k_wide_rz(uint64_t *out, uint32_t seed, int iters)
{
    uint64_t acc[K];
    for (int k = 0; k < K; ++k) acc[k] = (uint64_t)(threadIdx.x + k) * 0x9e3779b97f4a7c15ull + seed;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                uint64_t p;
                asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p) : "r"((uint32_t)acc[(k + 1) % K]), "r"((uint32_t)(acc[(k + 3) % K] >> 32)));
                acc[k] ^= p;
            }
        }
    }
    uint64_t r = 0;
    for (int k = 0; k < K; ++k) r ^= acc[k];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// E/F: field multiplication chains, two independent chains per thread
template<class F> __global__ void __launch_bounds__(128) k_fp_mul(uint64_t *out, const F *in, int iters)
{
    F a = in[threadIdx.x], b = in[threadIdx.x + 128], c = in[threadIdx.x + 256], d = in[threadIdx.x + 384];
    for (int it = 0; it < iters; ++it) {
        a = F::mul_inline(a, b);
        c = F::mul_inline(c, d);
        b = F::mul_inline(b, a);
        d = F::mul_inline(d, c);
    }
    a = a + b + c + d;
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = a.v[0];
}
// one dependent chain per thread / three independent chains per thread (ILP probe)
template<class F> __global__ void __launch_bounds__(128) k_fp_mul1(uint64_t *out, const F *in, int iters)
{
    F a = in[threadIdx.x], b = in[threadIdx.x + 128];
    for (int it = 0; it < iters; ++it) {
        a = F::mul_inline(a, b);
        b = F::mul_inline(b, a);
        a = F::mul_inline(a, b);
        b = F::mul_inline(b, a);
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = (a + b).v[0];
}
template<class F> __global__ void __launch_bounds__(128) k_fp_mul3(uint64_t *out, const F *in, int iters)
{
    F a = in[threadIdx.x], b = in[threadIdx.x + 128], c = in[threadIdx.x + 256], d = in[threadIdx.x + 384];
    F e = a + c, f = b + d;
    for (int it = 0; it < iters; ++it) {
        a = F::mul_inline(a, b);
        c = F::mul_inline(c, d);
        e = F::mul_inline(e, f);
        b = b + a;
        d = d + c;
        f = f + e;
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = (a + c + e).v[0];
}
template<class P> __global__ void __launch_bounds__(128) k_fpu_mul(uint64_t *out, const Fp<P> *in, int iters)
{
    typedef FpUOf<P> U;
    U a = U::from_mont32_reduced(in[threadIdx.x]), b = U::from_mont32_reduced(in[threadIdx.x + 128]);
    U c = U::from_mont32_reduced(in[threadIdx.x + 256]), d = U::from_mont32_reduced(in[threadIdx.x + 384]);
    for (int it = 0; it < iters; ++it) {
        a = a * b;
        c = c * d;
        b = b * a;
        d = d * c;
    }
    a = (a + b) + (c + d);
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = (uint64_t)(uint32_t)a.v[0];
}

template<class Fn> static double time_ms(Fn fn)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    fn(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(e0); fn(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    return best;
}

int main()
{
    cudaDeviceProp pr; CHECK(cudaGetDeviceProperties(&pr, 0));
    const int sms = pr.multiProcessorCount;
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("device %s, %d SMs, %.0f MHz max\n", pr.name, sms, clk / 1e3);
    uint64_t *out; CHECK(cudaMalloc(&out, (size_t)sms * 8 * 256 * 8));
    const int iters = 4096;
    const int grid = sms * 8;
    const double per_sm_clk = 1.0 / (sms * (clk * 1e3));
    {
        double ms = time_ms([&] { k_wide_acc<<<grid, 256>>>(out, 1, iters); });
        double ops = (double)grid * 256 * iters * 4 * K;
        printf("A IMAD.WIDE.U32 acc64 (varying operands): %.3e /s = %.1f per clk per SM\n", ops / ms * 1e3, ops / ms * 1e3 * per_sm_clk);
    }
    {
        double ms = time_ms([&] { k_wide_acc_s<<<grid, 256>>>(out, 1, iters); });
        double ops = (double)grid * 256 * iters * 4 * K;
        printf("B IMAD.WIDE (signed) acc64:               %.3e /s = %.1f per clk per SM\n", ops / ms * 1e3, ops / ms * 1e3 * per_sm_clk);
    }
    {
        double ms = time_ms([&] { k_lo<<<grid, 256>>>(out, 1, iters); });
        double ops = (double)grid * 256 * iters * 4 * K;
        printf("C IMAD (32-bit lo):                       %.3e /s = %.1f per clk per SM\n", ops / ms * 1e3, ops / ms * 1e3 * per_sm_clk);
    }
    {
        double ms = time_ms([&] { k_carry<<<grid, 256>>>(out, 1, iters); });
        double ops = (double)grid * 256 * iters * 2 * 2 * 4; // 4 wide multiply-adds per asm block
        printf("D IMAD.WIDE.U32.X carry chain:            %.3e /s = %.1f per clk per SM\n", ops / ms * 1e3, ops / ms * 1e3 * per_sm_clk);
    }
    {
        double ms = time_ms([&] { k_dfma<<<grid, 256>>>(out, 1, iters); });
        double ops = (double)grid * 256 * iters * 4 * K;
        printf("G DFMA (fp64 fma):                        %.3e /s = %.1f per clk per SM\n", ops / ms * 1e3, ops / ms * 1e3 * per_sm_clk);
    }
    {
        double ms = time_ms([&] { k_wide_rz<<<grid, 256>>>(out, 1, iters); });
        double ops = (double)grid * 256 * iters * 4 * K;
        printf("H IMAD.WIDE.U32 zero addend + 2 LOP3:     %.3e /s = %.1f per clk per SM\n", ops / ms * 1e3, ops / ms * 1e3 * per_sm_clk);
    }
    // field multiplications
    typedef alt_bn128_fq F;
    F *in; CHECK(cudaMalloc(&in, 512 * sizeof(F)));
    {
        F h[512];
        for (int i = 0; i < 512; ++i) { h[i] = F::one(); h[i].v[0] += i + 1; h[i].v[3] ^= 0x1234567u * i; }
        CHECK(cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice));
    }
    const int mit = 2048;
    for (int cta_per_sm : {1, 2, 4}) {
        const int g2 = sms * cta_per_sm;
        double ms = time_ms([&] { k_fp_mul1<F><<<g2, 128>>>(out, in, mit); });
        double muls = (double)g2 * 128 * mit * 4;
        printf("I Fp<254> ONE chain/thread,   %d CTA/SM x128 (%d warps/SMSP): %.2f Tmad/s\n", cta_per_sm, cta_per_sm, muls * 136 / ms * 1e3 / 1e12);
        ms = time_ms([&] { k_fp_mul3<F><<<g2, 128>>>(out, in, mit); });
        muls = (double)g2 * 128 * mit * 3;
        printf("J Fp<254> THREE chains/thread, %d CTA/SM x128 (%d warps/SMSP): %.2f Tmad/s\n", cta_per_sm, cta_per_sm, muls * 136 / ms * 1e3 / 1e12);
    }
    for (int cta_per_sm : {1, 2, 4, 8}) {
        const int g2 = sms * cta_per_sm;
        double ms = time_ms([&] { k_fp_mul<F><<<g2, 128>>>(out, in, mit); });
        double muls = (double)g2 * 128 * mit * 4;
        printf("E Fp<254> 8x32 Montgomery (IMAD.X), %d CTA/SM x128: %.3e modmul/s  (%.2f Tmad/s at 136/modmul)\n", cta_per_sm, muls / ms * 1e3, muls * 136 / ms * 1e3 / 1e12);
        ms = time_ms([&] { k_fpu_mul<alt_bn128_fq_params><<<g2, 128>>>(out, in, mit); });
        printf("F FpU<254> 9x29 unsaturated,        %d CTA/SM x128: %.3e modmul/s  (%.2f Tmad/s-equivalent)\n", cta_per_sm, muls / ms * 1e3, muls * 136 / ms * 1e3 / 1e12);
    }
    CHECK(cudaDeviceSynchronize());
    printf("done\n");
    return 0;
}
