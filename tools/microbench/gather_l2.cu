// gather_l2.cu — does the 126 MB L2 turn the second gather of the affine rounds into L2 hits if round 1 runs one table
// row (64 MiB) at a time? Stand-alone model of the memory side of k_aff_up_round (32 of every
// 64-byte point, one 32-byte prefix written per pair) and k_aff_round (all 64 bytes, prefix read back, 64-byte result
// written per pair) over a 15-row table of 2^20 points with random entries:
//   whole : up over all rows, then down over all rows (what the library does)
//   rows  : for each group of G rows: up(group) then down(group)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o gpurun_out/gather_l2 tools/microbench/gather_l2.cu
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

// MODE 0: plain ld.global; 1: ld.global.nc (__ldg, what the library's load_vec_ro emits); 2: ld.global.nc.L2::64B;
// 3: ld.global.L1::no_allocate; 4: ld.global.cg
template<int MODE> __device__ __forceinline__ uint4 ld16(const uint4 *p)
{
    uint4 v;
    if (MODE == 0) {
        v = *p;
    } else if (MODE == 1) {
        v = __ldg(p);
    } else if (MODE == 2) {
        asm volatile("ld.global.nc.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    } else if (MODE == 3) {
        asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    } else {
        asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    }
    return v;
}
template<int MODE>
__global__ void __launch_bounds__(128) k_up_mode(const uint4 *table, const uint32_t *entries, uint64_t npairs, uint4 *pex)
{
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    const uint32_t ia = entries[2 * p], ib = entries[2 * p + 1];
    const uint4 a0 = ld16<MODE>(table + (uint64_t)ia * 4), a1 = ld16<MODE>(table + (uint64_t)ia * 4 + 1);
    const uint4 b0 = ld16<MODE>(table + (uint64_t)ib * 4), b1 = ld16<MODE>(table + (uint64_t)ib * 4 + 1);
    pex[p * 2] = make_uint4(a0.x ^ b0.x, a0.y + b0.y, a0.z ^ b0.z, a0.w + b0.w);
    pex[p * 2 + 1] = make_uint4(a1.x ^ b1.x, a1.y + b1.y, a1.z ^ b1.z, a1.w + b1.w);
}
// the library's thread shape: one thread walks 16 consecutive pairs (4 blocks of 8 entries, consecutive threads own
// consecutive blocks), loads issued 4 pairs ahead
template<int MODE>
__global__ void __launch_bounds__(128, 4) k_up_owner(const uint4 *table, const uint32_t *entries, uint64_t nthreads, uint4 *pex)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nthreads) return;
    uint4 acc0 = make_uint4(1, 2, 3, 4), acc1 = acc0;
    for (int jj = 0; jj < 4; ++jj) {
        const uint64_t m = (uint64_t)jj * nthreads + t;   // block m: entries [8 m, 8 m + 8)
        uint4 x[4][4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t ia = entries[8 * m + 2 * u], ib = entries[8 * m + 2 * u + 1];
            x[u][0] = ld16<MODE>(table + (uint64_t)ia * 4);
            x[u][1] = ld16<MODE>(table + (uint64_t)ia * 4 + 1);
            x[u][2] = ld16<MODE>(table + (uint64_t)ib * 4);
            x[u][3] = ld16<MODE>(table + (uint64_t)ib * 4 + 1);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            acc0.x ^= x[u][0].x + x[u][2].x; acc0.y += x[u][0].y ^ x[u][2].y; acc0.z ^= x[u][0].z; acc0.w += x[u][2].w;
            acc1.x ^= x[u][1].x + x[u][3].x; acc1.y += x[u][1].y ^ x[u][3].y; acc1.z ^= x[u][1].z; acc1.w += x[u][3].w;
            pex[((uint64_t)u * 4 * nthreads + m) * 2] = acc0;       // element-major, as Pex
            pex[((uint64_t)u * 4 * nthreads + m) * 2 + 1] = acc1;
        }
    }
}
__global__ void __launch_bounds__(128) k_up(const uint4 *table, const uint32_t *entries, uint64_t first, uint64_t npairs, uint4 *pex)
{
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    const uint32_t ia = entries[first + 2 * p], ib = entries[first + 2 * p + 1];
    const uint4 a0 = table[(uint64_t)ia * 4], a1 = table[(uint64_t)ia * 4 + 1];   // x of a
    const uint4 b0 = table[(uint64_t)ib * 4], b1 = table[(uint64_t)ib * 4 + 1];   // x of b
    uint4 d0 = make_uint4(a0.x ^ b0.x, a0.y + b0.y, a0.z ^ b0.z, a0.w + b0.w);
    uint4 d1 = make_uint4(a1.x ^ b1.x, a1.y + b1.y, a1.z ^ b1.z, a1.w + b1.w);
    pex[(first / 2 + p) * 2] = d0;
    pex[(first / 2 + p) * 2 + 1] = d1;
}
__global__ void __launch_bounds__(128) k_down(const uint4 *table, const uint32_t *entries, uint64_t first, uint64_t npairs, const uint4 *pex,
                                             uint4 *out)
{
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    const uint32_t ia = entries[first + 2 * p], ib = entries[first + 2 * p + 1];
    uint4 acc[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint4 a = table[(uint64_t)ia * 4 + k], b = table[(uint64_t)ib * 4 + k];
        acc[k] = make_uint4(a.x ^ b.x, a.y + b.y, a.z ^ b.z, a.w + b.w);
    }
    const uint4 q0 = pex[(first / 2 + p) * 2], q1 = pex[(first / 2 + p) * 2 + 1];
    acc[0].x ^= q0.x;
    acc[1].y += q1.y;
#pragma unroll
    for (int k = 0; k < 4; ++k) out[(first / 2 + p) * 4 + k] = acc[k];
}

int main()
{
    const int R = 15;
    const uint64_t n = 1ull << 20, N = R * n;
    uint4 *table, *pex, *out;
    uint32_t *entries;
    CHECK(cudaMalloc(&table, N * 64));
    CHECK(cudaMalloc(&pex, N / 2 * 32));
    CHECK(cudaMalloc(&out, N / 2 * 64));
    CHECK(cudaMalloc(&entries, N * 4));
    CHECK(cudaMemset(table, 1, N * 64));
    std::vector<uint32_t> h(N);
    uint64_t s = 88172645463325252ull;
    for (int r = 0; r < R; ++r)
        for (uint64_t i = 0; i < n; ++i) {
            s ^= s << 13; s ^= s >> 7; s ^= s << 17;
            h[r * n + i] = (uint32_t)(r * n + (s % n));   // a random point of row r (sorted entries are window-major)
        }
    CHECK(cudaMemcpy(entries, h.data(), N * 4, cudaMemcpyHostToDevice));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    auto run = [&](int G, float *ms_out) -> int {   // G rows per group; G = R: the whole table at once
        for (int rep = 0; rep < 3; ++rep) {
            CHECK(cudaEventRecord(e0));
            for (int r0 = 0; r0 < R; r0 += G) {
                const int g = r0 + G <= R ? G : R - r0;
                const uint64_t first = r0 * n, np = g * n / 2;
                k_up<<<(unsigned)((np + 127) / 128), 128>>>(table, entries, first, np, pex);
                k_down<<<(unsigned)((np + 127) / 128), 128>>>(table, entries, first, np, pex, out);
            }
            CHECK(cudaEventRecord(e1));
            CHECK(cudaEventSynchronize(e1));
            CHECK(cudaEventElapsedTime(ms_out, e0, e1));
        }
        return 0;
    };
    // each pass alone, whole table
    float ms = 0;
    for (int rep = 0; rep < 3; ++rep) {
        CHECK(cudaEventRecord(e0));
        k_up<<<(unsigned)((N / 2 + 127) / 128), 128>>>(table, entries, 0, N / 2, pex);
        CHECK(cudaEventRecord(e1));
        CHECK(cudaEventSynchronize(e1));
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
    }
    printf("{\"up_whole_ms\": %.3f, ", ms);
    for (int rep = 0; rep < 3; ++rep) {
        CHECK(cudaEventRecord(e0));
        k_down<<<(unsigned)((N / 2 + 127) / 128), 128>>>(table, entries, 0, N / 2, pex, out);
        CHECK(cudaEventRecord(e1));
        CHECK(cudaEventSynchronize(e1));
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
    }
    printf("\"down_whole_ms\": %.3f", ms);
    // load flavours, one launch each (timed; `ncu --metrics dram__bytes_read.sum` gives the bytes)
#define MODE_RUN(K, MODE, GRID, ARG)                                                                \
    for (int rep = 0; rep < 2; ++rep) {                                                             \
        CHECK(cudaEventRecord(e0));                                                                 \
        K<MODE><<<(unsigned)(GRID), 128>>>(table, entries, ARG, pex);                               \
        CHECK(cudaEventRecord(e1));                                                                 \
        CHECK(cudaEventSynchronize(e1));                                                            \
        CHECK(cudaEventElapsedTime(&ms, e0, e1));                                                   \
    }                                                                                               \
    printf(", \"" #K "_mode" #MODE "_ms\": %.3f", ms);
    MODE_RUN(k_up_mode, 0, (N / 2 + 127) / 128, N / 2)
    MODE_RUN(k_up_mode, 1, (N / 2 + 127) / 128, N / 2)
    MODE_RUN(k_up_mode, 2, (N / 2 + 127) / 128, N / 2)
    MODE_RUN(k_up_mode, 3, (N / 2 + 127) / 128, N / 2)
    MODE_RUN(k_up_mode, 4, (N / 2 + 127) / 128, N / 2)
    {
        const uint64_t nth = N / 32;   // 4 blocks of 8 entries per thread
        MODE_RUN(k_up_owner, 0, (nth + 127) / 128, nth)
        MODE_RUN(k_up_owner, 1, (nth + 127) / 128, nth)
        MODE_RUN(k_up_owner, 4, (nth + 127) / 128, nth)
    }
    const int Gs[] = {15, 5, 3, 2, 1};
    for (int G : Gs) {
        if (run(G, &ms)) return 1;
        printf(", \"up_down_groups_of_%d_rows_ms\": %.3f", G, ms);
    }
    printf("}\n");
    return 0;
}
