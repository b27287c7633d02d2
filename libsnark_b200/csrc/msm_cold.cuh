// msm_cold.cuh — stages 6-7 of the MSM (bucket reduction, window sums): kernels and their
// launcher. Included only by msm_cold_<curve>_<group>.cu, which are compiled with
// -DB200_OUTLINE_MUL (see fp.cuh): these kernels are chains of full group additions executed by
// comparatively few warps, so compact code beats inlined code.
#pragma once
#include "msm.cuh"
#include "msm_affine.cuh"
#include "runtime.hpp"

namespace b200
{
namespace msm
{

// Warp-level exchange of a point (32-bit shuffles of every limb).
template<class F> __device__ __forceinline__ XYZZ<F> shfl_down_point(const XYZZ<F> &p, int delta)
{
    XYZZ<F> r;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(&p);
    uint32_t *dst = reinterpret_cast<uint32_t *>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(XYZZ<F>) / 4); ++i) {
        dst[i] = __shfl_down_sync(0xffffffffu, src[i], delta);
    }
    return r;
}

// 5c. Buckets that were cut into several tasks (more than SEG entries: heavy buckets of skewed
// scalar sets, and the narrow top window) get their partial sums merged here, one warp per
// bucket: lanes stride over the partials, then a shuffle tree. The total lands in the bucket's
// first task slot, which is the only one the reduction reads.
template<class F>
__global__ void __launch_bounds__(128) k_collapse(const uint32_t *multi_count, const uint32_t *multi_list,
                                                  const uint32_t *task_off, XYZZ<F> *partial)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nmulti = *multi_count;
    for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < nmulti; i += warps) {
        const uint32_t b = multi_list[i];
        const uint32_t t0 = task_off[b], t1 = task_off[b + 1];
        XYZZ<F> acc = XYZZ<F>::inf();
        for (uint32_t t = t0 + lane; t < t1; t += 32) {
            const XYZZ<F> s = load_vec(partial + t);
            add_cold(acc, s);
        }
        for (int delta = 16; delta >= 1; delta >>= 1) {
            const XYZZ<F> other = shfl_down_point(acc, delta);
            add_cold(acc, other);
        }
        if (lane == 0) {
            store_vec(partial + t0, acc);
        }
    }
}

// 6. One thread per CHUNK buckets (reduce_chunk_body), then a shared-memory tree over the CTA:
// block_out[blockIdx] = sum of the CTA's weighted chunk sums. g.rb threads, all in one window.
template<class F>
__global__ void __launch_bounds__(REDUCE_BLOCK) k_reduce_chunks(
    Geometry g, const uint32_t *task_off, const XYZZ<F> *partial, XYZZ<F> *block_out)
{
    extern __shared__ uint4 reduce_smem[];
    XYZZ<F> *sh = reinterpret_cast<XYZZ<F> *>(reduce_smem);
    const size_t gidx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    XYZZ<F> mine;
    reduce_chunk_body<F>(gidx, g, task_off, partial, &mine);
    store_vec(sh + threadIdx.x, mine);
    __syncthreads();
    for (int stride = (int)blockDim.x / 2; stride >= 1; stride >>= 1) {
        if ((int)threadIdx.x < stride) {
            XYZZ<F> a = load_vec(sh + threadIdx.x);
            const XYZZ<F> b = load_vec(sh + threadIdx.x + stride);
            add_cold(a, b);
            store_vec(sh + threadIdx.x, a);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        store_vec(block_out + blockIdx.x, load_vec(sh));
    }
}

// 7. One CTA per window sums the window's g.opw block results.
constexpr int WSUM_THREADS = 32;
template<class F>
__global__ void __launch_bounds__(WSUM_THREADS) k_window_sum(Geometry g, const XYZZ<F> *block_out, XYZZ<F> *windows)
{
    const int w = blockIdx.x;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (uint32_t i = threadIdx.x; i < g.opw; i += WSUM_THREADS) {
        const XYZZ<F> s = load_vec(block_out + (size_t)w * g.opw + i);
        add_cold(acc, s);
    }
    for (int delta = 16; delta >= 1; delta >>= 1) {
        const XYZZ<F> other = shfl_down_point(acc, delta);
        add_cold(acc, other);
    }
    if (threadIdx.x == 0) {
        store_vec(windows + w, acc);
    }
}

template<class F>
__global__ void __launch_bounds__(128) k_build_table(size_t n, int c, int W, Affine<F> *table, const uint8_t *inf, uint32_t *fail)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        build_table_body<F>(i, n, c, W, table, inf, fail);
    }
}

template<class F> int table_launch(size_t n, int c, int W, Affine<F> *table, const uint8_t *inf, uint32_t *fail)
{
    k_build_table<F><<<(unsigned)((n + 127) / 128), 128, 0, rt::stream()>>>(n, c, W, table, inf, fail);
    B200_LAUNCH_CHECK();
    return B200_OK;
}

// ---- the upper levels of an affine round's inversion tree (msm_affine.cuh): few threads, chains of dependent products ----
// level 1: elements = T1 (lv.n1 of them), threads = lv.n2; level 2: elements = T2 (lv.n2), threads = lv.n3.
template<class F>
__global__ void __launch_bounds__(128) k_aff_up(const uint32_t *total_entries, uint32_t rounds, int level, const F *E, F *Pre, F *T)
{
    const AffLevels lv = aff_levels(*total_entries, rounds);
    const uint32_t count = level == 1 ? lv.n1 : lv.n2, threads = level == 1 ? lv.n2 : lv.n3;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < threads) {
        aff_up_body<F>(t, count, threads, E, Pre, T);
    }
}
template<class F>
__global__ void __launch_bounds__(128) k_aff_down(const uint32_t *total_entries, uint32_t rounds, int level, F *E, const F *Pre, const F *Inv)
{
    const AffLevels lv = aff_levels(*total_entries, rounds);
    const uint32_t count = level == 1 ? lv.n1 : lv.n2, threads = level == 1 ? lv.n2 : lv.n3;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < threads) {
        aff_down_body<F>(t, count, threads, E, Pre, Inv);
    }
}
// One inversion per element of the top level, 32 threads per CTA so that the few hundred warps spread over all SMs.
template<class F> __global__ void __launch_bounds__(32) k_aff_top(const uint32_t *total_entries, uint32_t rounds, F *T3)
{
    const AffLevels lv = aff_levels(*total_entries, rounds);
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < lv.n3) {
        store_vec(T3 + t, load_vec(T3 + t).inverse_fast());
    }
}

// Inverts the n1 thread totals in T1 in place (the tree is the same size in every round: one element per level-0 thread).
template<class F>
int affine_tree_launch(const uint32_t *total_entries, uint32_t rounds, uint32_t max_entries, F *T1, F *P1, F *T2, F *P2, F *T3,
                       cudaStream_t s)
{
    const AffLevels mx = aff_levels(max_entries, rounds);
    const auto grid = [](uint32_t n, uint32_t b) { return (n + b - 1) / b; };
    k_aff_up<F><<<grid(mx.n2, 128), 128, 0, s>>>(total_entries, rounds, 1, T1, P1, T2);
    B200_LAUNCH_CHECK();
    k_aff_up<F><<<grid(mx.n3, 128), 128, 0, s>>>(total_entries, rounds, 2, T2, P2, T3);
    B200_LAUNCH_CHECK();
    k_aff_top<F><<<grid(mx.n3, 32), 32, 0, s>>>(total_entries, rounds, T3);
    B200_LAUNCH_CHECK();
    k_aff_down<F><<<grid(mx.n3, 128), 128, 0, s>>>(total_entries, rounds, 2, T2, P2, T3);
    B200_LAUNCH_CHECK();
    k_aff_down<F><<<grid(mx.n2, 128), 128, 0, s>>>(total_entries, rounds, 1, T1, P1, T2);
    B200_LAUNCH_CHECK();
    return B200_OK;
}

template<class F>
int reduce_launch(const Geometry &g, const uint32_t *task_off, const uint32_t *multi_count,
                  const uint32_t *multi_list, XYZZ<F> *partial, XYZZ<F> *block_out, XYZZ<F> *windows_dev, cudaStream_t s)
{
    k_collapse<F><<<148 * 4, 128, 0, s>>>(multi_count, multi_list, task_off, partial);
    B200_LAUNCH_CHECK();
    const size_t nchunks = (size_t)g.W * g.cpw;
    const size_t smem = (size_t)g.rb * sizeof(XYZZ<F>);
    static bool attr_set = false;
    if (!attr_set) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(k_reduce_chunks<F>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)(REDUCE_BLOCK * sizeof(XYZZ<F>))));
        attr_set = true;
    }
    k_reduce_chunks<F><<<(unsigned)(nchunks / g.rb), g.rb, smem, s>>>(g, task_off, partial, block_out);
    B200_LAUNCH_CHECK();
    k_window_sum<F><<<g.W, WSUM_THREADS, 0, s>>>(g, block_out, windows_dev);
    B200_LAUNCH_CHECK();
    return B200_OK;
}

} // namespace msm
} // namespace b200
