// msm_cold.cuh — stages 6-7 of the MSM (bucket reduction, window sums): kernels and their
// launcher. Included only by msm_cold_<curve>_<group>.cu, which are compiled with
// -DB200_OUTLINE_MUL (see fp.cuh): these kernels are chains of full group additions executed by
// comparatively few warps, so compact code beats inlined code.
#pragma once
#include <cstdlib>
#include <cstring>

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

// 6-7. Bucket reduction: window sum = sum_k (k + 1) S_k over the B buckets of a window (the reference's running sums,
// multiexp_accumulate_buckets, multiexp.tcc:91-125) as a two-level tree instead of a chain.
//
// Level 1 (k_reduce_rows): a CTA of RROW threads owns a ROW of RROW * cpt consecutive buckets. Thread t folds its cpt
// buckets serially (run_t = their sum, sum_t = their sum weighted 1..cpt: 2 additions per bucket), then the CTA forms
//     R = sum_t run_t,   Q0 = sum_t t * run_t,   T = sum_t sum_t
// without any chain longer than log2: a suffix scan inside groups of 16 threads (4 steps) turns t * run_t into a
// plain sum of suffix sums, which one tree over the CTA adds up (8 steps) while one half-warp repeats scan + tree on
// the 16 group totals. The row's weighted sum is  Q = T + cpt * Q0  (cpt a power of two: doublings).
// Level 2 (k_reduce_windows): one CTA per window does the same over the rows:  sum_h Q_h + (RROW cpt) sum_h h R_h.
// Every step is a full XYZZ addition per active thread; the dependent depth is ~30 additions per level where the chunked
// running sums needed 2 * chunk + a 16-bit double-and-add per thread (r1: 0.58 ms of a 3.9 ms MSM, 28-48 % of the
// wide-field MSMs).
template<class F> struct RowSmem {
    XYZZ<F> *v;  // [RROW]
    XYZZ<F> *gr; // [RGRP + 1] group totals; slot RGRP: their sum
    XYZZ<F> *q;  // [RROW] second operand (plain sum), only when WITH_SUM
};

// One point per thread (index t = threadIdx.x): R = sum_t P_t and Q0 = sum_t t * P_t; with WITH_SUM also T = sum_t S_t of a
// second point per thread, summed by a plain tree in the same steps (two independent additions per thread and step
// cost one addition's latency). Results are valid in thread 0.
template<class F, bool WITH_SUM>
__device__ __forceinline__ void cta_weighted_sum(const RowSmem<F> &sm, XYZZ<F> mine, XYZZ<F> qsum, XYZZ<F> &R, XYZZ<F> &Q0,
                                                 XYZZ<F> &T)
{
    const uint32_t l = threadIdx.x, b = l & (RGRP - 1);
    __syncthreads();
    store_vec(sm.v + l, mine);
    if (WITH_SUM) {
        store_vec(sm.q + l, qsum);
    }
    __syncthreads();
    // suffix scan inside every group of 16: mine = sum of the group's points at positions >= b
    uint32_t qd = RROW / 2; // distance of the next step of the plain tree
    for (uint32_t d = 1; d < RGRP; d <<= 1) {
        const bool take = b + d < RGRP;
        const bool qtake = WITH_SUM && l < qd;
        XYZZ<F> o, qo;
        if (take) {
            o = load_vec(sm.v + l + d);
        }
        if (qtake) {
            qo = load_vec(sm.q + l + qd);
        }
        __syncthreads();
        if (take) {
            add_ilp(mine, o);
            store_vec(sm.v + l, mine);
        }
        if (qtake) {
            add_ilp(qsum, qo);
            store_vec(sm.q + l, qsum);
        }
        qd >>= 1;
        __syncthreads();
    }
    // sum_t t P_t = sum over groups a of [ sum_{b >= 1} suffix_{a,b} + 16 a * total_a ]
    if (b == 0) {
        store_vec(sm.gr + (l >> 4), mine);
        mine = XYZZ<F>::inf();
        store_vec(sm.v + l, mine);
    }
    __syncthreads();
    // one tree over the CTA for the suffix sums; beside it the last half-warp runs scan + tree over the 16 group totals
    const bool gl = l >= RROW - RGRP; // lanes 16-31 of the last warp: idle in every tree step but the first
    const uint32_t gb = l - (RROW - RGRP);
    XYZZ<F> gm = XYZZ<F>::inf();
    if (gl) {
        gm = load_vec(sm.gr + gb);
    }
    int step = 0;
    for (uint32_t d = RROW / 2; d >= 1; d >>= 1, ++step) {
        XYZZ<F> o, qo;
        const bool tree = l < d;
        const bool qtake = WITH_SUM && qd >= 1 && l < qd;
        // group level: steps 0-3 suffix scan (distance 1, 2, 4, 8), steps 4-7 tree over positions >= 1 (distance 8, 4, 2, 1)
        const uint32_t gd = step < 4 ? (1u << step) : (8u >> (step - 4));
        const bool gtake = gl && (step < 4 ? gb + gd < RGRP : gb < gd);
        if (tree) {
            o = load_vec(sm.v + l + d);
        } else if (gtake) {
            o = load_vec(sm.gr + gb + gd);
        }
        if (qtake) {
            qo = load_vec(sm.q + l + qd);
        }
        __syncthreads();
        if (tree) {
            add_ilp(mine, o);
            store_vec(sm.v + l, mine);
        } else if (gtake) {
            add_ilp(gm, o);
            store_vec(sm.gr + gb, gm);
        }
        if (qtake) {
            add_ilp(qsum, qo);
            store_vec(sm.q + l, qsum);
        }
        if (gl && step == 3 && gb == 0) { // scan done: position 0 holds the sum of all groups; mask it out of the tree
            store_vec(sm.gr + RGRP, gm);
            gm = XYZZ<F>::inf();
            store_vec(sm.gr, gm);
        }
        qd >>= 1;
        __syncthreads();
    }
    if (l == 0) {
        R = load_vec(sm.gr + RGRP);
        XYZZ<F> wr = load_vec(sm.gr); // sum_a a * total_a
        for (int k = 0; k < 4; ++k) {
            dbl_ilp(wr);
        }
        add_ilp(mine, wr);
        Q0 = mine;
        T = qsum;
    }
}

// acc = 2^k * acc
template<class F> __device__ __forceinline__ void dbl_times(XYZZ<F> &acc, uint32_t k)
{
    for (uint32_t i = 0; i < k; ++i) {
        dbl_ilp(acc);
    }
}

template<class F>
__global__ void __launch_bounds__(RROW) k_reduce_rows(Geometry g, uint32_t lg_cpt, uint32_t rows, const uint32_t *task_off,
                                                      const XYZZ<F> *partial, XYZZ<F> *rowR, XYZZ<F> *rowQ)
{
    extern __shared__ uint4 reduce_smem[];
    RowSmem<F> sm;
    sm.v = reinterpret_cast<XYZZ<F> *>(reduce_smem);
    sm.gr = sm.v + RROW;
    sm.q = sm.gr + RGRP + 1;
    const uint32_t l = threadIdx.x;
    XYZZ<F> run, sum;
    reduce_thread_fold<F>(g, lg_cpt, blockIdx.x / rows, blockIdx.x % rows, l, task_off, partial, run, sum);
    XYZZ<F> R, Q0, T;
    if (lg_cpt) {
        cta_weighted_sum<F, true>(sm, run, sum, R, Q0, T);
    } else {
        cta_weighted_sum<F, false>(sm, run, sum, R, Q0, T);
        T = R; // one bucket per thread: sum_t = run_t
    }
    if (l == 0) {
        dbl_times(Q0, lg_cpt);
        add_ilp(T, Q0);
        store_vec(rowR + blockIdx.x, R);
        store_vec(rowQ + blockIdx.x, T);
    }
}

// One CTA per window: sum_h Q_h + (RROW cpt) sum_h h R_h over the window's rows (rows <= RROW).
template<class F>
__global__ void __launch_bounds__(RROW) k_reduce_windows(uint32_t lg_cpt, uint32_t rows, const XYZZ<F> *rowR, const XYZZ<F> *rowQ,
                                                         XYZZ<F> *windows)
{
    extern __shared__ uint4 reduce_smem[];
    RowSmem<F> sm;
    sm.v = reinterpret_cast<XYZZ<F> *>(reduce_smem);
    sm.gr = sm.v + RROW;
    sm.q = sm.gr + RGRP + 1;
    const uint32_t l = threadIdx.x, w = blockIdx.x;
    if (rows == 1) {
        if (l == 0) {
            store_vec(windows + w, load_vec(rowQ + w));
        }
        return;
    }
    XYZZ<F> r = XYZZ<F>::inf(), q = XYZZ<F>::inf();
    if (l < rows) {
        r = load_vec(rowR + (size_t)w * rows + l);
        q = load_vec(rowQ + (size_t)w * rows + l);
    }
    XYZZ<F> R, Q0, T;
    cta_weighted_sum<F, true>(sm, r, q, R, Q0, T);
    if (l == 0) {
        dbl_times(Q0, 8 + lg_cpt); // RROW = 2^8
        add_ilp(T, Q0);
        store_vec(windows + w, T);
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
__global__ void __launch_bounds__(128) k_aff_up(const uint32_t *total_entries, int round, int level, const F *E, F *Pre, F *T)
{
    const AffLevels lv = aff_levels(*total_entries, round);
    const uint32_t count = level == 1 ? lv.n1 : lv.n2, threads = level == 1 ? lv.n2 : lv.n3;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < threads) {
        aff_up_body<F>(t, count, threads, E, Pre, T);
    }
}
template<class F>
__global__ void __launch_bounds__(128) k_aff_down(const uint32_t *total_entries, int round, int level, F *E, const F *Pre, const F *Inv)
{
    const AffLevels lv = aff_levels(*total_entries, round);
    const uint32_t count = level == 1 ? lv.n1 : lv.n2, threads = level == 1 ? lv.n2 : lv.n3;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < threads) {
        aff_down_body<F>(t, count, threads, E, Pre, Inv);
    }
}
// One inversion per element of the top level, 32 threads per CTA so that the few hundred warps spread over all SMs.
template<class F> __global__ void __launch_bounds__(32) k_aff_top(const uint32_t *total_entries, int round, F *T3)
{
    const AffLevels lv = aff_levels(*total_entries, round);
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < lv.n3) {
        store_vec(T3 + t, load_vec(T3 + t).inverse_fast());
    }
}

template<class F>
int affine_tree_launch(const uint32_t *total_entries, int round, uint32_t max_entries, F *T1, F *P1, F *T2, F *P2, F *T3)
{
    cudaStream_t s = rt::stream();
    const AffLevels mx = aff_levels(max_entries, round);
    const auto grid = [](uint32_t n, uint32_t b) { return (n + b - 1) / b; };
    k_aff_up<F><<<grid(mx.n2, 128), 128, 0, s>>>(total_entries, round, 1, T1, P1, T2);
    B200_LAUNCH_CHECK();
    k_aff_up<F><<<grid(mx.n3, 128), 128, 0, s>>>(total_entries, round, 2, T2, P2, T3);
    B200_LAUNCH_CHECK();
    k_aff_top<F><<<grid(mx.n3, 32), 32, 0, s>>>(total_entries, round, T3);
    B200_LAUNCH_CHECK();
    k_aff_down<F><<<grid(mx.n3, 128), 128, 0, s>>>(total_entries, round, 2, T2, P2, T3);
    B200_LAUNCH_CHECK();
    k_aff_down<F><<<grid(mx.n2, 128), 128, 0, s>>>(total_entries, round, 1, T1, P1, T2);
    B200_LAUNCH_CHECK();
    return B200_OK;
}

template<class F>
int reduce_launch_tree(const Geometry &g, const uint32_t *task_off, const uint32_t *multi_count,
                  const uint32_t *multi_list, XYZZ<F> *partial, XYZZ<F> *block_out, XYZZ<F> *windows_dev)
{
    cudaStream_t s = rt::stream();
    k_collapse<F><<<148 * 4, 128, 0, s>>>(multi_count, multi_list, task_off, partial);
    B200_LAUNCH_CHECK();
    const uint32_t lg_cpt = pick_lg_cpt((uint64_t)g.W, g.B);
    const uint32_t per_row = (uint32_t)RROW << lg_cpt;
    const uint32_t rows = (g.B + per_row - 1) / per_row;
    const size_t smem1 = (size_t)(RROW + RGRP + 1) * sizeof(XYZZ<F>), smem2 = smem1 + (size_t)RROW * sizeof(XYZZ<F>);
    static bool attr_set = false;
    if (!attr_set) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(k_reduce_rows<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
        B200_CUDA_CHECK(cudaFuncSetAttribute(k_reduce_windows<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
        attr_set = true;
    }
    // block_out: rows * W points of R, then as many of Q (reduce_rows_capacity)
    XYZZ<F> *rowR = block_out, *rowQ = block_out + (size_t)g.W * rows;
    k_reduce_rows<F><<<(unsigned)(g.W * rows), RROW, lg_cpt ? smem2 : smem1, s>>>(g, lg_cpt, rows, task_off, partial, rowR, rowQ);
    B200_LAUNCH_CHECK();
    k_reduce_windows<F><<<g.W, RROW, smem2, s>>>(lg_cpt, rows, rowR, rowQ, windows_dev);
    B200_LAUNCH_CHECK();
    return B200_OK;
}

template<class F>
int reduce_launch_chunks(const Geometry &g, const uint32_t *task_off, const uint32_t *multi_count,
                  const uint32_t *multi_list, XYZZ<F> *partial, XYZZ<F> *block_out, XYZZ<F> *windows_dev)
{
    cudaStream_t s = rt::stream();
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


// Which formulation runs: B200_REDUCE=tree | chunks overrides (tuning knob).
template<class F>
int reduce_launch(const Geometry &g, const uint32_t *task_off, const uint32_t *multi_count,
                  const uint32_t *multi_list, XYZZ<F> *partial, XYZZ<F> *block_out, XYZZ<F> *windows_dev)
{
    static const char *mode = getenv("B200_REDUCE");
    const bool tree = mode ? strcmp(mode, "tree") == 0 : false;
    return tree ? reduce_launch_tree<F>(g, task_off, multi_count, multi_list, partial, block_out, windows_dev)
                : reduce_launch_chunks<F>(g, task_off, multi_count, multi_list, partial, block_out, windows_dev);
}

} // namespace msm
} // namespace b200
