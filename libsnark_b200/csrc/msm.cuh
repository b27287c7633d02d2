// msm.cuh — Pippenger bucket MSM, device side.
//
// Replaces libff::multi_exp<GroupT, FieldT, multi_exp_method_BDLO12_signed,
// multi_exp_base_form_special> (multiexp.tcc:643-688) and its inner loop
// multi_exp_implementation<...BDLO12_signed...>::multi_exp_inner / signed_digits_round
// (multiexp.tcc:507-633) with a different, GPU-shaped algorithm that yields the same group
// element (results are compared after normalisation):
//
//   1. prepare     scalar (Montgomery) -> standard form (libff as_bigint, fp.tcc:270-281),
//                  plus the constant K = sum_{w<W-1} 2^(c-1+cw): the base-2^c digits of
//                  t = s + K, minus 2^(c-1), are exactly libff's signed digits
//                  (field_get_signed_digit, field_utils.tcc:82-115), so every window can be
//                  extracted independently with shifts.
//   2. histogram   count entries per (window, |digit|-1) bucket            (global REDs)
//   3. scan        exclusive scan of counts -> bucket offsets; and of ceil(count/SEG)
//                  -> task offsets (a task = <= SEG consecutive entries of one bucket)
//   4. scatter     counting-sort scatter of (point index | sign) into bucket order
//   5. accumulate  one thread per task: XYZZ mixed adds of its entries; tasks are ordered
//                  by length inside each CTA so the warps run in lock-step
//   6. reduce      one thread per CHUNK buckets: running sums (the reference's
//                  multiexp_accumulate_buckets, multiexp.tcc:91-125, cut into chunks) plus
//                  a short double-and-add for the chunk's weight offset
//   7. window sum  tree sum of the chunk results of each window
//   8. fold        (host, host_fold.hpp) Horner over the windows (2^c * acc + R_w,
//                  multiexp.tcc:602-629) and normalisation to affine.
//
// Kernel bodies are written as HD functions of a flat thread index so that
// tests/host_emu can execute the same index logic on the CPU (test infrastructure only;
// the product launches the __global__ wrappers below).
#pragma once
#include "curves.cuh"
#include "msm_geometry.hpp"

namespace b200
{
namespace msm
{


template<int NL> struct Limbs {
    uint32_t v[NL];
};

// K = sum_{w < W-1} 2^(c-1 + c*w)   (the top window keeps its unsigned digit: it cannot be
// >= 2^(c-1) because W = ceil((bits+1)/c)).
template<int NL> inline Limbs<NL> make_offset_constant(const Geometry &g)
{
    Limbs<NL> k;
    for (int i = 0; i < NL; ++i) {
        k.v[i] = 0;
    }
    for (int w = 0; w < g.W - 1; ++w) {
        const int bit = g.c - 1 + g.c * w;
        if (bit < 32 * NL) {
            k.v[bit >> 5] |= 1u << (bit & 31);
        }
    }
    return k;
}

#if defined(__CUDA_ARCH__)
DEV uint32_t atomic_inc_u32(uint32_t *p) { return atomicAdd(p, 1u); }
#else
inline uint32_t atomic_inc_u32(uint32_t *p) { return (*p)++; }
#endif

// 16-byte vector copies between global memory and register structs.
template<class T> HD T load_vec(const T *p)
{
    static_assert(sizeof(T) % 16 == 0, "16-byte multiple");
    T r;
#if defined(__CUDA_ARCH__)
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&r);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(T) / 16); ++k) {
        d[k] = s[k];
    }
#else
    r = *p;
#endif
    return r;
}
template<class T> HD T load_vec_ro(const T *p)
{
    static_assert(sizeof(T) % 16 == 0, "16-byte multiple");
    T r;
#if defined(__CUDA_ARCH__)
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&r);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(T) / 16); ++k) {
        d[k] = __ldg(s + k);
    }
#else
    r = *p;
#endif
    return r;
}
// The same for a GATHER out of a table far larger than L2 (one base point per access, no neighbour wanted): the L2 is asked
// to fetch 64 bytes from DRAM instead of promoting the miss to the whole 128-byte line (ld.global.nc.L2::64B). Measured on
// the access pattern of the affine rounds (tools/microbench/gather_l2.cu): 0.95 -> 0.71 GB of DRAM reads for the same loads.
template<class T> HD T load_vec_gather(const T *p)
{
    static_assert(sizeof(T) % 16 == 0, "16-byte multiple");
    T r;
#if defined(__CUDA_ARCH__)
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&r);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(T) / 16); ++k) {
        asm volatile("ld.global.nc.L2::64B.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(d[k].x), "=r"(d[k].y), "=r"(d[k].z), "=r"(d[k].w)
                     : "l"(s + k));
    }
#else
    r = *p;
#endif
    return r;
}
template<class T> HD void store_vec(T *p, const T &v)
{
    static_assert(sizeof(T) % 16 == 0, "16-byte multiple");
#if defined(__CUDA_ARCH__)
    uint4 *d = reinterpret_cast<uint4 *>(p);
    const uint4 *s = reinterpret_cast<const uint4 *>(&v);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(T) / 16); ++k) {
        d[k] = s[k];
    }
#else
    *p = v;
#endif
}

// Signed digit of window w from t = s + K (see header). t points at the NL limbs of one scalar.
HD int32_t window_digit(const uint32_t *t, int NL, const Geometry &g, int w)
{
    const uint32_t o = (uint32_t)g.c * (uint32_t)w;
    const int limb = (int)(o >> 5);
    const int sh = (int)(o & 31);
    uint64_t v = t[limb];
    if (limb + 1 < NL) {
        v |= (uint64_t)t[limb + 1] << 32;
    }
    v >>= sh;
    if (w < g.W - 1) {
        v &= (1ull << g.c) - 1;
        return (int32_t)v - (int32_t)(1u << (g.c - 1));
    }
    return (int32_t)v; // top window: all remaining bits, 0 <= v <= 2^(c-1)
}

// ---- 1. prepare ---------------------------------------------------------------------------
template<class Fr>
HD void prepare_body(size_t i, const Fr *scalars, uint32_t *t, const Limbs<Fr::N> &K, const uint8_t *inf)
{
    constexpr int NL = Fr::N;
    Fr s = load_vec(scalars + i).from_mont();
    if (inf != nullptr && inf[i]) {
        s = Fr::zero(); // a base at infinity contributes nothing
    }
    Limbs<NL> o;
    o.v[0] = prim::add_cc(s.v[0], K.v[0]);
#pragma unroll
    for (int j = 1; j < NL - 1; ++j) {
        o.v[j] = prim::addc_cc(s.v[j], K.v[j]);
    }
    o.v[NL - 1] = prim::addc(s.v[NL - 1], K.v[NL - 1]);
    store_vec(reinterpret_cast<Limbs<NL> *>(t) + i, o);
}

// ---- 2. histogram / 4. scatter ---------------------------------------------------------------
HD void histogram_body(size_t i, const uint32_t *t, int NL, const Geometry &g, uint32_t *counts)
{
    const uint32_t *ti = t + i * NL;
    for (int w = 0; w < g.W; ++w) {
        const int32_t d = window_digit(ti, NL, g, w);
        if (d != 0) {
            const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
            atomic_inc_u32(counts + (size_t)w * g.bstride + (mag - 1));
        }
    }
}

HD void scatter_body(size_t i, const uint32_t *t, int NL, const Geometry &g, uint32_t *cursor, uint32_t *entries)
{
    const uint32_t *ti = t + i * NL;
    for (int w = 0; w < g.W; ++w) {
        const int32_t d = window_digit(ti, NL, g, w);
        if (d != 0) {
            const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
            const uint32_t pos = atomic_inc_u32(cursor + (size_t)w * g.bstride + (mag - 1));
            const uint32_t where = g.eshift ? ((uint32_t)i | ((uint32_t)w << g.eshift)) : ((uint32_t)i + (uint32_t)w * g.estride);
            entries[pos] = where | (d < 0 ? 0x80000000u : 0u);
        }
    }
}

// ---- 5a. tasks -------------------------------------------------------------------------------
// task = (start offset into entries, length <= SEG); tasks of bucket b occupy
// [task_off[b], task_off[b+1]).
struct Task {
    uint32_t start;
    uint32_t len;
};

HD void build_tasks_body(
    size_t b, const uint32_t *counts, const uint32_t *offsets, const uint32_t *task_off, Task *tasks,
    uint32_t *multi_count, uint32_t *multi_list, uint32_t collapse, uint32_t rounds = 0, uint32_t seg = SEG)
{
    // with affine rounds (msm_affine.cuh) the tasks are cut over the bucket's survivors
    const uint32_t cnt = padded_count(counts[b], rounds) >> rounds;
    const uint32_t first = task_off[b];
    const uint32_t off = offsets[b] >> rounds;
    for (uint32_t s = 0, done = 0; done < cnt; ++s, done += seg) {
        Task tk;
        tk.start = off + done;
        tk.len = (cnt - done) < seg ? (cnt - done) : seg;
        tasks[first + s] = tk;
    }
    if (cnt > collapse * seg) { // many tasks: their partial sums are merged by k_collapse
#if defined(__CUDA_ARCH__)
        multi_list[atomicAdd(multi_count, 1u)] = (uint32_t)b;
#else
        multi_list[(*multi_count)++] = (uint32_t)b;
#endif
    }
}

// ---- 5b. accumulate ---------------------------------------------------------------------------
template<class F>
HD void accumulate_task(const Task &tk, const uint32_t *entries, const Affine<F> *bases, XYZZ<F> *out)
{
    XYZZ<F> acc = XYZZ<F>::inf();
#if defined(B200_ACC_NO_PREFETCH)
    // wide fields: the next base is NOT held in registers across the addition (it would cost
    // 2 sizeof(F) / 4 registers of a budget that is already spilling); the other resident warps
    // cover the load latency instead
    for (uint32_t k = 0; k < tk.len; ++k) {
        const uint32_t e = entries[tk.start + k];
        Affine<F> cur = load_vec_gather(bases + (e & 0x7fffffffu));
        cur.y = cur.y.cneg((e >> 31) != 0);
        madd(acc, cur);
    }
#else
    uint32_t e = entries[tk.start];
    Affine<F> p = load_vec_gather(bases + (e & 0x7fffffffu));
    for (uint32_t k = 0; k < tk.len; ++k) {
        const bool negate = (e >> 31) != 0;
        Affine<F> cur = p;
        if (k + 1 < tk.len) { // prefetch the next base while this one is added
            e = entries[tk.start + k + 1];
            p = load_vec_gather(bases + (e & 0x7fffffffu));
        }
        cur.y = cur.y.cneg(negate);
        madd(acc, cur);
    }
#endif
    store_vec(out, acc);
}

// ---- 6. bucket reduce -------------------------------------------------------------------------
// Thread `gidx` owns buckets [gidx*CHUNK, (gidx+1)*CHUNK) (global ids; B is a multiple of
// CHUNK so a chunk never straddles windows). Output = sum_k weight(k) * S_k, where the
// weight of bucket k of the window is (local index + 1).
template<class F>
HD void reduce_chunk_body(
    size_t gidx, const Geometry &g, const uint32_t *task_off, const XYZZ<F> *partial, XYZZ<F> *chunk_out)
{
    // chunk_out: where this thread's weighted sum goes (a register/local struct in the kernel)
    XYZZ<F> run = XYZZ<F>::inf();
    XYZZ<F> sum = XYZZ<F>::inf();
    const size_t b0 = gidx * g.chunk;
    for (int k = (int)g.chunk - 1; k >= 0; --k) {
        const uint32_t t0 = task_off[b0 + k];
        const uint32_t t1 = task_off[b0 + k + 1];
        // up to g.collapse partial sums are added here; beyond that k_collapse has already
        // left the bucket's total in its first task slot
        const uint32_t tend = (t1 - t0 > g.collapse) ? t0 + 1 : t1;
        for (uint32_t t = t0; t < tend; ++t) {
            const XYZZ<F> s = load_vec(partial + t);
            add_cold(run, s);
        }
        add_cold(sum, run);
    }
    // sum = sum_k (k+1) S_k ; the chunk starts at local bucket m = (gidx % cpw) * CHUNK, so add m * run.
    const uint32_t m = (uint32_t)(gidx % g.cpw) * g.chunk;
    if (m != 0 && !run.is_inf()) {
        XYZZ<F> acc = run;
        int top = 31;
        while (((m >> top) & 1u) == 0) {
            --top;
        }
        for (int bit = top - 1; bit >= 0; --bit) {
            dbl_cold(acc);
            if ((m >> bit) & 1u) {
                add_cold(acc, run);
            }
        }
        add_cold(sum, acc);
    }
    *chunk_out = sum;
}

// Reference layout of a normalised point: (x, y, one) or zero() = (0, one, 0)
// (alt_bn128_g1.cpp:68-91; bls12_377 and bw6_761 use the same encoding).
template<class F> struct RefPoint {
    F x, y, z;
};
template<class F> HD RefPoint<F> normalise(const XYZZ<F> &p)
{
    RefPoint<F> o;
    Affine<F> a;
    if (to_affine(p, a)) {
        o.x = a.x;
        o.y = a.y;
        o.z = F::one();
    } else {
        o.x = F::zero();
        o.y = F::one();
        o.z = F::zero();
    }
    return o;
}

// Base conversion from the reference's (X, Y, Z) special form to device affine + infinity flag.
template<class F>
HD void convert_base_body(size_t i, const uint8_t *src, size_t stride, Affine<F> *dst, uint8_t *inf, uint32_t *any_inf)
{
    const RefPoint<F> *p = reinterpret_cast<const RefPoint<F> *>(src + i * stride);
    Affine<F> a;
    a.x = load_vec(&p->x);
    a.y = load_vec(&p->y);
    const F z = load_vec(&p->z);
    const bool is_inf = z.is_zero();
    if (is_inf) {
        a.x = F::zero();
        a.y = F::zero();
#if defined(__CUDA_ARCH__)
        atomicOr(any_inf, 1u);
#else
        *any_inf |= 1u;
#endif
    }
    inf[i] = is_inf ? 1 : 0;
    store_vec(dst + i, a);
}

// Precomputed tables: row w of the table holds [2^(c w)] P_i, affine, for every base i (row 0 = the bases).
// One thread per base walks up the rows: c doublings, one normalisation (a field inversion — this is a
// once-per-key cost). A multiple that is the point at infinity (possible only for a base outside the
// prime-order subgroup) cannot be stored as an affine point: it raises *fail and the caller drops the table.
template<class F>
HD void build_table_body(size_t i, size_t row, int c, int W, Affine<F> *table, const uint8_t *inf, uint32_t *fail)
{
    if (inf != nullptr && inf[i]) {
        Affine<F> z;
        z.x = F::zero();
        z.y = F::zero();
        for (int w = 1; w < W; ++w) {
            store_vec(table + (size_t)w * row + i, z);
        }
        return;
    }
    Affine<F> a = load_vec(table + i);
    for (int w = 1; w < W; ++w) {
        XYZZ<F> q = dbl_affine_cold(a);
        for (int k = 1; k < c; ++k) {
            dbl_cold(q);
        }
        if (!to_affine(q, a)) {
#if defined(__CUDA_ARCH__)
            atomicOr(fail, 1u);
#else
            *fail |= 1u;
#endif
            return;
        }
        store_vec(table + (size_t)w * row + i, a);
    }
}

#if defined(__CUDACC__)
// =================================== __global__ wrappers =====================================

template<class Fr>
__global__ void k_prepare(const Fr *scalars, uint32_t *t, size_t n, Limbs<Fr::N> K, const uint8_t *inf)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        prepare_body<Fr>(i, scalars, t, K, inf);
    }
}

static __global__ void k_histogram(const uint32_t *t, int NL, Geometry g, uint32_t *counts)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < g.n) {
        histogram_body(i, t, NL, g, counts);
    }
}

static __global__ void k_scatter(const uint32_t *t, int NL, Geometry g, uint32_t *cursor, uint32_t *entries)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < g.n) {
        scatter_body(i, t, NL, g, cursor, entries);
    }
}

// ---- 3. scan: counts -> offsets (exclusive) and ceil(count/SEG) -> task_off (exclusive) ------
// Outputs have nb + 1 entries; [nb] holds the totals. cursor = copy of offsets for the scatter.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint2 block_exclusive_scan2(uint2 v, uint2 *total)
{
    // exclusive scan of one uint2 per thread over a 256-thread block
    __shared__ uint2 warp_tot[SCAN_THREADS / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint2 inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t ax = __shfl_up_sync(0xffffffffu, inc.x, d);
        const uint32_t ay = __shfl_up_sync(0xffffffffu, inc.y, d);
        if (lane >= d) {
            inc.x += ax;
            inc.y += ay;
        }
    }
    if (lane == 31) {
        warp_tot[wid] = inc;
    }
    __syncthreads();
    uint2 base = make_uint2(0, 0), tot = make_uint2(0, 0);
#pragma unroll
    for (int k = 0; k < SCAN_THREADS / 32; ++k) {
        const uint2 wt = warp_tot[k];
        if (k < wid) {
            base.x += wt.x;
            base.y += wt.y;
        }
        tot.x += wt.x;
        tot.y += wt.y;
    }
    __syncthreads();
    *total = tot;
    return make_uint2(base.x + inc.x - v.x, base.y + inc.y - v.y);
}

// Both scans see a bucket's count padded to a multiple of 2^rounds (affine rounds, msm_affine.cuh; rounds = 0: as is).
__device__ __forceinline__ uint2 scan_item(uint32_t c, uint32_t rounds)
{
    const uint32_t p = padded_count(c, rounds);
    const uint32_t seg = rounds ? SEG_PTS : SEG; // Geometry::seg
    return make_uint2(p, ((p >> rounds) + seg - 1) / seg);
}

static __global__ void k_scan_tiles(const uint32_t *counts, uint32_t nb, uint2 *tile_sums, uint32_t rounds)
{
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint2 s = make_uint2(0, 0);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const uint32_t i = base + k;
        const uint2 it = scan_item(i < nb ? counts[i] : 0u, rounds);
        s.x += it.x;
        s.y += it.y;
    }
    uint2 tot;
    block_exclusive_scan2(s, &tot);
    if (threadIdx.x == 0) {
        tile_sums[blockIdx.x] = tot;
    }
}

static __global__ void k_scan_tile_sums(uint2 *tile_sums, uint32_t ntiles)
{
    // single block; exclusive scan in place, running carry across iterations
    uint2 carry = make_uint2(0, 0);
    for (uint32_t base = 0; base < ntiles; base += SCAN_THREADS) {
        const uint32_t i = base + threadIdx.x;
        const uint2 v = i < ntiles ? tile_sums[i] : make_uint2(0, 0);
        uint2 tot;
        const uint2 ex = block_exclusive_scan2(v, &tot);
        if (i < ntiles) {
            tile_sums[i] = make_uint2(carry.x + ex.x, carry.y + ex.y);
        }
        carry.x += tot.x;
        carry.y += tot.y;
    }
}

static __global__ void k_scan_finish(
    const uint32_t *counts, uint32_t nb, const uint2 *tile_sums, uint32_t *offsets, uint32_t *cursor,
    uint32_t *task_off, uint32_t rounds)
{
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t c[SCAN_ITEMS];
    uint2 s = make_uint2(0, 0);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const uint32_t i = base + k;
        c[k] = i < nb ? counts[i] : 0u;
        const uint2 it = scan_item(c[k], rounds);
        s.x += it.x;
        s.y += it.y;
    }
    uint2 tot;
    uint2 ex = block_exclusive_scan2(s, &tot);
    const uint2 tb = tile_sums[blockIdx.x];
    ex.x += tb.x;
    ex.y += tb.y;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const uint32_t i = base + k;
        if (i < nb) {
            offsets[i] = ex.x;
            cursor[i] = ex.x;
            task_off[i] = ex.y;
        }
        const uint2 it = scan_item(c[k], rounds);
        ex.x += it.x;
        ex.y += it.y;
        if (i + 1 == nb) { // totals
            offsets[nb] = ex.x;
            task_off[nb] = ex.y;
        }
    }
}

static __global__ void k_build_tasks(
    const uint32_t *counts, const uint32_t *offsets, const uint32_t *task_off, uint32_t nb, Task *tasks,
    uint32_t *multi_count, uint32_t *multi_list, uint32_t collapse, uint32_t rounds)
{
    const size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b < nb) {
        build_tasks_body(b, counts, offsets, task_off, tasks, multi_count, multi_list, collapse, rounds,
                         rounds ? (uint32_t)SEG_PTS : (uint32_t)SEG);
    }
}

// One thread per task. The ACC_THREADS tasks of a CTA are consecutive (so their entries are
// contiguous) and are redistributed over the threads in order of decreasing length, which
// keeps the trip counts inside each warp nearly equal.
constexpr int ACC_THREADS = 128;
// Resident CTAs per SM the accumulate kernel is compiled for (caps its registers): the wide fields and
// Fp2 otherwise take 222-255 registers and run two warps per scheduler (profiles/r1_ncu_accumulate_g2.md).
#ifndef B200_ACC_MIN_BLOCKS
#define B200_ACC_MIN_BLOCKS 1
#endif

template<class F>
__global__ void __launch_bounds__(ACC_THREADS, B200_ACC_MIN_BLOCKS) k_accumulate(
    const Task *tasks, const uint32_t *num_tasks, const uint32_t *entries, const Affine<F> *bases,
    XYZZ<F> *partial)
{
    __shared__ uint32_t hist[SEG + 1];
    __shared__ uint32_t order[ACC_THREADS];
    const uint32_t T = *num_tasks;
    const uint32_t base = blockIdx.x * ACC_THREADS;
    if (base >= T) {
        return;
    }
    const uint32_t tid = threadIdx.x;
    const uint32_t mine = base + tid;
    const uint32_t len = mine < T ? tasks[mine].len : 0u;
    if (tid <= SEG) {
        hist[tid] = 0;
    }
    __syncthreads();
    uint32_t rank = 0;
    if (len) {
        rank = atomicAdd(&hist[len], 1u);
    }
    __syncthreads();
    if (tid == 0) { // exclusive scan from the longest bin down
        uint32_t acc = 0;
        for (int l = SEG; l >= 1; --l) {
            const uint32_t c = hist[l];
            hist[l] = acc;
            acc += c;
        }
        hist[0] = acc; // number of live tasks in this CTA
    }
    __syncthreads();
    if (len) {
        order[hist[len] + rank] = tid;
    }
    __syncthreads();
    if (tid < hist[0]) {
        const uint32_t id = base + order[tid];
        const Task tk = tasks[id];
        accumulate_task<F>(tk, entries, bases, partial + id);
    }
}

template<class F>
__global__ void k_convert_bases(const uint8_t *src, size_t stride, size_t n, Affine<F> *dst, uint8_t *inf, uint32_t *any_inf)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        convert_base_body<F>(i, src, stride, dst, inf, any_inf);
    }
}

#endif // __CUDACC__

} // namespace msm
} // namespace b200
