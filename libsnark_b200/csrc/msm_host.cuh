// msm_host.cuh — host-side driver of the MSM pipeline in msm.cuh (stream-ordered launches,
// no host synchronisation between the stages; the only device->host transfer is the result).
#pragma once
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "host_fold.hpp"
#include "msm.cuh"
#include "msm_affine.cuh"
#include "runtime.hpp"

namespace b200
{
namespace msm
{

struct Workspace {
    rt::DevBuf t, counts, offsets, cursor, task_off, tile_sums, entries, tasks, partial, chunk_out, staging, any_inf, multi;
    rt::DevBuf aff_p0, aff_t1, aff_p1, aff_t2, aff_p2, aff_t3, aff_a, aff_b; // affine rounds (msm_affine.cuh)
};
inline Workspace &workspace() // of the calling thread's execution slot (runtime.hpp)
{
    static Workspace ws[rt::NUM_SLOTS];
    return ws[rt::slot()];
}

// Stages 6-7; defined in msm_cold.cuh and instantiated in msm_cold_<curve>_<group>.cu.
template<class F>
int reduce_launch(const Geometry &g, const uint32_t *task_off, const uint32_t *multi_count,
                  const uint32_t *multi_list, XYZZ<F> *partial, XYZZ<F> *block_out, XYZZ<F> *windows_dev, cudaStream_t s);

// Upper levels of an affine round's inversion tree; defined in msm_cold.cuh like reduce_launch.
template<class F>
int affine_tree_launch(const uint32_t *total_entries, uint32_t rounds, uint32_t max_entries, F *T1, F *P1, F *T2, F *P2, F *T3,
                       cudaStream_t s);

// Table rows 1..W-1 of a base set (build_table_body); defined in msm_cold.cuh like reduce_launch.
template<class F> int table_launch(size_t n, int c, int W, Affine<F> *table, const uint8_t *inf, uint32_t *fail);

inline unsigned grid_for(size_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }

template<class F>
int bases_upload(int curve, int group, const void *host, size_t stride, size_t n, rt::BaseSet *out)
{
    if (stride < sizeof(RefPoint<F>) || stride % 16 != 0) {
        return rt::set_error(B200_ERR_ARG, "base stride %zu invalid for a %zu-byte point", stride, sizeof(RefPoint<F>));
    }
    Workspace &ws = workspace();
    cudaStream_t s = rt::stream();
    const size_t bytes = n ? (n - 1) * stride + sizeof(RefPoint<F>) : 0;
    void *pts = nullptr;
    uint8_t *inf = nullptr;
    if (n) {
        if (ws.staging.reserve(bytes) || ws.any_inf.reserve(16)) {
            return B200_ERR_CUDA;
        }
        B200_CUDA_CHECK(cudaMalloc(&pts, n * sizeof(Affine<F>)));
        uint32_t any = 0;
        // every failure after the first allocation releases what was allocated (such failures are usually
        // out-of-memory, where a leak of the partial set makes things worse)
        const auto convert = [&]() -> int {
            B200_CUDA_CHECK(cudaMalloc((void **)&inf, n));
            if (const int rc = rt::staged_h2d(ws.staging.p, host, bytes, s)) { // the reference's vectors are pageable
                return rc;
            }
            B200_CUDA_CHECK(cudaMemsetAsync(ws.any_inf.p, 0, 4, s));
            k_convert_bases<F><<<grid_for(n, 256), 256, 0, s>>>(
                ws.staging.as<uint8_t>(), stride, n, static_cast<Affine<F> *>(pts), inf, ws.any_inf.as<uint32_t>());
            B200_LAUNCH_CHECK();
            B200_CUDA_CHECK(cudaMemcpyAsync(&any, ws.any_inf.p, 4, cudaMemcpyDeviceToHost, s));
            B200_CUDA_CHECK(cudaStreamSynchronize(s));
            return B200_OK;
        };
        if (const int rc = convert()) {
            cudaFree(pts);
            if (inf) {
                cudaFree(inf);
            }
            return rc;
        }
        if (!any) {
            cudaFree(inf);
            inf = nullptr;
        }
    }
    out->curve = curve;
    out->group = group;
    out->n = n;
    out->point_bytes = sizeof(Affine<F>);
    out->pts = pts;
    out->inf = inf;
    return B200_OK;
}

// b200_bases_precompute: grows the set's point array to W rows, row w = [2^(c w)] P, for the window width
// that is best once the reduction is paid for a single bucket set (table_window_bits). On failure (out of
// memory, a multiple at infinity) the set is left as it was, without tables.
template<class Fr, class F> int bases_precompute(rt::BaseSet *bs)
{
    if (bs->table_c || bs->n == 0) {
        return B200_OK;
    }
    const int c = table_window_bits(bs->n, Fr::params::BITS);
    const int W = (Fr::params::BITS + 1 + c - 1) / c;
    if ((uint64_t)W * bs->n >= 0x7fffffffull) {
        return rt::set_error(B200_ERR_ARG, "a table of %d rows of %zu points exceeds the 32-bit entry index", W, bs->n);
    }
    Workspace &ws = workspace();
    cudaStream_t s = rt::stream();
    if (ws.any_inf.reserve(16)) {
        return B200_ERR_CUDA;
    }
    Affine<F> *table = nullptr;
    B200_CUDA_CHECK(cudaMalloc((void **)&table, (size_t)W * bs->n * sizeof(Affine<F>)));
    cudaError_t e = cudaMemcpyAsync(table, bs->pts, bs->n * sizeof(Affine<F>), cudaMemcpyDeviceToDevice, s);
    if (e == cudaSuccess) {
        e = cudaMemsetAsync(ws.any_inf.p, 0, 4, s);
    }
    int rc = e == cudaSuccess ? table_launch<F>(bs->n, c, W, table, bs->inf, ws.any_inf.as<uint32_t>()) : B200_ERR_CUDA;
    uint32_t fail = 0;
    if (rc == B200_OK) {
        e = cudaMemcpyAsync(&fail, ws.any_inf.p, 4, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) {
            e = cudaStreamSynchronize(s);
        }
    }
    if (rc != B200_OK || e != cudaSuccess || fail) {
        cudaFree(table);
        if (rc != B200_OK) {
            return rc;
        }
        if (e != cudaSuccess) {
            return rt::set_error(B200_ERR_CUDA, "table build failed: %s", cudaGetErrorString(e));
        }
        return rt::set_error(B200_ERR_ARG, "a multiple 2^(c w) P of a base is the point at infinity: no table built");
    }
    cudaFree(bs->pts);
    bs->pts = table;
    bs->table_c = c;
    bs->table_W = W;
    return B200_OK;
}

// Geometry of the MSM over n points of `bs`: the table geometry when the set has tables, plus the affine rounds.
template<class Fr> Geometry run_geometry(const rt::BaseSet &bs, size_t n)
{
    static const int table_chunk = getenv("B200_TABLE_CHUNK") ? atoi(getenv("B200_TABLE_CHUNK")) : 0; // tuning knob
    Geometry g = bs.table_c ? make_table_geometry(n, Fr::params::BITS, bs.table_c, bs.n, table_chunk)
                            : make_geometry(n, Fr::params::BITS);
    // affine rounds ahead of the XYZZ accumulation; B200_AFFINE_ROUNDS = 0..4 overrides the choice (tuning knob)
    static const int forced = getenv("B200_AFFINE_ROUNDS") ? atoi(getenv("B200_AFFINE_ROUNDS")) : -1;
    const uint64_t ents = (uint64_t)g.W * n;
    const uint32_t r = forced >= 0 ? (uint32_t)(forced > 4 ? 4 : forced) : pick_rounds(ents, g.NB);
    set_rounds(g, r, ents);
    return g;
}
inline size_t max_entries_of(const Geometry &g, size_t n) { return (size_t)g.W * n + (size_t)g.NB * ((1u << g.rounds) - 1u); }
inline size_t max_tasks_of(const Geometry &g, size_t max_entries) { return (size_t)g.NB + (max_entries >> g.rounds) / g.seg + 1; }

// Stages 1-4 into the given buffers (a slot's workspace or a SharedSort).
struct SortBuffers {
    rt::DevBuf &t, &counts, &offsets, &cursor, &task_off, &tile_sums, &entries, &tasks, &multi;
};
template<class Fr>
int sort_launch(const Geometry &g, size_t n, const void *scalars_dev, const uint8_t *inf, SortBuffers b, cudaStream_t s)
{
    constexpr int NL = Fr::N;
    const uint32_t R = g.rounds;
    const size_t max_entries = max_entries_of(g, n), max_tasks = max_tasks_of(g, max_entries);
    const uint32_t ntiles = grid_for(g.NB, SCAN_TILE);
    if (max_entries >= 0x7fffffffull || max_tasks >= 0xffffffffull) {
        return rt::set_error(B200_ERR_ARG, "MSM of %zu points exceeds the 32-bit entry index", n);
    }
    if (b.t.reserve(n * NL * 4 + 16) || b.counts.reserve(((size_t)g.NB + 1) * 4) || b.offsets.reserve(((size_t)g.NB + 1) * 4) ||
        b.cursor.reserve(((size_t)g.NB + 1) * 4) || b.task_off.reserve(((size_t)g.NB + 1) * 4) ||
        b.tile_sums.reserve((size_t)ntiles * 8 + 8) || b.entries.reserve(max_entries * 4 + 16) ||
        b.tasks.reserve(max_tasks * sizeof(Task)) || b.multi.reserve(((size_t)g.NB + 4) * 4)) {
        return B200_ERR_CUDA;
    }
    uint32_t *t = b.t.as<uint32_t>(), *counts = b.counts.as<uint32_t>(), *offsets = b.offsets.as<uint32_t>();
    uint32_t *cursor = b.cursor.as<uint32_t>(), *task_off = b.task_off.as<uint32_t>(), *entries = b.entries.as<uint32_t>();
    uint2 *tile_sums = b.tile_sums.as<uint2>();
    Task *tasks = b.tasks.as<Task>();
    uint32_t *multi_count = b.multi.as<uint32_t>(), *multi_list = multi_count + 4;
    B200_CUDA_CHECK(cudaMemsetAsync(counts, 0, ((size_t)g.NB + 1) * 4, s));
    B200_CUDA_CHECK(cudaMemsetAsync(multi_count, 0, 16, s));
    const Limbs<NL> K = make_offset_constant<NL>(g);
    k_prepare<Fr><<<grid_for(n, 256), 256, 0, s>>>(static_cast<const Fr *>(scalars_dev), t, n, K, inf);
    B200_LAUNCH_CHECK();
    k_histogram<<<grid_for(n, 256), 256, 0, s>>>(t, NL, g, counts);
    B200_LAUNCH_CHECK();
    k_scan_tiles<<<ntiles, SCAN_THREADS, 0, s>>>(counts, g.NB, tile_sums, R);
    B200_LAUNCH_CHECK();
    k_scan_tile_sums<<<1, SCAN_THREADS, 0, s>>>(tile_sums, ntiles);
    B200_LAUNCH_CHECK();
    k_scan_finish<<<ntiles, SCAN_THREADS, 0, s>>>(counts, g.NB, tile_sums, offsets, cursor, task_off, R);
    B200_LAUNCH_CHECK();
    if (R) { // the padding of every bucket reads as NULL_ENTRY
        B200_CUDA_CHECK(cudaMemsetAsync(entries, 0xff, max_entries * 4, s));
    }
    k_scatter<<<grid_for(n, 256), 256, 0, s>>>(t, NL, g, cursor, entries);
    B200_LAUNCH_CHECK();
    k_build_tasks<<<grid_for(g.NB, 256), 256, 0, s>>>(counts, offsets, task_off, g.NB, tasks, multi_count, multi_list, g.collapse, R);
    B200_LAUNCH_CHECK();
    return B200_OK;
}

// The sort of n scalars over the geometry of `bs` (a set with tables), for several MSMs (rt::SharedSort).
template<class Fr> int msm_sort(const rt::BaseSet &bs, size_t n, const void *scalars_dev, rt::SharedSort *out, cudaStream_t s)
{
    Geometry g = run_geometry<Fr>(bs, n);
    int k = 1;
    while ((1ull << k) < n) {
        ++k;
    }
    if (!bs.table_c || g.rounds == 0 || bs.inf || (uint64_t)g.W << k >= 0x7fffffffull) {
        return rt::set_error(B200_ERR_ARG, "this MSM cannot run over a shared sort");
    }
    g.eshift = (uint32_t)k;
    if (!out->done) {
        B200_CUDA_CHECK(cudaEventCreateWithFlags(&out->done, cudaEventDisableTiming));
    }
    SortBuffers b{out->t, out->counts, out->offsets, out->cursor, out->task_off, out->tile_sums, out->entries, out->tasks, out->multi};
    if (int rc = sort_launch<Fr>(g, n, scalars_dev, nullptr, b, s)) {
        return rc;
    }
    B200_CUDA_CHECK(cudaEventRecord(out->done, s));
    out->g = g;
    out->max_entries = max_entries_of(g, n);
    out->max_tasks = max_tasks_of(g, out->max_entries);
    out->curve = bs.curve;
    return B200_OK;
}

// Runs stages 1-7 (5-7 over a shared sort); leaves W window sums (XYZZ<F>) at windows_dev.
template<class Fr, class F>
int msm_windows(const rt::BaseSet &bs, size_t first, size_t n, const void *scalars_dev, void *windows_dev, int *W_out, int *c_out,
                const rt::SharedSort *shared = nullptr, uint32_t skip = 0)
{
    // What the caller sees is always the plain geometry of n (b200_msm_num_windows / b200_msm_fold_windows).
    // Over a table the whole sum comes out as "window 0" and the other windows are the point at infinity,
    // which the Horner fold passes through unchanged.
    const Geometry pub = make_geometry(n, Fr::params::BITS);
    if (shared && (!bs.table_c || bs.table_c != shared->g.c || bs.inf || first != 0 || n + skip != shared->g.n)) {
        return rt::set_error(B200_ERR_ARG, "base set does not match the shared sort");
    }
    const Geometry g = shared ? shared->g : run_geometry<Fr>(bs, n);
    const int reduce_windows = bs.table_c ? 1 : g.W;
    const uint32_t R = g.rounds;
    *W_out = pub.W;
    *c_out = pub.c;
    Workspace &ws = workspace();
    cudaStream_t s = rt::stream();
    rt::MsmTiming &tm = rt::msm_timing();

    const size_t max_entries = shared ? shared->max_entries : max_entries_of(g, n);
    const size_t max_tasks = shared ? shared->max_tasks : max_tasks_of(g, max_entries);
    if (ws.partial.reserve(max_tasks * sizeof(XYZZ<F>)) ||
        ws.chunk_out.reserve(((size_t)reduce_windows * g.opw) * sizeof(XYZZ<F>))) {
        return B200_ERR_CUDA;
    }
    const AffLevels amax = aff_levels((uint32_t)max_entries, R);
    if (R && (ws.aff_p0.reserve((max_entries / 2 + 1) * sizeof(F)) || ws.aff_t1.reserve((size_t)amax.n1 * sizeof(F)) ||
              ws.aff_p1.reserve((size_t)amax.n1 * sizeof(F)) || ws.aff_t2.reserve((size_t)amax.n2 * sizeof(F)) ||
              ws.aff_p2.reserve((size_t)amax.n2 * sizeof(F)) || ws.aff_t3.reserve((size_t)amax.n3 * sizeof(F)) ||
              ws.aff_a.reserve((max_entries / 2 + 1) * sizeof(Affine<F>)) ||
              (R > 1 && ws.aff_b.reserve((max_entries / 4 + 1) * sizeof(Affine<F>))))) {
        return B200_ERR_CUDA;
    }
    const Affine<F> *bases = static_cast<const Affine<F> *>(bs.pts) + first;
    const uint8_t *inf = bs.inf ? bs.inf + first : nullptr;

    B200_CUDA_CHECK(cudaEventRecord(tm.ev[0], s));
    // 1-4: digits and counting sort
    if (shared) {
        B200_CUDA_CHECK(cudaStreamWaitEvent(s, shared->done, 0));
    } else {
        SortBuffers b{ws.t, ws.counts, ws.offsets, ws.cursor, ws.task_off, ws.tile_sums, ws.entries, ws.tasks, ws.multi};
        if (int rc = sort_launch<Fr>(g, n, scalars_dev, inf, b, s)) {
            return rc;
        }
    }
    // (const_cast: a shared sort is only read)
    rt::SharedSort *sh = const_cast<rt::SharedSort *>(shared);
    uint32_t *offsets = shared ? sh->offsets.as<uint32_t>() : ws.offsets.as<uint32_t>();
    uint32_t *task_off = shared ? sh->task_off.as<uint32_t>() : ws.task_off.as<uint32_t>();
    uint32_t *entries = shared ? sh->entries.as<uint32_t>() : ws.entries.as<uint32_t>();
    Task *tasks = shared ? sh->tasks.as<Task>() : ws.tasks.as<Task>();
    uint32_t *multi_count = shared ? sh->multi.as<uint32_t>() : ws.multi.as<uint32_t>();
    uint32_t *multi_list = multi_count + 4;
    XYZZ<F> *partial = ws.partial.as<XYZZ<F>>();
    XYZZ<F> *chunk_out = ws.chunk_out.as<XYZZ<F>>();
    B200_CUDA_CHECK(cudaEventRecord(tm.ev[1], s));
    // 5: bucket accumulation (grid sized for the task upper bound; surplus CTAs exit at once)
    if (R) {
        // affine rounds (msm_affine.cuh): round r halves every bucket
        const uint32_t *total = offsets + g.NB;
        Affine<F> *buf[2] = {ws.aff_a.as<Affine<F>>(), ws.aff_b.as<Affine<F>>()};
        F *Pex = ws.aff_p0.as<F>(), *T1 = ws.aff_t1.as<F>();
        const Affine<F> *prev = nullptr;
        SrcEntries<F> esrc{entries, bases};
        if (shared) {
            esrc.eshift = g.eshift;
            esrc.row = (uint32_t)bs.n;
            esrc.skip = skip;
        }
        const unsigned grid = grid_for(amax.n1, AFF_THREADS);
        for (uint32_t r = 1; r <= R; ++r) {
            Affine<F> *out = buf[(r - 1) & 1];
            const SrcPoints<F> psrc{prev};
            if (r == 1) {
                k_aff_up_round<F, SrcEntries<F>><<<grid, AFF_THREADS, 0, s>>>(total, R, r, esrc, Pex, T1);
            } else {
                k_aff_up_round<F, SrcPoints<F>><<<grid, AFF_THREADS, 0, s>>>(total, R, r, psrc, Pex, T1);
            }
            B200_LAUNCH_CHECK();
            if (int rc = affine_tree_launch<F>(total, R, (uint32_t)max_entries, T1, ws.aff_p1.as<F>(), ws.aff_t2.as<F>(),
                                               ws.aff_p2.as<F>(), ws.aff_t3.as<F>(), s)) {
                return rc;
            }
            if (r == 1) {
                k_aff_round<F, SrcEntries<F>><<<grid, AFF_THREADS, 0, s>>>(total, R, r, esrc, Pex, T1, out);
            } else {
                k_aff_round<F, SrcPoints<F>><<<grid, AFF_THREADS, 0, s>>>(total, R, r, psrc, Pex, T1, out);
            }
            B200_LAUNCH_CHECK();
            prev = out;
        }
        k_accumulate_pts<F><<<grid_for(max_tasks, ACC_THREADS), ACC_THREADS, 0, s>>>(tasks, task_off + g.NB, prev, partial);
    } else {
        k_accumulate<F><<<grid_for(max_tasks, ACC_THREADS), ACC_THREADS, 0, s>>>(
            tasks, task_off + g.NB, entries, bases, partial);
    }
    B200_LAUNCH_CHECK();
    B200_CUDA_CHECK(cudaEventRecord(tm.ev[2], s));
    // 6-7: bucket reduction
    {
        Geometry gr = g;
        gr.W = reduce_windows;
        const int rc = reduce_launch<F>(gr, task_off, multi_count, multi_list, partial, chunk_out,
                                         static_cast<XYZZ<F> *>(windows_dev), s);
        if (rc) {
            return rc;
        }
        if (pub.W > reduce_windows) { // XYZZ infinity is all zeros
            B200_CUDA_CHECK(cudaMemsetAsync(static_cast<XYZZ<F> *>(windows_dev) + reduce_windows, 0,
                                            (size_t)(pub.W - reduce_windows) * sizeof(XYZZ<F>), s));
        }
    }
    B200_CUDA_CHECK(cudaEventRecord(tm.ev[3], s));
    return B200_OK;
}

// Stage 8 on the host (host_fold.hpp). `windows` holds shards * W points; when on_device it is
// first copied back (2 KB for alt_bn128 G1). out: one normalised point, host memory.
template<class F> int msm_fold(const void *windows, int on_device, int W, int c, size_t shards, void *out)
{
    const size_t count = (size_t)W * shards;
    std::vector<XYZZ<F>> host(count);
    rt::MsmTiming &tm = rt::msm_timing();
    if (on_device) {
        cudaStream_t s = rt::stream();
        B200_CUDA_CHECK(cudaMemcpyAsync(host.data(), windows, count * sizeof(XYZZ<F>), cudaMemcpyDeviceToHost, s));
        B200_CUDA_CHECK(cudaEventRecord(tm.ev[4], s));
        B200_CUDA_CHECK(cudaStreamSynchronize(s));
        tm.valid = true;
    } else {
        memcpy(host.data(), windows, count * sizeof(XYZZ<F>));
    }
    const auto t0 = std::chrono::steady_clock::now();
    const RefPoint<F> r = host_fold<F>(host.data(), W, c, shards);
    tm.host_fold_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    memcpy(out, &r, sizeof(r));
    return B200_OK;
}

} // namespace msm

#define B200_DEFINE_GROUP_ENTRY(name, CURVE, GROUP)                                                            \
    int name##_bases_upload(int curve, int group, const void *host, size_t stride, size_t n,                  \
                            rt::BaseSet *out)                                                                  \
    {                                                                                                          \
        return msm::bases_upload<group_traits<CURVE, GROUP>::F>(curve, group, host, stride, n, out);           \
    }                                                                                                          \
    int name##_bases_precompute(rt::BaseSet *bs)                                                               \
    {                                                                                                          \
        return msm::bases_precompute<group_traits<CURVE, GROUP>::Fr, group_traits<CURVE, GROUP>::F>(bs);       \
    }                                                                                                          \
    int name##_msm_windows(const rt::BaseSet &bs, size_t first, size_t n, const void *scalars_dev,            \
                           void *windows_dev, int *W, int *c, const rt::SharedSort *shared, uint32_t skip)     \
    {                                                                                                          \
        return msm::msm_windows<group_traits<CURVE, GROUP>::Fr, group_traits<CURVE, GROUP>::F>(                \
            bs, first, n, scalars_dev, windows_dev, W, c, shared, skip);                                       \
    }                                                                                                          \
    int name##_msm_sort(const rt::BaseSet &bs, size_t n, const void *scalars_dev, rt::SharedSort *out,        \
                        cudaStream_t s)                                                                        \
    {                                                                                                          \
        return msm::msm_sort<group_traits<CURVE, GROUP>::Fr>(bs, n, scalars_dev, out, s);                      \
    }                                                                                                          \
    int name##_msm_fold(const void *windows, int on_device, int W, int c, size_t shards, void *out)           \
    {                                                                                                          \
        return msm::msm_fold<group_traits<CURVE, GROUP>::F>(windows, on_device, W, c, shards, out);            \
    }                                                                                                          \
    size_t name##_window_point_bytes() { return sizeof(XYZZ<group_traits<CURVE, GROUP>::F>); }                 \
    int name##_scalar_bits() { return group_traits<CURVE, GROUP>::Fr::params::BITS; }

} // namespace b200
