// msm_geometry.hpp — window / bucket geometry of the MSM (host-only arithmetic, shared by the
// kernels' launch code and the C-ABI's b200_msm_num_windows).
#pragma once
#include <cstdint>

namespace b200
{
namespace msm
{

constexpr int SEG = 64;   // max entries per accumulation task
constexpr int SEG_PTS = 16; // max survivors per task after affine rounds: few, short tasks balance badly over the SMs
constexpr int COLLAPSE_TASKS = 4; // floor of Geometry::collapse (see there)
constexpr int CHUNK = 8;  // buckets per reduce thread
constexpr int REDUCE_BLOCK = 128; // reduce threads per CTA (all in one window)

struct Geometry {
    uint32_t n;   // points
    int c;        // window width (bits)
    int W;        // number of windows
    uint32_t B;   // buckets per window = 2^(c-1)
    uint32_t NB;  // W * B
    uint32_t cpw; // chunks per window = B / CHUNK
    uint32_t rb;  // reduce CTA size = min(REDUCE_BLOCK, cpw)
    uint32_t opw; // reduce CTAs (= partial window sums) per window = cpw / rb
    // Buckets with more tasks than this are pre-merged by k_collapse (warp per bucket); up to this
    // many partial sums are added inline by the reduce thread. Scales with the AVERAGE tasks per bucket
    // so that the collapse kernel only ever sees outliers (at 2^24 with c = 17 the average bucket has
    // 4 tasks; a fixed threshold of 4 sent half a million buckets through it: 24 ms).
    uint32_t collapse;
    // Where window w's digits go: bucket w * bstride + |digit| - 1, entry value w * estride + point index.
    // Plain MSM: bstride = B, estride = 0 (W bucket sets over one base array). With precomputed tables
    // [2^(c w)] P (make_table_geometry): bstride = 0, estride = the table's row length — every window feeds
    // the SAME B buckets with the matching multiple of the base, and the W-fold reduction becomes one.
    uint32_t bstride, estride;
    uint32_t chunk; // buckets per reduce thread (CHUNK; smaller over tables, see make_table_geometry)
    // Batched-affine rounds ahead of the XYZZ accumulation (msm_affine.cuh): every bucket is padded to a multiple of
    // 2^rounds entries by the counting sort, `rounds` pairwise-addition rounds leave ceil(count / 2^rounds) survivors per
    // bucket at offsets[b] >> rounds, and the tasks of the accumulation are cut over the survivors. 0 = none.
    uint32_t rounds;
    uint32_t seg;   // task length: SEG entries, or SEG_PTS survivors when rounds > 0
    // Entry encoding. 0: the table index itself, point + window * estride. k > 0: point | window << k — a sort that does
    // not know the row length, so that MSMs over DIFFERENT base sets with the SAME scalars can share it (msm_host.cuh,
    // SharedSort): each decodes with its own row length.
    uint32_t eshift;
};

// Entries of a bucket after padding to a multiple of 2^rounds, and its survivors after the rounds.
inline
#if defined(__CUDACC__)
    __host__ __device__
#endif
    uint32_t padded_count(uint32_t count, uint32_t rounds)
{
    const uint32_t mask = (1u << rounds) - 1u;
    return (count + mask) & ~mask;
}

// How many affine rounds pay: a round over the survivors of the previous one saves ~3.7 of 10 field products per
// addition but costs a fixed ~0.1 ms of dependent small launches (the inversion tree), and the padding costs up to
// 2^rounds - 1 dead slots per bucket. Rounds while the average bucket still holds >= 16 entries, at most 3 (2); none for
// small MSMs, whose accumulation is shorter than the rounds' fixed cost.
inline uint32_t pick_rounds(uint64_t entries, uint64_t buckets)
{
    if (entries < (1ull << 21)) {
        return 0;
    }
    // a third round only for the large MSMs (2^20 points over tables): below, its inversion tree costs more than the
    // products it saves — standalone 2^18: 1.60 ms with two rounds, 1.74 with three; the 8-GPU prover, whose shard
    // MSMs have 2^17 points: 7.6 ms with two, 8.6 with three (profiles/r2_rounds_vs_size.jsonl, r2_prove_multi_rounds_n8.jsonl)
    const uint32_t max_r = entries >= (1ull << 23) ? 3 : 2;
    uint32_t r = 0;
    uint64_t avg = entries / (buckets ? buckets : 1);
    while (r < max_r && avg >= 16) {
        ++r;
        avg >>= 1;
    }
    return r;
}

// Buckets per reduce thread. The reduction is a serial chain of group additions per thread, run at a fraction of
// a warp per scheduler (ncu: 6 % occupancy). While the reduce threads number well under one wave of the machine,
// halving the chunk shortens every chain and costs nothing: over tables (one bucket set, 2^16 buckets) 8 -> 2
// buckets per thread takes alt_bn128 G1 from 0.76 to 0.57 ms and bw6_761 from 4.6 to 3.2 ms
// (profiles/r1s5_reduce_chunk_tune.log). Once the threads approach a full wave (4 reduce CTAs per SM) the kernel
// is throughput-bound and every extra thread's offset multiplication is paid in full: the plain 2^20 geometry
// (17 x 2^14 buckets) with 4 instead of 8 buckets per thread went from 0.56 to 0.75 ms. Hence: the smallest chunk
// that keeps the thread count under about half a wave.
constexpr uint64_t REDUCE_THREADS_TARGET = 40000;
inline uint32_t pick_chunk(uint64_t buckets)
{
    uint32_t chunk = CHUNK;
    while (chunk > 2 && buckets / (chunk / 2) <= REDUCE_THREADS_TARGET) {
        chunk /= 2;
    }
    return chunk;
}
inline void set_chunk(Geometry &g, uint32_t chunk)
{
    if (chunk == 0 || chunk > g.B || g.B % chunk != 0) {
        chunk = g.B < (uint32_t)CHUNK ? g.B : (uint32_t)CHUNK;
    }
    g.chunk = chunk;
    g.cpw = g.B / g.chunk;
    g.rb = g.cpw < (uint32_t)REDUCE_BLOCK ? g.cpw : (uint32_t)REDUCE_BLOCK;
    g.opw = g.cpw / g.rb;
}

// Window width: minimise  W(c) * (10 n + 100 * 2^(c-1))  field mults (10 per mixed add,
// ~100 per bucket in the reduction as measured: the reduction runs at low occupancy), c in [5, 20].
inline Geometry make_geometry(uint64_t n, int scalar_bits, int force_c = 0)
{
    Geometry g;
    g.n = (uint32_t)n;
    int best_c = 5;
    double best = 1e300;
    for (int c = 5; c <= 20; ++c) {
        const int W = (scalar_bits + 1 + c - 1) / c;
        const double cost = (double)W * (10.0 * (double)n + 100.0 * (double)(1u << (c - 1)));
        if (cost < best) {
            best = cost;
            best_c = c;
        }
    }
    g.c = force_c ? force_c : best_c;
    g.W = (scalar_bits + 1 + g.c - 1) / g.c;
    g.B = 1u << (g.c - 1);
    g.NB = (uint32_t)g.W * g.B;
    set_chunk(g, pick_chunk((uint64_t)g.W * g.B));
    const uint64_t avg_tasks = (n / g.B + SEG - 1) / SEG;
    g.collapse = (uint32_t)(2 * avg_tasks + 2 > (uint64_t)COLLAPSE_TASKS ? 2 * avg_tasks + 2 : (uint64_t)COLLAPSE_TASKS);
    g.bstride = g.B;
    g.estride = 0;
    g.rounds = 0;
    g.seg = SEG;
    g.eshift = 0;
    return g;
}

// Sets the number of affine rounds (and the collapse threshold, which counts tasks over the survivors).
inline void set_rounds(Geometry &g, uint32_t rounds, uint64_t entries)
{
    g.rounds = rounds;
    g.seg = rounds ? SEG_PTS : SEG;
    const uint64_t buckets = g.bstride ? (uint64_t)g.W * g.B : g.B;
    const uint64_t avg_tasks = ((entries / buckets >> rounds) + g.seg - 1) / g.seg;
    g.collapse = (uint32_t)(2 * avg_tasks + 2 > (uint64_t)COLLAPSE_TASKS ? 2 * avg_tasks + 2 : (uint64_t)COLLAPSE_TASKS);
}

// Window width for a base set with precomputed tables [2^(c w)] P, w < W: all W n entries land in one set
// of 2^(c-1) buckets, so the reduction is paid once: minimise 10 W(c) n + 100 * 2^(c-1).
inline int table_window_bits(uint64_t n, int scalar_bits)
{
    int best_c = 5;
    double best = 1e300;
    for (int c = 5; c <= 20; ++c) {
        const int W = (scalar_bits + 1 + c - 1) / c;
        const double cost = 10.0 * (double)W * (double)n + 100.0 * (double)(1u << (c - 1));
        if (cost < best) {
            best = cost;
            best_c = c;
        }
    }
    return best_c;
}

// Geometry of an MSM over n points of a table of row length `row` built for window width c.
// The digit extraction sees W windows; everything after the scatter sees ONE window of B buckets.
inline Geometry make_table_geometry(uint64_t n, int scalar_bits, int c, uint64_t row, int chunk = 0)
{
    Geometry g = make_geometry(n, scalar_bits, c);
    g.bstride = 0;
    g.estride = (uint32_t)row;
    g.NB = g.B;
    set_chunk(g, chunk > 0 ? (uint32_t)chunk : pick_chunk(g.B)); // ONE bucket set: W times fewer reduce threads
    const uint64_t avg_tasks = ((uint64_t)g.W * n / g.B + SEG - 1) / SEG;
    g.collapse = (uint32_t)(2 * avg_tasks + 2 > (uint64_t)COLLAPSE_TASKS ? 2 * avg_tasks + 2 : (uint64_t)COLLAPSE_TASKS);
    return g;
}

} // namespace msm
} // namespace b200
