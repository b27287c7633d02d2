// msm_affine.cuh — batched-affine rounds in front of the bucket accumulation.
//
// The bucket loop of libff's multi_exp_inner (multiexp.tcc:540-550, mixed_add into a Jacobian bucket, 11 field
// products) became an XYZZ mixed add in msm.cuh (10 products). The integer pipe is ~87 % busy in that kernel, so the
// only way left to go faster is to spend fewer products per addition: an AFFINE addition
//     lambda = (y2 - y1) / (x2 - x1),  x3 = lambda^2 - x1 - x2,  y3 = lambda (x1 - x3) - y1
// costs 2 products + 1 square + ONE INVERSION, and inversions batch: with Montgomery's trick n inversions cost one
// inversion + 3 (n - 1) products (libff uses the same trick in batch_invert, field_utils.tcc:330-350). One ROUND adds
// the sorted entries of every bucket pairwise — (e0 + e1), (e2 + e3), ... — halving every bucket; all pairs of the
// round, of all buckets, share one tree of batched inversions, so an addition costs ~6.3 products instead of 10.
// After `rounds` rounds (Geometry::rounds, 0-4) the survivors — 1/2^rounds of the entries — go through the XYZZ
// accumulation of msm.cuh as before. The sum of a bucket is a group element: the order of the additions cannot be
// seen in the result, which stays bit-identical to the reference's after normalisation.
//
// Layout. The counting sort pads every bucket to a multiple of 2^rounds entries (pad = NULL_ENTRY), so pair q of round
// r is simply elements (2q, 2q + 1) of the previous round's array and never straddles two buckets; bucket b's
// survivors sit at offsets[b] >> rounds. A pad, and the sum P + (-P), is the affine pair (0, 0) — not a point of
// y^2 = x^3 + b for b != 0 — which every later stage treats as the neutral element.
//
// Ownership. A BLOCK is 2^rounds consecutive entries (-> one survivor); thread t owns blocks t, t + T, t + 2T, ... (AFF_KB of
// them, T = threads) in EVERY round, and the rounds keep their arrays element-major (element e of block m at e * blocks + m):
// the threads of a warp — consecutive blocks — read and write consecutive addresses in every round, and the inversion tree
// above level 0 has the same size in all rounds. A thread's chain of a round = its pairs in the order the up pass
// multiplied them; the down pass walks it backwards. Pex[q] = the product of the chain's elements before pair q.
// One round = 7 launches: up (per pair: denominator d = x2 - x1, running products per thread) -> up, up (two more levels of
// running products, AFF_K elements per thread) -> top (one inversion per remaining element, binary Euclid,
// Fp::inverse_fast) -> down, down (inverses of every level's elements) -> round (inverse of each pair's d, the affine
// addition, result written).
// (Tried and dropped, profiles/r2_fused_rounds.jsonl: carrying the next round's up pass inside the down pass. The products
// are the same, only the re-read of the sums is saved, and holding two sums plus a second running product costs more in
// registers than that: alt_bn128 G1 2.74 -> 3.42 ms, G2 7.5 -> 12.4 ms.)
#pragma once
#include "msm.cuh"

namespace b200
{
namespace msm
{

constexpr int AFF_K = 16;                       // elements per thread at the upper levels of the inversion tree
constexpr int AFF_KB = 4;                       // blocks per thread at level 0 (2^(rounds - r) pairs each in round r)
constexpr uint32_t NULL_ENTRY = 0xffffffffu;    // padding of the entries array

struct AffLevels {
    uint32_t nblk, n1, n2, n3; // blocks; threads of level 0 (= elements of level 1), of level 1, of level 2
};
HD AffLevels aff_levels(uint32_t total_entries, uint32_t rounds)
{
    AffLevels lv;
    lv.nblk = total_entries >> rounds;
    lv.n1 = (lv.nblk + AFF_KB - 1) / AFF_KB;
    lv.n2 = (lv.n1 + AFF_K - 1) / AFF_K;
    lv.n3 = (lv.n2 + AFF_K - 1) / AFF_K;
    return lv;
}

template<class F> HD Affine<F> null_point()
{
    Affine<F> p;
    p.x = F::zero();
    p.y = F::zero();
    return p;
}
template<class F> HD bool is_null(const Affine<F> &p) { return p.x.is_zero() && p.y.is_zero(); }

// Where a round's inputs come from: round 1 reads (index | sign) entries and gathers the bases, later rounds read the
// previous round's points.
template<class F> struct SrcEntries {
    const uint32_t *entries;
    const Affine<F> *bases;
    // A shared sort (Geometry::eshift > 0): entry = point | window << eshift over the scalars of a LONGER vector whose
    // first `skip` points this MSM does not have (the prover's L query starts at variable num_inputs + 1): those read as
    // the neutral element; row = this set's table row length.
    uint32_t eshift = 0, row = 0, skip = 0;
    HD bool locate(uint32_t e, uint32_t &idx) const
    {
        if (e == NULL_ENTRY) {
            return false;
        }
        const uint32_t v = e & 0x7fffffffu;
        if (eshift == 0) {
            idx = v;
            return true;
        }
        const uint32_t pt = v & ((1u << eshift) - 1u);
        if (pt < skip) {
            return false;
        }
        idx = (v >> eshift) * row + (pt - skip);
        return true;
    }
    // element e of block m: the sorted entries are contiguous per block (per = 2^rounds of them)
    HD uint32_t index(uint32_t m, uint32_t e, uint32_t /*nblk*/, uint32_t per) const { return m * per + e; }
    HD Affine<F> load(uint32_t i) const
    {
        const uint32_t e = entries[i];
        uint32_t idx;
        if (!locate(e, idx)) {
            return null_point<F>();
        }
        Affine<F> p = load_vec_gather(bases + idx);
        p.y = p.y.cneg((e >> 31) != 0);
        return p;
    }
    // x only (the common case of the up pass needs nothing else); null reported through the flag
    HD F load_x(uint32_t i, bool &null) const
    {
        uint32_t idx;
        null = !locate(entries[i], idx);
        if (null) {
            return F::zero();
        }
        return load_vec_gather(&bases[idx].x);
    }
};
template<class F> struct SrcPoints {
    const Affine<F> *pts;
    // element e of block m: the rounds keep their arrays element-major (all blocks' element 0, then element 1, ...), so
    // that the threads of a warp — consecutive blocks — touch consecutive addresses
    HD uint32_t index(uint32_t m, uint32_t e, uint32_t nblk, uint32_t /*per*/) const { return e * nblk + m; }
    HD Affine<F> load(uint32_t i) const { return load_vec(pts + i); }
    HD F load_x(uint32_t i, bool &null) const
    {
        const F x = load_vec(&pts[i].x);
        null = false;
        if (x.is_zero()) { // (0, 0) is the neutral element; (0, y != 0) may be a point
            null = load_vec(&pts[i].y).is_zero();
        }
        return x;
    }
};

// What a pair needs inverted, and what kind of pair it is.
enum PairKind { PAIR_ADD = 0, PAIR_DBL = 1, PAIR_NULL = 2, PAIR_KEEP_A = 3, PAIR_KEEP_B = 4 };

template<class F> HD F pair_den(const Affine<F> &A, const Affine<F> &B, int &kind)
{
    const bool an = is_null(A), bn = is_null(B);
    if (an || bn) {
        kind = an ? (bn ? PAIR_NULL : PAIR_KEEP_B) : PAIR_KEEP_A;
        return F::one();
    }
    const F d = B.x - A.x;
    if (d.is_zero()) {
        if (A.y == B.y && !A.y.is_zero()) {
            kind = PAIR_DBL; // P + P: lambda = 3 x^2 / (2 y)
            return A.y.dbl();
        }
        kind = PAIR_NULL; // P + (-P)
        return F::one();
    }
    kind = PAIR_ADD;
    return d;
}

// The same denominator from the x coordinates alone where that is enough (up pass).
template<class F, class Src> HD F pair_den_x(const Src &src, uint32_t ia, uint32_t ib)
{
    bool an, bn;
    const F ax = src.load_x(ia, an);
    const F bx = src.load_x(ib, bn);
    if (an || bn) {
        return F::one();
    }
    const F d = bx - ax;
    if (d.is_zero()) { // rare: needs the y coordinates
        int kind;
        return pair_den(src.load(ia), src.load(ib), kind);
    }
    return d;
}

// ---- level 0, up: thread t multiplies the denominators of its pairs of round r ---------------------------------------
// Odd rounds ascending, even rounds descending (the down pass of a round walks the other way).
// The denominators do not depend on the running product: they are fetched AFF_U pairs ahead of the chain of products so
// that the gathers (round 1: two 32-byte x coordinates per pair out of a table far larger than L2) overlap it.
// Pair i of block m lives at i * nblk + m in Pex and in the rounds' output arrays (element-major, see SrcPoints).
constexpr int AFF_U = 4;
HD uint32_t aff_owned_blocks(uint32_t t, const AffLevels &lv)
{
    uint32_t nb = 0;
    while (nb < (uint32_t)AFF_KB && nb * lv.n1 + t < lv.nblk) {
        ++nb;
    }
    return nb;
}
template<class F, class Src>
HD void aff_up_round_body(uint32_t t, const AffLevels &lv, uint32_t rounds, uint32_t r, const Src &src, F *Pex, F *T1)
{
    const uint32_t L = 1u << (rounds - r), per = 2 * L; // pairs, elements per block in this round
    const bool asc = (r & 1u) != 0;
    const uint32_t nb = aff_owned_blocks(t, lv);
    F acc = F::one();
    bool started = false;
    for (uint32_t jj = 0; jj < nb; ++jj) {
        const uint32_t m = (asc ? jj : nb - 1 - jj) * lv.n1 + t;
        for (uint32_t i0 = 0; i0 < L; i0 += AFF_U) {
            F d[AFF_U];
#pragma unroll
            for (int u = 0; u < AFF_U; ++u) {
                const uint32_t i = asc ? i0 + u : L - 1 - (i0 + u);
                d[u] = i0 + u < L ? pair_den_x<F, Src>(src, src.index(m, 2 * i, lv.nblk, per), src.index(m, 2 * i + 1, lv.nblk, per))
                                  : F::one();
            }
#pragma unroll
            for (int u = 0; u < AFF_U; ++u) {
                if (i0 + u < L) {
                    const uint32_t i = asc ? i0 + u : L - 1 - (i0 + u);
                    store_vec(Pex + i * lv.nblk + m, acc);
                    acc = started ? acc * d[u] : d[u];
                    started = true;
                }
            }
        }
    }
    store_vec(T1 + t, acc);
}

// ---- upper levels, up: the same over an array E of `count` elements, `threads` = ceil(count / AFF_K) -----------------
template<class F> HD void aff_up_body(uint32_t t, uint32_t count, uint32_t threads, const F *E, F *Pre, F *T)
{
    F acc = F::one();
    for (int j0 = 0; j0 < AFF_K; j0 += AFF_U) {
        F e[AFF_U];
#pragma unroll
        for (int u = 0; u < AFF_U; ++u) {
            const uint32_t q = (uint32_t)(j0 + u) * threads + t;
            e[u] = q < count ? load_vec(E + q) : F::one();
        }
#pragma unroll
        for (int u = 0; u < AFF_U; ++u) {
            const uint32_t q = (uint32_t)(j0 + u) * threads + t;
            if (q < count) {
                acc = (j0 + u) ? acc * e[u] : e[u];
                store_vec(Pre + q, acc);
            }
        }
    }
    store_vec(T + t, acc);
}

// ---- upper levels, down: Inv[t] = inverse of thread t's total; E is overwritten by the inverses of its elements ----
template<class F> HD void aff_down_body(uint32_t t, uint32_t count, uint32_t threads, F *E, const F *Pre, const F *Inv)
{
    F inv = load_vec(Inv + t);
    for (int j0 = AFF_K - AFF_U; j0 >= 0; j0 -= AFF_U) {
        F e[AFF_U], prev[AFF_U];
#pragma unroll
        for (int u = AFF_U - 1; u >= 0; --u) {
            const int j = j0 + u;
            const uint32_t q = (uint32_t)j * threads + t;
            if (q < count && j) {
                e[u] = load_vec(E + q);
                prev[u] = load_vec(Pre + (q - threads));
            }
        }
#pragma unroll
        for (int u = AFF_U - 1; u >= 0; --u) {
            const int j = j0 + u;
            const uint32_t q = (uint32_t)j * threads + t;
            if (q >= count) {
                continue;
            }
            if (j == 0) {
                store_vec(E + q, inv);
            } else {
                store_vec(E + q, inv * prev[u]);
                inv = inv * e[u];
            }
        }
    }
}

// ---- level 0, down: inverse of each pair's denominator, then the affine addition -----------------------------------
template<class F> HD Affine<F> pair_add(const Affine<F> &A, const Affine<F> &B, int kind, const F &dinv)
{
    F num = B.y - A.y;
    if (kind == PAIR_DBL) {
        const F x2 = A.x.sqr();
        num = x2.dbl() + x2;
    }
    const F lam = num * dinv;
    Affine<F> r;
    r.x = lam.sqr() - A.x - B.x;
    r.y = lam * (A.x - r.x) - A.y;
    if (kind >= PAIR_NULL) {
        r = kind == PAIR_NULL ? null_point<F>() : (kind == PAIR_KEEP_A ? A : B);
    }
    return r;
}

// Round r of `rounds` for thread t: the down pass over its chain (Pex, inverse of its total in I1) with the affine additions.
template<class F, class Src>
HD void aff_round_body(uint32_t t, const AffLevels &lv, uint32_t rounds, uint32_t r, const Src &src, const F *Pex, const F *I1,
                       Affine<F> *out)
{
    const uint32_t L = 1u << (rounds - r), per = 2 * L; // pairs per block in this round (1 in the last)
    const bool desc = (r & 1u) != 0;                    // odd rounds were multiplied up ascending: walk them descending
    const uint32_t nb = aff_owned_blocks(t, lv);
    F inv = load_vec(I1 + t);
    for (uint32_t jj = 0; jj < nb; ++jj) {
        const uint32_t m = (desc ? nb - 1 - jj : jj) * lv.n1 + t;
        for (uint32_t ii = 0; ii < L; ++ii) {
            const uint32_t i = desc ? L - 1 - ii : ii;
            const Affine<F> A = src.load(src.index(m, 2 * i, lv.nblk, per)), B = src.load(src.index(m, 2 * i + 1, lv.nblk, per));
            int kind;
            const F d = pair_den(A, B, kind);
            const F dinv = inv * load_vec(Pex + i * lv.nblk + m);
            inv = inv * d;
            store_vec(out + i * lv.nblk + m, pair_add(A, B, kind, dinv));
        }
    }
}

// ---- the XYZZ accumulation over the survivors (points instead of entries; (0, 0) skipped) ---------------------------
template<class F> HD void accumulate_task_pts(const Task &tk, const Affine<F> *pts, XYZZ<F> *out)
{
    XYZZ<F> acc = XYZZ<F>::inf();
#if defined(B200_ACC_NO_PREFETCH)
    for (uint32_t k = 0; k < tk.len; ++k) {
        const Affine<F> cur = load_vec(pts + tk.start + k);
        if (!is_null(cur)) {
            madd(acc, cur);
        }
    }
#else
    Affine<F> p = load_vec(pts + tk.start);
    for (uint32_t k = 0; k < tk.len; ++k) {
        const Affine<F> cur = p;
        if (k + 1 < tk.len) { // the next survivor is fetched while this one is added
            p = load_vec(pts + tk.start + k + 1);
        }
        if (!is_null(cur)) {
            madd(acc, cur);
        }
    }
#endif
    store_vec(out, acc);
}

#if defined(__CUDACC__)
constexpr int AFF_THREADS = 128;
// Resident CTAs per SM the level-0 kernels are compiled for (caps their registers; per field, csrc/Makefile)
#ifndef B200_AFF_MIN_BLOCKS
#define B200_AFF_MIN_BLOCKS 1
#endif

template<class F, class Src>
__global__ void __launch_bounds__(AFF_THREADS, B200_AFF_MIN_BLOCKS) k_aff_up_round(const uint32_t *total_entries, uint32_t rounds,
                                                                                   uint32_t r, Src src, F *Pex, F *T1)
{
    const AffLevels lv = aff_levels(*total_entries, rounds);
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < lv.n1) {
        aff_up_round_body<F, Src>(t, lv, rounds, r, src, Pex, T1);
    }
}

template<class F, class Src>
__global__ void __launch_bounds__(AFF_THREADS, B200_AFF_MIN_BLOCKS) k_aff_round(const uint32_t *total_entries, uint32_t rounds,
                                                                                uint32_t r, Src src, const F *Pex, const F *I1,
                                                                                Affine<F> *out)
{
    const AffLevels lv = aff_levels(*total_entries, rounds);
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < lv.n1) {
        aff_round_body<F, Src>(t, lv, rounds, r, src, Pex, I1, out);
    }
}

// Survivors through the XYZZ path: same CTA-local ordering by task length as k_accumulate.
template<class F>
__global__ void __launch_bounds__(ACC_THREADS, B200_ACC_MIN_BLOCKS) k_accumulate_pts(
    const Task *tasks, const uint32_t *num_tasks, const Affine<F> *pts, XYZZ<F> *partial)
{
    const uint32_t T = *num_tasks;
    const uint32_t id = blockIdx.x * ACC_THREADS + threadIdx.x;
    if (id < T) {
        accumulate_task_pts<F>(tasks[id], pts, partial + id);
    }
}
#endif

} // namespace msm
} // namespace b200
