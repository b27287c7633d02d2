// tests/host_emu/emu.cpp — TEST INFRASTRUCTURE.
// Compiles the device field / curve templates (libsnark_b200/csrc/{fp,curve}.cuh) for the
// HOST through the carry-flag emulation in prims.cuh, so their arithmetic can be checked
// against the reference on a machine without a GPU. Not linked into the product library.
#include <cstdint>
#include <cstring>
#include <vector>

#include "curves.cuh"
#include "batch_exp.cuh"
#include "host_fold.hpp"
#include "msm.cuh"
#include "msm_affine.cuh"
#include "ntt.cuh"

using namespace b200;

template<class F> static int field_op(int op, const void *a_, const void *b_, void *o_, size_t n)
{
    const F *a = (const F *)a_;
    const F *b = (const F *)b_;
    F *o = (F *)o_;
    for (size_t i = 0; i < n; ++i) {
        switch (op) {
        case 0: o[i] = a[i] + b[i]; break;
        case 1: o[i] = a[i] - b[i]; break;
        case 2: o[i] = a[i] * b[i]; break;
        case 3: o[i] = a[i].sqr(); break;
        case 4: o[i] = a[i].neg(); break;
        case 5: o[i] = a[i].is_zero() ? a[i] : a[i].inverse(); break;
        default: return -1;
        }
    }
    return 0;
}

template<class F> static int fp_convert(int to_mont, const void *a_, void *o_, size_t n)
{
    const F *a = (const F *)a_;
    F *o = (F *)o_;
    for (size_t i = 0; i < n; ++i) {
        o[i] = to_mont ? a[i].to_mont() : a[i].from_mont();
    }
    return 0;
}

// Points travel in the reference's Jacobian (X,Y,Z) layout; Z in {0 -> infinity, one -> affine} on input.
// op: 0 = add (a, b affine or inf) via XYZZ add, 1 = madd (a += b), 2 = dbl, 3 = add via madd-chain check.
// out: affine X|Y|Z(one) or the all-zero-X, Y=one, Z=0 encoding is NOT produced here: out is (x, y, flag).
template<class F> static int group_op(int op, const void *a_, const void *b_, void *o_, size_t n)
{
    struct J {
        F x, y, z;
    };
    const J *a = (const J *)a_;
    const J *b = (const J *)b_;
    J *o = (J *)o_;
    for (size_t i = 0; i < n; ++i) {
        XYZZ<F> A = a[i].z.is_zero() ? XYZZ<F>::inf() : XYZZ<F>::from_affine(Affine<F>{a[i].x, a[i].y});
        XYZZ<F> B = b[i].z.is_zero() ? XYZZ<F>::inf() : XYZZ<F>::from_affine(Affine<F>{b[i].x, b[i].y});
        // de-normalise A so the general formulas are exercised: A <- 2A - A would need neg; use dbl+add paths.
        switch (op) {
        case 0: add(A, B); break;
        case 1:
            if (!b[i].z.is_zero()) {
                madd(A, Affine<F>{b[i].x, b[i].y});
            }
            break;
        case 2: dbl(A); break;
        case 3: {
            // (2A) + B through the general add with a non-trivial ZZ on the left
            dbl(A);
            add(A, B);
            break;
        }
        case 4: {
            // (2A) + B through madd with a non-trivial ZZ on the left
            dbl(A);
            if (!b[i].z.is_zero()) {
                madd(A, Affine<F>{b[i].x, b[i].y});
            }
            break;
        }
        case 5: {
            // A + (2B): non-trivial ZZ on the right
            dbl(B);
            add(A, B);
            break;
        }
        default: return -1;
        }
        Affine<F> r;
        if (to_affine(A, r)) {
            o[i].x = r.x;
            o[i].y = r.y;
            o[i].z = F::one();
        } else {
            o[i].x = F::zero();
            o[i].y = F::one();
            o[i].z = F::zero();
        }
    }
    return 0;
}

#define FIELD_DISPATCH(fn, ...)                                         \
    switch (curve * 4 + field) {                                        \
    case 0: return fn<alt_bn128_fr>(__VA_ARGS__);                       \
    case 1: return fn<alt_bn128_fq>(__VA_ARGS__);                       \
    case 4: return fn<bls12_377_fr>(__VA_ARGS__);                       \
    case 5: return fn<bls12_377_fq>(__VA_ARGS__);                       \
    case 8: return fn<bw6_761_fr>(__VA_ARGS__);                         \
    case 9: return fn<bw6_761_fq>(__VA_ARGS__);                         \
    }

extern "C" int emu_field_op(int curve, int field, int op, const void *a, const void *b, void *o, size_t n)
{
    FIELD_DISPATCH(field_op, op, a, b, o, n)
    if (curve == 0 && field == 2) return field_op<alt_bn128_fq2>(op, a, b, o, n);
    if (curve == 1 && field == 2) return field_op<bls12_377_fq2>(op, a, b, o, n);
    if (curve == 2 && field == 2) return field_op<bw6_761_fq>(op, a, b, o, n);
    return -100;
}

extern "C" int emu_fp_convert(int curve, int field, int to_mont, const void *a, void *o, size_t n)
{
    FIELD_DISPATCH(fp_convert, to_mont, a, o, n)
    return -100;
}

extern "C" int emu_group_op(int curve, int group, int op, const void *a, const void *b, void *o, size_t n)
{
    switch (curve * 4 + group) {
    case 1: return group_op<alt_bn128_fq>(op, a, b, o, n);
    case 2: return group_op<alt_bn128_fq2>(op, a, b, o, n);
    case 5: return group_op<bls12_377_fq>(op, a, b, o, n);
    case 6: return group_op<bls12_377_fq2>(op, a, b, o, n);
    case 9: return group_op<bw6_761_fq>(op, a, b, o, n);
    case 10: return group_op<bw6_761_fq>(op, a, b, o, n);
    }
    return -100;
}


// ---- MSM index logic (msm.cuh bodies) on the host -------------------------------------------
// Signed window digits as the device computes them (prepare_body: from_mont + offset constant K,
// then window_digit), for comparison with libff's field_get_signed_digit.
template<class Fr> static int signed_digits(const void *scalars, size_t n, int c, int64_t *out, int *W_out)
{
    msm::Geometry g = msm::make_geometry(n, Fr::params::BITS, c);
    *W_out = g.W;
    const msm::Limbs<Fr::N> K = msm::make_offset_constant<Fr::N>(g);
    std::vector<uint32_t> t(n * Fr::N);
    for (size_t i = 0; i < n; ++i) {
        msm::prepare_body<Fr>(i, (const Fr *)scalars, t.data(), K, nullptr);
        for (int w = 0; w < g.W; ++w) {
            out[i * g.W + w] = msm::window_digit(t.data() + i * Fr::N, Fr::N, g, w);
        }
    }
    return 0;
}

extern "C" int emu_signed_digits(int curve, const void *scalars, size_t n, int c, int64_t *out, int *W_out)
{
    switch (curve) {
    case 0: return signed_digits<alt_bn128_fr>(scalars, n, c, out, W_out);
    case 1: return signed_digits<bls12_377_fr>(scalars, n, c, out, W_out);
    case 2: return signed_digits<bw6_761_fr>(scalars, n, c, out, W_out);
    }
    return -100;
}

// The whole MSM pipeline of msm.cuh executed sequentially on the host (same bodies, same index
// logic: histogram -> scan -> scatter -> tasks -> accumulate -> reduce -> fold).
// table = true: the precomputed-table variant (b200_bases_precompute): build_table_body fills rows [2^(c w)] P,
// every window scatters into the one bucket set, the reduction runs over a single window.
// shared_skip >= 0: the sort is a SHARED one (rt::SharedSort, msm_host.cuh) over all n scalars with entries
// point | window << eshift, consumed by a base set that lacks the first shared_skip points (bases_ref holds n - shared_skip
// points, its table rows are that long): the prover's L query against the witness sort. Needs table and rounds > 0.
template<class Fr, class F>
static int msm_emulated(const void *bases_ref, size_t n, const void *scalars, int force_c, void *out, bool table = false,
                        int rounds = 0, long shared_skip = -1)
{
    using namespace msm;
    const bool shared = shared_skip >= 0;
    if (shared && (!table || rounds == 0 || (size_t)shared_skip >= n)) {
        return -9;
    }
    const size_t nb = shared ? n - (size_t)shared_skip : n; // points of the consuming base set
    const int tc = force_c ? force_c : table_window_bits(n, Fr::params::BITS);
    Geometry g = table ? make_table_geometry(n, Fr::params::BITS, tc, nb) : make_geometry(n, Fr::params::BITS, force_c);
    set_rounds(g, (uint32_t)rounds, (uint64_t)g.W * n);
    if (shared) {
        uint32_t k = 1;
        while ((1ull << k) < n) {
            ++k;
        }
        g.eshift = k;
    }
    const uint32_t R = g.rounds;
    const Limbs<Fr::N> K = make_offset_constant<Fr::N>(g);
    std::vector<Affine<F>> bases(table ? nb * g.W : nb);
    std::vector<uint8_t> inf(n);
    uint32_t any = 0;
    for (size_t i = 0; i < nb; ++i) {
        convert_base_body<F>(i, (const uint8_t *)bases_ref, sizeof(RefPoint<F>), bases.data(), inf.data(), &any);
    }
    if (shared && any) {
        return -9; // a shared sort cannot zero the scalars of one set's infinity bases
    }
    if (table) {
        uint32_t fail = 0;
        for (size_t i = 0; i < nb; ++i) {
            build_table_body<F>(i, nb, g.c, g.W, bases.data(), any ? inf.data() : nullptr, &fail);
        }
        if (fail) {
            return -7;
        }
    }
    std::vector<uint32_t> t(n * Fr::N), counts(g.NB + 1, 0), offsets(g.NB + 1), cursor(g.NB + 1), task_off(g.NB + 1);
    for (size_t i = 0; i < n; ++i) {
        prepare_body<Fr>(i, (const Fr *)scalars, t.data(), K, any ? inf.data() : nullptr);
    }
    for (size_t i = 0; i < n; ++i) {
        histogram_body(i, t.data(), Fr::N, g, counts.data());
    }
    uint32_t off = 0, toff = 0;
    for (uint32_t b = 0; b < g.NB; ++b) { // k_scan_*: counts padded to a multiple of 2^rounds
        offsets[b] = cursor[b] = off;
        task_off[b] = toff;
        const uint32_t pc = padded_count(counts[b], R);
        off += pc;
        toff += ((pc >> R) + g.seg - 1) / g.seg;
    }
    offsets[g.NB] = off;
    task_off[g.NB] = toff;
    std::vector<uint32_t> entries(off + 1, NULL_ENTRY), multi_list(g.NB + 1);
    uint32_t multi_count = 0;
    for (size_t i = 0; i < n; ++i) {
        scatter_body(i, t.data(), Fr::N, g, cursor.data(), entries.data());
    }
    std::vector<Task> tasks(toff + 1);
    for (size_t b = 0; b < g.NB; ++b) {
        build_tasks_body(b, counts.data(), offsets.data(), task_off.data(), tasks.data(), &multi_count, multi_list.data(), g.collapse, R, g.seg);
    }
    std::vector<XYZZ<F>> partial(toff + 1);
    if (R) {
        // the affine rounds of msm_affine.cuh, launch by launch, every "thread" in turn
        std::vector<Affine<F>> buf[2];
        const Affine<F> *prev = nullptr;
        const AffLevels lv = aff_levels(off, R);
        std::vector<F> Pex(off / 2 + 2, F::zero()), T1(lv.n1 + 1, F::zero());
        std::vector<F> P1(lv.n1 + 1), T2(lv.n2 + 1), P2(lv.n2 + 1), T3(lv.n3 + 1);
        SrcEntries<F> se{entries.data(), bases.data()};
        if (shared) {
            se.eshift = g.eshift;
            se.row = (uint32_t)nb;
            se.skip = (uint32_t)shared_skip;
        }
        for (uint32_t r = 1; r <= R; ++r) {
            std::vector<Affine<F>> &o = buf[(r - 1) & 1];
            o.assign((off >> r) + 1, null_point<F>());
            const SrcPoints<F> sp{prev};
            for (uint32_t t = 0; t < lv.n1; ++t) {
                if (r == 1) {
                    aff_up_round_body<F, SrcEntries<F>>(t, lv, R, r, se, Pex.data(), T1.data());
                } else {
                    aff_up_round_body<F, SrcPoints<F>>(t, lv, R, r, sp, Pex.data(), T1.data());
                }
            }
            for (uint32_t t = 0; t < lv.n2; ++t) {
                aff_up_body<F>(t, lv.n1, lv.n2, T1.data(), P1.data(), T2.data());
            }
            for (uint32_t t = 0; t < lv.n3; ++t) {
                aff_up_body<F>(t, lv.n2, lv.n3, T2.data(), P2.data(), T3.data());
            }
            for (uint32_t t = 0; t < lv.n3; ++t) {
                T3[t] = T3[t].inverse_fast();
            }
            for (uint32_t t = 0; t < lv.n3; ++t) {
                aff_down_body<F>(t, lv.n2, lv.n3, T2.data(), P2.data(), T3.data());
            }
            for (uint32_t t = 0; t < lv.n2; ++t) {
                aff_down_body<F>(t, lv.n1, lv.n2, T1.data(), P1.data(), T2.data());
            }
            for (uint32_t t = 0; t < lv.n1; ++t) {
                if (r == 1) {
                    aff_round_body<F, SrcEntries<F>>(t, lv, R, r, se, Pex.data(), T1.data(), o.data());
                } else {
                    aff_round_body<F, SrcPoints<F>>(t, lv, R, r, sp, Pex.data(), T1.data(), o.data());
                }
            }
            prev = o.data();
        }
        for (uint32_t k = 0; k < toff; ++k) {
            accumulate_task_pts<F>(tasks[k], prev, &partial[k]);
        }
    } else {
        for (uint32_t k = 0; k < toff; ++k) {
            accumulate_task<F>(tasks[k], entries.data(), bases.data(), &partial[k]);
        }
    }
    for (uint32_t k = 0; k < multi_count; ++k) { // k_collapse
        const uint32_t b = multi_list[k];
        XYZZ<F> acc = XYZZ<F>::inf();
        for (uint32_t q = task_off[b]; q < task_off[b + 1]; ++q) {
            add(acc, partial[q]);
        }
        partial[task_off[b]] = acc;
    }
    std::vector<XYZZ<F>> windows(g.W, XYZZ<F>::inf());
    const size_t reduce_windows = table ? 1 : (size_t)g.W;
    for (size_t gidx = 0; gidx < reduce_windows * g.cpw; ++gidx) {
        XYZZ<F> s;
        reduce_chunk_body<F>(gidx, g, task_off.data(), partial.data(), &s);
        add(windows[gidx / g.cpw], s);
    }
    const RefPoint<F> r = host_fold<F>(windows.data(), g.W, g.c, 1);
    memcpy(out, &r, sizeof(r));
    return 0;
}

extern "C" int emu_msm_shared(int curve, int group, const void *bases, size_t n, const void *scalars, int c, int rounds, long skip,
                              void *out)
{
    switch (curve * 4 + group) {
    case 1: return msm_emulated<alt_bn128_fr, alt_bn128_fq>(bases, n, scalars, c, out, true, rounds, skip);
    case 2: return msm_emulated<alt_bn128_fr, alt_bn128_fq2>(bases, n, scalars, c, out, true, rounds, skip);
    case 5: return msm_emulated<bls12_377_fr, bls12_377_fq>(bases, n, scalars, c, out, true, rounds, skip);
    }
    return -100;
}

extern "C" int emu_msm_table(int curve, int group, const void *bases, size_t n, const void *scalars, int c, void *out)
{
    switch (curve * 4 + group) {
    case 1: return msm_emulated<alt_bn128_fr, alt_bn128_fq>(bases, n, scalars, c, out, true);
    case 2: return msm_emulated<alt_bn128_fr, alt_bn128_fq2>(bases, n, scalars, c, out, true);
    case 5: return msm_emulated<bls12_377_fr, bls12_377_fq>(bases, n, scalars, c, out, true);
    case 6: return msm_emulated<bls12_377_fr, bls12_377_fq2>(bases, n, scalars, c, out, true);
    case 9:
    case 10: return msm_emulated<bw6_761_fr, bw6_761_fq>(bases, n, scalars, c, out, true);
    }
    return -100;
}

// batch_exp.cuh on the host: table build, one exp_body per scalar, normalise_body per group of K results.
template<class Fr, class F> static int batch_exp_emulated(const void *g_ref, const void *coeff, const void *scalars, size_t n, void *out)
{
    msm::RefPoint<F> g;
    memcpy(&g, g_ref, sizeof(g));
    std::vector<Affine<F>> table;
    if (!bexp::build_table<Fr, F>(g, table)) {
        return -7;
    }
    std::vector<XYZZ<F>> acc(n);
    for (size_t i = 0; i < n; ++i) {
        bexp::exp_body<Fr, F>(i, (const Fr *)scalars, (const Fr *)coeff, table.data(), acc.data());
    }
    const size_t K = bexp::norm_group<F>();
    for (size_t t = 0; t * K < n; ++t) {
        bexp::normalise_body<F>(t, n, acc.data(), (msm::RefPoint<F> *)out);
    }
    return 0;
}

// msm_geometry.hpp as the launch code sees it: {c, W, B, NB, cpw, rb, opw, collapse, bstride, estride, chunk}
extern "C" int emu_geometry(uint64_t n, int scalar_bits, int table, uint32_t *out)
{
    using namespace msm;
    const Geometry g = table ? make_table_geometry(n, scalar_bits, table_window_bits(n, scalar_bits), n)
                             : make_geometry(n, scalar_bits);
    const uint32_t v[11] = {(uint32_t)g.c, (uint32_t)g.W, g.B, g.NB, g.cpw, g.rb, g.opw, g.collapse, g.bstride, g.estride, g.chunk};
    memcpy(out, v, sizeof(v));
    return 0;
}

extern "C" int emu_batch_exp(int curve, int group, const void *g, const void *coeff, const void *scalars, size_t n, void *out)
{
    switch (curve * 4 + group) {
    case 1: return batch_exp_emulated<alt_bn128_fr, alt_bn128_fq>(g, coeff, scalars, n, out);
    case 2: return batch_exp_emulated<alt_bn128_fr, alt_bn128_fq2>(g, coeff, scalars, n, out);
    case 5: return batch_exp_emulated<bls12_377_fr, bls12_377_fq>(g, coeff, scalars, n, out);
    case 6: return batch_exp_emulated<bls12_377_fr, bls12_377_fq2>(g, coeff, scalars, n, out);
    case 9:
    case 10: return batch_exp_emulated<bw6_761_fr, bw6_761_fq>(g, coeff, scalars, n, out);
    }
    return -100;
}

extern "C" int emu_msm(int curve, int group, const void *bases, size_t n, const void *scalars, int force_c, void *out)
{
    switch (curve * 4 + group) {
    case 1: return msm_emulated<alt_bn128_fr, alt_bn128_fq>(bases, n, scalars, force_c, out);
    case 2: return msm_emulated<alt_bn128_fr, alt_bn128_fq2>(bases, n, scalars, force_c, out);
    case 5: return msm_emulated<bls12_377_fr, bls12_377_fq>(bases, n, scalars, force_c, out);
    case 6: return msm_emulated<bls12_377_fr, bls12_377_fq2>(bases, n, scalars, force_c, out);
    case 9:
    case 10: return msm_emulated<bw6_761_fr, bw6_761_fq>(bases, n, scalars, force_c, out);
    }
    return -100;
}

// The pipeline with `rounds` batched-affine rounds ahead of the XYZZ accumulation (msm_affine.cuh), plain or over tables.
extern "C" int emu_msm_affine(int curve, int group, const void *bases, size_t n, const void *scalars, int force_c, int table,
                              int rounds, void *out)
{
    switch (curve * 4 + group) {
    case 1: return msm_emulated<alt_bn128_fr, alt_bn128_fq>(bases, n, scalars, force_c, out, table != 0, rounds);
    case 2: return msm_emulated<alt_bn128_fr, alt_bn128_fq2>(bases, n, scalars, force_c, out, table != 0, rounds);
    case 5: return msm_emulated<bls12_377_fr, bls12_377_fq>(bases, n, scalars, force_c, out, table != 0, rounds);
    case 6: return msm_emulated<bls12_377_fr, bls12_377_fq2>(bases, n, scalars, force_c, out, table != 0, rounds);
    case 9:
    case 10: return msm_emulated<bw6_761_fr, bw6_761_fq>(bases, n, scalars, force_c, out, table != 0, rounds);
    }
    return -100;
}

// Fp::inverse_fast (binary Euclid) on the host emulation: op 6 of emu_field_op would do, kept separate for clarity.
template<class F> static int inv_fast(const void *a_, void *o_, size_t n)
{
    const F *a = (const F *)a_;
    F *o = (F *)o_;
    for (size_t i = 0; i < n; ++i) {
        o[i] = a[i].inverse_fast();
    }
    return 0;
}
extern "C" int emu_inverse_fast(int curve, int field, const void *a, void *o, size_t n)
{
    FIELD_DISPATCH(inv_fast, a, o, n)
    if (curve == 0 && field == 2) return inv_fast<alt_bn128_fq2>(a, o, n);
    if (curve == 1 && field == 2) return inv_fast<bls12_377_fq2>(a, o, n);
    if (curve == 2 && field == 2) return inv_fast<bw6_761_fq>(a, o, n);
    return -100;
}

// ---- the NTT pass kernel on the host ---------------------------------------------------------------------------------
// k_ntt_pass (ntt.cuh) run as plain loops: the CTA's shared-memory tile is a host array, the phases between barriers are
// called once with (tid, nthreads) = (0, 1) so that each covers the whole tile, and the passes, twiddle table, coset
// tables and constants come from the same plan_transform / pow_table_body / domain_consts_body the launcher uses
// (ntt.cu: get_domain, coset_table, run_transform, ntt_op_dev — mirrored here for basic radix-2 domains).
template<class Fr> static void emu_pass(const ntt::PassLaunch &pl, const Fr *in, Fr *out, const Fr *tw, const Fr *pre, const Fr *post,
                                        const Fr *post_const)
{
    const ntt::PassParams &pp = pl.pp;
    std::vector<ntt::Vec4> smem(pl.smem / sizeof(ntt::Vec4) + 1);
    for (unsigned b = 0; b < pl.blocks; ++b) {
        ntt::Tile<Fr> tile;
        tile.base = smem.data();
        tile.row_stride = (1u << pp.lgc) + 1;
        tile.plane = (1u << pp.deg) * tile.row_stride;
        const uint64_t index0 = (uint64_t)b << pp.lgc;
        ntt::load_phase<Fr>(0, 1, pp, index0, in, pre, tile);
        int r = 0;
        for (int st = 0; st < pp.nsteps; ++st) {
            const int q = pp.step_q[st];
            if (q == 3) {
                ntt::step_phase<Fr, 3>(0, 1, pp, r, index0, tw, tile);
            } else if (q == 2) {
                ntt::step_phase<Fr, 2>(0, 1, pp, r, index0, tw, tile);
            } else {
                ntt::step_phase<Fr, 1>(0, 1, pp, r, index0, tw, tile);
            }
            r += q;
        }
        ntt::store_phase<Fr>(0, 1, pp, index0, out, post, post_const, tile);
    }
}

// one transform of 2^lgn elements in place (run_transform of ntt.cu): twiddles and constants for the size, the passes
template<class Fr> struct EmuDomain {
    int lgn;
    std::vector<Fr> consts, pw, pg, pgi, tw;
    EmuDomain(int lgn_, const Fr &g) : lgn(lgn_), consts(8), pw(64), pg(64), pgi(64), tw(((size_t)1 << lgn_) / 2 + 1)
    {
        ntt::domain_consts_body<Fr>(lgn, &g, consts.data(), pw.data(), pg.data(), pgi.data());
        for (uint64_t e = 0; e < ((uint64_t)1 << lgn) / 2; ++e) {
            ntt::pow_table_body<Fr>(e, pw.data(), lgn - 1, nullptr, tw.data());
        }
    }
    void run(Fr *data, int inverse, const Fr *pre, const Fr *post, const Fr *post_const) const
    {
        const size_t m = (size_t)1 << lgn;
        ntt::PassLaunch plan[8];
        const int P = ntt::plan_transform(lgn, inverse, Fr::N, plan);
        std::vector<Fr> a(data, data + m), b(m);
        Fr *cur = a.data(), *nxt = b.data();
        for (int i = 0; i < P; ++i) {
            emu_pass<Fr>(plan[i], cur, nxt, tw.data(), i == 0 ? pre : nullptr, i == P - 1 ? post : nullptr,
                         i == P - 1 ? post_const : nullptr);
            std::swap(cur, nxt);
        }
        memcpy(data, cur, m * sizeof(Fr));
    }
};
template<class Fr> static Fr emu_generator(const void *g_)
{
    Fr g;
    if (g_) {
        memcpy(&g, g_, sizeof(Fr));
    } else {
        for (int i = 0; i < Fr::N; ++i) {
            g.v[i] = Fr::params::generator(i);
        }
    }
    return g;
}

// op: 0 FFT, 1 iFFT, 2 cosetFFT(g), 3 icosetFFT(g); in place on m = 2^lgn elements (ntt_op_dev, basic domain)
template<class Fr> static int emu_ntt_t(int op, void *data, size_t m, const void *g_)
{
    int lgn = 0;
    while (((size_t)1 << lgn) < m) {
        ++lgn;
    }
    if (m < 2 || ((size_t)1 << lgn) != m || lgn > Fr::params::S || op < 0 || op > 3) {
        return -2;
    }
    const EmuDomain<Fr> d(lgn, emu_generator<Fr>(g_));
    std::vector<Fr> table(m);
    const Fr *pre = nullptr, *post = nullptr, *post_const = nullptr;
    if (op == 2) { // a[i] *= g^i fused into the first pass's loads
        for (uint64_t e = 0; e < m; ++e) {
            ntt::pow_table_body<Fr>(e, d.pg.data(), lgn, nullptr, table.data());
        }
        pre = table.data();
    } else if (op == 3) { // g^-i / m fused into the last pass's stores
        for (uint64_t e = 0; e < m; ++e) {
            ntt::pow_table_body<Fr>(e, d.pgi.data(), lgn, &d.consts[1], table.data());
        }
        post = table.data();
    } else if (op == 1) {
        post_const = &d.consts[1];
    }
    d.run((Fr *)data, (op == 1 || op == 3) ? 1 : 0, pre, post, post_const);
    return 0;
}

// extended_radix2_domain of size m = 2 small (ext_op_dev / ext_fft / ext_ifft / ext_divide_by_z_dev of ntt.cu), any
// small = 2^k: op 0-3 as above, 4 = divide_by_Z_on_coset
template<class Fr> static int emu_ext_ntt_t(int op, void *data_, size_t m, const void *g_)
{
    int lgn = 0;
    while (((size_t)1 << lgn) < m) {
        ++lgn;
    }
    if (m < 4 || ((size_t)1 << lgn) != m || lgn - 1 > Fr::params::S || op < 0 || op > 4) {
        return -2;
    }
    const int lg_small = lgn - 1;
    const uint64_t small = m / 2;
    Fr *a = (Fr *)data_;
    Fr hc[4], shift;
    ntt::ext_domain_consts_host<Fr>(lg_small, hc, &shift);
    if (op == 4) {
        for (uint64_t i = 0; i < m; ++i) {
            ntt::ext_divz_body<Fr>(i, small, a, hc + 2);
        }
        return 0;
    }
    const int lg_lo_s = lg_small < ntt::EXT_LG_LO ? lg_small : ntt::EXT_LG_LO;
    const size_t ns = ((size_t)1 << lg_lo_s) + (small >> lg_lo_s);
    std::vector<Fr> stab(2 * ns);
    ntt::ext_pow_tables_host<Fr>(shift, small, lg_lo_s, stab.data());
    const ntt::PowTable<Fr> spow{stab.data(), stab.data() + ((size_t)1 << lg_lo_s), lg_lo_s};
    const ntt::PowTable<Fr> sinv{stab.data() + ns, stab.data() + ns + ((size_t)1 << lg_lo_s), lg_lo_s};
    const Fr g = emu_generator<Fr>(g_);
    const int lg_lo_g = lgn < ntt::EXT_LG_LO ? lgn : ntt::EXT_LG_LO;
    const size_t ng = ((size_t)1 << lg_lo_g) + (m >> lg_lo_g);
    std::vector<Fr> gtab(2 * ng);
    if (op >= 2) {
        ntt::ext_pow_tables_host<Fr>(g, m, lg_lo_g, gtab.data());
    }
    const ntt::PowTable<Fr> gpow{gtab.data(), gtab.data() + ((size_t)1 << lg_lo_g), lg_lo_g};
    const ntt::PowTable<Fr> ginv{gtab.data() + ng, gtab.data() + ng + ((size_t)1 << lg_lo_g), lg_lo_g};
    const EmuDomain<Fr> d(lg_small, g);
    const auto fft = [&]() {
        for (uint64_t i = 0; i < small; ++i) {
            ntt::ext_fft_pre_body<Fr>(i, small, a, spow, hc);
        }
        d.run(a, 0, nullptr, nullptr, nullptr);
        d.run(a + small, 0, nullptr, nullptr, nullptr);
    };
    const auto ifft = [&]() {
        d.run(a, 1, nullptr, nullptr, nullptr);
        d.run(a + small, 1, nullptr, nullptr, nullptr);
        for (uint64_t i = 0; i < small; ++i) {
            ntt::ext_ifft_post_body<Fr>(i, small, a, sinv, hc);
        }
    };
    switch (op) {
    case 0: fft(); break;
    case 1: ifft(); break;
    case 2:
        for (uint64_t i = 0; i < m; ++i) {
            ntt::mul_pow_body<Fr>(i, a, gpow);
        }
        fft();
        break;
    case 3:
        ifft();
        for (uint64_t i = 0; i < m; ++i) {
            ntt::mul_pow_body<Fr>(i, a, ginv);
        }
        break;
    }
    return 0;
}

// step_radix2_domain of size m = big + small (step_fft / step_ifft / step_op_dev / step_divide_by_z_dev of ntt.cu):
// op 0-3 as above, 4 = divide_by_Z_on_coset
template<class Fr> static int emu_step_ntt_t(int op, void *data_, size_t m, const void *g_)
{
    if (m < 3 || (m & (m - 1)) == 0 || op < 0 || op > 4) {
        return -2;
    }
    int lg_big = 0;
    while (((size_t)1 << (lg_big + 1)) <= m) {
        ++lg_big;
    }
    const uint64_t big = 1ull << lg_big, small = m - big;
    if ((small & (small - 1)) || lg_big + 1 > Fr::params::S) {
        return -2;
    }
    int lg_small = 0;
    while (((uint64_t)1 << lg_small) < small) {
        ++lg_small;
    }
    const uint64_t compr = big / small;
    Fr *a = (Fr *)data_;
    Fr hc[6];
    ntt::step_domain_consts_host<Fr>(lg_big, small, hc);
    const Fr g = emu_generator<Fr>(g_);
    const EmuDomain<Fr> d2(lg_big + 1, g), dbig(lg_big, g);
    const Fr *T2 = d2.tw.data(); // omega^e, e < big
    if (op == 4) {
        std::vector<Fr> zinv(compr);
        for (uint64_t k = 0; k < compr; ++k) {
            ntt::step_zinv_body<Fr>(k, hc + 1, zinv.data());
        }
        for (uint64_t i = 0; i < m; ++i) {
            ntt::step_divz_body<Fr>(i, big, compr, zinv.data(), hc + 4, a);
        }
        return 0;
    }
    std::vector<Fr> dd(big);
    const auto fold_rows = [&]() {
        for (uint64_t rows = compr; rows > 1; rows >>= 1) {
            const uint64_t half = (rows >> 1) * small;
            for (uint64_t t = 0; t < half; ++t) {
                ntt::fold_halve_body<Fr>(t, half, dd.data());
            }
        }
    };
    const auto small_transform = [&](Fr *v, int inverse) {
        if (lg_small >= 1) {
            const EmuDomain<Fr> ds(lg_small, g);
            ds.run(v, inverse, nullptr, nullptr, inverse ? &ds.consts[1] : nullptr);
        }
    };
    const auto fft = [&]() {
        for (uint64_t i = 0; i < big; ++i) {
            ntt::step_fft_pre_body<Fr>(i, big, small, T2, a, dd.data());
        }
        fold_rows();
        dbig.run(a, 0, nullptr, nullptr, nullptr);
        if (lg_small >= 1) {
            small_transform(dd.data(), 0);
        }
        memcpy(a + big, dd.data(), small * sizeof(Fr));
    };
    const auto ifft = [&]() {
        dbig.run(a, 1, nullptr, nullptr, &dbig.consts[1]);
        small_transform(a + big, 1);
        for (uint64_t i = 0; i < big; ++i) {
            ntt::step_ifft_tmp_body<Fr>(i, small, T2, a, dd.data());
        }
        fold_rows();
        for (uint64_t i = 0; i < small; ++i) {
            ntt::step_ifft_post_body<Fr>(i, big, T2, dd.data(), hc, a);
        }
    };
    std::vector<Fr> table(m);
    switch (op) {
    case 0: fft(); break;
    case 1: ifft(); break;
    case 2: // g^i, i < m, from the size-2big domain's powers of g (step_coset_tables)
        for (uint64_t e = 0; e < m; ++e) {
            ntt::pow_table_body<Fr>(e, d2.pg.data(), lg_big + 1, nullptr, table.data());
        }
        for (uint64_t i = 0; i < m; ++i) {
            ntt::mul_table_body<Fr>(i, a, table.data());
        }
        fft();
        break;
    case 3:
        ifft();
        for (uint64_t e = 0; e < m; ++e) {
            ntt::pow_table_body<Fr>(e, d2.pgi.data(), lg_big + 1, nullptr, table.data());
        }
        for (uint64_t i = 0; i < m; ++i) {
            ntt::mul_table_body<Fr>(i, a, table.data());
        }
        break;
    }
    return 0;
}

// The sparse R1CS evaluation (r1cs_eval_launch of ntt.cu): the row body for every output element, then the long rows
// by the launcher's own segment plan — per segment a sum of its terms (k_r1cs_eval_long), per row a sum of its partials
// (k_r1cs_long_finish); the CTA trees of those kernels are sums, and field addition is exact.
template<class Fr>
static int emu_r1cs_t(size_t nc, size_t ni, size_t m, const uint32_t *row_ptr, const uint32_t *col, const void *val_, const void *w_,
                      void *aA_, void *aB_, void *aC_)
{
    if (m < nc + ni + 1) {
        return -2;
    }
    const Fr *val = (const Fr *)val_, *w = (const Fr *)w_;
    Fr *aA = (Fr *)aA_, *aB = (Fr *)aB_, *aC = (Fr *)aC_;
    for (uint64_t t = 0; t < 3 * m; ++t) {
        ntt::r1cs_eval_body<Fr>(t, nc, ni, m, row_ptr, col, val, w, aA, aB, aC);
    }
    std::vector<uint32_t> long_rows, seg_ptr;
    std::vector<ntt::LongSeg> segs;
    ntt::r1cs_long_plan(row_ptr, 3 * nc, long_rows, seg_ptr, segs);
    std::vector<Fr> partials(segs.size() + 1);
    for (size_t k = 0; k < segs.size(); ++k) {
        Fr acc = Fr::zero();
        for (uint32_t e = segs[k].e0; e < segs[k].e1; ++e) {
            acc = acc + val[e] * w[col[e]];
        }
        partials[k] = acc;
    }
    for (size_t r = 0; r < long_rows.size(); ++r) {
        Fr acc = Fr::zero();
        for (uint32_t k = seg_ptr[r]; k < seg_ptr[r + 1]; ++k) {
            acc = acc + partials[k];
        }
        const uint64_t row = long_rows[r], mat = row / nc, i = row - mat * nc;
        (mat == 0 ? aA : (mat == 1 ? aB : aC))[i] = acc;
    }
    return (int)long_rows.size();
}
extern "C" int emu_r1cs_evaluate(int curve, size_t nc, size_t ni, size_t m, const uint32_t *row_ptr, const uint32_t *col,
                                 const void *val, const void *w, void *aA, void *aB, void *aC)
{
    switch (curve) {
    case 0: return emu_r1cs_t<alt_bn128_fr>(nc, ni, m, row_ptr, col, val, w, aA, aB, aC);
    case 1: return emu_r1cs_t<bls12_377_fr>(nc, ni, m, row_ptr, col, val, w, aA, aB, aC);
    case 2: return emu_r1cs_t<bw6_761_fr>(nc, ni, m, row_ptr, col, val, w, aA, aB, aC);
    }
    return -1;
}

extern "C" int emu_step_ntt(int curve, int op, void *data, size_t m, const void *g)
{
    switch (curve) {
    case 0: return emu_step_ntt_t<alt_bn128_fr>(op, data, m, g);
    case 1: return emu_step_ntt_t<bls12_377_fr>(op, data, m, g);
    case 2: return emu_step_ntt_t<bw6_761_fr>(op, data, m, g);
    }
    return -1;
}
extern "C" int emu_ntt(int curve, int op, void *data, size_t m, const void *g)
{
    switch (curve) {
    case 0: return emu_ntt_t<alt_bn128_fr>(op, data, m, g);
    case 1: return emu_ntt_t<bls12_377_fr>(op, data, m, g);
    case 2: return emu_ntt_t<bw6_761_fr>(op, data, m, g);
    }
    return -1;
}
extern "C" int emu_ext_ntt(int curve, int op, void *data, size_t m, const void *g)
{
    switch (curve) {
    case 0: return emu_ext_ntt_t<alt_bn128_fr>(op, data, m, g);
    case 1: return emu_ext_ntt_t<bls12_377_fr>(op, data, m, g);
    case 2: return emu_ext_ntt_t<bw6_761_fr>(op, data, m, g);
    }
    return -1;
}
