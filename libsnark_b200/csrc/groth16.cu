// groth16.cu — the orchestration of libsnark::r1cs_gg_ppzksnark_prover
// (libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.tcc:460-603) over
// device-resident data, behind the C-ABI: one call per proof.
//
//   w = (1 | primary | auxiliary)                       one H2D copy (pinned staging)
//   H = r1cs_to_qap_witness_map(cs, w, 0, 0, 0)         b200_witness_map_r1cs_dev, stays on the device
//   At = <A_query, w>   Bt = <B_query, w> (G2 and G1)   Ht = <H_query, H>   Lt = <L_query, w[ni+1..]>
//   g_A = alpha + At + r delta      g_B = beta + Bt + s delta (G1 and G2)
//   g_C = Ht + Lt + s g_A + r g_B(G1) - (r s) delta                          (.tcc:583-595)
//
// The five MSMs and the witness map run on the GPU; the closing handful of group operations and
// four scalar multiplications of single points run on the host (as in the reference, and for the
// same reason as host_fold.hpp: a serial chain on one point).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <future>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "curves.cuh"
#include "host_fold.hpp"
#include "msm_geometry.hpp"
#include "msm.cuh"
#include "runtime.hpp"

using namespace b200;

namespace
{

struct ProvingKey {
    int curve = -1;
    std::vector<uint8_t> alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2; // reference layout, affine
    b200_bases_t A = 0, B_g2 = 0, B_g1 = 0, H = 0, L = 0;
    b200_r1cs_t r1cs = 0;
    size_t nc = 0, ni = 0, nv = 0, m = 0, nB = 0;
    void *d_B_idx = nullptr; // device copy of B_query.indices when the query is sparse
    void *d_w = nullptr, *d_h = nullptr, *d_g = nullptr;
    void *h_w = nullptr; // pinned staging for (1 | primary | auxiliary)
    // sharded keys (b200_groth16_pk_create_shard): this rank's slice [first, first + count) of each
    // query, in the order A, B, H, L; tot = the whole query's size
    int rank = 0, world = 1;
    size_t tot[4] = {0, 0, 0, 0}, first[4] = {0, 0, 0, 0}, cnt[4] = {0, 0, 0, 0};
};

// The five MSMs of the prover in blob order, and the query (A, B, H, L) each one reads.
constexpr int NUM_MSM = 5;
constexpr int MSM_QUERY[NUM_MSM] = {0, 1, 1, 2, 3};
constexpr int MSM_GROUP[NUM_MSM] = {B200_G1, B200_G2, B200_G1, B200_G1, B200_G1};

void shard_range(size_t total, int world, int rank, size_t *first, size_t *count)
{
    const size_t per = (total + (size_t)world - 1) / (size_t)world;
    const size_t f = std::min(total, per * (size_t)rank);
    *first = f;
    *count = std::min(per, total - f);
}

// Where each MSM's window sums sit inside one rank's blob. Every rank's blob has the same size (the
// ranks all-gather them): a slot holds the largest window count any rank's shard of that query has.
struct BlobLayout {
    size_t off[NUM_MSM], total;
};
BlobLayout blob_layout(int curve, const size_t tot[4], int world)
{
    BlobLayout L;
    size_t at = 0;
    for (int k = 0; k < NUM_MSM; ++k) {
        int wmax = 0;
        for (int g = 0; g < world; ++g) {
            size_t f, c;
            shard_range(tot[MSM_QUERY[k]], world, g, &f, &c);
            if (c) {
                wmax = std::max(wmax, b200_msm_num_windows(curve, c));
            }
        }
        L.off[k] = at;
        at += (size_t)wmax * b200_msm_window_bytes(curve, MSM_GROUP[k]);
    }
    L.total = at ? at : 16;
    return L;
}

std::mutex g_mu;
std::map<uint64_t, ProvingKey> g_keys;
uint64_t g_next = 1;

double now_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

template<class F> XYZZ<F> load_ref(const void *p)
{
    msm::RefPoint<F> r;
    memcpy(&r, p, sizeof(r));
    if (r.z.is_zero()) {
        return XYZZ<F>::inf();
    }
    return XYZZ<F>::from_affine(Affine<F>{r.x, r.y});
}

// k * P, k a Montgomery-form scalar: left-to-right double-and-add on the host.
template<class Fr, class F> XYZZ<F> scalar_mul(const Fr &k_mont, const XYZZ<F> &p)
{
    const Fr k = k_mont.from_mont();
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int bit = Fr::params::BITS - 1; bit >= 0; --bit) {
        if (!acc.is_inf()) {
            dbl(acc);
        }
        if ((k.v[bit >> 5] >> (bit & 31)) & 1u) {
            add(acc, p);
        }
    }
    return acc;
}

struct KeyPoints {
    const void *alpha_g1, *beta_g1, *beta_g2, *delta_g1, *delta_g2;
};

// The scalar multiplications of the proof assembly that need only r, s and the key (.tcc:583-595): r delta,
// s delta (G1 and G2), (r s) delta. Serial double-and-add chains on one point each, side by side on host
// threads; the one-GPU prover starts them while the GPU is busy with the MSMs.
template<int CURVE> struct PreAssembled {
    XYZZ<typename group_traits<CURVE, GROUP_G1>::F> rd1, sd1, rsd1;
    XYZZ<typename group_traits<CURVE, GROUP_G2>::F> sd2;
};
template<int CURVE> PreAssembled<CURVE> pre_assemble(const KeyPoints &pk, const void *r_, const void *s_)
{
    typedef typename group_traits<CURVE, GROUP_G1>::Fr Fr;
    typedef typename group_traits<CURVE, GROUP_G1>::F F1;
    typedef typename group_traits<CURVE, GROUP_G2>::F F2;
    Fr r, s;
    memcpy(&r, r_, sizeof(Fr));
    memcpy(&s, s_, sizeof(Fr));
    const XYZZ<F1> delta1 = load_ref<F1>(pk.delta_g1);
    const XYZZ<F2> delta2 = load_ref<F2>(pk.delta_g2);
    auto f_sd2 = std::async(std::launch::async, [&]() { return scalar_mul<Fr, F2>(s, delta2); });
    auto f_rd1 = std::async(std::launch::async, [&]() { return scalar_mul<Fr, F1>(r, delta1); });
    auto f_rsd1 = std::async(std::launch::async, [&]() { return scalar_mul<Fr, F1>(r * s, delta1); });
    PreAssembled<CURVE> p;
    p.sd1 = scalar_mul<Fr, F1>(s, delta1);
    p.sd2 = f_sd2.get();
    p.rd1 = f_rd1.get();
    p.rsd1 = f_rsd1.get();
    return p;
}

template<int CURVE>
int assemble(const KeyPoints &pk, const void *r_, const void *s_, const void *At, const void *Bg2, const void *Bg1,
             const void *Ht, const void *Lt, void *proof_out, const PreAssembled<CURVE> *pre_given = nullptr)
{
    typedef typename group_traits<CURVE, GROUP_G1>::Fr Fr;
    typedef typename group_traits<CURVE, GROUP_G1>::F F1;
    typedef typename group_traits<CURVE, GROUP_G2>::F F2;
    Fr r, s;
    memcpy(&r, r_, sizeof(Fr));
    memcpy(&s, s_, sizeof(Fr));
    const PreAssembled<CURVE> pre = pre_given ? *pre_given : pre_assemble<CURVE>(pk, r_, s_);

    XYZZ<F1> gA = load_ref<F1>(pk.alpha_g1);
    add(gA, load_ref<F1>(At));
    add(gA, pre.rd1);

    XYZZ<F1> gB1 = load_ref<F1>(pk.beta_g1);
    add(gB1, load_ref<F1>(Bg1));
    add(gB1, pre.sd1);
    auto f_rgB1 = std::async(std::launch::async, [&]() { return scalar_mul<Fr, F1>(r, gB1); });
    const XYZZ<F1> sgA = scalar_mul<Fr, F1>(s, gA);

    XYZZ<F2> gB2 = load_ref<F2>(pk.beta_g2);
    add(gB2, load_ref<F2>(Bg2));
    add(gB2, pre.sd2);

    XYZZ<F1> gC = load_ref<F1>(Ht);
    add(gC, load_ref<F1>(Lt));
    add(gC, sgA);
    add(gC, f_rgB1.get());
    add(gC, neg(pre.rsd1));

    const msm::RefPoint<F1> oA = msm::normalise(gA), oC = msm::normalise(gC);
    const msm::RefPoint<F2> oB = msm::normalise(gB2);
    uint8_t *o = static_cast<uint8_t *>(proof_out);
    memcpy(o, &oA, sizeof(oA));
    memcpy(o + sizeof(oA), &oB, sizeof(oB));
    memcpy(o + sizeof(oA) + sizeof(oB), &oC, sizeof(oC));
    return B200_OK;
}

// Sum over the shards of one query: every rank's W window sums are folded by Horner's rule
// (host_fold.hpp); consecutive shards with the same window geometry share one fold.
template<int CURVE, int GROUP>
void fold_shards(size_t total, int world, const uint8_t *blobs, size_t blob_stride, size_t slot_off, void *out)
{
    typedef typename group_traits<CURVE, GROUP>::F F;
    typedef typename group_traits<CURVE, GROUP>::Fr Fr;
    XYZZ<F> sum = XYZZ<F>::inf();
    std::vector<XYZZ<F>> run;
    int run_W = 0, run_c = 0;
    size_t run_shards = 0;
    auto flush = [&]() {
        if (run_shards) {
            const msm::RefPoint<F> r = msm::host_fold<F>(run.data(), run_W, run_c, run_shards);
            add(sum, load_ref<F>(&r));
            run.clear();
            run_shards = 0;
        }
    };
    for (int g = 0; g < world; ++g) {
        size_t first, cnt;
        shard_range(total, world, g, &first, &cnt);
        if (!cnt) {
            continue;
        }
        const msm::Geometry geo = msm::make_geometry(cnt, Fr::params::BITS);
        if (run_shards && (geo.W != run_W || geo.c != run_c)) {
            flush();
        }
        run_W = geo.W;
        run_c = geo.c;
        const size_t at = run.size();
        run.resize(at + (size_t)geo.W);
        memcpy(run.data() + at, blobs + (size_t)g * blob_stride + slot_off, (size_t)geo.W * sizeof(XYZZ<F>));
        ++run_shards;
    }
    flush();
    const msm::RefPoint<F> r = msm::normalise(sum);
    memcpy(out, &r, sizeof(r));
}

// b200_groth16_assemble_prepare: the r, s-only scalar multiplications started ahead of the MSMs, keyed by the bytes of
// (curve, delta_g1, delta_g2, r, s); b200_groth16_assemble picks a matching one up instead of computing them at the end.
std::mutex g_pre_mu;
std::map<std::string, std::shared_future<std::vector<uint8_t>>> g_pre;
template<int CURVE> std::string pre_key(const KeyPoints &kp, const void *r, const void *s)
{
    typedef typename group_traits<CURVE, GROUP_G1>::Fr Fr;
    typedef typename group_traits<CURVE, GROUP_G1>::F F1;
    typedef typename group_traits<CURVE, GROUP_G2>::F F2;
    std::string k(1, (char)CURVE);
    k.append(static_cast<const char *>(kp.delta_g1), 3 * sizeof(F1));
    k.append(static_cast<const char *>(kp.delta_g2), 3 * sizeof(F2));
    k.append(static_cast<const char *>(r), sizeof(Fr));
    k.append(static_cast<const char *>(s), sizeof(Fr));
    return k;
}
template<int CURVE> int assemble_prepare(const void *delta_g1, const void *delta_g2, const void *r, const void *s)
{
    typedef typename group_traits<CURVE, GROUP_G1>::Fr Fr;
    typedef typename group_traits<CURVE, GROUP_G1>::F F1;
    typedef typename group_traits<CURVE, GROUP_G2>::F F2;
    KeyPoints kp = {nullptr, nullptr, nullptr, delta_g1, delta_g2};
    const std::string key = pre_key<CURVE>(kp, r, s);
    // the inputs are copied: the caller's buffers need not outlive this call
    std::vector<uint8_t> d1((const uint8_t *)delta_g1, (const uint8_t *)delta_g1 + 3 * sizeof(F1));
    std::vector<uint8_t> d2((const uint8_t *)delta_g2, (const uint8_t *)delta_g2 + 3 * sizeof(F2));
    std::vector<uint8_t> rs((const uint8_t *)r, (const uint8_t *)r + sizeof(Fr));
    rs.insert(rs.end(), (const uint8_t *)s, (const uint8_t *)s + sizeof(Fr));
    std::lock_guard<std::mutex> lk(g_pre_mu);
    if (g_pre.size() > 16) {
        g_pre.clear(); // abandoned preparations
    }
    g_pre[key] = std::async(std::launch::async, [d1, d2, rs]() {
                     const KeyPoints k2 = {nullptr, nullptr, nullptr, d1.data(), d2.data()};
                     const PreAssembled<CURVE> p = pre_assemble<CURVE>(k2, rs.data(), rs.data() + sizeof(Fr));
                     std::vector<uint8_t> out(sizeof(p));
                     memcpy(out.data(), &p, sizeof(p));
                     return out;
                 }).share();
    return B200_OK;
}
template<int CURVE> bool take_prepared(const KeyPoints &kp, const void *r, const void *s, PreAssembled<CURVE> *out)
{
    std::shared_future<std::vector<uint8_t>> f;
    {
        std::lock_guard<std::mutex> lk(g_pre_mu);
        auto it = g_pre.find(pre_key<CURVE>(kp, r, s));
        if (it == g_pre.end()) {
            return false;
        }
        f = it->second;
        g_pre.erase(it);
    }
    const std::vector<uint8_t> bytes = f.get();
    memcpy(out, bytes.data(), sizeof(*out));
    return true;
}

template<int CURVE>
int assemble_shards(const KeyPoints &kp, const size_t tot[4], int world, const void *blobs_, const void *r, const void *s,
                    void *proof_out)
{
    typedef typename group_traits<CURVE, GROUP_G1>::F F1;
    typedef typename group_traits<CURVE, GROUP_G2>::F F2;
    const BlobLayout L = blob_layout(CURVE, tot, world);
    const uint8_t *blobs = static_cast<const uint8_t *>(blobs_);
    msm::RefPoint<F1> At, Bg1, Ht, Lt;
    msm::RefPoint<F2> Bg2;
    // the five folds and the r, s-only scalar multiplications are independent: side by side on host threads
    auto f_pre = std::async(std::launch::async, [&]() {
        PreAssembled<CURVE> p;
        if (!take_prepared<CURVE>(kp, r, s, &p)) {
            p = pre_assemble<CURVE>(kp, r, s);
        }
        return p;
    });
    auto f_g2 = std::async(std::launch::async, [&]() { fold_shards<CURVE, GROUP_G2>(tot[1], world, blobs, L.total, L.off[1], &Bg2); });
    auto f_a = std::async(std::launch::async, [&]() { fold_shards<CURVE, GROUP_G1>(tot[0], world, blobs, L.total, L.off[0], &At); });
    auto f_b1 = std::async(std::launch::async, [&]() { fold_shards<CURVE, GROUP_G1>(tot[1], world, blobs, L.total, L.off[2], &Bg1); });
    auto f_h = std::async(std::launch::async, [&]() { fold_shards<CURVE, GROUP_G1>(tot[2], world, blobs, L.total, L.off[3], &Ht); });
    fold_shards<CURVE, GROUP_G1>(tot[3], world, blobs, L.total, L.off[4], &Lt);
    f_g2.get();
    f_a.get();
    f_b1.get();
    f_h.get();
    const PreAssembled<CURVE> pre = f_pre.get();
    return assemble<CURVE>(kp, r, s, &At, &Bg2, &Bg1, &Ht, &Lt, proof_out, &pre);
}

// ---- concurrent MSMs ----------------------------------------------------------------------------------
// The prover's five MSMs are independent of each other (only H waits for the witness map). Each runs its
// sort -> accumulate -> reduce pipeline in an execution slot of its own (runtime.hpp), driven by a host thread
// of its own, so that the GPU always has bucket-accumulation work of one MSM in flight while another is in a
// latency-bound stage. Slot plan: 0 (this thread) = witness map, then H; 1 = B.g2; 2 = A; 3 = L; 4 = B.g1
// (profiles/r2_prove_slot_plans.jsonl: 21.7 -> 20.8 ms against A and B.g1 sharing slot 2).
struct MsmJob {
    int k;             // blob order: 0 A, 1 B.g2, 2 B.g1, 3 H, 4 L
    b200_bases_t set;
    size_t n;
    const void *scalars; // device
    void *out;           // host point (windows_dev == false) or device window array
    bool windows_dev;
    const rt::SharedSort *shared = nullptr; // the sort of the witness shared by A, B.g2, B.g1 and L (share_witness_sort)
    uint32_t skip = 0;
};
struct JobResult {
    int rc = B200_OK;
    std::string err;
};
std::map<uint64_t, std::future<JobResult>> g_pending; // b200_groth16_shard_msm_begin .. _end, by key handle (under g_mu)
constexpr int MSM_SLOT[NUM_MSM] = {2, 1, 4, 0, 3}; // H must stay in slot 0: it runs on the thread that computed its scalars

// Tuning knobs (environment, read per call): B200_MSM_SLOTS = five slot numbers "A,Bg2,Bg1,H,L" overriding the plan
// above; B200_WM_FIRST = 0 starts the other slots before the witness map instead of after it.
void slot_plan(int plan[NUM_MSM], bool *wm_first)
{
    for (int k = 0; k < NUM_MSM; ++k) {
        plan[k] = MSM_SLOT[k];
    }
    if (const char *e = getenv("B200_MSM_SLOTS")) {
        int v[NUM_MSM];
        if (sscanf(e, "%d,%d,%d,%d,%d", &v[0], &v[1], &v[2], &v[3], &v[4]) == NUM_MSM) {
            for (int k = 0; k < NUM_MSM; ++k) {
                plan[k] = v[k] < 0 || v[k] >= rt::NUM_SLOTS ? 0 : v[k];
            }
            plan[3] = 0; // H waits for the scalars this thread computes: it cannot move to another slot's thread
        }
    }
    const char *w = getenv("B200_WM_FIRST");
    *wm_first = !(w && w[0] == '0');
}

int run_job(const MsmJob &j, double *ms5)
{
    const double t0 = now_ms();
    const int rc = j.shared        ? rt::msm_resident_shared(j.set, j.n, j.scalars, j.windows_dev ? nullptr : j.out,
                                                              j.windows_dev ? j.out : nullptr, j.shared, j.skip)
                   : j.windows_dev ? b200_msm_resident_windows_dev(j.set, 0, j.n, j.scalars, 1, j.out)
                                   : b200_msm_resident(j.set, 0, j.n, j.scalars, 1, j.out);
    if (ms5) {
        ms5[j.k] = now_ms() - t0;
    }
    return rc;
}

JobResult run_slot_jobs(int slot, std::vector<MsmJob> jobs, double *ms5)
{
    JobResult r;
    r.rc = rt::bind_thread(slot);
    for (size_t i = 0; i < jobs.size() && r.rc == B200_OK; ++i) {
        r.rc = run_job(jobs[i], ms5);
    }
    if (r.rc != B200_OK) {
        r.err = rt::last_error(); // thread-local: carry the text back to the calling thread
    }
    return r;
}

// A, B.g2, B.g1 and L all take the witness as their scalars (r1cs_gg_ppzksnark.tcc:539-581; L from variable
// num_inputs + 1 on): when the whole key is on this GPU, B is dense and the four sets have window tables of one width,
// their digits are sorted ONCE — on slot 1's stream, beside the witness map — and the four MSMs read that sort
// (rt::SharedSort). Returns with the jobs marked, or unchanged when the conditions do not hold. B200_SHARED_SORT=0 turns
// it off.
rt::SharedSort g_shared_sort; // one proof at a time uses it (under g_mu)
int share_witness_sort(const ProvingKey &pk, std::vector<MsmJob> &jobs)
{
    static const bool off = getenv("B200_SHARED_SORT") && getenv("B200_SHARED_SORT")[0] == '0';
    if (off || pk.world != 1 || pk.d_B_idx || jobs.size() != NUM_MSM) {
        return B200_OK;
    }
    const b200_bases_t sets[4] = {pk.A, pk.B_g2, pk.B_g1, pk.L};
    const size_t n[4] = {jobs[0].n, jobs[1].n, jobs[2].n, jobs[4].n};
    const uint32_t skip = (uint32_t)(pk.ni + 1);
    if (!rt::msm_shared_eligible(sets, n, skip)) {
        return B200_OK;
    }
    if (rt::msm_shared_sort(pk.A, n[0], pk.d_w, &g_shared_sort, rt::slot_stream(1)) != B200_OK) {
        return B200_OK; // e.g. too small for affine rounds: every MSM sorts for itself
    }
    for (int k : {0, 1, 2, 4}) {
        jobs[k].shared = &g_shared_sort;
        jobs[k].scalars = pk.d_w;
    }
    jobs[4].skip = skip;
    return B200_OK;
}

// Runs `jobs` (any subset of the five MSMs) over the slots; `before_h`, if set, runs on this thread ahead of the H
// job (the witness map). Returns the first failure.
template<class BeforeH> int run_msms(const std::vector<MsmJob> &jobs, double *ms5, BeforeH before_h, bool before_h_is_short = false)
{
    int plan[NUM_MSM];
    bool wm_first = true;
    slot_plan(plan, &wm_first);
    if (before_h_is_short) {
        wm_first = false; // a handful of launches (pointwise + one inverse transform): let the other slots start at once
    }
    std::vector<MsmJob> per_slot[rt::NUM_SLOTS];
    for (const MsmJob &j : jobs) {
        per_slot[plan[j.k]].push_back(j);
    }
    std::future<JobResult> fut[rt::NUM_SLOTS];
    // The witness map is integer-pipe work like the accumulations, so overlapping it buys nothing — and its ~25 short
    // launches each queue behind resident accumulate CTAs (measured: 13 ms instead of 2 ms). It runs alone, first.
    int rc = wm_first ? before_h() : B200_OK;
    for (int sl = 1; sl < rt::NUM_SLOTS && rc == B200_OK; ++sl) {
        if (!per_slot[sl].empty()) {
            fut[sl] = std::async(std::launch::async, run_slot_jobs, sl, per_slot[sl], ms5);
        }
    }
    if (!wm_first) {
        rc = before_h();
    }
    for (size_t i = 0; i < per_slot[0].size() && rc == B200_OK; ++i) {
        rc = run_job(per_slot[0][i], ms5);
    }
    std::string err = rc != B200_OK ? rt::last_error() : "";
    for (int sl = 1; sl < rt::NUM_SLOTS; ++sl) {
        if (fut[sl].valid()) {
            const JobResult r = fut[sl].get();
            if (rc == B200_OK && r.rc != B200_OK) {
                rc = r.rc;
                err = r.err;
            }
        }
    }
    return rc == B200_OK ? B200_OK : rt::set_error(rc, "%s", err.c_str());
}

ProvingKey *find_key(b200_groth16_pk_t handle)
{
    auto it = g_keys.find(handle);
    if (it == g_keys.end()) {
        rt::set_error(B200_ERR_HANDLE, "unknown proving-key handle %llu", (unsigned long long)handle);
        return nullptr;
    }
    return &it->second;
}

// (1 | primary | auxiliary) into the pinned staging buffer and on to the device, r1cs_gg_ppzksnark.tcc:525-529
int upload_witness(const ProvingKey &pk, const void *primary_input, const void *auxiliary_input)
{
    const size_t fr = b200_sizeof_fr(pk.curve);
    uint8_t *w = static_cast<uint8_t *>(pk.h_w);
    uint32_t one[32];
    for (int i = 0; i < (int)(fr / 4); ++i) {
        one[i] = pk.curve == B200_ALT_BN128   ? alt_bn128_fr::params::one(i)
                 : pk.curve == B200_BLS12_377 ? bls12_377_fr::params::one(i)
                                              : bw6_761_fr::params::one(i);
    }
    memcpy(w, one, fr);
    memcpy(w + fr, primary_input, pk.ni * fr);
    {
        // A caller that keeps its witness in page-locked memory (cudaHostAlloc / cudaHostRegister, e.g. a pinned torch
        // tensor) needs no staging: the bulk goes to the device straight from where it lies.
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, auxiliary_input) == cudaSuccess && at.type == cudaMemoryTypeHost) {
            cudaStream_t st0 = rt::stream();
            const size_t head0 = (1 + pk.ni) * fr;
            B200_CUDA_CHECK(cudaMemcpyAsync(pk.d_w, w, head0, cudaMemcpyHostToDevice, st0));
            B200_CUDA_CHECK(cudaMemcpyAsync(static_cast<uint8_t *>(pk.d_w) + head0, auxiliary_input, (pk.nv - pk.ni) * fr,
                                            cudaMemcpyHostToDevice, st0));
            B200_CUDA_CHECK(cudaStreamSynchronize(st0));
            return B200_OK;
        }
        cudaGetLastError(); // an unregistered pointer may leave an error behind on old runtimes
    }
    // The auxiliary input is the bulk (32 MiB at 2^20, 128 MiB at 2^22) and one core copies at ~10 GB/s: stage it into
    // the pinned buffer with a few threads, and send each slice to the device as soon as it is staged.
    const size_t head = (1 + pk.ni) * fr, bulk = (pk.nv - pk.ni) * fr;
    const size_t T = std::max<size_t>(1, std::min<size_t>(4, bulk >> 22));
    cudaStream_t st = rt::stream();
    const uint8_t *src = static_cast<const uint8_t *>(auxiliary_input);
    uint8_t *dst = w + head, *dev = static_cast<uint8_t *>(pk.d_w);
    std::vector<std::thread> workers;
    std::vector<size_t> cut(T + 1);
    for (size_t t = 0; t <= T; ++t) {
        cut[t] = t == T ? bulk : (bulk / T * t) & ~(size_t)63;
    }
    for (size_t t = 1; t < T; ++t) {
        workers.emplace_back([=]() { memcpy(dst + cut[t], src + cut[t], cut[t + 1] - cut[t]); });
    }
    memcpy(dst, src, cut[1]);
    cudaError_t e = cudaMemcpyAsync(dev, w, head + cut[1], cudaMemcpyHostToDevice, st);
    for (size_t t = 1; t < T; ++t) {
        workers[t - 1].join();
        if (e == cudaSuccess) {
            e = cudaMemcpyAsync(dev + head + cut[t], dst + cut[t], cut[t + 1] - cut[t], cudaMemcpyHostToDevice, st);
        }
    }
    if (e == cudaSuccess) {
        e = cudaStreamSynchronize(st);
    }
    if (e != cudaSuccess) {
        return rt::set_error(B200_ERR_CUDA, "witness upload failed: %s", cudaGetErrorString(e));
    }
    return B200_OK;
}

// The MSMs of this rank's shards selected by `mask` (bit k = MSM k in blob order A, B.g2, B.g1, H, L), run
// concurrently (run_msms); with_witness_map computes coefficients_for_H on this thread first.
int shard_msms(const ProvingKey &pk, unsigned mask, void *dev_blob, double *ms5, bool with_witness_map, double *wm_ms,
               void *dA = nullptr, const void *dB = nullptr, const void *dC = nullptr)
{
    const size_t fr = b200_sizeof_fr(pk.curve);
    const BlobLayout L = blob_layout(pk.curve, pk.tot, pk.world);
    const b200_bases_t sets[NUM_MSM] = {pk.A, pk.B_g2, pk.B_g1, pk.H, pk.L};
    const uint8_t *w = static_cast<const uint8_t *>(pk.d_w);
    const void *scB = w + pk.first[1] * fr;
    if (pk.d_B_idx && (mask & 6u) && pk.cnt[1]) {
        if (int rc = b200_gather_dev(pk.curve, pk.d_w, static_cast<const uint32_t *>(pk.d_B_idx), pk.cnt[1], pk.d_g)) {
            return rc;
        }
        scB = pk.d_g;
    }
    const void *sc[NUM_MSM] = {w + pk.first[0] * fr, scB, scB, static_cast<const uint8_t *>(pk.d_h) + pk.first[2] * fr,
                               w + (pk.ni + 1 + pk.first[3]) * fr};
    std::vector<MsmJob> jobs;
    for (int k = 0; k < NUM_MSM; ++k) {
        const size_t cnt = pk.cnt[MSM_QUERY[k]];
        if ((mask >> k & 1u) && cnt) {
            jobs.push_back(MsmJob{k, sets[k], cnt, sc[k], static_cast<uint8_t *>(dev_blob) + L.off[k], true});
        }
    }
    return run_msms(
        jobs, ms5,
        [&]() -> int {
            if (dA) { // the three chain results are here: pointwise step + icosetFFT -> coefficients_for_H
                return b200_witness_finish_dev(pk.curve, dA, dB, dC, pk.m, pk.d_h);
            }
            if (!with_witness_map) {
                return B200_OK;
            }
            const double t0 = now_ms();
            const int rc = b200_witness_map_r1cs_dev(pk.r1cs, pk.d_w, pk.m, pk.d_h);
            if (wm_ms) {
                *wm_ms = now_ms() - t0;
            }
            return rc;
        },
        dA != nullptr);
}

// r1cs_gg_ppzksnark_prover over a one-GPU key (b200_groth16_prove).
template<int CURVE>
int prove_impl(const ProvingKey &pk, const void *primary_input, const void *auxiliary_input, const void *r, const void *s,
               void *proof_out, double *stage_ms, bool witness_resident = false)
{
    const size_t fr = b200_sizeof_fr(pk.curve);
    const size_t g1 = b200_sizeof_group(pk.curve, B200_G1), g2 = b200_sizeof_group(pk.curve, B200_G2);
    const double t0 = now_ms();
    double st[7] = {0, 0, 0, 0, 0, 0, 0};
    const KeyPoints kp = {pk.alpha_g1.data(), pk.beta_g1.data(), pk.beta_g2.data(), pk.delta_g1.data(), pk.delta_g2.data()};
    // host work that needs only r, s and the key goes first, beside everything else
    auto f_pre = std::async(std::launch::async, [&]() { return pre_assemble<CURVE>(kp, r, s); });

    int rc;
    if (!witness_resident && (rc = upload_witness(pk, primary_input, auxiliary_input))) {
        return rc;
    }
    st[0] = now_ms() - t0;
    std::vector<uint8_t> At(g1), Bg1(g1), Ht(g1), Lt(g1), Bg2(g2);
    const void *scB = pk.d_w;
    if (pk.d_B_idx) {
        if ((rc = b200_gather_dev(pk.curve, pk.d_w, static_cast<const uint32_t *>(pk.d_B_idx), pk.nB, pk.d_g))) {
            return rc;
        }
        scB = pk.d_g;
    }
    // the five MSMs side by side (run_msms); this thread computes H's scalars first
    double msm_ms[NUM_MSM] = {0, 0, 0, 0, 0};
    std::vector<MsmJob> jobs = {
        {0, pk.A, pk.nv + 1, pk.d_w, At.data(), false},
        {1, pk.B_g2, pk.nB, scB, Bg2.data(), false},
        {2, pk.B_g1, pk.nB, scB, Bg1.data(), false},
        {3, pk.H, pk.m - 1, pk.d_h, Ht.data(), false},
        {4, pk.L, pk.nv - pk.ni, static_cast<const uint8_t *>(pk.d_w) + (pk.ni + 1) * fr, Lt.data(), false},
    };
    if ((rc = share_witness_sort(pk, jobs))) {
        return rc;
    }
    rc = run_msms(jobs, msm_ms, [&]() -> int {
        const double w0 = now_ms();
        const int wrc = b200_witness_map_r1cs_dev(pk.r1cs, pk.d_w, pk.m, pk.d_h);
        st[1] = now_ms() - w0;
        return wrc;
    });
    if (rc) {
        return rc;
    }
    // wall clock of each MSM inside its own slot; they overlap, so they do not add up to the total
    st[2] = msm_ms[0];
    st[3] = msm_ms[1] + msm_ms[2];
    st[4] = msm_ms[3];
    st[5] = msm_ms[4];
    const PreAssembled<CURVE> pre = f_pre.get();
    rc = assemble<CURVE>(kp, r, s, At.data(), Bg2.data(), Bg1.data(), Ht.data(), Lt.data(), proof_out, &pre);
    st[6] = now_ms() - t0;
    if (stage_ms) {
        // witness upload | witness map | A | B | H | L | total (the last includes the host assembly)
        memcpy(stage_ms, st, sizeof(st));
    }
    return rc;
}

} // namespace

// Releases the device / pinned buffers a ProvingKey owns (the base sets and the R1CS belong to the caller).
static void release_key_buffers(ProvingKey &pk)
{
    if (pk.d_w) {
        b200_dev_free(pk.d_w);
    }
    if (pk.d_h) {
        b200_dev_free(pk.d_h);
    }
    if (pk.d_g) {
        b200_dev_free(pk.d_g);
    }
    if (pk.d_B_idx) {
        b200_dev_free(pk.d_B_idx);
    }
    if (pk.h_w) {
        cudaFreeHost(pk.h_w);
    }
    pk.d_w = pk.d_h = pk.d_g = pk.d_B_idx = pk.h_w = nullptr;
}

// The key's buffers, all or nothing: a failure part-way (usually out of memory) releases what was allocated.
static int alloc_key_buffers(ProvingKey &pk, size_t fr, const uint32_t *B_indices, size_t nB_idx)
{
    int rc = B200_OK;
    if ((rc = b200_dev_alloc((pk.nv + 1) * fr, &pk.d_w)) || (rc = b200_dev_alloc((pk.m + 1) * fr, &pk.d_h))) {
        release_key_buffers(pk);
        return rc;
    }
    if (B_indices && nB_idx) {
        if ((rc = b200_dev_alloc(nB_idx * 4, &pk.d_B_idx)) || (rc = b200_memcpy_h2d(pk.d_B_idx, B_indices, nB_idx * 4)) ||
            (rc = b200_dev_alloc(nB_idx * fr, &pk.d_g))) {
            release_key_buffers(pk);
            return rc;
        }
    }
    const cudaError_t e = cudaHostAlloc(&pk.h_w, (pk.nv + 1) * fr, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        pk.h_w = nullptr;
        release_key_buffers(pk);
        return rt::set_error(B200_ERR_CUDA, "cudaHostAlloc(%zu) failed: %s", (pk.nv + 1) * fr, cudaGetErrorString(e));
    }
    return B200_OK;
}

extern "C" {

int b200_groth16_pk_create(int curve, const void *alpha_g1, const void *beta_g1, const void *beta_g2,
                           const void *delta_g1, const void *delta_g2, b200_bases_t A_query, b200_bases_t B_query_g2,
                           b200_bases_t B_query_g1, const uint32_t *B_indices, b200_bases_t H_query,
                           b200_bases_t L_query, b200_r1cs_t r1cs, size_t num_constraints, size_t num_inputs,
                           size_t num_variables, size_t domain_size, b200_groth16_pk_t *handle)
{
    if (!alpha_g1 || !beta_g1 || !beta_g2 || !delta_g1 || !delta_g2 || !handle) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    const size_t g1 = b200_sizeof_group(curve, B200_G1), g2 = b200_sizeof_group(curve, B200_G2);
    if (!g1 || !g2) {
        return rt::set_error(B200_ERR_ARG, "unknown curve %d", curve);
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    const size_t nv = num_variables, ni = num_inputs, m = domain_size;
    const size_t nB = b200_bases_size(B_query_g2);
    if (b200_bases_size(A_query) != nv + 1 || b200_bases_size(B_query_g1) != nB || nB > nv + 1 ||
        b200_bases_size(H_query) + 1 < m || b200_bases_size(L_query) != nv - ni || ni > nv ||
        m < num_constraints + ni + 1) {
        return rt::set_error(B200_ERR_ARG,
                             "query sizes do not match the constraint system (r1cs_gg_ppzksnark.tcc:506-511)");
    }
    ProvingKey pk;
    pk.curve = curve;
    pk.alpha_g1.assign((const uint8_t *)alpha_g1, (const uint8_t *)alpha_g1 + g1);
    pk.beta_g1.assign((const uint8_t *)beta_g1, (const uint8_t *)beta_g1 + g1);
    pk.beta_g2.assign((const uint8_t *)beta_g2, (const uint8_t *)beta_g2 + g2);
    pk.delta_g1.assign((const uint8_t *)delta_g1, (const uint8_t *)delta_g1 + g1);
    pk.delta_g2.assign((const uint8_t *)delta_g2, (const uint8_t *)delta_g2 + g2);
    pk.A = A_query;
    pk.B_g2 = B_query_g2;
    pk.B_g1 = B_query_g1;
    pk.H = H_query;
    pk.L = L_query;
    pk.r1cs = r1cs;
    pk.nc = num_constraints;
    pk.ni = ni;
    pk.nv = nv;
    pk.m = m;
    pk.nB = nB;
    const size_t fr = b200_sizeof_fr(curve);
    bool dense = true;
    if (B_indices) {
        for (size_t k = 0; k < nB; ++k) {
            if (B_indices[k] > nv || (k && B_indices[k] <= B_indices[k - 1])) {
                return rt::set_error(B200_ERR_ARG, "B_indices must be increasing and <= num_variables");
            }
            dense = dense && B_indices[k] == k;
        }
    } else if (nB != nv + 1) {
        return rt::set_error(B200_ERR_ARG, "a B query shorter than num_variables + 1 needs its index list");
    }
    if (const int rc = alloc_key_buffers(pk, fr, dense ? nullptr : B_indices, nB)) {
        return rc;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    *handle = g_next++;
    g_keys[*handle] = pk;
    return B200_OK;
}

int b200_groth16_pk_free(b200_groth16_pk_t handle)
{
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_keys.find(handle);
    if (it == g_keys.end()) {
        return rt::set_error(B200_ERR_HANDLE, "unknown proving-key handle %llu", (unsigned long long)handle);
    }
    release_key_buffers(it->second);
    g_keys.erase(it);
    return B200_OK;
}

int b200_groth16_prove(b200_groth16_pk_t handle, const void *primary_input, const void *auxiliary_input, const void *r,
                       const void *s, void *proof_out, double *stage_ms)
{
    if (!auxiliary_input || !r || !s || !proof_out) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_keys.find(handle);
    if (it == g_keys.end()) {
        return rt::set_error(B200_ERR_HANDLE, "unknown proving-key handle %llu", (unsigned long long)handle);
    }
    const ProvingKey &pk = it->second;
    if (pk.ni && !primary_input) {
        return rt::set_error(B200_ERR_ARG, "null primary input");
    }
    switch (pk.curve) {
    case B200_ALT_BN128: return prove_impl<CURVE_ALT_BN128>(pk, primary_input, auxiliary_input, r, s, proof_out, stage_ms);
    case B200_BLS12_377: return prove_impl<CURVE_BLS12_377>(pk, primary_input, auxiliary_input, r, s, proof_out, stage_ms);
    default: return prove_impl<CURVE_BW6_761>(pk, primary_input, auxiliary_input, r, s, proof_out, stage_ms);
    }
}

int b200_groth16_witness_upload(b200_groth16_pk_t handle, const void *primary_input, const void *auxiliary_input)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    if (!auxiliary_input || (pk->ni && !primary_input)) {
        return rt::set_error(B200_ERR_ARG, "null input");
    }
    return upload_witness(*pk, primary_input, auxiliary_input);
}

int b200_groth16_prove_resident(b200_groth16_pk_t handle, const void *r, const void *s, void *proof_out, double *stage_ms)
{
    if (!r || !s || !proof_out) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    if (pk->world != 1) {
        return rt::set_error(B200_ERR_ARG, "b200_groth16_prove_resident needs a whole (unsharded) key");
    }
    switch (pk->curve) {
    case B200_ALT_BN128: return prove_impl<CURVE_ALT_BN128>(*pk, nullptr, nullptr, r, s, proof_out, stage_ms, true);
    case B200_BLS12_377: return prove_impl<CURVE_BLS12_377>(*pk, nullptr, nullptr, r, s, proof_out, stage_ms, true);
    default: return prove_impl<CURVE_BW6_761>(*pk, nullptr, nullptr, r, s, proof_out, stage_ms, true);
    }
}

/* ---- sharded prover: one process per GPU, every query cut by base range ---------- */

void b200_shard_range(size_t total, int world, int rank, size_t *first, size_t *count)
{
    size_t f = 0, c = 0;
    if (world > 0 && rank >= 0 && rank < world) {
        shard_range(total, world, rank, &f, &c);
    }
    if (first) {
        *first = f;
    }
    if (count) {
        *count = c;
    }
}

size_t b200_groth16_shard_bytes(int curve, const size_t query_sizes[4], int world)
{
    if (!query_sizes || world < 1 || !b200_sizeof_fr(curve)) {
        return 0;
    }
    return blob_layout(curve, query_sizes, world).total;
}

int b200_groth16_pk_create_shard(int curve, const void *alpha_g1, const void *beta_g1, const void *beta_g2,
                                 const void *delta_g1, const void *delta_g2, b200_bases_t A_shard,
                                 b200_bases_t B_g2_shard, b200_bases_t B_g1_shard, const uint32_t *B_indices_shard,
                                 size_t B_total, b200_bases_t H_shard, b200_bases_t L_shard, b200_r1cs_t r1cs,
                                 size_t num_constraints, size_t num_inputs, size_t num_variables, size_t domain_size,
                                 int rank, int world, b200_groth16_pk_t *handle)
{
    if (!alpha_g1 || !beta_g1 || !beta_g2 || !delta_g1 || !delta_g2 || !handle) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    const size_t g1 = b200_sizeof_group(curve, B200_G1), g2 = b200_sizeof_group(curve, B200_G2);
    if (!g1 || !g2) {
        return rt::set_error(B200_ERR_ARG, "unknown curve %d", curve);
    }
    if (world < 1 || rank < 0 || rank >= world) {
        return rt::set_error(B200_ERR_ARG, "rank %d outside a world of %d", rank, world);
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    const size_t nv = num_variables, ni = num_inputs, m = domain_size;
    if (ni > nv || m < num_constraints + ni + 1 || B_total > nv + 1 || (!B_indices_shard && B_total != nv + 1)) {
        return rt::set_error(B200_ERR_ARG, "sizes do not describe a proving key (r1cs_gg_ppzksnark.tcc:506-511)");
    }
    ProvingKey pk;
    pk.curve = curve;
    pk.alpha_g1.assign((const uint8_t *)alpha_g1, (const uint8_t *)alpha_g1 + g1);
    pk.beta_g1.assign((const uint8_t *)beta_g1, (const uint8_t *)beta_g1 + g1);
    pk.beta_g2.assign((const uint8_t *)beta_g2, (const uint8_t *)beta_g2 + g2);
    pk.delta_g1.assign((const uint8_t *)delta_g1, (const uint8_t *)delta_g1 + g1);
    pk.delta_g2.assign((const uint8_t *)delta_g2, (const uint8_t *)delta_g2 + g2);
    pk.A = A_shard;
    pk.B_g2 = B_g2_shard;
    pk.B_g1 = B_g1_shard;
    pk.H = H_shard;
    pk.L = L_shard;
    pk.r1cs = r1cs;
    pk.nc = num_constraints;
    pk.ni = ni;
    pk.nv = nv;
    pk.m = m;
    pk.rank = rank;
    pk.world = world;
    pk.tot[0] = nv + 1;
    pk.tot[1] = B_total;
    pk.tot[2] = m - 1;
    pk.tot[3] = nv - ni;
    for (int q = 0; q < 4; ++q) {
        shard_range(pk.tot[q], world, rank, &pk.first[q], &pk.cnt[q]);
    }
    pk.nB = pk.cnt[1];
    if (b200_bases_size(A_shard) != pk.cnt[0] || b200_bases_size(B_g2_shard) != pk.cnt[1] ||
        b200_bases_size(B_g1_shard) != pk.cnt[1] || b200_bases_size(H_shard) != pk.cnt[2] ||
        b200_bases_size(L_shard) != pk.cnt[3]) {
        return rt::set_error(B200_ERR_ARG, "a query shard does not hold b200_shard_range(total, world = %d, rank = %d) points",
                             world, rank);
    }
    const size_t fr = b200_sizeof_fr(curve);
    if (B_indices_shard) {
        for (size_t k = 0; k < pk.cnt[1]; ++k) {
            if (B_indices_shard[k] > nv || (k && B_indices_shard[k] <= B_indices_shard[k - 1])) {
                return rt::set_error(B200_ERR_ARG, "B_indices must be increasing and <= num_variables");
            }
        }
    }
    int rc = B200_OK;
    if ((rc = alloc_key_buffers(pk, fr, B_indices_shard, pk.cnt[1]))) {
        return rc;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    *handle = g_next++;
    g_keys[*handle] = pk;
    return B200_OK;
}

int b200_groth16_shard_begin(b200_groth16_pk_t handle, const void *primary_input, const void *auxiliary_input)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    if (!auxiliary_input || (pk->ni && !primary_input)) {
        return rt::set_error(B200_ERR_ARG, "null input");
    }
    return upload_witness(*pk, primary_input, auxiliary_input);
}

int b200_groth16_shard_witness_copy(b200_groth16_pk_t handle, void *dev_buf, int to_key)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    if (!dev_buf) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    const size_t bytes = (pk->nv + 1) * b200_sizeof_fr(pk->curve);
    cudaStream_t st = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(to_key ? pk->d_w : dev_buf, to_key ? dev_buf : pk->d_w, bytes, cudaMemcpyDeviceToDevice, st));
    B200_CUDA_CHECK(cudaStreamSynchronize(st));
    return B200_OK;
}

int b200_groth16_shard_h_copy(b200_groth16_pk_t handle, void *dev_buf, size_t first, size_t count, int to_key)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    if (!dev_buf || first + count > pk->m + 1) {
        return rt::set_error(B200_ERR_ARG, "null pointer or range beyond the %zu coefficients of H", pk->m + 1);
    }
    const size_t fr = b200_sizeof_fr(pk->curve);
    uint8_t *h = static_cast<uint8_t *>(pk->d_h) + first * fr;
    cudaStream_t st = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(to_key ? (void *)h : dev_buf, to_key ? dev_buf : (void *)h, count * fr, cudaMemcpyDeviceToDevice, st));
    B200_CUDA_CHECK(cudaStreamSynchronize(st));
    return B200_OK;
}

int b200_groth16_shard_chain(b200_groth16_pk_t handle, int which, void *dev_out)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    return b200_witness_chain_r1cs_dev(pk->r1cs, pk->d_w, pk->m, which, dev_out);
}

int b200_groth16_shard_h(b200_groth16_pk_t handle, void *dA, const void *dB, const void *dC)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    if (!dA && !dB && !dC) {
        return b200_witness_map_r1cs_dev(pk->r1cs, pk->d_w, pk->m, pk->d_h);
    }
    return b200_witness_finish_dev(pk->curve, dA, dB, dC, pk->m, pk->d_h);
}

int b200_groth16_shard_msm(b200_groth16_pk_t handle, unsigned mask, void *dev_blob)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    if (!dev_blob) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    return shard_msms(*pk, mask, dev_blob, nullptr, false, nullptr);
}

int b200_groth16_shard_finish_msm(b200_groth16_pk_t handle, void *dA, const void *dB, const void *dC, unsigned mask,
                                  void *dev_blob)
{
    ProvingKey key;
    {
        std::lock_guard<std::mutex> lk(g_mu); // not held over the work: a begun msm_begin of the same key runs beside this
        const ProvingKey *pk = find_key(handle);
        if (!pk) {
            return B200_ERR_HANDLE;
        }
        key = *pk;
    }
    if (!dA || !dB || !dC || !dev_blob) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    return shard_msms(key, mask, dev_blob, nullptr, false, nullptr, dA, dB, dC);
}

// The MSMs that need only the witness (A, B.g2, B.g1, L) started in the background, so that the caller can run the
// collectives and the H computation of the same proof beside them; _end waits and reports.
int b200_groth16_shard_msm_begin(b200_groth16_pk_t handle, unsigned mask, void *dev_blob)
{
    if (!dev_blob || (mask & 8u) || !(mask & 0x17u)) {
        return rt::set_error(B200_ERR_ARG, "msm_begin takes a device blob and a mask over A, B.g2, B.g1, L (H waits for `h`)");
    }
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    if (g_pending.count(handle)) {
        return rt::set_error(B200_ERR_ARG, "msm_begin: the previous msm_begin of this key has not been ended");
    }
    const ProvingKey key = *pk;
    g_pending[handle] = std::async(std::launch::async, [key, mask, dev_blob]() {
        JobResult r;
        // a slot no plan uses: this thread launches the index gather of a sparse B query and runs the plan's slot-0 job
        // (B.g1) itself, beside the caller's own slot-0 work (the chains, then H)
        r.rc = rt::bind_thread(rt::NUM_SLOTS - 1);
        if (r.rc == B200_OK) {
            r.rc = shard_msms(key, mask, dev_blob, nullptr, false, nullptr);
        }
        if (r.rc != B200_OK) {
            r.err = rt::last_error();
        }
        return r;
    });
    return B200_OK;
}

int b200_groth16_shard_msm_end(b200_groth16_pk_t handle)
{
    std::future<JobResult> f;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_pending.find(handle);
        if (it == g_pending.end()) {
            return rt::set_error(B200_ERR_ARG, "msm_end without msm_begin");
        }
        f = std::move(it->second);
        g_pending.erase(it);
    }
    const JobResult r = f.get();
    return r.rc == B200_OK ? B200_OK : rt::set_error(r.rc, "%s", r.err.c_str());
}

int b200_groth16_prove_shard(b200_groth16_pk_t handle, const void *primary_input, const void *auxiliary_input,
                             void *dev_blob, double *stage_ms)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const ProvingKey *pk = find_key(handle);
    if (!pk) {
        return B200_ERR_HANDLE;
    }
    if (!auxiliary_input || !dev_blob || (pk->ni && !primary_input)) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    double st[7] = {0, 0, 0, 0, 0, 0, 0}, msm[NUM_MSM] = {0, 0, 0, 0, 0};
    const double t0 = now_ms();
    int rc = upload_witness(*pk, primary_input, auxiliary_input);
    if (rc) {
        return rc;
    }
    st[0] = now_ms() - t0;
    if ((rc = shard_msms(*pk, 0x1f, dev_blob, msm, true, &st[1]))) {
        return rc;
    }
    st[2] = msm[0];
    st[3] = msm[1] + msm[2];
    st[4] = msm[3];
    st[5] = msm[4];
    st[6] = now_ms() - t0;
    if (stage_ms) {
        memcpy(stage_ms, st, sizeof(st));
    }
    return B200_OK;
}

int b200_groth16_assemble_prepare(int curve, const void *delta_g1, const void *delta_g2, const void *r, const void *s)
{
    if (!delta_g1 || !delta_g2 || !r || !s) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    switch (curve) {
    case B200_ALT_BN128: return assemble_prepare<CURVE_ALT_BN128>(delta_g1, delta_g2, r, s);
    case B200_BLS12_377: return assemble_prepare<CURVE_BLS12_377>(delta_g1, delta_g2, r, s);
    case B200_BW6_761: return assemble_prepare<CURVE_BW6_761>(delta_g1, delta_g2, r, s);
    default: return rt::set_error(B200_ERR_ARG, "unknown curve %d", curve);
    }
}

int b200_groth16_assemble(int curve, const void *alpha_g1, const void *beta_g1, const void *beta_g2,
                          const void *delta_g1, const void *delta_g2, const size_t query_sizes[4], int world,
                          const void *blobs, const void *r, const void *s, void *proof_out)
{
    if (!alpha_g1 || !beta_g1 || !beta_g2 || !delta_g1 || !delta_g2 || !query_sizes || !blobs || !r || !s || !proof_out ||
        world < 1) {
        return rt::set_error(B200_ERR_ARG, "null pointer or empty world");
    }
    const KeyPoints kp = {alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2};
    switch (curve) {
    case B200_ALT_BN128: return assemble_shards<CURVE_ALT_BN128>(kp, query_sizes, world, blobs, r, s, proof_out);
    case B200_BLS12_377: return assemble_shards<CURVE_BLS12_377>(kp, query_sizes, world, blobs, r, s, proof_out);
    case B200_BW6_761: return assemble_shards<CURVE_BW6_761>(kp, query_sizes, world, blobs, r, s, proof_out);
    default: return rt::set_error(B200_ERR_ARG, "unknown curve %d", curve);
    }
}

} // extern "C"
