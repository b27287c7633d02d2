// libsnark_b200_overlay.hpp — the reference-side binding of the B200 drop-in.
//
// clearmatics/libsnark has no plugin mechanism on the Groth16 proving hot path: the operators are
// C++ templates instantiated in the caller's translation unit. This header is what a
// maintainer adds to interpose the CUDA library WITHOUT editing the reference: compile the
// translation units that instantiate the prover with
//
//     -include libsnark_b200/overlay/libsnark_b200_overlay.hpp  -Iinclude  -lsnark_b200
//
// (it must be seen before the first implicit instantiation of the templates below). It provides
// explicit specialisations that forward to the C-ABI of include/libsnark_b200.h for
//   libff::multi_exp<G, Fr, multi_exp_method_BDLO12_signed, multi_exp_base_form_special>
//                                       (libff/algebra/scalar_multiplication/multiexp.hpp:63-73)
//   libff::multi_exp_filter_one_zero<...same...>                       (multiexp.hpp:83-88)
//   libsnark::kc_multi_exp_with_mixed_addition<G2, G1, Fr, BDLO12_signed, special>
//                                       (libsnark/knowledge_commitment/kc_multiexp.hpp:32-44)
//   libfqfft::basic_radix2_domain<Fr>, extended_radix2_domain<Fr> and step_radix2_domain<Fr>::{FFT, iFFT, cosetFFT,
//   icosetFFT, divide_by_Z_on_coset}
//                                       (libfqfft/evaluation_domain/domains/basic_radix2_domain.hpp)
// for G in {G1, G2} of alt_bn128, bls12_377 and bw6_761. Every other instantiation (other methods,
// base forms, curves, domains) keeps the reference's CPU template. There is no CPU fallback inside
// the specialisations: a failing C-ABI call throws (std::runtime_error, or
// libfqfft::DomainSizeException for a size the basic radix-2 domain does not have).
//
// Proving-key residency: the reference passes host iterators on every call and gives no lifetime
// signal, so base vectors of 2^12 points or more are cached on the device keyed by (host pointer,
// stride, count, curve, group, a content fingerprint). The fingerprint hashes 256 points spread evenly
// over the vector (every point with B200_OVERLAY_FINGERPRINT=full or set_cache_fingerprint_full(true)),
// so a different key of the same shape at a reused address is uploaded afresh instead of being served
// stale; the constraint-system and index caches carry content fingerprints the same way. The cache is
// bounded (B200_OVERLAY_CACHE_BYTES of host-side vector bytes, default 16 GiB, least recently used
// evicted first). Smaller vectors (verifier accumulations, generator temporaries) go through the
// uncached b200_msm. libsnark_b200::release_cached_bases() drops the cache explicitly.
#ifndef LIBSNARK_B200_OVERLAY_HPP_
#define LIBSNARK_B200_OVERLAY_HPP_

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include <libff/algebra/curves/alt_bn128/alt_bn128_pp.hpp>
#include <libff/algebra/curves/bls12_377/bls12_377_pp.hpp>
#include <libff/algebra/curves/bw6_761/bw6_761_pp.hpp>
#include <libff/algebra/scalar_multiplication/multiexp.hpp>
#include <libff/common/profiling.hpp>
#include <libfqfft/evaluation_domain/domains/basic_radix2_domain.hpp>
#include <libfqfft/evaluation_domain/domains/extended_radix2_domain.hpp>
#include <libfqfft/evaluation_domain/domains/step_radix2_domain.hpp>
#include <libfqfft/evaluation_domain/get_evaluation_domain.hpp>
#include <libfqfft/tools/exceptions.hpp>
#include <libsnark/knowledge_commitment/kc_multiexp.hpp>
#include <libsnark/reductions/r1cs_to_qap/r1cs_to_qap.hpp>
#include <libsnark/reductions/r1cs_to_sap/r1cs_to_sap.hpp>
#include <libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.hpp>

#include "libsnark_b200.h"

namespace libsnark_b200
{

template<typename T> struct curve_of;
#define B200_OVERLAY_TRAIT(TYPE, CURVE, GROUP)                                                     \
    template<> struct curve_of<TYPE> {                                                             \
        static constexpr int curve = CURVE;                                                        \
        static constexpr int group = GROUP;                                                        \
    };
B200_OVERLAY_TRAIT(libff::alt_bn128_G1, B200_ALT_BN128, B200_G1)
B200_OVERLAY_TRAIT(libff::alt_bn128_G2, B200_ALT_BN128, B200_G2)
B200_OVERLAY_TRAIT(libff::bls12_377_G1, B200_BLS12_377, B200_G1)
B200_OVERLAY_TRAIT(libff::bls12_377_G2, B200_BLS12_377, B200_G2)
B200_OVERLAY_TRAIT(libff::bw6_761_G1, B200_BW6_761, B200_G1)
B200_OVERLAY_TRAIT(libff::bw6_761_G2, B200_BW6_761, B200_G2)
B200_OVERLAY_TRAIT(libff::alt_bn128_Fr, B200_ALT_BN128, 0)
B200_OVERLAY_TRAIT(libff::bls12_377_Fr, B200_BLS12_377, 0)
B200_OVERLAY_TRAIT(libff::bw6_761_Fr, B200_BW6_761, 0)
#undef B200_OVERLAY_TRAIT

inline void check(int rc, const char *what)
{
    if (rc == B200_OK) {
        return;
    }
    const std::string msg = std::string("libsnark_b200: ") + what + ": " + b200_last_error();
    if (rc == B200_ERR_DOMAIN) {
        throw libfqfft::DomainSizeException(msg);
    }
    throw std::runtime_error(msg);
}

// ---- content fingerprints of the cached host data ---------------------------------------------------
// 64-bit multiply-xorshift over 8-byte words (all hashed types are arrays of 64-bit limbs / size_t).
inline unsigned long long hash_words(unsigned long long h, const void *p, size_t bytes)
{
    const unsigned char *b = static_cast<const unsigned char *>(p);
    size_t i = 0;
    if (bytes >= 4096) { // long inputs (index arrays of MBs): four independent lanes, so the multiplies overlap
        unsigned long long l[4] = {h, h ^ 0x243F6A8885A308D3ull, h ^ 0x13198A2E03707344ull, h ^ 0xA4093822299F31D0ull};
        for (; i + 32 <= bytes; i += 32) {
            unsigned long long w[4];
            memcpy(w, b + i, 32);
            for (int k = 0; k < 4; ++k) {
                l[k] = (l[k] ^ w[k]) * 0x9E3779B97F4A7C15ull;
                l[k] ^= l[k] >> 29;
            }
        }
        for (int k = 0; k < 4; ++k) {
            h = (h ^ l[k]) * 0x9E3779B97F4A7C15ull;
            h ^= h >> 29;
        }
    }
    for (; i + 8 <= bytes; i += 8) {
        unsigned long long w;
        memcpy(&w, b + i, 8);
        h = (h ^ w) * 0x9E3779B97F4A7C15ull;
        h ^= h >> 29;
    }
    for (; i < bytes; ++i) {
        h = (h ^ b[i]) * 1099511628211ull;
    }
    return h;
}
inline bool &fingerprint_full_flag()
{
    static bool full = [] {
        const char *e = getenv("B200_OVERLAY_FINGERPRINT");
        return e != nullptr && strcmp(e, "full") == 0;
    }();
    return full;
}
// true: every cached vector / constraint system is hashed in full on every lookup (hundreds of MB per proof at
// 2^20 — a few ms on 16 cores' worth of memory bandwidth, single-threaded tens of ms); false: evenly spread samples.
inline void set_cache_fingerprint_full(bool full) { fingerprint_full_flag() = full; }
constexpr size_t FINGERPRINT_SAMPLES = 256;
constexpr size_t CACHE_MIN_POINTS = size_t(1) << 12; // smaller base vectors are one-off: not cached
constexpr size_t FINGERPRINT_ROW_TERMS = 64;          // terms hashed of one sampled linear combination

struct pk_cache;
inline void pk_cache_clear();

struct bases_cache {
    typedef std::tuple<const void *, size_t, size_t, int, int, unsigned long long> key_t;
    // A set that is seen a second time is a proving-key query, not a one-off vector: it then gets its window tables
    // (b200_bases_precompute: W x the device memory, 12-30 % faster MSMs). B200_OVERLAY_NO_TABLES=1 turns that off.
    struct entry {
        b200_bases_t handle;
        bool tables_tried;
        size_t bytes;             // host-side size of the vector (the budget's unit; tables multiply the device side)
        unsigned long long used;  // tick of the last lookup
    };
    std::map<key_t, entry> sets;
    unsigned long long tick = 0, pin = 0; // entries used after `pin` belong to the call in progress: never evicted
    size_t total_bytes = 0;
    void begin_group() { pin = tick; }    // call before the lookups of one operator call
    static bases_cache &get()
    {
        static bases_cache c;
        return c;
    }
    static size_t budget()
    {
        static const size_t b = getenv("B200_OVERLAY_CACHE_BYTES") ? (size_t)strtoull(getenv("B200_OVERLAY_CACHE_BYTES"), nullptr, 10)
                                                                     : (size_t(16) << 30);
        return b;
    }
    static unsigned long long fingerprint(const void *p, size_t stride, size_t n, size_t point_bytes)
    {
        unsigned long long h = 1469598103934665603ull ^ n;
        const unsigned char *b = static_cast<const unsigned char *>(p);
        if (n <= FINGERPRINT_SAMPLES || fingerprint_full_flag()) {
            for (size_t i = 0; i < n; ++i) {
                h = hash_words(h, b + i * stride, point_bytes);
            }
            return h;
        }
        for (size_t k = 0; k < FINGERPRINT_SAMPLES; ++k) { // first, last and 254 evenly spread points
            const size_t i = (size_t)((unsigned __int128)k * (n - 1) / (FINGERPRINT_SAMPLES - 1));
            h = hash_words(h, b + i * stride, point_bytes);
        }
        return h;
    }
    void evict_to(size_t limit)
    {
        bool any = false;
        while (total_bytes > limit) {
            auto lru = sets.end();
            for (auto it = sets.begin(); it != sets.end(); ++it) {
                if (it->second.used <= pin && (lru == sets.end() || it->second.used < lru->second.used)) {
                    lru = it;
                }
            }
            if (lru == sets.end()) {
                break; // everything left is in use by this call: over budget until it ends
            }
            if (!any) {
                pk_cache_clear(); // prover keys hold base handles: drop them before any handle dies
                any = true;
            }
            b200_bases_free(lru->second.handle);
            total_bytes -= lru->second.bytes;
            sets.erase(lru);
        }
    }
    b200_bases_t lookup(int curve, int group, const void *p, size_t stride, size_t n, size_t point_bytes)
    {
        const key_t key(p, stride, n, curve, group, fingerprint(p, stride, n, point_bytes));
        auto it = sets.find(key);
        if (it != sets.end()) {
            it->second.used = ++tick;
            if (!it->second.tables_tried) {
                it->second.tables_tried = true;
                static const bool no_tables = getenv("B200_OVERLAY_NO_TABLES") != nullptr;
                if (!no_tables && n >= (size_t(1) << 14)) {
                    (void)b200_bases_precompute(it->second.handle); // on failure the set simply stays table-free
                }
            }
            return it->second.handle;
        }
        // same address and shape, other contents: the old copy is stale for good
        for (auto st = sets.begin(); st != sets.end();) {
            const key_t &k = st->first;
            if (std::get<0>(k) == p && std::get<1>(k) == stride && std::get<3>(k) == curve && std::get<4>(k) == group) {
                pk_cache_clear();
                b200_bases_free(st->second.handle);
                total_bytes -= st->second.bytes;
                st = sets.erase(st);
            } else {
                ++st;
            }
        }
        const size_t bytes = n * point_bytes;
        evict_to(budget() > bytes ? budget() - bytes : 0);
        b200_bases_t h = 0;
        check(b200_bases_upload(curve, group, p, stride, n, &h), "b200_bases_upload");
        sets[key] = entry{h, false, bytes, ++tick};
        total_bytes += bytes;
        return h;
    }
    void clear()
    {
        for (auto &kv : sets) {
            b200_bases_free(kv.second.handle);
        }
        sets.clear();
        total_bytes = 0;
    }
};
// b200_groth16_pk_t per (A, B.g2, B.g1, H, L, R1CS) handle tuple (r1cs_gg_ppzksnark_prover_fused)
struct pk_cache {
    // the last field: fingerprint of the five single points and of B_query.indices (everything pk_create copies)
    typedef std::tuple<b200_bases_t, b200_bases_t, b200_bases_t, b200_bases_t, b200_bases_t, unsigned long long,
                       unsigned long long> key_t;
    std::map<key_t, b200_groth16_pk_t> keys;
    static pk_cache &get()
    {
        static pk_cache c;
        return c;
    }
    void clear()
    {
        for (auto &kv : keys) {
            b200_groth16_pk_free(kv.second);
        }
        keys.clear();
    }
};
inline void pk_cache_clear() { pk_cache::get().clear(); }
inline void release_cached_bases()
{
    pk_cache::get().clear(); // the keys refer to the base handles freed next
    bases_cache::get().clear();
}

// (coeff * v[i]) * g for every i, in special form: libff::batch_exp / batch_exp_with_coeff + batch_to_special
// (multiexp.tcc:879-974). The base is entry [0][1] of the reference's window table (get_window_table, :808-844).
template<typename GroupT, typename FieldT>
std::vector<GroupT> batch_exp_special(const libff::window_table<GroupT> &table, const FieldT *coeff, const FieldT *v, size_t n)
{
    std::vector<GroupT> res(n, GroupT::zero());
    if (n == 0) {
        return res;
    }
    if (table.empty() || table[0].size() < 2) {
        throw std::runtime_error("libsnark_b200: batch_exp needs a window table with at least two entries");
    }
    GroupT g = table[0][1];
    g.to_special();
    check(b200_batch_exp(curve_of<GroupT>::curve, curve_of<GroupT>::group, &g, coeff, v, n, res.data()), "b200_batch_exp");
    return res;
}

// libsnark::kc_batch_exp (kc_multiexp.tcc:147-239): zeros of v are skipped, the rest become
// (T1_coeff * v[i]) * g_T1 | (T2_coeff * v[i]) * g_T2, always in special form.
template<typename T1, typename T2, typename FieldT>
libsnark::knowledge_commitment_vector<T1, T2> kc_batch_exp_special(const libff::window_table<T1> &T1_table,
                                                                   const libff::window_table<T2> &T2_table,
                                                                   const FieldT &T1_coeff, const FieldT &T2_coeff,
                                                                   const std::vector<FieldT> &v)
{
    libsnark::knowledge_commitment_vector<T1, T2> res;
    res.domain_size_ = v.size();
    std::vector<FieldT> nz;
    for (size_t i = 0; i < v.size(); ++i) {
        if (!v[i].is_zero()) {
            nz.push_back(v[i]);
            res.indices.push_back(i);
        }
    }
    if (!libff::inhibit_profiling_info) {
        libff::print_indent();
        printf("Non-zero coordinate count: %zu/%zu (%0.2f%%)\n", nz.size(), v.size(), 100. * nz.size() / v.size());
    }
    const std::vector<T1> a = batch_exp_special<T1, FieldT>(T1_table, &T1_coeff, nz.data(), nz.size());
    const std::vector<T2> b = batch_exp_special<T2, FieldT>(T2_table, &T2_coeff, nz.data(), nz.size());
    res.values.reserve(nz.size());
    for (size_t k = 0; k < nz.size(); ++k) {
        res.values.emplace_back(libsnark::knowledge_commitment<T1, T2>(a[k], b[k]));
    }
    return res;
}

// sum_i scalar[i] * base[i] over contiguous host ranges (the reference's iterator pairs).
template<typename GroupT, typename FieldT>
GroupT msm_contiguous(const GroupT *bases, size_t n, const FieldT *scalars)
{
    GroupT out = GroupT::zero();
    if (n == 0) {
        return out;
    }
    typedef curve_of<GroupT> tr;
    if (n < CACHE_MIN_POINTS) { // a one-off vector (verifier accumulation, generator temporary): nothing kept on the device
        check(b200_msm(tr::curve, tr::group, bases, sizeof(GroupT), n, scalars, &out), "b200_msm");
        return out;
    }
    bases_cache::get().begin_group();
    const b200_bases_t h = bases_cache::get().lookup(tr::curve, tr::group, bases, sizeof(GroupT), n, sizeof(GroupT));
    check(b200_msm_resident(h, 0, n, scalars, 0, &out), "b200_msm_resident");
    return out;
}

// The three statistics lines of multi_exp_filter_one_zero (multiexp.tcc:737-751) / kc (kc_multiexp.tcc:91-105).
template<typename FieldT> void print_filter_stats(const FieldT *scalars, size_t n)
{
    if (libff::inhibit_profiling_info) {
        return;
    }
    size_t num_skip = 0, num_add = 0;
    const FieldT zero = FieldT::zero(), one = FieldT::one();
    for (size_t i = 0; i < n; ++i) {
        num_skip += scalars[i] == zero;
        num_add += scalars[i] == one;
    }
    const size_t num_other = n - num_skip - num_add;
    libff::print_indent();
    printf("* Elements of w skipped: %zu (%0.2f%%)\n", num_skip, 100. * num_skip / (n ? n : 1));
    libff::print_indent();
    printf("* Elements of w processed with special addition: %zu (%0.2f%%)\n", num_add, 100. * num_add / (n ? n : 1));
    libff::print_indent();
    printf("* Elements of w remaining: %zu (%0.2f%%)\n", num_other, 100. * num_other / (n ? n : 1));
}

template<typename T1, typename T2, typename FieldT>
libsnark::knowledge_commitment<T1, T2> kc_msm(
    const libsnark::knowledge_commitment_vector<T1, T2> &vec, const size_t min_idx, const size_t max_idx,
    typename std::vector<FieldT>::const_iterator scalar_start, typename std::vector<FieldT>::const_iterator scalar_end)
{
    typedef libsnark::knowledge_commitment<T1, T2> KC;
    (void)scalar_end;
    // kc_multiexp.tcc:40-44, :63: the stored entries with min_idx <= index < max_idx
    const auto lo = std::lower_bound(vec.indices.begin(), vec.indices.end(), min_idx);
    const auto hi = std::lower_bound(lo, vec.indices.end(), max_idx);
    const size_t first = lo - vec.indices.begin(), cnt = hi - lo;
    KC out(T1::zero(), T2::zero());
    if (cnt == 0) {
        return out;
    }
    std::vector<FieldT> gathered(cnt);
    for (size_t k = 0; k < cnt; ++k) {
        gathered[k] = *(scalar_start + (vec.indices[first + k] - min_idx));
    }
    print_filter_stats(gathered.data(), cnt);
    const unsigned char *vals = reinterpret_cast<const unsigned char *>(vec.values.data());
    const size_t n = vec.values.size();
    const size_t off2 = reinterpret_cast<const unsigned char *>(&vec.values[0].h) - vals;
    bases_cache &c = bases_cache::get();
    c.begin_group();
    const b200_bases_t hg = c.lookup(curve_of<T1>::curve, curve_of<T1>::group, vals, sizeof(KC), n, sizeof(T1));
    const b200_bases_t hh = c.lookup(curve_of<T2>::curve, curve_of<T2>::group, vals + off2, sizeof(KC), n, sizeof(T2));
    check(b200_msm_resident(hg, first, cnt, gathered.data(), 0, &out.g), "b200_msm_resident (kc.g)");
    check(b200_msm_resident(hh, first, cnt, gathered.data(), 0, &out.h), "b200_msm_resident (kc.h)");
    return out;
}

template<typename FieldT> void ntt(std::vector<FieldT> &a, size_t m, int op, const FieldT *g)
{
    if (a.size() != m) { // basic_radix2_domain.tcc:57, :65
        throw libfqfft::DomainSizeException("basic_radix2: expected a.size() == this->m");
    }
    check(b200_ntt(curve_of<FieldT>::curve, op, a.data(), m, g), "b200_ntt");
}

// The NTT chain of r1cs_to_qap_witness_map (r1cs_to_qap.tcc:252-330) as ONE device pipeline, for
// callers that can change one line of the prover: evaluations stay on the device between the seven
// transforms. Written against the reference's types; Groth16 case (d1 = d2 = d3 = 0).
template<typename FieldT>
libsnark::qap_witness<FieldT> r1cs_to_qap_witness_map_device(
    const libsnark::r1cs_constraint_system<FieldT> &cs, const libsnark::r1cs_primary_input<FieldT> &primary_input,
    const libsnark::r1cs_auxiliary_input<FieldT> &auxiliary_input, const FieldT &d1, const FieldT &d2, const FieldT &d3,
    bool force_pow_2_domain);
// older name (round 1), Groth16 case d1 = d2 = d3 = 0
template<typename FieldT>
libsnark::qap_witness<FieldT> r1cs_to_qap_witness_map_fused(
    const libsnark::r1cs_constraint_system<FieldT> &cs, const libsnark::r1cs_primary_input<FieldT> &primary_input,
    const libsnark::r1cs_auxiliary_input<FieldT> &auxiliary_input)
{
    return r1cs_to_qap_witness_map_device<FieldT>(cs, primary_input, auxiliary_input, FieldT::zero(), FieldT::zero(),
                                                  FieldT::zero(), false);
}

// ---------------------------------------------------------------------------------------------------
// Device-resident prover (SURVEY.md §8f N1 + N2): the orchestration of r1cs_gg_ppzksnark_prover
// (r1cs_gg_ppzksnark.tcc:460-603) with the witness uploaded ONCE and everything between the witness
// and the five MSM results kept on the device: sparse R1CS evaluation, the seven transforms, and the
// scalars of all MSMs (the witness itself and coefficients_for_H) never return to the host.
// The operator-level specialisations above stay available; this is the one-call alternative for
// callers that can change the line that invokes the prover.
struct device_buffer {
    void *p = nullptr;
    size_t cap = 0;
    void reserve(size_t bytes)
    {
        if (bytes <= cap) {
            return;
        }
        if (p) {
            b200_dev_free(p);
            p = nullptr;
        }
        check(b200_dev_alloc(bytes, &p), "b200_dev_alloc");
        cap = bytes;
    }
    ~device_buffer()
    {
        // process exit: the CUDA context may already be gone; leak rather than call into it
    }
};

struct r1cs_cache {
    struct entry {
        b200_r1cs_t handle = 0;
        size_t nc = 0, nv = 0;
        unsigned long long fp = 0;
    };
    std::map<const void *, entry> sets;
    static r1cs_cache &get()
    {
        static r1cs_cache c;
        return c;
    }
    // num_inputs and the terms (index, coefficient; up to FINGERPRINT_ROW_TERMS per linear combination) of up to
    // FINGERPRINT_SAMPLES evenly spread constraints — everything in full mode
    template<typename FieldT> static unsigned long long fingerprint(const libsnark::r1cs_constraint_system<FieldT> &cs)
    {
        const size_t nc = cs.num_constraints();
        unsigned long long h = 1469598103934665603ull ^ cs.num_inputs();
        const bool all = nc <= FINGERPRINT_SAMPLES || fingerprint_full_flag();
        const size_t samples = all ? nc : FINGERPRINT_SAMPLES;
        for (size_t k = 0; k < samples; ++k) {
            const size_t i = all ? k : (size_t)((unsigned __int128)k * (nc - 1) / (FINGERPRINT_SAMPLES - 1));
            const libsnark::linear_combination<FieldT> *lcs[3] = {&cs.constraints[i].a, &cs.constraints[i].b, &cs.constraints[i].c};
            for (const auto *lc : lcs) {
                const unsigned long long cnt = lc->terms.size();
                h = hash_words(h, &cnt, 8);
                // a row can be as long as the variable count (the closing constraint of the reference's own examples sums
                // every variable: 17 ms of hashing per proof at 2^20): of such a row, evenly spread terms only
                const bool all_terms = fingerprint_full_flag() || cnt <= FINGERPRINT_ROW_TERMS;
                const size_t nt = all_terms ? (size_t)cnt : FINGERPRINT_ROW_TERMS;
                for (size_t q = 0; q < nt; ++q) {
                    const auto &t = lc->terms[all_terms ? q : (size_t)((unsigned __int128)q * (cnt - 1) / (FINGERPRINT_ROW_TERMS - 1))];
                    const unsigned long long idx = t.index;
                    h = hash_words(h, &idx, 8);
                    h = hash_words(h, &t.coeff, sizeof(FieldT));
                }
            }
        }
        return h;
    }
    template<typename FieldT> b200_r1cs_t lookup(const libsnark::r1cs_constraint_system<FieldT> &cs)
    {
        const unsigned long long fp = fingerprint(cs);
        auto it = sets.find(&cs);
        if (it != sets.end() && it->second.nc == cs.num_constraints() && it->second.nv == cs.num_variables() &&
            it->second.fp == fp) {
            return it->second.handle;
        }
        if (it != sets.end()) { // same address, other contents
            pk_cache::get().clear();
            b200_r1cs_free(it->second.handle);
            sets.erase(it);
        }
        const size_t nc = cs.num_constraints();
        std::vector<uint32_t> row_ptr(3 * nc + 1, 0);
        size_t nnz = 0;
        for (size_t i = 0; i < nc; ++i) {
            nnz += cs.constraints[i].a.terms.size() + cs.constraints[i].b.terms.size() + cs.constraints[i].c.terms.size();
        }
        std::vector<uint32_t> col(nnz);
        std::vector<FieldT> val(nnz);
        size_t e = 0;
        for (int mat = 0; mat < 3; ++mat) {
            for (size_t i = 0; i < nc; ++i) {
                const libsnark::linear_combination<FieldT> &lc =
                    mat == 0 ? cs.constraints[i].a : (mat == 1 ? cs.constraints[i].b : cs.constraints[i].c);
                row_ptr[mat * nc + i] = (uint32_t)e;
                for (const auto &t : lc.terms) {
                    col[e] = (uint32_t)t.index;
                    val[e] = t.coeff;
                    ++e;
                }
            }
        }
        row_ptr[3 * nc] = (uint32_t)e;
        entry en;
        en.nc = nc;
        en.nv = cs.num_variables();
        en.fp = fp;
        check(b200_r1cs_upload(curve_of<FieldT>::curve, nc, cs.num_inputs(), cs.num_variables(), row_ptr.data(),
                               col.data(), val.data(), &en.handle),
              "b200_r1cs_upload");
        sets[&cs] = en;
        return en.handle;
    }
    void clear()
    {
        for (auto &kv : sets) {
            b200_r1cs_free(kv.second.handle);
        }
        sets.clear();
    }
};
inline void release_cached_r1cs()
{
    pk_cache::get().clear();
    r1cs_cache::get().clear();
}

// libsnark::r1cs_to_sap_witness_map (r1cs_to_sap.tcc:297-481; the witness map of r1cs_se_ppzksnark): the values of the
// constraints' three linear combinations come from the device's sparse evaluation (the reference evaluates a and b four
// times each per constraint on one core), the SAP vectors are formed on the host from them with additions only, and
// H = (A^2 - C) / Z runs as one device pipeline (b200_witness_map_h with B = A); the ZK patch 2 d1 coef(A) - d2 + d1^2 Z
// costs one more inverse transform. Exact field arithmetic: every coefficient equals the reference's.
template<typename FieldT>
libsnark::sap_witness<FieldT> r1cs_to_sap_witness_map_device(
    const libsnark::r1cs_constraint_system<FieldT> &cs, const libsnark::r1cs_primary_input<FieldT> &primary_input,
    const libsnark::r1cs_auxiliary_input<FieldT> &auxiliary_input, const FieldT &d1, const FieldT &d2)
{
    const size_t nc = cs.num_constraints(), ni = cs.num_inputs(), nv = cs.num_variables();
    if (primary_input.size() != ni || primary_input.size() + auxiliary_input.size() != nv) {
        throw std::runtime_error("libsnark_b200: witness size does not match the constraint system");
    }
    const std::shared_ptr<libfqfft::evaluation_domain<FieldT>> domain = libsnark::r1cs_to_sap_get_domain(cs);
    const size_t m = domain->m;
    libsnark::r1cs_variable_assignment<FieldT> full = primary_input;
    full.insert(full.end(), auxiliary_input.begin(), auxiliary_input.end());
    std::vector<FieldT> w(nv + 1);
    w[0] = FieldT::one();
    std::copy(full.begin(), full.end(), w.begin() + 1);
    const size_t mr = nc + ni + 1; // rows [0, nc) of ea, eb, ec are <a_i, w>, <b_i, w>, <c_i, w>
    std::vector<FieldT> ea(mr), eb(mr), ec(mr);
    check(b200_r1cs_evaluate(r1cs_cache::get().lookup(cs), w.data(), mr, ea.data(), eb.data(), ec.data()), "b200_r1cs_evaluate");
    const FieldT one = FieldT::one();
    full.reserve(nv + nc + ni);
    for (size_t i = 0; i < nc; ++i) { // :330-341, (a - b)^2
        const FieldT t = ea[i] - eb[i];
        full.push_back(t * t);
    }
    for (size_t i = 1; i <= ni; ++i) { // :342-351, (x_i - 1)^2
        const FieldT t = full[i - 1] - one;
        full.push_back(t * t);
    }
    auto times_four = [](const FieldT &x) {
        const FieldT two = x + x;
        return two + two;
    };
    std::vector<FieldT> aA(m, FieldT::zero()), aC(m, FieldT::zero());
    const size_t off = 2 * nc, evo = nv + 1, evo2 = nv + nc;
    for (size_t i = 0; i < nc; ++i) { // :357-363, :414-420
        aA[2 * i] = ea[i] + eb[i];
        aA[2 * i + 1] = ea[i] - eb[i];
        aC[2 * i] = times_four(ec[i]) + full[evo + i - 1];
        aC[2 * i + 1] = full[evo + i - 1];
    }
    aA[off] = one; // :367, :423
    aC[off] = one;
    for (size_t i = 1; i <= ni; ++i) { // :369-375, :425-433
        aA[off + 2 * i - 1] = full[i - 1] + one;
        aA[off + 2 * i] = full[i - 1] - one;
        aC[off + 2 * i - 1] = times_four(full[i - 1]) + full[evo2 + i - 1];
        aC[off + 2 * i] = full[evo2 + i - 1];
    }
    std::vector<FieldT> H(m + 1, FieldT::zero());
    check(b200_witness_map_h(curve_of<FieldT>::curve, aA.data(), aA.data(), aC.data(), m, H.data()), "b200_witness_map_h");
    if (!(d1.is_zero() && d2.is_zero())) { // :388-393 and :461-469
        check(b200_ntt(curve_of<FieldT>::curve, B200_IFFT, aA.data(), m, nullptr), "b200_ntt");
        const FieldT two_d1 = d1 + d1;
        for (size_t i = 0; i < m; ++i) {
            H[i] += two_d1 * aA[i];
        }
        H[0] -= d2;
        domain->add_poly_Z(d1 * d1, H);
    }
    return libsnark::sap_witness<FieldT>(nv + nc + ni, m, ni, d1, d2, full, std::move(H));
}

// libsnark::r1cs_to_qap_witness_map (r1cs_to_qap.tcc:209-344) on the device, ZK patch included: the constraint system
// as a cached CSR, the witness up, coefficients_for_H back (b200_witness_map_r1cs_zk). The reference's
// assert(cs.is_satisfied(...)) (a debug check that re-evaluates every constraint on one core) is not repeated.
template<typename FieldT>
libsnark::qap_witness<FieldT> r1cs_to_qap_witness_map_device(
    const libsnark::r1cs_constraint_system<FieldT> &cs, const libsnark::r1cs_primary_input<FieldT> &primary_input,
    const libsnark::r1cs_auxiliary_input<FieldT> &auxiliary_input, const FieldT &d1, const FieldT &d2, const FieldT &d3,
    bool force_pow_2_domain)
{
    const size_t min_n = cs.num_constraints() + cs.num_inputs() + 1; // :224-227
    const size_t m = libfqfft::get_evaluation_domain<FieldT>(force_pow_2_domain ? (size_t)1 << libff::log2(min_n) : min_n)->m;
    if (primary_input.size() != cs.num_inputs() || primary_input.size() + auxiliary_input.size() != cs.num_variables()) {
        throw std::runtime_error("libsnark_b200: witness size does not match the constraint system");
    }
    libsnark::r1cs_variable_assignment<FieldT> full = primary_input;
    full.insert(full.end(), auxiliary_input.begin(), auxiliary_input.end());
    std::vector<FieldT> w(full.size() + 1);
    w[0] = FieldT::one();
    std::copy(full.begin(), full.end(), w.begin() + 1);
    const FieldT d[3] = {d1, d2, d3};
    std::vector<FieldT> H(m + 1, FieldT::zero());
    check(b200_witness_map_r1cs_zk(r1cs_cache::get().lookup(cs), w.data(), m, d, H.data()), "b200_witness_map_r1cs_zk");
    return libsnark::qap_witness<FieldT>(cs.num_variables(), m, cs.num_inputs(), d1, d2, d3, full, std::move(H));
}

// device copy of a sparse query's index list (or nothing when the query is dense)
struct index_cache {
    struct entry {
        void *d_idx = nullptr;
        bool dense = true;
        size_t n = 0;
        unsigned long long fp = 0;
    };
    std::map<const void *, entry> sets;
    static index_cache &get()
    {
        static index_cache c;
        return c;
    }
    const entry &lookup(const std::vector<size_t> &indices)
    {
        // the whole array is hashed: 8 bytes per entry, about a millisecond at 2^20
        const unsigned long long fp = hash_words(1469598103934665603ull, indices.data(), indices.size() * sizeof(size_t));
        auto it = sets.find(indices.data());
        if (it != sets.end() && it->second.n == indices.size() && it->second.fp == fp) {
            return it->second;
        }
        if (it != sets.end()) {
            pk_cache::get().clear(); // a prover key may hold the old index list
            if (it->second.d_idx) {
                b200_dev_free(it->second.d_idx);
            }
            sets.erase(it);
        }
        entry en;
        en.n = indices.size();
        en.fp = fp;
        std::vector<uint32_t> idx(indices.size());
        for (size_t k = 0; k < indices.size(); ++k) {
            idx[k] = (uint32_t)indices[k];
            en.dense = en.dense && indices[k] == k;
        }
        if (!en.dense) {
            check(b200_dev_alloc(idx.size() * 4, &en.d_idx), "b200_dev_alloc");
            check(b200_memcpy_h2d(en.d_idx, idx.data(), idx.size() * 4), "b200_memcpy_h2d");
        }
        return sets[indices.data()] = en;
    }
};

template<typename ppT>
libsnark::r1cs_gg_ppzksnark_proof<ppT> r1cs_gg_ppzksnark_prover_fused(
    const libsnark::r1cs_gg_ppzksnark_proving_key<ppT> &pk,
    const libsnark::r1cs_gg_ppzksnark_primary_input<ppT> &primary_input,
    const libsnark::r1cs_gg_ppzksnark_auxiliary_input<ppT> &auxiliary_input, const libff::Fr<ppT> &r,
    const libff::Fr<ppT> &s, bool force_pow_2_domain = false)
{
    typedef libff::Fr<ppT> FrT;
    typedef libff::G1<ppT> G1T;
    typedef libff::G2<ppT> G2T;
    typedef libsnark::knowledge_commitment<G2T, G1T> KC;
    const int curve = curve_of<FrT>::curve;
    const auto &cs = pk.constraint_system;
    const size_t nv = cs.num_variables(), ni = cs.num_inputs();
    if (primary_input.size() != ni || primary_input.size() + auxiliary_input.size() != nv) {
        throw std::runtime_error("libsnark_b200: witness size does not match the constraint system");
    }
    const size_t min_n = cs.num_constraints() + ni + 1; // r1cs_to_qap.tcc:224-227
    const size_t m = libfqfft::get_evaluation_domain<FrT>(force_pow_2_domain ? (size_t)1 << libff::log2(min_n) : min_n)->m;
    static const bool trace = getenv("B200_OVERLAY_TRACE") != nullptr;
    long long t_prev = libff::get_nsec_time();
    auto lap = [&](const char *what) {
        if (trace) {
            const long long now = libff::get_nsec_time();
            fprintf(stderr, "[b200 fused prover] %-28s %9.3f ms\n", what, (now - t_prev) * 1e-6);
            t_prev = now;
        }
    };

    if (!trace) {
        // The whole prover behind one C-ABI call: the key's queries (from the caches below, with their window
        // tables from the second proof on), the CSR and the five single points become a b200_groth16_pk_t once;
        // b200_groth16_prove then runs the witness map and the five MSMs concurrently on the device.
        static const bool times = getenv("B200_OVERLAY_TIMES") != nullptr; // one line per proof on stderr
        const long long t_in = libff::get_nsec_time();
        bases_cache &bc = bases_cache::get();
        bc.begin_group();
        const auto &vec = pk.B_query;
        const unsigned char *vals = reinterpret_cast<const unsigned char *>(vec.values.data());
        const size_t off2 = vec.values.empty() ? 0 : reinterpret_cast<const unsigned char *>(&vec.values[0].h) - vals;
        const b200_bases_t hA = bc.lookup(curve, B200_G1, pk.A_query.data(), sizeof(G1T), pk.A_query.size(), sizeof(G1T));
        const b200_bases_t hB2 = bc.lookup(curve, B200_G2, vals, sizeof(KC), vec.values.size(), sizeof(G2T));
        const b200_bases_t hB1 = bc.lookup(curve, B200_G1, vals + off2, sizeof(KC), vec.values.size(), sizeof(G1T));
        const b200_bases_t hH = bc.lookup(curve, B200_G1, pk.H_query.data(), sizeof(G1T), pk.H_query.size(), sizeof(G1T));
        const b200_bases_t hL = bc.lookup(curve, B200_G1, pk.L_query.data(), sizeof(G1T), pk.L_query.size(), sizeof(G1T));
        const long long t_bases = libff::get_nsec_time();
        const b200_r1cs_t hr = r1cs_cache::get().lookup(cs);
        const long long t_r1cs = libff::get_nsec_time();
        std::map<pk_cache::key_t, b200_groth16_pk_t> &keys = pk_cache::get().keys;
        unsigned long long pfp = hash_words(1469598103934665603ull, &pk.alpha_g1, sizeof(G1T));
        pfp = hash_words(pfp, &pk.beta_g1, sizeof(G1T));
        pfp = hash_words(pfp, &pk.beta_g2, sizeof(G2T));
        pfp = hash_words(pfp, &pk.delta_g1, sizeof(G1T));
        pfp = hash_words(pfp, &pk.delta_g2, sizeof(G2T));
        // B_query.indices: 8 MiB at 2^20, 1.5 ms to hash in full on every proof — like the bases, evenly spread samples
        // (4096 of them) plus the length unless B200_OVERLAY_FINGERPRINT=full; a regenerated key changes every point above
        {
            const size_t ni_ = vec.indices.size(), samples = 4096;
            pfp = hash_words(pfp, &ni_, sizeof(ni_));
            if (ni_ <= samples || fingerprint_full_flag()) {
                pfp = hash_words(pfp, vec.indices.data(), ni_ * sizeof(size_t));
            } else {
                for (size_t k = 0; k < samples; ++k) {
                    const size_t i = (size_t)((unsigned __int128)k * (ni_ - 1) / (samples - 1));
                    pfp = hash_words(pfp, &vec.indices[i], sizeof(size_t));
                }
            }
        }
        const pk_cache::key_t key(hA, hB2, hB1, hH, hL, hr, pfp);
        const long long t_fp = libff::get_nsec_time();
        auto it = keys.find(key);
        if (it == keys.end()) {
            G1T alpha = pk.alpha_g1, beta1 = pk.beta_g1, delta1 = pk.delta_g1;
            G2T beta2 = pk.beta_g2, delta2 = pk.delta_g2;
            alpha.to_special();
            beta1.to_special();
            delta1.to_special();
            beta2.to_special();
            delta2.to_special();
            std::vector<uint32_t> idx(vec.indices.begin(), vec.indices.end());
            b200_groth16_pk_t h = 0;
            check(b200_groth16_pk_create(curve, &alpha, &beta1, &beta2, &delta1, &delta2, hA, hB2, hB1,
                                         idx.empty() ? nullptr : idx.data(), hH, hL, hr, cs.num_constraints(), ni, nv, m, &h),
                  "b200_groth16_pk_create");
            it = keys.emplace(key, h).first;
        }
        struct {
            G1T a;
            G2T b;
            G1T c;
        } out;
        static_assert(sizeof(out) == 2 * sizeof(G1T) + sizeof(G2T), "proof points are laid out back to back");
        double st[7] = {0, 0, 0, 0, 0, 0, 0};
        const long long t_key = libff::get_nsec_time();
        check(b200_groth16_prove(it->second, primary_input.data(), auxiliary_input.data(), &r, &s, &out, times ? st : nullptr),
              "b200_groth16_prove");
        if (times) {
            fprintf(stderr, "[b200 fused prover] cache lookups %.3f ms (bases %.3f, r1cs %.3f, key fingerprint %.3f, key %.3f), "
                            "b200_groth16_prove %.3f ms (upload %.3f, witness map %.3f, total %.3f)\n",
                    (t_key - t_in) * 1e-6, (t_bases - t_in) * 1e-6, (t_r1cs - t_bases) * 1e-6, (t_fp - t_r1cs) * 1e-6,
                    (t_key - t_fp) * 1e-6, (libff::get_nsec_time() - t_key) * 1e-6, st[0], st[1], st[6]);
        }
        return libsnark::r1cs_gg_ppzksnark_proof<ppT>(G1T(out.a), G2T(out.b), G1T(out.c));
    }
    // B200_OVERLAY_TRACE: the same proof stage by stage, one MSM after the other, with a time per stage on stderr.

    // (1 | primary | auxiliary): const_padded_assignment of r1cs_gg_ppzksnark.tcc:525-529
    static std::vector<FrT> w;
    w.resize(nv + 1);
    w[0] = FrT::one();
    std::copy(primary_input.begin(), primary_input.end(), w.begin() + 1);
    std::copy(auxiliary_input.begin(), auxiliary_input.end(), w.begin() + 1 + ni);

    static device_buffer d_w, d_h, d_g;
    d_w.reserve((nv + 1) * sizeof(FrT));
    d_h.reserve((m + 1) * sizeof(FrT));
    lap("assemble witness");
    check(b200_memcpy_h2d(d_w.p, w.data(), (nv + 1) * sizeof(FrT)), "b200_memcpy_h2d (witness)");
    lap("witness H2D");

    // r1cs_to_qap_witness_map with d1 = d2 = d3 = 0 (:497): coefficients_for_H stays on the device
    const b200_r1cs_t hr = r1cs_cache::get().lookup(cs);
    lap("R1CS lookup");
    check(b200_witness_map_r1cs_dev(hr, d_w.p, m, d_h.p), "b200_witness_map_r1cs_dev");
    lap("witness map (device)");

    bases_cache &bc = bases_cache::get();
    bc.begin_group();
    const FrT *dw = static_cast<const FrT *>(d_w.p);
    // A (:531-541): A_query[0 .. nv] x w
    G1T evaluation_At = G1T::zero();
    {
        const b200_bases_t h = bc.lookup(curve, B200_G1, pk.A_query.data(), sizeof(G1T), pk.A_query.size(), sizeof(G1T));
        lap("A lookup");
        check(b200_msm_resident(h, 0, nv + 1, dw, 1, &evaluation_At), "b200_msm_resident (A)");
        lap("A msm");
    }
    // B (:545-555): sparse knowledge-commitment query, entries with index in [0, nv]
    KC evaluation_Bt(G2T::zero(), G1T::zero());
    {
        const auto &vec = pk.B_query;
        const size_t cnt = std::lower_bound(vec.indices.begin(), vec.indices.end(), nv + 1) - vec.indices.begin();
        if (cnt) {
            const unsigned char *vals = reinterpret_cast<const unsigned char *>(vec.values.data());
            const size_t off2 = reinterpret_cast<const unsigned char *>(&vec.values[0].h) - vals;
            const b200_bases_t hg = bc.lookup(curve, B200_G2, vals, sizeof(KC), vec.values.size(), sizeof(G2T));
            const b200_bases_t hh = bc.lookup(curve, B200_G1, vals + off2, sizeof(KC), vec.values.size(), sizeof(G1T));
            const index_cache::entry &ix = index_cache::get().lookup(vec.indices);
            const void *sc = dw;
            if (!ix.dense) {
                d_g.reserve(cnt * sizeof(FrT));
                check(b200_gather_dev(curve, dw, static_cast<const uint32_t *>(ix.d_idx), cnt, d_g.p), "b200_gather_dev");
                sc = d_g.p;
            }
            lap("B lookup + gather");
            check(b200_msm_resident(hg, 0, cnt, sc, 1, &evaluation_Bt.g), "b200_msm_resident (B.g)");
            lap("B.g msm (G2)");
            check(b200_msm_resident(hh, 0, cnt, sc, 1, &evaluation_Bt.h), "b200_msm_resident (B.h)");
            lap("B.h msm (G1)");
        }
    }
    // H (:561-567): H_query[0 .. m-2] x coefficients_for_H
    G1T evaluation_Ht = G1T::zero();
    {
        const b200_bases_t h = bc.lookup(curve, B200_G1, pk.H_query.data(), sizeof(G1T), pk.H_query.size(), sizeof(G1T));
        check(b200_msm_resident(h, 0, m - 1, d_h.p, 1, &evaluation_Ht), "b200_msm_resident (H)");
        lap("H lookup + msm");
    }
    // L (:570-580): L_query x w[ni+1 .. nv]
    G1T evaluation_Lt = G1T::zero();
    {
        const b200_bases_t h = bc.lookup(curve, B200_G1, pk.L_query.data(), sizeof(G1T), pk.L_query.size(), sizeof(G1T));
        check(b200_msm_resident(h, 0, pk.L_query.size(), dw + ni + 1, 1, &evaluation_Lt), "b200_msm_resident (L)");
        lap("L lookup + msm");
    }
    // :583-595, unchanged
    const G1T g1_A = pk.alpha_g1 + evaluation_At + r * pk.delta_g1;
    const G1T g1_B = pk.beta_g1 + evaluation_Bt.h + s * pk.delta_g1;
    const G2T g2_B = pk.beta_g2 + evaluation_Bt.g + s * pk.delta_g2;
    const G1T g1_C = evaluation_Ht + evaluation_Lt + s * g1_A + r * g1_B - (r * s) * pk.delta_g1;
    lap("proof assembly");
    return libsnark::r1cs_gg_ppzksnark_proof<ppT>(G1T(g1_A), G2T(g2_B), G1T(g1_C));
}

// Drop-in for r1cs_gg_ppzksnark_prover(pk, primary, auxiliary): r, s drawn as the reference draws them (:499-500).
template<typename ppT>
libsnark::r1cs_gg_ppzksnark_proof<ppT> r1cs_gg_ppzksnark_prover_fused(
    const libsnark::r1cs_gg_ppzksnark_proving_key<ppT> &pk,
    const libsnark::r1cs_gg_ppzksnark_primary_input<ppT> &primary_input,
    const libsnark::r1cs_gg_ppzksnark_auxiliary_input<ppT> &auxiliary_input, bool force_pow_2_domain = false)
{
    const libff::Fr<ppT> r = libff::Fr<ppT>::random_element(), s = libff::Fr<ppT>::random_element();
    return r1cs_gg_ppzksnark_prover_fused<ppT>(pk, primary_input, auxiliary_input, r, s, force_pow_2_domain);
}

} // namespace libsnark_b200

// ---------------------------------------------------------------------------------------------------
// explicit specialisations
// ---------------------------------------------------------------------------------------------------
#define B200_OVERLAY_MSM(G, FR)                                                                                        \
    namespace libff                                                                                                    \
    {                                                                                                                  \
    template<>                                                                                                         \
    inline G multi_exp<G, FR, multi_exp_method_BDLO12_signed, multi_exp_base_form_special>(                            \
        std::vector<G>::const_iterator vec_start, std::vector<G>::const_iterator vec_end,                              \
        std::vector<FR>::const_iterator scalar_start, std::vector<FR>::const_iterator scalar_end, const size_t chunks) \
    {                                                                                                                  \
        (void)chunks;                                                                                                  \
        (void)scalar_end;                                                                                              \
        const size_t n = vec_end - vec_start;                                                                          \
        return libsnark_b200::msm_contiguous<G, FR>(n ? &*vec_start : nullptr, n, n ? &*scalar_start : nullptr);       \
    }                                                                                                                  \
    template<>                                                                                                         \
    inline G multi_exp_filter_one_zero<G, FR, multi_exp_method_BDLO12_signed, multi_exp_base_form_special>(            \
        std::vector<G>::const_iterator vec_start, std::vector<G>::const_iterator vec_end,                              \
        std::vector<FR>::const_iterator scalar_start, std::vector<FR>::const_iterator scalar_end, const size_t chunks) \
    {                                                                                                                  \
        (void)chunks;                                                                                                  \
        (void)scalar_end;                                                                                              \
        const size_t n = vec_end - vec_start;                                                                          \
        if (n) {                                                                                                       \
            libsnark_b200::print_filter_stats<FR>(&*scalar_start, n);                                                  \
        }                                                                                                              \
        return libsnark_b200::msm_contiguous<G, FR>(n ? &*vec_start : nullptr, n, n ? &*scalar_start : nullptr);       \
    }                                                                                                                  \
    }

// key generation: batch_exp (the 4-argument overload forwards to this one, multiexp.tcc:879-887),
// batch_exp_with_coeff. The results come back in special form, which every caller accepts (group equality is
// projective) and which r1cs_gg_ppzksnark_generator asks for anyway (batch_to_special, .tcc:353-386).
#define B200_OVERLAY_BATCH_EXP(G, FR)                                                                                  \
    namespace libff                                                                                                    \
    {                                                                                                                  \
    template<>                                                                                                         \
    inline std::vector<G> batch_exp<G, FR>(const size_t, const size_t, const window_table<G> &table,                  \
                                           const std::vector<FR> &v, size_t num_entries)                              \
    {                                                                                                                  \
        return libsnark_b200::batch_exp_special<G, FR>(table, nullptr, v.data(), num_entries);                         \
    }                                                                                                                  \
    template<>                                                                                                         \
    inline std::vector<G> batch_exp_with_coeff<G, FR>(const size_t, const size_t, const window_table<G> &table,       \
                                                      const FR &coeff, const std::vector<FR> &v)                      \
    {                                                                                                                  \
        return libsnark_b200::batch_exp_special<G, FR>(table, &coeff, v.data(), v.size());                             \
    }                                                                                                                  \
    }

#define B200_OVERLAY_KC(G2T, G1T, FR)                                                                                  \
    namespace libsnark                                                                                                 \
    {                                                                                                                  \
    template<>                                                                                                         \
    inline knowledge_commitment_vector<G2T, G1T> kc_batch_exp<G2T, G1T, FR>(                                           \
        const size_t, const size_t, const size_t, const libff::window_table<G2T> &T1_table,                            \
        const libff::window_table<G1T> &T2_table, const FR &T1_coeff, const FR &T2_coeff, const std::vector<FR> &v,   \
        const size_t, bool)                                                                                            \
    {                                                                                                                  \
        return libsnark_b200::kc_batch_exp_special<G2T, G1T, FR>(T1_table, T2_table, T1_coeff, T2_coeff, v);          \
    }                                                                                                                  \
    template<>                                                                                                         \
    inline knowledge_commitment<G2T, G1T> kc_multi_exp_with_mixed_addition<                                            \
        G2T, G1T, FR, libff::multi_exp_method_BDLO12_signed, libff::multi_exp_base_form_special>(                      \
        const knowledge_commitment_vector<G2T, G1T> &vec, const size_t min_idx, const size_t max_idx,                  \
        std::vector<FR>::const_iterator scalar_start, std::vector<FR>::const_iterator scalar_end, const size_t chunks) \
    {                                                                                                                  \
        (void)chunks;                                                                                                  \
        return libsnark_b200::kc_msm<G2T, G1T, FR>(vec, min_idx, max_idx, scalar_start, scalar_end);                   \
    }                                                                                                                  \
    }

#define B200_OVERLAY_DOMAIN(DOMAIN, FR)                                                                                \
    template<> inline void DOMAIN<FR>::FFT(std::vector<FR> &a)                                                         \
    {                                                                                                                  \
        libsnark_b200::ntt<FR>(a, this->m, B200_FFT, nullptr);                                                         \
    }                                                                                                                  \
    template<> inline void DOMAIN<FR>::iFFT(std::vector<FR> &a)                                                        \
    {                                                                                                                  \
        libsnark_b200::ntt<FR>(a, this->m, B200_IFFT, nullptr);                                                        \
    }                                                                                                                  \
    template<> inline void DOMAIN<FR>::cosetFFT(std::vector<FR> &a, const FR &g)                                       \
    {                                                                                                                  \
        libsnark_b200::ntt<FR>(a, this->m, B200_COSET_FFT, &g);                                                        \
    }                                                                                                                  \
    template<> inline void DOMAIN<FR>::icosetFFT(std::vector<FR> &a, const FR &g)                                      \
    {                                                                                                                  \
        libsnark_b200::ntt<FR>(a, this->m, B200_ICOSET_FFT, &g);                                                       \
    }                                                                                                                  \
    template<> inline void DOMAIN<FR>::divide_by_Z_on_coset(std::vector<FR> &P)                                        \
    {                                                                                                                  \
        libsnark_b200::check(                                                                                          \
            b200_divide_by_Z_on_coset(libsnark_b200::curve_of<FR>::curve, P.data(), this->m),                          \
            "b200_divide_by_Z_on_coset");                                                                              \
    }
// basic_radix2_domain (m = 2^k, k <= s), extended_radix2_domain (m = 2^(s+1), extended_radix2_domain.tcc:64-236) and
// step_radix2_domain (m = 2^k + 2^r, step_radix2_domain.tcc:65-276): the library tells them apart by m. A host that
// lowers FieldT::s / FieldT::root_of_unity must tell the library (b200_set_two_adicity).
#define B200_OVERLAY_NTT(FR)                                                                                           \
    namespace libfqfft                                                                                                 \
    {                                                                                                                  \
    B200_OVERLAY_DOMAIN(basic_radix2_domain, FR)                                                                       \
    B200_OVERLAY_DOMAIN(extended_radix2_domain, FR)                                                                    \
    B200_OVERLAY_DOMAIN(step_radix2_domain, FR)                                                                        \
    }

// The prover entry point itself, for its default instantiation (BDLO12_signed / special, r1cs_gg_ppzksnark.hpp:418-426):
// r1cs_gg_ppzksnark_prover(pk, primary, auxiliary[, force_pow_2_domain]) in unmodified caller code becomes the
// device-resident prover above — the key's queries and the CSR are uploaded and cached on first use, later proofs run
// the witness map and the five MSMs on the device from the witness alone. r and s are drawn as the reference draws
// them (r1cs_gg_ppzksnark.tcc:499-500). Define B200_OVERLAY_KEEP_REFERENCE_PROVER to keep the reference's body (its
// operators are then still the specialisations below).
// r1cs_to_qap_witness_map itself (every prover's first step; r1cs_to_qap.hpp) for the three scalar fields
#define B200_OVERLAY_WITNESS_MAP(FR)                                                                                   \
    namespace libsnark                                                                                                 \
    {                                                                                                                  \
    template<>                                                                                                         \
    inline qap_witness<FR> r1cs_to_qap_witness_map<FR>(                                                                \
        const r1cs_constraint_system<FR> &cs, const r1cs_primary_input<FR> &primary_input,                             \
        const r1cs_auxiliary_input<FR> &auxiliary_input, const FR &d1, const FR &d2, const FR &d3,                     \
        bool force_pow_2_domain)                                                                                       \
    {                                                                                                                  \
        return libsnark_b200::r1cs_to_qap_witness_map_device<FR>(cs, primary_input, auxiliary_input, d1, d2, d3,       \
                                                                 force_pow_2_domain);                                  \
    }                                                                                                                  \
    }

// r1cs_to_sap_witness_map (r1cs_se_ppzksnark's first step; r1cs_to_sap.hpp)
#define B200_OVERLAY_SAP_WITNESS_MAP(FR)                                                                               \
    namespace libsnark                                                                                                 \
    {                                                                                                                  \
    template<>                                                                                                         \
    inline sap_witness<FR> r1cs_to_sap_witness_map<FR>(                                                                \
        const r1cs_constraint_system<FR> &cs, const r1cs_primary_input<FR> &primary_input,                             \
        const r1cs_auxiliary_input<FR> &auxiliary_input, const FR &d1, const FR &d2)                                   \
    {                                                                                                                  \
        return libsnark_b200::r1cs_to_sap_witness_map_device<FR>(cs, primary_input, auxiliary_input, d1, d2);          \
    }                                                                                                                  \
    }

#ifndef B200_OVERLAY_KEEP_REFERENCE_PROVER
#define B200_OVERLAY_PROVER(PP)                                                                                        \
    namespace libsnark                                                                                                 \
    {                                                                                                                  \
    template<>                                                                                                         \
    inline r1cs_gg_ppzksnark_proof<PP>                                                                                 \
    r1cs_gg_ppzksnark_prover<PP, libff::multi_exp_method_BDLO12_signed, libff::multi_exp_base_form_special>(           \
        const r1cs_gg_ppzksnark_proving_key<PP> &pk, const r1cs_gg_ppzksnark_primary_input<PP> &primary_input,         \
        const r1cs_gg_ppzksnark_auxiliary_input<PP> &auxiliary_input, bool force_pow_2_domain)                         \
    {                                                                                                                  \
        return libsnark_b200::r1cs_gg_ppzksnark_prover_fused<PP>(pk, primary_input, auxiliary_input, force_pow_2_domain); \
    }                                                                                                                  \
    }
#else
#define B200_OVERLAY_PROVER(PP)
#endif

B200_OVERLAY_MSM(libff::alt_bn128_G1, libff::alt_bn128_Fr)
B200_OVERLAY_MSM(libff::alt_bn128_G2, libff::alt_bn128_Fr)
B200_OVERLAY_BATCH_EXP(libff::alt_bn128_G1, libff::alt_bn128_Fr)
B200_OVERLAY_BATCH_EXP(libff::alt_bn128_G2, libff::alt_bn128_Fr)
B200_OVERLAY_KC(libff::alt_bn128_G2, libff::alt_bn128_G1, libff::alt_bn128_Fr)
B200_OVERLAY_KC(libff::alt_bn128_G1, libff::alt_bn128_G1, libff::alt_bn128_Fr) // r1cs_ppzksnark's A and C queries (r1cs_ppzksnark.tcc:570-612)
B200_OVERLAY_NTT(libff::alt_bn128_Fr)
B200_OVERLAY_WITNESS_MAP(libff::alt_bn128_Fr)
B200_OVERLAY_SAP_WITNESS_MAP(libff::alt_bn128_Fr)
B200_OVERLAY_PROVER(libff::alt_bn128_pp)

B200_OVERLAY_MSM(libff::bls12_377_G1, libff::bls12_377_Fr)
B200_OVERLAY_MSM(libff::bls12_377_G2, libff::bls12_377_Fr)
B200_OVERLAY_BATCH_EXP(libff::bls12_377_G1, libff::bls12_377_Fr)
B200_OVERLAY_BATCH_EXP(libff::bls12_377_G2, libff::bls12_377_Fr)
B200_OVERLAY_KC(libff::bls12_377_G2, libff::bls12_377_G1, libff::bls12_377_Fr)
B200_OVERLAY_KC(libff::bls12_377_G1, libff::bls12_377_G1, libff::bls12_377_Fr) // r1cs_ppzksnark's A and C queries (r1cs_ppzksnark.tcc:570-612)
B200_OVERLAY_NTT(libff::bls12_377_Fr)
B200_OVERLAY_WITNESS_MAP(libff::bls12_377_Fr)
B200_OVERLAY_SAP_WITNESS_MAP(libff::bls12_377_Fr)
B200_OVERLAY_PROVER(libff::bls12_377_pp)

B200_OVERLAY_MSM(libff::bw6_761_G1, libff::bw6_761_Fr)
B200_OVERLAY_MSM(libff::bw6_761_G2, libff::bw6_761_Fr)
B200_OVERLAY_BATCH_EXP(libff::bw6_761_G1, libff::bw6_761_Fr)
B200_OVERLAY_BATCH_EXP(libff::bw6_761_G2, libff::bw6_761_Fr)
B200_OVERLAY_KC(libff::bw6_761_G2, libff::bw6_761_G1, libff::bw6_761_Fr)
B200_OVERLAY_KC(libff::bw6_761_G1, libff::bw6_761_G1, libff::bw6_761_Fr) // r1cs_ppzksnark's A and C queries (r1cs_ppzksnark.tcc:570-612)
B200_OVERLAY_NTT(libff::bw6_761_Fr)
B200_OVERLAY_WITNESS_MAP(libff::bw6_761_Fr)
B200_OVERLAY_SAP_WITNESS_MAP(libff::bw6_761_Fr)
B200_OVERLAY_PROVER(libff::bw6_761_pp)

#endif // LIBSNARK_B200_OVERLAY_HPP_
