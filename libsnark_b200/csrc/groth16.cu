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
#include <chrono>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

#include "curves.cuh"
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
};

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

template<int CURVE>
int assemble(const ProvingKey &pk, const void *r_, const void *s_, const void *At, const void *Bg2, const void *Bg1,
             const void *Ht, const void *Lt, void *proof_out)
{
    typedef typename group_traits<CURVE, GROUP_G1>::Fr Fr;
    typedef typename group_traits<CURVE, GROUP_G1>::F F1;
    typedef typename group_traits<CURVE, GROUP_G2>::F F2;
    Fr r, s;
    memcpy(&r, r_, sizeof(Fr));
    memcpy(&s, s_, sizeof(Fr));
    const XYZZ<F1> delta1 = load_ref<F1>(pk.delta_g1.data());
    const XYZZ<F2> delta2 = load_ref<F2>(pk.delta_g2.data());

    XYZZ<F1> gA = load_ref<F1>(pk.alpha_g1.data());
    add(gA, load_ref<F1>(At));
    add(gA, scalar_mul<Fr, F1>(r, delta1));

    XYZZ<F1> gB1 = load_ref<F1>(pk.beta_g1.data());
    add(gB1, load_ref<F1>(Bg1));
    add(gB1, scalar_mul<Fr, F1>(s, delta1));

    XYZZ<F2> gB2 = load_ref<F2>(pk.beta_g2.data());
    add(gB2, load_ref<F2>(Bg2));
    add(gB2, scalar_mul<Fr, F2>(s, delta2));

    XYZZ<F1> gC = load_ref<F1>(Ht);
    add(gC, load_ref<F1>(Lt));
    add(gC, scalar_mul<Fr, F1>(s, gA));
    add(gC, scalar_mul<Fr, F1>(r, gB1));
    add(gC, neg(scalar_mul<Fr, F1>(r * s, delta1)));

    const msm::RefPoint<F1> oA = msm::normalise(gA), oC = msm::normalise(gC);
    const msm::RefPoint<F2> oB = msm::normalise(gB2);
    uint8_t *o = static_cast<uint8_t *>(proof_out);
    memcpy(o, &oA, sizeof(oA));
    memcpy(o + sizeof(oA), &oB, sizeof(oB));
    memcpy(o + sizeof(oA) + sizeof(oB), &oC, sizeof(oC));
    return B200_OK;
}

} // namespace

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
    int rc = B200_OK;
    if ((rc = b200_dev_alloc((nv + 1) * fr, &pk.d_w)) || (rc = b200_dev_alloc((m + 1) * fr, &pk.d_h))) {
        return rc;
    }
    if (!dense) {
        if ((rc = b200_dev_alloc(nB * 4, &pk.d_B_idx)) || (rc = b200_memcpy_h2d(pk.d_B_idx, B_indices, nB * 4)) ||
            (rc = b200_dev_alloc(nB * fr, &pk.d_g))) {
            return rc;
        }
    }
    B200_CUDA_CHECK(cudaHostAlloc(&pk.h_w, (nv + 1) * fr, cudaHostAllocDefault));
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
    ProvingKey &pk = it->second;
    b200_dev_free(pk.d_w);
    b200_dev_free(pk.d_h);
    if (pk.d_g) {
        b200_dev_free(pk.d_g);
    }
    if (pk.d_B_idx) {
        b200_dev_free(pk.d_B_idx);
    }
    cudaFreeHost(pk.h_w);
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
    const size_t fr = b200_sizeof_fr(pk.curve);
    const size_t g1 = b200_sizeof_group(pk.curve, B200_G1), g2 = b200_sizeof_group(pk.curve, B200_G2);
    double t0 = now_ms(), t = t0, t1;
    double st[7] = {0, 0, 0, 0, 0, 0, 0};

    // (1 | primary | auxiliary), r1cs_gg_ppzksnark.tcc:525-529
    uint8_t *w = static_cast<uint8_t *>(pk.h_w);
    {
        uint32_t one[32];
        for (int i = 0; i < (int)(fr / 4); ++i) {
            one[i] = pk.curve == B200_ALT_BN128   ? alt_bn128_fr::params::one(i)
                     : pk.curve == B200_BLS12_377 ? bls12_377_fr::params::one(i)
                                                  : bw6_761_fr::params::one(i);
        }
        memcpy(w, one, fr);
        memcpy(w + fr, primary_input, pk.ni * fr);
        memcpy(w + (1 + pk.ni) * fr, auxiliary_input, (pk.nv - pk.ni) * fr);
    }
    int rc;
    if ((rc = b200_memcpy_h2d(pk.d_w, w, (pk.nv + 1) * fr))) {
        return rc;
    }
    t1 = now_ms(), st[0] = t1 - t, t = t1;
    if ((rc = b200_witness_map_r1cs_dev(pk.r1cs, pk.d_w, pk.m, pk.d_h))) {
        return rc;
    }
    t1 = now_ms(), st[1] = t1 - t, t = t1;

    std::vector<uint8_t> At(g1), Bg1(g1), Ht(g1), Lt(g1), Bg2(g2);
    if ((rc = b200_msm_resident(pk.A, 0, pk.nv + 1, pk.d_w, 1, At.data()))) {
        return rc;
    }
    t1 = now_ms(), st[2] = t1 - t, t = t1;
    const void *scB = pk.d_w;
    if (pk.d_B_idx) {
        if ((rc = b200_gather_dev(pk.curve, pk.d_w, static_cast<const uint32_t *>(pk.d_B_idx), pk.nB, pk.d_g))) {
            return rc;
        }
        scB = pk.d_g;
    }
    if ((rc = b200_msm_resident(pk.B_g2, 0, pk.nB, scB, 1, Bg2.data())) ||
        (rc = b200_msm_resident(pk.B_g1, 0, pk.nB, scB, 1, Bg1.data()))) {
        return rc;
    }
    t1 = now_ms(), st[3] = t1 - t, t = t1;
    if ((rc = b200_msm_resident(pk.H, 0, pk.m - 1, pk.d_h, 1, Ht.data()))) {
        return rc;
    }
    t1 = now_ms(), st[4] = t1 - t, t = t1;
    if ((rc = b200_msm_resident(pk.L, 0, pk.nv - pk.ni, static_cast<const uint8_t *>(pk.d_w) + (pk.ni + 1) * fr, 1,
                                Lt.data()))) {
        return rc;
    }
    t1 = now_ms(), st[5] = t1 - t, t = t1;

    switch (pk.curve) {
    case B200_ALT_BN128:
        rc = assemble<CURVE_ALT_BN128>(pk, r, s, At.data(), Bg2.data(), Bg1.data(), Ht.data(), Lt.data(), proof_out);
        break;
    case B200_BLS12_377:
        rc = assemble<CURVE_BLS12_377>(pk, r, s, At.data(), Bg2.data(), Bg1.data(), Ht.data(), Lt.data(), proof_out);
        break;
    default:
        rc = assemble<CURVE_BW6_761>(pk, r, s, At.data(), Bg2.data(), Bg1.data(), Ht.data(), Lt.data(), proof_out);
        break;
    }
    st[6] = now_ms() - t0;
    if (stage_ms) {
        // witness upload | witness map | A | B | H | L | total (the last includes the host assembly)
        memcpy(stage_ms, st, sizeof(st));
    }
    return rc;
}

} // extern "C"
