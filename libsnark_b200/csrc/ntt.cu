// ntt.cu — host driver and C-ABI of the NTT operators (ntt.cuh): cached per-size domains
// (twiddle and coset tables in HBM), pass planning, the evaluation_domain entry points and
// the fused witness-map chain of r1cs_to_qap_witness_map (r1cs_to_qap.tcc:252-330).
#include <algorithm>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "ntt.cuh"
#include "runtime.hpp"

using namespace b200;

namespace
{

constexpr int MAX_DEG = ntt::NTT_MAX_DEG;

inline unsigned grid_for(uint64_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }

template<class Fr> struct Domain {
    int lgn = 0;
    rt::DevBuf twiddles, consts, pows, g_dev;
    rt::DevBuf gpow, ginv_minv, gpow_minv; // coset tables for the generator in g_host
    uint32_t g_host[Fr::N];
    bool has_g = false;
    const Fr *c(int i) { return consts.template as<Fr>() + i; }
    const Fr *pows_w() { return pows.template as<Fr>(); }
    const Fr *pows_g() { return pows.template as<Fr>() + 64; }
    const Fr *pows_gi() { return pows.template as<Fr>() + 128; }
};

template<class Fr> std::map<int, std::unique_ptr<Domain<Fr>>> &domains()
{
    static std::map<int, std::unique_ptr<Domain<Fr>>> m;
    return m;
}

// (Re)computes the domain constants for generator g (host pointer to Fr::N u32) and drops
// stale coset tables.
template<class Fr> int domain_set_generator(Domain<Fr> &d, const uint32_t *g)
{
    cudaStream_t s = rt::stream();
    if (d.consts.reserve(8 * sizeof(Fr)) || d.pows.reserve(192 * sizeof(Fr)) || d.g_dev.reserve(sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    memcpy(d.g_host, g, sizeof(d.g_host));
    B200_CUDA_CHECK(cudaMemcpyAsync(d.g_dev.p, d.g_host, sizeof(Fr), cudaMemcpyHostToDevice, s));
    ntt::k_domain_consts<Fr><<<1, 32, 0, s>>>(d.lgn, d.g_dev.template as<Fr>(), d.consts.template as<Fr>(), d.pows.template as<Fr>(),
                                              d.pows.template as<Fr>() + 64, d.pows.template as<Fr>() + 128);
    B200_LAUNCH_CHECK();
    d.gpow.release();
    d.ginv_minv.release();
    d.gpow_minv.release();
    d.has_g = true;
    return B200_OK;
}

template<class Fr> int get_domain(int lgn, const uint32_t *g, Domain<Fr> **out)
{
    auto &m = domains<Fr>();
    auto it = m.find(lgn);
    if (it == m.end()) {
        std::unique_ptr<Domain<Fr>> d(new Domain<Fr>());
        d->lgn = lgn;
        uint32_t gen[Fr::N];
        for (int i = 0; i < Fr::N; ++i) {
            gen[i] = Fr::params::generator(i);
        }
        int rc = domain_set_generator(*d, g ? g : gen);
        if (rc) {
            return rc;
        }
        const uint64_t half = 1ull << (lgn - 1);
        if (d->twiddles.reserve(half * sizeof(Fr))) {
            return B200_ERR_CUDA;
        }
        ntt::k_pow_table<Fr><<<grid_for(half, 256), 256, 0, rt::stream()>>>(
            half, d->pows_w(), lgn - 1, nullptr, d->twiddles.template as<Fr>());
        B200_LAUNCH_CHECK();
        it = m.emplace(lgn, std::move(d)).first;
    } else if (g && memcmp(g, it->second->g_host, sizeof(uint32_t) * Fr::N) != 0) {
        int rc = domain_set_generator(*it->second, g);
        if (rc) {
            return rc;
        }
    }
    *out = it->second.get();
    return B200_OK;
}

enum CosetTable { T_GPOW, T_GINV_MINV, T_GPOW_MINV };
template<class Fr> int coset_table(Domain<Fr> &d, CosetTable which, const Fr **out)
{
    rt::DevBuf &buf = which == T_GPOW ? d.gpow : which == T_GINV_MINV ? d.ginv_minv : d.gpow_minv;
    const uint64_t n = 1ull << d.lgn;
    if (!buf.p) {
        if (buf.reserve(n * sizeof(Fr))) {
            return B200_ERR_CUDA;
        }
        const Fr *pows = which == T_GINV_MINV ? d.pows_gi() : d.pows_g();
        const Fr *scale = which == T_GPOW ? nullptr : d.c(1);
        ntt::k_pow_table<Fr><<<grid_for(n, 256), 256, 0, rt::stream()>>>(n, pows, d.lgn, scale, buf.template as<Fr>());
        B200_LAUNCH_CHECK();
    }
    *out = buf.template as<Fr>();
    return B200_OK;
}

// Runs one transform. Passes ping-pong between `in` and `tmp`; the last pass writes `out`.
// out may equal in (then an odd pass count ends with one device copy from tmp).
template<class Fr>
int run_transform(Domain<Fr> &d, int inverse, Fr *in, Fr *tmp, Fr *out, const Fr *pre_table, const Fr *post_table,
                  const Fr *post_const)
{
    const int lgn = d.lgn;
    ntt::PassLaunch plan[8];
    const int P = ntt::plan_transform(lgn, inverse, Fr::N, plan);
    cudaStream_t s = rt::stream();
    Fr *cur = in;
    for (int i = 0; i < P; ++i) {
        const ntt::PassParams &pp = plan[i].pp;
        const bool last = (i == P - 1);
        Fr *dst = last ? out : (cur == in ? tmp : in);
        bool copy_back = false;
        if (dst == cur) { // in-place request with an odd number of passes
            dst = tmp;
            copy_back = true;
        }
        const size_t smem = plan[i].smem;
        static size_t smem_attr = 0;
        if (smem > smem_attr) {
            B200_CUDA_CHECK(cudaFuncSetAttribute(ntt::k_ntt_pass<Fr>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)(smem > 48 * 1024 ? smem : 48 * 1024)));
            smem_attr = smem;
        }
        ntt::k_ntt_pass<Fr><<<plan[i].blocks, plan[i].threads, smem, s>>>(pp, cur, dst, d.twiddles.template as<Fr>(),
                                                                         i == 0 ? pre_table : nullptr, last ? post_table : nullptr,
                                                                         last ? post_const : nullptr);
        B200_LAUNCH_CHECK();
        if (copy_back) {
            B200_CUDA_CHECK(cudaMemcpyAsync(out, tmp, sizeof(Fr) << lgn, cudaMemcpyDeviceToDevice, s));
        }
        cur = dst;
    }
    return B200_OK;
}

int log2_exact(size_t m)
{
    if (m < 2 || (m & (m - 1))) {
        return -1;
    }
    int k = 0;
    while ((1ull << k) < m) {
        ++k;
    }
    return k;
}

rt::DevBuf g_bufA, g_bufB, g_bufC, g_bufS, g_bufH;
std::mutex g_mu;

// The 2-adicity the domain factory works with: the field's own (FieldT::s), or the lower value a host set through
// b200_set_two_adicity. libff keeps FieldT::s and FieldT::root_of_unity as mutable static parameters
// (alt_bn128_init.cpp:53-58 and siblings), and get_evaluation_domain.tcc:37-58 reads them: sizes up to 2^s are
// basic_radix2_domain, 2^(s+1) is extended_radix2_domain, 2^k + 2^r needs k + 1 <= s.
template<class Fr> int &two_adicity()
{
    static int s = Fr::params::S;
    return s;
}

template<class Fr> int ext_try_op(int op, Fr *data, Fr *tmp, Fr *out, size_t m, const void *g, bool *handled);
struct StepShape;
template<class Fr> bool step_shape(size_t m, StepShape *sh);
template<class Fr> int step_op_dev(int op, Fr *a, Fr *tmp, const StepShape &sh, const void *g);
template<class Fr> int step_try_op(int op, Fr *data, Fr *tmp, Fr *out, size_t m, const void *g, bool *handled);

template<class Fr> int ntt_op_dev(int op, Fr *data, Fr *tmp, Fr *out, size_t m, const void *g)
{
    const int lgn = log2_exact(m);
    if (lgn < 0) {
        bool handled = false;
        const int rc = step_try_op<Fr>(op, data, tmp, out, m, g, &handled);
        if (handled) {
            return rc;
        }
    }
    if (lgn == two_adicity<Fr>() + 1) {
        bool handled = false;
        const int rc = ext_try_op<Fr>(op, data, tmp, out, m, g, &handled);
        if (handled) {
            return rc;
        }
    }
    if (lgn < 0 || lgn > two_adicity<Fr>()) {
        // get_evaluation_domain / basic_radix2_domain ctor (basic_radix2_domain.tcc:25-52)
        return rt::set_error(B200_ERR_DOMAIN, "basic_radix2_domain: size %zu is not a power of two in [2, 2^%d]", m,
                             two_adicity<Fr>());
    }
    const bool coset = (op == B200_COSET_FFT || op == B200_ICOSET_FFT);
    if (coset && !g) {
        return rt::set_error(B200_ERR_ARG, "coset transform needs a generator");
    }
    Domain<Fr> *d = nullptr;
    int rc = get_domain<Fr>(lgn, coset ? static_cast<const uint32_t *>(g) : nullptr, &d);
    if (rc) {
        return rc;
    }
    const Fr *tbl = nullptr;
    switch (op) {
    case B200_FFT: return run_transform<Fr>(*d, 0, data, tmp, out, nullptr, nullptr, nullptr);
    case B200_IFFT: return run_transform<Fr>(*d, 1, data, tmp, out, nullptr, nullptr, d->c(1));
    case B200_COSET_FFT:
        if ((rc = coset_table(*d, T_GPOW, &tbl))) {
            return rc;
        }
        return run_transform<Fr>(*d, 0, data, tmp, out, tbl, nullptr, nullptr);
    case B200_ICOSET_FFT:
        if ((rc = coset_table(*d, T_GINV_MINV, &tbl))) {
            return rc;
        }
        return run_transform<Fr>(*d, 1, data, tmp, out, nullptr, tbl, nullptr);
    }
    return rt::set_error(B200_ERR_ARG, "unknown NTT op %d", op);
}

template<class Fr> int ntt_host(int op, void *data, size_t m, const void *g)
{
    const size_t bytes = m * sizeof(Fr);
    if (g_bufA.reserve(bytes) || g_bufS.reserve(bytes)) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(g_bufA.p, data, bytes, cudaMemcpyHostToDevice, s));
    int rc = ntt_op_dev<Fr>(op, g_bufA.as<Fr>(), g_bufS.as<Fr>(), g_bufA.as<Fr>(), m, g);
    if (rc) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(data, g_bufA.p, bytes, cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

template<class Fr> int ntt_dev(int op, void *dev_data, size_t m, const void *g)
{
    if (g_bufS.reserve(m * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    int rc = ntt_op_dev<Fr>(op, static_cast<Fr *>(dev_data), g_bufS.as<Fr>(), static_cast<Fr *>(dev_data), m, g);
    if (rc) {
        return rc;
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

template<class Fr> int step_divide_by_z_host(void *data, size_t m, bool *handled);
template<class Fr> int step_witness_map_try(Fr *dA, Fr *dB, Fr *dC, size_t m, Fr *dH, bool *handled);
template<class Fr> int ext_divide_by_z_host(void *data, size_t m);
template<class Fr> int ext_witness_map_dev(Fr *dA, Fr *dB, Fr *dC, size_t m, Fr *dH);

template<class Fr> int divide_by_z(void *data, size_t m)
{
    const int lgn = log2_exact(m);
    if (lgn < 0) {
        bool handled = false;
        const int rc = step_divide_by_z_host<Fr>(data, m, &handled);
        if (handled) {
            return rc;
        }
    }
    if (lgn == two_adicity<Fr>() + 1) {
        return ext_divide_by_z_host<Fr>(data, m);
    }
    if (lgn < 0 || lgn > two_adicity<Fr>()) {
        return rt::set_error(B200_ERR_DOMAIN, "basic_radix2_domain: bad size %zu", m);
    }
    Domain<Fr> *d = nullptr;
    uint32_t gen[Fr::N];
    for (int i = 0; i < Fr::N; ++i) {
        gen[i] = Fr::params::generator(i);
    }
    int rc = get_domain<Fr>(lgn, gen, &d);
    if (rc) {
        return rc;
    }
    const size_t bytes = m * sizeof(Fr);
    if (g_bufA.reserve(bytes)) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(g_bufA.p, data, bytes, cudaMemcpyHostToDevice, s));
    ntt::k_scale<Fr><<<grid_for(m, 256), 256, 0, s>>>(m, g_bufA.as<Fr>(), d->c(4));
    B200_LAUNCH_CHECK();
    B200_CUDA_CHECK(cudaMemcpyAsync(data, g_bufA.p, bytes, cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

// The basic radix-2 domain of size m with g = FieldT::multiplicative_generator (r1cs_to_qap.tcc:273).
template<class Fr> int witness_domain(size_t m, Domain<Fr> **d)
{
    const int lgn = log2_exact(m);
    if (lgn < 0 || lgn > two_adicity<Fr>()) {
        return rt::set_error(B200_ERR_DOMAIN, "basic_radix2_domain: bad size %zu", m);
    }
    uint32_t gen[Fr::N];
    for (int i = 0; i < Fr::N; ++i) {
        gen[i] = Fr::params::generator(i);
    }
    if (int rc = get_domain<Fr>(lgn, gen, d)) {
        return rc;
    }
    return g_bufS.reserve(m * sizeof(Fr)) ? B200_ERR_CUDA : B200_OK;
}

// One of the three chains of r1cs_to_qap.tcc:252-290: iFFT then cosetFFT, in place on X (m elements).
// The 1/m of the inverse and the g^i of the coset are one table multiplication fused into the
// inverse's last pass. With an odd pass count the inverse lands in S and the forward transform
// brings it back: no device copies.
template<class Fr> int witness_chain(Domain<Fr> &d, Fr *X)
{
    const Fr *gpow_minv = nullptr;
    if (int rc = coset_table(d, T_GPOW_MINV, &gpow_minv)) {
        return rc;
    }
    Fr *S = g_bufS.as<Fr>();
    const bool odd = ((d.lgn + MAX_DEG - 1) / MAX_DEG) & 1;
    Fr *mid = odd ? S : X, *other = odd ? X : S;
    if (int rc = run_transform<Fr>(d, 1, X, S, mid, nullptr, gpow_minv, nullptr)) {
        return rc;
    }
    return run_transform<Fr>(d, 0, mid, other, X, nullptr, nullptr, nullptr);
}

// r1cs_to_qap.tcc:293-330: (A*B - C) / Z on the coset, icosetFFT, coefficients_for_H[m] = 0. dA clobbered.
template<class Fr> int witness_finish(Domain<Fr> &d, size_t m, Fr *dA, const Fr *dB, const Fr *dC, Fr *dH)
{
    const Fr *ginv_minv = nullptr;
    if (int rc = coset_table(d, T_GINV_MINV, &ginv_minv)) {
        return rc;
    }
    cudaStream_t s = rt::stream();
    ntt::k_pointwise_h<Fr><<<grid_for(m, 256), 256, 0, s>>>(m, dA, dB, dC, d.c(4), dA);
    B200_LAUNCH_CHECK();
    if (int rc = run_transform<Fr>(d, 1, dA, g_bufS.as<Fr>(), dH, nullptr, ginv_minv, nullptr)) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemsetAsync(dH + m, 0, sizeof(Fr), s));
    return B200_OK;
}

// r1cs_to_qap.tcc:252-330 with d1 = d2 = d3 = 0. dA, dB, dC clobbered; dH gets m + 1 elements.
template<class Fr> int witness_map_dev(Fr *dA, Fr *dB, Fr *dC, size_t m, Fr *dH)
{
    if (log2_exact(m) < 0) {
        bool handled = false;
        const int rc = step_witness_map_try<Fr>(dA, dB, dC, m, dH, &handled);
        if (handled) {
            return rc;
        }
    }
    if (log2_exact(m) == two_adicity<Fr>() + 1) {
        return ext_witness_map_dev<Fr>(dA, dB, dC, m, dH);
    }
    Domain<Fr> *d = nullptr;
    int rc = witness_domain<Fr>(m, &d);
    if (rc) {
        return rc;
    }
    Fr *vecs[3] = {dA, dB, dC};
    for (int v = 0; v < 3; ++v) {
        if ((rc = witness_chain<Fr>(*d, vecs[v]))) {
            return rc;
        }
    }
    return witness_finish<Fr>(*d, m, dA, dB, dC, dH);
}

template<class Fr> int witness_map_host(const void *aA, const void *aB, const void *aC, size_t m, void *out_H)
{
    const size_t bytes = m * sizeof(Fr);
    if (g_bufA.reserve(bytes) || g_bufB.reserve(bytes) || g_bufC.reserve(bytes) || g_bufH.reserve(bytes + sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(g_bufA.p, aA, bytes, cudaMemcpyHostToDevice, s));
    B200_CUDA_CHECK(cudaMemcpyAsync(g_bufB.p, aB, bytes, cudaMemcpyHostToDevice, s));
    B200_CUDA_CHECK(cudaMemcpyAsync(g_bufC.p, aC, bytes, cudaMemcpyHostToDevice, s));
    int rc = witness_map_dev<Fr>(g_bufA.as<Fr>(), g_bufB.as<Fr>(), g_bufC.as<Fr>(), m, g_bufH.as<Fr>());
    if (rc) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(out_H, g_bufH.p, bytes + sizeof(Fr), cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

// ---- step_radix2_domain ---------------------------------------------------------------------------
// m = big + small (big = 2^k > small = 2^r): what get_evaluation_domain returns for every size that is
// not a power of two (get_evaluation_domain.tcc:41-58; step_radix2_domain.tcc:20-62).
struct StepShape {
    size_t m = 0, big = 0, small = 0, compr = 0;
    int lg_big = 0, lg_small = 0;
};
template<class Fr> bool step_shape(size_t m, StepShape *sh)
{
    if (m < 3 || (m & (m - 1)) == 0) {
        return false;
    }
    int lg = 0;
    while ((1ull << (lg + 1)) <= m) {
        ++lg;
    }
    const size_t big = 1ull << lg, small = m - big;
    if (small & (small - 1)) {
        return false;
    }
    if (lg + 1 > two_adicity<Fr>()) {
        return false;
    }
    sh->m = m;
    sh->big = big;
    sh->small = small;
    sh->compr = big / small;
    sh->lg_big = lg;
    sh->lg_small = 0;
    while ((1ull << sh->lg_small) < small) {
        ++sh->lg_small;
    }
    return true;
}

using ntt::host_pow;

template<class Fr> struct StepDomain {
    StepShape sh;
    Domain<Fr> *d2 = nullptr, *dbig = nullptr, *dsmall = nullptr; // sizes 2 big, big, small (null when small == 1)
    rt::DevBuf consts; // [0] 1/2  [1] A  [2] B  [3] w2s  [4] Z1^-1  [5] 1/small (when small == 1: one)
    rt::DevBuf zinv, gpow, ginv;
    bool has_zinv = false;
    uint32_t g_host[Fr::N];
    bool has_g = false;
    const Fr *c(int i) { return consts.template as<Fr>() + i; }
};
template<class Fr> std::map<size_t, std::unique_ptr<StepDomain<Fr>>> &step_domains()
{
    static std::map<size_t, std::unique_ptr<StepDomain<Fr>>> m;
    return m;
}

template<class Fr> int get_step_domain(const StepShape &sh, StepDomain<Fr> **out)
{
    auto &mp = step_domains<Fr>();
    auto it = mp.find(sh.m);
    if (it == mp.end()) {
        std::unique_ptr<StepDomain<Fr>> d(new StepDomain<Fr>());
        d->sh = sh;
        int rc;
        if ((rc = get_domain<Fr>(sh.lg_big + 1, nullptr, &d->d2)) || (rc = get_domain<Fr>(sh.lg_big, nullptr, &d->dbig))) {
            return rc;
        }
        if (sh.lg_small >= 1 && (rc = get_domain<Fr>(sh.lg_small, nullptr, &d->dsmall))) {
            return rc;
        }
        // constants on the host (a handful of exponentiations; step_radix2_domain.tcc:252-276)
        Fr hc[6];
        ntt::step_domain_consts_host<Fr>(sh.lg_big, sh.small, hc);
        if (d->consts.reserve(sizeof(hc))) {
            return B200_ERR_CUDA;
        }
        B200_CUDA_CHECK(cudaMemcpyAsync(d->consts.p, hc, sizeof(hc), cudaMemcpyHostToDevice, rt::stream()));
        B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
        it = mp.emplace(sh.m, std::move(d)).first;
    }
    *out = it->second.get();
    return B200_OK;
}

// g^i and g^-i, i < m, for the coset transforms (tables rebuilt when the generator changes)
template<class Fr> int step_coset_tables(StepDomain<Fr> &d, const uint32_t *g)
{
    if (d.has_g && memcmp(g, d.g_host, sizeof(d.g_host)) == 0) {
        return B200_OK;
    }
    // the size-2big radix-2 domain holds g^(2^b) and g^-(2^b) for b <= lg_big: enough for indices < 2 big
    Domain<Fr> *d2 = nullptr;
    int rc = get_domain<Fr>(d.sh.lg_big + 1, g, &d2);
    if (rc) {
        return rc;
    }
    if (d.gpow.reserve(d.sh.m * sizeof(Fr)) || d.ginv.reserve(d.sh.m * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    ntt::k_pow_table<Fr><<<grid_for(d.sh.m, 256), 256, 0, s>>>(d.sh.m, d2->pows_g(), d.sh.lg_big + 1, nullptr, d.gpow.template as<Fr>());
    B200_LAUNCH_CHECK();
    ntt::k_pow_table<Fr><<<grid_for(d.sh.m, 256), 256, 0, s>>>(d.sh.m, d2->pows_gi(), d.sh.lg_big + 1, nullptr, d.ginv.template as<Fr>());
    B200_LAUNCH_CHECK();
    memcpy(d.g_host, g, sizeof(d.g_host));
    d.has_g = true;
    return B200_OK;
}

rt::DevBuf g_stepD;

// sum over the compr rows of v (compr x small, row-major) into row 0
template<class Fr> int fold_rows(Fr *v, const StepShape &sh)
{
    for (size_t rows = sh.compr; rows > 1; rows >>= 1) {
        const uint64_t half = (rows >> 1) * sh.small;
        ntt::k_fold_halve<Fr><<<grid_for(half, 256), 256, 0, rt::stream()>>>(half, v);
        B200_LAUNCH_CHECK();
    }
    return B200_OK;
}

// step_radix2_domain::FFT (:65-113) in place on a (m elements, device); tmp: >= big elements
template<class Fr> int step_fft(StepDomain<Fr> &d, Fr *a, Fr *tmp)
{
    const StepShape &sh = d.sh;
    if (g_stepD.reserve(sh.big * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    Fr *dd = g_stepD.as<Fr>();
    cudaStream_t s = rt::stream();
    ntt::k_step_fft_pre<Fr><<<grid_for(sh.big, 256), 256, 0, s>>>(sh.big, sh.small, d.d2->twiddles.template as<Fr>(), a, dd);
    B200_LAUNCH_CHECK();
    int rc;
    if ((rc = fold_rows<Fr>(dd, sh)) || (rc = run_transform<Fr>(*d.dbig, 0, a, tmp, a, nullptr, nullptr, nullptr))) {
        return rc;
    }
    if (d.dsmall) {
        return run_transform<Fr>(*d.dsmall, 0, dd, tmp, a + sh.big, nullptr, nullptr, nullptr);
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(a + sh.big, dd, sizeof(Fr), cudaMemcpyDeviceToDevice, s));
    return B200_OK;
}

// step_radix2_domain::iFFT (:115-174) in place
template<class Fr> int step_ifft(StepDomain<Fr> &d, Fr *a, Fr *tmp)
{
    const StepShape &sh = d.sh;
    if (g_stepD.reserve(sh.big * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    Fr *dd = g_stepD.as<Fr>();
    cudaStream_t s = rt::stream();
    int rc;
    if ((rc = run_transform<Fr>(*d.dbig, 1, a, tmp, a, nullptr, nullptr, d.dbig->c(1)))) {
        return rc;
    }
    if (d.dsmall && (rc = run_transform<Fr>(*d.dsmall, 1, a + sh.big, tmp, a + sh.big, nullptr, nullptr, d.dsmall->c(1)))) {
        return rc;
    }
    const Fr *T2 = d.d2->twiddles.template as<Fr>();
    ntt::k_step_ifft_tmp<Fr><<<grid_for(sh.big, 256), 256, 0, s>>>(sh.big, sh.small, T2, a, dd);
    B200_LAUNCH_CHECK();
    if ((rc = fold_rows<Fr>(dd, sh))) {
        return rc;
    }
    ntt::k_step_ifft_post<Fr><<<grid_for(sh.small, 256), 256, 0, s>>>(sh.big, sh.small, T2, dd, d.c(0), a);
    B200_LAUNCH_CHECK();
    return B200_OK;
}

template<class Fr> int step_op_dev(int op, Fr *a, Fr *tmp, const StepShape &sh, const void *g)
{
    StepDomain<Fr> *d = nullptr;
    int rc = get_step_domain<Fr>(sh, &d);
    if (rc) {
        return rc;
    }
    const bool coset = (op == B200_COSET_FFT || op == B200_ICOSET_FFT);
    if (coset) {
        if (!g) {
            return rt::set_error(B200_ERR_ARG, "coset transform needs a generator");
        }
        if ((rc = step_coset_tables<Fr>(*d, static_cast<const uint32_t *>(g)))) {
            return rc;
        }
    }
    cudaStream_t s = rt::stream();
    switch (op) {
    case B200_FFT: return step_fft<Fr>(*d, a, tmp);
    case B200_IFFT: return step_ifft<Fr>(*d, a, tmp);
    case B200_COSET_FFT: // _multiply_by_coset then FFT (:176-181)
        ntt::k_mul_table<Fr><<<grid_for(sh.m, 256), 256, 0, s>>>(sh.m, a, d->gpow.template as<Fr>());
        B200_LAUNCH_CHECK();
        return step_fft<Fr>(*d, a, tmp);
    case B200_ICOSET_FFT: // iFFT then _multiply_by_coset(g^-1) (:183-188)
        if ((rc = step_ifft<Fr>(*d, a, tmp))) {
            return rc;
        }
        ntt::k_mul_table<Fr><<<grid_for(sh.m, 256), 256, 0, s>>>(sh.m, a, d->ginv.template as<Fr>());
        B200_LAUNCH_CHECK();
        return B200_OK;
    }
    return rt::set_error(B200_ERR_ARG, "unknown NTT op %d", op);
}

template<class Fr> int step_try_op(int op, Fr *data, Fr *tmp, Fr *out, size_t m, const void *g, bool *handled)
{
    StepShape sh;
    *handled = step_shape<Fr>(m, &sh);
    if (!*handled) {
        return B200_OK;
    }
    if (out != data) {
        return rt::set_error(B200_ERR_ARG, "step domain transforms run in place");
    }
    return step_op_dev<Fr>(op, data, tmp, sh, g);
}

// step_radix2_domain::divide_by_Z_on_coset (:252-276), coset = Fr::multiplicative_generator
template<class Fr> int step_divide_by_z_dev(StepDomain<Fr> &d, Fr *p)
{
    const StepShape &sh = d.sh;
    cudaStream_t s = rt::stream();
    if (!d.has_zinv) {
        if (d.zinv.reserve(sh.compr * sizeof(Fr))) {
            return B200_ERR_CUDA;
        }
        ntt::k_step_zinv<Fr><<<grid_for(sh.compr, 128), 128, 0, s>>>(sh.compr, d.c(1), d.zinv.template as<Fr>());
        B200_LAUNCH_CHECK();
        d.has_zinv = true;
    }
    ntt::k_step_divz<Fr><<<grid_for(sh.m, 256), 256, 0, s>>>(sh.big, sh.m, sh.compr, d.zinv.template as<Fr>(), d.c(4), p);
    B200_LAUNCH_CHECK();
    return B200_OK;
}

// r1cs_to_qap_witness_map (r1cs_to_qap.tcc:252-330, d1 = d2 = d3 = 0) over a step domain
template<class Fr> int step_witness_map_dev(Fr *dA, Fr *dB, Fr *dC, const StepShape &sh, Fr *dH)
{
    StepDomain<Fr> *d = nullptr;
    int rc = get_step_domain<Fr>(sh, &d);
    if (rc) {
        return rc;
    }
    uint32_t gen[Fr::N];
    for (int i = 0; i < Fr::N; ++i) {
        gen[i] = Fr::params::generator(i);
    }
    if (g_bufS.reserve(sh.big * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    Fr *S = g_bufS.as<Fr>();
    Fr *vecs[3] = {dA, dB, dC};
    for (int v = 0; v < 3; ++v) {
        if ((rc = step_op_dev<Fr>(B200_IFFT, vecs[v], S, sh, nullptr)) ||
            (rc = step_op_dev<Fr>(B200_COSET_FFT, vecs[v], S, sh, gen))) {
            return rc;
        }
    }
    cudaStream_t s = rt::stream();
    ntt::k_pointwise_abc<Fr><<<grid_for(sh.m, 256), 256, 0, s>>>(sh.m, dA, dB, dC, dH);
    B200_LAUNCH_CHECK();
    if ((rc = step_divide_by_z_dev<Fr>(*d, dH)) || (rc = step_op_dev<Fr>(B200_ICOSET_FFT, dH, S, sh, gen))) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemsetAsync(dH + sh.m, 0, sizeof(Fr), s));
    return B200_OK;
}

template<class Fr> int step_divide_by_z_host(void *data, size_t m, bool *handled)
{
    StepShape sh;
    *handled = step_shape<Fr>(m, &sh);
    if (!*handled) {
        return B200_OK;
    }
    StepDomain<Fr> *d = nullptr;
    int rc = get_step_domain<Fr>(sh, &d);
    if (rc) {
        return rc;
    }
    const size_t bytes = m * sizeof(Fr);
    if (g_bufA.reserve(bytes)) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(g_bufA.p, data, bytes, cudaMemcpyHostToDevice, s));
    if ((rc = step_divide_by_z_dev<Fr>(*d, g_bufA.as<Fr>()))) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(data, g_bufA.p, bytes, cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

template<class Fr> int step_witness_map_try(Fr *dA, Fr *dB, Fr *dC, size_t m, Fr *dH, bool *handled)
{
    StepShape sh;
    *handled = step_shape<Fr>(m, &sh);
    if (!*handled) {
        return B200_OK;
    }
    return step_witness_map_dev<Fr>(dA, dB, dC, sh, dH);
}

// ---- extended_radix2_domain -----------------------------------------------------------------------
// m = 2 small with small = 2^s, s the 2-adicity in force (two_adicity): what get_evaluation_domain returns for
// exactly that size (get_evaluation_domain.tcc:44-46; extended_radix2_domain.tcc:21-62). With the fields' own
// parameters only alt_bn128 reaches it with a vector that fits a GPU (2^29 elements, 16 GiB).
using ntt::EXT_LG_LO;

template<class Fr> struct ExtDomain {
    int lg_small = 0;
    size_t small = 0;
    Domain<Fr> *dsmall = nullptr;
    rt::DevBuf consts; // [0] shift^small  [1] (small (1 - shift^small))^-1  [2] Z0^-1  [3] Z1^-1
    rt::DevBuf stab;   // two-level tables of shift^i and shift^-i, i < small
    rt::DevBuf gtab;   // of g^i and g^-i, i < 2 small, for the generator in g_host
    int lg_lo_s = 0, lg_lo_g = 0;
    uint32_t g_host[Fr::N];
    bool has_g = false;
    const Fr *c(int i) { return consts.template as<Fr>() + i; }
    // table t (0: b^i, 1: b^-i) of a buffer laid out lo0 | hi0 | lo1 | hi1 for `count` indices
    static ntt::PowTable<Fr> table(rt::DevBuf &buf, int lg_lo, size_t count, int t)
    {
        const size_t nlo = (size_t)1 << lg_lo, nhi = count >> lg_lo;
        const Fr *base = buf.template as<Fr>() + t * (nlo + nhi);
        return ntt::PowTable<Fr>{base, base + nlo, lg_lo};
    }
};
template<class Fr> std::map<int, std::unique_ptr<ExtDomain<Fr>>> &ext_domains()
{
    static std::map<int, std::unique_ptr<ExtDomain<Fr>>> m;
    return m;
}

// the two-level tables of b and b^-1 (ntt::ext_pow_tables_host: 3 * 2^15 multiplications at most), uploaded
template<class Fr> int ext_pow_tables(const Fr &b, size_t count, int lg_lo, rt::DevBuf &buf)
{
    const size_t nlo = (size_t)1 << lg_lo, nhi = count >> lg_lo;
    std::vector<Fr> h(2 * (nlo + nhi));
    ntt::ext_pow_tables_host<Fr>(b, count, lg_lo, h.data());
    if (buf.reserve(h.size() * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(buf.p, h.data(), h.size() * sizeof(Fr), cudaMemcpyHostToDevice, rt::stream()));
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

template<class Fr> int get_ext_domain(int lg_small, ExtDomain<Fr> **out)
{
    auto &mp = ext_domains<Fr>();
    auto it = mp.find(lg_small);
    if (it == mp.end()) {
        std::unique_ptr<ExtDomain<Fr>> d(new ExtDomain<Fr>());
        d->lg_small = lg_small;
        d->small = (size_t)1 << lg_small;
        int rc = get_domain<Fr>(lg_small, nullptr, &d->dsmall);
        if (rc) {
            return rc;
        }
        Fr hc[4], shift;
        ntt::ext_domain_consts_host<Fr>(lg_small, hc, &shift);
        if (d->consts.reserve(sizeof(hc))) {
            return B200_ERR_CUDA;
        }
        B200_CUDA_CHECK(cudaMemcpyAsync(d->consts.p, hc, sizeof(hc), cudaMemcpyHostToDevice, rt::stream()));
        B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
        d->lg_lo_s = lg_small < EXT_LG_LO ? lg_small : EXT_LG_LO;
        if ((rc = ext_pow_tables<Fr>(shift, d->small, d->lg_lo_s, d->stab))) {
            return rc;
        }
        it = mp.emplace(lg_small, std::move(d)).first;
    }
    *out = it->second.get();
    return B200_OK;
}

template<class Fr> int ext_coset_tables(ExtDomain<Fr> &d, const uint32_t *g)
{
    if (d.has_g && memcmp(g, d.g_host, sizeof(d.g_host)) == 0) {
        return B200_OK;
    }
    Fr gv;
    memcpy(gv.v, g, sizeof(gv.v));
    if (gv.is_zero()) {
        return rt::set_error(B200_ERR_ARG, "coset generator is zero");
    }
    d.lg_lo_g = d.lg_small + 1 < EXT_LG_LO ? d.lg_small + 1 : EXT_LG_LO;
    if (int rc = ext_pow_tables<Fr>(gv, 2 * d.small, d.lg_lo_g, d.gtab)) {
        return rc;
    }
    memcpy(d.g_host, g, sizeof(d.g_host));
    d.has_g = true;
    return B200_OK;
}

// extended_radix2_domain::FFT (:64-91) in place on a (2 small elements); tmp: >= small elements
template<class Fr> int ext_fft(ExtDomain<Fr> &d, Fr *a, Fr *tmp)
{
    cudaStream_t s = rt::stream();
    ntt::k_ext_fft_pre<Fr><<<grid_for(d.small, 256), 256, 0, s>>>(d.small, a, ExtDomain<Fr>::table(d.stab, d.lg_lo_s, d.small, 0), d.c(0));
    B200_LAUNCH_CHECK();
    if (int rc = run_transform<Fr>(*d.dsmall, 0, a, tmp, a, nullptr, nullptr, nullptr)) {
        return rc;
    }
    return run_transform<Fr>(*d.dsmall, 0, a + d.small, tmp, a + d.small, nullptr, nullptr, nullptr);
}

// extended_radix2_domain::iFFT (:93-120) in place
template<class Fr> int ext_ifft(ExtDomain<Fr> &d, Fr *a, Fr *tmp)
{
    int rc;
    if ((rc = run_transform<Fr>(*d.dsmall, 1, a, tmp, a, nullptr, nullptr, nullptr)) ||
        (rc = run_transform<Fr>(*d.dsmall, 1, a + d.small, tmp, a + d.small, nullptr, nullptr, nullptr))) {
        return rc;
    }
    ntt::k_ext_ifft_post<Fr><<<grid_for(d.small, 256), 256, 0, rt::stream()>>>(
        d.small, a, ExtDomain<Fr>::table(d.stab, d.lg_lo_s, d.small, 1), d.c(0));
    B200_LAUNCH_CHECK();
    return B200_OK;
}

template<class Fr> int ext_op_dev(int op, Fr *a, Fr *tmp, int lg_small, const void *g)
{
    ExtDomain<Fr> *d = nullptr;
    int rc = get_ext_domain<Fr>(lg_small, &d);
    if (rc) {
        return rc;
    }
    const bool coset = (op == B200_COSET_FFT || op == B200_ICOSET_FFT);
    if (coset) {
        if (!g) {
            return rt::set_error(B200_ERR_ARG, "coset transform needs a generator");
        }
        if ((rc = ext_coset_tables<Fr>(*d, static_cast<const uint32_t *>(g)))) {
            return rc;
        }
    }
    const uint64_t m = 2 * d->small;
    cudaStream_t s = rt::stream();
    switch (op) {
    case B200_FFT: return ext_fft<Fr>(*d, a, tmp);
    case B200_IFFT: return ext_ifft<Fr>(*d, a, tmp);
    case B200_COSET_FFT: // _multiply_by_coset then FFT (:122-127)
        ntt::k_mul_pow<Fr><<<grid_for(m, 256), 256, 0, s>>>(m, a, ExtDomain<Fr>::table(d->gtab, d->lg_lo_g, m, 0));
        B200_LAUNCH_CHECK();
        return ext_fft<Fr>(*d, a, tmp);
    case B200_ICOSET_FFT: // iFFT then _multiply_by_coset(g^-1) (:129-134)
        if ((rc = ext_ifft<Fr>(*d, a, tmp))) {
            return rc;
        }
        ntt::k_mul_pow<Fr><<<grid_for(m, 256), 256, 0, s>>>(m, a, ExtDomain<Fr>::table(d->gtab, d->lg_lo_g, m, 1));
        B200_LAUNCH_CHECK();
        return B200_OK;
    }
    return rt::set_error(B200_ERR_ARG, "unknown NTT op %d", op);
}

template<class Fr> int ext_try_op(int op, Fr *data, Fr *tmp, Fr *out, size_t m, const void *g, bool *handled)
{
    const int lgn = log2_exact(m);
    *handled = lgn >= 2 && lgn == two_adicity<Fr>() + 1; // small = 1 has no radix-2 transform to run
    if (!*handled) {
        return B200_OK;
    }
    if (out != data) {
        return rt::set_error(B200_ERR_ARG, "extended domain transforms run in place");
    }
    return ext_op_dev<Fr>(op, data, tmp, lgn - 1, g);
}

template<class Fr> int ext_divide_by_z_dev(ExtDomain<Fr> &d, Fr *p)
{
    ntt::k_ext_divz<Fr><<<grid_for(2 * d.small, 256), 256, 0, rt::stream()>>>(d.small, p, d.c(2));
    B200_LAUNCH_CHECK();
    return B200_OK;
}

template<class Fr> int ext_divide_by_z_host(void *data, size_t m)
{
    ExtDomain<Fr> *d = nullptr;
    int rc = get_ext_domain<Fr>(log2_exact(m) - 1, &d);
    if (rc) {
        return rc;
    }
    const size_t bytes = m * sizeof(Fr);
    if (g_bufA.reserve(bytes)) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(g_bufA.p, data, bytes, cudaMemcpyHostToDevice, s));
    if ((rc = ext_divide_by_z_dev<Fr>(*d, g_bufA.as<Fr>()))) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(data, g_bufA.p, bytes, cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

// r1cs_to_qap_witness_map (r1cs_to_qap.tcc:252-330, d1 = d2 = d3 = 0) over an extended domain
template<class Fr> int ext_witness_map_dev(Fr *dA, Fr *dB, Fr *dC, size_t m, Fr *dH)
{
    const int lg_small = log2_exact(m) - 1;
    if (lg_small < 1) {
        return rt::set_error(B200_ERR_DOMAIN, "extended_radix2_domain: bad size %zu", m);
    }
    ExtDomain<Fr> *d = nullptr;
    int rc = get_ext_domain<Fr>(lg_small, &d);
    if (rc) {
        return rc;
    }
    uint32_t gen[Fr::N];
    for (int i = 0; i < Fr::N; ++i) {
        gen[i] = Fr::params::generator(i);
    }
    if (g_bufS.reserve(d->small * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    Fr *S = g_bufS.as<Fr>();
    Fr *vecs[3] = {dA, dB, dC};
    for (int v = 0; v < 3; ++v) {
        if ((rc = ext_op_dev<Fr>(B200_IFFT, vecs[v], S, lg_small, nullptr)) ||
            (rc = ext_op_dev<Fr>(B200_COSET_FFT, vecs[v], S, lg_small, gen))) {
            return rc;
        }
    }
    cudaStream_t s = rt::stream();
    ntt::k_pointwise_abc<Fr><<<grid_for(m, 256), 256, 0, s>>>(m, dA, dB, dC, dH);
    B200_LAUNCH_CHECK();
    if ((rc = ext_divide_by_z_dev<Fr>(*d, dH)) || (rc = ext_op_dev<Fr>(B200_ICOSET_FFT, dH, S, lg_small, gen))) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemsetAsync(dH + m, 0, sizeof(Fr), s));
    return B200_OK;
}

// ---- R1CS on the device -------------------------------------------------------------------------
struct R1csSet {
    int curve = -1;
    size_t nc = 0, ni = 0, nv = 0, nnz = 0, num_long = 0, num_segs = 0;
    uint32_t *row_ptr = nullptr, *col = nullptr, *long_rows = nullptr, *seg_ptr = nullptr;
    ntt::LongSeg *segs = nullptr;
    void *partials = nullptr;
    void *val = nullptr;
};
std::map<uint64_t, R1csSet> g_r1cs;
uint64_t g_next_r1cs = 1;

void free_r1cs(R1csSet &r)
{
    cudaFree(r.row_ptr);
    cudaFree(r.long_rows);
    cudaFree(r.seg_ptr);
    cudaFree(r.segs);
    cudaFree(r.partials);
    cudaFree(r.col);
    cudaFree(r.val);
    r = R1csSet();
}

template<class Fr>
int r1cs_upload(R1csSet &r, const uint32_t *row_ptr, const uint32_t *col, const void *val)
{
    cudaStream_t s = rt::stream();
    const size_t rows = 3 * r.nc + 1;
    B200_CUDA_CHECK(cudaMalloc((void **)&r.row_ptr, rows * 4));
    B200_CUDA_CHECK(cudaMalloc((void **)&r.col, (r.nnz ? r.nnz : 1) * 4));
    B200_CUDA_CHECK(cudaMalloc(&r.val, (r.nnz ? r.nnz : 1) * sizeof(Fr)));
    B200_CUDA_CHECK(cudaMemcpyAsync(r.row_ptr, row_ptr, rows * 4, cudaMemcpyHostToDevice, s));
    std::vector<uint32_t> long_rows, seg_ptr;
    std::vector<ntt::LongSeg> segs;
    ntt::r1cs_long_plan(row_ptr, rows - 1, long_rows, seg_ptr, segs);
    r.num_long = long_rows.size();
    r.num_segs = segs.size();
    if (r.num_long) {
        B200_CUDA_CHECK(cudaMalloc((void **)&r.long_rows, r.num_long * 4));
        B200_CUDA_CHECK(cudaMalloc((void **)&r.seg_ptr, seg_ptr.size() * 4));
        B200_CUDA_CHECK(cudaMalloc((void **)&r.segs, segs.size() * sizeof(ntt::LongSeg)));
        B200_CUDA_CHECK(cudaMalloc(&r.partials, segs.size() * sizeof(Fr)));
        B200_CUDA_CHECK(cudaMemcpyAsync(r.long_rows, long_rows.data(), r.num_long * 4, cudaMemcpyHostToDevice, s));
        B200_CUDA_CHECK(cudaMemcpyAsync(r.seg_ptr, seg_ptr.data(), seg_ptr.size() * 4, cudaMemcpyHostToDevice, s));
        B200_CUDA_CHECK(cudaMemcpyAsync(r.segs, segs.data(), segs.size() * sizeof(ntt::LongSeg), cudaMemcpyHostToDevice, s));
        B200_CUDA_CHECK(cudaStreamSynchronize(s)); // the host vectors go out of scope
    }
    if (r.nnz) {
        B200_CUDA_CHECK(cudaMemcpyAsync(r.col, col, r.nnz * 4, cudaMemcpyHostToDevice, s));
        B200_CUDA_CHECK(cudaMemcpyAsync(r.val, val, r.nnz * sizeof(Fr), cudaMemcpyHostToDevice, s));
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

template<class Fr> int r1cs_eval_launch(const R1csSet &r, const Fr *d_w, size_t m, Fr *aA, Fr *aB, Fr *aC)
{
    cudaStream_t s = rt::stream();
    ntt::k_r1cs_eval<Fr><<<grid_for(3 * m, 256), 256, 0, s>>>(
        r.nc, r.ni, m, r.row_ptr, r.col, static_cast<const Fr *>(r.val), d_w, aA, aB, aC);
    B200_LAUNCH_CHECK();
    if (r.num_long) {
        ntt::k_r1cs_eval_long<Fr><<<(unsigned)r.num_segs, ntt::R1CS_LONG_THREADS, 0, s>>>(
            r.segs, r.col, static_cast<const Fr *>(r.val), d_w, static_cast<Fr *>(r.partials));
        B200_LAUNCH_CHECK();
        ntt::k_r1cs_long_finish<Fr><<<(unsigned)r.num_long, ntt::R1CS_LONG_THREADS, 0, s>>>(
            r.nc, r.long_rows, r.seg_ptr, static_cast<const Fr *>(r.partials), aA, aB, aC);
        B200_LAUNCH_CHECK();
    }
    return B200_OK;
}

// d_w: (1 | primary | auxiliary), nv + 1 elements on the device. dH: m + 1 elements on the device.
template<class Fr> int witness_map_r1cs_dev(const R1csSet &r, const void *d_w, size_t m, void *dH)
{
    if (m < r.nc + r.ni + 1) {
        return rt::set_error(B200_ERR_DOMAIN, "domain size %zu < num_constraints + num_inputs + 1 = %zu", m,
                             r.nc + r.ni + 1);
    }
    const size_t bytes = m * sizeof(Fr);
    if (g_bufA.reserve(bytes) || g_bufB.reserve(bytes) || g_bufC.reserve(bytes)) {
        return B200_ERR_CUDA;
    }
    if (int rc = r1cs_eval_launch<Fr>(r, static_cast<const Fr *>(d_w), m, g_bufA.as<Fr>(), g_bufB.as<Fr>(), g_bufC.as<Fr>())) {
        return rc;
    }
    return witness_map_dev<Fr>(g_bufA.as<Fr>(), g_bufB.as<Fr>(), g_bufC.as<Fr>(), m, static_cast<Fr *>(dH));
}

// add_poly_Z(c, H) of the domain kind that serves m (basic_radix2_domain.tcc:109-115, extended_radix2_domain.tcc:
// 203-214, step_radix2_domain.tcc:235-245), as a list of single-coefficient additions
template<class Fr> int add_poly_z_fixups(size_t m, const Fr &c, ntt::CoeffFixups<Fr> *f)
{
    const Fr one = Fr::one();
    auto push = [&](uint64_t i, const Fr &v) {
        f->idx[f->n] = i;
        f->val[f->n] = v;
        ++f->n;
    };
    const int lgn = log2_exact(m);
    if (lgn >= 1 && lgn <= two_adicity<Fr>()) {
        push(m, c);
        push(0, c.neg());
        return B200_OK;
    }
    Fr g;
    for (int i = 0; i < Fr::N; ++i) {
        g.v[i] = Fr::params::generator(i);
    }
    if (lgn >= 2 && lgn == two_adicity<Fr>() + 1) {
        const size_t small = m / 2;
        const Fr ssm = host_pow(g.sqr(), small);
        push(m, c);
        push(small, (c * (ssm + one)).neg());
        push(0, c * ssm);
        return B200_OK;
    }
    StepShape sh;
    if (!step_shape<Fr>(m, &sh)) {
        return rt::set_error(B200_ERR_DOMAIN, "no radix-2 domain of size %zu", m);
    }
    Fr w;
    for (int i = 0; i < Fr::N; ++i) {
        w.v[i] = Fr::params::root_of_unity(i);
    }
    for (int i = Fr::params::S; i > sh.lg_big + 1; --i) {
        w = w.sqr(); // omega = get_root_of_unity(2 big), step_radix2_domain.tcc:57
    }
    const Fr wsm = host_pow(w, sh.small);
    push(m, c);
    push(sh.big, (c * wsm).neg());
    push(sh.small, c.neg());
    push(0, c * wsm);
    return B200_OK;
}

rt::DevBuf g_bufW, g_bufT;

// r1cs_to_qap_witness_map (r1cs_to_qap.tcc:209-344) with the ZK patch: coefficients_for_H =
// H + d2 coef(A) + d1 coef(B) - d3 + d1 d2 Z. The patch's only vector part, d2 coef(A) + d1 coef(B), is the inverse
// transform of d2 aA + d1 aB (the transform is linear), taken before the chains overwrite the evaluations.
// w_host: (1 | primary | auxiliary); d123: d1, d2, d3 (host, Montgomery); out_H: m + 1 elements (host).
template<class Fr> int witness_map_r1cs_zk_host(const R1csSet &r, const void *w_host, size_t m, const void *d123, void *out_H)
{
    if (m < r.nc + r.ni + 1) {
        return rt::set_error(B200_ERR_DOMAIN, "domain size %zu < num_constraints + num_inputs + 1 = %zu", m,
                             r.nc + r.ni + 1);
    }
    Fr d[3];
    memcpy(d, d123, sizeof(d));
    const bool zk = !(d[0].is_zero() && d[1].is_zero() && d[2].is_zero());
    const size_t bytes = m * sizeof(Fr);
    if (g_bufA.reserve(bytes) || g_bufB.reserve(bytes) || g_bufC.reserve(bytes) || g_bufH.reserve(bytes + sizeof(Fr)) ||
        g_bufW.reserve((r.nv + 1) * sizeof(Fr)) || g_bufS.reserve(bytes) || (zk && g_bufT.reserve(bytes))) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    Fr *A = g_bufA.as<Fr>(), *B = g_bufB.as<Fr>(), *Cc = g_bufC.as<Fr>(), *H = g_bufH.as<Fr>(), *T = g_bufT.as<Fr>();
    B200_CUDA_CHECK(cudaMemcpyAsync(g_bufW.p, w_host, (r.nv + 1) * sizeof(Fr), cudaMemcpyHostToDevice, s));
    int rc = r1cs_eval_launch<Fr>(r, g_bufW.as<Fr>(), m, A, B, Cc);
    if (rc) {
        return rc;
    }
    ntt::CoeffFixups<Fr> fix;
    fix.n = 0;
    if (zk) {
        ntt::k_lincomb2<Fr><<<grid_for(m, 256), 256, 0, s>>>(m, A, B, d[1], d[0], T);
        B200_LAUNCH_CHECK();
        if ((rc = ntt_op_dev<Fr>(B200_IFFT, T, g_bufS.as<Fr>(), T, m, nullptr))) {
            return rc;
        }
        fix.idx[0] = 0;
        fix.val[0] = d[2].neg(); // coefficients_for_H[0] -= d3, :270
        fix.n = 1;
        if ((rc = add_poly_z_fixups<Fr>(m, d[0] * d[1], &fix))) {
            return rc;
        }
    }
    if ((rc = witness_map_dev<Fr>(A, B, Cc, m, H))) {
        return rc;
    }
    if (zk) {
        ntt::k_add_vec<Fr><<<grid_for(m, 256), 256, 0, s>>>(m, H, T);
        B200_LAUNCH_CHECK();
        ntt::k_coeff_fixups<Fr><<<1, 32, 0, s>>>(fix, H);
        B200_LAUNCH_CHECK();
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(out_H, H, bytes + sizeof(Fr), cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

template<class Fr> int r1cs_eval_host(const R1csSet &r, const void *w_host, size_t m, void *aA, void *aB, void *aC)
{
    const size_t bytes = m * sizeof(Fr);
    if (g_bufA.reserve(bytes) || g_bufB.reserve(bytes) || g_bufC.reserve(bytes) || g_bufH.reserve((r.nv + 1) * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    B200_CUDA_CHECK(cudaMemcpyAsync(g_bufH.p, w_host, (r.nv + 1) * sizeof(Fr), cudaMemcpyHostToDevice, s));
    if (int rc = r1cs_eval_launch<Fr>(r, g_bufH.as<Fr>(), m, g_bufA.as<Fr>(), g_bufB.as<Fr>(), g_bufC.as<Fr>())) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(aA, g_bufA.p, bytes, cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaMemcpyAsync(aB, g_bufB.p, bytes, cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaMemcpyAsync(aC, g_bufC.p, bytes, cudaMemcpyDeviceToHost, s));
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

template<class Fr> int gather_dev(const void *d_src, const uint32_t *d_idx, size_t n, void *d_dst)
{
    ntt::k_gather<Fr><<<grid_for(n ? n : 1, 256), 256, 0, rt::stream()>>>(
        n, static_cast<const Fr *>(d_src), d_idx, static_cast<Fr *>(d_dst));
    B200_LAUNCH_CHECK();
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

// iFFT then cosetFFT(multiplicative_generator) of one vector, in place, on whichever radix-2 kind serves m: the fused
// two-transform chain on a basic domain, the two operators one after the other on a step or extended domain.
template<class Fr> int chain_any_domain(Fr *X, size_t m)
{
    const int lgn = log2_exact(m);
    if (lgn >= 1 && lgn <= two_adicity<Fr>()) {
        Domain<Fr> *d = nullptr;
        if (int rc = witness_domain<Fr>(m, &d)) {
            return rc;
        }
        return witness_chain<Fr>(*d, X);
    }
    uint32_t gen[Fr::N];
    for (int i = 0; i < Fr::N; ++i) {
        gen[i] = Fr::params::generator(i);
    }
    if (g_bufS.reserve(m * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    if (int rc = ntt_op_dev<Fr>(B200_IFFT, X, g_bufS.as<Fr>(), X, m, nullptr)) {
        return rc;
    }
    return ntt_op_dev<Fr>(B200_COSET_FFT, X, g_bufS.as<Fr>(), X, m, gen);
}

// (A B - C) / Z on the coset, icosetFFT, H[m] = 0 on whichever kind serves m (r1cs_to_qap.tcc:293-330). dA clobbered.
template<class Fr> int finish_any_domain(Fr *dA, const Fr *dB, const Fr *dC, size_t m, Fr *dH)
{
    const int lgn = log2_exact(m);
    if (lgn >= 1 && lgn <= two_adicity<Fr>()) {
        Domain<Fr> *d = nullptr;
        if (int rc = witness_domain<Fr>(m, &d)) {
            return rc;
        }
        return witness_finish<Fr>(*d, m, dA, dB, dC, dH);
    }
    uint32_t gen[Fr::N];
    for (int i = 0; i < Fr::N; ++i) {
        gen[i] = Fr::params::generator(i);
    }
    if (g_bufS.reserve(m * sizeof(Fr))) {
        return B200_ERR_CUDA;
    }
    cudaStream_t s = rt::stream();
    ntt::k_pointwise_abc<Fr><<<grid_for(m, 256), 256, 0, s>>>(m, dA, dB, dC, dH);
    B200_LAUNCH_CHECK();
    int rc;
    if (lgn >= 2 && lgn == two_adicity<Fr>() + 1) {
        ExtDomain<Fr> *d = nullptr;
        if ((rc = get_ext_domain<Fr>(lgn - 1, &d)) || (rc = ext_divide_by_z_dev<Fr>(*d, dH))) {
            return rc;
        }
    } else {
        StepShape sh;
        if (!step_shape<Fr>(m, &sh)) {
            return rt::set_error(B200_ERR_DOMAIN, "no radix-2 domain of size %zu", m);
        }
        StepDomain<Fr> *d = nullptr;
        if ((rc = get_step_domain<Fr>(sh, &d)) || (rc = step_divide_by_z_dev<Fr>(*d, dH))) {
            return rc;
        }
    }
    if ((rc = ntt_op_dev<Fr>(B200_ICOSET_FFT, dH, g_bufS.as<Fr>(), dH, m, gen))) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemsetAsync(dH + m, 0, sizeof(Fr), s));
    return B200_OK;
}

// One chain of the witness map on its own (multi-GPU: the A, B and C chains run on different GPUs,
// SURVEY.md §8e): evaluations of the constraint rows from d_w, iFFT, cosetFFT -> d_out (m elements).
template<class Fr> int witness_chain_r1cs_dev(const R1csSet &r, const void *d_w, size_t m, int which, void *d_out)
{
    if (m < r.nc + r.ni + 1) {
        return rt::set_error(B200_ERR_DOMAIN, "domain size %zu < num_constraints + num_inputs + 1 = %zu", m,
                             r.nc + r.ni + 1);
    }
    const size_t bytes = m * sizeof(Fr);
    if (g_bufA.reserve(bytes) || g_bufB.reserve(bytes) || g_bufC.reserve(bytes)) {
        return B200_ERR_CUDA;
    }
    if (int rc = r1cs_eval_launch<Fr>(r, static_cast<const Fr *>(d_w), m, g_bufA.as<Fr>(), g_bufB.as<Fr>(), g_bufC.as<Fr>())) {
        return rc;
    }
    Fr *X = which == 0 ? g_bufA.as<Fr>() : which == 1 ? g_bufB.as<Fr>() : g_bufC.as<Fr>();
    if (int rc = chain_any_domain<Fr>(X, m)) {
        return rc;
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(d_out, X, bytes, cudaMemcpyDeviceToDevice, rt::stream()));
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

template<class Fr> int witness_finish_dev(void *dA, const void *dB, const void *dC, size_t m, void *dH)
{
    if (int rc = finish_any_domain<Fr>(static_cast<Fr *>(dA), static_cast<const Fr *>(dB), static_cast<const Fr *>(dC), m,
                                       static_cast<Fr *>(dH))) {
        return rc;
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

#define DISPATCH_FR(curve, call)                                                                   \
    switch (curve) {                                                                               \
    case B200_ALT_BN128: {                                                                         \
        typedef alt_bn128_fr Fr;                                                                   \
        return call;                                                                               \
    }                                                                                              \
    case B200_BLS12_377: {                                                                         \
        typedef bls12_377_fr Fr;                                                                   \
        return call;                                                                               \
    }                                                                                              \
    case B200_BW6_761: {                                                                           \
        typedef bw6_761_fr Fr;                                                                     \
        return call;                                                                               \
    }                                                                                              \
    default: return rt::set_error(B200_ERR_ARG, "unknown curve %d", curve);                        \
    }

} // namespace

template<class Fr> int set_two_adicity(int s)
{
    if (s < 0 || s > Fr::params::S) {
        return rt::set_error(B200_ERR_ARG, "two-adicity %d outside [1, %d] (0 restores the field's own)", s, Fr::params::S);
    }
    two_adicity<Fr>() = s ? s : Fr::params::S;
    return B200_OK;
}

// get_evaluation_domain.tcc:37-58 for the sizes it accepts as they are: which radix-2 kind serves m
template<class Fr> int domain_kind(size_t m)
{
    const int lgn = log2_exact(m);
    if (lgn >= 1 && lgn <= two_adicity<Fr>()) {
        return B200_DOMAIN_BASIC;
    }
    if (lgn >= 2 && lgn == two_adicity<Fr>() + 1) {
        return B200_DOMAIN_EXTENDED;
    }
    StepShape sh;
    return step_shape<Fr>(m, &sh) ? B200_DOMAIN_STEP : B200_DOMAIN_NONE;
}

extern "C" {

int b200_domain_kind(int curve, size_t m)
{
    std::lock_guard<std::mutex> lk(g_mu);
    switch (curve) {
    case B200_ALT_BN128: return domain_kind<alt_bn128_fr>(m);
    case B200_BLS12_377: return domain_kind<bls12_377_fr>(m);
    case B200_BW6_761: return domain_kind<bw6_761_fr>(m);
    }
    return B200_DOMAIN_NONE;
}

int b200_ntt(int curve, int op, void *data, size_t m, const void *g)
{
    if (!data) {
        return rt::set_error(B200_ERR_ARG, "null data");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    DISPATCH_FR(curve, ntt_host<Fr>(op, data, m, g));
}

int b200_ntt_dev(int curve, int op, void *dev_data, size_t m, const void *g)
{
    if (!dev_data) {
        return rt::set_error(B200_ERR_ARG, "null data");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    DISPATCH_FR(curve, ntt_dev<Fr>(op, dev_data, m, g));
}

int b200_set_two_adicity(int curve, int s)
{
    std::lock_guard<std::mutex> lk(g_mu);
    DISPATCH_FR(curve, set_two_adicity<Fr>(s));
}

int b200_divide_by_Z_on_coset(int curve, void *data, size_t m)
{
    if (!data) {
        return rt::set_error(B200_ERR_ARG, "null data");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    DISPATCH_FR(curve, divide_by_z<Fr>(data, m));
}

int b200_witness_map_h(int curve, const void *aA, const void *aB, const void *aC, size_t m, void *out_H)
{
    if (!aA || !aB || !aC || !out_H) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    DISPATCH_FR(curve, witness_map_host<Fr>(aA, aB, aC, m, out_H));
}

int b200_witness_map_h_dev(int curve, void *dA, void *dB, void *dC, size_t m, void *dH)
{
    if (!dA || !dB || !dC || !dH) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    int rc;
    switch (curve) {
    case B200_ALT_BN128:
        rc = witness_map_dev<alt_bn128_fr>((alt_bn128_fr *)dA, (alt_bn128_fr *)dB, (alt_bn128_fr *)dC, m, (alt_bn128_fr *)dH);
        break;
    case B200_BLS12_377:
        rc = witness_map_dev<bls12_377_fr>((bls12_377_fr *)dA, (bls12_377_fr *)dB, (bls12_377_fr *)dC, m, (bls12_377_fr *)dH);
        break;
    case B200_BW6_761:
        rc = witness_map_dev<bw6_761_fr>((bw6_761_fr *)dA, (bw6_761_fr *)dB, (bw6_761_fr *)dC, m, (bw6_761_fr *)dH);
        break;
    default: return rt::set_error(B200_ERR_ARG, "unknown curve %d", curve);
    }
    if (rc) {
        return rc;
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

int b200_r1cs_upload(int curve, size_t num_constraints, size_t num_inputs, size_t num_variables,
                     const uint32_t *row_ptr, const uint32_t *col, const void *val, b200_r1cs_t *handle)
{
    if (!row_ptr || !handle || num_inputs > num_variables) {
        return rt::set_error(B200_ERR_ARG, "bad R1CS arguments");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    const size_t nnz = row_ptr[3 * num_constraints];
    if (row_ptr[0] != 0 || (nnz && (!col || !val)) || 3 * (num_constraints + num_inputs + 1) >= 0xffffffffull) {
        return rt::set_error(B200_ERR_ARG, "bad R1CS CSR");
    }
    for (size_t i = 0; i < 3 * num_constraints; ++i) {
        if (row_ptr[i] > row_ptr[i + 1]) {
            return rt::set_error(B200_ERR_ARG, "row_ptr not monotone at %zu", i);
        }
    }
    for (size_t e = 0; e < nnz; ++e) {
        if (col[e] > num_variables) {
            return rt::set_error(B200_ERR_ARG, "column %u > num_variables %zu", col[e], num_variables);
        }
    }
    std::lock_guard<std::mutex> lk(g_mu);
    R1csSet r;
    r.curve = curve;
    r.nc = num_constraints;
    r.ni = num_inputs;
    r.nv = num_variables;
    r.nnz = nnz;
    int rc;
    switch (curve) {
    case B200_ALT_BN128: rc = r1cs_upload<alt_bn128_fr>(r, row_ptr, col, val); break;
    case B200_BLS12_377: rc = r1cs_upload<bls12_377_fr>(r, row_ptr, col, val); break;
    case B200_BW6_761: rc = r1cs_upload<bw6_761_fr>(r, row_ptr, col, val); break;
    default: return rt::set_error(B200_ERR_ARG, "unknown curve %d", curve);
    }
    if (rc) {
        free_r1cs(r);
        return rc;
    }
    *handle = g_next_r1cs++;
    g_r1cs[*handle] = r;
    return B200_OK;
}

int b200_r1cs_free(b200_r1cs_t handle)
{
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_r1cs.find(handle);
    if (it == g_r1cs.end()) {
        return rt::set_error(B200_ERR_HANDLE, "unknown R1CS handle %llu", (unsigned long long)handle);
    }
    cudaStreamSynchronize(rt::stream());
    free_r1cs(it->second);
    g_r1cs.erase(it);
    return B200_OK;
}

int b200_r1cs_evaluate(b200_r1cs_t handle, const void *w, size_t m, void *aA, void *aB, void *aC)
{
    if (!w || !aA || !aB || !aC) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_r1cs.find(handle);
    if (it == g_r1cs.end()) {
        return rt::set_error(B200_ERR_HANDLE, "unknown R1CS handle %llu", (unsigned long long)handle);
    }
    const R1csSet &r = it->second;
    if (m < r.nc + r.ni + 1) {
        return rt::set_error(B200_ERR_DOMAIN, "m = %zu < num_constraints + num_inputs + 1", m);
    }
    DISPATCH_FR(r.curve, r1cs_eval_host<Fr>(r, w, m, aA, aB, aC));
}

int b200_witness_map_r1cs_zk(b200_r1cs_t handle, const void *w, size_t m, const void *d123, void *out_H)
{
    if (!w || !d123 || !out_H) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_r1cs.find(handle);
    if (it == g_r1cs.end()) {
        return rt::set_error(B200_ERR_HANDLE, "unknown R1CS handle %llu", (unsigned long long)handle);
    }
    const R1csSet &r = it->second;
    DISPATCH_FR(r.curve, witness_map_r1cs_zk_host<Fr>(r, w, m, d123, out_H));
}

int b200_witness_map_r1cs_dev(b200_r1cs_t handle, const void *d_w, size_t m, void *dH)
{
    if (!d_w || !dH) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_r1cs.find(handle);
    if (it == g_r1cs.end()) {
        return rt::set_error(B200_ERR_HANDLE, "unknown R1CS handle %llu", (unsigned long long)handle);
    }
    const R1csSet &r = it->second;
    int rc;
    switch (r.curve) {
    case B200_ALT_BN128: rc = witness_map_r1cs_dev<alt_bn128_fr>(r, d_w, m, dH); break;
    case B200_BLS12_377: rc = witness_map_r1cs_dev<bls12_377_fr>(r, d_w, m, dH); break;
    default: rc = witness_map_r1cs_dev<bw6_761_fr>(r, d_w, m, dH); break;
    }
    if (rc) {
        return rc;
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

int b200_witness_chain_r1cs_dev(b200_r1cs_t handle, const void *d_w, size_t m, int which, void *d_out)
{
    if (!d_w || !d_out || which < 0 || which > 2) {
        return rt::set_error(B200_ERR_ARG, "null pointer or chain index outside 0..2");
    }
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_r1cs.find(handle);
    if (it == g_r1cs.end()) {
        return rt::set_error(B200_ERR_HANDLE, "unknown R1CS handle %llu", (unsigned long long)handle);
    }
    const R1csSet &r = it->second;
    DISPATCH_FR(r.curve, witness_chain_r1cs_dev<Fr>(r, d_w, m, which, d_out));
}

int b200_witness_finish_dev(int curve, void *dA, const void *dB, const void *dC, size_t m, void *dH)
{
    if (!dA || !dB || !dC || !dH) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    DISPATCH_FR(curve, witness_finish_dev<Fr>(dA, dB, dC, m, dH));
}

int b200_gather_dev(int curve, const void *d_src, const uint32_t *d_idx, size_t n, void *d_dst)
{
    if (n && (!d_src || !d_idx || !d_dst)) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    DISPATCH_FR(curve, gather_dev<Fr>(d_src, d_idx, n, d_dst));
}

} // extern "C"

// ---- b200_ntt_dev_noalloc: the radix-2 transforms without any device allocation ------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Everything b200_ntt_dev keeps in
// grow-only cudaMalloc buffers — the Stockham ping-pong buffer, the twiddle table, the domain constants and the
// coset table — lives in ONE file-scope device array of a fixed size instead. Every GPU that loads the library
// holds these bytes until the process exits. All calls share the array: calls that overlap on the GPU race.
namespace
{

constexpr int NOALLOC_MAX_LOG = B200_NTT_NOALLOC_MAX_LOG;  // m <= 2^16 elements per call
constexpr size_t NOALLOC_MAX_FR = 48;                      // bytes of the widest Fr (bw6_761: 6 x u64)
constexpr size_t NOALLOC_SMALL = 256 * NOALLOC_MAX_FR;     // consts[8] | g[1] | pows[192], padded
// tmp[m] | twiddles[m / 2] | coset table[m] | small
constexpr size_t NOALLOC_SCRATCH_BYTES = ((size_t)5 << (NOALLOC_MAX_LOG - 1)) * NOALLOC_MAX_FR + NOALLOC_SMALL;

static_assert(NOALLOC_SCRATCH_BYTES == B200_NTT_NOALLOC_SCRATCH_BYTES, "header and scratch disagree");

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

// What the array holds from earlier calls on one device (the tables are rebuilt only when the key changes).
struct NoallocState {
    unsigned char *base = nullptr; // device address of g_noalloc_scratch on this device (cudaGetSymbolAddress)
    int curve = -1, lgn = 0;
    bool has_g = false;
    uint32_t g[NOALLOC_MAX_FR / 4];
    int coset_which = -1; // CosetTable currently in the coset slot, -1 none
};
constexpr int NOALLOC_MAX_DEVICES = 64;
NoallocState g_noalloc[NOALLOC_MAX_DEVICES];

// g arrives by value (kernel parameter space), so no host-to-device copy is needed for it
template<class Fr> __global__ void k_noalloc_consts(int lgn, Fr g, Fr *g_slot, Fr *consts, Fr *pows)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        ntt::gstore(g_slot, g);
        ntt::domain_consts_body<Fr>(lgn, g_slot, consts, pows, pows + 64, pows + 128);
    }
}

template<class Fr> int ntt_dev_noalloc(NoallocState &st, int curve, int op, void *dev_data, size_t m, const void *g)
{
    const int lgn = log2_exact(m);
    if (lgn < 0 || lgn > two_adicity<Fr>()) {
        return rt::set_error(B200_ERR_DOMAIN, "basic_radix2_domain: size %zu is not a power of two in [2, 2^%d]", m,
                             two_adicity<Fr>());
    }
    if (lgn > NOALLOC_MAX_LOG) {
        return rt::set_error(B200_ERR_ARG,
                             "b200_ntt_dev_noalloc: m = %zu exceeds the static scratch (at most 2^%d elements; use "
                             "b200_ntt_dev for larger transforms)",
                             m, NOALLOC_MAX_LOG);
    }
    const bool coset = (op == B200_COSET_FFT || op == B200_ICOSET_FFT);
    if (op < B200_FFT || op > B200_ICOSET_FFT) {
        return rt::set_error(B200_ERR_ARG, "unknown NTT op %d", op);
    }
    if (coset && !g) {
        return rt::set_error(B200_ERR_ARG, "coset transform needs a generator");
    }
    static_assert(sizeof(Fr) <= NOALLOC_MAX_FR, "scratch is sized for 48-byte scalars");
    cudaStream_t s = rt::stream();
    Fr *tmp = reinterpret_cast<Fr *>(st.base);
    Fr *twiddles = reinterpret_cast<Fr *>(st.base + ((size_t)2 << (NOALLOC_MAX_LOG - 1)) * NOALLOC_MAX_FR);
    Fr *coset_tbl = reinterpret_cast<Fr *>(st.base + ((size_t)3 << (NOALLOC_MAX_LOG - 1)) * NOALLOC_MAX_FR);
    Fr *small = reinterpret_cast<Fr *>(st.base + ((size_t)5 << (NOALLOC_MAX_LOG - 1)) * NOALLOC_MAX_FR);
    Fr *consts = small, *g_slot = small + 8, *pows = small + 16;

    uint32_t gen[Fr::N];
    if (coset) {
        memcpy(gen, g, sizeof(gen));
    } else if (st.curve == curve && st.has_g) {
        memcpy(gen, st.g, sizeof(gen)); // plain transforms do not care which generator the constants carry
    } else {
        for (int i = 0; i < Fr::N; ++i) {
            gen[i] = Fr::params::generator(i);
        }
    }
    const bool new_domain = st.curve != curve || st.lgn != lgn;
    const bool new_g = new_domain || !st.has_g || memcmp(gen, st.g, sizeof(gen)) != 0;
    if (new_g) {
        st.curve = -1; // stays invalid if a launch below fails
        Fr gv;
        memcpy(gv.v, gen, sizeof(gen));
        k_noalloc_consts<Fr><<<1, 32, 0, s>>>(lgn, gv, g_slot, consts, pows);
        B200_LAUNCH_CHECK();
        st.coset_which = -1;
    }
    if (new_domain) {
        const uint64_t half = 1ull << (lgn - 1);
        ntt::k_pow_table<Fr><<<grid_for(half, 256), 256, 0, s>>>(half, pows, lgn - 1, nullptr, twiddles);
        B200_LAUNCH_CHECK();
    }
    if (new_g) {
        st.curve = curve;
        st.lgn = lgn;
        st.has_g = true;
        memcpy(st.g, gen, sizeof(gen));
    }
    if (coset) {
        const int which = op == B200_COSET_FFT ? T_GPOW : T_GINV_MINV;
        if (st.coset_which != which) {
            st.coset_which = -1;
            ntt::k_pow_table<Fr><<<grid_for(m, 256), 256, 0, s>>>(m, which == T_GPOW ? pows + 64 : pows + 128, lgn,
                                                                  which == T_GPOW ? nullptr : consts + 1, coset_tbl);
            B200_LAUNCH_CHECK();
            st.coset_which = which;
        }
    }
    // run_transform reads only lgn and the twiddle pointer of a Domain; this one owns nothing
    Domain<Fr> d;
    d.lgn = lgn;
    d.twiddles.p = twiddles;
    Fr *data = static_cast<Fr *>(dev_data);
    int rc = B200_OK;
    switch (op) {
    case B200_FFT: rc = run_transform<Fr>(d, 0, data, tmp, data, nullptr, nullptr, nullptr); break;
    case B200_IFFT: rc = run_transform<Fr>(d, 1, data, tmp, data, nullptr, nullptr, consts + 1); break;
    case B200_COSET_FFT: rc = run_transform<Fr>(d, 0, data, tmp, data, coset_tbl, nullptr, nullptr); break;
    case B200_ICOSET_FFT: rc = run_transform<Fr>(d, 1, data, tmp, data, nullptr, coset_tbl, nullptr); break;
    }
    if (rc) {
        return rc;
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

} // namespace

extern "C" int b200_ntt_dev_noalloc(int curve, int op, void *dev_data, size_t m, const void *g)
{
    if (!dev_data) {
        return rt::set_error(B200_ERR_ARG, "null data");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    int dev = -1;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) {
        return rt::set_error(B200_ERR_ARG, "device %d outside [0, %d)", dev, NOALLOC_MAX_DEVICES);
    }
    std::lock_guard<std::mutex> lk(g_mu); // also guards run_transform's shared-memory attribute cache
    NoallocState &st = g_noalloc[dev];
    if (!st.base) {
        // In host code g_noalloc_scratch is only a stand-in for the symbol: the address a kernel may use on THIS
        // device comes from the runtime, once per device, with that device current.
        void *p = nullptr;
        B200_CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        st.base = static_cast<unsigned char *>(p);
    }
    DISPATCH_FR(curve, ntt_dev_noalloc<Fr>(st, curve, op, dev_data, m, g));
}
