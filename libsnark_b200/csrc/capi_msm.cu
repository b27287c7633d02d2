// capi_msm.cu — C-ABI entry points for the MSM operators (include/libsnark_b200.h).
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <shared_mutex>
#include <vector>

#include "curve_params.cuh"
#include "msm_geometry.hpp"
#include "runtime.hpp"

using namespace b200;

namespace
{

struct GroupVt {
    int (*bases_upload)(int, int, const void *, size_t, size_t, rt::BaseSet *);
    int (*bases_precompute)(rt::BaseSet *);
    int (*msm_windows)(const rt::BaseSet &, size_t, size_t, const void *, void *, int *, int *, const rt::SharedSort *, uint32_t);
    int (*msm_sort)(const rt::BaseSet &, size_t, const void *, rt::SharedSort *, cudaStream_t);
    int (*msm_fold)(const void *, int, int, int, size_t, void *);
    size_t (*window_point_bytes)();
    int (*scalar_bits)();
};
#define GROUP_VT(name) {name##_bases_upload, name##_bases_precompute, name##_msm_windows, name##_msm_sort, name##_msm_fold, name##_window_point_bytes, name##_scalar_bits}

struct CurveVt {
    GroupVt g[2]; // [0] = G1, [1] = G2
    size_t fr_bytes, fq_bytes;
    bool g2_is_fq2;
    const GroupVt &grp(int group) const { return g[group == B200_G2 ? 1 : 0]; }
    int scalar_bits() const { return g[0].scalar_bits(); }
    size_t window_point_bytes(int group) const { return grp(group).window_point_bytes(); }
};

const CurveVt *vt(int curve)
{
    static const CurveVt tbl[3] = {
        {{GROUP_VT(alt_bn128_g1), GROUP_VT(alt_bn128_g2)}, 32, 32, true},
        {{GROUP_VT(bls12_377_g1), GROUP_VT(bls12_377_g2)}, 32, 48, true},
        {{GROUP_VT(bw6_761_g1), GROUP_VT(bw6_761_g1)}, 48, 96, false},
    };
    if (curve < 0 || curve > 2) {
        rt::set_error(B200_ERR_ARG, "unknown curve %d", curve);
        return nullptr;
    }
    return &tbl[curve];
}

size_t group_bytes(const CurveVt *v, int group) { return 3 * v->fq_bytes * ((group == B200_G2 && v->g2_is_fq2) ? 2 : 1); }

std::mutex g_mu; // guards g_sets
// Lifetime of the sets' device arrays: every MSM over a resident set holds this shared for the whole (synchronous)
// call, b200_bases_free and b200_bases_precompute — which free or replace the arrays — hold it exclusively.
// Lock order everywhere: g_life, then the slot's mutex, then g_mu.
std::shared_mutex g_life;
std::map<uint64_t, rt::BaseSet> g_sets;
uint64_t g_next_handle = 1;

// Per execution slot (runtime.hpp): the buffers of an MSM call and the lock that serialises the calls of one slot.
struct SlotState {
    std::mutex mu;
    rt::DevBuf scalars, windows, out;
};
SlotState g_slot[rt::NUM_SLOTS];
#define g_scalars g_slot[rt::slot()].scalars
#define g_windows g_slot[rt::slot()].windows
#define g_out g_slot[rt::slot()].out

// A copy of the set's descriptor (the device arrays stay owned by the handle; the caller must not free a set
// that an MSM is still running over).
bool find_set(b200_bases_t handle, rt::BaseSet *out)
{
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_sets.find(handle);
    if (it == g_sets.end()) {
        rt::set_error(B200_ERR_HANDLE, "unknown base-set handle %llu", (unsigned long long)handle);
        return false;
    }
    *out = it->second;
    return true;
}

// The reference's zero(): (0, one, 0).
void write_zero_point(int curve, int group, void *out)
{
    const CurveVt *v = vt(curve);
    const size_t fq = v->fq_bytes;
    const size_t coord = group_bytes(v, group) / 3;
    memset(out, 0, 3 * coord);
    uint32_t *y = reinterpret_cast<uint32_t *>(static_cast<uint8_t *>(out) + coord);
    for (size_t i = 0; i < fq / 4; ++i) {
        switch (curve) {
        case B200_ALT_BN128: y[i] = alt_bn128_fq_params::one((int)i); break;
        case B200_BLS12_377: y[i] = bls12_377_fq_params::one((int)i); break;
        default: y[i] = bw6_761_fq_params::one((int)i); break;
        }
    }
}

int msm_resident_impl(const rt::BaseSet &bs, size_t first, size_t n, const void *scalars, int on_device,
                      void *out, void *out_windows, void *dev_out_windows = nullptr, const rt::SharedSort *shared = nullptr,
                      uint32_t skip = 0)
{
    const CurveVt *v = vt(bs.curve);
    if (first + n > bs.n) {
        return rt::set_error(B200_ERR_ARG, "range [%zu, %zu) exceeds the %zu resident bases", first, first + n, bs.n);
    }
    const int bits = v->scalar_bits();
    if (n == 0) {
        if (out) {
            write_zero_point(bs.curve, bs.group, out);
        }
        if (out_windows) {
            return rt::set_error(B200_ERR_ARG, "windows of an empty MSM are undefined");
        }
        return B200_OK;
    }
    (void)bits;
    cudaStream_t s = rt::stream();
    const void *sdev = scalars;
    if (!on_device) {
        if (g_scalars.reserve(n * v->fr_bytes)) {
            return B200_ERR_CUDA;
        }
        if (const int rc = rt::staged_h2d(g_scalars.p, scalars, n * v->fr_bytes, s)) {
            return rc;
        }
        sdev = g_scalars.p;
    }
    const size_t wbytes = v->window_point_bytes(bs.group);
    if (g_windows.reserve(64 * wbytes * 2) || g_out.reserve(1024)) {
        return B200_ERR_CUDA;
    }
    int W = 0, c = 0;
    int rc = v->grp(bs.group).msm_windows(bs, first, n, sdev, g_windows.p, &W, &c, shared, skip);
    if (rc) {
        return rc;
    }
    if (out_windows) {
        B200_CUDA_CHECK(cudaMemcpyAsync(out_windows, g_windows.p, (size_t)W * wbytes, cudaMemcpyDeviceToHost, s));
    }
    if (dev_out_windows) {
        B200_CUDA_CHECK(cudaMemcpyAsync(dev_out_windows, g_windows.p, (size_t)W * wbytes, cudaMemcpyDeviceToDevice, s));
    }
    if (out) {
        rc = v->grp(bs.group).msm_fold(g_windows.p, 1, W, c, 1, out);
        if (rc) {
            return rc;
        }
    } else {
        // window sums only (sharded callers fold later): close the timing record here
        rt::MsmTiming &tm = rt::msm_timing();
        B200_CUDA_CHECK(cudaEventRecord(tm.ev[4], s));
        tm.host_fold_ms = 0;
        tm.valid = true;
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(s));
    return B200_OK;
}

void free_set(rt::BaseSet &bs)
{
    if (bs.pts) {
        cudaFree(bs.pts);
    }
    if (bs.inf) {
        cudaFree(bs.inf);
    }
    bs = rt::BaseSet();
}

template<int N> bool limbs_equal_one(const uint32_t *s, uint32_t (*one)(int))
{
    for (int i = 0; i < N; ++i) {
        if (s[i] != one(i)) {
            return false;
        }
    }
    return true;
}

void scalar_stats(int curve, const void *scalars, size_t n, size_t *stats)
{
    const CurveVt *v = vt(curve);
    const size_t nl = v->fr_bytes / 4;
    uint32_t one[12];
    for (size_t i = 0; i < nl; ++i) {
        one[i] = curve == B200_ALT_BN128   ? alt_bn128_fr_params::one((int)i)
                 : curve == B200_BLS12_377 ? bls12_377_fr_params::one((int)i)
                                           : bw6_761_fr_params::one((int)i);
    }
    size_t zeros = 0, ones = 0;
    const uint32_t *s = static_cast<const uint32_t *>(scalars);
    for (size_t i = 0; i < n; ++i, s += nl) {
        uint32_t orr = 0, diff = 0;
        for (size_t j = 0; j < nl; ++j) {
            orr |= s[j];
            diff |= s[j] ^ one[j];
        }
        zeros += orr == 0;
        ones += diff == 0;
    }
    stats[0] = zeros;
    stats[1] = ones;
    stats[2] = n - zeros - ones;
}

} // namespace

namespace b200
{
namespace rt
{

bool msm_shared_eligible(const b200_bases_t sets[4], const size_t n[4], uint32_t skip)
{
    BaseSet bs[4];
    for (int k = 0; k < 4; ++k) {
        if (!find_set(sets[k], &bs[k])) {
            return false;
        }
    }
    for (int k = 0; k < 4; ++k) {
        if (bs[k].curve != bs[0].curve || !bs[k].table_c || bs[k].table_c != bs[0].table_c || bs[k].inf || n[k] > bs[k].n ||
            n[k] == 0) {
            return false;
        }
    }
    return n[1] == n[0] && n[2] == n[0] && n[3] + skip == n[0];
}

int msm_shared_sort(b200_bases_t set, size_t n, const void *scalars_dev, SharedSort *out, cudaStream_t s)
{
    std::shared_lock<std::shared_mutex> life(g_life);
    BaseSet bs;
    if (!find_set(set, &bs)) {
        return B200_ERR_HANDLE;
    }
    return vt(bs.curve)->grp(bs.group).msm_sort(bs, n, scalars_dev, out, s);
}

int msm_resident_shared(b200_bases_t handle, size_t n, const void *scalars_dev, void *out_host, void *dev_out_windows,
                        const SharedSort *shared, uint32_t skip)
{
    std::shared_lock<std::shared_mutex> life(g_life);
    BaseSet bs;
    if (!find_set(handle, &bs)) {
        return B200_ERR_HANDLE;
    }
    if ((!out_host && !dev_out_windows) || !scalars_dev || n == 0) {
        return set_error(B200_ERR_ARG, "null pointer or empty range");
    }
    std::lock_guard<std::mutex> lk(g_slot[rt::slot()].mu);
    return msm_resident_impl(bs, 0, n, scalars_dev, 1, out_host, nullptr, dev_out_windows, shared, skip);
}

} // namespace rt
} // namespace b200

extern "C" {

size_t b200_sizeof_fr(int curve)
{
    const CurveVt *v = vt(curve);
    return v ? v->fr_bytes : 0;
}

size_t b200_sizeof_group(int curve, int group)
{
    const CurveVt *v = vt(curve);
    if (!v || (group != B200_G1 && group != B200_G2)) {
        return 0;
    }
    return group_bytes(v, group);
}

int b200_bases_upload(int curve, int group, const void *bases, size_t base_stride, size_t n, b200_bases_t *handle)
{
    const CurveVt *v = vt(curve);
    if (!v) {
        return B200_ERR_ARG;
    }
    if ((group != B200_G1 && group != B200_G2) || !handle || (n && !bases)) {
        return rt::set_error(B200_ERR_ARG, "bad group / null pointer");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    std::lock_guard<std::mutex> slot_lk(g_slot[rt::slot()].mu); // the upload stages through the slot's workspace
    std::lock_guard<std::mutex> lk(g_mu);
    rt::BaseSet bs;
    const int rc = v->grp(group).bases_upload(curve, group, bases, base_stride, n, &bs);
    if (rc) {
        return rc;
    }
    *handle = g_next_handle++;
    g_sets[*handle] = bs;
    return B200_OK;
}

int b200_bases_free(b200_bases_t handle)
{
    std::unique_lock<std::shared_mutex> life(g_life);
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_sets.find(handle);
    if (it == g_sets.end()) {
        return rt::set_error(B200_ERR_HANDLE, "unknown base-set handle %llu", (unsigned long long)handle);
    }
    cudaDeviceSynchronize(); // an MSM of any slot may still be reading the set
    free_set(it->second);
    g_sets.erase(it);
    return B200_OK;
}

int b200_bases_precompute(b200_bases_t handle)
{
    std::unique_lock<std::shared_mutex> life(g_life); // the set's point array is replaced: no MSM may be reading it
    std::lock_guard<std::mutex> slot_lk(g_slot[rt::slot()].mu);
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_sets.find(handle);
    if (it == g_sets.end()) {
        return rt::set_error(B200_ERR_HANDLE, "unknown base-set handle %llu", (unsigned long long)handle);
    }
    const CurveVt *v = vt(it->second.curve);
    return v->grp(it->second.group).bases_precompute(&it->second);
}

int b200_bases_table_rows(b200_bases_t handle)
{
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_sets.find(handle);
    return it == g_sets.end() ? 0 : it->second.table_W;
}

int b200_bases_table_window_bits(b200_bases_t handle)
{
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_sets.find(handle);
    return it == g_sets.end() ? 0 : it->second.table_c;
}

size_t b200_bases_size(b200_bases_t handle)
{
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_sets.find(handle);
    return it == g_sets.end() ? 0 : it->second.n;
}

int b200_msm_resident(b200_bases_t handle, size_t first, size_t n, const void *scalars, int scalars_on_device,
                      void *out)
{
    std::shared_lock<std::shared_mutex> life(g_life);
    rt::BaseSet bs;
    if (!find_set(handle, &bs)) {
        return B200_ERR_HANDLE;
    }
    if (!out || (n && !scalars)) {
        return rt::set_error(B200_ERR_ARG, "null pointer");
    }
    std::lock_guard<std::mutex> lk(g_slot[rt::slot()].mu);
    return msm_resident_impl(bs, first, n, scalars, scalars_on_device, out, nullptr);
}

int b200_msm_num_windows(int curve, size_t n)
{
    const CurveVt *v = vt(curve);
    if (!v) {
        return B200_ERR_ARG;
    }
    return msm::make_geometry(n, v->scalar_bits()).W;
}

int b200_msm_window_bits(int curve, size_t n)
{
    const CurveVt *v = vt(curve);
    if (!v) {
        return B200_ERR_ARG;
    }
    return msm::make_geometry(n, v->scalar_bits()).c;
}

size_t b200_msm_window_bytes(int curve, int group)
{
    const CurveVt *v = vt(curve);
    return v ? v->window_point_bytes(group) : 0;
}

int b200_msm_resident_windows(b200_bases_t handle, size_t first, size_t n, const void *scalars,
                              int scalars_on_device, void *out_windows)
{
    std::shared_lock<std::shared_mutex> life(g_life);
    rt::BaseSet bs;
    if (!find_set(handle, &bs)) {
        return B200_ERR_HANDLE;
    }
    if (!out_windows || !scalars || n == 0) {
        return rt::set_error(B200_ERR_ARG, "null pointer or empty range");
    }
    std::lock_guard<std::mutex> lk(g_slot[rt::slot()].mu);
    return msm_resident_impl(bs, first, n, scalars, scalars_on_device, nullptr, out_windows);
}

int b200_msm_resident_windows_dev(b200_bases_t handle, size_t first, size_t n, const void *scalars,
                                  int scalars_on_device, void *dev_out_windows)
{
    std::shared_lock<std::shared_mutex> life(g_life);
    rt::BaseSet bs;
    if (!find_set(handle, &bs)) {
        return B200_ERR_HANDLE;
    }
    if (!dev_out_windows || !scalars || n == 0) {
        return rt::set_error(B200_ERR_ARG, "null pointer or empty range");
    }
    std::lock_guard<std::mutex> lk(g_slot[rt::slot()].mu);
    return msm_resident_impl(bs, first, n, scalars, scalars_on_device, nullptr, nullptr, dev_out_windows);
}

int b200_msm_fold_windows(int curve, int group, size_t n_per_shard, const void *windows, size_t num_shards, void *out)
{
    const CurveVt *v = vt(curve);
    if (!v) {
        return B200_ERR_ARG;
    }
    if (!windows || !out || num_shards == 0 || (group != B200_G1 && group != B200_G2)) {
        return rt::set_error(B200_ERR_ARG, "bad argument");
    }
    const msm::Geometry geo = msm::make_geometry(n_per_shard, v->scalar_bits());
    return v->grp(group).msm_fold(windows, 0, geo.W, geo.c, num_shards, out);
}

int b200_msm(int curve, int group, const void *bases, size_t base_stride, size_t n, const void *scalars, void *out)
{
    const CurveVt *v = vt(curve);
    if (!v) {
        return B200_ERR_ARG;
    }
    if ((group != B200_G1 && group != B200_G2) || !out || (n && (!bases || !scalars))) {
        return rt::set_error(B200_ERR_ARG, "bad group / null pointer");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    if (n == 0) {
        write_zero_point(curve, group, out);
        return B200_OK;
    }
    std::lock_guard<std::mutex> lk(g_slot[rt::slot()].mu);
    rt::BaseSet bs;
    int rc = v->grp(group).bases_upload(curve, group, bases, base_stride, n, &bs);
    if (rc) {
        return rc;
    }
    rc = msm_resident_impl(bs, 0, n, scalars, 0, out, nullptr);
    cudaStreamSynchronize(rt::stream());
    free_set(bs);
    return rc;
}

int b200_msm_filter_one_zero(int curve, int group, const void *bases, size_t base_stride, size_t n,
                             const void *scalars, void *out, size_t *stats)
{
    if (stats && vt(curve) && scalars) {
        scalar_stats(curve, scalars, n, stats);
    }
    return b200_msm(curve, group, bases, base_stride, n, scalars, out);
}

int b200_kc_msm(int curve, const size_t *indices, const void *values, size_t nnz, size_t min_idx, size_t max_idx,
                const void *scalars, size_t num_scalars, void *out_g2, void *out_g1, size_t *stats)
{
    const CurveVt *v = vt(curve);
    if (!v) {
        return B200_ERR_ARG;
    }
    if (!out_g2 || !out_g1 || (nnz && (!indices || !values)) || max_idx < min_idx ||
        max_idx - min_idx > num_scalars || (num_scalars && !scalars)) {
        return rt::set_error(B200_ERR_ARG, "bad argument");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    // kc_multiexp.tcc:40-44 / :63: entries with min_idx <= index < max_idx
    const size_t *lo = std::lower_bound(indices, indices + nnz, min_idx);
    const size_t *hi = std::lower_bound(lo, indices + nnz, max_idx);
    const size_t first = (size_t)(lo - indices), cnt = (size_t)(hi - lo);
    const size_t g2b = group_bytes(v, B200_G2), g1b = group_bytes(v, B200_G1), stride = g2b + g1b;
    std::vector<uint8_t> gathered(cnt * v->fr_bytes);
    const uint8_t *sc = static_cast<const uint8_t *>(scalars);
    for (size_t k = 0; k < cnt; ++k) {
        memcpy(gathered.data() + k * v->fr_bytes, sc + (indices[first + k] - min_idx) * v->fr_bytes, v->fr_bytes);
    }
    if (stats) {
        scalar_stats(curve, gathered.data(), cnt, stats);
    }
    const uint8_t *vals = static_cast<const uint8_t *>(values) + first * stride;
    int rc = b200_msm(curve, B200_G2, vals, stride, cnt, gathered.data(), out_g2);
    if (rc) {
        return rc;
    }
    return b200_msm(curve, B200_G1, vals + g2b, stride, cnt, gathered.data(), out_g1);
}

} // extern "C"
