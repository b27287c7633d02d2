// runtime.hpp — process-wide state of the library (one process per GPU): the bound device,
// the compute stream, the last-error string, the launch counter and grow-only device buffers.
#pragma once
#include <cstdarg>
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/libsnark_b200.h"
#include "msm_geometry.hpp"

namespace b200
{
namespace rt
{

int set_error(int code, const char *fmt, ...);
const char *last_error();
bool ensure_init();         // true when a device is bound (b200_init called or lazily device 0)
// Execution slots: a slot = one stream + one set of MSM workspaces + one timing record. A host thread works in
// exactly one slot (0 unless it called bind_thread); the prover runs its independent MSMs in different slots on
// different host threads so that the latency-bound stages of one (sort, bucket reduction, host fold) overlap the
// integer-pipe-bound bucket accumulation of another. Slot 0 has the highest stream priority (it carries the
// witness map, which the H MSM waits for).
constexpr int NUM_SLOTS = 6;
int slot();                 // the calling host thread's slot
int bind_thread(int slot);  // call first in a new host thread: selects the slot and the library's device
cudaStream_t stream();      // the compute stream of the calling thread's slot
cudaStream_t slot_stream(int slot);
void count_launch(int n = 1);

#define B200_CUDA_CHECK(expr)                                                                      \
    do {                                                                                           \
        cudaError_t e__ = (expr);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            return b200::rt::set_error(B200_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,             \
                                       cudaGetErrorString(e__), __FILE__, __LINE__);               \
        }                                                                                          \
    } while (0)

#define B200_LAUNCH_CHECK()                                                                        \
    do {                                                                                           \
        b200::rt::count_launch();                                                                  \
        cudaError_t e__ = cudaGetLastError();                                                      \
        if (e__ != cudaSuccess) {                                                                  \
            return b200::rt::set_error(B200_ERR_CUDA, "kernel launch failed: %s (%s:%d)",          \
                                       cudaGetErrorString(e__), __FILE__, __LINE__);               \
        }                                                                                          \
    } while (0)

// Grow-only device buffer (cudaMalloc is a synchronising, millisecond-scale call: workspaces
// are kept between calls).
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes);
    void release();
    template<class T> T *as() { return static_cast<T *>(p); }
};

// Host-to-device copy of a buffer that may lie in PAGEABLE memory (the reference hands over std::vector storage). The
// driver stages a pageable cudaMemcpyAsync through one bounce buffer on one core — 5-12 GB/s depending on the box; this
// stages it through a page-locked buffer of the calling slot with a few host threads and sends every slice on as soon as
// it is staged, returning when the copy is complete (the staging buffer is reused by the next call). Page-locked sources
// (cudaHostAlloc / cudaHostRegister) and small copies are one stream-ordered cudaMemcpyAsync, not waited for.
int staged_h2d(void *dev_dst, const void *host_src, size_t bytes, cudaStream_t s);

// A device-resident base set (proving-key query) in device affine form.
struct BaseSet {
    int curve = -1, group = 0;
    size_t n = 0;
    size_t point_bytes = 0; // sizeof(Affine<F>)
    void *pts = nullptr;    // Affine<F>[n]
    uint8_t *inf = nullptr; // n flags, or nullptr when no base is the point at infinity
    // b200_bases_precompute: pts then holds table_W rows of n points, row w = [2^(table_c w)] P (msm.cuh)
    int table_c = 0, table_W = 0;
};

// Stages 1-4 of an MSM (digits, counting sort, tasks: everything that depends on the scalars alone) computed once and
// read by several MSMs over different base sets: the prover's A, B.g2, B.g1 and L queries all take the witness as their
// scalars (r1cs_gg_ppzksnark.tcc:539-581), so one sort serves four MSMs. Filled by <group>_msm_sort, consumed through
// the `shared` argument of <group>_msm_windows. The buffers belong to this object; `done` is recorded on the sorting
// stream after the last sort kernel, and every consumer's stream waits for it.
struct SharedSort {
    DevBuf t, counts, offsets, cursor, task_off, tile_sums, entries, tasks, multi;
    msm::Geometry g{};   // table geometry incl. rounds and eshift; g.n = the scalar count
    size_t max_entries = 0, max_tasks = 0;
    int curve = -1;
    cudaEvent_t done = nullptr;
};

struct MsmTiming {
    cudaEvent_t ev[5]; // start | sorted | accumulated | reduced | windows on host
    double host_fold_ms = 0;
    bool valid = false;
};
MsmTiming &msm_timing();

// capi_msm.cu, for the prover (groth16.cu). sets = {A, B.g2, B.g1, L}, n = their MSM lengths: true when one sort of the
// n[0] witness scalars can serve all four (same curve, tables of one window width, no base at infinity, B dense,
// n[3] = n[0] - skip). msm_shared_sort runs it on stream s over the geometry of sets[0];
// msm_resident_shared is b200_msm_resident / b200_msm_resident_windows_dev over such a sort.
bool msm_shared_eligible(const b200_bases_t sets[4], const size_t n[4], uint32_t skip);
int msm_shared_sort(b200_bases_t set, size_t n, const void *scalars_dev, SharedSort *out, cudaStream_t s);
// out_host (one folded point, as b200_msm_resident) or dev_out_windows (as b200_msm_resident_windows_dev), the other null
int msm_resident_shared(b200_bases_t handle, size_t n, const void *scalars_dev, void *out_host, void *dev_out_windows,
                        const SharedSort *shared, uint32_t skip);

} // namespace rt

// Implemented once per (curve, group) in msm_<curve>_<group>.cu. bw6_761's G2 is over Fq
// like its G1 (bw6_761_g2.hpp), so both groups share the g1 instance.
#define B200_DECLARE_GROUP_ENTRY(name)                                                             \
    int name##_bases_upload(int curve, int group, const void *host, size_t stride, size_t n,      \
                            rt::BaseSet *out);                                                     \
    int name##_bases_precompute(rt::BaseSet *bs);                                                  \
    int name##_msm_windows(const rt::BaseSet &bs, size_t first, size_t n, const void *scalars_dev, \
                           void *windows_dev, int *W, int *c, const rt::SharedSort *shared,       \
                           uint32_t skip);                                                         \
    int name##_msm_sort(const rt::BaseSet &bs, size_t n, const void *scalars_dev,                  \
                        rt::SharedSort *out, cudaStream_t s);                                      \
    int name##_msm_fold(const void *windows, int on_device, int W, int c, size_t shards, void *out); \
    size_t name##_window_point_bytes();                                                            \
    int name##_scalar_bits();

B200_DECLARE_GROUP_ENTRY(alt_bn128_g1)
B200_DECLARE_GROUP_ENTRY(alt_bn128_g2)
B200_DECLARE_GROUP_ENTRY(bls12_377_g1)
B200_DECLARE_GROUP_ENTRY(bls12_377_g2)
B200_DECLARE_GROUP_ENTRY(bw6_761_g1)

} // namespace b200
