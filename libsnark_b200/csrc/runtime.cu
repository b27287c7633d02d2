// runtime.cu — device binding, stream, error string, workspace buffers, device-memory C-ABI.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "runtime.hpp"

namespace b200
{
namespace rt
{

namespace
{
thread_local char g_err[512] = "";
int g_device = -1;
cudaStream_t g_streams[NUM_SLOTS] = {};
#define g_stream g_streams[0]
thread_local int g_slot = 0;
std::atomic<uint64_t> g_launches{0};
MsmTiming g_timing[NUM_SLOTS];
cudaEvent_t g_timer[2] = {nullptr, nullptr};
} // namespace

int set_error(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
const char *last_error() { return g_err; }

static int do_init(int device)
{
    if (g_device == device && g_stream) {
        return B200_OK;
    }
    if (g_stream) {
        // one process per GPU: the streams, events and every workspace live on the device bound first
        return set_error(B200_ERR_ARG, "b200_init(%d): this process is already bound to device %d (one process per GPU)",
                         device, g_device);
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        return set_error(B200_ERR_CUDA, "no CUDA device: %s (this library has no CPU fallback)",
                         e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= count) {
        return set_error(B200_ERR_ARG, "device %d out of range [0, %d)", device, count);
    }
    B200_CUDA_CHECK(cudaSetDevice(device));
    if (!g_stream) {
        int least = 0, greatest = 0;
        B200_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        for (int k = 0; k < NUM_SLOTS; ++k) {
            // numerically lower = higher priority: slot 0 highest, slot 1 next (the prover's G2 MSM, whose long
            // bucket reduction should not be the last thing running), the others lowest
            const int prio = k == 0 ? greatest : (k == 1 && greatest + 1 <= least) ? greatest + 1 : least;
            B200_CUDA_CHECK(cudaStreamCreateWithPriority(&g_streams[k], cudaStreamNonBlocking, prio));
            for (int i = 0; i < 5; ++i) {
                B200_CUDA_CHECK(cudaEventCreate(&g_timing[k].ev[i]));
            }
        }
    }
    g_device = device;
    return B200_OK;
}

// Lazy binding (a compute entry point called before b200_init): the CUDA current device of the calling thread,
// which is 0 unless the caller selected another one.
bool ensure_init()
{
    if (g_stream && g_device >= 0) {
        return true;
    }
    int cur = 0;
    if (cudaGetDevice(&cur) != cudaSuccess) {
        cur = 0;
    }
    return do_init(cur) == B200_OK;
}
cudaStream_t stream() { return g_streams[g_slot]; }
cudaStream_t slot_stream(int slot) { return g_streams[slot < 0 || slot >= NUM_SLOTS ? 0 : slot]; }
int slot() { return g_slot; }
int bind_thread(int slot)
{
    if (slot < 0 || slot >= NUM_SLOTS) {
        return set_error(B200_ERR_ARG, "slot %d outside [0, %d)", slot, NUM_SLOTS);
    }
    if (!ensure_init()) {
        return B200_ERR_CUDA;
    }
    B200_CUDA_CHECK(cudaSetDevice(g_device));
    g_slot = slot;
    return B200_OK;
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
MsmTiming &msm_timing() { return g_timing[g_slot]; }

int DevBuf::reserve(size_t bytes)
{
    if (bytes <= cap) {
        return B200_OK;
    }
    if (p) {
        // stream-ordered users of the old buffer must be done before it is freed
        cudaStreamSynchronize(stream());
        cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    const size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) {
        p = nullptr;
        return set_error(B200_ERR_CUDA, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
    }
    cap = want;
    return B200_OK;
}
void DevBuf::release()
{
    if (p) {
        cudaFree(p);
    }
    p = nullptr;
    cap = 0;
}

namespace
{
struct PinnedBuf {
    void *p = nullptr;
    size_t cap = 0;
};
PinnedBuf g_pinned[NUM_SLOTS];
constexpr size_t STAGE_MIN = size_t(4) << 20;    // below this a pageable copy is as good
constexpr size_t STAGE_CAP = size_t(128) << 20;  // page-locked bytes held per slot that stages large copies
constexpr size_t STAGE_THREADS = 4;
} // namespace

int staged_h2d(void *dev_dst, const void *host_src, size_t bytes, cudaStream_t s)
{
    if (bytes == 0) {
        return B200_OK;
    }
    bool direct = bytes < STAGE_MIN;
    if (!direct) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, host_src) == cudaSuccess && at.type == cudaMemoryTypeHost) {
            direct = true;
        }
        cudaGetLastError(); // an unregistered pointer may leave an error behind on old runtimes
    }
    PinnedBuf &pb = g_pinned[slot()];
    if (!direct && pb.cap < (bytes < STAGE_CAP ? bytes : STAGE_CAP)) {
        if (pb.p) {
            cudaFreeHost(pb.p);
            pb.p = nullptr;
            pb.cap = 0;
        }
        const size_t want = bytes < STAGE_CAP ? bytes + bytes / 8 : STAGE_CAP;
        if (cudaHostAlloc(&pb.p, want, cudaHostAllocDefault) != cudaSuccess) {
            cudaGetLastError();
            pb.p = nullptr;
            direct = true; // no page-locked memory to be had: the plain copy still works
        } else {
            pb.cap = want;
        }
    }
    if (direct) { // stream-ordered, as before: the caller's kernels queue up behind it
        B200_CUDA_CHECK(cudaMemcpyAsync(dev_dst, host_src, bytes, cudaMemcpyHostToDevice, s));
        return B200_OK;
    }
    const uint8_t *src = static_cast<const uint8_t *>(host_src);
    uint8_t *dst = static_cast<uint8_t *>(dev_dst), *pin = static_cast<uint8_t *>(pb.p);
    for (size_t done = 0; done < bytes;) {
        const size_t part = bytes - done < pb.cap ? bytes - done : pb.cap;
        size_t cut[STAGE_THREADS + 1];
        for (size_t t = 0; t <= STAGE_THREADS; ++t) {
            cut[t] = t == STAGE_THREADS ? part : (part / STAGE_THREADS * t) & ~(size_t)255;
        }
        std::vector<std::thread> workers;
        for (size_t t = 1; t < STAGE_THREADS; ++t) {
            workers.emplace_back([=]() { memcpy(pin + cut[t], src + done + cut[t], cut[t + 1] - cut[t]); });
        }
        memcpy(pin, src + done, cut[1]);
        cudaError_t e = cudaMemcpyAsync(dst + done, pin, cut[1], cudaMemcpyHostToDevice, s);
        for (size_t t = 1; t < STAGE_THREADS; ++t) {
            workers[t - 1].join();
            if (e == cudaSuccess) {
                e = cudaMemcpyAsync(dst + done + cut[t], pin + cut[t], cut[t + 1] - cut[t], cudaMemcpyHostToDevice, s);
            }
        }
        if (e == cudaSuccess) {
            e = cudaStreamSynchronize(s); // the staging buffer is about to be overwritten / handed to the next call
        }
        if (e != cudaSuccess) {
            return set_error(B200_ERR_CUDA, "staged host-to-device copy failed: %s", cudaGetErrorString(e));
        }
        done += part;
    }
    return B200_OK;
}

} // namespace rt
} // namespace b200

using namespace b200;

extern "C" {

int b200_version(void) { return 100; }

int b200_init(int device) { return rt::do_init(device); }

int b200_device_count(void)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) {
        return 0;
    }
    return count;
}

const char *b200_last_error(void) { return rt::last_error(); }

int b200_dev_alloc(size_t bytes, void **dev_ptr)
{
    if (!dev_ptr) {
        return rt::set_error(B200_ERR_ARG, "null dev_ptr");
    }
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    B200_CUDA_CHECK(cudaMalloc(dev_ptr, bytes ? bytes : 16));
    return B200_OK;
}

int b200_dev_free(void *dev_ptr)
{
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    B200_CUDA_CHECK(cudaFree(dev_ptr));
    return B200_OK;
}

int b200_memcpy_h2d(void *dev_dst, const void *host_src, size_t bytes)
{
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(dev_dst, host_src, bytes, cudaMemcpyHostToDevice, rt::stream()));
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

int b200_memcpy_d2h(void *host_dst, const void *dev_src, size_t bytes)
{
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    B200_CUDA_CHECK(cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, rt::stream()));
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

int b200_sync(void)
{
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    return B200_OK;
}

int b200_timer_start(void)
{
    if (!rt::ensure_init()) {
        return B200_ERR_CUDA;
    }
    if (!rt::g_timer[0]) {
        B200_CUDA_CHECK(cudaEventCreate(&rt::g_timer[0]));
        B200_CUDA_CHECK(cudaEventCreate(&rt::g_timer[1]));
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(rt::stream()));
    B200_CUDA_CHECK(cudaEventRecord(rt::g_timer[0], rt::stream()));
    return B200_OK;
}

int b200_timer_stop(double *ms)
{
    if (!ms || !rt::g_timer[0]) {
        return rt::set_error(B200_ERR_ARG, "timer not started");
    }
    B200_CUDA_CHECK(cudaEventRecord(rt::g_timer[1], rt::stream()));
    B200_CUDA_CHECK(cudaEventSynchronize(rt::g_timer[1]));
    float f = 0;
    B200_CUDA_CHECK(cudaEventElapsedTime(&f, rt::g_timer[0], rt::g_timer[1]));
    *ms = f;
    return B200_OK;
}

uint64_t b200_launch_count(void) { return rt::g_launches.load(); }

int b200_last_msm_timing(double *ms5)
{
    rt::MsmTiming &tm = rt::msm_timing();
    if (!ms5 || !tm.valid) {
        return rt::set_error(B200_ERR_ARG, "no MSM has run yet");
    }
    B200_CUDA_CHECK(cudaEventSynchronize(tm.ev[4]));
    for (int i = 0; i < 3; ++i) {
        float ms = 0;
        B200_CUDA_CHECK(cudaEventElapsedTime(&ms, tm.ev[i], tm.ev[i + 1]));
        ms5[i] = ms;
    }
    float tot = 0;
    B200_CUDA_CHECK(cudaEventElapsedTime(&tot, tm.ev[0], tm.ev[4]));
    ms5[3] = tm.host_fold_ms;
    ms5[4] = tot + tm.host_fold_ms;
    return B200_OK;
}

} // extern "C"
