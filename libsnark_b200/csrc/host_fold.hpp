// host_fold.hpp — the O(scalar bits) tail of an MSM on the host: sum the per-window results of
// all shards, fold them by Horner's rule (result = 2^c * result + R_w, the reference's
// multiexp.tcc:602-629 loop), normalise to affine (to_affine_coordinates, alt_bn128_g1.cpp:68-82).
//
// This is ~(bits + W) group operations on ONE point: a serial dependency chain that a single
// GPU thread runs ~10x slower than a CPU core, so it stays on the host like the reference's own
// r * delta_g1 scalar multiplications. It is not a fallback for any device
// work: the buckets, their reduction and the window sums all come from the GPU.
#pragma once
#include "curves.cuh"
#include "msm.cuh"

namespace b200
{
namespace msm
{

template<class F> inline RefPoint<F> host_fold(const XYZZ<F> *windows, int W, int c, size_t num_shards)
{
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int w = W - 1; w >= 0; --w) {
        if (!acc.is_inf()) {
            for (int k = 0; k < c; ++k) {
                dbl(acc);
            }
        }
        for (size_t s = 0; s < num_shards; ++s) {
            // over precomputed tables only window 0 carries a sum; adding the point at infinity would still cost a
            // whole addition formula on the host (W * shards of them: 1.5 ms of an 8-GPU proof's assembly)
            if (!windows[s * W + w].is_inf()) {
                add(acc, windows[s * W + w]);
            }
        }
    }
    return normalise(acc);
}

} // namespace msm
} // namespace b200
