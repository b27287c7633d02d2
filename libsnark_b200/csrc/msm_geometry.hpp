// msm_geometry.hpp — window / bucket geometry of the MSM (host-only arithmetic, shared by the
// kernels' launch code and the C-ABI's b200_msm_num_windows).
#pragma once
#include <cstdint>

namespace b200
{
namespace msm
{

constexpr int SEG = 64;   // max entries per accumulation task
constexpr int CHUNK = 16; // buckets per reduce thread

struct Geometry {
    uint32_t n;   // points
    int c;        // window width (bits)
    int W;        // number of windows
    uint32_t B;   // buckets per window = 2^(c-1)
    uint32_t NB;  // W * B
    uint32_t cpw; // chunks per window = B / CHUNK
};

// Window width: minimise  W(c) * (10 n + 42 * 2^(c-1))  field mults (10 per mixed add,
// ~42 per bucket in the reduction), c in [5, 20].
inline Geometry make_geometry(uint64_t n, int scalar_bits, int force_c = 0)
{
    Geometry g;
    g.n = (uint32_t)n;
    int best_c = 5;
    double best = 1e300;
    for (int c = 5; c <= 20; ++c) {
        const int W = (scalar_bits + 1 + c - 1) / c;
        const double cost = (double)W * (10.0 * (double)n + 42.0 * (double)(1u << (c - 1)));
        if (cost < best) {
            best = cost;
            best_c = c;
        }
    }
    g.c = force_c ? force_c : best_c;
    g.W = (scalar_bits + 1 + g.c - 1) / g.c;
    g.B = 1u << (g.c - 1);
    g.NB = (uint32_t)g.W * g.B;
    g.cpw = g.B / CHUNK;
    return g;
}

} // namespace msm
} // namespace b200
