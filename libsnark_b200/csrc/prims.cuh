// prims.cuh — 32-bit carry-chain primitives.
//
// Device: inline PTX (add.cc / addc / mad.lo.cc / madc.hi.cc ...). ptxas fuses a
// mad.lo.cc + madc.hi.cc pair on the same operands into one IMAD.WIDE.U32 with
// carry, which is what the Montgomery kernels are built from.
// Host (no __CUDA_ARCH__): a bit-exact emulation with a thread-local carry flag,
// used ONLY by the CPU unit tests of the field / curve templates (tests/host_emu)
// — it is not a product path: every product entry point launches CUDA kernels.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define HD __host__ __device__ __forceinline__
#define DEV __device__ __forceinline__
#else
#define HD inline
#define DEV inline
#endif

namespace b200
{
namespace prim
{

#if defined(__CUDA_ARCH__)

DEV uint32_t add_cc(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
DEV uint32_t addc_cc(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
DEV uint32_t addc(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
DEV uint32_t sub_cc(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
DEV uint32_t subc_cc(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
DEV uint32_t subc(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
DEV uint32_t mul_lo(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
DEV uint32_t mul_hi(uint32_t a, uint32_t b)
{
    uint32_t r;
    asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
DEV uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
DEV uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
DEV uint32_t mad_hi_cc(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm volatile("mad.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
DEV uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
DEV uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

#else // ---------------- host emulation (unit tests only) -------------------

inline uint32_t &cf_()
{
    static thread_local uint32_t cf = 0;
    return cf;
}
inline uint32_t add3_(uint32_t a, uint32_t b, uint32_t cin, bool set)
{
    const uint64_t t = (uint64_t)a + b + cin;
    if (set) {
        cf_() = (uint32_t)(t >> 32);
    }
    return (uint32_t)t;
}
inline uint32_t sub3_(uint32_t a, uint32_t b, uint32_t bin, bool set)
{
    const uint64_t t = (uint64_t)a - b - bin;
    if (set) {
        cf_() = (uint32_t)((t >> 32) & 1);
    }
    return (uint32_t)t;
}
inline uint32_t add_cc(uint32_t a, uint32_t b) { return add3_(a, b, 0, true); }
inline uint32_t addc_cc(uint32_t a, uint32_t b) { return add3_(a, b, cf_(), true); }
inline uint32_t addc(uint32_t a, uint32_t b) { return add3_(a, b, cf_(), false); }
inline uint32_t sub_cc(uint32_t a, uint32_t b) { return sub3_(a, b, 0, true); }
inline uint32_t subc_cc(uint32_t a, uint32_t b) { return sub3_(a, b, cf_(), true); }
inline uint32_t subc(uint32_t a, uint32_t b) { return sub3_(a, b, cf_(), false); }
inline uint32_t mul_lo(uint32_t a, uint32_t b) { return (uint32_t)((uint64_t)a * b); }
inline uint32_t mul_hi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
inline uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { return add3_(mul_lo(a, b), c, 0, true); }
inline uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { return add3_(mul_lo(a, b), c, cf_(), true); }
inline uint32_t mad_hi_cc(uint32_t a, uint32_t b, uint32_t c) { return add3_(mul_hi(a, b), c, 0, true); }
inline uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { return add3_(mul_hi(a, b), c, cf_(), true); }
inline uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { return add3_(mul_hi(a, b), c, cf_(), false); }

#endif

} // namespace prim
} // namespace b200
