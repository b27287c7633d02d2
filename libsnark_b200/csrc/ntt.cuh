// ntt.cuh — radix-2 NTT over Fr, device side.
//
// Replaces libfqfft::basic_radix2_domain<FieldT>::{FFT, iFFT, cosetFFT, icosetFFT}
// (basic_radix2_domain.tcc:54-88) and the kernels under them, _basic_serial_radix2_FFT /
// _basic_parallel_radix2_FFT (basic_radix2_domain_aux.tcc:44-172) and _multiply_by_coset
// (:174-183). Natural order in, natural order out, same values (the transform is a
// function of its input, so any correct algorithm is bit-identical).
//
// Algorithm: a size-n = 2^lgn transform is cut into passes of DEG <= 9 stage bits
// (Stockham autosort: no bit-reversal pass). With p = 2^lgp the product of the radices of the
// earlier passes, R = 2^DEG and t = n / R, pass input x and output y:
//     for index in [0, t), k = index mod p:
//         U[j] = sum_i x[index + i t] * w^i * omega_R^(i j),   w = omega_(pR)^k
//         y[(index - k) R + k + j p] = U[j]
// The inter-pass twiddle w^i is merged into the butterfly twiddles: loading x[index + i t]
// at local row bitrev(i) and running DEG decimation-in-time stages, the butterfly of stage
// s (half-size 2^s) at local offset j < 2^s uses the *global* twiddle
//     omega_(p 2^(s+1))^(k + j p) = omega_n^e,   e = (k + j p) << (lgn - lgp - s - 1) < n/2,
// so a transform costs exactly (n/2) lgn field multiplications and reads one table
// T[e] = omega_n^e, e < n/2 (precomputed in HBM). The inverse uses omega^-e = -T[n/2 - e].
//
// One CTA owns C consecutive `index` values (columns; C * sizeof(Fr) contiguous bytes per
// row on both the gather and the scatter side) and keeps the R x C tile in shared memory.
// Each thread holds 8 elements in registers and runs up to 3 stages (radix-8) between
// shared-memory exchanges. Coset scaling (a[i] *= g^i) and the 1/m of the inverse are
// fused into the first pass's loads / the last pass's stores from precomputed tables.
#pragma once
#include <vector>

#include "curves.cuh"

namespace b200
{
namespace ntt
{

struct alignas(16) Vec4 {
    uint32_t x, y, z, w;
};

struct PassParams {
    int lgn;      // log2 of the transform size
    int lgp;      // log2 of the product of the earlier passes' radices
    int deg;      // stage bits of this pass
    int lgc;      // log2 of the columns per CTA
    int inverse;  // 0: omega, 1: omega^-1
    int nsteps;   // register steps
    int step_q[4];
};

// How a size-2^lgn transform is cut into passes (host side; shared by ntt.cu's launcher and the host emulation of the
// kernel, tests/host_emu): at most NTT_MAX_DEG stage bits per pass, spread evenly; 8 columns per CTA where the size
// allows; register steps of radix 8 (4 + 4 for a remainder of four stage bits).
constexpr int NTT_MAX_DEG = 8;
struct PassLaunch {
    PassParams pp;
    unsigned threads, blocks;
    size_t smem;
};
inline int plan_transform(int lgn, int inverse, int fr_limbs32, PassLaunch *out /* >= 8 entries */)
{
    const int P = (lgn + NTT_MAX_DEG - 1) / NTT_MAX_DEG;
    int lgp = 0;
    for (int i = 0; i < P; ++i) {
        PassParams &pp = out[i].pp;
        pp.lgn = lgn;
        pp.lgp = lgp;
        pp.deg = lgn / P + (i < lgn % P ? 1 : 0);
        pp.lgc = lgn - pp.deg < 3 ? lgn - pp.deg : 3;
        pp.inverse = inverse;
        pp.nsteps = 0;
        for (int rem = pp.deg; rem > 0;) {
            const int q = rem == 4 ? 2 : (rem >= 3 ? 3 : rem);
            pp.step_q[pp.nsteps++] = q;
            rem -= q;
        }
        const uint64_t tile_elems = 1ull << (pp.deg + pp.lgc);
        out[i].threads = (unsigned)(tile_elems / 8);
        if (out[i].threads < 32) {
            out[i].threads = 32;
        }
        out[i].blocks = (unsigned)(1ull << (lgn - pp.deg - pp.lgc));
        out[i].smem = (size_t)(fr_limbs32 / 4) * 16 * ((size_t)1 << pp.deg) * ((1u << pp.lgc) + 1);
        lgp += pp.deg;
    }
    return P;
}

HD uint32_t bitrev(uint32_t v, int bits)
{
    uint32_t r = 0;
    for (int i = 0; i < bits; ++i) {
        r |= ((v >> i) & 1u) << (bits - 1 - i);
    }
    return r;
}

// shared-memory tile: NPL planes of 16-byte quarter-elements, row stride C + 1 (bank spreading)
template<class Fr> struct Tile {
    static constexpr int NPL = Fr::N / 4;
    Vec4 *base;
    uint32_t row_stride; // C + 1
    uint32_t plane;      // R * (C + 1)
    HD Fr load(uint32_t row, uint32_t col) const
    {
        Fr r;
        const uint32_t o = row * row_stride + col;
#pragma unroll
        for (int p = 0; p < NPL; ++p) {
            const Vec4 q = base[p * plane + o];
            r.v[4 * p + 0] = q.x;
            r.v[4 * p + 1] = q.y;
            r.v[4 * p + 2] = q.z;
            r.v[4 * p + 3] = q.w;
        }
        return r;
    }
    HD void store(uint32_t row, uint32_t col, const Fr &v) const
    {
        const uint32_t o = row * row_stride + col;
#pragma unroll
        for (int p = 0; p < NPL; ++p) {
            Vec4 q;
            q.x = v.v[4 * p + 0];
            q.y = v.v[4 * p + 1];
            q.z = v.v[4 * p + 2];
            q.w = v.v[4 * p + 3];
            base[p * plane + o] = q;
        }
    }
};

template<class Fr> HD Fr gload(const Fr *p)
{
    Fr r;
    const Vec4 *s = reinterpret_cast<const Vec4 *>(p);
#pragma unroll
    for (int k = 0; k < Fr::N / 4; ++k) {
        const Vec4 q = s[k];
        r.v[4 * k + 0] = q.x;
        r.v[4 * k + 1] = q.y;
        r.v[4 * k + 2] = q.z;
        r.v[4 * k + 3] = q.w;
    }
    return r;
}
template<class Fr> HD void gstore(Fr *p, const Fr &v)
{
    Vec4 *d = reinterpret_cast<Vec4 *>(p);
#pragma unroll
    for (int k = 0; k < Fr::N / 4; ++k) {
        Vec4 q;
        q.x = v.v[4 * k + 0];
        q.y = v.v[4 * k + 1];
        q.z = v.v[4 * k + 2];
        q.w = v.v[4 * k + 3];
        d[k] = q;
    }
}

// ---- phase A: gather the tile (rows land bit-reversed) ---------------------------------------
template<class Fr>
HD void load_phase(uint32_t tid, uint32_t nthreads, const PassParams &pp, uint64_t index0, const Fr *in,
                   const Fr *pre_table, const Tile<Fr> &tile)
{
    const uint32_t C = 1u << pp.lgc, R = 1u << pp.deg;
    const uint64_t t = 1ull << (pp.lgn - pp.deg);
    for (uint32_t idx = tid; idx < R * C; idx += nthreads) {
        const uint32_t col = idx & (C - 1), i = idx >> pp.lgc;
        const uint64_t gpos = index0 + col + (uint64_t)i * t;
        Fr x = gload(in + gpos);
        if (pre_table != nullptr) {
            x = x * gload(pre_table + gpos);
        }
        tile.store(bitrev(i, pp.deg), col, x);
    }
}

// ---- phase B: Q stages on 2^Q rows {base + a 2^r} held in registers -----------------------------
template<class Fr, int Q>
HD void radix_group(Fr *v, const PassParams &pp, int r, uint32_t lo, uint64_t k, const Fr *twiddles)
{
    const uint64_t half_n = 1ull << (pp.lgn - 1);
#pragma unroll
    for (int q = 0; q < Q; ++q) {
        const int sh = pp.lgn - pp.lgp - (r + q) - 1;
        const uint64_t e0 = (k + ((uint64_t)lo << pp.lgp)) << sh;
#pragma unroll
        for (int jq = 0; jq < (1 << q); ++jq) {
            const uint64_t e = e0 + ((uint64_t)jq << (pp.lgn - q - 1));
            const bool trivial = (e == 0);
            Fr tw;
            if (!trivial) {
                tw = gload(twiddles + (pp.inverse ? half_n - e : e));
            }
#pragma unroll
            for (int h = 0; h < (1 << (Q - q - 1)); ++h) {
                const int a = (h << (q + 1)) | jq;
                const int b = a | (1 << q);
                const Fr y = trivial ? v[b] : v[b] * tw;
                const Fr x = v[a];
                if (pp.inverse && !trivial) { // omega^-e = -T[n/2 - e]
                    v[a] = x - y;
                    v[b] = x + y;
                } else {
                    v[a] = x + y;
                    v[b] = x - y;
                }
            }
        }
    }
}

template<class Fr, int Q>
HD void step_phase(uint32_t tid, uint32_t nthreads, const PassParams &pp, int r, uint64_t index0, const Fr *twiddles,
                   const Tile<Fr> &tile)
{
    const uint32_t C = 1u << pp.lgc, R = 1u << pp.deg;
    const uint32_t ngroups = (R >> Q) << pp.lgc;
    const uint64_t pmask = (1ull << pp.lgp) - 1;
    for (uint32_t gam = tid; gam < ngroups; gam += nthreads) {
        const uint32_t col = gam & (C - 1), gi = gam >> pp.lgc;
        const uint32_t lo = gi & ((1u << r) - 1), hi = gi >> r;
        const uint32_t base = (hi << (r + Q)) | lo;
        Fr v[1 << Q];
#pragma unroll
        for (int a = 0; a < (1 << Q); ++a) {
            v[a] = tile.load(base + ((uint32_t)a << r), col);
        }
        radix_group<Fr, Q>(v, pp, r, lo, (index0 + col) & pmask, twiddles);
#pragma unroll
        for (int a = 0; a < (1 << Q); ++a) {
            tile.store(base + ((uint32_t)a << r), col, v[a]);
        }
    }
}

// ---- phase C: scatter ---------------------------------------------------------------------------
template<class Fr>
HD void store_phase(uint32_t tid, uint32_t nthreads, const PassParams &pp, uint64_t index0, Fr *out,
                    const Fr *post_table, const Fr *post_const, const Tile<Fr> &tile)
{
    const uint32_t C = 1u << pp.lgc, R = 1u << pp.deg;
    const uint64_t p = 1ull << pp.lgp;
    for (uint32_t idx = tid; idx < R * C; idx += nthreads) {
        uint32_t col, j;
        uint64_t dst;
        if (pp.lgp >= pp.lgc) {
            col = idx & (C - 1);
            j = idx >> pp.lgc;
            const uint64_t k = (index0 + col) & (p - 1);
            dst = ((index0 + col - k) << pp.deg) + k + ((uint64_t)j << pp.lgp);
        } else { // p < C: runs of R * p contiguous outputs per block of p columns
            const uint32_t kk = idx & (uint32_t)(p - 1);
            j = (idx >> pp.lgp) & (R - 1);
            const uint32_t chi = idx >> (pp.lgp + pp.deg);
            col = (chi << pp.lgp) + kk;
            dst = ((index0 + ((uint64_t)chi << pp.lgp)) << pp.deg) + ((uint64_t)j << pp.lgp) + kk;
        }
        Fr x = tile.load(j, col);
        if (post_table != nullptr) {
            x = x * gload(post_table + dst);
        } else if (post_const != nullptr) {
            x = x * gload(post_const);
        }
        gstore(out + dst, x);
    }
}

// out[e] = scale * prod_{b : bit b of e} pows[b]   (pows[b] = base^(2^b)); used for the twiddle
// table (base = omega_n) and the coset tables (base = g or g^-1, scale = 1 or m^-1).
template<class Fr> HD void pow_table_body(uint64_t e, const Fr *pows, int nbits, const Fr *scale, Fr *out)
{
    Fr acc = scale ? gload(scale) : Fr::one();
    for (int b = 0; b < nbits; ++b) {
        if ((e >> b) & 1ull) {
            acc = acc * gload(pows + b);
        }
    }
    gstore(out + e, acc);
}

// Domain constants, computed by one thread: consts layout (Fr each):
//   [0] omega_n  [1] m^-1  [2] g  [3] g^-1  [4] (g^m - 1)^-1  [5] one
//   pows_w[b] = omega^(2^b), pows_g[b] = g^(2^b), pows_gi[b] = g^-(2^b),  b < lgn
template<class Fr> HD void domain_consts_body(int lgn, const Fr *g_in, Fr *consts, Fr *pows_w, Fr *pows_g, Fr *pows_gi)
{
    Fr w;
#pragma unroll
    for (int i = 0; i < Fr::N; ++i) {
        w.v[i] = Fr::params::root_of_unity(i);
    }
    for (int i = Fr::params::S; i > lgn; --i) { // get_root_of_unity, field_utils.tcc:190-206
        w = w.sqr();
    }
    Fr m = Fr::zero();
    m.v[lgn >> 5] = 1u << (lgn & 31);
    const Fr minv = m.to_mont().inverse();
    const Fr g = gload(g_in);
    const Fr gi = g.is_zero() ? g : g.inverse();
    Fr pw = w, pg = g, pgi = gi;
    for (int b = 0; b < lgn; ++b) {
        gstore(pows_w + b, pw);
        gstore(pows_g + b, pg);
        gstore(pows_gi + b, pgi);
        pw = pw.sqr();
        pg = pg.sqr();
        pgi = pgi.sqr();
    }
    // after lgn squarings pg = g^m: Z(g x) = g^m - 1 on the coset (basic_radix2_domain.tcc:117-126)
    const Fr z = pg - Fr::one();
    gstore(consts + 0, w);
    gstore(consts + 1, minv);
    gstore(consts + 2, g);
    gstore(consts + 3, gi);
    gstore(consts + 4, z.is_zero() ? z : z.inverse());
    gstore(consts + 5, Fr::one());
}

// ---- sparse R1CS evaluation: the row body and the host-side plan for long rows (kernels further down) -------------------
constexpr uint32_t R1CS_LONG_ROW = 64;
constexpr int R1CS_LONG_THREADS = 256;
constexpr uint32_t R1CS_LONG_SEG = 2048; // terms per CTA of the long-row kernel
// Rows with more than R1CS_LONG_ROW terms (a sum over all variables is a common last constraint) are cut into segments
// of R1CS_LONG_SEG terms: one CTA per segment writes a partial; a second launch adds the partials of each row.
struct LongSeg {
    uint32_t row, e0, e1, slot; // slot: index of the row in the long-row list
};
// rows = 3 nc (the a-, b-, c-combinations); seg_ptr[k] .. seg_ptr[k+1]: the segments of long row k
inline void r1cs_long_plan(const uint32_t *row_ptr, size_t rows, std::vector<uint32_t> &long_rows, std::vector<uint32_t> &seg_ptr,
                           std::vector<LongSeg> &segs)
{
    for (size_t i = 0; i < rows; ++i) {
        if (row_ptr[i + 1] - row_ptr[i] > R1CS_LONG_ROW) {
            seg_ptr.push_back((uint32_t)segs.size());
            for (uint32_t e = row_ptr[i]; e < row_ptr[i + 1]; e += R1CS_LONG_SEG) {
                const uint32_t e1 = e + R1CS_LONG_SEG < row_ptr[i + 1] ? e + R1CS_LONG_SEG : row_ptr[i + 1];
                segs.push_back(LongSeg{(uint32_t)i, e, e1, (uint32_t)long_rows.size()});
            }
            long_rows.push_back((uint32_t)i);
        }
    }
    seg_ptr.push_back((uint32_t)segs.size());
}
// Element t of the 3 m outputs (r1cs_to_qap.tcc:236-248, :293-296; linear_combination::evaluate, variable.tcc:288-299):
// rows [0, nc) of A, [nc, 2nc) of B, [2nc, 3nc) of C in one CSR; column 0 is the constant 1, column k the k-th variable,
// so w = (1 | primary | auxiliary) is indexed directly. Also the input-consistency rows aA[nc + i] = w[i], i <= num_inputs
// (:240-243) and the zero padding up to m. A long row is left to the segment kernels.
template<class Fr>
HD void r1cs_eval_body(uint64_t t, uint64_t nc, uint64_t ni, uint64_t m, const uint32_t *row_ptr, const uint32_t *col, const Fr *val,
                       const Fr *w, Fr *aA, Fr *aB, Fr *aC)
{
    const uint64_t mat = t / m, i = t - mat * m;
    Fr *dst = mat == 0 ? aA : (mat == 1 ? aB : aC);
    Fr acc = Fr::zero();
    if (i < nc) {
        const uint64_t row = mat * nc + i;
        const uint32_t e0 = row_ptr[row], e1 = row_ptr[row + 1];
        if (e1 - e0 > R1CS_LONG_ROW) {
            return; // the segment kernels own this row
        }
        for (uint32_t e = e0; e < e1; ++e) {
            acc = acc + gload(val + e) * gload(w + col[e]);
        }
    } else if (mat == 0 && i - nc <= ni) {
        acc = gload(w + (i - nc));
    }
    gstore(dst + i, acc);
}

// ---- step_radix2_domain: element-wise steps and host-side constants (kernels further down) ------------------------------
// m = big + small, big = 2^k > small = 2^r; T2[e] = omega^e, e < big, is the twiddle table of the radix-2 domain of size
// 2 big (omega = its root of unity). Line numbers: step_radix2_domain.tcc.
// FFT, :73-88: c[i] = a[i] + a[i+big] (i < small) else a[i]; d[i] = omega^i * (a[i] - a[i+big] or a[i])
template<class Fr> HD void step_fft_pre_body(uint64_t i, uint64_t big, uint64_t small, const Fr *T2, Fr *a, Fr *d)
{
    const Fr x = gload(a + i);
    Fr c = x, dd = x;
    if (i < small) {
        const Fr y = gload(a + i + big);
        c = x + y;
        dd = x - y;
    }
    gstore(a + i, c);
    gstore(d + i, dd * gload(T2 + i));
}
// one halving step of e[i] = sum_j d[i + j small] (:90-98): v[t] += v[t + half]
template<class Fr> HD void fold_halve_body(uint64_t t, uint64_t half, Fr *v) { gstore(v + t, gload(v + t) + gload(v + t + half)); }
// iFFT, :135-150: tmp[i] = U0[i] omega^i for i >= small, 0 below (the j = 0 term is not subtracted)
template<class Fr> HD void step_ifft_tmp_body(uint64_t i, uint64_t small, const Fr *T2, const Fr *u0, Fr *tmp)
{
    gstore(tmp + i, i < small ? Fr::zero() : gload(u0 + i) * gload(T2 + i));
}
// iFFT, :152-173: U1[i] = (U1[i] - s[i]) omega^-i; a[i] = (U0[i] + U1[i]) / 2; a[big+i] = (U0[i] - U1[i]) / 2
template<class Fr> HD void step_ifft_post_body(uint64_t i, uint64_t big, const Fr *T2, const Fr *s, const Fr *over_two, Fr *a)
{
    const Fr u0 = gload(a + i);
    Fr u1 = gload(a + big + i) - gload(s + i);
    if (i != 0) {
        u1 = u1 * gload(T2 + (big - i)).neg(); // omega^big = -1
    }
    const Fr h = gload(over_two);
    gstore(a + i, (u0 + u1) * h);
    gstore(a + big + i, (u0 - u1) * h);
}
// a[i] *= table[i]
template<class Fr> HD void mul_table_body(uint64_t i, Fr *a, const Fr *table) { gstore(a + i, gload(a + i) * gload(table + i)); }
// divide_by_Z_on_coset, :252-276: zinv[k] = (A w2s^k - B)^-1, k < compr (the factor has period compr in i)
template<class Fr> HD void step_zinv_body(uint64_t k, const Fr *abw /* A, B, w2s */, Fr *zinv)
{
    Fr elt = Fr::one(), base = gload(abw + 2);
    for (uint64_t e = k; e; e >>= 1) {
        if (e & 1) {
            elt = elt * base;
        }
        base = base.sqr();
    }
    gstore(zinv + k, (gload(abw) * elt - gload(abw + 1)).inverse());
}
template<class Fr> HD void step_divz_body(uint64_t i, uint64_t big, uint64_t compr, const Fr *zinv, const Fr *z1inv, Fr *p)
{
    const Fr f = i < big ? gload(zinv + (i & (compr - 1))) : gload(z1inv);
    gstore(p + i, gload(p + i) * f);
}

// ---- extended_radix2_domain: element-wise steps and host-side constants (kernels further down) --------------------------
// Powers b^i of a fixed element from a two-level table: b^i = lo[i mod 2^lg_lo] * hi[i >> lg_lo].
template<class Fr> struct PowTable {
    const Fr *lo, *hi;
    int lg_lo;
    HD Fr at(uint64_t i) const { return gload(lo + (i & ((1ull << lg_lo) - 1))) * gload(hi + (i >> lg_lo)); }
};
// FFT, extended_radix2_domain.tcc:64-91: a0[i] = a[i] + a[i+small], a1[i] = shift^i (a[i] + shift^small a[i+small])
template<class Fr> HD void ext_fft_pre_body(uint64_t i, uint64_t small, Fr *a, const PowTable<Fr> &spow, const Fr *ssm)
{
    const Fr x = gload(a + i), y = gload(a + i + small);
    gstore(a + i, x + y);
    gstore(a + i + small, spow.at(i) * (x + gload(ssm) * y));
}
// iFFT, :93-120, on the unnormalised inverse transforms a0 | a1: a[i] = sconst (shift^-i a1[i] - shift^small a0[i]),
// a[i+small] = sconst (a0[i] - shift^-i a1[i]); cst = { shift^small, (small (1 - shift^small))^-1 }
template<class Fr> HD void ext_ifft_post_body(uint64_t i, uint64_t small, Fr *a, const PowTable<Fr> &sinv, const Fr *cst)
{
    const Fr a0 = gload(a + i), t = sinv.at(i) * gload(a + i + small);
    const Fr sc = gload(cst + 1);
    gstore(a + i, sc * (t - gload(cst) * a0));
    gstore(a + i + small, sc * (a0 - t));
}
// _multiply_by_coset (basic_radix2_domain_aux.tcc:174-183): a[i] *= g^i
template<class Fr> HD void mul_pow_body(uint64_t i, Fr *a, const PowTable<Fr> &gpow) { gstore(a + i, gload(a + i) * gpow.at(i)); }
// divide_by_Z_on_coset, :216-236: one factor per half
template<class Fr> HD void ext_divz_body(uint64_t i, uint64_t small, Fr *p, const Fr *zinv) { gstore(p + i, gload(p + i) * gload(zinv + (i >= small))); }

template<class Fr> inline Fr host_pow(Fr base, uint64_t e)
{
    Fr acc = Fr::one();
    for (; e; e >>= 1) {
        if (e & 1) {
            acc = acc * base;
        }
        base = base.sqr();
    }
    return acc;
}
// step domain constants (step_radix2_domain.tcc:252-276): [0] 1/2  [1] A = g^small Z0  [2] B = omega^small Z0
// [3] omega^(2 small)  [4] Z1^-1  [5] one;  Z0 = g^big - 1, omega = the root of unity of order 2 big, g = multiplicative_generator
template<class Fr> inline void step_domain_consts_host(int lg_big, uint64_t small, Fr hc[6])
{
    const uint64_t big = 1ull << lg_big;
    Fr w, g;
    for (int i = 0; i < Fr::N; ++i) {
        w.v[i] = Fr::params::root_of_unity(i);
        g.v[i] = Fr::params::generator(i);
    }
    for (int i = Fr::params::S; i > lg_big + 1; --i) {
        w = w.sqr(); // omega = get_root_of_unity(2 big)
    }
    const Fr one = Fr::one();
    const Fr z0 = host_pow(g, big) - one;
    const Fr w_small = host_pow(w, small);
    const Fr gw = g * w;
    const Fr z1 = (host_pow(gw, big) - one) * (host_pow(gw, small) - w_small);
    hc[0] = (one + one).inverse();
    hc[1] = host_pow(g, small) * z0;
    hc[2] = w_small * z0;
    hc[3] = host_pow(w, 2 * small);
    hc[4] = z1.inverse();
    hc[5] = one;
}
// [0] shift^small  [1] (small (1 - shift^small))^-1 (:107)  [2] Z0^-1  [3] Z1^-1 (:224-229), small = 2^lg_small,
// shift = libff::coset_shift() = multiplicative_generator^2 (field_utils.tcc:152-155); *shift_out = shift
template<class Fr> inline void ext_domain_consts_host(int lg_small, Fr hc[4], Fr *shift_out)
{
    Fr g;
    for (int i = 0; i < Fr::N; ++i) {
        g.v[i] = Fr::params::generator(i); // FieldT::multiplicative_generator
    }
    const uint64_t small = 1ull << lg_small;
    const Fr one = Fr::one();
    const Fr shift = g.sqr();
    const Fr ssm = host_pow(shift, small), gsm = host_pow(g, small);
    Fr small_f = Fr::zero();
    small_f.v[lg_small >> 5] = 1u << (lg_small & 31);
    hc[0] = ssm;
    hc[1] = (small_f.to_mont() * (one - ssm)).inverse();
    hc[2] = ((gsm - one) * (gsm - ssm)).inverse();
    hc[3] = ((gsm * ssm - one) * (gsm * ssm - ssm)).inverse();
    *shift_out = shift;
}
// lo[j] = b^j (j < 2^lg_lo), hi[k] = b^(k 2^lg_lo) (k < count >> lg_lo), for b (t = 0) and b^-1 (t = 1), laid out
// lo0 | hi0 | lo1 | hi1 in h (2 * (2^lg_lo + (count >> lg_lo)) elements)
template<class Fr> inline void ext_pow_tables_host(const Fr &b, size_t count, int lg_lo, Fr *h)
{
    const size_t nlo = (size_t)1 << lg_lo, nhi = count >> lg_lo;
    const Fr bases[2] = {b, b.inverse()};
    for (int t = 0; t < 2; ++t) {
        Fr *lo = h + t * (nlo + nhi), *hi = lo + nlo;
        lo[0] = Fr::one();
        for (size_t j = 1; j < nlo; ++j) {
            lo[j] = lo[j - 1] * bases[t];
        }
        const Fr step = lo[nlo - 1] * bases[t]; // b^(2^lg_lo)
        hi[0] = Fr::one();
        for (size_t k = 1; k < nhi; ++k) {
            hi[k] = hi[k - 1] * step;
        }
    }
}
constexpr int EXT_LG_LO = 15;

#if defined(__CUDACC__)

template<class Fr>
__global__ void __launch_bounds__(512) k_ntt_pass(PassParams pp, const Fr *in, Fr *out, const Fr *twiddles,
                                                  const Fr *pre_table, const Fr *post_table, const Fr *post_const)
{
    extern __shared__ uint4 ntt_smem[];
    Tile<Fr> tile;
    tile.base = reinterpret_cast<Vec4 *>(ntt_smem);
    tile.row_stride = (1u << pp.lgc) + 1;
    tile.plane = (1u << pp.deg) * tile.row_stride;
    const uint32_t tid = threadIdx.x, nth = blockDim.x;
    const uint64_t index0 = (uint64_t)blockIdx.x << pp.lgc;
    load_phase<Fr>(tid, nth, pp, index0, in, pre_table, tile);
    __syncthreads();
    int r = 0;
    for (int s = 0; s < pp.nsteps; ++s) {
        const int q = pp.step_q[s];
        if (q == 3) {
            step_phase<Fr, 3>(tid, nth, pp, r, index0, twiddles, tile);
        } else if (q == 2) {
            step_phase<Fr, 2>(tid, nth, pp, r, index0, twiddles, tile);
        } else {
            step_phase<Fr, 1>(tid, nth, pp, r, index0, twiddles, tile);
        }
        r += q;
        __syncthreads();
    }
    store_phase<Fr>(tid, nth, pp, index0, out, post_table, post_const, tile);
}

template<class Fr>
__global__ void k_pow_table(uint64_t count, const Fr *pows, int nbits, const Fr *scale, Fr *out)
{
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < count) {
        pow_table_body<Fr>(e, pows, nbits, scale, out);
    }
}

template<class Fr> __global__ void k_domain_consts(int lgn, const Fr *g_in, Fr *consts, Fr *pows_w, Fr *pows_g, Fr *pows_gi)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        domain_consts_body<Fr>(lgn, g_in, consts, pows_w, pows_g, pows_gi);
    }
}

// h[i] = (a[i] * b[i] - c[i]) * zinv      (r1cs_to_qap.tcc:286-288, :310-312, :315)
template<class Fr> __global__ void k_pointwise_h(uint64_t m, const Fr *a, const Fr *b, const Fr *c, const Fr *zinv, Fr *h)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) {
        const Fr z = gload(zinv);
        const Fr v = (gload(a + i) * gload(b + i) - gload(c + i)) * z;
        gstore(h + i, v);
    }
}

// a[i] *= c   (divide_by_Z_on_coset, basic_radix2_domain.tcc:117-126)
template<class Fr> __global__ void k_scale(uint64_t m, Fr *a, const Fr *c)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) {
        gstore(a + i, gload(a + i) * gload(c));
    }
}

// ---- step_radix2_domain (libfqfft/evaluation_domain/domains/step_radix2_domain.tcc) ---------------
// Domain of size m = big + small, big = 2^k > small = 2^r: the factory's choice for every size that
// is not a power of two (get_evaluation_domain.tcc:41-58). A transform is two radix-2 transforms
// (sizes big and small) plus the element-wise steps below; T2[e] = omega^e, e < big, is the
// twiddle table of the radix-2 domain of size 2 big (omega = its root of unity).

// (the element-wise steps themselves: step_*_body above)
template<class Fr> __global__ void k_step_fft_pre(uint64_t big, uint64_t small, const Fr *T2, Fr *a, Fr *d)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < big) {
        step_fft_pre_body<Fr>(i, big, small, T2, a, d);
    }
}
template<class Fr> __global__ void k_fold_halve(uint64_t half, Fr *v)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < half) {
        fold_halve_body<Fr>(t, half, v);
    }
}
template<class Fr> __global__ void k_step_ifft_tmp(uint64_t big, uint64_t small, const Fr *T2, const Fr *u0, Fr *tmp)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < big) {
        step_ifft_tmp_body<Fr>(i, small, T2, u0, tmp);
    }
}
template<class Fr>
__global__ void k_step_ifft_post(uint64_t big, uint64_t small, const Fr *T2, const Fr *s, const Fr *over_two, Fr *a)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < small) {
        step_ifft_post_body<Fr>(i, big, T2, s, over_two, a);
    }
}
template<class Fr> __global__ void k_mul_table(uint64_t n, Fr *a, const Fr *table)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        mul_table_body<Fr>(i, a, table);
    }
}
template<class Fr> __global__ void k_step_zinv(uint64_t compr, const Fr *abw /* A, B, w2s */, Fr *zinv)
{
    const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < compr) {
        step_zinv_body<Fr>(k, abw, zinv);
    }
}
template<class Fr>
__global__ void k_step_divz(uint64_t big, uint64_t m, uint64_t compr, const Fr *zinv, const Fr *z1inv, Fr *p)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) {
        step_divz_body<Fr>(i, big, compr, zinv, z1inv, p);
    }
}
// h[i] = a[i] b[i] - c[i]
template<class Fr> __global__ void k_pointwise_abc(uint64_t m, const Fr *a, const Fr *b, const Fr *c, Fr *h)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) {
        gstore(h + i, gload(a + i) * gload(b + i) - gload(c + i));
    }
}

// ---- extended_radix2_domain (libfqfft/evaluation_domain/domains/extended_radix2_domain.tcc) --------
// Domain of size m = 2 small, small = 2^s the largest radix-2 subgroup of the field: the subgroup H of order
// small and its coset shift * H, shift = coset_shift() = multiplicative_generator^2. A transform is two radix-2
// transforms of size small plus the element-wise steps above (ext_*_body). The two-level power tables are L2-resident:
// two multiplications per element instead of a table as long as the vector, which at 2^29 elements would be 16 GiB.
template<class Fr> __global__ void k_ext_fft_pre(uint64_t small, Fr *a, PowTable<Fr> spow, const Fr *ssm)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < small) {
        ext_fft_pre_body<Fr>(i, small, a, spow, ssm);
    }
}
template<class Fr> __global__ void k_ext_ifft_post(uint64_t small, Fr *a, PowTable<Fr> sinv, const Fr *cst)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < small) {
        ext_ifft_post_body<Fr>(i, small, a, sinv, cst);
    }
}
template<class Fr> __global__ void k_mul_pow(uint64_t n, Fr *a, PowTable<Fr> gpow)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        mul_pow_body<Fr>(i, a, gpow);
    }
}
template<class Fr> __global__ void k_ext_divz(uint64_t small, Fr *p, const Fr *zinv /* Z0^-1, Z1^-1 */)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 2 * small) {
        ext_divz_body<Fr>(i, small, p, zinv);
    }
}

// ---- the ZK patch of r1cs_to_qap_witness_map (r1cs_to_qap.tcc:262-271, :327-333) -----------------------------------
// t[i] = ca a[i] + cb b[i]: the evaluations of d2 A + d1 B, whose inverse transform gives d2 coef(A) + d1 coef(B)
template<class Fr> __global__ void k_lincomb2(uint64_t m, const Fr *a, const Fr *b, Fr ca, Fr cb, Fr *t)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) {
        gstore(t + i, ca * gload(a + i) + cb * gload(b + i));
    }
}
// h[i] += t[i]
template<class Fr> __global__ void k_add_vec(uint64_t m, Fr *h, const Fr *t)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) {
        gstore(h + i, gload(h + i) + gload(t + i));
    }
}
// the handful of single coefficients -d3 and add_poly_Z(d1 d2) touch: h[idx[k]] += val[k], in order (indices may repeat)
template<class Fr> struct CoeffFixups {
    uint64_t idx[6];
    Fr val[6];
    int n;
};
template<class Fr> __global__ void k_coeff_fixups(CoeffFixups<Fr> f, Fr *h)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (int k = 0; k < f.n; ++k) {
            gstore(h + f.idx[k], gload(h + f.idx[k]) + f.val[k]);
        }
    }
}

// Sparse R1CS evaluation: one thread per output element (r1cs_eval_body above); rows of R1CS matrices hold a handful of terms.
template<class Fr>
__global__ void k_r1cs_eval(uint64_t nc, uint64_t ni, uint64_t m, const uint32_t *row_ptr, const uint32_t *col,
                            const Fr *val, const Fr *w, Fr *aA, Fr *aB, Fr *aC)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < 3 * m) {
        r1cs_eval_body<Fr>(t, nc, ni, m, row_ptr, col, val, w, aA, aB, aC);
    }
}

// Long rows: one CTA per segment (strided partial sums, shared-memory tree), then one CTA per row over its partials.
template<class Fr> __device__ Fr cta_tree_sum(Fr acc, Fr *part)
{
    part[threadIdx.x] = acc;
    __syncthreads();
    for (int half = R1CS_LONG_THREADS / 2; half > 0; half >>= 1) {
        if ((int)threadIdx.x < half) {
            part[threadIdx.x] = part[threadIdx.x] + part[threadIdx.x + half];
        }
        __syncthreads();
    }
    return part[0];
}
template<class Fr>
__global__ void __launch_bounds__(R1CS_LONG_THREADS)
    k_r1cs_eval_long(const LongSeg *segs, const uint32_t *col, const Fr *val, const Fr *w, Fr *partials)
{
    __shared__ Fr part[R1CS_LONG_THREADS];
    const LongSeg sg = segs[blockIdx.x];
    Fr acc = Fr::zero();
    for (uint32_t e = sg.e0 + threadIdx.x; e < sg.e1; e += R1CS_LONG_THREADS) {
        acc = acc + gload(val + e) * gload(w + col[e]);
    }
    const Fr total = cta_tree_sum(acc, part);
    if (threadIdx.x == 0) {
        gstore(partials + blockIdx.x, total);
    }
}
// seg_ptr[k] .. seg_ptr[k+1]: the segments of long row k (consecutive in `partials`)
template<class Fr>
__global__ void __launch_bounds__(R1CS_LONG_THREADS)
    k_r1cs_long_finish(uint64_t nc, const uint32_t *long_rows, const uint32_t *seg_ptr, const Fr *partials, Fr *aA, Fr *aB,
                       Fr *aC)
{
    __shared__ Fr part[R1CS_LONG_THREADS];
    const uint32_t s0 = seg_ptr[blockIdx.x], s1 = seg_ptr[blockIdx.x + 1];
    Fr acc = Fr::zero();
    for (uint32_t k = s0 + threadIdx.x; k < s1; k += R1CS_LONG_THREADS) {
        acc = acc + gload(partials + k);
    }
    const Fr total = cta_tree_sum(acc, part);
    if (threadIdx.x == 0) {
        const uint64_t row = long_rows[blockIdx.x];
        const uint64_t mat = row / nc, i = row - mat * nc;
        Fr *dst = mat == 0 ? aA : (mat == 1 ? aB : aC);
        gstore(dst + i, total);
    }
}

// dst[i] = src[idx[i]]
template<class Fr> __global__ void k_gather(uint64_t n, const Fr *src, const uint32_t *idx, Fr *dst)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        gstore(dst + i, gload(src + idx[i]));
    }
}

#endif // __CUDACC__

} // namespace ntt
} // namespace b200
