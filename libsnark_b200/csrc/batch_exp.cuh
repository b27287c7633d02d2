// batch_exp.cuh — fixed-base batch exponentiation: out[i] = (coeff * v[i]) * g for one base g and many scalars.
//
// Replaces, for key generation, libff::batch_exp / batch_exp_with_coeff
// (multiexp.tcc:879-946; windowed_exp :846-876 over get_window_table :808-844) followed by batch_to_special
// (:949-974, batched inversion field_utils.tcc:330-350), and through them libsnark::kc_batch_exp
// (kc_multiexp.tcc:147-239). Same idea as the reference — a table of the multiples d * 2^(c w) * g, one table
// addition per window — with a fixed 8-bit window (the reference tunes the window to the vector length because
// its table is built on one core; here the table is 32 x 255 points and costs milliseconds), affine table
// entries so that every addition is a mixed addition, and the normalisation of all results by Montgomery's
// simultaneous inversion in groups of K per thread. The result of s * g is a group element and its special
// form is unique, so outputs are bit-identical to the reference's after batch_to_special.
//
// Kernel bodies are HD functions of a flat index so that tests/host_emu runs the same logic on the CPU.
#pragma once
#include <vector>

#include "curves.cuh"
#include "msm.cuh"

namespace b200
{
namespace bexp
{

constexpr int WBITS = 8;
constexpr int PER_WINDOW = (1 << WBITS) - 1; // table entries per window: d = 1 .. 255
template<class Fr> constexpr int num_windows() { return (Fr::params::BITS + WBITS - 1) / WBITS; }
// results normalised per thread with one shared inversion
template<class F> constexpr int norm_group() { return sizeof(F) <= 64 ? 8 : 4; }

// Host: table[w * 255 + d - 1] = d * 2^(8 w) * g (affine). g in special form, not infinity.
// Returns false when some multiple is the point at infinity (g of small order): no table.
template<class Fr, class F> bool build_table(const msm::RefPoint<F> &g, std::vector<Affine<F>> &table)
{
    const int W = num_windows<Fr>();
    std::vector<XYZZ<F>> t((size_t)W * PER_WINDOW);
    Affine<F> outer_a;
    outer_a.x = g.x;
    outer_a.y = g.y;
    for (int w = 0; w < W; ++w) {
        XYZZ<F> acc = XYZZ<F>::from_affine(outer_a);
        t[(size_t)w * PER_WINDOW] = acc;
        for (int d = 2; d <= PER_WINDOW; ++d) {
            madd(acc, outer_a);
            t[(size_t)w * PER_WINDOW + d - 1] = acc;
        }
        madd(acc, outer_a); // 2^8 * outer
        if (!to_affine(acc, outer_a) && w + 1 < W) {
            return false;
        }
    }
    // simultaneous inversion of all ZZ * ZZZ
    const size_t m = t.size();
    std::vector<F> pre(m);
    F run = F::one();
    for (size_t k = 0; k < m; ++k) {
        if (t[k].is_inf()) {
            return false;
        }
        pre[k] = run;
        run = run * (t[k].zz * t[k].zzz);
    }
    F inv = run.inverse();
    table.resize(m);
    for (size_t k = m; k-- > 0;) {
        const F zi = inv * pre[k];
        inv = inv * (t[k].zz * t[k].zzz);
        table[k].x = t[k].x * (zi * t[k].zzz);
        table[k].y = t[k].y * (zi * t[k].zz);
    }
    return true;
}

// out[i] = (coeff * v[i]) * g as an XYZZ point: one mixed addition per non-zero 8-bit digit.
template<class Fr, class F>
HD void exp_body(size_t i, const Fr *scalars, const Fr *coeff, const Affine<F> *table, XYZZ<F> *out)
{
    Fr s = msm::load_vec(scalars + i);
    if (coeff != nullptr) {
        s = msm::load_vec(coeff) * s; // coeff * v[i], batch_exp_with_coeff (multiexp.tcc:935)
    }
    s = s.from_mont(); // as_bigint (windowed_exp, multiexp.tcc:854)
    XYZZ<F> acc = XYZZ<F>::inf();
    constexpr int W = num_windows<Fr>();
    for (int w = 0; w < W; ++w) {
        const int bit = w * WBITS;
        const uint32_t d = (s.v[bit >> 5] >> (bit & 31)) & (uint32_t)PER_WINDOW;
        if (d != 0) {
            const Affine<F> p = msm::load_vec_ro(table + (size_t)w * PER_WINDOW + (d - 1));
            madd(acc, p);
        }
    }
    msm::store_vec(out + i, acc);
}

// batch_to_special: points [t K, (t + 1) K) of `in` normalised with one inversion (Montgomery's trick);
// the point at infinity becomes the reference's zero().to_special() = (0, one, 0).
template<class F> HD void normalise_body(size_t t, size_t n, const XYZZ<F> *in, msm::RefPoint<F> *out)
{
    constexpr int K = norm_group<F>();
    F pre[K];
    F run = F::one();
    const size_t base = t * K;
    for (int k = 0; k < K; ++k) {
        pre[k] = run;
        if (base + k < n) {
            const XYZZ<F> p = msm::load_vec(in + base + k);
            if (!p.is_inf()) {
                run = run * (p.zz * p.zzz);
            }
        }
    }
    F inv = run.inverse();
    for (int k = K - 1; k >= 0; --k) {
        if (base + k >= n) {
            continue;
        }
        const XYZZ<F> p = msm::load_vec(in + base + k);
        msm::RefPoint<F> o;
        if (p.is_inf()) {
            o.x = F::zero();
            o.y = F::one();
            o.z = F::zero();
        } else {
            const F zi = inv * pre[k];
            inv = inv * (p.zz * p.zzz);
            o.x = p.x * (zi * p.zzz);
            o.y = p.y * (zi * p.zz);
            o.z = F::one();
        }
        msm::store_vec(out + base + k, o);
    }
}

} // namespace bexp
} // namespace b200
