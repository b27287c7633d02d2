// curves.cuh — concrete field / group types for the three curves in scope
// (alt_bn128 = "BN254", bls12_377, bw6_761; SURVEY.md §8 size table).
#pragma once
#include "curve.cuh"
#include "curve_params.cuh"

namespace b200
{

enum curve_id { CURVE_ALT_BN128 = 0, CURVE_BLS12_377 = 1, CURVE_BW6_761 = 2 };
enum group_id { GROUP_G1 = 1, GROUP_G2 = 2 };

typedef Fp<alt_bn128_fr_params> alt_bn128_fr;
typedef Fp<alt_bn128_fq_params> alt_bn128_fq;
typedef Fp<bls12_377_fr_params> bls12_377_fr;
typedef Fp<bls12_377_fq_params> bls12_377_fq;
typedef Fp<bw6_761_fr_params> bw6_761_fr;
typedef Fp<bw6_761_fq_params> bw6_761_fq;

// alt_bn128 Fq2 = Fq[u]/(u^2 + 1)  (non_residue = -1, alt_bn128_init.cpp:138-140)
struct alt_bn128_fq2_params {
    typedef alt_bn128_fq base;
    HD static base mul_by_nr(const base &x) { return x.neg(); }
};
typedef Fp2<alt_bn128_fq2_params> alt_bn128_fq2;

// bls12_377 Fq2 = Fq[u]/(u^2 + 5)  (non_residue = -5, bls12_377_init.cpp:174)
struct bls12_377_fq2_params {
    typedef bls12_377_fq base;
    HD static base mul_by_nr(const base &x)
    {
        const base x4 = x.dbl().dbl();
        return (x4 + x).neg();
    }
};
typedef Fp2<bls12_377_fq2_params> bls12_377_fq2;

// (curve, group) -> scalar field, coordinate field.
template<int CURVE, int GROUP> struct group_traits;
template<> struct group_traits<CURVE_ALT_BN128, GROUP_G1> {
    typedef alt_bn128_fr Fr;
    typedef alt_bn128_fq F;
};
template<> struct group_traits<CURVE_ALT_BN128, GROUP_G2> {
    typedef alt_bn128_fr Fr;
    typedef alt_bn128_fq2 F;
};
template<> struct group_traits<CURVE_BLS12_377, GROUP_G1> {
    typedef bls12_377_fr Fr;
    typedef bls12_377_fq F;
};
template<> struct group_traits<CURVE_BLS12_377, GROUP_G2> {
    typedef bls12_377_fr Fr;
    typedef bls12_377_fq2 F;
};
// bw6_761: G1 and G2 are both over Fq (bw6_761_g2.hpp: base field bw6_761_Fq).
template<> struct group_traits<CURVE_BW6_761, GROUP_G1> {
    typedef bw6_761_fr Fr;
    typedef bw6_761_fq F;
};
template<> struct group_traits<CURVE_BW6_761, GROUP_G2> {
    typedef bw6_761_fr Fr;
    typedef bw6_761_fq F;
};

} // namespace b200
