// msm_cold_bls12_377_g1.cu — bucket reduction / window sums for bls12_377 G1 (compiled with
// -DB200_OUTLINE_MUL, see msm_cold.cuh).
#include "msm_cold.cuh"

namespace b200
{
namespace msm
{
typedef group_traits<CURVE_BLS12_377, GROUP_G1>::F F_;
template int reduce_launch<F_>(const Geometry &, const uint32_t *, const uint32_t *, const uint32_t *, XYZZ<F_> *,
                               XYZZ<F_> *, XYZZ<F_> *, cudaStream_t);
template int affine_tree_launch<F_>(const uint32_t *, uint32_t, uint32_t, F_ *, F_ *, F_ *, F_ *, F_ *, cudaStream_t);
template int table_launch<F_>(size_t, int, int, Affine<F_> *, const uint8_t *, uint32_t *);
} // namespace msm
} // namespace b200
