// msm_alt_bn128_g1.cu — instantiates the MSM pipeline (msm.cuh / msm_host.cuh) for alt_bn128 G1.
#include "msm_host.cuh"

namespace b200
{
B200_DEFINE_GROUP_ENTRY(alt_bn128_g1, CURVE_ALT_BN128, GROUP_G1)
} // namespace b200
