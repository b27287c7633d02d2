// msm_bls12_377_g1.cu — instantiates the MSM pipeline (msm.cuh / msm_host.cuh) for bls12_377 G1.
#include "msm_host.cuh"

namespace b200
{
B200_DEFINE_GROUP_ENTRY(bls12_377_g1, CURVE_BLS12_377, GROUP_G1)
} // namespace b200
