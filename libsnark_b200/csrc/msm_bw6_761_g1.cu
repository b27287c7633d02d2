// msm_bw6_761_g1.cu — instantiates the MSM pipeline (msm.cuh / msm_host.cuh) for bw6_761 G1.
#include "msm_host.cuh"

namespace b200
{
B200_DEFINE_GROUP_ENTRY(bw6_761_g1, CURVE_BW6_761, GROUP_G1)
} // namespace b200
