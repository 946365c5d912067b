// G2 bucket accumulation (see msm_acc.cuh)
#include "msm_acc.cuh"
namespace ctg {
template int accumulate_max_blocks_per_sm<Fq2>();
template void accumulate_launch<Fq2>(int, cudaStream_t, const Affine<Fq2>*, const uint32_t*, const uint32_t*, uint32_t, XYZZ<Fq2>*, XYZZ<Fq2>*, uint32_t*);
}  // namespace ctg
