// G1 bucket accumulation (see msm_acc.cuh)
#include "msm_acc.cuh"
namespace ctg {
template int accumulate_max_blocks_per_sm<Fq>();
template void accumulate_launch<Fq>(int, cudaStream_t, const Affine<Fq>*, const uint32_t*, const uint32_t*, uint32_t, XYZZ<Fq>*, XYZZ<Fq>*, uint32_t*);
}  // namespace ctg
