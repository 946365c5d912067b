// G1 bucket accumulation (see msm_acc.cuh)
#include "msm_acc.cuh"
namespace ctg {
template int accumulate_max_blocks_per_sm<Fq>();
template void accumulate_launch<Fq>(int, cudaStream_t, bool, const Affine<Fq>*, const uint32_t*, const uint32_t*, uint32_t, XYZZ<Fq>*, XYZZ<Fq>*, uint32_t*);
template int ba_max_blocks_per_sm<Fq>();
template void ba_scan_launch<Fq>(cudaStream_t, const uint32_t*, uint32_t*, uint32_t, uint32_t*);
template void ba_level_launch<Fq>(int, cudaStream_t, bool, const Affine<Fq>*, const uint32_t*, const Affine<Fq>*, const uint32_t*, const uint32_t*, uint32_t, Affine<Fq>*);
}  // namespace ctg
