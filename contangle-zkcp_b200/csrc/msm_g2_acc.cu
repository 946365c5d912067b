// G2 bucket accumulation (see msm_acc.cuh)
#include "msm_acc.cuh"
namespace ctg {
template int accumulate_max_blocks_per_sm<Fq2>();
template void accumulate_launch<Fq2>(int, cudaStream_t, bool, const Affine<Fq2>*, const uint32_t*, const uint32_t*, uint32_t, XYZZ<Fq2>*, XYZZ<Fq2>*, uint32_t*);
template int ba_max_blocks_per_sm<Fq2>();
template void ba_scan_launch<Fq2>(cudaStream_t, const uint32_t*, uint32_t*, uint32_t, uint32_t*);
template void ba_level_launch<Fq2>(int, cudaStream_t, bool, const Affine<Fq2>*, const uint32_t*, const Affine<Fq2>*, const uint32_t*, const uint32_t*, uint32_t, Affine<Fq2>*);
}  // namespace ctg
