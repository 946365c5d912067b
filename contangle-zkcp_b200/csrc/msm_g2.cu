// G2 (Fq2) instantiation of the MSM pipeline (see msm.cuh).
#include <algorithm>

#include "msm.cuh"

namespace ctg {

int32_t msm_g2_run(ctg_ctx* ctx, MsmWorkspace& ws, const Bases& bases, const uint32_t* d_scalars, uint64_t n, bool scalars_mont, G2XYZZ* d_out,
                   cudaStream_t st, float* stage_ms, uint32_t batch, uint64_t vec_stride, const MsmWorkspace* sorted) {
  return MsmHost<Fq2>::run(ctx, ws, bases, d_scalars, n, scalars_mont, d_out, st, stage_ms, batch, vec_stride, sorted);
}
int32_t g2_precompute(ctg_ctx* ctx, Bases& b, cudaStream_t st) { return MsmHost<Fq2>::precompute(ctx, b, st); }
int32_t g2_to_affine_dev(ctg_ctx* ctx, const G2XYZZ* d_in, G2Affine* d_out, int count, cudaStream_t st) {
  return MsmHost<Fq2>::to_affine(ctx, d_in, d_out, count, st);
}
int32_t g2_progression_dev(ctg_ctx* ctx, const G2Affine& g, const Fr& a, const Fr& b, uint64_t n, G2Affine* d_out, cudaStream_t st) {
  return MsmHost<Fq2>::progression(ctx, g, a, b, n, d_out, st);
}
int32_t g2_combine_dev(ctg_ctx* ctx, const G2XYZZ* d_in, uint32_t k, G2Affine* d_out_affine, G2XYZZ* d_out_xyzz, cudaStream_t st) {
  combine_kernel<Fq2><<<1, 32, 0, st>>>(d_in, k, d_out_affine, d_out_xyzz);
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}

}  // namespace ctg
