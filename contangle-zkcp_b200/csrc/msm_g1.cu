// G1 instantiation of the MSM pipeline (see msm.cuh) + plan selection.
#include <algorithm>

#include "msm.cuh"

namespace ctg {

// Window size: minimise  windows * n  (mixed additions)  +  3.5 * sets * 2^(c-1)  (bucket reduction,
// two full additions per bucket plus the chunk offsets).  Windows are signed, so c bits need
// 2^(c-1) buckets; c <= 16 keeps a bucket set at 2^15 entries.
MsmPlan msm_make_plan(uint64_t n, int copies_wanted, int c_override) {
  MsmPlan best{};
  double best_cost = -1;
  for (int c = 2; c <= 16; c++) {
    if (c_override > 0 && c != c_override) continue;
    MsmPlan p;
    p.c = c;
    p.nwin = msm_num_windows(c);
    p.copies = copies_wanted <= 0 ? 1 : std::min(copies_wanted, p.nwin);
    p.nsets = (p.nwin + p.copies - 1) / p.copies;
    p.copies = (p.nwin + p.nsets - 1) / p.nsets;  // drop copies no window would use
    p.nbuckets_per_set = 1u << (c - 1);
    double cost = (double)p.nwin * (double)(n ? n : 1) + 3.5 * p.nsets * (double)p.nbuckets_per_set + 40.0 * c * (p.nsets - 1);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = p; }
  }
  return best;
}

int32_t msm_g1_run(ctg_ctx* ctx, MsmWorkspace& ws, const Bases& bases, const uint32_t* d_scalars, uint64_t n, bool scalars_mont, G1XYZZ* d_out,
                   cudaStream_t st, float* stage_ms) {
  return MsmHost<Fq>::run(ctx, ws, bases, d_scalars, n, scalars_mont, d_out, st, stage_ms);
}
int32_t g1_precompute(ctg_ctx* ctx, Bases& b, cudaStream_t st) { return MsmHost<Fq>::precompute(ctx, b, st); }
int32_t g1_to_affine_dev(ctg_ctx* ctx, const G1XYZZ* d_in, G1Affine* d_out, int count, cudaStream_t st) {
  return MsmHost<Fq>::to_affine(ctx, d_in, d_out, count, st);
}
int32_t g1_progression_dev(ctg_ctx* ctx, const G1Affine& g, const Fr& a, const Fr& b, uint64_t n, G1Affine* d_out, cudaStream_t st) {
  return MsmHost<Fq>::progression(ctx, g, a, b, n, d_out, st);
}
int32_t g1_combine_dev(ctg_ctx* ctx, const G1XYZZ* d_in, uint32_t k, G1Affine* d_out_affine, G1XYZZ* d_out_xyzz, cudaStream_t st) {
  combine_kernel<Fq><<<1, 32, 0, st>>>(d_in, k, d_out_affine, d_out_xyzz);
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}

void msm_ws_free(MsmWorkspace& ws) {
  void* ptrs[] = {ws.counts, ws.cursor, ws.entries, ws.buckets, ws.part_pt, ws.part_gb, ws.chunk, ws.result};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  ws = MsmWorkspace();
}

}  // namespace ctg
