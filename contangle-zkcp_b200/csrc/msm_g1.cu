// G1 instantiation of the MSM pipeline (see msm.cuh) + plan selection.
#include <algorithm>

#include "msm.cuh"

namespace ctg {

// Plan selection.  Cost model in microseconds, constants measured on B200 (profiles/r01_*):
//   bucket accumulation   windows * n mixed additions at ~2.5 G/s,
//   bucket reduction      halving trees: ~(c + 1) dependent additions + (c - 2) doublings at 12-20 us each (latency-bound),
//                         plus a throughput term when a set has more than 2^16 buckets (chunked running sums first),
//   window combine        c * (sets - 1) sequential doublings at ~12 us each (one thread).
// Windows are signed, so c bits need 2^(c-1) buckets.  c <= 21 keeps the bucket-index scan within 1023 tiles.
// copies_wanted = 0: keep as many precomputed multiples 2^(c G j) P as fit a 12 GiB table.
MsmPlan msm_make_plan(uint64_t n, int copies_wanted, int c_override, size_t point_bytes, bool batch_throughput) {
  MsmPlan best{};
  double best_cost = -1;
  const double budget = 12.0 * 1024 * 1024 * 1024;
  const uint64_t nn = n ? n : 1;
  for (int c = 2; c <= 21; c++) {
    if (c_override > 0 && c != c_override) continue;
    MsmPlan p;
    p.c = c;
    p.nwin = msm_num_windows(c);
    int cap = (int)std::min<double>(p.nwin, std::max(1.0, budget / ((double)nn * (double)point_bytes)));
    if (cap > 32) cap = 32;  // precompute_kernel keeps at most 32 multiples per base
    while ((uint64_t)cap * nn >= (1ull << 31) && cap > 1) cap--;
    p.copies = copies_wanted <= 0 ? cap : std::min(copies_wanted, p.nwin);
    if (p.copies > 32) p.copies = 32;
    p.nsets = (p.nwin + p.copies - 1) / p.copies;
    p.copies = (p.nwin + p.nsets - 1) / p.nsets;  // drop copies no window would use
    p.nbuckets_per_set = 1u << (c - 1);
    const double nb = (double)p.nsets * p.nbuckets_per_set;
    if (c_override <= 0 && (nb > 2097152.0 || (double)nn * p.nwin >= 4294967296.0)) continue;
    double red = p.nbuckets_per_set > 65536u ? 0.0012 * nb : 0.0003 * nb;   // chunked running sums only above 2^16 buckets per set
    double lat = 32.0 * c + 12.0 * c * (p.nsets - 1);
    if (batch_throughput) {
      // The plan of the batch twin of a small key (prove.cu pk_upload): ctg_prove_batch reduces the buckets of a whole chunk of
      // proofs with one launch per level, so the serial depth is paid once per chunk while the per-bucket work (zero fill,
      // fix-up, two additions, scan) is paid once per proof.  Measured on B200 with 1020 PoE-shaped proofs per call
      // (tools/run_once.py batch, CBITS sweep): 0.003 us per bucket; c = 12 .. 13 gives 0.47 ms per proof against 0.61 ms with
      // the latency-optimal window -- which a lone proof keeps (the same sweep costs a lone 2^13 proof 5.8 -> 7.7 ms).
      red = 0.003 * nb;
      lat *= 0.1;
    }
    double cost = 0.0004 * p.nwin * (double)nn + red + lat;
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = p; }
  }
  return best;
}

int32_t msm_g1_run(ctg_ctx* ctx, MsmWorkspace& ws, const Bases& bases, const uint32_t* d_scalars, uint64_t n, bool scalars_mont, G1XYZZ* d_out,
                   cudaStream_t st, float* stage_ms, uint32_t batch, uint64_t vec_stride, const MsmWorkspace* sorted) {
  return MsmHost<Fq>::run(ctx, ws, bases, d_scalars, n, scalars_mont, d_out, st, stage_ms, batch, vec_stride, sorted);
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
  void* ptrs[] = {ws.counts, ws.cursor, ws.entries, ws.buckets, ws.part_pt, ws.part_gb, ws.chunk, ws.result, ws.tile_sums,
                  ws.ba_pts[0], ws.ba_pts[1], ws.ba_starts[0], ws.ba_starts[1]};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  if (ws.ev_sorted) cudaEventDestroy(ws.ev_sorted);
  ws = MsmWorkspace();
}

}  // namespace ctg
