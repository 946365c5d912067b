// G1 instantiation of the MSM pipeline (see msm.cuh) + plan selection.
#include <algorithm>
#include <cstdlib>
#include <cmath>

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
    // Buckets cut by slice boundaries (msm_fixup_*): the sorted entry list is cut into slices of 16 entries (at most 296 x 128 of
    // them), so a bucket holding q slices is stitched from q partial sums by ONE thread (q <= 8) or one warp (q / 32 serial additions
    // per lane + a 5-level tree), at the latency of a lone group addition (~30 us G1, ~165 us G2).  Narrow windows on small inputs make
    // q huge: MulChain 2^12 chose c = 4 (16 buckets of 16 K entries each) and spent 6 of its 8.2 ms there.  Measured on B200 with and
    // without this term (tools/round2/r2z2.sh, r2z3.sh): proof 2^12 8.2 -> 3.0 ms, 2^8 1.9 ms; with queries of 8.3 K points the PoE-shaped
    // proof over a 2^13 domain gains (3.67 -> 3.30 ms) and the one over a 2^14 domain loses as much (3.74 -> 4.12 ms), and MulChain 2^14
    // loses (4.97 -> 5.5 ms: the term over-estimates the warp kernel there) -- so it only applies below 2^13 points, where it is a pure
    // win.  Not for the batch twin, whose window was measured directly.  CTG_PLAN_FIXUP_TERM=0 turns it off.
    double fix = 0.0;
    static int use_fix = -1;
    if (use_fix < 0) { const char* e = getenv("CTG_PLAN_FIXUP_TERM"); use_fix = !(e && *e == '0'); }
    if (!batch_throughput && use_fix && nn < 8192) {
      const double t_add = point_bytes > sizeof(G1Affine) ? 165.0 : 30.0;
      const double slices = std::min(296.0 * 128.0, std::max(1.0, (double)nn * p.nwin / 16.0));
      const double q = slices / nb;
      fix = t_add * (q <= 1.0 ? 1.0 : (q <= 8.0 ? q : std::ceil(q / 32.0) + 5.0));
    }
    double cost = 0.0004 * p.nwin * (double)nn + red + lat + fix;
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
