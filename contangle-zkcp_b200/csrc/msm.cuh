// Pippenger multi-scalar multiplication with signed-digit windows, templated on the base
// field (Fq -> G1, Fq2 -> G2).  Replaces ark-ec 0.3.0 VariableBaseMSM::multi_scalar_mul
// (msm/variable_base.rs), called five times per proof by ark-groth16's create_proof, which
// the reference reaches from Groth16::prove at lib/src/zk/encryption.rs:76.
//
// Pipeline (all on one stream):
//   1. msm_count     scalar -> signed digits, histogram of (bucket set, |digit|-1)    [warp-aggregated atomics]
//   2. msm_scan      exclusive prefix sum -> bucket start offsets
//   3. msm_scatter   counting sort: entries[] = point index | sign << 31, grouped by bucket
//   4. msm_accumulate  the sorted entry list is cut into equal slices, one per thread; each thread
//                    walks its slice adding affine bases into an XYZZ accumulator (8M + 2S) and
//                    emits a bucket sum whenever the bucket id changes.  Work per thread is
//                    identical whatever the scalar distribution (zeros, ones, bytes, uniform).
//   5. msm_fixup     buckets cut by a slice boundary: their partial sums are added by one warp each
//   6. msm_halve x log2(B), msm_reduce_final   sum_b (b+1) * bucket[b] by halving trees of minimal serial depth
//   7. msm_horner  sum_g 2^(c g) S_g over the bucket sets (skipped when there is one set)
// Windows w = j * G + g use the precomputed multiple 2^(c G j) * P (copy j of the base table),
// so only G = ceil(W / copies) bucket sets exist; with copies = W there is a single set and no
// doubling chain at the end.
#pragma once
#include <type_traits>
#include <cstdlib>
#include "ctg_internal.h"
#include <cstdlib>

#include "msm_digits.cuh"

namespace ctg {

static constexpr uint32_t kNoBucket = 0xffffffffu;
static constexpr uint32_t kHeadBit = 0x80000000u;

template <class F> __device__ __forceinline__ void ld_affine(Affine<F>& p, const Affine<F>* src) {
  constexpr int NW = sizeof(Affine<F>) / 16;
  const uint4* q = reinterpret_cast<const uint4*>(src);
  uint4* d = reinterpret_cast<uint4*>(&p);
#pragma unroll
  for (int i = 0; i < NW; i++) d[i] = __ldg(q + i);
}
template <class T> __device__ __forceinline__ void copy16(T& dst, const T& src) {
  constexpr int NW = sizeof(T) / 16;
  const uint4* q = reinterpret_cast<const uint4*>(&src);
  uint4* d = reinterpret_cast<uint4*>(&dst);
#pragma unroll
  for (int i = 0; i < NW; i++) d[i] = q[i];
}

__device__ __forceinline__ Fr load_scalar(const uint32_t* scalars, uint64_t i, int mont) {
  const uint4* q = reinterpret_cast<const uint4*>(scalars + 8 * i);
  uint4 a = __ldg(q), b = __ldg(q + 1);
  Fr s;
  s.v[0] = a.x; s.v[1] = a.y; s.v[2] = a.z; s.v[3] = a.w;
  s.v[4] = b.x; s.v[5] = b.y; s.v[6] = b.z; s.v[7] = b.w;
  return mont ? s.from_mont() : s;
}

// ---- 1. histogram -------------------------------------------------------------------------
// blockIdx.y = index of the scalar vector in a batch (several MSMs over the SAME bases, one bucket-set group each):
// vector p reads scalars + p * vec_stride and owns the bucket ids [p * nsets * nbuckets_per_set, (p + 1) * ...)
template <class F> __global__ void msm_count_kernel(const uint32_t* __restrict__ scalars, uint64_t n, uint64_t vec_stride, int mont, MsmPlan plan, uint32_t* counts) {
  scalars += (uint64_t)blockIdx.y * vec_stride * 8;
  const uint32_t key_base = blockIdx.y * (uint32_t)plan.nsets * plan.nbuckets_per_set;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n32 = (n + 31) & ~31ull;
  const unsigned lane = threadIdx.x & 31;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += stride) {
    const bool valid = i < n;
    uint32_t k[8];
    if (valid) { Fr s = load_scalar(scalars, i, mont);
#pragma unroll
      for (int j = 0; j < 8; j++) k[j] = s.v[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; j++) k[j] = 0;
    }
    uint32_t carry = 0;
    for (int w = 0; w < plan.nwin; w++) {
      int32_t d = msm_signed_digit(k, plan.c, w, plan.nwin, carry);
      uint32_t key = kNoBucket;
      if (d != 0) key = key_base + (uint32_t)(w % plan.nsets) * plan.nbuckets_per_set + (uint32_t)((d < 0 ? -d : d) - 1);
      unsigned peers = __match_any_sync(0xffffffffu, key);
      if (key != kNoBucket && lane == (unsigned)(__ffs(peers) - 1)) atomicAdd(&counts[key], (uint32_t)__popc(peers));
    }
  }
}

// ---- 2. exclusive scan of the bucket sizes: per-tile sums, scan of the tile sums, per-tile scan ----
static constexpr int kScanTile = 2048;  // entries per CTA (256 threads x 8)
template <class F> __global__ void __launch_bounds__(256) msm_scan_tile_sum_kernel(const uint32_t* __restrict__ counts, uint32_t nb, uint32_t* tile_sums) {
  __shared__ uint32_t sh[256];
  const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * 8;
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) if (base + k < nb) s += counts[base + k];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int off = 128; off >= 1; off >>= 1) {
    if ((int)threadIdx.x < off) sh[threadIdx.x] += sh[threadIdx.x + off];
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = sh[0];
}
// one CTA: exclusive scan of up to 1024 tile sums in place; tile_sums[ntiles] = grand total
template <class F> __global__ void __launch_bounds__(1024) msm_scan_tiles_kernel(uint32_t* tile_sums, uint32_t ntiles) {
  __shared__ uint32_t sh[1024];
  const uint32_t tid = threadIdx.x;
  const uint32_t v = tid < ntiles ? tile_sums[tid] : 0;
  sh[tid] = v;
  __syncthreads();
  for (uint32_t off = 1; off < 1024; off <<= 1) {
    uint32_t x = tid >= off ? sh[tid - off] : 0;
    __syncthreads();
    sh[tid] += x;
    __syncthreads();
  }
  if (tid < ntiles) tile_sums[tid] = sh[tid] - v;
  if (tid == 1023) tile_sums[ntiles] = sh[1023];
}
template <class F> __global__ void __launch_bounds__(256) msm_scan_apply_kernel(uint32_t* counts, uint32_t* cursor, uint32_t nb, const uint32_t* __restrict__ tile_sums,
                                                                               uint32_t ntiles) {
  __shared__ uint32_t sh[256];
  const uint32_t tid = threadIdx.x;
  const uint32_t base = blockIdx.x * kScanTile + tid * 8;
  uint32_t c[8], s = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) { c[k] = base + k < nb ? counts[base + k] : 0; s += c[k]; }
  sh[tid] = s;
  __syncthreads();
  for (uint32_t off = 1; off < 256; off <<= 1) {
    uint32_t x = tid >= off ? sh[tid - off] : 0;
    __syncthreads();
    sh[tid] += x;
    __syncthreads();
  }
  uint32_t run = tile_sums[blockIdx.x] + sh[tid] - s;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    if (base + k < nb) { counts[base + k] = run; cursor[base + k] = run; }
    run += c[k];
  }
  if (blockIdx.x == 0 && tid == 0) counts[nb] = tile_sums[ntiles];
}

// ---- 3. scatter ---------------------------------------------------------------------------
template <class F> __global__ void msm_scatter_kernel(const uint32_t* __restrict__ scalars, uint64_t n, uint64_t vec_stride, uint64_t copy_stride, int mont, MsmPlan plan,
                                                      uint32_t* cursor, uint32_t* entries) {
  scalars += (uint64_t)blockIdx.y * vec_stride * 8;
  const uint32_t key_base = blockIdx.y * (uint32_t)plan.nsets * plan.nbuckets_per_set;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t n32 = (n + 31) & ~31ull;
  const unsigned lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += stride) {
    const bool valid = i < n;
    uint32_t k[8];
    if (valid) { Fr s = load_scalar(scalars, i, mont);
#pragma unroll
      for (int j = 0; j < 8; j++) k[j] = s.v[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; j++) k[j] = 0;
    }
    uint32_t carry = 0;
    for (int w = 0; w < plan.nwin; w++) {
      int32_t d = msm_signed_digit(k, plan.c, w, plan.nwin, carry);
      uint32_t key = kNoBucket;
      if (d != 0) key = key_base + (uint32_t)(w % plan.nsets) * plan.nbuckets_per_set + (uint32_t)((d < 0 ? -d : d) - 1);
      unsigned peers = __match_any_sync(0xffffffffu, key);
      int leader = __ffs(peers) - 1;
      uint32_t base = 0;
      if (key != kNoBucket && (int)lane == leader) base = atomicAdd(&cursor[key], (uint32_t)__popc(peers));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (key != kNoBucket) {
        uint32_t copy = (uint32_t)(w / plan.nsets);
        entries[base + __popc(peers & lt_mask)] = (uint32_t)(copy * copy_stride + i) | (d < 0 ? 0x80000000u : 0u);
      }
    }
  }
}

// ---- 4. bucket accumulation (dominant kernel): msm_acc.cuh, compiled in its own translation unit
template <class F> int accumulate_max_blocks_per_sm();
template <class F>
void accumulate_launch(int blocks, cudaStream_t st, bool direct, const Affine<F>* bases, const uint32_t* entries, const uint32_t* starts, uint32_t nb,
                       XYZZ<F>* buckets, XYZZ<F>* part_pt, uint32_t* part_gb);
// batched-affine tree levels in front of it (msm_acc.cuh, stage 4a)
template <class F> int ba_max_blocks_per_sm();
template <class F> void ba_scan_launch(cudaStream_t st, const uint32_t* in_starts, uint32_t* out_starts, uint32_t nb, uint32_t* tile_sums);
template <class F>
void ba_level_launch(int blocks, cudaStream_t st, bool first, const Affine<F>* bases, const uint32_t* entries, const Affine<F>* in_pts,
                     const uint32_t* in_starts, const uint32_t* out_starts, uint32_t nb, Affine<F>* out_pts);
// Number of tree levels for an MSM with `entries` bucket entries, `nb` buckets and `threads` level-kernel threads.
// A level pays while a thread still has >= 24 pairs to share an inversion over and buckets still hold several points -- and only
// over Fq2 (G2).  Measured on B200 (profiles/r01_batched_affine.txt): level 1 must gather every operand TWICE (denominators,
// then the additions), and random 96-byte gathers from a multi-GB table top out at 12-15 G/s (tools/gather_probe.cu), so for
// G1 (6 Fq products per addition) the level is gather-bound and no faster than the XYZZ kernel it replaces, while for G2
// (17 Fq products per addition against 28) it wins.  ctg_ctx_set_ba_levels / CTG_BA_LEVELS force a count for either group.
inline int ba_num_levels(uint64_t entries, uint64_t nb, uint64_t threads, size_t field_bytes) {
  static int forced = -2;
  if (forced == -2) { const char* e = getenv("CTG_BA_LEVELS"); forced = e ? atoi(e) : -1; }
  if (forced >= 0) return forced > 6 ? 6 : forced;
  if (field_bytes <= 48) return 0;
  int L = 0;
  uint64_t m = entries;
  while (L < 3 && (m / 2) / threads >= 24 && m / nb >= 3) { m = m / 2 + nb / 2; L++; }
  return L;
}

// ---- 5. buckets cut by slice boundaries -------------------------------------------------------
template <class F> __device__ __forceinline__ XYZZ<F> warp_sum(XYZZ<F> acc, XYZZ<F>* sh /* 32 per warp */, unsigned lane) {
  copy16(sh[lane], acc);
  __syncwarp();
  for (int off = 16; off >= 1; off >>= 1) {
    if ((int)lane < off) {
      XYZZ<F> o;
      copy16(o, sh[lane + off]);
      acc = xyzz_add(acc, o);
      copy16(sh[lane], acc);
    }
    __syncwarp();
  }
  return acc;
}

// 5a. one thread per run head: runs that end within kShortRun slots (the common case: a bucket cut once) are summed
//     sequentially; longer runs (huge buckets of skewed scalar distributions) get kLongBit for the warp kernel below.
static constexpr uint32_t kLongBit = 0x40000000u;
static constexpr uint32_t kTagMask = 0x3fffffffu;
static constexpr uint32_t kShortRun = 8;
template <class F> __global__ void __launch_bounds__(128) msm_fixup_short_kernel(const XYZZ<F>* part_pt, uint32_t* part_gb, uint32_t nslots, XYZZ<F>* buckets) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots) return;
  const uint32_t tag = part_gb[s];
  if (tag == kNoBucket || !(tag & kHeadBit)) return;
  const uint32_t gb = tag & kTagMask;
  XYZZ<F> acc;
  copy16(acc, part_pt[s]);
  uint32_t j = s + 1;
  const uint32_t lim = (s + 1 + kShortRun < nslots) ? s + 1 + kShortRun : nslots;
  bool ended = false;
  for (; j < lim; j++) {
    const uint32_t g = part_gb[j];
    if (g == kNoBucket) continue;
    if ((g & kTagMask) != gb) { ended = true; break; }
    XYZZ<F> o;
    copy16(o, part_pt[j]);
    acc = xyzz_add(acc, o);
  }
  if (ended || j >= nslots) copy16(buckets[gb], acc);
  else part_gb[s] = tag | kLongBit;
}

// 5b. long runs: one warp per run, lanes stride over the partials, then a shared-memory tree
template <class F> __global__ void __launch_bounds__(64) msm_fixup_kernel(const XYZZ<F>* part_pt, const uint32_t* __restrict__ part_gb, uint32_t nslots,
                                                                         XYZZ<F>* buckets) {
  __shared__ XYZZ<F> sh[2][32];
  const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const uint32_t slot0 = (blockIdx.x * 2 + wib) * 32;
  if (slot0 >= nslots) return;
  const uint32_t my = slot0 + lane;
  const uint32_t tag = my < nslots ? part_gb[my] : kNoBucket;
  unsigned heads = __ballot_sync(0xffffffffu, tag != kNoBucket && (tag & kLongBit));
  while (heads) {
    const int h = __ffs(heads) - 1;
    heads &= heads - 1;
    const uint32_t gb = __shfl_sync(0xffffffffu, tag, h) & kTagMask;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (uint32_t base = slot0 + h;; base += 32) {
      const uint32_t s = base + lane;
      const uint32_t g = s < nslots ? part_gb[s] : kNoBucket;
      const bool mine = g != kNoBucket && (g & kTagMask) == gb;
      if (mine) {
        XYZZ<F> o;
        copy16(o, part_pt[s]);
        acc = xyzz_add(acc, o);
      }
      const bool stop = s >= nslots || (g != kNoBucket && (g & kTagMask) != gb);
      if (__any_sync(0xffffffffu, stop)) break;
    }
    acc = warp_sum(acc, sh[wib], lane);
    if (lane == 0) copy16(buckets[gb], acc);
    __syncwarp();
  }
}

// ---- 6. bucket reduction ----------------------------------------------------------------------
// S = sum_b (b + 1) * B[b] over the 2^(c-1) buckets of a set.  A single group operation has a latency of ~15-20 us
// (14 dependent 381-bit products on one thread), so the reduction is organised for minimal SERIAL DEPTH:
//   S = total + sum_k 2^k * O_k,   O_k = sum of the odd-indexed items of B^(k),   B^(k+1)[i] = B^(k)[2i] + B^(k)[2i+1]
// (because sum_b b B[b] = 2 * sum_i i B'[i] + sum of odd items).  One launch of msm_halve_kernel performs one halving
// step of the main array AND of every O_k tree that already exists, so after log2(N) + 1 launches every O_k and the
// total are known; msm_reduce_final_kernel then lets thread k double O_k k times and tree-adds the results.
// Depth: log2(N) + 1 additions, <= c - 2 doublings, 5 additions -- instead of the O(L * levels) of running sums.
// Sets with more than 2^16 buckets are first shrunk by msm_reduce_chunk_kernel (running sums over chunks of L = 2^logL
// buckets, throughput-bound): sum_b b B[b] = sum_t wt_t + L * sum_t t run_t.
template <class F, int TB> __global__ void __launch_bounds__(TB) msm_reduce_chunk_kernel(const XYZZ<F>* __restrict__ items, uint64_t set_stride, int logL,
                                                                                        XYZZ<F>* run_out, XYZZ<F>* wt_out, uint32_t T) {
  const uint32_t set = blockIdx.y;
  const uint32_t t = blockIdx.x * TB + threadIdx.x;
  if (t >= T) return;
  const XYZZ<F>* it = items + (uint64_t)set * set_stride + ((uint64_t)t << logL);
  XYZZ<F> run = XYZZ<F>::inf(), wt = XYZZ<F>::inf();
  for (int j = (1 << logL) - 1; j >= 1; j--) {   // wt = sum_j j * item[j]
    XYZZ<F> o;
    copy16(o, it[j]);
    run = xyzz_add(run, o);
    wt = xyzz_add(wt, run);
  }
  XYZZ<F> o;
  copy16(o, it[0]);
  run = xyzz_add(run, o);
  copy16(run_out[(uint64_t)set * T + t], run);
  copy16(wt_out[(uint64_t)set * T + t], wt);
}

static constexpr int kMaxHalveSegs = 24;
struct HalveSeg {
  const void* src;
  void* dst;
  uint32_t src_set_stride, dst_set_stride;  // in points
  uint32_t step, off;                       // operands: src[(2i) * step + off], src[(2i + 1) * step + off]
  uint32_t out_count, first_thread;
  uint32_t copy;                            // 1: dst[0] = src[off] (a tree that starts with a single item)
};
struct HalveArgs { int nseg; HalveSeg s[kMaxHalveSegs]; };

template <class F> __global__ void __launch_bounds__(64) msm_halve_kernel(HalveArgs a) {
  const uint32_t t = blockIdx.x * 64 + threadIdx.x;
  const uint32_t set = blockIdx.y;
  int k = -1;
#pragma unroll 1
  for (int i = 0; i < a.nseg; i++)
    if (t >= a.s[i].first_thread && t - a.s[i].first_thread < a.s[i].out_count) k = i;
  if (k < 0) return;
  const HalveSeg& s = a.s[k];
  const uint32_t i = t - s.first_thread;
  const XYZZ<F>* src = reinterpret_cast<const XYZZ<F>*>(s.src) + (uint64_t)set * s.src_set_stride;
  XYZZ<F>* dst = reinterpret_cast<XYZZ<F>*>(s.dst) + (uint64_t)set * s.dst_set_stride;
  XYZZ<F> x;
  copy16(x, src[(uint64_t)(2 * i) * s.step + s.off]);
  if (!s.copy) {
    XYZZ<F> y;
    copy16(y, src[(uint64_t)(2 * i + 1) * s.step + s.off]);
    x = xyzz_add(x, y);
  }
  copy16(dst[i], x);
}

// finals[set][slot]: slot j < nplain are added as they are; slot nplain + k is doubled (k + shift) times first.
template <class F> __global__ void __launch_bounds__(32) msm_reduce_final_kernel(const XYZZ<F>* finals, int nslots, int nplain, int shift, XYZZ<F>* set_out) {
  __shared__ XYZZ<F> sh[32];
  const unsigned lane = threadIdx.x;
  XYZZ<F> acc = XYZZ<F>::inf();
  if ((int)lane < nslots) {
    copy16(acc, finals[(uint64_t)blockIdx.x * nslots + lane]);
    if ((int)lane >= nplain && !acc.is_inf()) {
      const int d = (int)lane - nplain + shift;
      for (int i = 0; i < d; i++) acc = xyzz_dbl(acc);
    }
  }
  acc = warp_sum(acc, sh, lane);
  if (lane == 0) copy16(set_out[blockIdx.x], acc);
}

}  // namespace ctg
#include "msm_pair.cuh"   // lane-paired G2 versions of the two latency-bound kernels above
namespace ctg {
// sets with more buckets than 2^this are first shrunk by msm_reduce_chunk_kernel (CTG_REDUCE_CHUNK_LOG, default 19)
inline int reduce_chunk_log() {
  static int v = -1;
  // measured on B200 (tools/round2/r2u.sh): without the chunk step the first halving levels are wide enough to be throughput-bound by
  // themselves -- G2 2^20 reduce 4.0 -> 3.3 ms, G1 2^22 1.6 -> 1.4 ms -- so the default only chunks sets above 2^19 buckets
  if (v < 0) { const char* e = getenv("CTG_REDUCE_CHUNK_LOG"); v = e ? atoi(e) : 19; if (v < 10) v = 10; if (v > 20) v = 20; }
  return v;
}
template <class F> struct PairOpsOf;
template <> struct PairOpsOf<Fq> { typedef G1PairOps type; };
template <> struct PairOpsOf<Fq2> { typedef G2PairOps type; };
// CTG_G1_PAIR=0 / CTG_G2_PAIR=0 fall back to one thread per group operation (A/B runs)
template <class F> inline bool pair_reduce() {
  static int v = -1;
  if (v < 0) { const char* e = getenv(std::is_same<F, Fq2>::value ? "CTG_G2_PAIR" : "CTG_G1_PAIR"); v = !(e && *e == '0'); }
  return v != 0;
}

// ---- 7. window combine --------------------------------------------------------------------------
// sum_g 2^(c g) S_g, single thread (G - 1) * c doublings; G = 1 when every window is precomputed
template <class F> __global__ void msm_horner_kernel(const XYZZ<F>* sets, int nsets, int c, XYZZ<F>* out, uint32_t batch) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;   // one thread per scalar vector of the batch
  if (v >= batch) return;
  sets += (size_t)v * nsets;
  out += v;
  XYZZ<F> acc;
  copy16(acc, sets[nsets - 1]);
  for (int g = nsets - 2; g >= 0; g--) {
    for (int k = 0; k < c; k++) acc = xyzz_dbl(acc);
    XYZZ<F> o;
    copy16(o, sets[g]);
    acc = xyzz_add(acc, o);
  }
  copy16(*out, acc);
}

// ---- small utilities ----------------------------------------------------------------------------
template <class F> __global__ void to_affine_kernel(const XYZZ<F>* in, Affine<F>* out, int count) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  XYZZ<F> p;
  copy16(p, in[i]);
  Affine<F> a = xyzz_to_affine(p);
  copy16(out[i], a);
}

// out = affine(sum_i in[i]); one warp
template <class F> __global__ void __launch_bounds__(32) combine_kernel(const XYZZ<F>* in, uint32_t k, Affine<F>* out_affine, XYZZ<F>* out_xyzz) {
  __shared__ XYZZ<F> sh[32];
  const unsigned lane = threadIdx.x;
  XYZZ<F> acc = XYZZ<F>::inf();
  for (uint32_t i = lane; i < k; i += 32) {
    XYZZ<F> o;
    copy16(o, in[i]);
    acc = xyzz_add(acc, o);
  }
  acc = warp_sum(acc, sh, lane);
  if (lane == 0) {
    if (out_xyzz) copy16(*out_xyzz, acc);
    if (out_affine) {
      Affine<F> a = xyzz_to_affine(acc);
      copy16(*out_affine, a);
    }
  }
}

// P_i = (a + i*b) * g for i in [0, n): each thread owns CH consecutive points
template <class F, int CH> __global__ void __launch_bounds__(128) progression_kernel(Affine<F> g, Affine<F> step, Fr a, Fr b, uint64_t n, Affine<F>* out) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t i0 = t * CH;
  if (i0 >= n) return;
  Fr idx = Fr::zero();
  idx.v[0] = (uint32_t)i0;
  idx.v[1] = (uint32_t)(i0 >> 32);
  Fr k = (a + b * idx.to_mont()).from_mont();
  XYZZ<F> pts[CH];
  XYZZ<F> acc = xyzz_mul(XYZZ<F>::from_affine(g), k.v);
  F prod[CH];
  F run = F::one();
#pragma unroll 1
  for (int j = 0; j < CH; j++) {
    pts[j] = acc;
    prod[j] = run;                        // product of the denominators before j
    if (!acc.is_inf()) run = run * (acc.ZZ * acc.ZZZ);
    acc = xyzz_madd(acc, step);
  }
  F inv = run.inv_bingcd();
#pragma unroll 1
  for (int j = CH - 1; j >= 0; j--) {
    if (i0 + j >= n) { if (!pts[j].is_inf()) inv = inv * (pts[j].ZZ * pts[j].ZZZ); continue; }
    Affine<F> o = Affine<F>::inf();
    if (!pts[j].is_inf()) {
      F t = inv * prod[j];                // 1 / (ZZ * ZZZ)
      inv = inv * (pts[j].ZZ * pts[j].ZZZ);
      o.x = pts[j].X * (t * pts[j].ZZZ);
      o.y = pts[j].Y * (t * pts[j].ZZ);
    }
    copy16(out[i0 + j], o);
  }
}

// copy j+1 = 2^(shift) * copy j, for j < copies-1; thread per base point; batch inversion per thread
template <class F, int MAXC> __global__ void __launch_bounds__(128) precompute_kernel(Affine<F>* pts, uint64_t n, int copies, int shift) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Affine<F> p0;
  copy16(p0, pts[i]);
  XYZZ<F> cur = XYZZ<F>::from_affine(p0);
  XYZZ<F> tab[MAXC];
  F prod[MAXC];
  F run = F::one();
#pragma unroll 1
  for (int j = 1; j < copies; j++) {
    for (int k = 0; k < shift; k++) cur = xyzz_dbl(cur);
    tab[j] = cur;
    prod[j] = run;
    if (!cur.is_inf()) run = run * (cur.ZZ * cur.ZZZ);
  }
  F inv = run.inv_bingcd();
#pragma unroll 1
  for (int j = copies - 1; j >= 1; j--) {
    Affine<F> o = Affine<F>::inf();
    if (!tab[j].is_inf()) {
      F t = inv * prod[j];
      inv = inv * (tab[j].ZZ * tab[j].ZZZ);
      o.x = tab[j].X * (t * tab[j].ZZZ);
      o.y = tab[j].Y * (t * tab[j].ZZ);
    }
    copy16(pts[(uint64_t)j * n + i], o);
  }
}

// ---- host orchestration -------------------------------------------------------------------------
template <class F> struct MsmHost {

  static int32_t ensure_ws(ctg_ctx* ctx, MsmWorkspace& ws, const MsmPlan& plan, uint64_t n, int max_threads, int ba_levels, uint32_t batch) {
    const size_t nb = (size_t)batch * plan.nsets * plan.nbuckets_per_set;
    const size_t ne = (size_t)n * plan.nwin * batch;
    if (nb + 1 > ws.cap_nb) {
      if (ws.counts) cudaFree(ws.counts);
      if (ws.cursor) cudaFree(ws.cursor);
      CTG_CUDA(ctx, cudaMalloc(&ws.counts, (nb + 1) * sizeof(uint32_t)));
      CTG_CUDA(ctx, cudaMalloc(&ws.cursor, (nb + 1) * sizeof(uint32_t)));
      ws.cap_nb = nb + 1;
    }
    if (ne > ws.cap_entries) {
      if (ws.entries) cudaFree(ws.entries);
      CTG_CUDA(ctx, cudaMalloc(&ws.entries, (ne ? ne : 1) * sizeof(uint32_t)));
      ws.cap_entries = ne;
    }
    const size_t bb = nb * sizeof(XYZZ<F>);
    if (bb > ws.cap_bucket_bytes) {
      if (ws.buckets) cudaFree(ws.buckets);
      CTG_CUDA(ctx, cudaMalloc(&ws.buckets, bb));
      ws.cap_bucket_bytes = bb;
    }
    const size_t pb = (size_t)2 * max_threads * sizeof(XYZZ<F>);
    if (pb > ws.cap_part_bytes) {
      if (ws.part_pt) cudaFree(ws.part_pt);
      if (ws.part_gb) cudaFree(ws.part_gb);
      CTG_CUDA(ctx, cudaMalloc(&ws.part_pt, pb));
      CTG_CUDA(ctx, cudaMalloc(&ws.part_gb, (size_t)2 * max_threads * sizeof(uint32_t)));
      ws.cap_part_bytes = pb;
    }
    // reduction scratch: chunk sums (2 * 2^16) + halving trees (main, weights, O_k: < 4 N points in all) per set
    const size_t red_cap = (size_t)1 << reduce_chunk_log();
    const size_t red_n = plan.nbuckets_per_set > red_cap ? red_cap : plan.nbuckets_per_set;
    const size_t cb = ((size_t)batch * plan.nsets * (6 * red_n + 64) + 256) * sizeof(XYZZ<F>);
    if (cb > ws.cap_chunk_bytes) {
      if (ws.chunk) cudaFree(ws.chunk);
      CTG_CUDA(ctx, cudaMalloc(&ws.chunk, cb));
      ws.cap_chunk_bytes = cb;
    }
    if (ba_levels > 0) {
      // level l holds at most ne / 2^l + nb points (every bucket rounds up once per level)
      const size_t need[2] = {(ne / 2 + nb + 1) * sizeof(Affine<F>), (ne / 4 + nb + 1) * sizeof(Affine<F>)};
      for (int k = 0; k < 2; k++) {
        if (k == 1 && ba_levels < 2) break;
        if (need[k] > ws.cap_ba_bytes[k]) {
          if (ws.ba_pts[k]) cudaFree(ws.ba_pts[k]);
          ws.ba_pts[k] = nullptr; ws.cap_ba_bytes[k] = 0;
          if (cudaMalloc(&ws.ba_pts[k], need[k]) != cudaSuccess) { cudaGetLastError(); return fail(ctx, CTG_ERR_OOM, "msm: out of device memory for the batched-affine level arrays"); }
          ws.cap_ba_bytes[k] = need[k];
        }
      }
      if (nb + 1 > ws.cap_ba_nb) {
        for (int k = 0; k < 2; k++) {
          if (ws.ba_starts[k]) cudaFree(ws.ba_starts[k]);
          CTG_CUDA(ctx, cudaMalloc(&ws.ba_starts[k], (nb + 1) * sizeof(uint32_t)));
        }
        ws.cap_ba_nb = nb + 1;
      }
    }
    if (!ws.result) CTG_CUDA(ctx, cudaMalloc(&ws.result, sizeof(XYZZ<Fq2>)));
    if (!ws.ev_sorted) CTG_CUDA(ctx, cudaEventCreateWithFlags(&ws.ev_sorted, cudaEventDisableTiming));
    if (!ws.tile_sums) CTG_CUDA(ctx, cudaMalloc(&ws.tile_sums, 1025 * sizeof(uint32_t)));
    return CTG_OK;
  }

  // batch > 1: `batch` scalar vectors (vector p at d_scalars + p * vec_stride elements) against the same bases; d_out[batch].
  // One sort, one accumulate pass and one reduction serve the whole batch: vector p owns its own group of bucket sets.
  // sorted != nullptr: another MSM over the SAME scalars and an identical plan / copy stride already sorted them into
  // sorted->counts / ->entries (the caller has made `st` wait for sorted->ev_sorted); the digit sort is skipped.
  static int32_t run(ctg_ctx* ctx, MsmWorkspace& ws, const Bases& bases, const uint32_t* d_scalars, uint64_t n, bool mont, XYZZ<F>* d_out,
                     cudaStream_t st, float* stage_ms, uint32_t batch = 1, uint64_t vec_stride = 0, const MsmWorkspace* sorted = nullptr) {
    const MsmPlan& plan = bases.plan;
    if (n > bases.n) return fail(ctx, CTG_ERR_INVALID, "msm: more scalars than bases");
    if (batch < 1) return fail(ctx, CTG_ERR_INVALID, "msm: empty batch");
    const uint64_t cstride = bases.stride ? bases.stride : bases.n;
    if (cstride * plan.copies >= (1ull << 31)) return fail(ctx, CTG_ERR_INVALID, "msm: n * copies must be < 2^31");
    if ((uint64_t)n * plan.nwin * batch >= (1ull << 32)) return fail(ctx, CTG_ERR_INVALID, "msm: n * windows (* batch) must be < 2^32");
    if (n == 0) {
      CTG_CUDA(ctx, cudaMemsetAsync(d_out, 0, (size_t)batch * sizeof(XYZZ<F>), st));
      return CTG_OK;
    }
    // accumulate launch geometry
    static int occ = 0, ba_occ = 0;
    if (!occ) {
      occ = accumulate_max_blocks_per_sm<F>();
      if (occ < 1) occ = 1;
      ba_occ = ba_max_blocks_per_sm<F>();
      if (ba_occ < 1) ba_occ = 1;
    }
    const int max_blocks = ctx->sm_count * occ;
    const int max_threads = max_blocks * 128;
    if ((uint64_t)batch * plan.nsets * plan.nbuckets_per_set > 1023ull * kScanTile) return fail(ctx, CTG_ERR_INVALID, "msm: too many buckets for the scan (c or the batch too large)");
    const uint32_t nb = batch * (uint32_t)plan.nsets * plan.nbuckets_per_set;
    const int ba_blocks = ctx->sm_count * ba_occ;
    const int ba_levels = ctx->ba_levels >= 0 ? (ctx->ba_levels > 6 ? 6 : ctx->ba_levels) : ba_num_levels((uint64_t)n * plan.nwin * batch, nb, (uint64_t)ba_blocks * 128, sizeof(F));
    CTG_TRY(ensure_ws(ctx, ws, plan, n, max_threads, ba_levels, batch));
    XYZZ<F>* buckets = reinterpret_cast<XYZZ<F>*>(ws.buckets);
    XYZZ<F>* part_pt = reinterpret_cast<XYZZ<F>*>(ws.part_pt);
    XYZZ<F>* chunk = reinterpret_cast<XYZZ<F>*>(ws.chunk);
    const bool timed = stage_ms != nullptr;
    if (timed && !ctx->events_ready) {
      for (auto& e : ctx->ev) CTG_CUDA(ctx, cudaEventCreate(&e));
      ctx->events_ready = true;
    }
    if (timed) CTG_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));

    const uint32_t* starts = ws.counts;      // exclusive bucket offsets, [nb + 1]
    const uint32_t* entries = ws.entries;
    if (sorted) {
      starts = sorted->counts;
      entries = sorted->entries;
    } else {
      CTG_CUDA(ctx, cudaMemsetAsync(ws.counts, 0, (size_t)(nb + 1) * sizeof(uint32_t), st));
      const int dig_x = (int)std::min<uint64_t>((n + 255) / 256, std::max<uint64_t>(1, (uint64_t)ctx->sm_count * 8 / batch));
      const dim3 dig_grid(dig_x, batch);
      msm_count_kernel<F><<<dig_grid, 256, 0, st>>>(d_scalars, n, vec_stride, mont ? 1 : 0, plan, ws.counts);
      CTG_KERNEL_CHECK(ctx);
      const uint32_t ntiles = (nb + kScanTile - 1) / kScanTile;
      msm_scan_tile_sum_kernel<F><<<ntiles, 256, 0, st>>>(ws.counts, nb, ws.tile_sums);
      CTG_KERNEL_CHECK(ctx);
      msm_scan_tiles_kernel<F><<<1, 1024, 0, st>>>(ws.tile_sums, ntiles);
      CTG_KERNEL_CHECK(ctx);
      msm_scan_apply_kernel<F><<<ntiles, 256, 0, st>>>(ws.counts, ws.cursor, nb, ws.tile_sums, ntiles);
      CTG_KERNEL_CHECK(ctx);
      msm_scatter_kernel<F><<<dig_grid, 256, 0, st>>>(d_scalars, n, vec_stride, cstride, mont ? 1 : 0, plan, ws.cursor, ws.entries);
      CTG_KERNEL_CHECK(ctx);
      CTG_CUDA(ctx, cudaEventRecord(ws.ev_sorted, st));
    }
    if (timed) CTG_CUDA(ctx, cudaEventRecord(ctx->ev[1], st));

    CTG_CUDA(ctx, cudaMemsetAsync(ws.buckets, 0, (size_t)nb * sizeof(XYZZ<F>), st));
    // 4a. batched-affine tree levels: level l reads level l-1 (level 1 the bases through the entry list)
    const uint32_t* cur_starts = starts;
    const Affine<F>* cur_pts = nullptr;
    for (int l = 1; l <= ba_levels; l++) {
      uint32_t* out_starts = ws.ba_starts[(l - 1) & 1];
      Affine<F>* out_pts = reinterpret_cast<Affine<F>*>(ws.ba_pts[(l - 1) & 1]);
      ba_scan_launch<F>(st, cur_starts, out_starts, nb, ws.tile_sums);
      ctx->launches += 2;
      CTG_KERNEL_CHECK(ctx);
      ba_level_launch<F>(ba_blocks, st, l == 1, reinterpret_cast<const Affine<F>*>(bases.pts), entries, cur_pts, cur_starts, out_starts, nb, out_pts);
      CTG_KERNEL_CHECK(ctx);
      cur_starts = out_starts;
      cur_pts = out_pts;
    }
    // 4b. XYZZ accumulation of what is left
    uint64_t want_threads = ((((uint64_t)n * plan.nwin * batch) >> ba_levels) + 15) / 16;
    int blocks = (int)std::min<uint64_t>((want_threads + 127) / 128, (uint64_t)max_blocks);
    if (blocks < 1) blocks = 1;
    const uint32_t nslots = (uint32_t)blocks * 128 * 2;
    if (ba_levels > 0) accumulate_launch<F>(blocks, st, true, cur_pts, nullptr, cur_starts, nb, buckets, part_pt, ws.part_gb);
    else accumulate_launch<F>(blocks, st, false, reinterpret_cast<const Affine<F>*>(bases.pts), entries, starts, nb, buckets, part_pt, ws.part_gb);
    CTG_KERNEL_CHECK(ctx);
    msm_fixup_short_kernel<F><<<(nslots + 127) / 128, 128, 0, st>>>(part_pt, ws.part_gb, nslots, buckets);
    CTG_KERNEL_CHECK(ctx);
    msm_fixup_kernel<F><<<(nslots / 32 + 1) / 2, 64, 0, st>>>(part_pt, ws.part_gb, nslots, buckets);
    CTG_KERNEL_CHECK(ctx);
    if (timed) CTG_CUDA(ctx, cudaEventRecord(ctx->ev[2], st));

    // bucket reduction (see section 6)
    {
      const int nsets = plan.nsets * (int)batch;   // every vector of a batch has its own group of plan.nsets bucket sets
      XYZZ<F>* pool = chunk;   // bump allocator over the reduction scratch, in points
      size_t used = 0;
      auto take = [&](size_t pts) { XYZZ<F>* p = pool + used; used += pts * (size_t)nsets; return p; };
      const XYZZ<F>* cur = buckets;
      uint32_t cur_stride = plan.nbuckets_per_set, N = plan.nbuckets_per_set;
      int shift = 0;
      const XYZZ<F>* wt_src = nullptr;
      uint32_t wt_count = 0;
      const uint32_t chunk_cap = 1u << reduce_chunk_log();
      if (N > chunk_cap) {   // shrink to chunk_cap items with chunked running sums
        shift = 0;
        while ((N >> shift) > chunk_cap) shift++;
        const uint32_t T = N >> shift;
        XYZZ<F>* run = take(T);
        XYZZ<F>* wt = take(T);
        msm_reduce_chunk_kernel<F, 64><<<dim3((T + 63) / 64, nsets), 64, 0, st>>>(cur, cur_stride, shift, run, wt, T);
        CTG_KERNEL_CHECK(ctx);
        cur = run; cur_stride = T; N = T;
        wt_src = wt; wt_count = T;
      }
      // slots: 0 = total, 1 = sum of chunk weights (or infinity), 2 + k = O_k
      int nO = 0;
      while ((1u << nO) < N) nO++;          // log2(N): O_0 .. O_{log2 N - 1}
      const int nslots = 2 + nO;
      if (nslots > 32) return fail(ctx, CTG_ERR_STATE, "msm: too many reduction slots");
      XYZZ<F>* finals = take(nslots);
      CTG_CUDA(ctx, cudaMemsetAsync(finals, 0, (size_t)nslots * nsets * sizeof(XYZZ<F>), st));
      struct Tree { const XYZZ<F>* src; uint32_t stride, count, step, off; int slot; };
      std::vector<Tree> trees;
      trees.push_back({cur, cur_stride, N, 1, 0, 0});                       // main array -> total
      if (wt_src) trees.push_back({wt_src, wt_count, wt_count, 1, 0, 1});   // plain sum of the chunk weights
      int next_O = 0;
      while (!trees.empty()) {
        HalveArgs args{};
        uint32_t threads = 0;
        std::vector<Tree> next;
        for (Tree& tr : trees) {
          if (args.nseg >= kMaxHalveSegs) return fail(ctx, CTG_ERR_STATE, "msm: too many reduction segments");
          HalveSeg& sg = args.s[args.nseg++];
          sg.src = tr.src; sg.src_set_stride = tr.stride; sg.step = tr.step; sg.off = tr.off; sg.first_thread = threads;
          if (tr.count == 1) {            // single item: copy it to its slot
            sg.copy = 1; sg.out_count = 1; sg.dst = finals + tr.slot; sg.dst_set_stride = nslots;
          } else {
            sg.copy = 0; sg.out_count = tr.count / 2;
            if (sg.out_count == 1) { sg.dst = finals + tr.slot; sg.dst_set_stride = nslots; }
            else {
              XYZZ<F>* d = take(sg.out_count);
              sg.dst = d; sg.dst_set_stride = sg.out_count;
              next.push_back({d, sg.out_count, sg.out_count, 1, 0, tr.slot});
            }
          }
          threads += sg.out_count;
          // the main array spawns the tree of the odd items of the level it is reading now
          if (tr.slot == 0 && tr.count >= 2) {
            next.push_back({tr.src, tr.stride, tr.count / 2, 2 * tr.step, tr.off + tr.step, 2 + next_O});
            next_O++;
          }
        }
        if ((used + 64) * sizeof(XYZZ<F>) > ws.cap_chunk_bytes) return fail(ctx, CTG_ERR_STATE, "msm: reduction scratch too small");
        // lane pairs on every level, wide ones too: measured for G2 at 2^20 (2^19 buckets, no chunk step) 3.3 ms all-paired against
        // 4.3 ms with one thread per addition on the levels above 16 K outputs (the 255-register single-thread kernel holds half the warps)
        if (pair_reduce<F>()) msm_halve_pair_kernel<typename PairOpsOf<F>::type><<<dim3((2 * threads + 127) / 128, nsets), 128, 0, st>>>(args);
        else msm_halve_kernel<F><<<dim3((threads + 63) / 64, nsets), 64, 0, st>>>(args);
        CTG_KERNEL_CHECK(ctx);
        trees.swap(next);
      }
      XYZZ<F>* set_out = plan.nsets == 1 ? d_out : take(1);
      if (pair_reduce<F>()) msm_reduce_final_pair_kernel<typename PairOpsOf<F>::type><<<nsets, 64, 0, st>>>(finals, nslots, 2, shift, set_out);
      else msm_reduce_final_kernel<F><<<nsets, 32, 0, st>>>(finals, nslots, 2, shift, set_out);
      CTG_KERNEL_CHECK(ctx);
      if (plan.nsets > 1) {
        msm_horner_kernel<F><<<(batch + 31) / 32, 32, 0, st>>>(set_out, plan.nsets, plan.c, d_out, batch);
        CTG_KERNEL_CHECK(ctx);
      }
    }
    if (timed) {
      CTG_CUDA(ctx, cudaEventRecord(ctx->ev[3], st));
      CTG_CUDA(ctx, cudaEventSynchronize(ctx->ev[3]));
      for (int i = 0; i < 3; i++) CTG_CUDA(ctx, cudaEventElapsedTime(&stage_ms[i], ctx->ev[i], ctx->ev[i + 1]));
    }
    return CTG_OK;
  }

  static int32_t precompute(ctg_ctx* ctx, Bases& b, cudaStream_t st) {
    if (b.plan.copies <= 1) return CTG_OK;
    if (b.plan.copies > 32) return fail(ctx, CTG_ERR_INVALID, "bases: at most 32 precomputed copies");
    const int shift = b.plan.c * b.plan.nsets;
    precompute_kernel<F, 32><<<(unsigned)((b.n + 127) / 128), 128, 0, st>>>(reinterpret_cast<Affine<F>*>(b.pts), b.n, b.plan.copies, shift);
    CTG_KERNEL_CHECK(ctx);
    return CTG_OK;
  }

  static int32_t to_affine(ctg_ctx* ctx, const XYZZ<F>* in, Affine<F>* out, int count, cudaStream_t st) {
    to_affine_kernel<F><<<(count + 31) / 32, 32, 0, st>>>(in, out, count);
    CTG_KERNEL_CHECK(ctx);
    return CTG_OK;
  }

  static int32_t progression(ctg_ctx* ctx, const Affine<F>& g, const Fr& a, const Fr& b, uint64_t n, Affine<F>* d_out, cudaStream_t st) {
    constexpr int CH = 16;
    Fr bc = b.from_mont();
    Affine<F> step = xyzz_to_affine(xyzz_mul(XYZZ<F>::from_affine(g), bc.v));  // host-side (emulated) one-off
    const uint64_t threads = (n + CH - 1) / CH;
    progression_kernel<F, CH><<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(g, step, a, b, n, d_out);
    CTG_KERNEL_CHECK(ctx);
    return CTG_OK;
  }
};

}  // namespace ctg
