// Bucket accumulation of the MSM pipeline (step 4 of msm.cuh) -- the dominant kernels of the prover.
// Kept in their own translation units (msm_g1_acc.cu / msm_g2_acc.cu) so that they compile in parallel
// with the rest.
//
// Two stages:
//  4a. batched-affine tree levels (ba_level_kernel).  Within a bucket the points are summed pairwise,
//      level by level, in AFFINE coordinates: P3 = P1 + P2 costs one inversion, 2M + 1S.  The
//      inversions of a thread's batch of pairs are shared (Montgomery's trick: 3M per pair + one
//      binary-GCD inversion per batch, fp.cuh inv_bingcd), so an addition costs 5M + 1S instead of
//      the 8M + 2S of the mixed XYZZ addition.  Every level halves the number of points per bucket;
//      level l reads the compact array of level l-1 (level 1 gathers the bases through the sorted
//      entry list) and writes a compact array whose bucket offsets are a prefix sum of ceil(cnt/2).
//  4b. msm_accumulate_kernel: what is left (cnt / 2^L points per bucket) is added into XYZZ buckets
//      by threads that walk equal slices of the (remaining) sorted list, as before.
#pragma once
#include "msm.cuh"

namespace ctg {

// ---- offsets of the next tree level: out_starts = exclusive scan of ceil(cnt_in / 2) --------------------
template <class F> __global__ void __launch_bounds__(256) ba_scan_tile_sum_kernel(const uint32_t* __restrict__ in_starts, uint32_t nb, uint32_t* tile_sums) {
  __shared__ uint32_t sh[256];
  const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * 8;
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < 8; k++)
    if (base + k < nb) s += (in_starts[base + k + 1] - in_starts[base + k] + 1) >> 1;
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int off = 128; off >= 1; off >>= 1) {
    if ((int)threadIdx.x < off) sh[threadIdx.x] += sh[threadIdx.x + off];
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = sh[0];
}
template <class F> __global__ void __launch_bounds__(256) ba_scan_apply_kernel(const uint32_t* __restrict__ in_starts, uint32_t* out_starts, uint32_t nb,
                                                                              const uint32_t* __restrict__ tile_sums, uint32_t ntiles) {
  __shared__ uint32_t sh[256];
  const uint32_t tid = threadIdx.x;
  const uint32_t base = blockIdx.x * kScanTile + tid * 8;
  uint32_t c[8], s = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) { c[k] = base + k < nb ? (in_starts[base + k + 1] - in_starts[base + k] + 1) >> 1 : 0; s += c[k]; }
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
    if (base + k < nb) out_starts[base + k] = run;
    run += c[k];
  }
  if (blockIdx.x == 0 && tid == 0) out_starts[nb] = tile_sums[ntiles];
}

// ---- compact mixed addition: the field products are CALLS into one out-of-line copy each instead of ten inlined 400-instruction
// bodies, so the accumulate loop is ~20 KB of SASS instead of ~64 KB and stays resident in the instruction caches
// (profiles/r01_msm_accumulate_g1_ncu.txt: sm__icc_request_hit_rate 84.7 %, no_instruction stalls 0.46 per issue with the inlined body).
template <class F> __device__ __noinline__ F fmul_call(const F a, const F b) { return a * b; }
template <class F> __device__ __noinline__ F fmul2_sub_call(const F a, const F b, const F c, const F d) { return F::mul2_sub(a, b, c, d); }
template <class F> __device__ __forceinline__ XYZZ<F> xyzz_madd_compact(const XYZZ<F>& acc, const Affine<F>& q_in, bool neg) {
  if (q_in.is_inf()) return acc;
  Affine<F> q = q_in;
  if (neg) q.y = q.y.neg();
  if (acc.is_inf()) return {q.x, q.y, F::one(), F::one()};
  F U2 = fmul_call(q.x, acc.ZZ), S2 = fmul_call(q.y, acc.ZZZ);
  F P = U2 - acc.X, R = S2 - acc.Y;
  if (P.is_zero()) {
    if (R.is_zero()) return xyzz_dbl_affine(q);
    return XYZZ<F>::inf();
  }
  F PP = fmul_call(P, P), PPP = fmul_call(P, PP), Q = fmul_call(acc.X, PP);
  XYZZ<F> r;
  r.X = fmul_call(R, R) - PPP - Q.dbl();
  r.Y = fmul2_sub_call(R, Q - r.X, acc.Y, PPP);
  r.ZZ = fmul_call(acc.ZZ, PP);
  r.ZZZ = fmul_call(acc.ZZZ, PPP);
  return r;
}

// field product for the big kernels: inlined for Fq, a call for Fq2 (whose inlined body is three times the size)
template <class F> struct AccCfg { static constexpr bool kCompact = sizeof(F) > 48; };
template <class F> __device__ __forceinline__ F fm(const F& a, const F& b) {
  if (AccCfg<F>::kCompact) return fmul_call(a, b);
  return a * b;
}

// ---- 4a. one level of the pairwise affine tree -----------------------------------------------------
// Pair classification (recomputed in both passes from the operands, no mults):
//   kCopy1 / kCopy2 : the other operand is missing (odd count) or the point at infinity -> result = the present one
//   kInf            : P2 == -P1 -> infinity;      kDbl : P2 == P1 -> tangent, denominator 2 y1;     kAdd : chord, denominator x2 - x1
enum { kAdd = 0, kDbl = 1, kCopy1 = 2, kCopy2 = 3, kInf = 4 };

template <class F> __device__ __forceinline__ int ba_classify(const Affine<F>& p1, const Affine<F>& p2, bool has2, F& d) {
  d = F::one();
  if (!has2 || p2.is_inf()) return kCopy1;
  if (p1.is_inf()) return kCopy2;
  F dx = p2.x - p1.x;
  if (!dx.is_zero()) { d = dx; return kAdd; }
  if (p1.y.eq(p2.y)) {
    F y2 = p1.y.dbl();
    if (y2.is_zero()) return kInf;   // 2-torsion point: cannot occur on BLS12-381 (odd group orders); never divide by zero
    d = y2;
    return kDbl;
  }
  return kInf;
}

template <class F, bool FIRST, int BMAX>
__global__ void __launch_bounds__(128) ba_level_kernel(const Affine<F>* __restrict__ bases, const uint32_t* __restrict__ entries,
                                                       const Affine<F>* __restrict__ in_pts, const uint32_t* __restrict__ in_starts,
                                                       const uint32_t* __restrict__ out_starts, uint32_t nb, Affine<F>* __restrict__ out_pts) {
  const uint32_t T = gridDim.x * blockDim.x;
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t O = out_starts[nb];
  const uint32_t K = (O + T - 1) / T;
  const uint64_t lo64 = (uint64_t)t * K;
  const uint32_t olo = lo64 < O ? (uint32_t)lo64 : O;
  const uint32_t ohi = (lo64 + K < O) ? (uint32_t)(lo64 + K) : O;
  if (olo >= ohi) return;
  // bucket containing output slot olo: largest gb with out_starts[gb] <= olo
  uint32_t a = 0, b = nb;
  while (b - a > 1) {
    uint32_t mid = (a + b) >> 1;
    if (out_starts[mid] <= olo) a = mid; else b = mid;
  }
  uint32_t gb = a;
  uint32_t ob = out_starts[gb], oe = out_starts[gb + 1], ib = in_starts[gb], ie = in_starts[gb + 1];

  auto load_in = [&](uint32_t i, Affine<F>& p) {
    if (FIRST) {
      const uint32_t e = entries[i];
      ld_affine(p, bases + (e & 0x7fffffffu));
      if (e >> 31) p.y = p.y.neg();     // -(0, 0) stays the point at infinity
    } else {
      ld_affine(p, in_pts + i);
    }
  };
  auto load_pair = [&](uint32_t o, Affine<F>& p1, Affine<F>& p2) -> bool {
    const uint32_t i0 = ib + 2 * (o - ob);
    const bool has2 = i0 + 1 < ie;
    load_in(i0, p1);
    if (has2) load_in(i0 + 1, p2); else p2 = Affine<F>::inf();
    return has2;
  };

  F prefix[BMAX];
  for (uint32_t blo = olo; blo < ohi; blo += BMAX) {
    const uint32_t bhi = blo + BMAX < ohi ? blo + BMAX : ohi;
    // ---- forward: running product of the denominators
    F run = F::one();
    for (uint32_t o = blo; o < bhi; o++) {
      while (o >= oe) { gb++; ob = oe; oe = out_starts[gb + 1]; ib = ie; ie = in_starts[gb + 1]; }
      Affine<F> p1, p2;
      const bool has2 = load_pair(o, p1, p2);
      F d;
      ba_classify(p1, p2, has2, d);
      prefix[o - blo] = run;
      run = fm(run, d);
    }
    F inv = run.inv_bingcd();
    // ---- backward: peel the inverses off, finish the additions
    for (uint32_t o = bhi; o-- > blo;) {
      while (o < ob) { gb--; oe = ob; ob = out_starts[gb]; ie = ib; ib = in_starts[gb]; }
      Affine<F> p1, p2;
      const bool has2 = load_pair(o, p1, p2);
      F d;
      const int kind = ba_classify(p1, p2, has2, d);
      const F dinv = fm(inv, prefix[o - blo]);
      inv = fm(inv, d);
      Affine<F> r;
      if (kind == kAdd || kind == kDbl) {
        F num;
        if (kind == kAdd) num = p2.y - p1.y;
        else { F xx = p1.x.sqr(); num = xx.dbl() + xx; }
        const F lam = fm(num, dinv);
        r.x = fm(lam, lam) - p1.x - p2.x;
        r.y = fm(lam, p1.x - r.x) - p1.y;
      } else if (kind == kCopy1) r = p1;
      else if (kind == kCopy2) r = p2;
      else r = Affine<F>::inf();
      copy16(out_pts[o], r);
    }
    // the backward walk may have stepped below the first bucket of the next batch: the forward walk re-advances from here
  }
}

// ---- 4b. bucket accumulation over what the tree left ------------------------------------------------
// DIRECT = false: the list is entries[] (point index | sign << 31) into bases[]; DIRECT = true: the list is the
// compact point array of the last tree level (bases = that array, entries unused).
template <class F, bool DIRECT, bool COMPACT>
__global__ void __launch_bounds__(128) msm_accumulate_kernel(const Affine<F>* __restrict__ bases, const uint32_t* __restrict__ entries,
                                                             const uint32_t* __restrict__ starts, uint32_t nb, XYZZ<F>* buckets,
                                                             XYZZ<F>* part_pt, uint32_t* part_gb) {
  const uint32_t T = gridDim.x * blockDim.x;
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t M = starts[nb];
  const uint32_t K = (M + T - 1) / T;
  const uint64_t lo64 = (uint64_t)t * K;
  const uint32_t lo = lo64 < M ? (uint32_t)lo64 : M;
  const uint32_t hi = (lo64 + K < M) ? (uint32_t)(lo64 + K) : M;
  part_gb[2 * t] = kNoBucket;
  part_gb[2 * t + 1] = kNoBucket;
  if (lo >= hi) return;
  // bucket containing entry `lo`: largest gb with starts[gb] <= lo
  uint32_t a = 0, b = nb;
  while (b - a > 1) {
    uint32_t mid = (a + b) >> 1;
    if (starts[mid] <= lo) a = mid; else b = mid;
  }
  uint32_t gb = a;
  uint32_t bucket_begin = starts[gb], bucket_end = starts[gb + 1];
  uint32_t seg_end = bucket_end < hi ? bucket_end : hi;
  int slot = 0;
  XYZZ<F> acc = XYZZ<F>::inf();

  auto flush = [&]() {
    if (bucket_begin >= lo && bucket_end <= hi) {
      copy16(buckets[gb], acc);
    } else {
      copy16(part_pt[2 * t + slot], acc);
      part_gb[2 * t + slot] = gb | (bucket_begin >= lo ? kHeadBit : 0u);
      slot++;
    }
  };

  uint32_t e = DIRECT ? lo : entries[lo];
  Affine<F> P;
  ld_affine(P, bases + (e & 0x7fffffffu));
  for (uint32_t pos = lo; pos < hi; pos++) {
    uint32_t e_next = e;
    Affine<F> P_next;
    if (pos + 1 < hi) e_next = DIRECT ? pos + 1 : entries[pos + 1];
    ld_affine(P_next, bases + (e_next & 0x7fffffffu));
    if (pos == seg_end) {
      flush();
      acc = XYZZ<F>::inf();
      do { gb++; bucket_begin = bucket_end; bucket_end = starts[gb + 1]; } while (bucket_end <= pos);
      seg_end = bucket_end < hi ? bucket_end : hi;
    }
    if (COMPACT) acc = xyzz_madd_compact(acc, P, !DIRECT && (e >> 31) != 0);
    else acc = xyzz_madd(acc, P, !DIRECT && (e >> 31) != 0);
    e = e_next;
    P = P_next;
  }
  flush();
}

template <class F> inline bool acc_compact() {   // default per field; CTG_ACC_COMPACT=0/1 overrides (sweeps)
  static int v = -1;
  if (v < 0) { const char* e = getenv("CTG_ACC_COMPACT"); v = e ? atoi(e) : (AccCfg<F>::kCompact ? 1 : 0); }
  return v != 0;
}
static constexpr int kBaBatch = 64;   // pairs per shared inversion (local-memory prefix array of 64 field elements per thread)

template <class F> int accumulate_max_blocks_per_sm() {
  int occ = 0;
  if (acc_compact<F>()) { if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, msm_accumulate_kernel<F, false, true>, 128, 0) != cudaSuccess) return 1; }
  else if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, msm_accumulate_kernel<F, false, false>, 128, 0) != cudaSuccess) return 1;
  return occ;
}
template <class F>
void accumulate_launch(int blocks, cudaStream_t st, bool direct, const Affine<F>* bases, const uint32_t* entries, const uint32_t* starts, uint32_t nb,
                       XYZZ<F>* buckets, XYZZ<F>* part_pt, uint32_t* part_gb) {
  if (acc_compact<F>()) {
    if (direct) msm_accumulate_kernel<F, true, true><<<blocks, 128, 0, st>>>(bases, entries, starts, nb, buckets, part_pt, part_gb);
    else msm_accumulate_kernel<F, false, true><<<blocks, 128, 0, st>>>(bases, entries, starts, nb, buckets, part_pt, part_gb);
  } else {
    if (direct) msm_accumulate_kernel<F, true, false><<<blocks, 128, 0, st>>>(bases, entries, starts, nb, buckets, part_pt, part_gb);
    else msm_accumulate_kernel<F, false, false><<<blocks, 128, 0, st>>>(bases, entries, starts, nb, buckets, part_pt, part_gb);
  }
}
template <class F> int ba_max_blocks_per_sm() {
  int o1 = 0, o2 = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o1, ba_level_kernel<F, true, kBaBatch>, 128, 0) != cudaSuccess) return 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, ba_level_kernel<F, false, kBaBatch>, 128, 0) != cudaSuccess) return 1;
  return o1 < o2 ? o1 : o2;
}
template <class F>
void ba_scan_launch(cudaStream_t st, const uint32_t* in_starts, uint32_t* out_starts, uint32_t nb, uint32_t* tile_sums) {
  const uint32_t ntiles = (nb + kScanTile - 1) / kScanTile;
  ba_scan_tile_sum_kernel<F><<<ntiles, 256, 0, st>>>(in_starts, nb, tile_sums);
  msm_scan_tiles_kernel<F><<<1, 1024, 0, st>>>(tile_sums, ntiles);
  ba_scan_apply_kernel<F><<<ntiles, 256, 0, st>>>(in_starts, out_starts, nb, tile_sums, ntiles);
}
template <class F>
void ba_level_launch(int blocks, cudaStream_t st, bool first, const Affine<F>* bases, const uint32_t* entries, const Affine<F>* in_pts,
                     const uint32_t* in_starts, const uint32_t* out_starts, uint32_t nb, Affine<F>* out_pts) {
  if (first) ba_level_kernel<F, true, kBaBatch><<<blocks, 128, 0, st>>>(bases, entries, in_pts, in_starts, out_starts, nb, out_pts);
  else ba_level_kernel<F, false, kBaBatch><<<blocks, 128, 0, st>>>(bases, entries, in_pts, in_starts, out_starts, nb, out_pts);
}

}  // namespace ctg
