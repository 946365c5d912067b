// Bucket accumulation kernel of the MSM pipeline (step 4 of msm.cuh) -- the dominant kernel of the
// prover.  Kept in its own translation unit (msm_g1_acc.cu / msm_g2_acc.cu) so that it compiles
// in parallel with the rest.
#pragma once
#include "msm.cuh"

namespace ctg {

// ---- 4. bucket accumulation (dominant kernel) ------------------------------------------------
template <class F, bool PREFETCH>
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

  uint32_t e = entries[lo];
  Affine<F> P;
  ld_affine(P, bases + (e & 0x7fffffffu));
  for (uint32_t pos = lo; pos < hi; pos++) {
    uint32_t e_next = e;
    Affine<F> P_next;
    if (PREFETCH) {
      if (pos + 1 < hi) e_next = entries[pos + 1];
      ld_affine(P_next, bases + (e_next & 0x7fffffffu));
    }
    if (pos == seg_end) {
      flush();
      acc = XYZZ<F>::inf();
      do { gb++; bucket_begin = bucket_end; bucket_end = starts[gb + 1]; } while (bucket_end <= pos);
      seg_end = bucket_end < hi ? bucket_end : hi;
    }
    acc = xyzz_madd(acc, P, (e >> 31) != 0);
    if (PREFETCH) {
      e = e_next;
      P = P_next;
    } else if (pos + 1 < hi) {
      e = entries[pos + 1];
      ld_affine(P, bases + (e & 0x7fffffffu));
    }
  }
  flush();
}


template <class F> int accumulate_max_blocks_per_sm() {
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, msm_accumulate_kernel<F, true>, 128, 0) != cudaSuccess) return 1;
  return occ;
}
template <class F>
void accumulate_launch(int blocks, cudaStream_t st, const Affine<F>* bases, const uint32_t* entries, const uint32_t* starts, uint32_t nb,
                       XYZZ<F>* buckets, XYZZ<F>* part_pt, uint32_t* part_gb) {
  msm_accumulate_kernel<F, true><<<blocks, 128, 0, st>>>(bases, entries, starts, nb, buckets, part_pt, part_gb);
}

}  // namespace ctg
