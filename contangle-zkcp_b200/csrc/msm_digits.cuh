// Signed-digit window recoding of a canonical 256-bit scalar (BLS12-381 Fr, < 2^255).
// With c-bit windows and nwin = ceil(256 / c) windows the top window always has a spare
// bit, so every digit lies in [-2^(c-1), 2^(c-1)] and no carry leaves the top window:
//   k = sum_w digit_w * 2^(c*w).
// Bucket index = |digit| - 1 (2^(c-1) buckets per window, half of arkworks' unsigned
// 2^c - 1; ark-ec 0.3.0 msm/variable_base.rs is the function this path replaces).
#pragma once
#include "ptx.cuh"

namespace ctg {

CTG_HD int msm_num_windows(int c) { return (256 + c - 1) / c; }

// Must be called for w = 0, 1, ..., nwin-1 in order with carry starting at 0.
CTG_HD int32_t msm_signed_digit(const uint32_t* k, int c, int w, int nwin, uint32_t& carry) {
  (void)nwin;
  const int bit = w * c;
  const int limb = bit >> 5, off = bit & 31;
  uint32_t lo = limb < 8 ? k[limb] : 0u;
  uint32_t hi = limb + 1 < 8 ? k[limb + 1] : 0u;
  uint64_t two = ((uint64_t)hi << 32) | lo;
  uint32_t raw = (uint32_t)(two >> off) & ((1u << c) - 1u);
  uint32_t d = raw + carry;
  if (d > (1u << (c - 1))) {
    carry = 1;
    return (int32_t)d - (int32_t)(1u << c);
  }
  carry = 0;
  return (int32_t)d;
}

}  // namespace ctg
