// Lane-paired G2 group operations for the LATENCY-bound tail of the bucket reduction (msm.cuh section 6).
//
// A lone thread needs ~165 us for one G2 XYZZ addition: 14 Fq2 products, each two dependent lazy dot products over Fq, on one
// carry chain.  The halving levels and the closing doublings of the reduction are ~35 such operations in sequence whatever the
// MSM size (profiles/r02_g2_full_ncu.txt: every msm_halve_kernel<Fq2> launch takes 163 us).  Here an Fq2 element lives on an
// even / odd LANE PAIR -- parity 0 holds c0, parity 1 holds c1 -- so that the two dot products of an Fq2 product
//     c0 = a0 b0 - a1 b1        c1 = a0 b1 + a1 b0
// run side by side, each lane fetching its partner's operands with twelve 32-bit shuffles.  The generic xyzz_add / xyzz_dbl
// (curve.cuh) run unchanged over this type: every branch they take depends on is_zero(), which is agreed on by the pair.
// Results are the same field elements, hence bit-identical points.
#pragma once
#include "curve.cuh"

namespace ctg {

struct Fq2L {
  Fq v;   // this lane's component

#if defined(__CUDA_ARCH__)
  __device__ __forceinline__ static unsigned parity() { return threadIdx.x & 1u; }
  __device__ __forceinline__ static unsigned pair_mask() { return 3u << (threadIdx.x & 30u); }
  __device__ __forceinline__ Fq partner() const {
    Fq p;
    const unsigned m = pair_mask();
#pragma unroll
    for (int i = 0; i < Fq::N; i++) p.v[i] = __shfl_xor_sync(m, v.v[i], 1);
    return p;
  }
  __device__ __forceinline__ static Fq sel(bool c, const Fq& a, const Fq& b) {
    Fq r;
#pragma unroll
    for (int i = 0; i < Fq::N; i++) r.v[i] = c ? a.v[i] : b.v[i];
    return r;
  }
#else   // never executed on the host; the generic curve code is __host__ __device__ and must compile for both
  static unsigned parity() { return 0; }
  static unsigned pair_mask() { return 0; }
  Fq partner() const { return v; }
  static Fq sel(bool c, const Fq& a, const Fq& b) { return c ? a : b; }
#endif

  CTG_HD static Fq2L zero() { return {Fq::zero()}; }
  CTG_HD static Fq2L one() { return {parity() ? Fq::zero() : Fq::one()}; }
  CTG_HD bool is_zero() const {
    const bool z = v.is_zero();
#if defined(__CUDA_ARCH__)
    const int pz = __shfl_xor_sync(pair_mask(), (int)z, 1);   // unconditionally: both lanes must reach the shuffle
    return z && pz != 0;
#else
    return z;
#endif
  }
  CTG_HD friend Fq2L operator+(const Fq2L& a, const Fq2L& b) { return {a.v + b.v}; }
  CTG_HD friend Fq2L operator-(const Fq2L& a, const Fq2L& b) { return {a.v - b.v}; }
  CTG_HD Fq2L neg() const { return {v.neg()}; }
  CTG_HD Fq2L dbl() const { return {v.dbl()}; }
  // parity 0: a0 b0 + a1 (-b1);  parity 1 (own = a1, b1; partner = a0, b0): a1 b0 + a0 b1
  CTG_HD friend Fq2L operator*(const Fq2L& a, const Fq2L& b) {
    const bool odd = parity() != 0;
    const Fq pa = a.partner(), pb = b.partner();
    const Fq y0 = sel(odd, pb, b.v);
    const Fq y1 = sel(odd, b.v, pb.neg_lazy());
    return {Fq::mul2_add(a.v, y0, pa, y1)};
  }
  // parity 0: (c0 + c1)(c0 - c1);  parity 1: (2 c0) c1
  CTG_HD Fq2L sqr() const {
    const bool odd = parity() != 0;
    const Fq p = partner();
    const Fq x = sel(odd, p + p, v + p);
    const Fq y = sel(odd, v, v - p);
    return {x * y};
  }
  // a b - c d:  c0 = a0 b0 - a1 b1 - c0 d0 + c1 d1,  c1 = a0 b1 + a1 b0 - c0 d1 - c1 d0 -- one 4-term lazy dot product per lane
  CTG_HD static Fq2L mul2_sub(const Fq2L& a, const Fq2L& b, const Fq2L& c, const Fq2L& d) {
    const bool odd = parity() != 0;
    const Fq pa = a.partner(), pb = b.partner(), pc = c.partner(), pd = d.partner();
    const Fq nd = d.v.neg_lazy();
    const Fq y0 = sel(odd, pb, b.v);
    const Fq y1 = sel(odd, b.v, pb.neg_lazy());
    const Fq y2 = sel(odd, pd.neg_lazy(), nd);
    const Fq y3 = sel(odd, nd, pd);
    const Fq* x[4] = {&a.v, &pa, &c.v, &pc};
    const Fq* y[4] = {&y0, &y1, &y2, &y3};
    return {Fq::dot<4>(x, y)};
  }
};

typedef XYZZ<Fq2L> G2XYZZL;

#if defined(__CUDACC__)
// this lane's half of a G2 XYZZ point in memory (X.c0, X.c1, Y.c0, ... : component `par` of each coordinate)
__device__ __forceinline__ void ld_half(G2XYZZL& p, const G2XYZZ* src, unsigned par) {
  const Fq* f = reinterpret_cast<const Fq*>(src);
  copy16(p.X.v, f[0 + par]);
  copy16(p.Y.v, f[2 + par]);
  copy16(p.ZZ.v, f[4 + par]);
  copy16(p.ZZZ.v, f[6 + par]);
}
__device__ __forceinline__ void st_half(G2XYZZ* dst, const G2XYZZL& p, unsigned par) {
  Fq* f = reinterpret_cast<Fq*>(dst);
  copy16(f[0 + par], p.X.v);
  copy16(f[2 + par], p.Y.v);
  copy16(f[4 + par], p.ZZ.v);
  copy16(f[6 + par], p.ZZZ.v);
}

// msm_halve_kernel over lane pairs: thread t handles output t / 2
template <int TAG> __global__ void __launch_bounds__(128) msm_halve_pair_kernel(HalveArgs a) {   // a template only so that every translation unit may include this header
  const uint32_t gt = blockIdx.x * 128 + threadIdx.x;
  const uint32_t t = gt >> 1, par = gt & 1u;
  const uint32_t set = blockIdx.y;
  int k = -1;
#pragma unroll 1
  for (int i = 0; i < a.nseg; i++)
    if (t >= a.s[i].first_thread && t - a.s[i].first_thread < a.s[i].out_count) k = i;
  if (k < 0) return;   // both lanes of a pair leave together
  const HalveSeg& s = a.s[k];
  const uint32_t i = t - s.first_thread;
  const G2XYZZ* src = reinterpret_cast<const G2XYZZ*>(s.src) + (uint64_t)set * s.src_set_stride;
  G2XYZZ* dst = reinterpret_cast<G2XYZZ*>(s.dst) + (uint64_t)set * s.dst_set_stride;
  G2XYZZL x;
  ld_half(x, src + ((uint64_t)(2 * i) * s.step + s.off), par);
  if (!s.copy) {
    G2XYZZL y;
    ld_half(y, src + ((uint64_t)(2 * i + 1) * s.step + s.off), par);
    x = xyzz_add(x, y);
  }
  st_half(dst + i, x, par);
}

// msm_reduce_final_kernel over lane pairs: 64 threads per bucket set, slot = thread / 2
template <int TAG> __global__ void __launch_bounds__(64) msm_reduce_final_pair_kernel(const G2XYZZ* finals, int nslots, int nplain, int shift, G2XYZZ* set_out) {
  __shared__ G2XYZZ sh[32];
  const unsigned slot = threadIdx.x >> 1, par = threadIdx.x & 1u;
  G2XYZZL acc = G2XYZZL::inf();
  if ((int)slot < nslots) {
    ld_half(acc, finals + ((uint64_t)blockIdx.x * nslots + slot), par);
    if ((int)slot >= nplain && !acc.is_inf()) {
      const int d = (int)slot - nplain + shift;
      for (int i = 0; i < d; i++) acc = xyzz_dbl(acc);
    }
  }
  st_half(&sh[slot], acc, par);
  __syncthreads();
  for (unsigned off = 16; off >= 1; off >>= 1) {
    if (slot < off) {
      G2XYZZL o;
      ld_half(o, &sh[slot + off], par);
      acc = xyzz_add(acc, o);
    }
    __syncthreads();
    if (slot < off) st_half(&sh[slot], acc, par);
    __syncthreads();
  }
  if (slot == 0) st_half(set_out + blockIdx.x, acc, par);
}
#endif

}  // namespace ctg
