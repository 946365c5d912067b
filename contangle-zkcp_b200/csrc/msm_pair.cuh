// Lane-paired group operations (G2, and at the end of the file G1) for the LATENCY-bound tail of the bucket reduction (msm.cuh section 6).
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

// ---- G1: the independent products of ONE group operation split over the lane pair -------------------------------------------
// Fq has no components to split, but an XYZZ addition is 14 products of which at most 2 depend on each other at a time.  Both
// lanes hold the whole point; every step is one Fq product per lane with lane-selected operands (the pair runs ONE instruction
// stream), results cross by shuffles: 7 product steps per addition instead of 14, 5 per doubling instead of 9.  Same formulas
// (add-2008-s, dbl-2008-s-1), every value fully reduced, hence the same XYZZ coordinates as xyzz_add / xyzz_dbl.
struct G1PairOps {
  typedef G1XYZZ Mem;
  typedef G1XYZZ Reg;
  __device__ __forceinline__ static Fq xchg(const Fq& mine) { return Fq2L{mine}.partner(); }
  __device__ __forceinline__ static Reg inf() { return G1XYZZ::inf(); }
  __device__ __forceinline__ static bool is_inf(const Reg& p) { return p.is_inf(); }   // both lanes hold the same point
  __device__ __forceinline__ static void ld(Reg& p, const Mem* src, unsigned) { copy16(p, *src); }
  __device__ __forceinline__ static void st(Mem* dst, const Reg& p, unsigned par) {
    Fq* f = reinterpret_cast<Fq*>(dst);
    if (par == 0) { copy16(f[0], p.X); copy16(f[1], p.Y); }
    else { copy16(f[2], p.ZZ); copy16(f[3], p.ZZZ); }
  }
  // each lane contributes two coordinates (s1, s2); the even lane owns (X, Y) when `xy_on_even`, else (ZZ, ZZZ)
  __device__ __forceinline__ static Reg gather(const Fq& s1, const Fq& s2, bool odd, bool xy_on_even) {
    const Fq r1 = xchg(s1), r2 = xchg(s2);
    const bool mine_is_xy = odd != xy_on_even;
    return Reg{Fq2L::sel(mine_is_xy, s1, r1), Fq2L::sel(mine_is_xy, s2, r2), Fq2L::sel(mine_is_xy, r1, s1), Fq2L::sel(mine_is_xy, r2, s2)};
  }
  //                                                                                  even lane           | odd lane
  __device__ __noinline__ static Reg dbl(const Reg& p) {
    if (p.is_inf()) return p;
    const bool odd = Fq2L::parity() != 0;
    const Fq U = p.Y.dbl();
    const Fq t1 = Fq2L::sel(odd, p.X, U).sqr();                                     // V = U^2            | xx = X^2
    const Fq V = Fq2L::sel(odd, xchg(t1), t1);
    const Fq M = t1.dbl() + t1;                                                     //                    | M = 3 xx
    const Fq t2 = Fq2L::sel(odd, p.X, U) * V;                                       // W = U V            | S = X V
    const Fq t3 = Fq2L::sel(odd, M, V) * Fq2L::sel(odd, M, p.ZZ);                   // ZZ3 = V ZZ         | M^2
    const Fq X3 = t3 - t2.dbl();                                                    //                    | X3 = M^2 - 2 S
    const Fq W = Fq2L::sel(odd, xchg(t2), t2);
    const Fq t4 = Fq2L::sel(odd, M, W) * Fq2L::sel(odd, t2 - X3, p.ZZZ);            // ZZZ3 = W ZZZ       | M (S - X3)
    const Fq t5 = W * p.Y;                                                          // (unused)           | W Y
    const Fq Y3 = t4 - t5;                                                          //                    | Y3
    return gather(Fq2L::sel(odd, X3, t3), Fq2L::sel(odd, Y3, t4), odd, false);
  }
  __device__ __noinline__ static Reg add(const Reg& a, const Reg& b) {
    if (a.is_inf()) return b;
    if (b.is_inf()) return a;
    const bool odd = Fq2L::parity() != 0;
    const Fq m1 = Fq2L::sel(odd, b.X, a.X) * Fq2L::sel(odd, a.ZZ, b.ZZ);            // U1 = aX bZZ        | U2 = bX aZZ
    const Fq m2 = Fq2L::sel(odd, b.Y, a.Y) * Fq2L::sel(odd, a.ZZZ, b.ZZZ);          // S1 = aY bZZZ       | S2 = bY aZZZ
    const Fq o1 = xchg(m1), o2 = xchg(m2);
    const Fq U1 = Fq2L::sel(odd, o1, m1), U2 = Fq2L::sel(odd, m1, o1), S1 = Fq2L::sel(odd, o2, m2), S2 = Fq2L::sel(odd, m2, o2);
    const Fq P = U2 - U1, R = S2 - S1;
    if (P.is_zero()) {
      if (R.is_zero()) return dbl(a);
      return G1XYZZ::inf();
    }
    const Fq t3 = Fq2L::sel(odd, R, P).sqr();                                       // PP = P^2           | RR = R^2
    const Fq o3 = xchg(t3);
    const Fq PP = Fq2L::sel(odd, o3, t3), RR = Fq2L::sel(odd, t3, o3);
    const Fq t4 = Fq2L::sel(odd, a.ZZ, P) * Fq2L::sel(odd, b.ZZ, PP);               // PPP = P PP         | ZZ12 = aZZ bZZ
    const Fq t5 = Fq2L::sel(odd, a.ZZZ, U1) * Fq2L::sel(odd, b.ZZZ, PP);            // Q = U1 PP          | ZZZ12 = aZZZ bZZZ
    const Fq PPP = Fq2L::sel(odd, xchg(t4), t4);
    const Fq X3 = RR - PPP - t5.dbl();                                              // X3                 | (unused)
    const Fq t6 = Fq2L::sel(odd, t4, R) * Fq2L::sel(odd, PP, t5 - X3);              // R (Q - X3)         | ZZ3 = ZZ12 PP
    const Fq t7 = Fq2L::sel(odd, t5, S1) * PPP;                                     // S1 PPP             | ZZZ3 = ZZZ12 PPP
    const Fq Y3 = t6 - t7;                                                          // Y3                 | (unused)
    return gather(Fq2L::sel(odd, t6, X3), Fq2L::sel(odd, t7, Y3), odd, true);
  }
};
struct G2PairOps {
  typedef G2XYZZ Mem;
  typedef G2XYZZL Reg;
  __device__ __forceinline__ static Reg inf() { return G2XYZZL::inf(); }
  __device__ __forceinline__ static bool is_inf(const Reg& p) { return p.is_inf(); }
  __device__ __forceinline__ static void ld(Reg& p, const Mem* src, unsigned par) { ld_half(p, src, par); }
  __device__ __forceinline__ static void st(Mem* dst, const Reg& p, unsigned par) { st_half(dst, p, par); }
  __device__ __forceinline__ static Reg add(const Reg& a, const Reg& b) { return xyzz_add(a, b); }
  __device__ __forceinline__ static Reg dbl(const Reg& a) { return xyzz_dbl(a); }
};

// msm_halve_kernel over lane pairs: thread t handles output t / 2
template <class Ops> __global__ void __launch_bounds__(128) msm_halve_pair_kernel(HalveArgs a) {
  typedef typename Ops::Mem Mem;
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
  const Mem* src = reinterpret_cast<const Mem*>(s.src) + (uint64_t)set * s.src_set_stride;
  Mem* dst = reinterpret_cast<Mem*>(s.dst) + (uint64_t)set * s.dst_set_stride;
  typename Ops::Reg x;
  Ops::ld(x, src + ((uint64_t)(2 * i) * s.step + s.off), par);
  if (!s.copy) {
    typename Ops::Reg y;
    Ops::ld(y, src + ((uint64_t)(2 * i + 1) * s.step + s.off), par);
    x = Ops::add(x, y);
  }
  Ops::st(dst + i, x, par);
}

// msm_reduce_final_kernel over lane pairs: 64 threads per bucket set, slot = thread / 2
template <class Ops> __global__ void __launch_bounds__(64) msm_reduce_final_pair_kernel(const typename Ops::Mem* finals, int nslots, int nplain, int shift, typename Ops::Mem* set_out) {
  __shared__ typename Ops::Mem sh[32];
  const unsigned slot = threadIdx.x >> 1, par = threadIdx.x & 1u;
  typename Ops::Reg acc = Ops::inf();
  if ((int)slot < nslots) {
    Ops::ld(acc, finals + ((uint64_t)blockIdx.x * nslots + slot), par);
    if ((int)slot >= nplain && !Ops::is_inf(acc)) {
      const int d = (int)slot - nplain + shift;
      for (int i = 0; i < d; i++) acc = Ops::dbl(acc);
    }
  }
  Ops::st(&sh[slot], acc, par);
  __syncthreads();
  for (unsigned off = 16; off >= 1; off >>= 1) {
    if (slot < off) {
      typename Ops::Reg o;
      Ops::ld(o, &sh[slot + off], par);
      acc = Ops::add(acc, o);
    }
    __syncthreads();
    if (slot < off) Ops::st(&sh[slot], acc, par);
    __syncthreads();
  }
  if (slot == 0) Ops::st(set_out + blockIdx.x, acc, par);
}
#endif

}  // namespace ctg
