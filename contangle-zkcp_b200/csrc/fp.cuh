// Montgomery prime-field arithmetic on 32-bit limbs for BLS12-381 Fr (8 limbs, R = 2^256)
// and Fq (12 limbs, R = 2^384) -- the same R arkworks uses with 64-bit limbs, so a host
// Fp256/Fp384 memory image is bit-identical to the device limb array.
// Replaces: ark-ff 0.3.0 fields/models/{fp_256,fp_384}.rs (mul_assign / add_assign /
// sub_assign / into_repr), reached from the reference at lib/src/zk/encryption.rs:76.
//
// mont_mul keeps the running sum T as two interleaved carry chains:
//   E holds limbs 0..N-1, O holds limbs 1..N, T = E + 2^32 * O.
// A row adds a*b_i: the products a_j*b_i with even j tile E without overlap, those with
// odd j tile O, so each is one clean lo/hi carry chain (IMAD.WIDE.X after ptxas pairing).
// The Montgomery step adds m*p the same way, leaving E[0] = 0; dividing by 2^32 then
// swaps the roles (O becomes the limb-0 chain) and the old E, now two limbs out of phase,
// is re-aligned for free inside the next row's multiply-add (register renaming once
// unrolled).  All values stay < 2^(32(N+1)) because T < 2p before each row, so neither
// chain can carry out of its top limb.
#pragma once
#include "ptx.cuh"

namespace ctg {

template <int N> struct Limbs { uint32_t v[N]; };

struct FrParams {
  static constexpr int N = 8;
  static constexpr uint32_t INV = 0xffffffffu;  // -r^-1 mod 2^32
  CTG_HD static constexpr uint32_t mod(int i) {
    constexpr uint32_t m[8] = {0x00000001u, 0xffffffffu, 0xfffe5bfeu, 0x53bda402u, 0x09a1d805u, 0x3339d808u, 0x299d7d48u, 0x73eda753u};
    return m[i];
  }
  CTG_HD static constexpr uint32_t one(int i) {  // R mod r
    constexpr uint32_t m[8] = {0xfffffffeu, 0x00000001u, 0x00034802u, 0x5884b7fau, 0xecbc4ff5u, 0x998c4fefu, 0xacc5056fu, 0x1824b159u};
    return m[i];
  }
  CTG_HD static constexpr uint32_t r2(int i) {  // R^2 mod r
    constexpr uint32_t m[8] = {0xf3f29c6du, 0xc999e990u, 0x87925c23u, 0x2b6cedcbu, 0x7254398fu, 0x05d31496u, 0x9f59ff11u, 0x0748d9d9u};
    return m[i];
  }
  CTG_HD static constexpr uint32_t inv_c(int i) {  // 2^(2 * 18) * R^3 mod r: closing constant of inv_bingcd (18 outer iterations)
    constexpr uint32_t m[8] = {0xcb9d63b7u, 0x6e19d729u, 0x5f8be0ccu, 0xf8f3f232u, 0xa5edc8e5u, 0xbab78932u, 0xc66409cfu, 0x4a73978cu};
    return m[i];
  }
};

#if defined(__CUDACC__)
// -r^-1 mod 2^32 of Fr (= 2^32 - 1) as a constant-bank word.  Fp<FrParams>::redc_row forms m = E[0] * this word with a real IMAD on
// the device: when m is produced by the ALU instead (the strength-reduced negation), ptxas 12.9 splits EVERY mad.lo.cc / madc.hi.cc
// pair of the row into IMAD.X + IMAD.HI.U32.X, and an IMAD.HI costs the FMA pipe as much as the paired IMAD.WIDE (two issue slots,
// profiles/r01_pipe_probe.txt): 304 slots per Montgomery product instead of 233.
static __constant__ uint32_t kFrNegInvBank = 0xffffffffu;
#endif
struct FqParams {
  static constexpr int N = 12;
  static constexpr uint32_t INV = 0xfffcfffdu;  // -p^-1 mod 2^32
  CTG_HD static constexpr uint32_t mod(int i) {
    constexpr uint32_t m[12] = {0xffffaaabu, 0xb9feffffu, 0xb153ffffu, 0x1eabfffeu, 0xf6b0f624u, 0x6730d2a0u,
                                0xf38512bfu, 0x64774b84u, 0x434bacd7u, 0x4b1ba7b6u, 0x397fe69au, 0x1a0111eau};
    return m[i];
  }
  CTG_HD static constexpr uint32_t one(int i) {  // R mod p
    constexpr uint32_t m[12] = {0x0002fffdu, 0x76090000u, 0xc40c0002u, 0xebf4000bu, 0x53c758bau, 0x5f489857u,
                                0x70525745u, 0x77ce5853u, 0xa256ec6du, 0x5c071a97u, 0xfa80e493u, 0x15f65ec3u};
    return m[i];
  }
  CTG_HD static constexpr uint32_t r2(int i) {  // R^2 mod p
    constexpr uint32_t m[12] = {0x1c341746u, 0xf4df1f34u, 0x09d104f1u, 0x0a76e6a6u, 0x4c95b6d5u, 0x8de5476cu,
                                0x939d83c0u, 0x67eb88a9u, 0xb519952du, 0x9a793e85u, 0x92cae3aau, 0x11988fe5u};
    return m[i];
  }
  CTG_HD static constexpr uint32_t inv_c(int i) {  // 2^(2 * 26) * R^3 mod p: closing constant of inv_bingcd (26 outer iterations)
    constexpr uint32_t m[12] = {0x8c5d1a2eu, 0x42957fa8u, 0x24cc85b1u, 0xa6507a62u, 0x39990a58u, 0xd8ca1b85u,
                                0xac2f63fcu, 0xe65a054fu, 0xe1f35417u, 0x4ca0897au, 0xa0365c54u, 0x032748cbu};
    return m[i];
  }
};

template <class P> struct alignas(16) Fp {
  static constexpr int N = P::N;
  uint32_t v[N];

  CTG_HD static Fp zero() { Fp r;
#pragma unroll
    for (int i = 0; i < N; i++) r.v[i] = 0;
    return r; }
  CTG_HD static Fp one() { Fp r;
#pragma unroll
    for (int i = 0; i < N; i++) r.v[i] = P::one(i);
    return r; }
  CTG_HD bool is_zero() const { uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < N; i++) o |= v[i];
    return o == 0; }
  CTG_HD bool eq(const Fp& b) const { uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < N; i++) o |= v[i] ^ b.v[i];
    return o == 0; }

  // r = a - p if a >= p else a   (a < 2p)
  CTG_HD static void reduce_once(uint32_t* a) {
    uint32_t t[N];
    t[0] = ptx::sub_cc(a[0], P::mod(0));
#pragma unroll
    for (int i = 1; i < N; i++) t[i] = ptx::subc_cc(a[i], P::mod(i));
    uint32_t borrow = ptx::subc(0, 0);  // 0xffffffff when a < p
#pragma unroll
    for (int i = 0; i < N; i++) a[i] = borrow ? a[i] : t[i];
  }

  CTG_HD friend Fp operator+(const Fp& a, const Fp& b) {
    Fp r;
    r.v[0] = ptx::add_cc(a.v[0], b.v[0]);
#pragma unroll
    for (int i = 1; i < N - 1; i++) r.v[i] = ptx::addc_cc(a.v[i], b.v[i]);
    r.v[N - 1] = ptx::addc(a.v[N - 1], b.v[N - 1]);  // 2p < 2^(32N): no carry out
    reduce_once(r.v);
    return r;
  }
  CTG_HD friend Fp operator-(const Fp& a, const Fp& b) {
    Fp r;
    r.v[0] = ptx::sub_cc(a.v[0], b.v[0]);
#pragma unroll
    for (int i = 1; i < N; i++) r.v[i] = ptx::subc_cc(a.v[i], b.v[i]);
    uint32_t mask = ptx::subc(0, 0);  // all ones on borrow
    r.v[0] = ptx::add_cc(r.v[0], P::mod(0) & mask);
#pragma unroll
    for (int i = 1; i < N - 1; i++) r.v[i] = ptx::addc_cc(r.v[i], P::mod(i) & mask);
    r.v[N - 1] = ptx::addc(r.v[N - 1], P::mod(N - 1) & mask);
    return r;
  }
  CTG_HD Fp neg() const { return zero() - *this; }
  CTG_HD Fp dbl() const { return *this + *this; }

  // --- Montgomery product ------------------------------------------------------------
  // add m*p into (E at limbs 0.., O at limbs 1..); leaves E[0] == 0
  CTG_HD static void redc_row(uint32_t* E, uint32_t* O) {
    if constexpr (P::mod(0) == 1u && P::mod(1) == 0xffffffffu && P::INV == 0xffffffffu && N % 2 == 0) {
      // Fr: the modulus ends in 2^64 - 2^32 + 1 and -r^-1 = 2^32 - 1.  Limb 0 of m * r is m itself (one add that clears E[0] and
      // carries into limb 1); limb 1 is multiplied by the all-ones word FROM THE CONSTANT BANK so that it stays a paired product.
      // 56 IMAD.WIDE + 8 IMAD per Montgomery product for this half and three carry-chain adds per row.  (Spelling the two low limbs
      // out with adds alone saves 8 IMAD.WIDE but puts five dependent ALU instructions in front of every row: slower, measured.)
#if defined(__CUDA_ARCH__)
      const uint32_t allones = kFrNegInvBank;                // see kFrNegInvBank: keeps the pairs below as IMAD.WIDE
#else
      const uint32_t allones = 0xffffffffu;
#endif
      const uint32_t m = ptx::mul_lo(E[0], allones);
      O[0] = ptx::mad_lo_cc(allones, m, O[0]);
      O[1] = ptx::madc_hi_cc(allones, m, O[1]);
#pragma unroll
      for (int j = 3; j < N; j += 2) {
        O[j - 1] = ptx::madc_lo_cc(P::mod(j), m, O[j - 1]);
        if (j == N - 1) O[j] = ptx::madc_hi(P::mod(j), m, O[j]);
        else O[j] = ptx::madc_hi_cc(P::mod(j), m, O[j]);
      }
      E[0] = ptx::add_cc(E[0], m);
      E[1] = ptx::addc_cc(E[1], 0u);
      E[2] = ptx::madc_lo_cc(P::mod(2), m, E[2]);
      E[3] = ptx::madc_hi_cc(P::mod(2), m, E[3]);
#pragma unroll
      for (int j = 4; j < N; j += 2) {
        E[j] = ptx::madc_lo_cc(P::mod(j), m, E[j]);
        E[j + 1] = ptx::madc_hi_cc(P::mod(j), m, E[j + 1]);
      }
      O[N - 1] = ptx::addc(O[N - 1], 0);
      return;
    }
    const uint32_t m = ptx::mul_lo(E[0], P::INV);
    O[0] = ptx::mad_lo_cc(P::mod(1), m, O[0]);
    O[1] = ptx::madc_hi_cc(P::mod(1), m, O[1]);
#pragma unroll
    for (int j = 3; j < N; j += 2) {
      O[j - 1] = ptx::madc_lo_cc(P::mod(j), m, O[j - 1]);
      if (j == N - 1) O[j] = ptx::madc_hi(P::mod(j), m, O[j]);
      else O[j] = ptx::madc_hi_cc(P::mod(j), m, O[j]);
    }
    E[0] = ptx::mad_lo_cc(P::mod(0), m, E[0]);
    E[1] = ptx::madc_hi_cc(P::mod(0), m, E[1]);
#pragma unroll
    for (int j = 2; j < N; j += 2) {
      E[j] = ptx::madc_lo_cc(P::mod(j), m, E[j]);
      E[j + 1] = ptx::madc_hi_cc(P::mod(j), m, E[j + 1]);
    }
    O[N - 1] = ptx::addc(O[N - 1], 0);
  }
  // E: limbs 0..N-1 (aligned); Oold: previous limb-0 chain, entry k sits at limb k-1.
  CTG_HD static void mad_row(uint32_t* E, uint32_t* O, const uint32_t* a, uint32_t bi) {
    E[0] = ptx::add_cc(E[0], O[1]);  // orphan limb; its carry enters limb 1 = head of the O chain
#pragma unroll
    for (int j = 1; j < N - 1; j += 2) {
      O[j - 1] = ptx::madc_lo_cc(a[j], bi, O[j + 1]);
      O[j] = ptx::madc_hi_cc(a[j], bi, O[j + 2]);
    }
    O[N - 2] = ptx::madc_lo_cc(a[N - 1], bi, 0);
    O[N - 1] = ptx::madc_hi(a[N - 1], bi, 0);
    E[0] = ptx::mad_lo_cc(a[0], bi, E[0]);
    E[1] = ptx::madc_hi_cc(a[0], bi, E[1]);
#pragma unroll
    for (int j = 2; j < N; j += 2) {
      E[j] = ptx::madc_lo_cc(a[j], bi, E[j]);
      E[j + 1] = ptx::madc_hi_cc(a[j], bi, E[j + 1]);
    }
    O[N - 1] = ptx::addc(O[N - 1], 0);
  }

  // E/O += x * y without re-alignment (same limb tiling as redc_row); used by sum-of-products rows
  CTG_HD static void acc_row(uint32_t* E, uint32_t* O, const uint32_t* x, uint32_t y) {
    O[0] = ptx::mad_lo_cc(x[1], y, O[0]);
    O[1] = ptx::madc_hi_cc(x[1], y, O[1]);
#pragma unroll
    for (int j = 3; j < N; j += 2) {
      O[j - 1] = ptx::madc_lo_cc(x[j], y, O[j - 1]);
      if (j == N - 1) O[j] = ptx::madc_hi(x[j], y, O[j]);
      else O[j] = ptx::madc_hi_cc(x[j], y, O[j]);
    }
    E[0] = ptx::mad_lo_cc(x[0], y, E[0]);
    E[1] = ptx::madc_hi_cc(x[0], y, E[1]);
#pragma unroll
    for (int j = 2; j < N; j += 2) {
      E[j] = ptx::madc_lo_cc(x[j], y, E[j]);
      E[j + 1] = ptx::madc_hi_cc(x[j], y, E[j + 1]);
    }
    O[N - 1] = ptx::addc(O[N - 1], 0);
  }

  // (sum_k a[k] * b[k]) / R mod p with ONE Montgomery reduction ("lazy reduction"): every row adds the K partial
  // products before the reduction step.  Needs (K + 1) * p < 2^(32 N) so that the running sum fits its limbs and
  // p * (1 + K p / R) < 2p for the single final subtraction: true for Fq (p / R ~ 0.10) up to K = 8, not for Fr.
  template <int K> CTG_HD static Fp dot(const Fp* const* a, const Fp* const* b) {
    static_assert(N == 12 && K >= 1 && K <= 8, "lazy sums of products are only safe for the 381-bit field");
    uint32_t ev[N], od[N];
#pragma unroll
    for (int j = 0; j < N; j += 2) {
      ev[j] = ptx::mul_lo(a[0]->v[j], b[0]->v[0]);
      ev[j + 1] = ptx::mul_hi(a[0]->v[j], b[0]->v[0]);
      od[j] = ptx::mul_lo(a[0]->v[j + 1], b[0]->v[0]);
      od[j + 1] = ptx::mul_hi(a[0]->v[j + 1], b[0]->v[0]);
    }
#pragma unroll
    for (int k = 1; k < K; k++) acc_row(ev, od, a[k]->v, b[k]->v[0]);
    redc_row(ev, od);
#pragma unroll
    for (int i = 1; i < N; i += 2) {
      mad_row(od, ev, a[0]->v, b[0]->v[i]);
#pragma unroll
      for (int k = 1; k < K; k++) acc_row(od, ev, a[k]->v, b[k]->v[i]);
      redc_row(od, ev);
      if (i + 1 < N) {
        mad_row(ev, od, a[0]->v, b[0]->v[i + 1]);
#pragma unroll
        for (int k = 1; k < K; k++) acc_row(ev, od, a[k]->v, b[k]->v[i + 1]);
        redc_row(ev, od);
      }
    }
    Fp r;
    r.v[0] = ptx::add_cc(ev[0], od[1]);
#pragma unroll
    for (int k = 1; k < N - 1; k++) r.v[k] = ptx::addc_cc(ev[k], od[k + 1]);
    r.v[N - 1] = ptx::addc(ev[N - 1], 0);
    reduce_once(r.v);
    return r;
  }
  // a*b + c*d and a*b - c*d with one reduction (p - d stands in for -d; d = 0 gives p, still within the bounds)
  CTG_HD static Fp mul2_add(const Fp& a, const Fp& b, const Fp& c, const Fp& d) {
    const Fp* x[2] = {&a, &c};
    const Fp* y[2] = {&b, &d};
    return dot<2>(x, y);
  }
  CTG_HD Fp neg_lazy() const {  // p - x in [1, p]
    Fp r;
    r.v[0] = ptx::sub_cc(P::mod(0), v[0]);
#pragma unroll
    for (int i = 1; i < N - 1; i++) r.v[i] = ptx::subc_cc(P::mod(i), v[i]);
    r.v[N - 1] = ptx::subc(P::mod(N - 1), v[N - 1]);
    return r;
  }
  CTG_HD static Fp mul2_sub(const Fp& a, const Fp& b, const Fp& c, const Fp& d) {
    const Fp nd = d.neg_lazy();
    return mul2_add(a, b, c, nd);
  }

  CTG_HD friend Fp operator*(const Fp& a, const Fp& b) {
    uint32_t ev[N], od[N];
#pragma unroll
    for (int j = 0; j < N; j += 2) {
      ev[j] = ptx::mul_lo(a.v[j], b.v[0]);
      ev[j + 1] = ptx::mul_hi(a.v[j], b.v[0]);
      od[j] = ptx::mul_lo(a.v[j + 1], b.v[0]);
      od[j + 1] = ptx::mul_hi(a.v[j + 1], b.v[0]);
    }
    redc_row(ev, od);
#pragma unroll
    for (int i = 1; i < N; i += 2) {
      mad_row(od, ev, a.v, b.v[i]);
      redc_row(od, ev);
      if (i + 1 < N) {
        mad_row(ev, od, a.v, b.v[i + 1]);
        redc_row(ev, od);
      }
    }
    // N even: the last row ran with E = od, O = ev.  Result limb k = ev[k] + od[k+1].
    Fp r;
    r.v[0] = ptx::add_cc(ev[0], od[1]);
#pragma unroll
    for (int k = 1; k < N - 1; k++) r.v[k] = ptx::addc_cc(ev[k], od[k + 1]);
    r.v[N - 1] = ptx::addc(ev[N - 1], 0);
    reduce_once(r.v);
    return r;
  }
  // Row of the dedicated squaring: mad_row with the multiplicand vector v and the products of positions j < FIRST left out
  // (their entries of v are zero); the odd chain still re-aligns O, the even chain simply starts later.
  template <int FIRST> CTG_HD static void sqr_row(uint32_t* E, uint32_t* O, const uint32_t* v, uint32_t bi) {
    E[0] = ptx::add_cc(E[0], O[1]);
#pragma unroll
    for (int j = 1; j < N - 1; j += 2) {
      if (j < FIRST) {
        O[j - 1] = ptx::addc_cc(O[j + 1], 0);
        O[j] = ptx::addc_cc(O[j + 2], 0);
      } else {
        O[j - 1] = ptx::madc_lo_cc(v[j], bi, O[j + 1]);
        O[j] = ptx::madc_hi_cc(v[j], bi, O[j + 2]);
      }
    }
    O[N - 2] = ptx::madc_lo_cc(v[N - 1], bi, 0);      // FIRST <= N - 1: the top product is always there
    O[N - 1] = ptx::madc_hi(v[N - 1], bi, 0);
    constexpr int E0 = (FIRST + 1) & ~1;              // first even position with a product
    if (E0 < N) {
      E[E0] = ptx::mad_lo_cc(v[E0], bi, E[E0]);
      E[E0 + 1] = ptx::madc_hi_cc(v[E0], bi, E[E0 + 1]);
#pragma unroll
      for (int j = E0 + 2; j < N; j += 2) {
        E[j] = ptx::madc_lo_cc(v[j], bi, E[j]);
        E[j + 1] = ptx::madc_hi_cc(v[j], bi, E[j + 1]);
      }
      O[N - 1] = ptx::addc(O[N - 1], 0);
    }
  }
  template <int I> CTG_HD static void sqr_rows(uint32_t* ev, uint32_t* od, const uint32_t* a, const uint32_t* d) {
    if constexpr (I < N) {
      // row I multiplies a[I] by (a[I] at position I, 2 * (a >> 32 (I + 1)) above it): v[I] = a[I], v[I + 1] = a[I + 1] << 1,
      // v[j] = limb j of 2a for j > I + 1; positions below I hold nothing
      uint32_t v[N];
#pragma unroll
      for (int j = 0; j < N; j++) v[j] = j < I ? 0u : (j == I ? a[I] : (j == I + 1 ? a[j] << 1 : d[j]));
      if (I & 1) { sqr_row<I>(od, ev, v, a[I]); redc_row(od, ev); }
      else { sqr_row<I>(ev, od, v, a[I]); redc_row(ev, od); }
      sqr_rows<I + 1>(ev, od, a, d);
    }
  }
  // a^2 R^-1: the Montgomery product with the symmetric half of the partial products left out.  a^2 = sum_i a_i 2^(32 i) *
  // (a_i 2^(32 i) + 2 * sum_{j > i} a_j 2^(32 j)), so row i of the interleaved product multiplies a_i by the limbs of 2a above
  // position i (and by a_i itself at position i) and skips positions below i: N (N - 1) / 2 of the N^2 products become carry
  // adds on the ALU pipe (Fq: 222 instead of 288 IMAD.WIDE).  Needs the top bit of the top limb clear (2a fits N limbs).
  // The early rows carry the doubled products, so the running sum reaches 3p between rows: only for moduli with 3p < 2^(32 N)
  // (Fq: p ~ 0.10 * 2^384).  Fr (r ~ 0.45 * 2^256) keeps the plain product.
  CTG_HD Fp sqr() const {
    if constexpr (!(P::mod(N - 1) < 0x55555555u && N % 2 == 0)) return *this * *this;
    uint32_t d[N], ev[N], od[N];
    d[0] = v[0] << 1;
#pragma unroll
    for (int j = 1; j < N; j++) d[j] = (v[j] << 1) | (v[j - 1] >> 31);
    // row 0: a[0] * (a[0], a[1] << 1, d[2], ...)
    uint32_t w[N];
#pragma unroll
    for (int j = 0; j < N; j++) w[j] = j == 0 ? v[0] : (j == 1 ? v[1] << 1 : d[j]);
#pragma unroll
    for (int j = 0; j < N; j += 2) {
      ev[j] = ptx::mul_lo(w[j], v[0]);
      ev[j + 1] = ptx::mul_hi(w[j], v[0]);
      od[j] = ptx::mul_lo(w[j + 1], v[0]);
      od[j + 1] = ptx::mul_hi(w[j + 1], v[0]);
    }
    redc_row(ev, od);
    sqr_rows<1>(ev, od, v, d);
    Fp r;
    r.v[0] = ptx::add_cc(ev[0], od[1]);
#pragma unroll
    for (int k = 1; k < N - 1; k++) r.v[k] = ptx::addc_cc(ev[k], od[k + 1]);
    r.v[N - 1] = ptx::addc(ev[N - 1], 0);
    reduce_once(r.v);
    return r;
  }

  CTG_HD Fp from_mont() const {  // canonical integer image (ark-ff into_repr)
    Fp o = zero();
    o.v[0] = 1;
    return *this * o;
  }
  CTG_HD Fp to_mont() const {
    Fp r2;
#pragma unroll
    for (int i = 0; i < N; i++) r2.v[i] = P::r2(i);
    return *this * r2;
  }
  // a^(p-2); cold path (final to-affine only)
  CTG_HD_COLD Fp inv() const {
    uint32_t e[N];
    e[0] = ptx::sub_cc(P::mod(0), 2);
#pragma unroll
    for (int i = 1; i < N; i++) e[i] = ptx::subc_cc(P::mod(i), 0);
    Fp acc = one();
    for (int i = N * 32 - 1; i >= 0; i--) {
      acc = acc.sqr();
      if ((e[i >> 5] >> (i & 31)) & 1) acc = acc * *this;
    }
    return acc;
  }

  // a^-1 by binary GCD with 30-step inner loops on 62-bit approximations (T. Pornin, "Optimized Binary GCD for Modular
  // Inversion", ePrint 2020/972, Algorithm 2, restated for 32-bit limbs).  About 12x cheaper than the Fermat ladder above and
  // almost entirely shifts / adds, i.e. off the IMAD pipe; branch-free, so the lanes of a warp stay in lockstep.  Used by the
  // batched-affine bucket accumulation (one inversion per batch of point additions).  inv_bingcd(0) = 0.
  //   invariants (y = this image, m = modulus):  a * 2^(30 i) == u_nodiv * y,  b * 2^(30 i) == v_nodiv * y   (mod m)
  //   u, v are divided by 2^32 per outer iteration (one Montgomery row), so after I iterations y^-1 = v * 2^(2 I); the closing
  //   product with P::inv_c = 2^(2 I) R^3 also turns (x R)^-1 into the Montgomery image x^-1 R.
  CTG_HD_COLD Fp inv_bingcd() const {
    constexpr int J = 30;                                  // inner steps per outer iteration
    constexpr int ITER = (2 * 32 * N - 1 + J - 1) / J;     // 2 len(m) - 1 steps suffice; 32 N >= len(m)
    uint32_t a[N], b[N], u[N], vv[N];
#pragma unroll
    for (int i = 0; i < N; i++) { a[i] = v[i]; b[i] = P::mod(i); u[i] = 0; vv[i] = 0; }
    u[0] = 1;
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
      // ---- 62-bit approximations: low 30 bits + the 32 bits below the top of max(len a, len b, 62)
      uint32_t top = 0, tl = 0;
#pragma unroll
      for (int i = 0; i < N; i++) {
        const uint32_t w = a[i] | b[i];
        const bool nz = w != 0;
        top = nz ? w : top;
        tl = nz ? (uint32_t)i : tl;
      }
#if defined(__CUDA_ARCH__)
      uint32_t n = 32u * tl + 32u - (uint32_t)__clz((int)top);
#else
      uint32_t n = top ? 32u * tl + 32u - (uint32_t)__builtin_clz(top) : 0u;
#endif
      if (n < 62u) n = 62u;
      const uint32_t p = n - 32u, w0 = p >> 5, r = p & 31u;
      uint32_t a0 = 0, a1 = 0, b0 = 0, b1 = 0;
#pragma unroll
      for (int i = 0; i < N; i++) {
        const uint32_t m0 = (uint32_t)i == w0 ? 0xffffffffu : 0u;
        const uint32_t m1 = (uint32_t)i == w0 + 1 ? 0xffffffffu : 0u;
        a0 |= a[i] & m0; b0 |= b[i] & m0;
        a1 |= a[i] & m1; b1 |= b[i] & m1;
      }
      const uint32_t ha = (uint32_t)(((((uint64_t)a1) << 32) | a0) >> r);
      const uint32_t hb = (uint32_t)(((((uint64_t)b1) << 32) | b0) >> r);
      uint64_t xa = (uint64_t)(a[0] & 0x3fffffffu) | ((uint64_t)ha << 30);
      uint64_t xb = (uint64_t)(b[0] & 0x3fffffffu) | ((uint64_t)hb << 30);
      // ---- 30 binary-GCD steps on the approximations, recording the transition matrix
      uint32_t f0 = 1, g0 = 0, f1 = 0, g1 = 1;   // two's complement int32
#pragma unroll 2
      for (int j = 0; j < J; j++) {
        const uint64_t odd = 0ull - (xa & 1ull);
        const uint64_t sw = odd & (xa < xb ? ~0ull : 0ull);
        const uint64_t t = (xa ^ xb) & sw;
        xa ^= t; xb ^= t;
        const uint32_t sw32 = (uint32_t)sw, odd32 = (uint32_t)odd;
        const uint32_t tf = (f0 ^ f1) & sw32, tg = (g0 ^ g1) & sw32;
        f0 ^= tf; f1 ^= tf; g0 ^= tg; g1 ^= tg;
        xa -= xb & odd;
        f0 -= f1 & odd32; g0 -= g1 & odd32;
        xa >>= 1;
        f1 <<= 1; g1 <<= 1;
      }
      // ---- (a, b) <- (f0 a + g0 b, f1 a + g1 b) / 2^30, made non-negative
      int64_t F0 = (int32_t)f0, G0 = (int32_t)g0, F1 = (int32_t)f1, G1 = (int32_t)g1;
      {
        int64_t ca = 0, cb = 0;
        uint32_t na[N + 1], nb[N + 1];
#pragma unroll
        for (int i = 0; i < N; i++) {
          const int64_t ai = (int64_t)(uint64_t)a[i], bi = (int64_t)(uint64_t)b[i];
          ca += F0 * ai + G0 * bi;
          cb += F1 * ai + G1 * bi;
          na[i] = (uint32_t)ca; nb[i] = (uint32_t)cb;
          ca >>= 32; cb >>= 32;
        }
        na[N] = (uint32_t)ca; nb[N] = (uint32_t)cb;
        const uint32_t sa = ca < 0 ? 0xffffffffu : 0u, sb = cb < 0 ? 0xffffffffu : 0u;
        // shift right by 30, then conditional two's-complement negation: x -> (x ^ s) - s
        uint32_t bora = 0, borb = 0;   // running "+1" carries of the negation
        bora = sa & 1u; borb = sb & 1u;
#pragma unroll
        for (int i = 0; i < N; i++) {
          const uint32_t xa32 = ((na[i] >> 30) | (na[i + 1] << 2)) ^ sa;
          const uint32_t xb32 = ((nb[i] >> 30) | (nb[i + 1] << 2)) ^ sb;
          const uint64_t ta = (uint64_t)xa32 + bora, tb = (uint64_t)xb32 + borb;
          a[i] = (uint32_t)ta; b[i] = (uint32_t)tb;
          bora = (uint32_t)(ta >> 32); borb = (uint32_t)(tb >> 32);
        }
        if (sa) { F0 = -F0; G0 = -G0; }
        if (sb) { F1 = -F1; G1 = -G1; }
      }
      // ---- (u, v) <- (F0 u + G0 v, F1 u + G1 v) / 2^32 mod m, factors made non-negative by negating the operand
      {
        uint32_t nu[N], nv[N];
        {
          uint32_t bw = 0;
          uint32_t zu = 0, zv = 0;
#pragma unroll
          for (int i = 0; i < N; i++) { zu |= u[i]; zv |= vv[i]; }
          uint64_t t;
          bw = 0;
#pragma unroll
          for (int i = 0; i < N; i++) { t = (uint64_t)P::mod(i) - u[i] - bw; nu[i] = zu ? (uint32_t)t : 0u; bw = (uint32_t)(t >> 63); }
          bw = 0;
#pragma unroll
          for (int i = 0; i < N; i++) { t = (uint64_t)P::mod(i) - vv[i] - bw; nv[i] = zv ? (uint32_t)t : 0u; bw = (uint32_t)(t >> 63); }
        }
        const bool nF0 = F0 < 0, nG0 = G0 < 0, nF1 = F1 < 0, nG1 = G1 < 0;
        const uint64_t f0u = (uint64_t)(nF0 ? -F0 : F0), g0u = (uint64_t)(nG0 ? -G0 : G0);
        const uint64_t f1u = (uint64_t)(nF1 ? -F1 : F1), g1u = (uint64_t)(nG1 ? -G1 : G1);
        uint32_t tu[N + 1], tv[N + 1];
        uint64_t cu = 0, cv = 0;
#pragma unroll
        for (int i = 0; i < N; i++) {
          const uint64_t ua = nF0 ? nu[i] : u[i], va = nG0 ? nv[i] : vv[i];
          const uint64_t ub = nF1 ? nu[i] : u[i], vb = nG1 ? nv[i] : vv[i];
          cu += f0u * ua + g0u * va;     // < 2^30 * 2^32 + carry: no overflow
          cv += f1u * ub + g1u * vb;
          tu[i] = (uint32_t)cu; tv[i] = (uint32_t)cv;
          cu >>= 32; cv >>= 32;
        }
        tu[N] = (uint32_t)cu; tv[N] = (uint32_t)cv;
        // one Montgomery row each: (t + q m) / 2^32 < 2 m, then a conditional subtraction
        const uint32_t qu = tu[0] * P::INV, qv = tv[0] * P::INV;
        cu = ((uint64_t)tu[0] + (uint64_t)qu * P::mod(0)) >> 32;
        cv = ((uint64_t)tv[0] + (uint64_t)qv * P::mod(0)) >> 32;
#pragma unroll
        for (int i = 1; i < N; i++) {
          cu += (uint64_t)tu[i] + (uint64_t)qu * P::mod(i);
          cv += (uint64_t)tv[i] + (uint64_t)qv * P::mod(i);
          u[i - 1] = (uint32_t)cu; vv[i - 1] = (uint32_t)cv;
          cu >>= 32; cv >>= 32;
        }
        u[N - 1] = (uint32_t)(cu + tu[N]);
        vv[N - 1] = (uint32_t)(cv + tv[N]);
        reduce_once(u);
        reduce_once(vv);
      }
    }
    Fp res, c;
#pragma unroll
    for (int i = 0; i < N; i++) { res.v[i] = vv[i]; c.v[i] = P::inv_c(i); }
    return res * c;
  }
};

typedef Fp<FrParams> Fr;
typedef Fp<FqParams> Fq;

// Fq2 = Fq[u]/(u^2 + 1)   (ark-ff fp2.rs with ark-bls12-381 NONRESIDUE = -1)
struct Fq2 {
  Fq c0, c1;
  CTG_HD static Fq2 zero() { return {Fq::zero(), Fq::zero()}; }
  CTG_HD static Fq2 one() { return {Fq::one(), Fq::zero()}; }
  CTG_HD bool is_zero() const { return c0.is_zero() && c1.is_zero(); }
  CTG_HD bool eq(const Fq2& b) const { return c0.eq(b.c0) && c1.eq(b.c1); }
  CTG_HD friend Fq2 operator+(const Fq2& a, const Fq2& b) { return {a.c0 + b.c0, a.c1 + b.c1}; }
  CTG_HD friend Fq2 operator-(const Fq2& a, const Fq2& b) { return {a.c0 - b.c0, a.c1 - b.c1}; }
  CTG_HD Fq2 neg() const { return {c0.neg(), c1.neg()}; }
  CTG_HD Fq2 dbl() const { return {c0.dbl(), c1.dbl()}; }
  // schoolbook with lazy reduction: 4 products, 2 reductions, one negation
  CTG_HD friend Fq2 operator*(const Fq2& a, const Fq2& b) {
    return {Fq::mul2_sub(a.c0, b.c0, a.c1, b.c1), Fq::mul2_add(a.c0, b.c1, a.c1, b.c0)};
  }
  // a*b - c*d over Fq2: 8 products, 2 reductions
  CTG_HD static Fq2 mul2_sub(const Fq2& a, const Fq2& b, const Fq2& c, const Fq2& d) {
    // c0 = a0 b0 - a1 b1 - c0 d0 + c1 d1 ; c1 = a0 b1 + a1 b0 - c0 d1 - c1 d0
    const Fq nb1 = b.c1.neg_lazy(), nd0 = d.c0.neg_lazy(), nd1 = d.c1.neg_lazy();
    const Fq* x0[4] = {&a.c0, &a.c1, &c.c0, &c.c1};
    const Fq* y0[4] = {&b.c0, &nb1, &nd0, &d.c1};
    const Fq* y1[4] = {&b.c1, &b.c0, &nd1, &nd0};
    return {Fq::dot<4>(x0, y0), Fq::dot<4>(x0, y1)};
  }
  CTG_HD Fq2 sqr() const {
    Fq t = c0 * c1;
    return {(c0 + c1) * (c0 - c1), t + t};
  }
  CTG_HD Fq2 inv() const {
    Fq n = (c0.sqr() + c1.sqr()).inv();
    return {c0 * n, (c1 * n).neg()};
  }
  CTG_HD Fq2 inv_bingcd() const {
    Fq n = (c0.sqr() + c1.sqr()).inv_bingcd();
    return {c0 * n, (c1 * n).neg()};
  }
};

}  // namespace ctg
