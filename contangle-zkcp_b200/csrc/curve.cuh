// Short-Weierstrass group arithmetic for BLS12-381 G1 (over Fq) and G2 (over Fq2), a = 0.
// Replaces: ark-ec 0.3.0 models/short_weierstrass_jacobian.rs (add_assign_mixed,
// double_in_place, add_assign, into_affine), reached from lib/src/zk/encryption.rs:76.
// arkworks accumulates in Jacobian coordinates; only the affine value of a result is
// observable, so buckets here use extended Jacobian ("XYZZ": x = X/ZZ, y = Y/ZZZ,
// ZZ^3 = ZZZ^2), whose mixed addition costs 8M + 2S and never needs a square root of ZZ.
// The point at infinity is ZZ = 0 (projective) or the all-zero image (affine; (0,0) is
// not on y^2 = x^3 + b for b != 0).
#pragma once
#include "fp.cuh"

namespace ctg {

template <class F> struct Affine {
  F x, y;
  CTG_HD bool is_inf() const { return x.is_zero() && y.is_zero(); }
  CTG_HD static Affine inf() { return {F::zero(), F::zero()}; }
};

template <class F> struct XYZZ {
  F X, Y, ZZ, ZZZ;
  CTG_HD static XYZZ inf() { return {F::zero(), F::zero(), F::zero(), F::zero()}; }
  CTG_HD bool is_inf() const { return ZZ.is_zero(); }
  CTG_HD static XYZZ from_affine(const Affine<F>& p) {
    if (p.is_inf()) return inf();
    return {p.x, p.y, F::one(), F::one()};
  }
  CTG_HD XYZZ neg() const { return {X, Y.neg(), ZZ, ZZZ}; }
};

// 2*(x, y) for an affine, non-infinity point (mdbl-2008-s-1, a = 0)
template <class F> CTG_HD_COLD XYZZ<F> xyzz_dbl_affine(const Affine<F>& p) {
  F U = p.y.dbl(), V = U.sqr(), W = U * V, S = p.x * V;
  F xx = p.x.sqr(), M = xx.dbl() + xx;
  XYZZ<F> r;
  r.X = M.sqr() - S.dbl();
  r.Y = F::mul2_sub(M, S - r.X, W, p.y);
  r.ZZ = V;
  r.ZZZ = W;
  return r;
}

// dbl-2008-s-1, a = 0
template <class F> CTG_HD_COLD XYZZ<F> xyzz_dbl(const XYZZ<F>& p) {
  if (p.is_inf()) return p;
  F U = p.Y.dbl(), V = U.sqr(), W = U * V, S = p.X * V;
  F xx = p.X.sqr(), M = xx.dbl() + xx;
  XYZZ<F> r;
  r.X = M.sqr() - S.dbl();
  r.Y = F::mul2_sub(M, S - r.X, W, p.Y);
  r.ZZ = V * p.ZZ;
  r.ZZZ = W * p.ZZZ;
  return r;
}

// acc + q, q affine (madd-2008-s).  `neg` adds -q.  Handles every special case
// (acc or q at infinity, q == acc, q == -acc), the rare ones on a divergent slow path.
template <class F> CTG_HD XYZZ<F> xyzz_madd(const XYZZ<F>& acc, const Affine<F>& q_in, bool neg = false) {
  if (q_in.is_inf()) return acc;
  Affine<F> q = q_in;
  if (neg) q.y = q.y.neg();
  if (acc.is_inf()) return {q.x, q.y, F::one(), F::one()};
  F U2 = q.x * acc.ZZ, S2 = q.y * acc.ZZZ;
  F P = U2 - acc.X, R = S2 - acc.Y;
  if (P.is_zero()) {
    if (R.is_zero()) return xyzz_dbl_affine(q);
    return XYZZ<F>::inf();
  }
  F PP = P.sqr(), PPP = P * PP, Q = acc.X * PP;
  XYZZ<F> r;
  r.X = R.sqr() - PPP - Q.dbl();
  r.Y = F::mul2_sub(R, Q - r.X, acc.Y, PPP);  // one reduction for the two products
  r.ZZ = acc.ZZ * PP;
  r.ZZZ = acc.ZZZ * PPP;
  return r;
}

// add-2008-s
template <class F> CTG_HD_COLD XYZZ<F> xyzz_add(const XYZZ<F>& a, const XYZZ<F>& b) {
  if (a.is_inf()) return b;
  if (b.is_inf()) return a;
  F U1 = a.X * b.ZZ, U2 = b.X * a.ZZ, S1 = a.Y * b.ZZZ, S2 = b.Y * a.ZZZ;
  F P = U2 - U1, R = S2 - S1;
  if (P.is_zero()) {
    if (R.is_zero()) return xyzz_dbl(a);
    return XYZZ<F>::inf();
  }
  F PP = P.sqr(), PPP = P * PP, Q = U1 * PP;
  XYZZ<F> r;
  r.X = R.sqr() - PPP - Q.dbl();
  r.Y = F::mul2_sub(R, Q - r.X, S1, PPP);
  r.ZZ = a.ZZ * b.ZZ * PP;
  r.ZZZ = a.ZZZ * b.ZZZ * PPP;
  return r;
}

template <class F> CTG_HD_COLD Affine<F> xyzz_to_affine(const XYZZ<F>& p) {
  if (p.is_inf()) return Affine<F>::inf();
  // 1/ZZZ gives both: 1/ZZ = ZZ^2 / ZZZ^2 * ... keep it simple: one inversion of ZZ*ZZZ
  F t = (p.ZZ * p.ZZZ).inv_bingcd();   // binary GCD: about 12x cheaper than the Fermat ladder, and this sits at the tail of every proof
  F izz = t * p.ZZZ, izzz = t * p.ZZ;
  return {p.X * izz, p.Y * izzz};
}

// k * p, k given as 8 little-endian 32-bit words (canonical integer), left-to-right.
template <class F> CTG_HD_COLD XYZZ<F> xyzz_mul(const XYZZ<F>& p, const uint32_t* k) {
  XYZZ<F> acc = XYZZ<F>::inf();
  int top = 255;
  while (top >= 0 && !((k[top >> 5] >> (top & 31)) & 1)) top--;
  for (int i = top; i >= 0; i--) {
    acc = xyzz_dbl(acc);
    if ((k[i >> 5] >> (i & 31)) & 1) acc = xyzz_add(acc, p);
  }
  return acc;
}

typedef Affine<Fq> G1Affine;
typedef Affine<Fq2> G2Affine;
typedef XYZZ<Fq> G1XYZZ;
typedef XYZZ<Fq2> G2XYZZ;

}  // namespace ctg
