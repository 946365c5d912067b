// Host build of the device math headers (ptx.cuh emulates the PTX carry chain on the
// CPU) so that the exact field / curve code the kernels run can be checked against the
// oracle without a GPU.  Test-only; not part of the product library.
#include "fp.cuh"
#include "curve.cuh"
#include "msm_digits.cuh"
using namespace ctg;
extern "C" {
void hm_fr_mul(const Fr* a, const Fr* b, Fr* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i] * b[i]; }
void hm_fr_add(const Fr* a, const Fr* b, Fr* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i] + b[i]; }
void hm_fr_sub(const Fr* a, const Fr* b, Fr* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i] - b[i]; }
void hm_fq_mul(const Fq* a, const Fq* b, Fq* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i] * b[i]; }
void hm_fq_add(const Fq* a, const Fq* b, Fq* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i] + b[i]; }
void hm_fq_sub(const Fq* a, const Fq* b, Fq* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i] - b[i]; }
void hm_fq2_mul(const Fq2* a, const Fq2* b, Fq2* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i] * b[i]; }
void hm_fr_sqr(const Fr* a, Fr* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i].sqr(); }
void hm_fq_sqr(const Fq* a, Fq* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i].sqr(); }
void hm_fq2_sqr(const Fq2* a, Fq2* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i].sqr(); }
// lazy-reduction sums of products: o = a*b + c*d, o = a*b - c*d (Fq), o = a*b - c*d (Fq2)
void hm_fq_mul2_add(const Fq* a, const Fq* b, const Fq* c, const Fq* d, Fq* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = Fq::mul2_add(a[i], b[i], c[i], d[i]); }
void hm_fq_mul2_sub(const Fq* a, const Fq* b, const Fq* c, const Fq* d, Fq* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = Fq::mul2_sub(a[i], b[i], c[i], d[i]); }
void hm_fq2_mul2_sub(const Fq2* a, const Fq2* b, const Fq2* c, const Fq2* d, Fq2* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = Fq2::mul2_sub(a[i], b[i], c[i], d[i]); }
void hm_fr_from_mont(const Fr* a, Fr* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i].from_mont(); }
void hm_fr_to_mont(const Fr* a, Fr* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i].to_mont(); }
void hm_fq_inv(const Fq* a, Fq* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i].inv(); }
void hm_fq_inv_bingcd(const Fq* a, Fq* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i].inv_bingcd(); }
void hm_fr_inv_bingcd(const Fr* a, Fr* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i].inv_bingcd(); }
void hm_fq2_inv_bingcd(const Fq2* a, Fq2* o, uint64_t n) { for (uint64_t i = 0; i < n; i++) o[i] = a[i].inv_bingcd(); }
// sum of points with per-point sign through the XYZZ mixed addition, then affine
void hm_g1_sum(const G1Affine* p, const uint8_t* neg, uint64_t n, G1Affine* out) {
  G1XYZZ acc = G1XYZZ::inf();
  for (uint64_t i = 0; i < n; i++) acc = xyzz_madd(acc, p[i], neg[i] != 0);
  *out = xyzz_to_affine(acc);
}
void hm_g2_sum(const G2Affine* p, const uint8_t* neg, uint64_t n, G2Affine* out) {
  G2XYZZ acc = G2XYZZ::inf();
  for (uint64_t i = 0; i < n; i++) acc = xyzz_madd(acc, p[i], neg[i] != 0);
  *out = xyzz_to_affine(acc);
}
// tree of full additions (exercises xyzz_add / xyzz_dbl incl. equal operands)
void hm_g1_tree(const G1Affine* p, uint64_t n, G1Affine* out) {
  G1XYZZ buf[64];
  for (uint64_t i = 0; i < n; i++) buf[i] = G1XYZZ::from_affine(p[i]);
  for (uint64_t m = n; m > 1; m = (m + 1) / 2)
    for (uint64_t i = 0; i < m / 2; i++) buf[i] = xyzz_add(buf[2 * i], buf[2 * i + 1]), (void)0;
  // (only valid for n a power of two)
  *out = xyzz_to_affine(buf[0]);
}
void hm_g1_mul(const G1Affine* p, const uint32_t* k, G1Affine* out) { *out = xyzz_to_affine(xyzz_mul(G1XYZZ::from_affine(*p), k)); }
void hm_g2_mul(const G2Affine* p, const uint32_t* k, G2Affine* out) { *out = xyzz_to_affine(xyzz_mul(G2XYZZ::from_affine(*p), k)); }
// signed-digit recoding used by the MSM kernels: digits[w] in [-2^(c-1), 2^(c-1)]
void hm_signed_digits(const uint32_t* k, int c, int nwin, int32_t* digits) {
  uint32_t carry = 0;
  for (int w = 0; w < nwin; w++) digits[w] = msm_signed_digit(k, c, w, nwin, carry);
}
}
