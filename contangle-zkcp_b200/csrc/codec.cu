// arkworks canonical wire formats for the prover's inputs and outputs.
// Replaces, for BLS12-381:
//   ProvingKey::deserialize_unchecked / serialize_unchecked of circuit.zkey   lib/src/utils.rs:85-110
//   ark_to_bytes(vk) (VerifyingKey::serialize, compressed)                    lib/src/utils.rs:94
//   ark_to_bytes(proof) / ark_from_bytes(proof) (Proof::serialize, 192 B)     circuits-ark/src/utils.rs:12-22,
//                                                                             lib/src/zk/encryption.rs:78-84,135-141
// ark-serialize 0.3 / ark-ec 0.3 short_weierstrass_jacobian.rs conventions restated here (no copy of the crates on
// this box; "parity unpinned" like the rest of the path):
//   * Fq: 48 bytes little-endian canonical (non-Montgomery); reading rejects values >= p;
//   * Fq2: c0 then c1;
//   * SWFlags live in the two top bits of the LAST byte of the element they ride on: 0x80 = "y is the larger of
//     {y, -y}" (Fq: canonical integer order; Fq2: c1 first, then c0), 0x40 = point at infinity, both = invalid;
//   * affine compressed: x with flags (infinity: x = 0 | 0x40); uncompressed / unchecked: x, then y with flags
//     (infinity: x = 0, y = 1 | 0x40 -- GroupAffine::zero() is (0, 1, infinity)); a set 0x80 is ignored on reading;
//   * Vec<T>: u64 little-endian length, then the items; structs field by field in declaration order:
//     ProvingKey { vk { alpha_g1, beta_g2, gamma_g2, delta_g2, gamma_abc_g1 }, beta_g1, delta_g1, a_query, b_g1_query,
//     b_g2_query, h_query, l_query };  Proof { a, b, c }.
// The bulk conversions (canonical bytes <-> packed Montgomery affine, 5 queries of up to 2^24 points) run on the
// device: a key image is copied to HBM as bytes and decoded there, straight into the layout ctg_pk_upload takes.
#include <cstring>

#include "ctg_internal.h"

namespace ctg {

static constexpr uint8_t kFlagLarger = 0x80, kFlagInfinity = 0x40;

// ---- device: canonical bytes <-> Montgomery limbs -----------------------------------------------------
// (p - 1) / 2 < y  <=>  y is the larger of {y, p - y}
CTG_HD bool fq_is_larger_half(const Fq& canon) {
  // y > p - y  <=>  2y > p  (p odd): compare 2y with p limb-wise from the top
  uint32_t carry = 0, d[13];
#pragma unroll
  for (int i = 0; i < 12; i++) { d[i] = (canon.v[i] << 1) | carry; carry = canon.v[i] >> 31; }
  d[12] = carry;
  if (d[12]) return true;
  for (int i = 11; i >= 0; i--) {
    const uint32_t m = FqParams::mod(i);
    if (d[i] != m) return d[i] > m;
  }
  return false;
}
CTG_HD bool fq_canon_lt_p(const Fq& canon) {
  for (int i = 11; i >= 0; i--) {
    const uint32_t m = FqParams::mod(i);
    if (canon.v[i] != m) return canon.v[i] < m;
  }
  return false;
}
CTG_HD bool y_is_larger(const Fq& y_canon) { return fq_is_larger_half(y_canon); }
CTG_HD bool y_is_larger(const Fq2& y_canon) { return y_canon.c1.is_zero() ? fq_is_larger_half(y_canon.c0) : fq_is_larger_half(y_canon.c1); }

// coordinates of a point as NC = 2 (G1) or 4 (G2) base-field elements: x[, x.c1], y[, y.c1]
template <class F> struct Coords;
template <> struct Coords<Fq> {
  static constexpr int NC = 2;
  CTG_HD static Fq& get(Affine<Fq>& p, int k) { return k == 0 ? p.x : p.y; }
};
template <> struct Coords<Fq2> {
  static constexpr int NC = 4;
  CTG_HD static Fq& get(Affine<Fq2>& p, int k) { return k == 0 ? p.x.c0 : k == 1 ? p.x.c1 : k == 2 ? p.y.c0 : p.y.c1; }
};

// uncompressed bytes -> packed Montgomery affine.  err[0] is set to 1 on a non-canonical coordinate or bad flags.
template <class F> __global__ void __launch_bounds__(128) decode_uncompressed_kernel(const uint32_t* __restrict__ src, uint64_t count, Affine<F>* __restrict__ out, int* err) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  constexpr int NC = Coords<F>::NC;
  const uint32_t* s = src + i * (NC * 12);
  Affine<F> p;
  bool bad = false;
  uint32_t flags = 0;
#pragma unroll
  for (int k = 0; k < NC; k++) {
    Fq c;
#pragma unroll
    for (int j = 0; j < 12; j++) c.v[j] = s[k * 12 + j];
    if (k == NC - 1) { flags = c.v[11] >> 30; c.v[11] &= 0x3fffffffu; }
    bad |= !fq_canon_lt_p(c);
    Coords<F>::get(p, k) = c.to_mont();
  }
  if (flags == 3) bad = true;
  if (flags & 1) p = Affine<F>::inf();  // bit 30 of the last word = 0x40 of the last byte
  if (bad) atomicExch(err, 1);
  out[i] = p;
}

// packed Montgomery affine -> uncompressed (compressed = 0) or compressed (x only, with the y-order flag) bytes
template <class F> __global__ void __launch_bounds__(128) encode_kernel(const Affine<F>* __restrict__ in, uint64_t count, uint32_t* __restrict__ dst, int compressed) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  constexpr int NC = Coords<F>::NC;
  Affine<F> p = in[i];
  const bool inf = p.is_inf();
  Fq c[NC];
#pragma unroll
  for (int k = 0; k < NC; k++) c[k] = Coords<F>::get(p, k).from_mont();
  if (compressed) {
    uint32_t* d = dst + i * (NC / 2 * 12);
    F y;
    if constexpr (NC == 2) y = c[1]; else y = F{c[2], c[3]};
    const uint32_t flag = inf ? (uint32_t)kFlagInfinity << 24 : (y_is_larger(y) ? (uint32_t)kFlagLarger << 24 : 0u);
#pragma unroll
    for (int k = 0; k < NC / 2; k++)
#pragma unroll
      for (int j = 0; j < 12; j++) d[k * 12 + j] = (inf ? 0u : c[k].v[j]) | ((k == NC / 2 - 1 && j == 11) ? flag : 0u);
  } else {
    uint32_t* d = dst + i * (NC * 12);
#pragma unroll
    for (int k = 0; k < NC; k++)
#pragma unroll
      for (int j = 0; j < 12; j++) {
        uint32_t w = inf ? 0u : c[k].v[j];
        if (inf && k == NC / 2 && j == 0) w = 1u;                                   // y = 1 (y.c0 = 1 for G2)
        if (inf && k == NC - 1 && j == 11) w |= (uint32_t)kFlagInfinity << 24;
        d[k * 12 + j] = w;
      }
  }
}

// ---- host: key images ----------------------------------------------------------------------------------
static constexpr uint64_t kG1U = 96, kG2U = 192, kG1C = 48, kG2C = 96;

template <class F> static int32_t decode_range(ctg_ctx* ctx, const uint8_t* d_bytes, uint64_t off, uint64_t count, Affine<F>* d_out, int* d_err, cudaStream_t st) {
  if (!count) return CTG_OK;
  decode_uncompressed_kernel<F><<<(unsigned)((count + 127) / 128), 128, 0, st>>>(reinterpret_cast<const uint32_t*>(d_bytes + off), count, d_out, d_err);
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}
template <class F> static int32_t encode_range(ctg_ctx* ctx, const Affine<F>* d_in, uint64_t count, uint8_t* d_bytes, uint64_t off, int compressed, cudaStream_t st) {
  if (!count) return CTG_OK;
  encode_kernel<F><<<(unsigned)((count + 127) / 128), 128, 0, st>>>(d_in, count, reinterpret_cast<uint32_t*>(d_bytes + off), compressed);
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}

struct DevBytes {
  void* p = nullptr;
  ~DevBytes() { if (p) cudaFree(p); }
};

static uint64_t pk_image_size(uint64_t ni, uint64_t m, uint64_t n, uint64_t nw) {
  return kG1U + 3 * kG2U + 8 + ni * kG1U + 2 * kG1U + 8 + m * kG1U + 8 + m * kG1U + 8 + m * kG2U + 8 + (n - 1) * kG1U + 8 + nw * kG1U;
}
static uint64_t vk_image_size(uint64_t ni) { return kG1C + 3 * kG2C + 8 + ni * kG1C; }

static int32_t params_deserialize_unchecked(ctg_ctx* ctx, const uint8_t* buf, uint64_t len, ctg_params** out) {
  // walk the length prefixes on the host
  auto need = [&](uint64_t off, uint64_t bytes) { return off <= len && bytes <= len - off; };
  auto rd64 = [&](uint64_t off) { uint64_t v; memcpy(&v, buf + off, 8); return v; };
  const char* trunc = "pk_deserialize: truncated key image";
  uint64_t off = 0;
  const uint64_t o_alpha = off; off += kG1U;
  const uint64_t o_beta2 = off; off += kG2U;
  const uint64_t o_gamma2 = off; off += kG2U;
  const uint64_t o_delta2 = off; off += kG2U;
  if (!need(off, 8)) return fail(ctx, CTG_ERR_INVALID, trunc);
  const uint64_t ni = rd64(off); off += 8;
  if (ni > len / kG1U) return fail(ctx, CTG_ERR_INVALID, trunc);
  const uint64_t o_abc = off; off += ni * kG1U;
  const uint64_t o_beta1 = off; off += kG1U;
  const uint64_t o_delta1 = off; off += kG1U;
  uint64_t cnt[5], o_q[5];
  const uint64_t item[5] = {kG1U, kG1U, kG2U, kG1U, kG1U};  // a, b_g1, b_g2, h, l
  for (int k = 0; k < 5; k++) {
    if (!need(off, 8)) return fail(ctx, CTG_ERR_INVALID, trunc);
    cnt[k] = rd64(off); off += 8;
    if (cnt[k] > len / item[k]) return fail(ctx, CTG_ERR_INVALID, trunc);
    o_q[k] = off; off += cnt[k] * item[k];
  }
  if (off > len) return fail(ctx, CTG_ERR_INVALID, trunc);
  if (off != len) return fail(ctx, CTG_ERR_INVALID, "pk_deserialize: trailing bytes after the key");
  const uint64_t m = cnt[0];
  if (ni == 0 || m < ni || cnt[1] != m || cnt[2] != m || cnt[4] != m - ni)
    return fail(ctx, CTG_ERR_INVALID, "pk_deserialize: query lengths are inconsistent (a = b_g1 = b_g2 = m, l = m - num_instance)");
  const uint64_t n = cnt[3] + 1;
  if (n & (n - 1)) return fail(ctx, CTG_ERR_INVALID, "pk_deserialize: h_query length + 1 is not a power of two");
  if (n < ni) return fail(ctx, CTG_ERR_INVALID, "pk_deserialize: domain smaller than num_instance");

  cudaStream_t st = ctx->stream;
  DevBytes img, err;
  if (cudaMalloc(&img.p, len) != cudaSuccess || cudaMalloc(&err.p, sizeof(int)) != cudaSuccess) { cudaGetLastError(); return fail(ctx, CTG_ERR_OOM, "pk_deserialize: cudaMalloc failed"); }
  CTG_CUDA(ctx, cudaMemsetAsync(err.p, 0, sizeof(int), st));
  CTG_CUDA(ctx, cudaMemcpyAsync(img.p, buf, len, cudaMemcpyHostToDevice, st));
  ctg_params* p = nullptr;
  CTG_TRY(params_alloc(ctx, ni, m - ni, n - ni, n, &p));
  struct Guard { ctg_params* p; ~Guard() { if (p) params_free(p); } } guard{p};
  const uint8_t* d = (const uint8_t*)img.p;
  int* d_err = (int*)err.p;
  CTG_TRY(decode_range<Fq>(ctx, d, o_alpha, 1, p->singles_g1 + 0, d_err, st));
  CTG_TRY(decode_range<Fq>(ctx, d, o_beta1, 1, p->singles_g1 + 1, d_err, st));
  CTG_TRY(decode_range<Fq>(ctx, d, o_delta1, 1, p->singles_g1 + 2, d_err, st));
  CTG_TRY(decode_range<Fq2>(ctx, d, o_beta2, 1, p->singles_g2 + 0, d_err, st));
  CTG_TRY(decode_range<Fq2>(ctx, d, o_gamma2, 1, p->singles_g2 + 1, d_err, st));
  CTG_TRY(decode_range<Fq2>(ctx, d, o_delta2, 1, p->singles_g2 + 2, d_err, st));
  CTG_TRY(decode_range<Fq>(ctx, d, o_abc, ni, p->gamma_abc_g1, d_err, st));
  CTG_TRY(decode_range<Fq>(ctx, d, o_q[0], m, p->a_query, d_err, st));
  CTG_TRY(decode_range<Fq>(ctx, d, o_q[1], m, p->b_g1_query, d_err, st));
  CTG_TRY(decode_range<Fq2>(ctx, d, o_q[2], m, p->b_g2_query, d_err, st));
  CTG_TRY(decode_range<Fq>(ctx, d, o_q[3], n - 1, p->h_query, d_err, st));
  CTG_TRY(decode_range<Fq>(ctx, d, o_q[4], m - ni, p->l_query, d_err, st));
  int bad = 0;
  CTG_CUDA(ctx, cudaMemcpyAsync(&bad, d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
  CTG_CUDA(ctx, cudaStreamSynchronize(st));
  if (bad) return fail(ctx, CTG_ERR_INVALID, "pk_deserialize: non-canonical field element or invalid point flags (arkworks: SerializationError::InvalidData)");
  guard.p = nullptr;
  *out = p;
  return CTG_OK;
}

static int32_t params_serialize(ctg_ctx* ctx, const ctg_params* p, int vk_compressed, uint8_t* out, uint64_t cap, uint64_t* written) {
  const uint64_t size = vk_compressed ? vk_image_size(p->ni) : pk_image_size(p->ni, p->m, p->n, p->nw);
  if (written) *written = size;
  if (!out) return CTG_OK;  // size query
  if (cap < size) return fail(ctx, CTG_ERR_INVALID, "serialize: output buffer too small");
  cudaStream_t st = ctx->stream;
  DevBytes img;
  if (cudaMalloc(&img.p, size) != cudaSuccess) { cudaGetLastError(); return fail(ctx, CTG_ERR_OOM, "serialize: cudaMalloc failed"); }
  uint8_t* d = (uint8_t*)img.p;
  uint64_t off = 0;
  auto put_len = [&](uint64_t v) { cudaError_t e = cudaMemcpyAsync(d + off, &v, 8, cudaMemcpyHostToDevice, st); off += 8; return e; };
  // lengths are staged through a small host array that must outlive the async copies
  if (vk_compressed) {
    CTG_TRY(encode_range<Fq>(ctx, p->singles_g1 + 0, 1, d, off, 1, st)); off += kG1C;
    for (int k = 0; k < 3; k++) { CTG_TRY(encode_range<Fq2>(ctx, p->singles_g2 + k, 1, d, off, 1, st)); off += kG2C; }
    CTG_CUDA(ctx, put_len(p->ni));
    CTG_CUDA(ctx, cudaStreamSynchronize(st));
    CTG_TRY(encode_range<Fq>(ctx, p->gamma_abc_g1, p->ni, d, off, 1, st)); off += p->ni * kG1C;
  } else {
    CTG_TRY(encode_range<Fq>(ctx, p->singles_g1 + 0, 1, d, off, 0, st)); off += kG1U;
    for (int k = 0; k < 3; k++) { CTG_TRY(encode_range<Fq2>(ctx, p->singles_g2 + k, 1, d, off, 0, st)); off += kG2U; }
    CTG_CUDA(ctx, put_len(p->ni)); CTG_CUDA(ctx, cudaStreamSynchronize(st));
    CTG_TRY(encode_range<Fq>(ctx, p->gamma_abc_g1, p->ni, d, off, 0, st)); off += p->ni * kG1U;
    CTG_TRY(encode_range<Fq>(ctx, p->singles_g1 + 1, 1, d, off, 0, st)); off += kG1U;
    CTG_TRY(encode_range<Fq>(ctx, p->singles_g1 + 2, 1, d, off, 0, st)); off += kG1U;
    CTG_CUDA(ctx, put_len(p->m)); CTG_CUDA(ctx, cudaStreamSynchronize(st));
    CTG_TRY(encode_range<Fq>(ctx, p->a_query, p->m, d, off, 0, st)); off += p->m * kG1U;
    CTG_CUDA(ctx, put_len(p->m)); CTG_CUDA(ctx, cudaStreamSynchronize(st));
    CTG_TRY(encode_range<Fq>(ctx, p->b_g1_query, p->m, d, off, 0, st)); off += p->m * kG1U;
    CTG_CUDA(ctx, put_len(p->m)); CTG_CUDA(ctx, cudaStreamSynchronize(st));
    CTG_TRY(encode_range<Fq2>(ctx, p->b_g2_query, p->m, d, off, 0, st)); off += p->m * kG2U;
    CTG_CUDA(ctx, put_len(p->n - 1)); CTG_CUDA(ctx, cudaStreamSynchronize(st));
    CTG_TRY(encode_range<Fq>(ctx, p->h_query, p->n - 1, d, off, 0, st)); off += (p->n - 1) * kG1U;
    CTG_CUDA(ctx, put_len(p->nw)); CTG_CUDA(ctx, cudaStreamSynchronize(st));
    CTG_TRY(encode_range<Fq>(ctx, p->l_query, p->nw, d, off, 0, st)); off += p->nw * kG1U;
  }
  if (off != size) return fail(ctx, CTG_ERR_STATE, "serialize: internal size mismatch");
  CTG_CUDA(ctx, cudaMemcpyAsync(out, d, size, cudaMemcpyDeviceToHost, st));
  CTG_CUDA(ctx, cudaStreamSynchronize(st));
  return CTG_OK;
}

// ---- host: proofs (3 points; no device involved) ---------------------------------------------------------
static void put_fq(uint8_t* dst, const Fq& canon) { memcpy(dst, canon.v, 48); }  // little-endian host
static bool get_fq(const uint8_t* src, Fq* canon, uint8_t* flags /* may be null: no flag bits expected */) {
  memcpy(canon->v, src, 48);
  if (flags) { *flags = (uint8_t)((canon->v[11] >> 24) & 0xc0); canon->v[11] &= 0x3fffffffu; }
  return fq_canon_lt_p(*canon);
}
static Fq fq_pow(const Fq& b, const uint32_t* e, int words) {
  Fq acc = Fq::one();
  for (int i = words * 32 - 1; i >= 0; i--) {
    acc = acc.sqr();
    if ((e[i >> 5] >> (i & 31)) & 1) acc = acc * b;
  }
  return acc;
}
// sqrt in Fq, p = 3 mod 4: a^((p + 1) / 4); false if a is not a square
static bool fq_sqrt(const Fq& a, Fq* out) {
  uint32_t e[12];
  uint64_t carry = 1;
  for (int i = 0; i < 12; i++) { uint64_t t = (uint64_t)FqParams::mod(i) + carry; e[i] = (uint32_t)t; carry = t >> 32; }
  for (int i = 0; i < 12; i++) e[i] = (e[i] >> 2) | (i + 1 < 12 ? e[i + 1] << 30 : 0u);
  Fq r = fq_pow(a, e, 12);
  if (!r.sqr().eq(a)) return false;
  *out = r;
  return true;
}
// sqrt in Fq2 = Fq[u]/(u^2 + 1) through the norm: with s = sqrt(a0^2 + a1^2), x0^2 = (a0 +- s)/2, x1 = a1 / (2 x0)
static bool fq2_sqrt(const Fq2& a, Fq2* out) {
  if (a.is_zero()) { *out = Fq2::zero(); return true; }
  Fq2 r;
  if (a.c1.is_zero()) {
    Fq t;
    if (fq_sqrt(a.c0, &t)) r = {t, Fq::zero()};
    else if (fq_sqrt(a.c0.neg(), &t)) r = {Fq::zero(), t};
    else return false;
  } else {
    Fq s;
    if (!fq_sqrt(a.c0.sqr() + a.c1.sqr(), &s)) return false;
    const Fq half = (Fq::one() + Fq::one()).inv();
    Fq x0;
    if (!fq_sqrt((a.c0 + s) * half, &x0) && !fq_sqrt((a.c0 - s) * half, &x0)) return false;
    r = {x0, a.c1 * (x0 + x0).inv()};
  }
  if (!r.sqr().eq(a)) return false;
  *out = r;
  return true;
}
static Fq curve_b(const Fq*) { Fq b = Fq::zero(); b.v[0] = 4; return b.to_mont(); }
static Fq2 curve_b(const Fq2*) { Fq b = curve_b((const Fq*)nullptr); return {b, b}; }
static bool field_sqrt(const Fq& a, Fq* o) { return fq_sqrt(a, o); }
static bool field_sqrt(const Fq2& a, Fq2* o) { return fq2_sqrt(a, o); }
static Fq canon_of(const Fq& a) { return a.from_mont(); }
static Fq2 canon_of(const Fq2& a) { return {a.c0.from_mont(), a.c1.from_mont()}; }
static const uint32_t kFrModulus[8] = {0x00000001u, 0xffffffffu, 0xfffe5bfeu, 0x53bda402u, 0x09a1d805u, 0x3339d808u, 0x299d7d48u, 0x73eda753u};

template <class F> static void point_serialize_compressed(const Affine<F>& p_in, uint8_t* dst) {
  constexpr int NC = Coords<F>::NC;
  Affine<F> p = p_in;
  memset(dst, 0, NC / 2 * 48);
  if (p.is_inf()) { dst[NC / 2 * 48 - 1] = kFlagInfinity; return; }
  for (int k = 0; k < NC / 2; k++) put_fq(dst + 48 * k, Coords<F>::get(p, k).from_mont());
  if (y_is_larger(canon_of(p.y))) dst[NC / 2 * 48 - 1] |= kFlagLarger;
}
// GroupAffine::deserialize: get_point_from_x(x, greatest) then the subgroup check; 0 ok, -1 invalid
template <class F> static int point_deserialize_compressed(const uint8_t* src, int check_subgroup, Affine<F>* out) {
  constexpr int NC = Coords<F>::NC;
  Affine<F> p = Affine<F>::inf();
  uint8_t flags = 0;
  for (int k = 0; k < NC / 2; k++) {
    Fq c;
    if (!get_fq(src + 48 * k, &c, k == NC / 2 - 1 ? &flags : nullptr)) return -1;
    Coords<F>::get(p, k) = c.to_mont();
  }
  if (flags == (kFlagLarger | kFlagInfinity)) return -1;
  if (flags & kFlagInfinity) { *out = Affine<F>::inf(); return 0; }
  F y;
  if (!field_sqrt(p.x.sqr() * p.x + curve_b((const F*)nullptr), &y)) return -1;
  if (y_is_larger(canon_of(y)) != ((flags & kFlagLarger) != 0)) y = y.neg();
  p.y = y;
  if (check_subgroup && !xyzz_mul(XYZZ<F>::from_affine(p), kFrModulus).is_inf()) return -1;
  *out = p;
  return 0;
}

}  // namespace ctg

// ---- C ABI -------------------------------------------------------------------------------------------------
using namespace ctg;
extern "C" {

int32_t ctg_proof_serialize(const ctg_proof* proof, uint8_t* out192) {
  if (!proof || !out192) return CTG_ERR_INVALID;
  G1Affine a, c; G2Affine b;
  memcpy(&a, &proof->a, sizeof(a)); memcpy(&b, &proof->b, sizeof(b)); memcpy(&c, &proof->c, sizeof(c));
  point_serialize_compressed<Fq>(a, out192);
  point_serialize_compressed<Fq2>(b, out192 + 48);
  point_serialize_compressed<Fq>(c, out192 + 144);
  return CTG_OK;
}
int32_t ctg_proof_deserialize(const uint8_t* in192, int check_subgroup, ctg_proof* out) {
  if (!in192 || !out) return CTG_ERR_INVALID;
  G1Affine a, c; G2Affine b;
  if (point_deserialize_compressed<Fq>(in192, check_subgroup, &a) || point_deserialize_compressed<Fq2>(in192 + 48, check_subgroup, &b) ||
      point_deserialize_compressed<Fq>(in192 + 144, check_subgroup, &c))
    return CTG_ERR_INVALID;
  memcpy(&out->a, &a, sizeof(a)); memcpy(&out->b, &b, sizeof(b)); memcpy(&out->c, &c, sizeof(c));
  return CTG_OK;
}

int32_t ctg_params_deserialize_unchecked(ctg_ctx* ctx, const uint8_t* key_image, uint64_t len, ctg_params** out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  if (!key_image) return fail(ctx, CTG_ERR_INVALID, "pk_deserialize: null buffer");
  return params_deserialize_unchecked(ctx, key_image, len, out);
}
int32_t ctg_params_serialize_unchecked(ctg_ctx* ctx, const ctg_params* params, uint8_t* out, uint64_t capacity, uint64_t* written) {
  if (!ctx) return CTG_ERR_INVALID;
  if (!params) return fail(ctx, CTG_ERR_INVALID, "serialize: null key");
  return params_serialize(ctx, params, 0, out, capacity, written);
}
int32_t ctg_params_serialize_vk(ctg_ctx* ctx, const ctg_params* params, uint8_t* out, uint64_t capacity, uint64_t* written) {
  if (!ctx) return CTG_ERR_INVALID;
  if (!params) return fail(ctx, CTG_ERR_INVALID, "serialize: null key");
  return params_serialize(ctx, params, 1, out, capacity, written);
}

}  // extern "C"
