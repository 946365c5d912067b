// Bulk scalar multiplication on a twisted Edwards curve -x^2 + y^2 = 1 + d x^2 y^2 over BLS12-381 Fr (JubJub when
// d = -10240/10241): the group part of the reference's hashed-ElGamal encryption, for many randomness values at once.
// SURVEY.md §8(f) row 4; replaces, in bulk,
//   /root/reference circuits-ark/src/encryption.rs:160-165   c1 = G.mul(r);  p_r = pk.mul(r);  p_ra = p_r.into_affine()
//   /root/reference circuits-ark/src/encryption.rs:126-137   keygen: pk = G.mul(sk)
// ark-ec ^0.3.0 (twisted_edwards_extended.rs, not in the tree -- parity unpinned) multiplies by double-and-add over the
// scalar's BigInteger bits; affine results are unique group elements, so any correct group law gives the same bytes.
// Here: extended coordinates (X : Y : T : Z) with a = -1 (Hisil-Wong-Carter-Dawson: doubling 4M + 4S, mixed addition 7M + one
// product by 2d; both complete because d is a non-square), 256 doublings, one binary-GCD inversion per result.
// TIMING: like ark-ec's `mul`, both kernels are VARIABLE-TIME in the scalar (the plain kernel branches on each bit, the table
// kernel skips zero digits and indexes its table by the digit): do not run them on a GPU shared with an adversary who can time
// kernels.  Scratch that held scalars, assignments or trapdoor powers is zeroed before it is freed or handed back.
// One thread per scalar; integer-pipe bound (about 3 100 Fr products per scalar, 96 bytes in and 64 out).  With one shared base
// and thousands of scalars (bulk encrypt: G and pk) a window table replaces the doublings: <= 32 additions per scalar.
#include <string>
#include <vector>

#include "ctg_internal.h"

namespace ctg {

struct TePoint { Fr x, y, t, z; };

__device__ __forceinline__ TePoint te_dbl(const TePoint& p) {
  const Fr a = p.x.sqr(), b = p.y.sqr(), c = p.z.sqr().dbl();
  const Fr e = (p.x + p.y).sqr() - a - b;
  const Fr g = b - a;          // D + B with D = -A
  const Fr f = g - c;
  const Fr h = (a + b).neg();  // D - B
  return TePoint{e * f, g * h, e * h, f * g};
}

// q affine: (x2, y2), t2d = 2 d x2 y2
__device__ __forceinline__ TePoint te_add_mixed(const TePoint& p, const Fr& x2, const Fr& y2, const Fr& t2d) {
  const Fr a = (p.y - p.x) * (y2 - x2), b = (p.y + p.x) * (y2 + x2);
  const Fr c = p.t * t2d, dd = p.z.dbl();
  const Fr e = b - a, f = dd - c, g = dd + c, h = b + a;
  return TePoint{e * f, g * h, e * h, f * g};
}

__device__ __forceinline__ Fr te_load(const Fr* p) {
  const uint4* q = reinterpret_cast<const uint4*>(p);
  uint4 a = __ldg(q), b = __ldg(q + 1);
  Fr r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}

// out[i] = scalars[i] * points[shared ? 0 : i]   (affine in, affine out; scalars: 256-bit little-endian integers)
__global__ void __launch_bounds__(128) te_mul_kernel(Fr d2, const Fr* __restrict__ points, int shared, const uint32_t* __restrict__ scalars, uint64_t n,
                                                     Fr* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Fr* pp = points + (shared ? 0 : 2 * i);
  const Fr x2 = te_load(pp), y2 = te_load(pp + 1);
  const Fr t2d = x2 * y2 * d2;
  uint32_t k[8];
  {
    const uint4* q = reinterpret_cast<const uint4*>(scalars + 8 * i);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    k[0] = a.x; k[1] = a.y; k[2] = a.z; k[3] = a.w; k[4] = b.x; k[5] = b.y; k[6] = b.z; k[7] = b.w;
  }
  TePoint acc{Fr::zero(), Fr::one(), Fr::zero(), Fr::one()};
  for (int w = 7; w >= 0; w--) {
    const uint32_t kw = k[w];
    for (int b = 31; b >= 0; b--) {
      acc = te_dbl(acc);
      if ((kw >> b) & 1) acc = te_add_mixed(acc, x2, y2, t2d);
    }
  }
  const Fr zi = acc.z.inv_bingcd();
  const Fr xo = acc.x * zi, yo = acc.y * zi;
  uint4* o = reinterpret_cast<uint4*>(out + 2 * i);
  o[0] = make_uint4(xo.v[0], xo.v[1], xo.v[2], xo.v[3]);
  o[1] = make_uint4(xo.v[4], xo.v[5], xo.v[6], xo.v[7]);
  o[2] = make_uint4(yo.v[0], yo.v[1], yo.v[2], yo.v[3]);
  o[3] = make_uint4(yo.v[4], yo.v[5], yo.v[6], yo.v[7]);
}

// Shared base, many scalars: a table of j * 2^(8w) * P (w < 32, 0 < j < 256; built by the kernel above, 8 160 scalar
// multiplications) turns every product into at most 32 mixed additions and no doubling.  The table (510 KB) lives in L2.
static constexpr int kTeWindows = 32, kTeDigits = 255;
static constexpr uint64_t kTeTableFrom = 4096;     // below this many scalars the table does not pay for itself

__global__ void __launch_bounds__(128) te_mul_table_kernel(Fr d2, const Fr* __restrict__ table, const uint32_t* __restrict__ scalars, uint64_t n,
                                                           Fr* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  TePoint acc{Fr::zero(), Fr::one(), Fr::zero(), Fr::one()};
  for (int w = 0; w < kTeWindows; w++) {
    const uint32_t digit = (__ldg(scalars + 8 * i + (w >> 2)) >> ((w & 3) * 8)) & 255u;
    if (digit) {
      const Fr* e = table + 2 * ((uint64_t)w * kTeDigits + digit - 1);
      const Fr x2 = te_load(e), y2 = te_load(e + 1);
      acc = te_add_mixed(acc, x2, y2, x2 * y2 * d2);
    }
  }
  const Fr zi = acc.z.inv_bingcd();
  const Fr xo = acc.x * zi, yo = acc.y * zi;
  uint4* o = reinterpret_cast<uint4*>(out + 2 * i);
  o[0] = make_uint4(xo.v[0], xo.v[1], xo.v[2], xo.v[3]);
  o[1] = make_uint4(xo.v[4], xo.v[5], xo.v[6], xo.v[7]);
  o[2] = make_uint4(yo.v[0], yo.v[1], yo.v[2], yo.v[3]);
  o[3] = make_uint4(yo.v[4], yo.v[5], yo.v[6], yo.v[7]);
}

static int32_t te_mul_shared_table(ctg_ctx* ctx, const Fr& d2, const Fr* d_base, const uint32_t* d_scalars, uint64_t n, Fr* d_out, cudaStream_t st) {
  const uint64_t entries = (uint64_t)kTeWindows * kTeDigits;
  std::vector<uint32_t> ks(entries * 8, 0u);
  for (int w = 0; w < kTeWindows; w++)
    for (int j = 1; j <= kTeDigits; j++) ks[((uint64_t)w * kTeDigits + j - 1) * 8 + (w >> 2)] = (uint32_t)j << ((w & 3) * 8);
  char* buf = nullptr;
  const size_t tab_bytes = entries * 2 * sizeof(Fr), ks_bytes = entries * 32;
  CTG_CUDA(ctx, cudaMallocAsync((void**)&buf, tab_bytes + ks_bytes, st));
  Fr* d_table = (Fr*)buf;
  uint32_t* d_ks = (uint32_t*)(buf + tab_bytes);
  cudaError_t e = cudaMemcpyAsync(d_ks, ks.data(), ks_bytes, cudaMemcpyHostToDevice, st);   // pageable source: staged before the call returns
  if (e == cudaSuccess) {
    te_mul_kernel<<<(unsigned)((entries + 127) / 128), 128, 0, st>>>(d2, d_base, 1, d_ks, entries, d_table);
    te_mul_table_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d2, d_table, d_scalars, n, d_out);
    ctx->launches += 2;
    e = cudaGetLastError();
  }
  cudaFreeAsync(buf, st);
  if (e != cudaSuccess) return fail(ctx, CTG_ERR_CUDA, std::string("te_mul: ") + cudaGetErrorString(e));
  return CTG_OK;
}

static int32_t te_mul_dev(ctg_ctx* ctx, const Fr& d, const Fr* d_points, int shared, const uint32_t* d_scalars, uint64_t n, Fr* d_out, cudaStream_t st) {
  if (n == 0) return CTG_OK;
  if (shared && n >= kTeTableFrom) return te_mul_shared_table(ctx, d + d, d_points, d_scalars, n, d_out, st);
  const uint64_t blocks = (n + 127) / 128;
  te_mul_kernel<<<(unsigned)blocks, 128, 0, st>>>(d + d, d_points, shared, d_scalars, n, d_out);
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}

}  // namespace ctg

using namespace ctg;

extern "C" {

int32_t ctg_te_mul_batch_dev(ctg_ctx* ctx, const ctg_fr* d, const void* d_points, uint64_t n_points, const void* d_scalars, uint64_t n, void* d_out) {
  if (!ctx || !d) return ctx ? fail(ctx, CTG_ERR_INVALID, "te_mul: null coefficient") : CTG_ERR_INVALID;
  if (n_points != 1 && n_points != n) return fail(ctx, CTG_ERR_INVALID, "te_mul: n_points must be 1 (shared base) or n");
  if (n && (!d_points || !d_scalars || !d_out)) return fail(ctx, CTG_ERR_INVALID, "te_mul: null buffer");
  Fr dd;
  memcpy(&dd, d, sizeof(Fr));
  return te_mul_dev(ctx, dd, (const Fr*)d_points, n_points == 1, (const uint32_t*)d_scalars, n, (Fr*)d_out, ctx->stream);
}

int32_t ctg_te_mul_batch(ctg_ctx* ctx, const ctg_fr* d, const ctg_fr* points, uint64_t n_points, const uint32_t* scalars, uint64_t n, ctg_fr* out) {
  if (!ctx || !d) return ctx ? fail(ctx, CTG_ERR_INVALID, "te_mul: null coefficient") : CTG_ERR_INVALID;
  if (n_points != 1 && n_points != n) return fail(ctx, CTG_ERR_INVALID, "te_mul: n_points must be 1 (shared base) or n");
  if (n == 0) return CTG_OK;
  if (!points || !scalars || !out) return fail(ctx, CTG_ERR_INVALID, "te_mul: null buffer");
  const size_t pb = (size_t)n_points * 2 * sizeof(Fr), sb = (size_t)n * 32, ob = (size_t)n * 2 * sizeof(Fr);
  CTG_TRY(ensure_scratch(ctx, pb + sb + ob));
  char* base = (char*)ctx->scratch;
  Fr* d_points = (Fr*)base;
  uint32_t* d_scalars = (uint32_t*)(base + pb);
  Fr* d_out = (Fr*)(base + pb + sb);
  CTG_CUDA(ctx, cudaMemcpyAsync(d_points, points, pb, cudaMemcpyHostToDevice, ctx->stream));
  CTG_CUDA(ctx, cudaMemcpyAsync(d_scalars, scalars, sb, cudaMemcpyHostToDevice, ctx->stream));
  Fr dd;
  memcpy(&dd, d, sizeof(Fr));
  CTG_TRY(te_mul_dev(ctx, dd, d_points, n_points == 1, d_scalars, n, d_out, ctx->stream));
  CTG_CUDA(ctx, cudaMemcpyAsync(out, d_out, ob, cudaMemcpyDeviceToHost, ctx->stream));
  CTG_CUDA(ctx, cudaMemsetAsync(d_scalars, 0, sb, ctx->stream));   // the scalars are secrets (sk of keygen, r of encrypt): not left in scratch
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}

}  // extern "C"
