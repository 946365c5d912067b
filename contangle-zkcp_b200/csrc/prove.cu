// The Groth16 prover after constraint synthesis, on one GPU.
// Replaces ark-groth16 0.3.0 create_proof (prover.rs) and R1CStoQAP::witness_map
// (r1cs_to_qap.rs) as reached from the reference's Groth16::prove calls
// (lib/src/zk/encryption.rs:76, verifiable_encryption.rs:92, sample_entries.rs:86,
// property.rs:133), plus the device-resident replacement of the host ProvingKey cache
// (lib/src/utils.rs:104-110, lib/src/zk/encryption.rs:22,31-32).
//
// Device layout of the key (built once by ctg_pk_upload):
//   a_ext  = a_query    ++ [alpha_g1, delta_g1, O]      scalars z ++ [1, r, s]  ->  A
//   b1_ext = b_g1_query ++ [beta_g1,  O, delta_g1]      scalars z ++ [1, r, s]  ->  B (in G1)
//   b2_ext = b_g2_query ++ [beta_g2,  O, delta_g2]      scalars z ++ [1, r, s]  ->  B (in G2)
//   l      = l_query                                     scalars z[num_instance..]
//   h_ext  : h_ext[bitrev(i)] = h_query[i] (i < n-1), h_ext[n-1] = delta_g1
//            scalars = h in bit-reversed order with slot n-1 (whose coefficient is always 0)
//            overwritten by -(r s)                                              ->  H - rs delta
// so r*delta, s*delta, alpha, beta and -rs*delta ride along in the MSMs (z[0] = 1 already
// selects query[0]).  Only s*A and r*B1 remain as scalar multiplications of results.
#include <algorithm>
#include <cstdlib>

#include "ctg_internal.h"

namespace ctg {

// ---- witness map kernels --------------------------------------------------------------------
// out[i] = <M_i, z> for i < nc; for the A matrix rows nc..nc+ni copy z[j] (input consistency); else 0
// blockIdx.y = proof of a batch: assignment z + y * z_stride, output out + y * n (0 strides for a single proof)
__global__ void spmv_kernel(const uint64_t* __restrict__ row_ptr, const uint32_t* __restrict__ col, const Fr* __restrict__ val,
                            const Fr* __restrict__ z, Fr* __restrict__ out, uint64_t nc, uint64_t ni, uint64_t n, int is_a, uint64_t z_stride) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  z += (uint64_t)blockIdx.y * z_stride;
  out += (uint64_t)blockIdx.y * n;
  Fr acc = Fr::zero();
  if (i < nc) {
    const uint64_t lo = row_ptr[i], hi = row_ptr[i + 1];
    for (uint64_t k = lo; k < hi; k++) acc = acc + val[k] * z[col[k]];
  } else if (is_a && i < nc + ni) {
    acc = z[i - nc];
  }
  out[i] = acc;
}
// a[i] = (a[i] * b[i] - c[i]) * zinv       (vanishing polynomial is the constant g^n - 1 on the coset)
__global__ void pointwise_kernel(Fr* __restrict__ a, const Fr* __restrict__ b, const Fr* __restrict__ c, Fr zinv, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  a[i] = (a[i] * b[i] - c[i]) * zinv;
}
__global__ void set_tail_kernel(Fr* z_ext, uint64_t m, Fr r, Fr s) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    z_ext[m] = Fr::one();
    z_ext[m + 1] = r;
    z_ext[m + 2] = s;
  }
}
__global__ void set_h_tail_kernel(Fr* h, uint64_t n, Fr r, Fr s) {
  if (threadIdx.x == 0 && blockIdx.x == 0) h[n - 1] = (r * s).neg();
}
// batched forms: proof p of the batch has its assignment at z_ext + p * (m + 3), its h at h + p * n and (r, s) at rs[2p], rs[2p + 1]
__global__ void set_tail_batch_kernel(Fr* z_ext, uint64_t m, const Fr* __restrict__ rs, uint32_t batch) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= batch) return;
  Fr* t = z_ext + (uint64_t)p * (m + 3) + m;
  t[0] = Fr::one();
  t[1] = rs[2 * p];
  t[2] = rs[2 * p + 1];
}
__global__ void set_h_tail_batch_kernel(Fr* h, uint64_t n, const Fr* __restrict__ rs, uint32_t batch) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= batch) return;
  h[(uint64_t)p * n + (n - 1)] = (rs[2 * p] * rs[2 * p + 1]).neg();
}
__global__ void bitrev_gather_fr_kernel(const Fr* __restrict__ in, Fr* __restrict__ out, int log_n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (1ull << log_n)) return;
  uint64_t j = log_n == 0 ? 0 : (uint64_t)(__brev((uint32_t)i) >> (32 - log_n));
  out[i] = in[j];
}
// dst[j - lo] = src[bitrev(j)] for j in [lo, lo + cnt) with bitrev(j) < count (count = n - 1), 96-byte points
__global__ void bitrev_gather_g1_kernel(const G1Affine* __restrict__ src, G1Affine* __restrict__ dst, uint64_t count, int log_n, uint64_t lo, uint64_t cnt) {
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= cnt) return;
  const uint64_t j = lo + t;
  const uint64_t i = log_n == 0 ? 0 : (uint64_t)(__brev((uint32_t)j) >> (32 - log_n));
  if (i >= count) return;
  const uint4* s = reinterpret_cast<const uint4*>(src + i);
  uint4* d = reinterpret_cast<uint4*>(dst + t);
#pragma unroll
  for (int k = 0; k < 6; k++) d[k] = s[k];
}
// sums the per-rank partial results of a sharded proof: in[k][0..3] G1 (A, B1, L, H), in2[k] G2
__global__ void __launch_bounds__(160) combine_partials_kernel(const G1XYZZ* g1, const G2XYZZ* g2, uint32_t k, G1XYZZ* out_g1, G2XYZZ* out_g2) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane != 0) return;
  if (warp < 4) {
    G1XYZZ acc = g1[warp];
    for (uint32_t i = 1; i < k; i++) acc = xyzz_add(acc, g1[(size_t)i * 4 + warp]);
    out_g1[warp] = acc;
  } else {
    G2XYZZ acc = g2[0];
    for (uint32_t i = 1; i < k; i++) acc = xyzz_add(acc, g2[i]);
    *out_g2 = acc;
  }
}

// ---- final assembly ---------------------------------------------------------------------------
// in: A, B1, L, H' (G1 XYZZ), B2 (G2 XYZZ), r, s (Montgomery).  C = s*A + r*B1 + L + H'.
struct FinalizeIn { const G1XYZZ *a, *b1, *l, *h; const G2XYZZ* b2; Fr r, s; };
__global__ void __launch_bounds__(96) finalize_kernel(FinalizeIn in, ctg_proof* out) {
  __shared__ G1XYZZ sh[2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    if (warp == 0) {
      Fr k = in.s.from_mont();
      sh[0] = xyzz_mul(*in.a, k.v);
    } else if (warp == 1) {
      Fr k = in.r.from_mont();
      sh[1] = xyzz_mul(*in.b1, k.v);
    } else {
      G2Affine b = xyzz_to_affine(*in.b2);
      *reinterpret_cast<G2Affine*>(&out->b) = b;
    }
  }
  __syncthreads();
  if (lane == 0) {
    if (warp == 0) {
      G1XYZZ c = xyzz_add(sh[0], sh[1]);
      c = xyzz_add(c, *in.l);
      c = xyzz_add(c, *in.h);
      G1Affine ca = xyzz_to_affine(c);
      *reinterpret_cast<G1Affine*>(&out->c) = ca;
    } else if (warp == 1) {
      G1Affine aa = xyzz_to_affine(*in.a);
      *reinterpret_cast<G1Affine*>(&out->a) = aa;
    }
  }
}

// The two scalar multiplications of the assembly, s * A and r * B1, are 255 dependent doublings each (about 1.4 ms for a lone
// thread).  A one-GPU proof therefore launches each on the stream of ITS MSM as soon as that MSM is done -- they then run under
// the G2 MSM, which finishes last -- and the closing kernel only adds four points and converts three to affine.
__global__ void __launch_bounds__(32) scalar_mul_g1_kernel(const G1XYZZ* p, Fr k_mont, G1XYZZ* out) {
  if (threadIdx.x == 0) {
    Fr k = k_mont.from_mont();
    *out = xyzz_mul(*p, k.v);
  }
}
struct FinalizeSumIn { const G1XYZZ *a, *sa, *rb1, *l, *h; const G2XYZZ* b2; };
__global__ void __launch_bounds__(96) finalize_sum_kernel(FinalizeSumIn in, ctg_proof* out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane != 0) return;
  if (warp == 0) {
    G1XYZZ c = xyzz_add(*in.sa, *in.rb1);
    c = xyzz_add(c, *in.l);
    c = xyzz_add(c, *in.h);
    G1Affine ca = xyzz_to_affine(c);
    *reinterpret_cast<G1Affine*>(&out->c) = ca;
  } else if (warp == 1) {
    G1Affine aa = xyzz_to_affine(*in.a);
    *reinterpret_cast<G1Affine*>(&out->a) = aa;
  } else {
    G2Affine b = xyzz_to_affine(*in.b2);
    *reinterpret_cast<G2Affine*>(&out->b) = b;
  }
}

// one block per proof of a batch: results laid out a[p], b1[p], l[p], h[p], b2[p]; (r, s) at rs[2p], rs[2p + 1]
struct FinalizeBatchIn { const G1XYZZ *a, *b1, *l, *h; const G2XYZZ* b2; const Fr* rs; };
__global__ void __launch_bounds__(96) finalize_batch_kernel(FinalizeBatchIn in, ctg_proof* out) {
  __shared__ G1XYZZ sh[2];
  const uint32_t p = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    if (warp == 0) {
      Fr k = in.rs[2 * p + 1].from_mont();
      sh[0] = xyzz_mul(in.a[p], k.v);
    } else if (warp == 1) {
      Fr k = in.rs[2 * p].from_mont();
      sh[1] = xyzz_mul(in.b1[p], k.v);
    } else {
      G2Affine b = xyzz_to_affine(in.b2[p]);
      *reinterpret_cast<G2Affine*>(&out[p].b) = b;
    }
  }
  __syncthreads();
  if (lane == 0) {
    if (warp == 0) {
      G1XYZZ c = xyzz_add(sh[0], sh[1]);
      c = xyzz_add(c, in.l[p]);
      c = xyzz_add(c, in.h[p]);
      G1Affine ca = xyzz_to_affine(c);
      *reinterpret_cast<G1Affine*>(&out[p].c) = ca;
    } else if (warp == 1) {
      G1Affine aa = xyzz_to_affine(in.a[p]);
      *reinterpret_cast<G1Affine*>(&out[p].a) = aa;
    }
  }
}

// ---- host: key and R1CS residency -------------------------------------------------------------
static void free_bases(Bases& b) { if (b.pts) cudaFree(b.pts); b.pts = nullptr; }

template <class A>
static int32_t make_bases(ctg_ctx* ctx, Bases& out, int group, uint64_t n, int copies, int c_bits, bool batch_throughput = false) {
  out.group = group;
  out.n = n;
  out.plan = msm_make_plan(n, copies, c_bits, sizeof(A), batch_throughput);
  out.pts = nullptr;
  if (n * (uint64_t)out.plan.copies >= (1ull << 31)) return fail(ctx, CTG_ERR_INVALID, "pk_upload: query length * precomputed copies must be < 2^31 (ask for fewer copies)");
  size_t bytes = (size_t)out.plan.copies * (n ? n : 1) * sizeof(A);
  cudaError_t e = cudaMalloc(&out.pts, bytes);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, CTG_ERR_OOM, "bases: cudaMalloc failed (try fewer precomputed copies)"); }
  CTG_CUDA(ctx, cudaMemsetAsync(out.pts, 0, bytes, ctx->stream));
  return CTG_OK;
}
static int32_t finish_bases(ctg_ctx* ctx, Bases& b) { if (b.n == 0) return CTG_OK;  /* an empty query (e.g. l_query of a circuit without witnesses) */ return b.group == 1 ? g1_precompute(ctx, b, ctx->stream) : g2_precompute(ctx, b, ctx->stream); }

// positions [lo, lo + cnt) of (query ++ extras[0..3)) -> dst; extras may hold nullptr (= point at infinity, dst is pre-zeroed)
template <class A>
static cudaError_t upload_ext_range(A* dst, const A* query, uint64_t m, const A* const extras[3], uint64_t lo, uint64_t cnt, cudaStream_t st) {
  const uint64_t hi = lo + cnt;
  if (lo < m) {
    const uint64_t q_hi = hi < m ? hi : m;
    cudaError_t e = cudaMemcpyAsync(dst, query + lo, (q_hi - lo) * sizeof(A), cudaMemcpyDefault, st);
    if (e != cudaSuccess) return e;
  }
  for (int k = 0; k < 3; k++) {
    const uint64_t pos = m + k;
    if (pos >= lo && pos < hi && extras[k]) {
      cudaError_t e = cudaMemcpyAsync(dst + (pos - lo), extras[k], sizeof(A), cudaMemcpyDefault, st);
      if (e != cudaSuccess) return e;
    }
  }
  return cudaSuccess;
}

// positions [total * lo_num / den, total * hi_num / den) -- equal shards are (index, index + 1, count); weighted shards pass
// cumulative weights, so that ranks which also run NTT chains can take fewer points
static void shard_range(uint64_t total, uint64_t lo_num, uint64_t hi_num, uint64_t den, uint64_t* lo, uint64_t* cnt) {
  const uint64_t a = (uint64_t)((unsigned __int128)total * lo_num / den), b = (uint64_t)((unsigned __int128)total * hi_num / den);
  *lo = a;
  *cnt = b - a;
}

static int32_t pk_upload_impl(ctg_ctx* ctx, const ctg_pk_view* v, uint64_t ni, uint64_t nc, int copies, int c_bits, uint64_t lo_num, uint64_t hi_num, uint64_t den, bool tp, ctg_pk** out) {
  if (!v || !out) return fail(ctx, CTG_ERR_INVALID, "pk_upload: null argument");
  if (den < 1 || lo_num > hi_num || hi_num > den) return fail(ctx, CTG_ERR_INVALID, "pk_upload: bad shard range (need 0 <= lo <= hi <= den, den >= 1)");
  const int shard_index = lo_num == 0 ? 0 : 1, shard_count = (lo_num == 0 && hi_num == den) ? 1 : 2;   // only "whole key or not" matters downstream
  const uint64_t m = v->a_len;
  if (m < ni || m == 0 || v->b_g1_len != m || v->b_g2_len != m) return fail(ctx, CTG_ERR_INVALID, "pk_upload: a/b query lengths must equal num_instance + num_witness");
  const uint64_t nw = m - ni;
  if (v->l_len != nw) return fail(ctx, CTG_ERR_INVALID, "pk_upload: l_query length must equal num_witness");
  uint64_t n = 1; int log_n = 0;
  while (n < nc + ni) { n <<= 1; log_n++; }
  if (v->h_len + 1 != n) return fail(ctx, CTG_ERR_INVALID, "pk_upload: h_query length must be domain size - 1");
  ctg_pk* pk = new ctg_pk();
  pk->m = m; pk->ni = ni; pk->nw = nw; pk->n = n; pk->log_n = log_n;
  pk->shard_index = shard_index; pk->shard_count = shard_count;
  shard_range(m + 3, lo_num, hi_num, den, &pk->ext_lo, &pk->ext_cnt);
  shard_range(nw, lo_num, hi_num, den, &pk->l_lo, &pk->l_cnt);
  shard_range(n, lo_num, hi_num, den, &pk->h_lo, &pk->h_cnt);
  cudaStream_t st = ctx->stream;
  auto cleanup = [&](int32_t rc) { free_bases(pk->a_ext); free_bases(pk->b1_ext); free_bases(pk->b2_ext); free_bases(pk->l); free_bases(pk->h_ext); delete pk; return rc; };
  memset(&pk->a_ext, 0, sizeof(Bases)); memset(&pk->b1_ext, 0, sizeof(Bases)); memset(&pk->b2_ext, 0, sizeof(Bases));
  memset(&pk->l, 0, sizeof(Bases)); memset(&pk->h_ext, 0, sizeof(Bases));
  int32_t rc;
  if ((rc = make_bases<G1Affine>(ctx, pk->a_ext, 1, pk->ext_cnt, copies, c_bits, tp)) != CTG_OK) return cleanup(rc);
  if ((rc = make_bases<G1Affine>(ctx, pk->b1_ext, 1, pk->ext_cnt, copies, c_bits, tp)) != CTG_OK) return cleanup(rc);
  if ((rc = make_bases<G2Affine>(ctx, pk->b2_ext, 2, pk->ext_cnt, copies, c_bits, tp)) != CTG_OK) return cleanup(rc);
  if ((rc = make_bases<G1Affine>(ctx, pk->l, 1, pk->l_cnt, copies, c_bits, tp)) != CTG_OK) return cleanup(rc);
  if ((rc = make_bases<G1Affine>(ctx, pk->h_ext, 1, pk->h_cnt, copies, c_bits, tp)) != CTG_OK) return cleanup(rc);
#define PK_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { ctx->last_error = std::string(#call) + ": " + cudaGetErrorString(e__); return cleanup(CTG_ERR_CUDA); } } while (0)
  const G1Affine* a_extra[3] = {(const G1Affine*)v->alpha_g1, (const G1Affine*)v->delta_g1, nullptr};
  const G1Affine* b1_extra[3] = {(const G1Affine*)v->beta_g1, nullptr, (const G1Affine*)v->delta_g1};
  const G2Affine* b2_extra[3] = {(const G2Affine*)v->beta_g2, nullptr, (const G2Affine*)v->delta_g2};
  PK_CUDA(upload_ext_range((G1Affine*)pk->a_ext.pts, (const G1Affine*)v->a_query, m, a_extra, pk->ext_lo, pk->ext_cnt, st));
  PK_CUDA(upload_ext_range((G1Affine*)pk->b1_ext.pts, (const G1Affine*)v->b_g1_query, m, b1_extra, pk->ext_lo, pk->ext_cnt, st));
  PK_CUDA(upload_ext_range((G2Affine*)pk->b2_ext.pts, (const G2Affine*)v->b_g2_query, m, b2_extra, pk->ext_lo, pk->ext_cnt, st));
  if (pk->l_cnt) PK_CUDA(cudaMemcpyAsync(pk->l.pts, (const G1Affine*)v->l_query + pk->l_lo, pk->l_cnt * sizeof(G1Affine), cudaMemcpyDefault, st));
  // h_query: stage natural order in scratch, gather this shard's bit-reversed positions; position n-1 holds delta_g1
  if (n > 1 && pk->h_cnt) {
    if ((rc = ensure_scratch(ctx, (n - 1) * sizeof(G1Affine))) != CTG_OK) return cleanup(rc);
    PK_CUDA(cudaMemcpyAsync(ctx->scratch, v->h_query, (n - 1) * sizeof(G1Affine), cudaMemcpyDefault, st));
    bitrev_gather_g1_kernel<<<(unsigned)((pk->h_cnt + 255) / 256), 256, 0, st>>>((const G1Affine*)ctx->scratch, (G1Affine*)pk->h_ext.pts, n - 1, log_n, pk->h_lo, pk->h_cnt);
    ctx->launches++;
  }
  if (n - 1 >= pk->h_lo && n - 1 < pk->h_lo + pk->h_cnt)
    PK_CUDA(cudaMemcpyAsync((G1Affine*)pk->h_ext.pts + (n - 1 - pk->h_lo), v->delta_g1, sizeof(G1Affine), cudaMemcpyDefault, st));
  for (Bases* b : {&pk->a_ext, &pk->b1_ext, &pk->b2_ext, &pk->l, &pk->h_ext})
    if ((rc = finish_bases(ctx, *b)) != CTG_OK) return cleanup(rc);
  PK_CUDA(cudaStreamSynchronize(st));
#undef PK_CUDA
  *out = pk;
  return CTG_OK;
}

void pk_free(ctg_pk* pk);
int32_t pk_upload(ctg_ctx* ctx, const ctg_pk_view* v, uint64_t ni, uint64_t nc, int copies, int c_bits, uint64_t lo_num, uint64_t hi_num, uint64_t den, ctg_pk** out) {
  CTG_TRY(pk_upload_impl(ctx, v, ni, nc, copies, c_bits, lo_num, hi_num, den, false, out));
  ctg_pk* pk = *out;
  // a whole small key with library-chosen windows also gets its batch twin (see ctg_pk::batch_twin); CTG_BATCH_TWIN=0 turns it off
  static int twin = -1;
  if (twin < 0) { const char* e = getenv("CTG_BATCH_TWIN"); twin = !(e && *e == '0'); }
  if (twin && pk->shard_count == 1 && c_bits == 0 && copies == 0 && pk->log_n >= 1 && pk->log_n <= 15) {
    const int32_t rc = pk_upload_impl(ctx, v, ni, nc, 0, 0, lo_num, hi_num, den, true, &pk->batch_twin);
    if (rc != CTG_OK) { pk->batch_twin = nullptr; pk_free(pk); *out = nullptr; return rc; }
  }
  return CTG_OK;
}

void pk_free(ctg_pk* pk) {
  if (!pk) return;
  pk_free(pk->batch_twin);
  free_bases(pk->a_ext); free_bases(pk->b1_ext); free_bases(pk->b2_ext); free_bases(pk->l); free_bases(pk->h_ext);
  delete pk;
}

int32_t r1cs_upload(ctg_ctx* ctx, uint64_t ni, uint64_t nw, uint64_t nc, const ctg_csr* const mats[3], ctg_r1cs** out) {
  if (ni == 0) return fail(ctx, CTG_ERR_INVALID, "r1cs_upload: num_instance includes the constant 1 and must be >= 1");
  ctg_r1cs* r = new ctg_r1cs();
  for (int k = 0; k < 3; k++) { r->row_ptr[k] = nullptr; r->col[k] = nullptr; r->val[k] = nullptr; r->nnz[k] = 0; }
  r->ni = ni; r->nw = nw; r->nc = nc;
  uint64_t n = 1; int log_n = 0;
  while (n < nc + ni) { n <<= 1; log_n++; }
  r->n = n; r->log_n = log_n;
  auto cleanup = [&](int32_t rc) { for (int k = 0; k < 3; k++) { cudaFree(r->row_ptr[k]); cudaFree(r->col[k]); cudaFree(r->val[k]); } delete r; return rc; };
  if (log_n > 28) { ctx->last_error = "r1cs_upload: domain larger than 2^28"; return cleanup(CTG_ERR_INVALID); }
  for (int k = 0; k < 3; k++) {
    const ctg_csr* m = mats[k];
    if (!m || !m->row_ptr) { ctx->last_error = "r1cs_upload: null matrix"; return cleanup(CTG_ERR_INVALID); }
    // the SpMV kernel trusts the structure: row_ptr[0] = 0, non-decreasing, row_ptr[nc] = nnz entries in col / val
    if (m->row_ptr[0] != 0) { ctx->last_error = "r1cs_upload: row_ptr[0] must be 0"; return cleanup(CTG_ERR_INVALID); }
    for (uint64_t i = 0; i < nc; i++)
      if (m->row_ptr[i + 1] < m->row_ptr[i]) { ctx->last_error = "r1cs_upload: row_ptr must be non-decreasing"; return cleanup(CTG_ERR_INVALID); }
    const uint64_t nnz = m->row_ptr[nc];
    if (nnz && (!m->col || !m->val)) { ctx->last_error = "r1cs_upload: null col / val arrays"; return cleanup(CTG_ERR_INVALID); }
    r->nnz[k] = nnz;
    for (uint64_t i = 0; i < nnz; i++)
      if (m->col[i] >= ni + nw) { ctx->last_error = "r1cs_upload: column index out of range"; return cleanup(CTG_ERR_INVALID); }
#define R_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { ctx->last_error = std::string(#call) + ": " + cudaGetErrorString(e__); return cleanup(CTG_ERR_CUDA); } } while (0)
    R_CUDA(cudaMalloc(&r->row_ptr[k], (nc + 1) * sizeof(uint64_t)));
    R_CUDA(cudaMalloc(&r->col[k], (nnz ? nnz : 1) * sizeof(uint32_t)));
    R_CUDA(cudaMalloc(&r->val[k], (nnz ? nnz : 1) * sizeof(Fr)));
    R_CUDA(cudaMemcpyAsync(r->row_ptr[k], m->row_ptr, (nc + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
    if (nnz) {
      R_CUDA(cudaMemcpyAsync(r->col[k], m->col, nnz * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
      R_CUDA(cudaMemcpyAsync(r->val[k], m->val, nnz * sizeof(Fr), cudaMemcpyHostToDevice, ctx->stream));
    }
  }
  R_CUDA(cudaStreamSynchronize(ctx->stream));
#undef R_CUDA
  *out = r;
  return CTG_OK;
}

void r1cs_free(ctg_r1cs* r) {
  if (!r) return;
  for (int k = 0; k < 3; k++) { cudaFree(r->row_ptr[k]); cudaFree(r->col[k]); cudaFree(r->val[k]); }
  delete r;
}

// ---- host: witness map ------------------------------------------------------------------------
// d_z: m (+3) Montgomery elements on the device.  d_abc: 3 * n scratch.  Leaves h (bit-reversed,
// Montgomery) in d_abc[0..n).
// chain k (0 = a, 1 = b, 2 = c): SpMV, then ifft + coset_fft fused -> coset evaluations in v (n elements, natural order)
int32_t witness_chain_dev(ctg_ctx* ctx, const ctg_r1cs* r, const Fr* d_z, Fr* v, int k, cudaStream_t st) {
  const uint64_t n = r->n;
  spmv_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(r->row_ptr[k], r->col[k], r->val[k], d_z, v, r->nc, r->ni, n, k == 0, 0);
  CTG_KERNEL_CHECK(ctx);
  return ntt_ifft_then_coset_fft(ctx, v, r->log_n, st);
}
// (a b - c) / Z on the coset, then coset_ifft: h (bit-reversed, Montgomery) in d_abc[0..n)
int32_t witness_finish_dev(ctg_ctx* ctx, const ctg_r1cs* r, Fr* d_abc, cudaStream_t st) {
  const uint64_t n = r->n;
  NttDomain* dom;
  CTG_TRY(ntt_get_domain(ctx, r->log_n, &dom));
  pointwise_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(d_abc, d_abc + n, d_abc + 2 * n, dom->zinv, n);
  CTG_KERNEL_CHECK(ctx);
  return ntt_coset_ifft_bitrev(ctx, d_abc, r->log_n, st);
}
static int32_t witness_map_dev(ctg_ctx* ctx, const ctg_r1cs* r, const Fr* d_z, Fr* d_abc, cudaStream_t st) {
  NttDomain* dom;
  CTG_TRY(ntt_get_domain(ctx, r->log_n, &dom));  // tables are built (and synchronised) before any chain is enqueued
  for (int k = 0; k < 3; k++) CTG_TRY(witness_chain_dev(ctx, r, d_z, d_abc + (uint64_t)k * r->n, k, st));
  return witness_finish_dev(ctx, r, d_abc, st);
}

int32_t witness_map_host(ctg_ctx* ctx, const ctg_r1cs* r, const ctg_fr* z, ctg_fr* h_out) {
  const uint64_t n = r->n, m = r->ni + r->nw;
  cudaStream_t st = ctx->stream;
  CTG_TRY(ensure_scratch(ctx, (m + 4 * n) * sizeof(Fr)));
  Fr* d_z = (Fr*)ctx->scratch;
  Fr* d_abc = d_z + m;
  CTG_CUDA(ctx, cudaMemcpyAsync(d_z, z, m * sizeof(Fr), cudaMemcpyHostToDevice, st));
  CTG_TRY(witness_map_dev(ctx, r, d_z, d_abc, st));
  Fr* d_nat = d_abc + 3 * n;
  bitrev_gather_fr_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_abc, d_nat, r->log_n);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaMemcpyAsync(h_out, d_nat, n * sizeof(Fr), cudaMemcpyDeviceToHost, st));
  CTG_CUDA(ctx, cudaStreamSynchronize(st));
  return CTG_OK;
}

// ---- host: prover lanes -----------------------------------------------------------------------
int32_t lane_init(ctg_ctx* ctx, int idx) {
  ProveLane& L = ctx->lanes[idx];
  if (L.ready) { if (idx == 0) L.st[0] = ctx->stream; return CTG_OK; }
  int lo_pri = 0, hi_pri = 0;
  CTG_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));   // numerically lower = more urgent
  for (int i = 0; i < 5; i++) {
    if (idx == 0 && i == 0) { L.st[0] = ctx->stream; continue; }
    CTG_CUDA(ctx, cudaStreamCreateWithFlags(&L.st[i], cudaStreamNonBlocking));
  }
  CTG_CUDA(ctx, cudaStreamCreateWithPriority(&L.st_hi, cudaStreamNonBlocking, hi_pri));
  CTG_CUDA(ctx, cudaEventCreateWithFlags(&L.ev_z, cudaEventDisableTiming));
  CTG_CUDA(ctx, cudaEventCreateWithFlags(&L.ev_zfull, cudaEventDisableTiming));
  CTG_CUDA(ctx, cudaEventCreateWithFlags(&L.ev_wm, cudaEventDisableTiming));
  for (auto& e : L.ev_done) CTG_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  for (auto& e : L.ev_t) CTG_CUDA(ctx, cudaEventCreate(&e));
  L.ready = true;
  return CTG_OK;
}
void lanes_destroy(ctg_ctx* ctx) {
  for (int idx = 0; idx < kMaxLanes; idx++) {
    ProveLane& L = ctx->lanes[idx];
    if (!L.ready) continue;
    for (int i = 0; i < 5; i++) {
      if (L.st[i]) cudaStreamSynchronize(L.st[i]);
      if (!(idx == 0 && i == 0) && L.st[i]) cudaStreamDestroy(L.st[i]);
      msm_ws_free(L.ws[i]);
    }
    if (L.scratch) { cudaMemset(L.scratch, 0, L.scratch_bytes); cudaFree(L.scratch); }   // the assignment z and (r, s) lived here
    if (L.st_hi) { cudaStreamSynchronize(L.st_hi); cudaStreamDestroy(L.st_hi); }
    cudaEventDestroy(L.ev_zfull); cudaEventDestroy(L.ev_wm);
    cudaEventDestroy(L.ev_z);
    for (auto& e : L.ev_done) cudaEventDestroy(e);
    for (auto& e : L.ev_t) cudaEventDestroy(e);
    L = ProveLane();
  }
  if (ctx->pinned_proofs) cudaFreeHost(ctx->pinned_proofs);
  ctx->pinned_proofs = nullptr;
}
int32_t lane_scratch(ctg_ctx* ctx, ProveLane& L, size_t bytes) {
  if (bytes <= L.scratch_bytes) return CTG_OK;
  if (L.scratch) {
    for (int i = 0; i < 5; i++) CTG_CUDA(ctx, cudaStreamSynchronize(L.st[i]));
    CTG_CUDA(ctx, cudaStreamSynchronize(L.st_hi));
    cudaMemset(L.scratch, 0, L.scratch_bytes);
    cudaFree(L.scratch);
    L.scratch = nullptr; L.scratch_bytes = 0;
  }
  const size_t want = bytes + bytes / 8 + 4096;
  if (cudaMalloc(&L.scratch, want) != cudaSuccess) { cudaGetLastError(); return fail(ctx, CTG_ERR_OOM, "out of device memory for prover scratch"); }
  L.scratch_bytes = want;
  return CTG_OK;
}

// two tables whose MSMs over the same scalars produce the same sorted entry list
static bool same_plan(const Bases& a, const Bases& b) {
  return a.plan.c == b.plan.c && a.plan.nwin == b.plan.nwin && a.plan.copies == b.plan.copies && a.plan.nsets == b.plan.nsets &&
         a.n == b.n && (a.stride ? a.stride : a.n) == (b.stride ? b.stride : b.n);
}

// scratch of a lane: z_ext (m+3) | a,b,c (3n, unless the caller owns that buffer) | results: 4 G1 XYZZ + 1 G2 XYZZ | proof
int32_t lane_bufs(ctg_ctx* ctx, ProveLane& L, uint64_t m, uint64_t n, Fr* d_abc_ext, LaneBufs* out) {
  const size_t fr_elems = m + 3 + (d_abc_ext ? 0 : 3 * n);
  const size_t bytes = fr_elems * sizeof(Fr) + 6 * sizeof(G1XYZZ) + sizeof(G2XYZZ) + sizeof(ctg_proof) + 512;
  CTG_TRY(lane_scratch(ctx, L, bytes));
  out->d_z = (Fr*)L.scratch;
  Fr* after = out->d_z + (m + 3);
  out->d_abc = d_abc_ext ? d_abc_ext : after;
  char* tail = (char*)(d_abc_ext ? after : after + 3 * n);
  tail = (char*)(((uintptr_t)tail + 63) & ~(uintptr_t)63);
  out->d_g1 = (G1XYZZ*)tail;
  out->d_g2 = (G2XYZZ*)(out->d_g1 + 4);
  out->d_proof = (ctg_proof*)(out->d_g2 + 1);
  out->d_early = (G1XYZZ*)(((uintptr_t)(out->d_proof + 1) + 63) & ~(uintptr_t)63);   // s * A, r * B1
  return CTG_OK;
}

// Enqueues one proof (or one shard's partial sums) on a lane; no host synchronisation.
//   d_proof_out != nullptr : full proof (shard_count must be 1), written to device memory d_proof_out
//   d_partials_out != nullptr : this shard's partial sums [A, B1, L, H] (4 G1 XYZZ) then B2 (G2 XYZZ), 1152 bytes
static int32_t prove_enqueue(ctg_ctx* ctx, int lane_idx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const Fr& r_m, const Fr& s_m,
                             ctg_proof* host_proof_out, void* host_partials_out, bool timed,
                             int phase = 0 /* 0 = whole proof, 1 = up to the coset evaluations, 2 = the rest */, uint32_t chain_mask = 7,
                             Fr* d_abc_ext = nullptr /* caller-owned 3n buffer for the a, b, c chains (multi-GPU exchange) */,
                             bool early_mul = true /* s * A and r * B1 on the streams of their MSMs (not for the partial-sum ABI) */) {
  CTG_TRY(lane_init(ctx, lane_idx));
  ProveLane& L = ctx->lanes[lane_idx];
  const uint64_t n = r->n, m = pk->m;
  LaneBufs B;
  CTG_TRY(lane_bufs(ctx, L, m, n, d_abc_ext, &B));
  Fr* d_z = B.d_z;
  Fr* d_abc = B.d_abc;
  G1XYZZ* d_g1 = B.d_g1;           // [0]=A [1]=B1 [2]=L [3]=H
  G2XYZZ* d_g2 = B.d_g2;
  ctg_proof* d_proof = B.d_proof;
  cudaStream_t s0 = L.st[0];
  cudaStream_t sw = L.st_hi;   // witness map + H MSM
  int ne = 0;
  auto mark = [&](cudaStream_t st) { if (timed && ne < 16) cudaEventRecord(L.ev_t[ne], st); ne++; };
  // An alternative schedule, OFF by default: the witness map first, alone, and all five MSMs behind it (CTG_WM_FIRST_MAX_LOG = largest
  // log2 domain it applies to).  Idea: beside four MSMs' long-lived CTAs the short witness-map kernels wait for SM slots however high
  // their stream priority (2^16: 3.2 ms instead of < 1 ms alone).  Measured on B200 it loses all the same -- 2^12 8.2 -> 8.5 ms, 2^16
  // 5.9 -> 6.3 ms, 2^18 14.3 -> 14.4 ms, 2^20 44.5 -> 44.0 ms: the critical path of a small proof is the G2 MSM, which then starts later.
  static int wm_first_max_log = -1;
  if (wm_first_max_log < 0) { const char* e = getenv("CTG_WM_FIRST_MAX_LOG"); wm_first_max_log = e ? atoi(e) : 0; }
  const bool wm_first = phase == 0 && !host_partials_out && pk->shard_count == 1 && r->log_n <= wm_first_max_log;
  // assignment-only MSMs on their own streams
  struct Job { int stream; const Bases* b; const Fr* sc; uint64_t cnt; int g1_slot; };
  const Job jobs[4] = {{1, &pk->l, d_z + pk->ni + pk->l_lo, pk->l_cnt, 2},
                       {2, &pk->a_ext, d_z + pk->ext_lo, pk->ext_cnt, 0},
                       {3, &pk->b1_ext, d_z + pk->ext_lo, pk->ext_cnt, 1},
                       {4, &pk->b2_ext, d_z + pk->ext_lo, pk->ext_cnt, -1}};
  // Queue order = L, A, B1, B2.  (Queueing the G2 MSM first on a raised-priority stream, so that its latency-bound tail runs under
  // the G1 accumulate kernels, was measured: 2^16 7.8 -> 7.5 ms, but 2^22 154 -> 159 ms and the fused batch 2 045 -> 1 441 proofs/s:
  // the G2 kernels then delay the sorts of the other MSMs.  Timing events keep their slots: 2 + 2 * index in `jobs`.)
  const int order[4] = {0, 1, 2, 3};
  auto enqueue_jobs = [&](cudaEvent_t gate) -> int32_t {
  for (int oi = 0; oi < 4; oi++) {
    const Job& j = jobs[order[oi]];
    ne = 2 + 2 * order[oi];
    cudaStream_t st = L.st[j.stream];
    CTG_CUDA(ctx, cudaStreamWaitEvent(st, L.ev_z, 0));
    if (gate) CTG_CUDA(ctx, cudaStreamWaitEvent(st, gate, 0));
    mark(st);                                                                 // 2,4,6,8
    if (j.g1_slot >= 0) {
      // A and B1 multiply the same scalars under the same plan: B1 reuses A's sorted entry list (one digit sort fewer per proof)
      const MsmWorkspace* sorted = nullptr;
      if (j.stream == 3 && j.cnt > 0 && same_plan(pk->a_ext, pk->b1_ext)) { sorted = &L.ws[2]; CTG_CUDA(ctx, cudaStreamWaitEvent(st, L.ws[2].ev_sorted, 0)); }
      CTG_TRY(msm_g1_run(ctx, L.ws[j.stream], *j.b, (const uint32_t*)j.sc, j.cnt, true, d_g1 + j.g1_slot, st, nullptr, 1, 0, sorted));
    } else CTG_TRY(msm_g2_run(ctx, L.ws[j.stream], *j.b, (const uint32_t*)j.sc, j.cnt, true, d_g2, st, nullptr));
    mark(st);                                                                 // 3,5,7,9
    if (early_mul && !host_partials_out && (j.g1_slot == 0 || j.g1_slot == 1)) {   // s * A / r * B1 under the remaining MSMs
      scalar_mul_g1_kernel<<<1, 32, 0, st>>>(d_g1 + j.g1_slot, j.g1_slot == 0 ? s_m : r_m, B.d_early + j.g1_slot);
      CTG_KERNEL_CHECK(ctx);
    }
    CTG_CUDA(ctx, cudaEventRecord(L.ev_done[j.stream], st));
  }
  return CTG_OK;
  };
  if (phase != 2) {
  mark(s0);                                                                   // 0
  // Upload order: first the part of z this key's A/B1/B2/L MSMs read (for a point-range shard, 1/N of it), so that they start
  // at once; then the rest, which only the witness map reads -- and only if this rank computes a chain at all.
  uint64_t s_lo = pk->ext_lo < m ? pk->ext_lo : m, s_hi = pk->ext_lo + pk->ext_cnt < m ? pk->ext_lo + pk->ext_cnt : m;
  if (pk->l_cnt) {
    if (pk->ni + pk->l_lo < s_lo) s_lo = pk->ni + pk->l_lo;
    if (pk->ni + pk->l_lo + pk->l_cnt > s_hi) s_hi = pk->ni + pk->l_lo + pk->l_cnt;
  }
  if (s_hi < s_lo) s_hi = s_lo;
  const ctg_fr* zh = z;
  if (s_hi > s_lo) CTG_CUDA(ctx, cudaMemcpyAsync(d_z + s_lo, zh + s_lo, (s_hi - s_lo) * sizeof(Fr), cudaMemcpyHostToDevice, s0));
  set_tail_kernel<<<1, 32, 0, s0>>>(d_z, m, r_m, s_m);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaEventRecord(L.ev_z, s0));
  if (phase != 1 || chain_mask != 0) {
    if (s_lo > 0) CTG_CUDA(ctx, cudaMemcpyAsync(d_z, zh, s_lo * sizeof(Fr), cudaMemcpyHostToDevice, s0));
    if (s_hi < m) CTG_CUDA(ctx, cudaMemcpyAsync(d_z + s_hi, zh + s_hi, (m - s_hi) * sizeof(Fr), cudaMemcpyHostToDevice, s0));
  }
  CTG_CUDA(ctx, cudaEventRecord(L.ev_zfull, s0));
  mark(s0);                                                                   // 1
  if (!wm_first) CTG_TRY(enqueue_jobs(nullptr));
  ne = 10;
  // witness map on the lane's main stream: the coset evaluations of a, b, c (the chains this rank owns) ...
  if (phase != 1 || chain_mask != 0) {   // tables are built (and synchronised) before any chain is enqueued
    NttDomain* dom;
    CTG_TRY(ntt_get_domain(ctx, r->log_n, &dom));
  }
  CTG_CUDA(ctx, cudaStreamWaitEvent(sw, L.ev_zfull, 0));
  for (int k = 0; k < 3; k++)
    if (chain_mask & (1u << k)) CTG_TRY(witness_chain_dev(ctx, r, d_z, d_abc + (uint64_t)k * n, k, sw));
  if (phase == 1) {   // the caller synchronises s0 and then exchanges the chains
    CTG_CUDA(ctx, cudaEventRecord(L.ev_wm, sw));
    CTG_CUDA(ctx, cudaStreamWaitEvent(s0, L.ev_wm, 0));
    return CTG_OK;
  }
  } else {
    ne = 10;
    CTG_CUDA(ctx, cudaEventRecord(L.ev_zfull, s0));    // the exchanged chains were written on (or synchronised with) s0
    CTG_CUDA(ctx, cudaStreamWaitEvent(sw, L.ev_zfull, 0));
  }
  // ... then (a b - c)/Z, the coset_ifft and the H MSM
  CTG_TRY(witness_finish_dev(ctx, r, d_abc, sw));
  set_h_tail_kernel<<<1, 32, 0, sw>>>(d_abc, n, r_m, s_m);
  CTG_KERNEL_CHECK(ctx);
  mark(sw);                                                                   // 10
  if (wm_first) {   // the four assignment MSMs start behind the witness map (see above); ev_t[15] is free of timing duty
    CTG_CUDA(ctx, cudaEventRecord(L.ev_t[15], sw));
    CTG_TRY(enqueue_jobs(L.ev_t[15]));
    ne = 11;
  }
  CTG_TRY(msm_g1_run(ctx, L.ws[0], pk->h_ext, (const uint32_t*)(d_abc + pk->h_lo), pk->h_cnt, true, d_g1 + 3, sw, nullptr));
  mark(sw);                                                                   // 11
  CTG_CUDA(ctx, cudaEventRecord(L.ev_wm, sw));
  CTG_CUDA(ctx, cudaStreamWaitEvent(s0, L.ev_wm, 0));
  for (int i = 1; i < 5; i++) CTG_CUDA(ctx, cudaStreamWaitEvent(s0, L.ev_done[i], 0));
  if (host_partials_out) {
    CTG_CUDA(ctx, cudaMemcpyAsync(host_partials_out, d_g1, 4 * sizeof(G1XYZZ) + sizeof(G2XYZZ), cudaMemcpyDeviceToHost, s0));
  } else {
    if (phase == 0) {
      FinalizeSumIn fin{d_g1 + 0, B.d_early + 0, B.d_early + 1, d_g1 + 2, d_g1 + 3, d_g2};
      finalize_sum_kernel<<<1, 96, 0, s0>>>(fin, d_proof);
    } else {   // the MSMs of phase 1 were enqueued by another call: assemble from the sums
      FinalizeIn fin{d_g1 + 0, d_g1 + 1, d_g1 + 2, d_g1 + 3, d_g2, r_m, s_m};
      finalize_kernel<<<1, 96, 0, s0>>>(fin, d_proof);
    }
    CTG_KERNEL_CHECK(ctx);
    CTG_CUDA(ctx, cudaMemcpyAsync(host_proof_out, d_proof, sizeof(ctg_proof), cudaMemcpyDeviceToHost, s0));
  }
  mark(s0);                                                                   // 12
  return CTG_OK;
}

static void lane_timings(ProveLane& L, ctg_timings* tm) {
  memset(tm, 0, sizeof(*tm));
  auto el = [&](int a, int b) { float ms = 0; cudaEventElapsedTime(&ms, L.ev_t[a], L.ev_t[b]); return ms; };
  tm->h2d_ms = el(0, 1);
  tm->msm_l_ms = el(2, 3); tm->msm_a_ms = el(4, 5); tm->msm_b_g1_ms = el(6, 7); tm->msm_b_g2_ms = el(8, 9);
  tm->witness_map_ms = el(1, 10);
  tm->msm_h_ms = el(10, 11);
  tm->finalize_ms = el(11, 12);   // includes waiting for the other streams
  tm->total_ms = el(0, 12);
}

static int32_t check_prove_args(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r) {
  if (pk->ni != r->ni || pk->nw != r->nw || pk->n != r->n) return fail(ctx, CTG_ERR_INVALID, "prove: proving key and R1CS shapes differ");
  return CTG_OK;
}

int32_t prove(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const ctg_fr* rr, const ctg_fr* ss, ctg_proof* out, ctg_timings* tm) {
  if (!pk || !r || !z || !rr || !ss || !out) return fail(ctx, CTG_ERR_INVALID, "prove: null argument");
  CTG_TRY(check_prove_args(ctx, pk, r));
  if (pk->shard_count != 1) return fail(ctx, CTG_ERR_STATE, "prove: the key is a shard; use ctg_prove_partial + ctg_prove_combine");
  Fr r_m, s_m;
  memcpy(&r_m, rr, sizeof(Fr));
  memcpy(&s_m, ss, sizeof(Fr));
  CTG_TRY(prove_enqueue(ctx, 0, pk, r, z, r_m, s_m, out, nullptr, tm != nullptr));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->lanes[0].st[0]));
  if (tm) lane_timings(ctx->lanes[0], tm);
  return CTG_OK;
}

// ---- batched prover: many proofs of ONE circuit under ONE key (the reference's production shape: a key-encryption proof per
// buyer request, lib/src/seller.rs:136-159, and per-chunk proofs "computed in parallel", docs/protocol.md:56) ------------------
// A chunk of up to kBatchMax proofs goes through the pipeline TOGETHER: one SpMV launch per matrix and one launch per NTT pass
// for all of them (blockIdx.y = proof), and each of the five MSMs is ONE digit sort + ONE accumulate pass + ONE reduction over
// the chunk's scalar vectors against the shared base table (bucket-set group = proof; msm.cuh run(batch)).  The latency-bound
// tails that dominate a lone small proof (bucket reduction, inversions) become wide, throughput-bound launches, and the launch
// count per proof drops from ~100 to ~170 / chunk.  A and B1 share one digit sort (same scalars, same plan).
static constexpr uint32_t kBatchMax = 256;
static uint32_t batch_chunk(const ctg_pk* pk, const ctg_r1cs* r, uint32_t n_proofs) {
  uint64_t c = std::min<uint64_t>(n_proofs, kBatchMax);
  for (const Bases* b : {&pk->a_ext, &pk->b1_ext, &pk->b2_ext, &pk->l, &pk->h_ext}) {
    if (b->n == 0) continue;
    const uint64_t per_nb = (uint64_t)b->plan.nsets * b->plan.nbuckets_per_set, per_ne = b->n * (uint64_t)b->plan.nwin;
    c = std::min<uint64_t>(c, (1023ull * 2048) / per_nb);          // the bucket-offset scan handles 1023 tiles of 2048 buckets
    c = std::min<uint64_t>(c, ((1ull << 28) - 1) / per_ne);        // <= 2^28 sorted entries (1 GiB) per MSM
  }
  (void)r;
  return (uint32_t)std::max<uint64_t>(c, 1);
}

static int32_t prove_batch_fused(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, uint32_t B, const ctg_fr* z_batch, const ctg_fr* rs_batch, ctg_proof* staged) {
  CTG_TRY(lane_init(ctx, 0));
  ProveLane& L = ctx->lanes[0];
  const uint64_t n = r->n, m = pk->m, zs = m + 3;
  // scratch: z_ext [B][m+3] | rs [B][2] | a, b, c [3][B][n] | results G1 [4][B], G2 [B] | proofs [B]
  const size_t fr_elems = (size_t)B * zs + 2 * (size_t)B + 3 * (size_t)B * n;
  const size_t bytes = fr_elems * sizeof(Fr) + (size_t)B * (4 * sizeof(G1XYZZ) + sizeof(G2XYZZ) + sizeof(ctg_proof)) + 512;
  CTG_TRY(lane_scratch(ctx, L, bytes));
  Fr* d_z = (Fr*)L.scratch;
  Fr* d_rs = d_z + (size_t)B * zs;
  Fr* d_abc = d_rs + 2 * (size_t)B;
  char* tail = (char*)(d_abc + 3 * (size_t)B * n);
  tail = (char*)(((uintptr_t)tail + 63) & ~(uintptr_t)63);
  G1XYZZ* d_g1 = (G1XYZZ*)tail;                 // [0..B) = A, [B..2B) = B1, [2B..3B) = L, [3B..4B) = H
  G2XYZZ* d_g2 = (G2XYZZ*)(d_g1 + 4 * (size_t)B);
  ctg_proof* d_proof = (ctg_proof*)(d_g2 + B);
  cudaStream_t s0 = L.st[0], sw = L.st_hi;
  NttDomain* dom;
  CTG_TRY(ntt_get_domain(ctx, r->log_n, &dom));
  CTG_CUDA(ctx, cudaMemcpy2DAsync(d_z, zs * sizeof(Fr), z_batch, m * sizeof(Fr), m * sizeof(Fr), B, cudaMemcpyHostToDevice, s0));
  CTG_CUDA(ctx, cudaMemcpyAsync(d_rs, rs_batch, 2 * (size_t)B * sizeof(Fr), cudaMemcpyHostToDevice, s0));
  set_tail_batch_kernel<<<(B + 127) / 128, 128, 0, s0>>>(d_z, m, d_rs, B);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaEventRecord(L.ev_z, s0));
  struct Job { int stream; const Bases* b; const Fr* sc; uint64_t cnt; int g1_slot; };
  const Job jobs[4] = {{1, &pk->l, d_z + pk->ni, pk->l_cnt, 2}, {2, &pk->a_ext, d_z, pk->ext_cnt, 0}, {3, &pk->b1_ext, d_z, pk->ext_cnt, 1}, {4, &pk->b2_ext, d_z, pk->ext_cnt, -1}};
  const bool share = same_plan(pk->a_ext, pk->b1_ext) && pk->ext_cnt > 0;
  for (int oi : {0, 1, 2, 3}) {
    const Job& j = jobs[oi];
    cudaStream_t st = L.st[j.stream];
    CTG_CUDA(ctx, cudaStreamWaitEvent(st, L.ev_z, 0));
    if (j.g1_slot >= 0) {
      const MsmWorkspace* sorted = nullptr;
      if (j.stream == 3 && share) { sorted = &L.ws[2]; CTG_CUDA(ctx, cudaStreamWaitEvent(st, L.ws[2].ev_sorted, 0)); }
      CTG_TRY(msm_g1_run(ctx, L.ws[j.stream], *j.b, (const uint32_t*)j.sc, j.cnt, true, d_g1 + (size_t)j.g1_slot * B, st, nullptr, B, zs, sorted));
    } else {
      CTG_TRY(msm_g2_run(ctx, L.ws[j.stream], *j.b, (const uint32_t*)j.sc, j.cnt, true, d_g2, st, nullptr, B, zs));
    }
    CTG_CUDA(ctx, cudaEventRecord(L.ev_done[j.stream], st));
  }
  // witness map of the whole chunk on the high-priority stream: chains k = a, b, c are stored [k][proof][n]
  CTG_CUDA(ctx, cudaStreamWaitEvent(sw, L.ev_z, 0));
  for (int k = 0; k < 3; k++) {
    spmv_kernel<<<dim3((unsigned)((n + 127) / 128), B), 128, 0, sw>>>(r->row_ptr[k], r->col[k], r->val[k], d_z, d_abc + (size_t)k * B * n, r->nc, r->ni, n, k == 0, zs);
    CTG_KERNEL_CHECK(ctx);
  }
  if (r->log_n == 0) return fail(ctx, CTG_ERR_STATE, "prove_batch: internal: the fused path needs a domain of at least 2");
  CTG_TRY(ntt_ifft_then_coset_fft(ctx, d_abc, r->log_n, sw, 3 * B));
  const uint64_t tot = (uint64_t)B * n;
  pointwise_kernel<<<(unsigned)((tot + 127) / 128), 128, 0, sw>>>(d_abc, d_abc + tot, d_abc + 2 * tot, dom->zinv, tot);
  CTG_KERNEL_CHECK(ctx);
  CTG_TRY(ntt_coset_ifft_bitrev(ctx, d_abc, r->log_n, sw, B));
  set_h_tail_batch_kernel<<<(B + 127) / 128, 128, 0, sw>>>(d_abc, n, d_rs, B);
  CTG_KERNEL_CHECK(ctx);
  CTG_TRY(msm_g1_run(ctx, L.ws[0], pk->h_ext, (const uint32_t*)d_abc, pk->h_cnt, true, d_g1 + 3 * (size_t)B, sw, nullptr, B, n));
  CTG_CUDA(ctx, cudaEventRecord(L.ev_wm, sw));
  CTG_CUDA(ctx, cudaStreamWaitEvent(s0, L.ev_wm, 0));
  for (int i = 1; i < 5; i++) CTG_CUDA(ctx, cudaStreamWaitEvent(s0, L.ev_done[i], 0));
  FinalizeBatchIn fin{d_g1, d_g1 + B, d_g1 + 2 * (size_t)B, d_g1 + 3 * (size_t)B, d_g2, d_rs};
  finalize_batch_kernel<<<B, 96, 0, s0>>>(fin, d_proof);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaMemcpyAsync(staged, d_proof, (size_t)B * sizeof(ctg_proof), cudaMemcpyDeviceToHost, s0));
  return CTG_OK;
}

// Small circuits (domain <= 2^17): chunks through the fused pipeline above.  Large ones saturate the GPU proof by proof; they
// are pipelined over two stream sets so that one proof's latency-bound tails overlap the next one's kernels.
int32_t prove_batch(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, uint32_t n_proofs, const ctg_fr* z_batch, const ctg_fr* rs_batch, ctg_proof* out) {
  CTG_TRY(check_prove_args(ctx, pk, r));
  if (pk->shard_count != 1) return fail(ctx, CTG_ERR_STATE, "prove_batch: the key is a shard");
  if (n_proofs == 0) return CTG_OK;
  const size_t need = (size_t)n_proofs * sizeof(ctg_proof);
  if (need > ctx->pinned_proofs_bytes) {
    if (ctx->pinned_proofs) cudaFreeHost(ctx->pinned_proofs);
    ctx->pinned_proofs = nullptr; ctx->pinned_proofs_bytes = 0;
    CTG_CUDA(ctx, cudaMallocHost(&ctx->pinned_proofs, need));
    ctx->pinned_proofs_bytes = need;
  }
  ctg_proof* staged = (ctg_proof*)ctx->pinned_proofs;
  const uint64_t m = pk->m;
  static int fused_max_log = -1;
  if (fused_max_log < 0) { const char* e = getenv("CTG_BATCH_FUSED_MAX_LOG_N"); fused_max_log = e ? atoi(e) : 17; }
  if (r->log_n >= 1 && r->log_n <= fused_max_log && n_proofs > 1) {
    const ctg_pk* bpk = pk->batch_twin && n_proofs >= 8 ? pk->batch_twin : pk;   // the tables laid out for batches (same points, wider windows)
    const uint32_t chunk = batch_chunk(bpk, r, n_proofs);
    for (uint32_t lo = 0; lo < n_proofs; lo += chunk) {
      const uint32_t cnt = std::min(chunk, n_proofs - lo);
      CTG_TRY(prove_batch_fused(ctx, bpk, r, cnt, z_batch + (size_t)lo * m, rs_batch + 2 * (size_t)lo, staged + lo));
    }
    CTG_CUDA(ctx, cudaStreamSynchronize(ctx->lanes[0].st[0]));
    memcpy(out, staged, need);
    return CTG_OK;
  }
  const int lanes = r->n <= (1u << 18) ? kMaxLanes : 2;
  for (uint32_t i = 0; i < n_proofs; i++) {
    Fr r_m, s_m;
    memcpy(&r_m, rs_batch + 2 * (size_t)i, sizeof(Fr));
    memcpy(&s_m, rs_batch + 2 * (size_t)i + 1, sizeof(Fr));
    CTG_TRY(prove_enqueue(ctx, (int)(i % lanes), pk, r, z_batch + (size_t)i * m, r_m, s_m, staged + i, nullptr, false));
  }
  for (int l = 0; l < lanes; l++)
    if (ctx->lanes[l].ready) CTG_CUDA(ctx, cudaStreamSynchronize(ctx->lanes[l].st[0]));
  memcpy(out, staged, need);
  return CTG_OK;
}

// One shard's partial sums (multi-GPU prove): out = [A, B1, L, H] as G1 XYZZ then B2 as G2 XYZZ (1152 bytes, host)
int32_t prove_partial(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const ctg_fr* rr, const ctg_fr* ss, void* out_partials, ctg_timings* tm) {
  if (!pk || !r || !z || !rr || !ss || !out_partials) return fail(ctx, CTG_ERR_INVALID, "prove_partial: null argument");
  CTG_TRY(check_prove_args(ctx, pk, r));
  Fr r_m, s_m;
  memcpy(&r_m, rr, sizeof(Fr));
  memcpy(&s_m, ss, sizeof(Fr));
  CTG_TRY(prove_enqueue(ctx, 0, pk, r, z, r_m, s_m, nullptr, out_partials, tm != nullptr));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->lanes[0].st[0]));
  if (tm) lane_timings(ctx->lanes[0], tm);
  return CTG_OK;
}

// Multi-GPU witness map: phase 1 computes only the chains in chain_mask into the caller's buffer; the caller exchanges the
// chains between ranks (NCCL broadcast over NVLink); phase 2 finishes.  Both phases use lane 0.
int32_t prove_partial_begin(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const ctg_fr* rr, const ctg_fr* ss, uint32_t chain_mask, void* d_abc) {
  if (!pk || !r || !z || !rr || !ss || !d_abc) return fail(ctx, CTG_ERR_INVALID, "prove_partial_begin: null argument");
  if (chain_mask > 7) return fail(ctx, CTG_ERR_INVALID, "prove_partial_begin: chain_mask has bits beyond a, b, c");
  CTG_TRY(check_prove_args(ctx, pk, r));
  Fr r_m, s_m;
  memcpy(&r_m, rr, sizeof(Fr));
  memcpy(&s_m, ss, sizeof(Fr));
  CTG_TRY(prove_enqueue(ctx, 0, pk, r, z, r_m, s_m, nullptr, nullptr, false, 1, chain_mask, (Fr*)d_abc, false));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->lanes[0].st[0]));   // the chains are complete: safe to hand to a collective on any stream
  return CTG_OK;
}
int32_t prove_partial_finish(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* rr, const ctg_fr* ss, void* d_abc, void* out_partials) {
  if (!pk || !r || !rr || !ss || !d_abc || !out_partials) return fail(ctx, CTG_ERR_INVALID, "prove_partial_finish: null argument");
  CTG_TRY(check_prove_args(ctx, pk, r));
  if (!ctx->lanes[0].ready) return fail(ctx, CTG_ERR_STATE, "prove_partial_finish without prove_partial_begin");
  Fr r_m, s_m;
  memcpy(&r_m, rr, sizeof(Fr));
  memcpy(&s_m, ss, sizeof(Fr));
  CTG_TRY(prove_enqueue(ctx, 0, pk, r, nullptr, r_m, s_m, nullptr, out_partials, false, 2, 0, (Fr*)d_abc));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->lanes[0].st[0]));
  return CTG_OK;
}

// Sum of k shards' partials, then the usual assembly: C = s*A + r*B1 + L + H', affine (A, B2, C)
int32_t prove_combine(ctg_ctx* ctx, const void* host_partials, uint32_t k, const ctg_fr* rr, const ctg_fr* ss, ctg_proof* out) {
  if (!host_partials || !rr || !ss || !out || k == 0) return fail(ctx, CTG_ERR_INVALID, "prove_combine: bad argument");
  const size_t one = 4 * sizeof(G1XYZZ) + sizeof(G2XYZZ);
  CTG_TRY(ensure_scratch(ctx, (size_t)(k + 1) * one + sizeof(ctg_proof) + 256));
  cudaStream_t st = ctx->stream;
  // device layout: g1[k][4] | g2[k] | sums: g1[4], g2 | proof
  G1XYZZ* d_g1 = (G1XYZZ*)ctx->scratch;
  G2XYZZ* d_g2 = (G2XYZZ*)(d_g1 + (size_t)4 * k);
  G1XYZZ* d_s1 = (G1XYZZ*)(d_g2 + k);
  G2XYZZ* d_s2 = (G2XYZZ*)(d_s1 + 4);
  ctg_proof* d_proof = (ctg_proof*)(d_s2 + 1);
  for (uint32_t i = 0; i < k; i++) {
    const char* src = (const char*)host_partials + (size_t)i * one;
    CTG_CUDA(ctx, cudaMemcpyAsync(d_g1 + (size_t)4 * i, src, 4 * sizeof(G1XYZZ), cudaMemcpyHostToDevice, st));
    CTG_CUDA(ctx, cudaMemcpyAsync(d_g2 + i, src + 4 * sizeof(G1XYZZ), sizeof(G2XYZZ), cudaMemcpyHostToDevice, st));
  }
  combine_partials_kernel<<<1, 160, 0, st>>>(d_g1, d_g2, k, d_s1, d_s2);
  CTG_KERNEL_CHECK(ctx);
  Fr r_m, s_m;
  memcpy(&r_m, rr, sizeof(Fr));
  memcpy(&s_m, ss, sizeof(Fr));
  FinalizeIn fin{d_s1 + 0, d_s1 + 1, d_s1 + 2, d_s1 + 3, d_s2, r_m, s_m};
  finalize_kernel<<<1, 96, 0, st>>>(fin, d_proof);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaMemcpyAsync(out, d_proof, sizeof(ctg_proof), cudaMemcpyDeviceToHost, st));
  CTG_CUDA(ctx, cudaStreamSynchronize(st));
  return CTG_OK;
}

uint64_t pk_num_vars(const ctg_pk* pk) { return pk->parts.empty() ? pk->m : pk->parts[0]->m; }

// ---- multi-device proof: one process, one worker thread per device, peer copies over NVLink ---------------------------
// Phases (multi.cu runs A on every device, then B on device 0, then C on every device, then D on device 0; the host only
// orders the ENQUEUE of the phases -- cudaStreamWaitEvent needs its event recorded first -- and synchronises once at the end):
//   A  upload the part of z this device reads, start its A / B1 / B2 / L MSMs, compute the witness-map chains it owns
//   B  device 0 pulls the chains it does not own from their owners (n * 32 B each), runs (a b - c)/Z and the coset_ifft ONCE
//   C  every device pulls ITS slice of h (n / N * 32 B) from device 0, runs its H MSM, and pushes 1152 B of partial sums to device 0
//   D  device 0 sums the partials and assembles (A, B, C)
// Bytes over NVLink per proof: 2-3 chains + one vector of h in slices, instead of broadcasting three chains to every device.
int32_t prove_group_a(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const Fr& r_m, const Fr& s_m, uint32_t chain_mask, Fr* d_abc,
                      cudaEvent_t ev_done) {
  CTG_TRY(prove_enqueue(ctx, 0, pk, r, z, r_m, s_m, nullptr, nullptr, false, 1, chain_mask, d_abc));
  CTG_CUDA(ctx, cudaEventRecord(ev_done, ctx->lanes[0].st[0]));   // s0 is ordered after the chains (prove_enqueue phase 1)
  return CTG_OK;
}
int32_t prove_group_b(Group* g, const ctg_r1cs* r0, const Fr& r_m, const Fr& s_m) {
  ctg_ctx* ctx = g->ctx[0];
  ProveLane& L = ctx->lanes[0];
  cudaStream_t s0 = L.st[0], sw = L.st_hi;
  const uint64_t n = r0->n;
  for (int k = 0; k < 3; k++) {
    const int o = g->chain_owner[k];
    if (o == 0) continue;
    CTG_CUDA(ctx, cudaStreamWaitEvent(s0, g->ev_a[o], 0));
    CTG_CUDA(ctx, cudaMemcpyPeerAsync(g->d_abc[0] + (uint64_t)k * n, g->ctx[0]->device, g->d_abc[o] + (uint64_t)k * n, g->ctx[o]->device, n * sizeof(Fr), s0));
  }
  CTG_CUDA(ctx, cudaEventRecord(L.ev_zfull, s0));
  CTG_CUDA(ctx, cudaStreamWaitEvent(sw, L.ev_zfull, 0));
  CTG_TRY(witness_finish_dev(ctx, r0, g->d_abc[0], sw));
  set_h_tail_kernel<<<1, 32, 0, sw>>>(g->d_abc[0], n, r_m, s_m);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaEventRecord(g->ev_h, sw));
  return CTG_OK;
}
// slots of a device's partial record: [0] = A_d, [1] = B1_d -> C_d, [2] = L_d, [3] = H_d (G1 XYZZ), then B2_d (G2 XYZZ)
__global__ void __launch_bounds__(32) group_share_kernel(const G1XYZZ* early /* s A_d, r B1_d */, G1XYZZ* g1) {
  if (threadIdx.x != 0) return;
  G1XYZZ c = xyzz_add(early[0], early[1]);
  c = xyzz_add(c, g1[2]);
  c = xyzz_add(c, g1[3]);
  g1[1] = c;
}
// device 0: A = sum_d A_d, C = sum_d C_d, B = sum_d B2_d by pairwise trees (three warps, up to 8 shards), then to affine
template <class P> __device__ __forceinline__ P shard_tree_sum(const char* parts, size_t offset, uint32_t k, P* sh, unsigned lane) {
  auto part = [&](uint32_t i) { return i < k ? *reinterpret_cast<const P*>(parts + (size_t)i * kPartialBytes + offset) : P::inf(); };
  if (lane < 4) sh[lane] = xyzz_add(part(lane), part(lane + 4));
  __syncwarp();
  if (lane < 2) sh[lane] = xyzz_add(sh[lane], sh[lane + 2]);
  __syncwarp();
  P total = P::inf();
  if (lane == 0) total = xyzz_add(sh[0], sh[1]);
  return total;
}
__global__ void __launch_bounds__(96) group_finalize_kernel(const char* parts, uint32_t k, ctg_proof* out) {
  __shared__ G1XYZZ sh1[2][4];
  __shared__ G2XYZZ sh2[4];
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < 2) {
    G1XYZZ t = shard_tree_sum<G1XYZZ>(parts, warp * sizeof(G1XYZZ), k, sh1[warp], lane);
    if (lane == 0) {
      G1Affine a = xyzz_to_affine(t);
      *reinterpret_cast<G1Affine*>(warp == 0 ? (void*)&out->a : (void*)&out->c) = a;
    }
  } else {
    G2XYZZ t = shard_tree_sum<G2XYZZ>(parts, 4 * sizeof(G1XYZZ), k, sh2, lane);
    if (lane == 0) {
      G2Affine b = xyzz_to_affine(t);
      *reinterpret_cast<G2Affine*>(&out->b) = b;
    }
  }
}
int32_t prove_group_c(Group* g, int d, const ctg_pk* pk, const ctg_r1cs* r) {
  ctg_ctx* ctx = g->ctx[d];
  ProveLane& L = ctx->lanes[0];
  cudaStream_t s0 = L.st[0], sw = L.st_hi;
  LaneBufs B;
  CTG_TRY(lane_bufs(ctx, L, pk->m, r->n, d == 0 ? g->d_abc[0] : g->d_hslice[d], &B));
  const Fr* h_scalars = g->d_abc[0] + pk->h_lo;
  if (d != 0) {
    CTG_CUDA(ctx, cudaStreamWaitEvent(sw, g->ev_h, 0));
    if (pk->h_cnt)
      CTG_CUDA(ctx, cudaMemcpyPeerAsync(g->d_hslice[d], ctx->device, g->d_abc[0] + pk->h_lo, g->ctx[0]->device, pk->h_cnt * sizeof(Fr), sw));
    h_scalars = g->d_hslice[d];
  }
  CTG_TRY(msm_g1_run(ctx, L.ws[0], pk->h_ext, (const uint32_t*)h_scalars, pk->h_cnt, true, B.d_g1 + 3, sw, nullptr));
  CTG_CUDA(ctx, cudaEventRecord(L.ev_wm, sw));
  CTG_CUDA(ctx, cudaStreamWaitEvent(s0, L.ev_wm, 0));
  for (int i = 1; i < 5; i++) CTG_CUDA(ctx, cudaStreamWaitEvent(s0, L.ev_done[i], 0));
  // C is linear in the shards: C = sum_d (s A_d + r B1_d + L_d + H_d).  s A_d and r B1_d were started by phase A on the streams of
  // their MSMs (prove_enqueue early_mul), so each device closes ITS share here and device 0 only adds points.
  group_share_kernel<<<1, 32, 0, s0>>>(B.d_early, B.d_g1);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaMemcpyPeerAsync((char*)g->d_parts0 + (size_t)d * kPartialBytes, g->ctx[0]->device, B.d_g1, ctx->device, kPartialBytes, s0));
  CTG_CUDA(ctx, cudaEventRecord(g->ev_part[d], s0));
  return CTG_OK;
}
// the k-th shard's partials sit at parts + k * 1152: [A, B1, L, H] G1 XYZZ then B2 G2 XYZZ
__global__ void __launch_bounds__(160) combine_partials_packed_kernel(const char* parts, uint32_t k, G1XYZZ* out_g1, G2XYZZ* out_g2) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane != 0) return;
  if (warp < 4) {
    G1XYZZ acc = reinterpret_cast<const G1XYZZ*>(parts)[warp];
    for (uint32_t i = 1; i < k; i++) acc = xyzz_add(acc, reinterpret_cast<const G1XYZZ*>(parts + (size_t)i * kPartialBytes)[warp]);
    out_g1[warp] = acc;
  } else {
    G2XYZZ acc = *reinterpret_cast<const G2XYZZ*>(parts + 4 * sizeof(G1XYZZ));
    for (uint32_t i = 1; i < k; i++) acc = xyzz_add(acc, *reinterpret_cast<const G2XYZZ*>(parts + (size_t)i * kPartialBytes + 4 * sizeof(G1XYZZ)));
    *out_g2 = acc;
  }
}
int32_t prove_group_d(Group* g, const Fr& r_m, const Fr& s_m, ctg_proof* out) {
  ctg_ctx* ctx = g->ctx[0];
  cudaStream_t s0 = ctx->lanes[0].st[0];
  for (int d = 1; d < g->n; d++) CTG_CUDA(ctx, cudaStreamWaitEvent(s0, g->ev_part[d], 0));
  char* base = (char*)g->d_parts0;
  G1XYZZ* d_s1 = (G1XYZZ*)(base + (size_t)g->n * kPartialBytes);
  G2XYZZ* d_s2 = (G2XYZZ*)(d_s1 + 4);
  ctg_proof* d_proof = (ctg_proof*)(d_s2 + 1);
  (void)r_m; (void)s_m;   // already inside every device's C_d
  group_finalize_kernel<<<1, 96, 0, s0>>>(base, (uint32_t)g->n, d_proof);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaMemcpyAsync(g->h_proof, d_proof, sizeof(ctg_proof), cudaMemcpyDeviceToHost, s0));
  CTG_CUDA(ctx, cudaStreamSynchronize(s0));
  memcpy(out, g->h_proof, sizeof(ctg_proof));
  return CTG_OK;
}

}  // namespace ctg
