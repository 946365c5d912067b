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
#include "ctg_internal.h"

struct ctg_pk {
  uint64_t m, ni, nw, n;
  int log_n;
  ctg::Bases a_ext, b1_ext, b2_ext, l, h_ext;
};
struct ctg_r1cs {
  uint64_t ni, nw, nc, n;
  int log_n;
  uint64_t* row_ptr[3];
  uint32_t* col[3];
  ctg::Fr* val[3];
  uint64_t nnz[3];
};

namespace ctg {

// ---- witness map kernels --------------------------------------------------------------------
// out[i] = <M_i, z> for i < nc; for the A matrix rows nc..nc+ni copy z[j] (input consistency); else 0
__global__ void spmv_kernel(const uint64_t* __restrict__ row_ptr, const uint32_t* __restrict__ col, const Fr* __restrict__ val,
                            const Fr* __restrict__ z, Fr* __restrict__ out, uint64_t nc, uint64_t ni, uint64_t n, int is_a) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
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
__global__ void bitrev_gather_fr_kernel(const Fr* __restrict__ in, Fr* __restrict__ out, int log_n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (1ull << log_n)) return;
  uint64_t j = log_n == 0 ? 0 : (uint64_t)(__brev((uint32_t)i) >> (32 - log_n));
  out[i] = in[j];
}
// dst[bitrev(i)] = src[i] for i < count  (count = n - 1), 96-byte points
__global__ void bitrev_scatter_g1_kernel(const G1Affine* __restrict__ src, G1Affine* __restrict__ dst, uint64_t count, int log_n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  uint64_t j = log_n == 0 ? 0 : (uint64_t)(__brev((uint32_t)i) >> (32 - log_n));
  const uint4* s = reinterpret_cast<const uint4*>(src + i);
  uint4* d = reinterpret_cast<uint4*>(dst + j);
#pragma unroll
  for (int k = 0; k < 6; k++) d[k] = s[k];
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

// ---- host -----------------------------------------------------------------------------------
static void free_bases(Bases& b) { if (b.pts) cudaFree(b.pts); b.pts = nullptr; }

template <class A>
static int32_t make_bases(ctg_ctx* ctx, Bases& out, int group, uint64_t n, int copies, int c_bits) {
  out.group = group;
  out.n = n;
  out.plan = msm_make_plan(n, copies, c_bits, sizeof(A));
  out.pts = nullptr;
  size_t bytes = (size_t)out.plan.copies * (n ? n : 1) * sizeof(A);
  cudaError_t e = cudaMalloc(&out.pts, bytes);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, CTG_ERR_OOM, "bases: cudaMalloc failed (try fewer precomputed copies)"); }
  CTG_CUDA(ctx, cudaMemsetAsync(out.pts, 0, bytes, ctx->stream));
  return CTG_OK;
}
static int32_t finish_bases(ctg_ctx* ctx, Bases& b) { return b.group == 1 ? g1_precompute(ctx, b, ctx->stream) : g2_precompute(ctx, b, ctx->stream); }

int32_t pk_upload(ctg_ctx* ctx, const ctg_pk_view* v, uint64_t ni, uint64_t nc, int copies, int c_bits, ctg_pk** out) {
  if (!v || !out) return fail(ctx, CTG_ERR_INVALID, "pk_upload: null argument");
  const uint64_t m = v->a_len;
  if (m < ni || m == 0 || v->b_g1_len != m || v->b_g2_len != m) return fail(ctx, CTG_ERR_INVALID, "pk_upload: a/b query lengths must equal num_instance + num_witness");
  const uint64_t nw = m - ni;
  if (v->l_len != nw) return fail(ctx, CTG_ERR_INVALID, "pk_upload: l_query length must equal num_witness");
  uint64_t n = 1; int log_n = 0;
  while (n < nc + ni) { n <<= 1; log_n++; }
  if (v->h_len + 1 != n) return fail(ctx, CTG_ERR_INVALID, "pk_upload: h_query length must be domain size - 1");
  ctg_pk* pk = new ctg_pk();
  pk->m = m; pk->ni = ni; pk->nw = nw; pk->n = n; pk->log_n = log_n;
  cudaStream_t st = ctx->stream;
  auto cleanup = [&](int32_t rc) { free_bases(pk->a_ext); free_bases(pk->b1_ext); free_bases(pk->b2_ext); free_bases(pk->l); free_bases(pk->h_ext); delete pk; return rc; };
  memset(&pk->a_ext, 0, sizeof(Bases)); memset(&pk->b1_ext, 0, sizeof(Bases)); memset(&pk->b2_ext, 0, sizeof(Bases));
  memset(&pk->l, 0, sizeof(Bases)); memset(&pk->h_ext, 0, sizeof(Bases));
  int32_t rc;
  if ((rc = make_bases<G1Affine>(ctx, pk->a_ext, 1, m + 3, copies, c_bits)) != CTG_OK) return cleanup(rc);
  if ((rc = make_bases<G1Affine>(ctx, pk->b1_ext, 1, m + 3, copies, c_bits)) != CTG_OK) return cleanup(rc);
  if ((rc = make_bases<G2Affine>(ctx, pk->b2_ext, 2, m + 3, copies, c_bits)) != CTG_OK) return cleanup(rc);
  if ((rc = make_bases<G1Affine>(ctx, pk->l, 1, nw, copies, c_bits)) != CTG_OK) return cleanup(rc);
  if ((rc = make_bases<G1Affine>(ctx, pk->h_ext, 1, n, copies, c_bits)) != CTG_OK) return cleanup(rc);
#define PK_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { ctx->last_error = std::string(#call) + ": " + cudaGetErrorString(e__); return cleanup(CTG_ERR_CUDA); } } while (0)
  G1Affine* a = (G1Affine*)pk->a_ext.pts; G1Affine* b1 = (G1Affine*)pk->b1_ext.pts; G2Affine* b2 = (G2Affine*)pk->b2_ext.pts;
  PK_CUDA(cudaMemcpyAsync(a, v->a_query, m * sizeof(G1Affine), cudaMemcpyHostToDevice, st));
  PK_CUDA(cudaMemcpyAsync(a + m, v->alpha_g1, sizeof(G1Affine), cudaMemcpyHostToDevice, st));
  PK_CUDA(cudaMemcpyAsync(a + m + 1, v->delta_g1, sizeof(G1Affine), cudaMemcpyHostToDevice, st));
  PK_CUDA(cudaMemcpyAsync(b1, v->b_g1_query, m * sizeof(G1Affine), cudaMemcpyHostToDevice, st));
  PK_CUDA(cudaMemcpyAsync(b1 + m, v->beta_g1, sizeof(G1Affine), cudaMemcpyHostToDevice, st));
  PK_CUDA(cudaMemcpyAsync(b1 + m + 2, v->delta_g1, sizeof(G1Affine), cudaMemcpyHostToDevice, st));
  PK_CUDA(cudaMemcpyAsync(b2, v->b_g2_query, m * sizeof(G2Affine), cudaMemcpyHostToDevice, st));
  PK_CUDA(cudaMemcpyAsync(b2 + m, v->beta_g2, sizeof(G2Affine), cudaMemcpyHostToDevice, st));
  PK_CUDA(cudaMemcpyAsync(b2 + m + 2, v->delta_g2, sizeof(G2Affine), cudaMemcpyHostToDevice, st));
  if (nw) PK_CUDA(cudaMemcpyAsync(pk->l.pts, v->l_query, nw * sizeof(G1Affine), cudaMemcpyHostToDevice, st));
  // h_query: stage natural order in scratch, scatter to bit-reversed slots
  if (n > 1) {
    if ((rc = ensure_scratch(ctx, (n - 1) * sizeof(G1Affine))) != CTG_OK) return cleanup(rc);
    PK_CUDA(cudaMemcpyAsync(ctx->scratch, v->h_query, (n - 1) * sizeof(G1Affine), cudaMemcpyHostToDevice, st));
    bitrev_scatter_g1_kernel<<<(unsigned)((n - 1 + 255) / 256), 256, 0, st>>>((const G1Affine*)ctx->scratch, (G1Affine*)pk->h_ext.pts, n - 1, log_n);
    ctx->launches++;
  }
  PK_CUDA(cudaMemcpyAsync((G1Affine*)pk->h_ext.pts + (n - 1), v->delta_g1, sizeof(G1Affine), cudaMemcpyHostToDevice, st));
  for (Bases* b : {&pk->a_ext, &pk->b1_ext, &pk->b2_ext, &pk->l, &pk->h_ext})
    if ((rc = finish_bases(ctx, *b)) != CTG_OK) return cleanup(rc);
  PK_CUDA(cudaStreamSynchronize(st));
#undef PK_CUDA
  *out = pk;
  return CTG_OK;
}

void pk_free(ctg_pk* pk) {
  if (!pk) return;
  free_bases(pk->a_ext); free_bases(pk->b1_ext); free_bases(pk->b2_ext); free_bases(pk->l); free_bases(pk->h_ext);
  delete pk;
}

int32_t r1cs_upload(ctg_ctx* ctx, uint64_t ni, uint64_t nw, uint64_t nc, const ctg_csr* const mats[3], ctg_r1cs** out) {
  if (ni == 0) return fail(ctx, CTG_ERR_INVALID, "r1cs_upload: num_instance includes the constant 1 and must be >= 1");
  ctg_r1cs* r = new ctg_r1cs();
  memset(r, 0, sizeof(*r));
  r->ni = ni; r->nw = nw; r->nc = nc;
  uint64_t n = 1; int log_n = 0;
  while (n < nc + ni) { n <<= 1; log_n++; }
  r->n = n; r->log_n = log_n;
  auto cleanup = [&](int32_t rc) { for (int k = 0; k < 3; k++) { cudaFree(r->row_ptr[k]); cudaFree(r->col[k]); cudaFree(r->val[k]); } delete r; return rc; };
  if (log_n > 28) { ctx->last_error = "r1cs_upload: domain larger than 2^28"; return cleanup(CTG_ERR_INVALID); }
  for (int k = 0; k < 3; k++) {
    const ctg_csr* m = mats[k];
    if (!m || !m->row_ptr) { ctx->last_error = "r1cs_upload: null matrix"; return cleanup(CTG_ERR_INVALID); }
    const uint64_t nnz = m->row_ptr[nc];
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

// d_z: m (+3) Montgomery elements on the device.  d_abc: 3 * n scratch.  Leaves h (bit-reversed,
// Montgomery) in d_abc[0..n).
static int32_t witness_map_dev(ctg_ctx* ctx, const ctg_r1cs* r, const Fr* d_z, Fr* d_abc, cudaStream_t st) {
  const uint64_t n = r->n;
  NttDomain* dom;
  CTG_TRY(ntt_get_domain(ctx, r->log_n, &dom));
  Fr* v[3] = {d_abc, d_abc + n, d_abc + 2 * n};
  const unsigned blocks = (unsigned)((n + 127) / 128);
  for (int k = 0; k < 3; k++) {
    spmv_kernel<<<blocks, 128, 0, st>>>(r->row_ptr[k], r->col[k], r->val[k], d_z, v[k], r->nc, r->ni, n, k == 0);
    CTG_KERNEL_CHECK(ctx);
    CTG_TRY(ntt_ifft_then_coset_fft(ctx, v[k], r->log_n, st));
  }
  pointwise_kernel<<<blocks, 128, 0, st>>>(v[0], v[1], v[2], dom->zinv, n);
  CTG_KERNEL_CHECK(ctx);
  return ntt_coset_ifft_bitrev(ctx, v[0], r->log_n, st);
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

int32_t prove(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const ctg_fr* rr, const ctg_fr* ss, ctg_proof* out, ctg_timings* tm) {
  if (!pk || !r || !z || !rr || !ss || !out) return fail(ctx, CTG_ERR_INVALID, "prove: null argument");
  if (pk->ni != r->ni || pk->nw != r->nw || pk->n != r->n) return fail(ctx, CTG_ERR_INVALID, "prove: proving key and R1CS shapes differ");
  const uint64_t n = r->n, m = pk->m;
  cudaStream_t st = ctx->stream;
  // scratch: z_ext (m+3) | a,b,c (3n) | results: 4 G1 XYZZ + 1 G2 XYZZ | proof
  const size_t fr_elems = m + 3 + 3 * n;
  const size_t bytes = fr_elems * sizeof(Fr) + 4 * sizeof(G1XYZZ) + sizeof(G2XYZZ) + sizeof(ctg_proof) + 256;
  CTG_TRY(ensure_scratch(ctx, bytes));
  Fr* d_z = (Fr*)ctx->scratch;
  Fr* d_abc = d_z + (m + 3);
  char* tail = (char*)(d_abc + 3 * n);
  tail = (char*)(((uintptr_t)tail + 63) & ~(uintptr_t)63);
  G1XYZZ* d_g1 = (G1XYZZ*)tail;           // [0]=A [1]=B1 [2]=L [3]=H
  G2XYZZ* d_g2 = (G2XYZZ*)(d_g1 + 4);
  ctg_proof* d_proof = (ctg_proof*)(d_g2 + 1);
  Fr r_m, s_m;
  memcpy(&r_m, rr, sizeof(Fr));
  memcpy(&s_m, ss, sizeof(Fr));
  const bool timed = tm != nullptr;
  if (!ctx->events_ready) {
    for (auto& e : ctx->ev) CTG_CUDA(ctx, cudaEventCreate(&e));
    ctx->events_ready = true;
  }
  cudaEvent_t* ev = ctx->ev + 4;  // ev[0..3] belong to the MSM stage timers
  int ne = 0;
  auto mark = [&]() { if (timed && ne < 10) cudaEventRecord(ev[ne++], st); };
  mark();
  CTG_CUDA(ctx, cudaMemcpyAsync(d_z, z, m * sizeof(Fr), cudaMemcpyHostToDevice, st));
  set_tail_kernel<<<1, 32, 0, st>>>(d_z, m, r_m, s_m);
  CTG_KERNEL_CHECK(ctx);
  mark();
  CTG_TRY(witness_map_dev(ctx, r, d_z, d_abc, st));
  set_h_tail_kernel<<<1, 32, 0, st>>>(d_abc, n, r_m, s_m);
  CTG_KERNEL_CHECK(ctx);
  mark();
  MsmWorkspace& ws = ctx->ws[0];
  CTG_TRY(msm_g1_run(ctx, ws, pk->h_ext, (const uint32_t*)d_abc, n, true, d_g1 + 3, st, nullptr));
  mark();
  CTG_TRY(msm_g1_run(ctx, ws, pk->l, (const uint32_t*)(d_z + pk->ni), pk->nw, true, d_g1 + 2, st, nullptr));
  mark();
  CTG_TRY(msm_g1_run(ctx, ws, pk->a_ext, (const uint32_t*)d_z, m + 3, true, d_g1 + 0, st, nullptr));
  mark();
  CTG_TRY(msm_g1_run(ctx, ws, pk->b1_ext, (const uint32_t*)d_z, m + 3, true, d_g1 + 1, st, nullptr));
  mark();
  CTG_TRY(msm_g2_run(ctx, ws, pk->b2_ext, (const uint32_t*)d_z, m + 3, true, d_g2, st, nullptr));
  mark();
  FinalizeIn fin{d_g1 + 0, d_g1 + 1, d_g1 + 2, d_g1 + 3, d_g2, r_m, s_m};
  finalize_kernel<<<1, 96, 0, st>>>(fin, d_proof);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaMemcpyAsync(out, d_proof, sizeof(ctg_proof), cudaMemcpyDeviceToHost, st));
  mark();
  CTG_CUDA(ctx, cudaStreamSynchronize(st));
  if (timed) {
    memset(tm, 0, sizeof(*tm));
    float* dst[8] = {&tm->h2d_ms, &tm->witness_map_ms, &tm->msm_h_ms, &tm->msm_l_ms, &tm->msm_a_ms, &tm->msm_b_g1_ms, &tm->msm_b_g2_ms, &tm->finalize_ms};
    for (int i = 0; i + 1 < ne && i < 8; i++) cudaEventElapsedTime(dst[i], ev[i], ev[i + 1]);
    cudaEventElapsedTime(&tm->total_ms, ev[0], ev[ne - 1]);
  }
  return CTG_OK;
}

}  // namespace ctg

namespace ctg {
uint64_t pk_num_vars(const ctg_pk* pk) { return pk->m; }
}  // namespace ctg
