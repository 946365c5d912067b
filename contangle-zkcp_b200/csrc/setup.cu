// Groth16 key generation on the device.
// Replaces ark-groth16 0.3.0 generator.rs::generate_parameters as reached from the reference's
// Groth16::setup calls (lib/src/zk/encryption.rs:169, sample_entries.rs:141, property.rs:192):
//   1. Lagrange coefficients of the radix-2 domain at tau (ark-poly evaluate_all_lagrange_coefficients),
//   2. the QAP polynomials at tau: u_j = sum_i A[i][j] L_i (+ L_{nc + j} for instance columns), v_j, w_j,
//   3. the query scalars  u_j | v_j | (beta u_j + alpha v_j + w_j)/gamma or /delta | tau^i t(tau)/delta,
//   4. fixed-base batch multiplication by the G1 / G2 generators (arkworks: FixedBaseMSM::multi_scalar_mul).
// The trapdoor (alpha, beta, gamma, delta, tau) and the two generators are drawn by the caller (the Rust
// shim, in arkworks' order, so the RNG stream matches) and passed in; nothing here is random.
// Every output is the unique affine image of its scalar multiple, so "bit-exact" constrains the scalars
// and the group law only; the window/table schedule below is free.
#include <algorithm>
#include <cstring>

#include "ctg_internal.h"

namespace ctg {

static constexpr int kWarp = 32;

// ---- 1. Lagrange coefficients ---------------------------------------------------------------
// L_i(tau) = (tau^n - 1)/n * w^i / (tau - w^i); tw holds w^k for k < n/2 (w^(k + n/2) = -w^k).
__global__ void __launch_bounds__(128) lagrange_kernel(const Fr* __restrict__ tw, uint64_t n, Fr tau, Fr zt_over_n, Fr* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t half = n > 1 ? n / 2 : 1;
  Fr wi = i < half ? tw[i] : tw[i - half].neg();
  Fr d = tau - wi;
  out[i] = zt_over_n * wi * d.inv_bingcd();
}

// ---- 2. QAP polynomials at tau: column sums of a CSR matrix through its transposed index -------
// warp per column: out[j] = sum_{k in col j} val[src[k]] * L[row[k]]  (+ L[nc + j] for instance columns of A)
__device__ __forceinline__ Fr warp_sum_fr(Fr acc) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    Fr o;
#pragma unroll
    for (int k = 0; k < 8; k++) o.v[k] = __shfl_down_sync(0xffffffffu, acc.v[k], off);
    acc = acc + o;
  }
  return acc;
}
__global__ void __launch_bounds__(128) qap_columns_kernel(const uint64_t* __restrict__ col_ptr, const uint32_t* __restrict__ row,
                                                          const uint32_t* __restrict__ src, const Fr* __restrict__ val, const Fr* __restrict__ L,
                                                          uint64_t m, uint64_t nc, uint64_t ni, int is_a, Fr* __restrict__ out) {
  const uint64_t j = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / kWarp;
  const unsigned lane = threadIdx.x & (kWarp - 1);
  if (j >= m) return;
  Fr acc = Fr::zero();
  const uint64_t lo = col_ptr[j], hi = col_ptr[j + 1];
  for (uint64_t k = lo + lane; k < hi; k += kWarp) acc = acc + val[src[k]] * L[row[k]];
  acc = warp_sum_fr(acc);
  if (lane == 0) {
    if (is_a && j < ni) acc = acc + L[nc + j];
    out[j] = acc;
  }
}

// ---- 3. query scalars --------------------------------------------------------------------------
// abc[j] = (beta u_j + alpha v_j + w_j) * (j < ni ? 1/gamma : 1/delta)
__global__ void __launch_bounds__(128) abc_kernel(const Fr* __restrict__ u, const Fr* __restrict__ v, const Fr* __restrict__ w, uint64_t m, uint64_t ni,
                                                  Fr alpha, Fr beta, Fr ginv, Fr dinv, Fr* __restrict__ out) {
  const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  Fr t = beta * u[j] + alpha * v[j] + w[j];
  out[j] = t * (j < ni ? ginv : dinv);
}

// ---- 4. fixed-base batch multiplication ----------------------------------------------------------
// table[w << c | d] = d * 2^(c w) * G (d = 0 is the all-zero image = infinity); out[i] = k_i * G, affine.
template <class F> __device__ __forceinline__ Affine<F> to_affine_fast(const XYZZ<F>& p) {
  if (p.is_inf()) return Affine<F>::inf();
  F t = (p.ZZ * p.ZZZ).inv_bingcd();
  F izz = t * p.ZZZ, izzz = t * p.ZZ;
  return {p.X * izz, p.Y * izzz};
}
template <class F> __global__ void __launch_bounds__(128) fixed_base_kernel(const Affine<F>* __restrict__ table, int c, int nwin,
                                                                            const Fr* __restrict__ scalars, uint64_t n, Affine<F>* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Fr k = scalars[i].from_mont();
  XYZZ<F> acc = XYZZ<F>::inf();
#pragma unroll 1
  for (int w = 0; w < nwin; w++) {
    const int bit = w * c;
    if (bit >= 256) break;
    uint64_t word = k.v[bit >> 5];
    if ((bit >> 5) + 1 < 8) word |= (uint64_t)k.v[(bit >> 5) + 1] << 32;
    const uint32_t d = (uint32_t)(word >> (bit & 31)) & ((1u << c) - 1);
    if (d == 0) continue;
    Affine<F> q = table[((uint64_t)w << c) | d];
    acc = xyzz_madd(acc, q);
  }
  out[i] = to_affine_fast(acc);
}

// ---- host orchestration ------------------------------------------------------------------------
struct DevBuf {  // frees on scope exit
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  template <class T> T* as() { return reinterpret_cast<T*>(p); }
};
#define SETUP_ALLOC(ctx, buf, bytes)                                                                       \
  do {                                                                                                     \
    if (cudaMalloc(&(buf).p, (bytes) ? (bytes) : 16) != cudaSuccess) {                                      \
      cudaGetLastError();                                                                                  \
      return fail(ctx, CTG_ERR_OOM, "generate_parameters: cudaMalloc failed");                             \
    }                                                                                                      \
  } while (0)

// column sums of one matrix: transposes the index on the host (counting sort by column), gathers on the device
static int32_t qap_columns(ctg_ctx* ctx, const ctg_csr* M, uint64_t nc, uint64_t ni, uint64_t m, const Fr* d_L, int is_a, Fr* d_out, cudaStream_t st) {
  if (!M || !M->row_ptr) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: null matrix");
  if (M->row_ptr[0] != 0) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: row_ptr[0] must be 0");
  for (uint64_t i = 0; i < nc; i++)
    if (M->row_ptr[i + 1] < M->row_ptr[i]) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: row_ptr must be non-decreasing");
  const uint64_t nnz = M->row_ptr[nc];
  if (nnz >= (1ull << 32)) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: more than 2^32 - 1 non-zeros in one matrix");
  if (nnz && (!M->col || !M->val)) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: null matrix arrays");
  std::vector<uint64_t> col_ptr(m + 1, 0);
  for (uint64_t k = 0; k < nnz; k++) {
    if (M->col[k] >= m) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: column index out of range");
    col_ptr[M->col[k] + 1]++;
  }
  for (uint64_t j = 0; j < m; j++) col_ptr[j + 1] += col_ptr[j];
  std::vector<uint32_t> row(nnz ? nnz : 1), src(nnz ? nnz : 1);
  {
    std::vector<uint64_t> cur(col_ptr.begin(), col_ptr.end() - 1);
    for (uint64_t i = 0; i < nc; i++)
      for (uint64_t k = M->row_ptr[i]; k < M->row_ptr[i + 1]; k++) {
        const uint64_t pos = cur[M->col[k]]++;
        row[pos] = (uint32_t)i;
        src[pos] = (uint32_t)k;
      }
  }
  DevBuf d_ptr, d_row, d_src, d_val;
  SETUP_ALLOC(ctx, d_ptr, (m + 1) * sizeof(uint64_t));
  SETUP_ALLOC(ctx, d_row, nnz * sizeof(uint32_t));
  SETUP_ALLOC(ctx, d_src, nnz * sizeof(uint32_t));
  SETUP_ALLOC(ctx, d_val, nnz * sizeof(Fr));
  CTG_CUDA(ctx, cudaMemcpyAsync(d_ptr.p, col_ptr.data(), (m + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  if (nnz) {
    CTG_CUDA(ctx, cudaMemcpyAsync(d_row.p, row.data(), nnz * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    CTG_CUDA(ctx, cudaMemcpyAsync(d_src.p, src.data(), nnz * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    CTG_CUDA(ctx, cudaMemcpyAsync(d_val.p, M->val, nnz * sizeof(Fr), cudaMemcpyHostToDevice, st));
  }
  const uint64_t threads = m * kWarp;
  qap_columns_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(d_ptr.as<uint64_t>(), d_row.as<uint32_t>(), d_src.as<uint32_t>(), d_val.as<Fr>(),
                                                                        d_L, m, nc, ni, is_a, d_out);
  CTG_KERNEL_CHECK(ctx);
  CTG_CUDA(ctx, cudaStreamSynchronize(st));  // the host vectors and DevBufs die here
  return CTG_OK;
}

// table of d * 2^(c w) * g for d < 2^c, w < nwin, built with the progression kernel (msm.cuh)
template <class A, class ProgFn>
static int32_t build_table(ctg_ctx* ctx, const A& g, int c, int nwin, A* d_table, ProgFn prog, cudaStream_t st) {
  Fr b = Fr::one();
  const Fr shift = fr_from_u64(1ull << c);
  for (int w = 0; w < nwin; w++) {
    CTG_TRY(prog(ctx, g, Fr::zero(), b, 1ull << c, d_table + ((uint64_t)w << c), st));
    b = b * shift;
  }
  return CTG_OK;
}
template <class F>
static int32_t fixed_base(ctg_ctx* ctx, const Affine<F>* d_table, int c, int nwin, const Fr* d_scalars, uint64_t count, Affine<F>* d_out, cudaStream_t st) {
  if (!count) return CTG_OK;
  fixed_base_kernel<F><<<(unsigned)((count + 127) / 128), 128, 0, st>>>(d_table, c, nwin, d_scalars, count, d_out);
  CTG_KERNEL_CHECK(ctx);
  return CTG_OK;
}

void params_free(ctg_params* p) {
  if (!p) return;
  void* ptrs[] = {p->a_query, p->b_g1_query, p->h_query, p->l_query, p->gamma_abc_g1, p->b_g2_query, p->singles_g1, p->singles_g2};
  for (void* q : ptrs)
    if (q) cudaFree(q);
  delete p;
}

int32_t params_alloc(ctg_ctx* ctx, uint64_t ni, uint64_t nw, uint64_t nc, uint64_t n, ctg_params** out) {
  ctg_params* p = new ctg_params();
  const uint64_t m = ni + nw;
  p->ni = ni; p->nw = nw; p->nc = nc; p->m = m; p->n = n;
  auto alloc = [&](void** dst, size_t bytes) { if (cudaMalloc(dst, bytes ? bytes : 16) != cudaSuccess) { cudaGetLastError(); return false; } return true; };
  if (!alloc((void**)&p->a_query, m * sizeof(G1Affine)) || !alloc((void**)&p->b_g1_query, m * sizeof(G1Affine)) ||
      !alloc((void**)&p->b_g2_query, m * sizeof(G2Affine)) || !alloc((void**)&p->h_query, (n - 1) * sizeof(G1Affine)) ||
      !alloc((void**)&p->l_query, nw * sizeof(G1Affine)) || !alloc((void**)&p->gamma_abc_g1, ni * sizeof(G1Affine)) ||
      !alloc((void**)&p->singles_g1, 3 * sizeof(G1Affine)) || !alloc((void**)&p->singles_g2, 3 * sizeof(G2Affine))) {
    params_free(p);
    return fail(ctx, CTG_ERR_OOM, "params: cudaMalloc failed");
  }
  *out = p;
  return CTG_OK;
}

int32_t generate_parameters(ctg_ctx* ctx, uint64_t ni, uint64_t nw, uint64_t nc, const ctg_csr* const mats[3], const ctg_trapdoor* td, ctg_params** out) {
  if (!td || !out) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: null argument");
  if (ni == 0) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: num_instance includes the constant 1 and must be >= 1");
  const uint64_t m = ni + nw;
  uint64_t n = 1; int log_n = 0;
  while (n < nc + ni) { n <<= 1; log_n++; }
  if (log_n > 28) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: domain larger than 2^28");
  if (m >= (1ull << 32)) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: too many variables");
  Fr alpha, beta, gamma, delta, tau;
  memcpy(&alpha, &td->alpha, sizeof(Fr)); memcpy(&beta, &td->beta, sizeof(Fr)); memcpy(&gamma, &td->gamma, sizeof(Fr));
  memcpy(&delta, &td->delta, sizeof(Fr)); memcpy(&tau, &td->tau, sizeof(Fr));
  G1Affine g1; G2Affine g2;
  memcpy(&g1, &td->g1, sizeof(g1)); memcpy(&g2, &td->g2, sizeof(g2));
  if (gamma.is_zero() || delta.is_zero()) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: gamma and delta must be invertible");
  // t(tau) = tau^n - 1; arkworks samples tau outside the domain, so it is never 0 there
  Fr zt = tau;
  for (int i = 0; i < log_n; i++) zt = zt.sqr();
  zt = zt - Fr::one();
  if (zt.is_zero()) return fail(ctx, CTG_ERR_INVALID, "generate_parameters: tau lies in the evaluation domain");
  const Fr ginv = gamma.inv(), dinv = delta.inv();
  cudaStream_t st = ctx->stream;
  NttDomain* dom;
  CTG_TRY(ntt_get_domain(ctx, log_n, &dom));

  ctg_params* p = nullptr;
  CTG_TRY(params_alloc(ctx, ni, nw, nc, n, &p));
  struct Guard { ctg_params* p; ~Guard() { if (p) params_free(p); } } guard{p};

  // scalars
  DevBuf d_L, d_u, d_v, d_w, d_abc, d_hs, d_single;
  SETUP_ALLOC(ctx, d_L, n * sizeof(Fr));
  SETUP_ALLOC(ctx, d_u, m * sizeof(Fr));
  SETUP_ALLOC(ctx, d_v, m * sizeof(Fr));
  SETUP_ALLOC(ctx, d_w, m * sizeof(Fr));
  SETUP_ALLOC(ctx, d_abc, m * sizeof(Fr));
  SETUP_ALLOC(ctx, d_hs, n * sizeof(Fr));
  SETUP_ALLOC(ctx, d_single, 6 * sizeof(Fr));
  lagrange_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(dom->tw_fwd, n, tau, zt * dom->n_inv, d_L.as<Fr>());
  CTG_KERNEL_CHECK(ctx);
  Fr* uvw[3] = {d_u.as<Fr>(), d_v.as<Fr>(), d_w.as<Fr>()};
  for (int k = 0; k < 3; k++) CTG_TRY(qap_columns(ctx, mats[k], nc, ni, m, d_L.as<Fr>(), k == 0, uvw[k], st));
  abc_kernel<<<(unsigned)((m + 127) / 128), 128, 0, st>>>(uvw[0], uvw[1], uvw[2], m, ni, alpha, beta, ginv, dinv, d_abc.as<Fr>());
  CTG_KERNEL_CHECK(ctx);
  CTG_TRY(fr_pow_table(ctx, d_hs.as<Fr>(), n - 1, tau, zt * dinv, st));  // tau^i t(tau)/delta, i < n - 1
  const Fr singles[6] = {alpha, beta, delta, beta, gamma, delta};
  CTG_CUDA(ctx, cudaMemcpyAsync(d_single.p, singles, sizeof(singles), cudaMemcpyHostToDevice, st));

  // fixed-base multiplications: 16-bit windows for real sizes, 8-bit for toy circuits
  const int c = (m + n) >= (1u << 14) ? 16 : 8;
  const int nwin = (255 + c - 1) / c;
  {
    DevBuf t1;
    SETUP_ALLOC(ctx, t1, ((size_t)nwin << c) * sizeof(G1Affine));
    CTG_TRY(build_table<G1Affine>(ctx, g1, c, nwin, t1.as<G1Affine>(), g1_progression_dev, st));
    CTG_TRY(fixed_base<Fq>(ctx, t1.as<G1Affine>(), c, nwin, uvw[0], m, p->a_query, st));
    CTG_TRY(fixed_base<Fq>(ctx, t1.as<G1Affine>(), c, nwin, uvw[1], m, p->b_g1_query, st));
    CTG_TRY(fixed_base<Fq>(ctx, t1.as<G1Affine>(), c, nwin, d_abc.as<Fr>(), ni, p->gamma_abc_g1, st));
    CTG_TRY(fixed_base<Fq>(ctx, t1.as<G1Affine>(), c, nwin, d_abc.as<Fr>() + ni, nw, p->l_query, st));
    CTG_TRY(fixed_base<Fq>(ctx, t1.as<G1Affine>(), c, nwin, d_hs.as<Fr>(), n - 1, p->h_query, st));
    CTG_TRY(fixed_base<Fq>(ctx, t1.as<G1Affine>(), c, nwin, d_single.as<Fr>(), 3, p->singles_g1, st));
    CTG_CUDA(ctx, cudaStreamSynchronize(st));
  }
  {
    DevBuf t2;
    SETUP_ALLOC(ctx, t2, ((size_t)nwin << c) * sizeof(G2Affine));
    CTG_TRY(build_table<G2Affine>(ctx, g2, c, nwin, t2.as<G2Affine>(), g2_progression_dev, st));
    CTG_TRY(fixed_base<Fq2>(ctx, t2.as<G2Affine>(), c, nwin, uvw[1], m, p->b_g2_query, st));
    CTG_TRY(fixed_base<Fq2>(ctx, t2.as<G2Affine>(), c, nwin, d_single.as<Fr>() + 3, 3, p->singles_g2, st));
    CTG_CUDA(ctx, cudaStreamSynchronize(st));
  }
  guard.p = nullptr;
  *out = p;
  return CTG_OK;
}

}  // namespace ctg

// ---- C ABI ----------------------------------------------------------------------------------------
using namespace ctg;
extern "C" {

int32_t ctg_generate_parameters(ctg_ctx* ctx, uint64_t num_instance, uint64_t num_witness, uint64_t num_constraints, const ctg_csr* a, const ctg_csr* b,
                                const ctg_csr* c, const ctg_trapdoor* trapdoor, ctg_params** out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  CTG_CUDA(ctx, cudaSetDevice(ctx->device));
  const ctg_csr* mats[3] = {a, b, c};
  try {
    return generate_parameters(ctx, num_instance, num_witness, num_constraints, mats, trapdoor, out);
  } catch (const std::bad_alloc&) {
    return fail(ctx, CTG_ERR_OOM, "generate_parameters: host allocation failed");
  }
}

int32_t ctg_params_pk_view(ctg_ctx* ctx, const ctg_params* p, ctg_pk_view* out, const ctg_g1_affine** gamma_abc_g1, uint64_t* gamma_abc_len) {
  if (!ctx) return CTG_ERR_INVALID;
  if (!p || !out) return fail(ctx, CTG_ERR_INVALID, "params_pk_view: null argument");
  out->alpha_g1 = (const ctg_g1_affine*)p->singles_g1;
  out->beta_g1 = (const ctg_g1_affine*)(p->singles_g1 + 1);
  out->delta_g1 = (const ctg_g1_affine*)(p->singles_g1 + 2);
  out->beta_g2 = (const ctg_g2_affine*)p->singles_g2;
  out->delta_g2 = (const ctg_g2_affine*)(p->singles_g2 + 2);
  out->a_query = (const ctg_g1_affine*)p->a_query;       out->a_len = p->m;
  out->b_g1_query = (const ctg_g1_affine*)p->b_g1_query; out->b_g1_len = p->m;
  out->b_g2_query = (const ctg_g2_affine*)p->b_g2_query; out->b_g2_len = p->m;
  out->h_query = (const ctg_g1_affine*)p->h_query;       out->h_len = p->n - 1;
  out->l_query = (const ctg_g1_affine*)p->l_query;       out->l_len = p->nw;
  if (gamma_abc_g1) *gamma_abc_g1 = (const ctg_g1_affine*)p->gamma_abc_g1;
  if (gamma_abc_len) *gamma_abc_len = p->ni;
  return CTG_OK;
}

int32_t ctg_params_download(ctg_ctx* ctx, const ctg_params* p, const ctg_params_out* dst) {
  if (!ctx) return CTG_ERR_INVALID;
  if (!p || !dst) return fail(ctx, CTG_ERR_INVALID, "params_download: null argument");
  CTG_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  struct Item { void* h; const void* d; size_t bytes; };
  const Item items[] = {
      {dst->alpha_g1, p->singles_g1, sizeof(G1Affine)},     {dst->beta_g1, p->singles_g1 + 1, sizeof(G1Affine)},
      {dst->delta_g1, p->singles_g1 + 2, sizeof(G1Affine)}, {dst->beta_g2, p->singles_g2, sizeof(G2Affine)},
      {dst->gamma_g2, p->singles_g2 + 1, sizeof(G2Affine)}, {dst->delta_g2, p->singles_g2 + 2, sizeof(G2Affine)},
      {dst->gamma_abc_g1, p->gamma_abc_g1, p->ni * sizeof(G1Affine)},
      {dst->a_query, p->a_query, p->m * sizeof(G1Affine)},  {dst->b_g1_query, p->b_g1_query, p->m * sizeof(G1Affine)},
      {dst->b_g2_query, p->b_g2_query, p->m * sizeof(G2Affine)},
      {dst->h_query, p->h_query, (p->n - 1) * sizeof(G1Affine)},
      {dst->l_query, p->l_query, p->nw * sizeof(G1Affine)},
  };
  for (const Item& it : items)
    if (it.h && it.bytes) CTG_CUDA(ctx, cudaMemcpyAsync(it.h, it.d, it.bytes, cudaMemcpyDeviceToHost, st));
  CTG_CUDA(ctx, cudaStreamSynchronize(st));
  return CTG_OK;
}

void ctg_params_free(ctg_ctx* ctx, ctg_params* p) {
  if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
  params_free(p);
}

}  // extern "C"
