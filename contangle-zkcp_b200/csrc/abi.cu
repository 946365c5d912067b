// C ABI of libctg_groth16.so (include/ctg_groth16.h): context management, host<->device
// staging and argument checking around the kernels in ntt.cu / msm*.cu / prove.cu / fields.cu.
// No CPU fallback exists anywhere below this file: every compute entry point launches CUDA
// kernels on the context's device or fails with an error code.
#include <algorithm>
#include <cstring>

#include "ctg_internal.h"

struct ctg_pk;
struct ctg_r1cs;

namespace ctg {
int32_t field_op_dev(ctg_ctx* ctx, int field, int op, const void* a, const void* b, void* out, size_t n, cudaStream_t st);
size_t field_elem_bytes(int field);
int32_t imad_probe(ctg_ctx* ctx, double* imad_per_s, double* fq_mul_per_s);
int32_t pk_upload(ctg_ctx* ctx, const ctg_pk_view* v, uint64_t ni, uint64_t nc, int copies, int c_bits, uint64_t lo_num, uint64_t hi_num, uint64_t den, ctg_pk** out);
int32_t prove_batch(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, uint32_t n_proofs, const ctg_fr* z_batch, const ctg_fr* rs_batch, ctg_proof* out);
int32_t prove_partial(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const ctg_fr* rr, const ctg_fr* ss, void* out_partials, ctg_timings* tm);
int32_t prove_partial_begin(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const ctg_fr* rr, const ctg_fr* ss, uint32_t chain_mask, void* d_abc);
int32_t prove_partial_finish(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* rr, const ctg_fr* ss, void* d_abc, void* out_partials);
int32_t prove_combine(ctg_ctx* ctx, const void* host_partials, uint32_t k, const ctg_fr* rr, const ctg_fr* ss, ctg_proof* out);
void lanes_destroy(ctg_ctx* ctx);
void pk_free(ctg_pk* pk);
int32_t r1cs_upload(ctg_ctx* ctx, uint64_t ni, uint64_t nw, uint64_t nc, const ctg_csr* const mats[3], ctg_r1cs** out);
void r1cs_free(ctg_r1cs* r);
int32_t witness_map_host(ctg_ctx* ctx, const ctg_r1cs* r, const ctg_fr* z, ctg_fr* h_out);
int32_t prove(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const ctg_fr* rr, const ctg_fr* ss, ctg_proof* out, ctg_timings* tm);
uint64_t pk_num_vars(const ctg_pk* pk);
// multi.cu
int32_t group_create(ctg_ctx* primary, const int* device_ids, int n);
void group_destroy(ctg_ctx* primary);
uint64_t group_launches(const Group* g);
int32_t group_pk_upload(ctg_ctx* ctx, const ctg_pk_view* v, uint64_t ni, uint64_t nc, int copies, int c_bits, ctg_pk** out);
void group_pk_free(ctg_ctx* ctx, ctg_pk* pk);
int32_t group_r1cs_upload(ctg_ctx* ctx, uint64_t ni, uint64_t nw, uint64_t nc, const ctg_csr* const mats[3], ctg_r1cs** out);
void group_r1cs_free(ctg_ctx* ctx, ctg_r1cs* r);
int32_t group_prove(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const ctg_fr* rr, const ctg_fr* ss, ctg_proof* out, ctg_timings* tm);
int32_t group_bases_new(ctg_ctx* ctx, int group, uint64_t n, const std::function<int32_t(int, uint64_t, uint64_t, ctg_bases**)>& make, ctg_bases** out);
void group_bases_free(ctg_ctx* ctx, ctg_bases* h);
int32_t group_msm(ctg_ctx* ctx, const ctg_bases* bases, int group, const void* scalars, bool from_host, uint64_t n, uint32_t flags, void* out_affine, void* d_out_xyzz);

int32_t fail(ctg_ctx* ctx, int32_t code, const std::string& msg) {
  if (ctx) ctx->last_error = msg;
  return code;
}
int32_t ensure_scratch(ctg_ctx* ctx, size_t bytes) {
  if (bytes <= ctx->scratch_bytes) return CTG_OK;
  if (ctx->scratch) {
    CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    cudaMemset(ctx->scratch, 0, ctx->scratch_bytes);   // it held assignments, randomness or trapdoor powers
    cudaFree(ctx->scratch);
    ctx->scratch = nullptr;
    ctx->scratch_bytes = 0;
  }
  size_t want = bytes + bytes / 8 + 4096;
  cudaError_t e = cudaMalloc(&ctx->scratch, want);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, CTG_ERR_OOM, "out of device memory for scratch"); }
  ctx->scratch_bytes = want;
  return CTG_OK;
}
int32_t ensure_pinned(ctg_ctx* ctx, size_t bytes) {
  if (bytes <= ctx->pinned_bytes) return CTG_OK;
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  ctx->pinned = nullptr;
  ctx->pinned_bytes = 0;
  CTG_CUDA(ctx, cudaMallocHost(&ctx->pinned, bytes));
  ctx->pinned_bytes = bytes;
  return CTG_OK;
}
static ctg_timings g_last_timings;  // per process is enough: one in-flight call per context
}  // namespace ctg

using namespace ctg;

static thread_local std::string g_null_ctx_error = "null context";

static int32_t msm_check(ctg_ctx* ctx, const ctg_bases* bases, int group, uint64_t n) {
  if (!ctx) return CTG_ERR_INVALID;
  if (!bases) return fail(ctx, CTG_ERR_INVALID, "msm: null bases");
  if (!bases->parts.empty()) return fail(ctx, CTG_ERR_STATE, "msm: the base table belongs to a multi-device context");
  if (bases->b.group != group) return fail(ctx, CTG_ERR_INVALID, "msm: base table is for the other group");
  if (n > bases->b.n) return fail(ctx, CTG_ERR_INVALID, "msm: more scalars than bases");
  return CTG_OK;
}
// ---- split MSM ---------------------------------------------------------------------------------------
// A large MSM is cut into point-range parts, each a complete MSM (sort, accumulate, reduce) on its own stream and workspace;
// the partial sums are added at the end.  The accumulate kernels fill the GPU one after the other, while the other parts'
// scalar uploads (host entry points), bucket sorts (HBM / atomics) and latency-bound bucket reductions run beside them.
// The same overlap the prover gets from running its five MSMs on five streams.
static int32_t split_init(ctg_ctx* ctx) {
  if (ctx->split_start) return CTG_OK;
  for (int i = 0; i < ctg_ctx::kSplit; i++) {
    CTG_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->split_st[i], cudaStreamNonBlocking));
    CTG_CUDA(ctx, cudaEventCreateWithFlags(&ctx->split_done[i], cudaEventDisableTiming));
  }
  CTG_CUDA(ctx, cudaMalloc(&ctx->split_part, ctg_ctx::kSplit * sizeof(G2XYZZ)));
  CTG_CUDA(ctx, cudaEventCreateWithFlags(&ctx->split_start, cudaEventDisableTiming));
  return CTG_OK;
}
static void split_destroy(ctg_ctx* ctx) {
  for (int i = 0; i < ctg_ctx::kSplit; i++) {
    if (ctx->split_st[i]) { cudaStreamSynchronize(ctx->split_st[i]); cudaStreamDestroy(ctx->split_st[i]); }
    if (ctx->split_done[i]) cudaEventDestroy(ctx->split_done[i]);
    msm_ws_free(ctx->split_ws[i]);
  }
  if (ctx->split_start) cudaEventDestroy(ctx->split_start);
  if (ctx->split_part) cudaFree(ctx->split_part);
}
static int split_parts(ctg_ctx* ctx, uint64_t n, uint32_t flags, bool from_host) {
  if (flags & CTG_MSM_TIMED) return 1;   // stage timings are only meaningful for the serial schedule
  if (ctx->msm_split >= 1) return n >= (uint64_t)ctx->msm_split ? ctx->msm_split : 1;
  // measured on B200 at 2^22 (tools/split_sweep.py): every extra part costs ~0.9 ms (a second bucket reduction and fix-up that only
  // partly hide under the other part's accumulate kernel); two parts pay when there is an upload to hide, none otherwise
  if (n < (1ull << 21) || !from_host) return 1;
  return 2;
}
// d_scalars: device buffer of n scalars; when host_scalars != nullptr it is filled from there first (part by part).
template <class XY>
static int32_t msm_any(ctg_ctx* ctx, const ctg_bases* bases, int group, void* d_scalars, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, XY* d_out) {
  const bool mont = (flags & CTG_SCALARS_MONT) != 0;
  const int parts = split_parts(ctx, n, flags, host_scalars != nullptr);
  auto run = [&](MsmWorkspace& ws, const Bases& b, const uint32_t* sc, uint64_t cnt, void* out, cudaStream_t st, float* tm) -> int32_t {
    return group == 1 ? msm_g1_run(ctx, ws, b, sc, cnt, mont, (G1XYZZ*)out, st, tm) : msm_g2_run(ctx, ws, b, sc, cnt, mont, (G2XYZZ*)out, st, tm);
  };
  if (parts <= 1) {
    if (host_scalars && n) CTG_CUDA(ctx, cudaMemcpyAsync(d_scalars, host_scalars, n * sizeof(Fr), cudaMemcpyHostToDevice, ctx->stream));
    float tm[3];
    const bool timed = (flags & CTG_MSM_TIMED) != 0;
    int32_t rc = run(ctx->ws[0], bases->b, (const uint32_t*)d_scalars, n, d_out, ctx->stream, timed ? tm : nullptr);
    if (rc == CTG_OK && timed) { g_last_timings.msm_sort_ms = tm[0]; g_last_timings.msm_accumulate_ms = tm[1]; g_last_timings.msm_reduce_ms = tm[2]; }
    return rc;
  }
  CTG_TRY(split_init(ctx));
  CTG_CUDA(ctx, cudaEventRecord(ctx->split_start, ctx->stream));
  const size_t pb = group == 1 ? sizeof(G1Affine) : sizeof(G2Affine);
  const uint64_t stride = bases->b.stride ? bases->b.stride : bases->b.n;
  for (int p = 0; p < parts; p++) {
    const uint64_t lo = n * (uint64_t)p / parts, hi = n * (uint64_t)(p + 1) / parts, cnt = hi - lo;
    cudaStream_t st = ctx->split_st[p];
    CTG_CUDA(ctx, cudaStreamWaitEvent(st, ctx->split_start, 0));
    Fr* sc = (Fr*)d_scalars + lo;
    if (host_scalars && cnt) CTG_CUDA(ctx, cudaMemcpyAsync(sc, (const Fr*)host_scalars + lo, cnt * sizeof(Fr), cudaMemcpyHostToDevice, st));
    Bases view = bases->b;
    view.pts = (char*)bases->b.pts + lo * pb;
    view.n = cnt;
    view.stride = stride;
    CTG_TRY(run(ctx->split_ws[p], view, (const uint32_t*)sc, cnt, (XY*)ctx->split_part + p, st, nullptr));
    CTG_CUDA(ctx, cudaEventRecord(ctx->split_done[p], st));
    CTG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->split_done[p], 0));
  }
  if (group == 1) return g1_combine_dev(ctx, (const G1XYZZ*)ctx->split_part, (uint32_t)parts, nullptr, (G1XYZZ*)d_out, ctx->stream);
  return g2_combine_dev(ctx, (const G2XYZZ*)ctx->split_part, (uint32_t)parts, nullptr, (G2XYZZ*)d_out, ctx->stream);
}

template <class XY, class AF>
static int32_t msm_host(ctg_ctx* ctx, const ctg_bases* bases, int group, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, void* out_affine,
                        void* out_xyzz) {
  CTG_TRY(msm_check(ctx, bases, group, n));
  if (n && !host_scalars) return fail(ctx, CTG_ERR_INVALID, "msm: null scalars");
  const size_t sb = n * sizeof(Fr);
  CTG_TRY(ensure_scratch(ctx, sb + sizeof(XY) + sizeof(AF) + 128));
  char* d = (char*)ctx->scratch;
  XY* d_res = (XY*)(d + ((sb + 63) & ~(size_t)63));
  AF* d_aff = (AF*)(d_res + 1);
  CTG_TRY(msm_any<XY>(ctx, bases, group, d, host_scalars, n, flags, d_res));
  if (out_xyzz) CTG_CUDA(ctx, cudaMemcpyAsync(out_xyzz, d_res, sizeof(XY), cudaMemcpyDeviceToHost, ctx->stream));
  if (out_affine) {
    if (group == 1) CTG_TRY(g1_to_affine_dev(ctx, (const G1XYZZ*)d_res, (G1Affine*)d_aff, 1, ctx->stream));
    else CTG_TRY(g2_to_affine_dev(ctx, (const G2XYZZ*)d_res, (G2Affine*)d_aff, 1, ctx->stream));
    CTG_CUDA(ctx, cudaMemcpyAsync(out_affine, d_aff, sizeof(AF), cudaMemcpyDeviceToHost, ctx->stream));
  }
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}

extern "C" {

int32_t ctg_ctx_create(const int* device_ids, int n_devices, ctg_ctx** out) {
  if (!out) return CTG_ERR_INVALID;
  *out = nullptr;
  if (n_devices < 1 || n_devices > kMaxDevices || !device_ids) return CTG_ERR_INVALID;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return CTG_ERR_CUDA;  // no device: no fallback
  if (device_ids[0] < 0 || device_ids[0] >= count) return CTG_ERR_INVALID;
  if (cudaSetDevice(device_ids[0]) != cudaSuccess) return CTG_ERR_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device_ids[0]) != cudaSuccess) return CTG_ERR_CUDA;
  if (prop.major != 10) return CTG_ERR_STATE;  // kernels are built for sm_100a only
  ctg_ctx* ctx = new ctg_ctx();
  ctx->device = device_ids[0];
  ctx->sm_count = prop.multiProcessorCount;
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return CTG_ERR_CUDA; }
  ctx->own_stream = true;
  if (n_devices > 1) {   // one process, several GPUs: children + worker threads + peer access (multi.cu)
    const int32_t rc = group_create(ctx, device_ids, n_devices);
    if (rc != CTG_OK) { ctg_ctx_destroy(ctx); return rc; }
  }
  *out = ctx;
  return CTG_OK;
}
int32_t ctg_ctx_device_count(ctg_ctx* ctx) { return !ctx ? 0 : (ctx->group ? ctx->group->n : 1); }

void ctg_ctx_destroy(ctg_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->group) group_destroy(ctx);
  for (auto& kv : ctx->domains) {
    cudaFree(kv.second.tw_fwd); cudaFree(kv.second.tw_inv); cudaFree(kv.second.gpow); cudaFree(kv.second.gpow_n); cudaFree(kv.second.gipow_n);
  }
  lanes_destroy(ctx);
  for (auto& w : ctx->ws) msm_ws_free(w);
  split_destroy(ctx);
  if (ctx->scratch) { cudaMemset(ctx->scratch, 0, ctx->scratch_bytes); cudaFree(ctx->scratch); }   // secrets do not outlive the context in HBM
  if (ctx->pinned) { memset(ctx->pinned, 0, ctx->pinned_bytes); cudaFreeHost(ctx->pinned); }
  if (ctx->events_ready) for (auto& e : ctx->ev) cudaEventDestroy(e);
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char* ctg_last_error(ctg_ctx* ctx) { return ctx ? ctx->last_error.c_str() : g_null_ctx_error.c_str(); }

int32_t ctg_ctx_set_stream(ctg_ctx* ctx, void* cuda_stream) {
  if (!ctx) return CTG_ERR_INVALID;
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  ctx->stream = (cudaStream_t)cuda_stream;
  ctx->own_stream = false;
  return CTG_OK;
}
int32_t ctg_sync(ctg_ctx* ctx) {
  if (!ctx) return CTG_ERR_INVALID;
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}
uint64_t ctg_launch_count(ctg_ctx* ctx) { return ctx ? ctx->launches + (ctx->group ? group_launches(ctx->group) : 0) : 0; }
int32_t ctg_last_timings(ctg_ctx* ctx, ctg_timings* out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  *out = g_last_timings;
  return CTG_OK;
}

int32_t ctg_dev_alloc(ctg_ctx* ctx, size_t bytes, void** d_out) {
  if (!ctx || !d_out) return CTG_ERR_INVALID;
  cudaError_t e = cudaMalloc(d_out, bytes ? bytes : 1);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, CTG_ERR_OOM, "cudaMalloc failed"); }
  return CTG_OK;
}
int32_t ctg_dev_free(ctg_ctx* ctx, void* d_ptr) {
  if (!ctx) return CTG_ERR_INVALID;
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  CTG_CUDA(ctx, cudaFree(d_ptr));
  return CTG_OK;
}
int32_t ctg_h2d(ctg_ctx* ctx, void* d_dst, const void* h_src, size_t bytes) {
  if (!ctx) return CTG_ERR_INVALID;
  CTG_CUDA(ctx, cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}
int32_t ctg_d2h(ctg_ctx* ctx, void* h_dst, const void* d_src, size_t bytes) {
  if (!ctx) return CTG_ERR_INVALID;
  CTG_CUDA(ctx, cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}

// ---- fields -----------------------------------------------------------------------------------
int32_t ctg_field_op_batch(ctg_ctx* ctx, int field, int op, const void* a, const void* b, void* out, size_t n) {
  if (!ctx) return CTG_ERR_INVALID;
  const size_t eb = field_elem_bytes(field);
  if (!eb) return fail(ctx, CTG_ERR_INVALID, "unknown field");
  if (n == 0) return CTG_OK;
  if (!a || !b || !out) return fail(ctx, CTG_ERR_INVALID, "null buffer");
  CTG_TRY(ensure_scratch(ctx, 3 * n * eb));
  char* d = (char*)ctx->scratch;
  CTG_CUDA(ctx, cudaMemcpyAsync(d, a, n * eb, cudaMemcpyHostToDevice, ctx->stream));
  CTG_CUDA(ctx, cudaMemcpyAsync(d + n * eb, b, n * eb, cudaMemcpyHostToDevice, ctx->stream));
  CTG_TRY(field_op_dev(ctx, field, op, d, d + n * eb, d + 2 * n * eb, n, ctx->stream));
  CTG_CUDA(ctx, cudaMemcpyAsync(out, d + 2 * n * eb, n * eb, cudaMemcpyDeviceToHost, ctx->stream));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}
int32_t ctg_fr_mul_batch(ctg_ctx* ctx, const ctg_fr* a, const ctg_fr* b, ctg_fr* out, size_t n) {
  return ctg_field_op_batch(ctx, CTG_FIELD_FR, CTG_OP_MUL, a, b, out, n);
}
int32_t ctg_fq_mul_batch(ctg_ctx* ctx, const ctg_fq* a, const ctg_fq* b, ctg_fq* out, size_t n) {
  return ctg_field_op_batch(ctx, CTG_FIELD_FQ, CTG_OP_MUL, a, b, out, n);
}
int32_t ctg_imad_peak_probe(ctg_ctx* ctx, double* imad_per_s, double* fq_mul_per_s) {
  if (!ctx) return CTG_ERR_INVALID;
  return imad_probe(ctx, imad_per_s, fq_mul_per_s);
}

// ---- NTT --------------------------------------------------------------------------------------
int32_t ctg_ntt_fr_dev(ctg_ctx* ctx, void* d_data, uint32_t log_n, int inverse, int coset) {
  if (!ctx || !d_data) return ctx ? fail(ctx, CTG_ERR_INVALID, "ntt: null buffer") : CTG_ERR_INVALID;
  return ntt_run(ctx, (Fr*)d_data, (int)log_n, inverse, coset, ctx->stream);
}
int32_t ctg_ntt_fr(ctg_ctx* ctx, ctg_fr* host_data, uint32_t log_n, int inverse, int coset) {
  if (!ctx || !host_data) return ctx ? fail(ctx, CTG_ERR_INVALID, "ntt: null buffer") : CTG_ERR_INVALID;
  if (log_n > 28) return fail(ctx, CTG_ERR_INVALID, "ntt: log_n out of range [0, 28]");
  const size_t bytes = ((size_t)1 << log_n) * sizeof(Fr);
  CTG_TRY(ensure_scratch(ctx, bytes));
  CTG_CUDA(ctx, cudaMemcpyAsync(ctx->scratch, host_data, bytes, cudaMemcpyHostToDevice, ctx->stream));
  CTG_TRY(ntt_run(ctx, (Fr*)ctx->scratch, (int)log_n, inverse, coset, ctx->stream));
  CTG_CUDA(ctx, cudaMemcpyAsync(host_data, ctx->scratch, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}

// ---- bases ------------------------------------------------------------------------------------
static int32_t bases_new(ctg_ctx* ctx, int group, uint64_t n, int copies, int c_bits, ctg_bases** out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  if (c_bits != 0 && (c_bits < 2 || c_bits > 21)) return fail(ctx, CTG_ERR_INVALID, "bases: c_bits must be 0 or in [2, 21]");
  if (copies < 0) return fail(ctx, CTG_ERR_INVALID, "bases: copies must be >= 0");
  ctg_bases* h = new ctg_bases();
  h->b.group = group;
  h->b.n = n;
  h->b.plan = msm_make_plan(n, copies, c_bits, group == 1 ? sizeof(G1Affine) : sizeof(G2Affine));
  if (n * (uint64_t)h->b.plan.copies >= (1ull << 31)) { delete h; return fail(ctx, CTG_ERR_INVALID, "bases: n * copies must be < 2^31"); }
  const size_t eb = group == 1 ? sizeof(G1Affine) : sizeof(G2Affine);
  const size_t bytes = (size_t)h->b.plan.copies * (n ? n : 1) * eb;
  cudaError_t e = cudaMalloc(&h->b.pts, bytes);
  if (e != cudaSuccess) { cudaGetLastError(); delete h; return fail(ctx, CTG_ERR_OOM, "bases: out of device memory (fewer copies?)"); }
  *out = h;
  return CTG_OK;
}
static int32_t bases_finish(ctg_ctx* ctx, ctg_bases* h) {
  if (h->b.n == 0) { CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); return CTG_OK; }   // an empty table (e.g. a trailing shard of a tiny query)
  int32_t rc = h->b.group == 1 ? g1_precompute(ctx, h->b, ctx->stream) : g2_precompute(ctx, h->b, ctx->stream);
  if (rc != CTG_OK) return rc;
  CTG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return CTG_OK;
}
// one device: the table of one context
static int32_t bases_upload_one(ctg_ctx* ctx, int group, const void* host_points, uint64_t n, int copies, int c_bits, ctg_bases** out) {
  CTG_TRY(bases_new(ctx, group, n, copies, c_bits, out));
  const size_t eb = group == 1 ? sizeof(G1Affine) : sizeof(G2Affine);
  if (n) CTG_CUDA(ctx, cudaMemcpyAsync((*out)->b.pts, host_points, n * eb, cudaMemcpyDefault, ctx->stream));
  return bases_finish(ctx, *out);
}
static int32_t bases_progression_one(ctg_ctx* ctx, int group, const void* gen, const ctg_fr* a, const ctg_fr* b, uint64_t n, int copies, int c_bits, ctg_bases** out) {
  CTG_TRY(bases_new(ctx, group, n, copies, c_bits, out));
  Fr fa, fb;
  memcpy(&fa, a, sizeof(fa)); memcpy(&fb, b, sizeof(fb));
  if (n && group == 1) { G1Affine gg; memcpy(&gg, gen, sizeof(gg)); CTG_TRY(g1_progression_dev(ctx, gg, fa, fb, n, (G1Affine*)(*out)->b.pts, ctx->stream)); }
  if (n && group == 2) { G2Affine gg; memcpy(&gg, gen, sizeof(gg)); CTG_TRY(g2_progression_dev(ctx, gg, fa, fb, n, (G2Affine*)(*out)->b.pts, ctx->stream)); }
  return bases_finish(ctx, *out);
}
// multi-device context: every device keeps the point range [n d / N, n (d + 1) / N) as an ordinary table of its own context
static int32_t group_bases_upload(ctg_ctx* ctx, int group, const void* host_points, uint64_t n, int copies, int c_bits, ctg_bases** out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  const size_t eb = group == 1 ? sizeof(G1Affine) : sizeof(G2Affine);
  Group* g = ctx->group;
  return group_bases_new(ctx, group, n, [=](int d, uint64_t lo, uint64_t cnt, ctg_bases** part) -> int32_t {
    const void* src = (const char*)host_points + lo * eb;
    return bases_upload_one(g->ctx[d], group, src, cnt, copies, c_bits, part);   // not the public entry point: ctx[0] is the multi-device context itself
  }, out);
}
// P_i = (a + i b) g: device d generates its own range from a + lo b
static int32_t group_bases_progression(ctg_ctx* ctx, int group, const void* gen, const ctg_fr* a, const ctg_fr* b, uint64_t n, int copies, int c_bits, ctg_bases** out) {
  if (!ctx || !out || !gen || !a || !b) return ctx ? fail(ctx, CTG_ERR_INVALID, "progression: null argument") : CTG_ERR_INVALID;
  Group* g = ctx->group;
  Fr fa, fb;
  memcpy(&fa, a, sizeof(fa)); memcpy(&fb, b, sizeof(fb));
  return group_bases_new(ctx, group, n, [=](int d, uint64_t lo, uint64_t cnt, ctg_bases** part) -> int32_t {
    Fr idx = Fr::zero();
    idx.v[0] = (uint32_t)lo; idx.v[1] = (uint32_t)(lo >> 32);
    const Fr a_d = fa + fb * idx.to_mont();
    ctg_fr aw; memcpy(&aw, &a_d, sizeof(aw));
    return bases_progression_one(g->ctx[d], group, gen, &aw, b, cnt, copies, c_bits, part);
  }, out);
}
int32_t ctg_bases_upload_g1(ctg_ctx* ctx, const ctg_g1_affine* host_points, uint64_t n, int copies, int c_bits, ctg_bases** out) {
  if (ctx && ctx->group) return group_bases_upload(ctx, 1, host_points, n, copies, c_bits, out);
  return bases_upload_one(ctx, 1, host_points, n, copies, c_bits, out);
}
int32_t ctg_bases_upload_g2(ctg_ctx* ctx, const ctg_g2_affine* host_points, uint64_t n, int copies, int c_bits, ctg_bases** out) {
  if (ctx && ctx->group) return group_bases_upload(ctx, 2, host_points, n, copies, c_bits, out);
  return bases_upload_one(ctx, 2, host_points, n, copies, c_bits, out);
}
int32_t ctg_bases_progression_g1(ctg_ctx* ctx, const ctg_g1_affine* g, const ctg_fr* a, const ctg_fr* b, uint64_t n, int copies, int c_bits, ctg_bases** out) {
  if (!g || !a || !b) return ctx ? fail(ctx, CTG_ERR_INVALID, "progression: null argument") : CTG_ERR_INVALID;
  if (ctx && ctx->group) return group_bases_progression(ctx, 1, g, a, b, n, copies, c_bits, out);
  return bases_progression_one(ctx, 1, g, a, b, n, copies, c_bits, out);
}
int32_t ctg_bases_progression_g2(ctg_ctx* ctx, const ctg_g2_affine* g, const ctg_fr* a, const ctg_fr* b, uint64_t n, int copies, int c_bits, ctg_bases** out) {
  if (!g || !a || !b) return ctx ? fail(ctx, CTG_ERR_INVALID, "progression: null argument") : CTG_ERR_INVALID;
  if (ctx && ctx->group) return group_bases_progression(ctx, 2, g, a, b, n, copies, c_bits, out);
  return bases_progression_one(ctx, 2, g, a, b, n, copies, c_bits, out);
}
int32_t ctg_bases_download(ctg_ctx* ctx, const ctg_bases* bases, uint64_t first, uint64_t count, void* host_points) {
  if (!ctx || !bases || !host_points) return CTG_ERR_INVALID;
  if (!bases->parts.empty()) {   // multi-device table: points [first, first + count) of the original (copy 0) order, part by part
    if (!ctx->group) return fail(ctx, CTG_ERR_STATE, "bases_download: the table belongs to a multi-device context");
    if (first + count > bases->b.n) return fail(ctx, CTG_ERR_INVALID, "bases_download: range out of bounds");
    const size_t eb = bases->b.group == 1 ? sizeof(G1Affine) : sizeof(G2Affine);
    for (int d = 0; d < ctx->group->n; d++) {
      const uint64_t lo = std::max<uint64_t>(first, bases->part_lo[d]), hi = std::min<uint64_t>(first + count, bases->part_lo[d + 1]);
      if (hi <= lo) continue;
      ctg_ctx* c = ctx->group->ctx[d];
      const ctg_bases* part = bases->parts[d];
      char* dst = (char*)host_points + (lo - first) * eb;
      const uint64_t off = lo - bases->part_lo[d], cnt = hi - lo;
      CTG_TRY(group_run_on(ctx->group, d, [=]() { return ctg_bases_download(c, part, off, cnt, dst); }));
    }
    return CTG_OK;
  }
  const uint64_t total = bases->b.n * (uint64_t)bases->b.plan.copies;
  if (first + count > total) return fail(ctx, CTG_ERR_INVALID, "bases_download: range out of bounds");
  const size_t eb = bases->b.group == 1 ? sizeof(G1Affine) : sizeof(G2Affine);
  return ctg_d2h(ctx, host_points, (const char*)bases->b.pts + first * eb, count * eb);
}
void ctg_bases_free(ctg_ctx* ctx, ctg_bases* bases) {
  if (!bases) return;
  if (!bases->parts.empty()) { if (ctx && ctx->group) { cudaStreamSynchronize(ctx->stream); group_bases_free(ctx, bases); } return; }
  if (ctx) cudaStreamSynchronize(ctx->stream);
  cudaFree(bases->b.pts);
  delete bases;
}
uint64_t ctg_bases_len(const ctg_bases* bases) { return bases ? bases->b.n : 0; }
int32_t ctg_ctx_set_ba_levels(ctg_ctx* ctx, int levels) {
  if (!ctx) return CTG_ERR_INVALID;
  if (levels < -1 || levels > 6) return fail(ctx, CTG_ERR_INVALID, "ba_levels must be in [-1, 6]");
  ctx->ba_levels = levels;
  return CTG_OK;
}
int32_t ctg_ctx_set_msm_split(ctg_ctx* ctx, int parts) {
  if (!ctx) return CTG_ERR_INVALID;
  if (parts != -1 && (parts < 1 || parts > ctg_ctx::kSplit)) return fail(ctx, CTG_ERR_INVALID, "msm_split must be -1 or in [1, 4]");
  ctx->msm_split = parts;
  return CTG_OK;
}
int32_t ctg_bases_plan(const ctg_bases* bases, ctg_msm_plan* out) {
  if (!bases || !out) return CTG_ERR_INVALID;
  const MsmPlan& p = bases->b.plan;
  out->window_bits = p.c; out->windows = p.nwin; out->copies = p.copies; out->bucket_sets = p.nsets; out->buckets_per_set = p.nbuckets_per_set;
  return CTG_OK;
}

// ---- MSM --------------------------------------------------------------------------------------
int32_t ctg_msm_g1_dev(ctg_ctx* ctx, const ctg_bases* bases, const void* d_scalars, uint64_t n, uint32_t flags, void* d_out_xyzz) {
  if (ctx && ctx->group && bases) return group_msm(ctx, bases, 1, d_scalars, false, n, flags, nullptr, d_out_xyzz);
  CTG_TRY(msm_check(ctx, bases, 1, n));
  return msm_any<G1XYZZ>(ctx, bases, 1, const_cast<void*>(d_scalars), nullptr, n, flags, (G1XYZZ*)d_out_xyzz);
}
int32_t ctg_msm_g2_dev(ctg_ctx* ctx, const ctg_bases* bases, const void* d_scalars, uint64_t n, uint32_t flags, void* d_out_xyzz) {
  if (ctx && ctx->group && bases) return group_msm(ctx, bases, 2, d_scalars, false, n, flags, nullptr, d_out_xyzz);
  CTG_TRY(msm_check(ctx, bases, 2, n));
  return msm_any<G2XYZZ>(ctx, bases, 2, const_cast<void*>(d_scalars), nullptr, n, flags, (G2XYZZ*)d_out_xyzz);
}

int32_t ctg_msm_g1(ctg_ctx* ctx, const ctg_bases* bases, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, ctg_g1_affine* out) {
  if (!out) return ctx ? fail(ctx, CTG_ERR_INVALID, "msm: null output") : CTG_ERR_INVALID;
  if (ctx && ctx->group && bases) return group_msm(ctx, bases, 1, host_scalars, true, n, flags, out, nullptr);
  return msm_host<G1XYZZ, G1Affine>(ctx, bases, 1, host_scalars, n, flags, out, nullptr);
}
int32_t ctg_msm_g2(ctg_ctx* ctx, const ctg_bases* bases, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, ctg_g2_affine* out) {
  if (!out) return ctx ? fail(ctx, CTG_ERR_INVALID, "msm: null output") : CTG_ERR_INVALID;
  if (ctx && ctx->group && bases) return group_msm(ctx, bases, 2, host_scalars, true, n, flags, out, nullptr);
  return msm_host<G2XYZZ, G2Affine>(ctx, bases, 2, host_scalars, n, flags, out, nullptr);
}
int32_t ctg_msm_g1_partial(ctg_ctx* ctx, const ctg_bases* bases, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, ctg_g1_xyzz* out) {
  if (!out) return ctx ? fail(ctx, CTG_ERR_INVALID, "msm: null output") : CTG_ERR_INVALID;
  return msm_host<G1XYZZ, G1Affine>(ctx, bases, 1, host_scalars, n, flags, nullptr, out);
}
int32_t ctg_msm_g2_partial(ctg_ctx* ctx, const ctg_bases* bases, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, ctg_g2_xyzz* out) {
  if (!out) return ctx ? fail(ctx, CTG_ERR_INVALID, "msm: null output") : CTG_ERR_INVALID;
  return msm_host<G2XYZZ, G2Affine>(ctx, bases, 2, host_scalars, n, flags, nullptr, out);
}

int32_t ctg_host_alloc(ctg_ctx* ctx, size_t bytes, void** h_out) {
  if (!ctx || !h_out) return CTG_ERR_INVALID;
  *h_out = nullptr;
  if (cudaMallocHost(h_out, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return fail(ctx, CTG_ERR_OOM, "host_alloc: cudaMallocHost failed"); }
  return CTG_OK;
}
int32_t ctg_host_free(ctg_ctx* ctx, void* h_ptr) {
  if (!ctx) return CTG_ERR_INVALID;
  if (h_ptr) CTG_CUDA(ctx, cudaFreeHost(h_ptr));
  return CTG_OK;
}
int32_t ctg_prepare_inputs(ctg_ctx* ctx, const ctg_bases* gamma_abc_g1, const ctg_fr* public_inputs, uint64_t n_inputs, ctg_g1_affine* out) {
  if (!ctx || !gamma_abc_g1 || !out || (n_inputs && !public_inputs)) return ctx ? fail(ctx, CTG_ERR_INVALID, "prepare_inputs: null argument") : CTG_ERR_INVALID;
  if (gamma_abc_g1->b.group != 1 || gamma_abc_g1->b.n != n_inputs + 1)
    return fail(ctx, CTG_ERR_INVALID, "prepare_inputs: gamma_abc_g1 must hold exactly n_inputs + 1 G1 points (malformed verifying key)");
  // scalars 1 || public_inputs select gamma_abc_g1[0] and weigh the rest: one MSM over the whole table
  CTG_TRY(ensure_pinned(ctx, (n_inputs + 1) * sizeof(Fr)));
  Fr* h = (Fr*)ctx->pinned;
  h[0] = Fr::one();
  if (n_inputs) memcpy(h + 1, public_inputs, n_inputs * sizeof(Fr));
  return msm_host<G1XYZZ, G1Affine>(ctx, gamma_abc_g1, 1, (const ctg_fr*)h, n_inputs + 1, CTG_SCALARS_MONT, out, nullptr);
}

int32_t ctg_g1_combine_dev(ctg_ctx* ctx, const void* d_partials, uint32_t k, void* d_out_affine) {
  if (!ctx || !d_partials || !d_out_affine) return CTG_ERR_INVALID;
  return g1_combine_dev(ctx, (const G1XYZZ*)d_partials, k, (G1Affine*)d_out_affine, nullptr, ctx->stream);
}
int32_t ctg_g2_combine_dev(ctg_ctx* ctx, const void* d_partials, uint32_t k, void* d_out_affine) {
  if (!ctx || !d_partials || !d_out_affine) return CTG_ERR_INVALID;
  return g2_combine_dev(ctx, (const G2XYZZ*)d_partials, k, (G2Affine*)d_out_affine, nullptr, ctx->stream);
}
int32_t ctg_g1_combine(ctg_ctx* ctx, const ctg_g1_xyzz* host_partials, uint32_t k, ctg_g1_affine* out) {
  if (!ctx || !host_partials || !out) return CTG_ERR_INVALID;
  CTG_TRY(ensure_scratch(ctx, (size_t)k * sizeof(G1XYZZ) + sizeof(G1Affine) + 64));
  char* d = (char*)ctx->scratch;
  G1Affine* d_out = (G1Affine*)(d + (size_t)k * sizeof(G1XYZZ));
  CTG_CUDA(ctx, cudaMemcpyAsync(d, host_partials, (size_t)k * sizeof(G1XYZZ), cudaMemcpyHostToDevice, ctx->stream));
  CTG_TRY(g1_combine_dev(ctx, (const G1XYZZ*)d, k, d_out, nullptr, ctx->stream));
  return ctg_d2h(ctx, out, d_out, sizeof(G1Affine));
}
int32_t ctg_g2_combine(ctg_ctx* ctx, const ctg_g2_xyzz* host_partials, uint32_t k, ctg_g2_affine* out) {
  if (!ctx || !host_partials || !out) return CTG_ERR_INVALID;
  CTG_TRY(ensure_scratch(ctx, (size_t)k * sizeof(G2XYZZ) + sizeof(G2Affine) + 64));
  char* d = (char*)ctx->scratch;
  G2Affine* d_out = (G2Affine*)(d + (size_t)k * sizeof(G2XYZZ));
  CTG_CUDA(ctx, cudaMemcpyAsync(d, host_partials, (size_t)k * sizeof(G2XYZZ), cudaMemcpyHostToDevice, ctx->stream));
  CTG_TRY(g2_combine_dev(ctx, (const G2XYZZ*)d, k, d_out, nullptr, ctx->stream));
  return ctg_d2h(ctx, out, d_out, sizeof(G2Affine));
}

// ---- proving key / R1CS / prover --------------------------------------------------------------
int32_t ctg_pk_upload(ctg_ctx* ctx, const ctg_pk_view* host_pk, uint64_t num_instance, uint64_t num_constraints, int copies, int c_bits, ctg_pk** out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  if (c_bits != 0 && (c_bits < 2 || c_bits > 21)) return fail(ctx, CTG_ERR_INVALID, "pk_upload: c_bits must be 0 or in [2, 21]");
  if (ctx->group) return group_pk_upload(ctx, host_pk, num_instance, num_constraints, copies, c_bits, out);   // one point-range shard per device
  return pk_upload(ctx, host_pk, num_instance, num_constraints, copies, c_bits, 0, 1, 1, out);
}
int32_t ctg_pk_upload_shard(ctg_ctx* ctx, const ctg_pk_view* host_pk, uint64_t num_instance, uint64_t num_constraints, int copies, int c_bits,
                            int shard_index, int shard_count, ctg_pk** out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  if (c_bits != 0 && (c_bits < 2 || c_bits > 21)) return fail(ctx, CTG_ERR_INVALID, "pk_upload: c_bits must be 0 or in [2, 21]");
  if (shard_count < 1 || shard_index < 0 || shard_index >= shard_count) return fail(ctx, CTG_ERR_INVALID, "pk_upload: bad shard index / count");
  return pk_upload(ctx, host_pk, num_instance, num_constraints, copies, c_bits, (uint64_t)shard_index, (uint64_t)shard_index + 1, (uint64_t)shard_count, out);
}
int32_t ctg_pk_upload_range(ctg_ctx* ctx, const ctg_pk_view* host_pk, uint64_t num_instance, uint64_t num_constraints, int copies, int c_bits,
                            uint64_t lo_num, uint64_t hi_num, uint64_t den, ctg_pk** out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  if (c_bits != 0 && (c_bits < 2 || c_bits > 21)) return fail(ctx, CTG_ERR_INVALID, "pk_upload: c_bits must be 0 or in [2, 21]");
  return pk_upload(ctx, host_pk, num_instance, num_constraints, copies, c_bits, lo_num, hi_num, den, out);
}
void ctg_pk_free(ctg_ctx* ctx, ctg_pk* pk) {
  if (!pk) return;
  if (ctx) cudaStreamSynchronize(ctx->stream);
  if (!pk->parts.empty()) { if (ctx && ctx->group) group_pk_free(ctx, pk); return; }
  pk_free(pk);
}
int32_t ctg_r1cs_upload(ctg_ctx* ctx, uint64_t num_instance, uint64_t num_witness, uint64_t num_constraints, const ctg_csr* a, const ctg_csr* b,
                        const ctg_csr* c, ctg_r1cs** out) {
  if (!ctx || !out) return CTG_ERR_INVALID;
  const ctg_csr* mats[3] = {a, b, c};
  if (ctx->group) return group_r1cs_upload(ctx, num_instance, num_witness, num_constraints, mats, out);
  return r1cs_upload(ctx, num_instance, num_witness, num_constraints, mats, out);
}
void ctg_r1cs_free(ctg_ctx* ctx, ctg_r1cs* r1cs) {
  if (!r1cs) return;
  if (ctx) cudaStreamSynchronize(ctx->stream);
  if (!r1cs->parts.empty()) { if (ctx && ctx->group) group_r1cs_free(ctx, r1cs); return; }
  r1cs_free(r1cs);
}
int32_t ctg_witness_map(ctg_ctx* ctx, const ctg_r1cs* r1cs, const ctg_fr* z, ctg_fr* h_out) {
  if (!ctx || !r1cs || !z || !h_out) return ctx ? fail(ctx, CTG_ERR_INVALID, "witness_map: null argument") : CTG_ERR_INVALID;
  if (!r1cs->parts.empty()) r1cs = r1cs->parts[0];   // multi-device context: the witness map alone runs on the primary device
  return witness_map_host(ctx, r1cs, z, h_out);
}
int32_t ctg_prove(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r1cs, const ctg_fr* z, const ctg_fr* r, const ctg_fr* s, ctg_proof* out,
                  ctg_timings* timings) {
  if (!ctx) return CTG_ERR_INVALID;
  ctg_timings local;
  if (ctx->group) {   // sharded over the context's devices; the result is the same proof
    int32_t rc = group_prove(ctx, pk, r1cs, z, r, s, out, timings ? &local : nullptr);
    if (rc == CTG_OK && timings) { *timings = local; g_last_timings = local; }
    return rc;
  }
  if (pk && !pk->parts.empty()) return fail(ctx, CTG_ERR_STATE, "prove: the key belongs to a multi-device context");
  int32_t rc = prove(ctx, pk, r1cs, z, r, s, out, timings ? &local : nullptr);
  if (rc == CTG_OK && timings) {
    local.msm_sort_ms = g_last_timings.msm_sort_ms; local.msm_accumulate_ms = g_last_timings.msm_accumulate_ms; local.msm_reduce_ms = g_last_timings.msm_reduce_ms;
    *timings = local;
    g_last_timings = local;
  }
  return rc;
}
int32_t ctg_prove_batch(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r1cs, uint32_t n_proofs, const ctg_fr* z_batch, const ctg_fr* rs_batch,
                        ctg_proof* out) {
  if (!ctx || !pk || !r1cs || !z_batch || !rs_batch || !out) return ctx ? fail(ctx, CTG_ERR_INVALID, "prove_batch: null argument") : CTG_ERR_INVALID;
  if (!pk->parts.empty() || !r1cs->parts.empty()) {   // a sharded key: the proofs of the batch run one after the other, each over all devices
    if (!ctx->group) return fail(ctx, CTG_ERR_STATE, "prove_batch: the key belongs to a multi-device context");
    const uint64_t m = pk_num_vars(pk);
    for (uint32_t i = 0; i < n_proofs; i++)
      CTG_TRY(group_prove(ctx, pk, r1cs, z_batch + (size_t)i * m, rs_batch + 2 * (size_t)i, rs_batch + 2 * (size_t)i + 1, out + i, nullptr));
    return CTG_OK;
  }
  return prove_batch(ctx, pk, r1cs, n_proofs, z_batch, rs_batch, out);
}
int32_t ctg_prove_partial(ctg_ctx* ctx, const ctg_pk* pk_shard, const ctg_r1cs* r1cs, const ctg_fr* z, const ctg_fr* r, const ctg_fr* s,
                          ctg_proof_partial* out, ctg_timings* timings) {
  if (!ctx) return CTG_ERR_INVALID;
  ctg_timings local;
  if ((pk_shard && !pk_shard->parts.empty()) || (r1cs && !r1cs->parts.empty())) return fail(ctx, CTG_ERR_STATE, "prove_partial: handles of a multi-device context; call ctg_prove");
  int32_t rc = prove_partial(ctx, pk_shard, r1cs, z, r, s, out, timings ? &local : nullptr);
  if (rc == CTG_OK && timings) { *timings = local; g_last_timings = local; }
  return rc;
}
int32_t ctg_prove_partial_begin(ctg_ctx* ctx, const ctg_pk* pk_shard, const ctg_r1cs* r1cs, const ctg_fr* z, const ctg_fr* r, const ctg_fr* s,
                                uint32_t chain_mask, void* d_abc) {
  if (!ctx) return CTG_ERR_INVALID;
  if ((pk_shard && !pk_shard->parts.empty()) || (r1cs && !r1cs->parts.empty())) return fail(ctx, CTG_ERR_STATE, "prove_partial: handles of a multi-device context; call ctg_prove");
  return prove_partial_begin(ctx, pk_shard, r1cs, z, r, s, chain_mask, d_abc);
}
int32_t ctg_prove_partial_finish(ctg_ctx* ctx, const ctg_pk* pk_shard, const ctg_r1cs* r1cs, const ctg_fr* r, const ctg_fr* s, void* d_abc,
                                 ctg_proof_partial* out) {
  if (!ctx) return CTG_ERR_INVALID;
  if ((pk_shard && !pk_shard->parts.empty()) || (r1cs && !r1cs->parts.empty())) return fail(ctx, CTG_ERR_STATE, "prove_partial: handles of a multi-device context; call ctg_prove");
  return prove_partial_finish(ctx, pk_shard, r1cs, r, s, d_abc, out);
}
int32_t ctg_prove_combine(ctg_ctx* ctx, const ctg_proof_partial* partials, uint32_t k, const ctg_fr* r, const ctg_fr* s, ctg_proof* out) {
  if (!ctx) return CTG_ERR_INVALID;
  return prove_combine(ctx, partials, k, r, s, out);
}

}  // extern "C"
