// Multi-device contexts: ctg_ctx_create(device_ids, n_devices > 1) -- ONE process drives several GPUs of one box, so that a
// Rust host (the reference's Seller is one process: lib/src/seller.rs:81-86, 129-159) gets a sharded proof from the same
// ctg_prove / ctg_msm_* calls it makes on one GPU.  No collective library: the exchange steps are cudaMemcpyPeerAsync over
// NVLink / NVSwitch between streams ordered by cross-device events, and there is no host round trip between the phases.
//
//   proving key   ctg_pk_upload on such a context keeps a point-range shard of every query on each device (1/N of the key)
//   R1CS          replicated on the (up to three) devices that compute witness-map chains
//   ctg_prove     SURVEY 8(e): the five MSMs split by point range; the three NTT chains a, b, c on their owner devices; the
//                 pointwise step and the last coset_ifft once, on device 0; h handed out in slices; 1152 B of partial sums per
//                 device summed on device 0 (prove.cu prove_group_a..d)
//   ctg_msm_g1/g2 the base table is sharded the same way; scalars go to each device in slices; 192 / 384 B partials
//
// Every device has a worker thread that binds the device once; the caller's thread only hands out the per-device enqueue work
// (so eight devices' ~100 launches each are issued in parallel) and orders the phases.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "ctg_internal.h"

namespace ctg {
static bool trace_on() { static int v = -1; if (v < 0) { const char* e = getenv("CTG_TRACE"); v = e && *e == '1'; } return v != 0; }
#define CTG_TRACE(...) do { if (trace_on()) { fprintf(stderr, "[ctg multi] " __VA_ARGS__); fputc('\n', stderr); fflush(stderr); } } while (0)
int32_t pk_upload(ctg_ctx* ctx, const ctg_pk_view* v, uint64_t ni, uint64_t nc, int copies, int c_bits, uint64_t lo_num, uint64_t hi_num, uint64_t den, ctg_pk** out);
void pk_free(ctg_pk* pk);
int32_t r1cs_upload(ctg_ctx* ctx, uint64_t ni, uint64_t nw, uint64_t nc, const ctg_csr* const mats[3], ctg_r1cs** out);
void r1cs_free(ctg_r1cs* r);

// ---- worker threads ---------------------------------------------------------------------------------------------------
static void worker_main(DeviceWorker* w) {
  cudaSetDevice(w->device);
  std::unique_lock<std::mutex> lk(w->mu);
  for (;;) {
    w->cv.wait(lk, [&] { return w->has_job || w->stop; });
    if (w->stop) return;
    std::function<int32_t()> job = std::move(w->job);
    w->has_job = false;
    lk.unlock();
    const int32_t rc = job();
    lk.lock();
    w->rc = rc;
    w->done = true;
    w->cv.notify_all();
  }
}
static void worker_submit(DeviceWorker* w, std::function<int32_t()> fn) {
  std::lock_guard<std::mutex> lk(w->mu);
  w->job = std::move(fn);
  w->has_job = true;
  w->done = false;
  w->cv.notify_all();
}
static int32_t worker_wait(DeviceWorker* w) {
  std::unique_lock<std::mutex> lk(w->mu);
  w->cv.wait(lk, [&] { return w->done; });
  return w->rc;
}
int32_t group_run(Group* g, const std::function<int32_t(int)>& fn) {
  CTG_TRACE("group_run: submit to %d workers", g->n);
  for (int d = 0; d < g->n; d++) worker_submit(g->worker[d], [&fn, d] { return fn(d); });
  int32_t first = CTG_OK;
  for (int d = 0; d < g->n; d++) {
    const int32_t rc = worker_wait(g->worker[d]);
    CTG_TRACE("group_run: worker %d returned %d", d, rc);
    if (rc != CTG_OK && first == CTG_OK) {
      first = rc;
      if (d != 0) g->ctx[0]->last_error = "device " + std::to_string(g->ctx[d]->device) + ": " + g->ctx[d]->last_error;
    }
  }
  return first;
}
int32_t group_run_on(Group* g, int d, const std::function<int32_t()>& fn) {
  CTG_TRACE("group_run_on %d", d);
  worker_submit(g->worker[d], fn);
  const int32_t rc = worker_wait(g->worker[d]);
  CTG_TRACE("group_run_on %d returned %d", d, rc);
  if (rc != CTG_OK && d != 0) g->ctx[0]->last_error = "device " + std::to_string(g->ctx[d]->device) + ": " + g->ctx[d]->last_error;
  return rc;
}

// ---- life cycle ---------------------------------------------------------------------------------------------------------
// primary: the context already created on device_ids[0]; children are created here through the public constructor
int32_t group_create(ctg_ctx* primary, const int* device_ids, int n) {
  if (n < 2 || n > kMaxDevices) return fail(primary, CTG_ERR_INVALID, "ctx_create: between 1 and 8 devices");
  // a device id may repeat: the shards then share that GPU (no speed-up; it lets a one-GPU box run the multi-device path in tests)
  Group* g = new Group();
  g->n = n;
  g->ctx[0] = primary;
  auto cleanup = [&](int32_t rc) {
    for (int d = 1; d < n; d++) if (g->ctx[d]) ctg_ctx_destroy(g->ctx[d]);
    cudaSetDevice(primary->device);
    delete g;
    return rc;
  };
  for (int d = 1; d < n; d++) {
    const int32_t rc = ctg_ctx_create(&device_ids[d], 1, &g->ctx[d]);
    if (rc != CTG_OK) return cleanup(rc);
  }
  // peer access in both directions between every pair (NVSwitch: all-to-all)
  for (int a = 0; a < n; a++) {
    cudaSetDevice(g->ctx[a]->device);
    for (int b = 0; b < n; b++) {
      if (a == b || g->ctx[a]->device == g->ctx[b]->device) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, g->ctx[a]->device, g->ctx[b]->device);
      if (!can) { primary->last_error = "ctx_create: the devices have no peer access to each other"; return cleanup(CTG_ERR_STATE); }
      cudaError_t e = cudaDeviceEnablePeerAccess(g->ctx[b]->device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { primary->last_error = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); return cleanup(CTG_ERR_CUDA); }
      cudaGetLastError();
    }
  }
  cudaSetDevice(primary->device);
  for (int d = 0; d < n; d++) {
    DeviceWorker* w = new DeviceWorker();
    w->device = g->ctx[d]->device;
    w->th = std::thread(worker_main, w);
    g->worker[d] = w;
  }
  // chain owners: device 0 finishes the witness map, so with four or more devices the chains go to devices 1..3
  if (n >= 4) { g->chain_owner[0] = 1; g->chain_owner[1] = 2; g->chain_owner[2] = 3; }
  else if (n == 3) { g->chain_owner[0] = 1; g->chain_owner[1] = 2; g->chain_owner[2] = 0; }
  else { g->chain_owner[0] = 0; g->chain_owner[1] = 1; g->chain_owner[2] = 1; }
  int32_t rc = group_run(g, [g](int d) -> int32_t {
    ctg_ctx* c = g->ctx[d];
    CTG_CUDA(c, cudaEventCreateWithFlags(&g->ev_a[d], cudaEventDisableTiming));
    CTG_CUDA(c, cudaEventCreateWithFlags(&g->ev_part[d], cudaEventDisableTiming));
    CTG_CUDA(c, cudaEventCreateWithFlags(&g->ev_msm[d], cudaEventDisableTiming));
    if (d == 0) {
      CTG_CUDA(c, cudaEventCreateWithFlags(&g->ev_h, cudaEventDisableTiming));
      CTG_CUDA(c, cudaEventCreateWithFlags(&g->ev_msm_go, cudaEventDisableTiming));
      CTG_CUDA(c, cudaEventCreate(&g->ev_t0));
      CTG_CUDA(c, cudaEventCreate(&g->ev_t1));
      CTG_CUDA(c, cudaMalloc(&g->d_parts0, (size_t)g->n * kPartialBytes + 4 * sizeof(G1XYZZ) + sizeof(G2XYZZ) + sizeof(ctg_proof) + 256));
      CTG_CUDA(c, cudaMalloc(&g->d_msm_parts0, (size_t)(g->n + 1) * sizeof(G2XYZZ) + sizeof(G2Affine) + 256));
      CTG_CUDA(c, cudaMallocHost(&g->h_proof, sizeof(ctg_proof) + sizeof(G2Affine)));
    }
    return CTG_OK;
  });
  if (rc != CTG_OK) { primary->group = g; return rc; }   // ctg_ctx_destroy tears it down
  primary->group = g;
  return CTG_OK;
}

static void group_free_abc(Group* g) {
  for (int d = 0; d < g->n; d++) {
    Fr* abc = g->d_abc[d]; Fr* hs = g->d_hslice[d];
    if (abc || hs) group_run_on(g, d, [abc, hs]() -> int32_t { cudaDeviceSynchronize(); if (abc) cudaFree(abc); if (hs) cudaFree(hs); return CTG_OK; });
    g->d_abc[d] = nullptr; g->d_hslice[d] = nullptr;
  }
  g->abc_n = 0;
}

void group_destroy(ctg_ctx* primary) {
  Group* g = primary->group;
  if (!g) return;
  group_free_abc(g);
  group_run(g, [g](int d) -> int32_t {
    cudaDeviceSynchronize();
    if (g->ev_a[d]) cudaEventDestroy(g->ev_a[d]);
    if (g->ev_part[d]) cudaEventDestroy(g->ev_part[d]);
    if (g->ev_msm[d]) cudaEventDestroy(g->ev_msm[d]);
    if (g->d_scalars[d]) cudaFree(g->d_scalars[d]);
    if (d == 0) {
      if (g->ev_h) cudaEventDestroy(g->ev_h);
      if (g->ev_msm_go) cudaEventDestroy(g->ev_msm_go);
      if (g->ev_t0) cudaEventDestroy(g->ev_t0);
      if (g->ev_t1) cudaEventDestroy(g->ev_t1);
      if (g->d_parts0) cudaFree(g->d_parts0);
      if (g->d_msm_parts0) cudaFree(g->d_msm_parts0);
      if (g->h_proof) cudaFreeHost(g->h_proof);
    }
    return CTG_OK;
  });
  for (int d = 0; d < g->n; d++) {
    DeviceWorker* w = g->worker[d];
    if (!w) continue;
    { std::lock_guard<std::mutex> lk(w->mu); w->stop = true; w->cv.notify_all(); }
    w->th.join();
    delete w;
  }
  for (int d = 1; d < g->n; d++) ctg_ctx_destroy(g->ctx[d]);
  cudaSetDevice(primary->device);
  primary->group = nullptr;
  delete g;
}

uint64_t group_launches(const Group* g) {
  uint64_t t = 0;
  for (int d = 1; d < g->n; d++) t += g->ctx[d]->launches;
  return t;
}

// ---- proving key, R1CS --------------------------------------------------------------------------------------------------
int32_t group_pk_upload(ctg_ctx* ctx, const ctg_pk_view* v, uint64_t ni, uint64_t nc, int copies, int c_bits, ctg_pk** out) {
  Group* g = ctx->group;
  ctg_pk* pk = new ctg_pk();
  pk->parts.assign(g->n, nullptr);
  int32_t rc = group_run(g, [&](int d) { return pk_upload(g->ctx[d], v, ni, nc, copies, c_bits, (uint64_t)d, (uint64_t)d + 1, (uint64_t)g->n, &pk->parts[d]); });
  if (rc != CTG_OK) {
    for (int d = 0; d < g->n; d++) { ctg_pk* p = pk->parts[d]; if (p) group_run_on(g, d, [p]() -> int32_t { pk_free(p); return CTG_OK; }); }
    delete pk;
    return rc;
  }
  const ctg_pk* p0 = pk->parts[0];
  pk->m = p0->m; pk->ni = p0->ni; pk->nw = p0->nw; pk->n = p0->n; pk->log_n = p0->log_n;
  pk->shard_index = 0; pk->shard_count = 1;
  memset(&pk->a_ext, 0, sizeof(Bases)); memset(&pk->b1_ext, 0, sizeof(Bases)); memset(&pk->b2_ext, 0, sizeof(Bases));
  memset(&pk->l, 0, sizeof(Bases)); memset(&pk->h_ext, 0, sizeof(Bases));
  *out = pk;
  return CTG_OK;
}
void group_pk_free(ctg_ctx* ctx, ctg_pk* pk) {
  Group* g = ctx->group;
  for (int d = 0; d < (int)pk->parts.size(); d++) {
    ctg_pk* p = pk->parts[d];
    ctg_ctx* c = g->ctx[d];
    if (p) group_run_on(g, d, [p, c]() -> int32_t { cudaStreamSynchronize(c->stream); pk_free(p); return CTG_OK; });
  }
  delete pk;
}

int32_t group_r1cs_upload(ctg_ctx* ctx, uint64_t ni, uint64_t nw, uint64_t nc, const ctg_csr* const mats[3], ctg_r1cs** out) {
  Group* g = ctx->group;
  ctg_r1cs* r = new ctg_r1cs();
  for (int k = 0; k < 3; k++) { r->row_ptr[k] = nullptr; r->col[k] = nullptr; r->val[k] = nullptr; r->nnz[k] = 0; }
  r->parts.assign(g->n, nullptr);
  bool owner[kMaxDevices] = {};
  owner[0] = true;   // device 0 also serves ctg_witness_map and validates the structure
  for (int k = 0; k < 3; k++) owner[g->chain_owner[k]] = true;
  int32_t rc = group_run(g, [&](int d) -> int32_t { return owner[d] ? r1cs_upload(g->ctx[d], ni, nw, nc, mats, &r->parts[d]) : CTG_OK; });
  if (rc != CTG_OK) {
    for (int d = 0; d < g->n; d++) { ctg_r1cs* p = r->parts[d]; if (p) group_run_on(g, d, [p]() -> int32_t { r1cs_free(p); return CTG_OK; }); }
    delete r;
    return rc;
  }
  const ctg_r1cs* r0 = r->parts[0];
  r->ni = r0->ni; r->nw = r0->nw; r->nc = r0->nc; r->n = r0->n; r->log_n = r0->log_n;
  *out = r;
  return CTG_OK;
}
void group_r1cs_free(ctg_ctx* ctx, ctg_r1cs* r) {
  Group* g = ctx->group;
  for (int d = 0; d < (int)r->parts.size(); d++) {
    ctg_r1cs* p = r->parts[d];
    ctg_ctx* c = g->ctx[d];
    if (p) group_run_on(g, d, [p, c]() -> int32_t { cudaStreamSynchronize(c->stream); r1cs_free(p); return CTG_OK; });
  }
  delete r;
}

// ---- sharded proof ------------------------------------------------------------------------------------------------------
static int32_t group_ensure_abc(Group* g, const ctg_pk* pk, uint64_t n) {
  if (g->abc_n == n) {
    bool ok = true;
    for (int d = 1; d < g->n; d++) if (!g->d_hslice[d]) ok = false;
    if (ok) return CTG_OK;
  }
  group_free_abc(g);
  bool owner[kMaxDevices] = {};
  owner[0] = true;
  for (int k = 0; k < 3; k++) owner[g->chain_owner[k]] = true;
  int32_t rc = group_run(g, [&](int d) -> int32_t {
    ctg_ctx* c = g->ctx[d];
    if (owner[d] && cudaMalloc(&g->d_abc[d], 3 * n * sizeof(Fr)) != cudaSuccess) { cudaGetLastError(); return fail(c, CTG_ERR_OOM, "prove: out of device memory for the witness-map chains"); }
    if (d != 0) {
      const uint64_t cnt = pk->parts[d]->h_cnt ? pk->parts[d]->h_cnt : 1;
      if (cudaMalloc(&g->d_hslice[d], cnt * sizeof(Fr)) != cudaSuccess) { cudaGetLastError(); return fail(c, CTG_ERR_OOM, "prove: out of device memory for the slice of h"); }
    }
    return CTG_OK;
  });
  if (rc == CTG_OK) g->abc_n = n;
  return rc;
}

int32_t group_prove(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const ctg_fr* rr, const ctg_fr* ss, ctg_proof* out, ctg_timings* tm) {
  Group* g = ctx->group;
  if (!pk || !r || !z || !rr || !ss || !out) return fail(ctx, CTG_ERR_INVALID, "prove: null argument");
  if ((int)pk->parts.size() != g->n || (int)r->parts.size() != g->n)
    return fail(ctx, CTG_ERR_STATE, "prove: on a multi-device context the key and the R1CS must have been uploaded through the same context");
  if (pk->ni != r->ni || pk->nw != r->nw || pk->n != r->n) return fail(ctx, CTG_ERR_INVALID, "prove: proving key and R1CS shapes differ");
  CTG_TRY(group_ensure_abc(g, pk, r->n));
  Fr r_m, s_m;
  memcpy(&r_m, rr, sizeof(Fr));
  memcpy(&s_m, ss, sizeof(Fr));
  uint32_t mask[kMaxDevices] = {};
  for (int k = 0; k < 3; k++) mask[g->chain_owner[k]] |= 1u << k;
  const auto t0 = std::chrono::steady_clock::now();
  // A: every device -- its slice of z, its four assignment MSMs, the chains it owns
  CTG_TRY(group_run(g, [&](int d) -> int32_t {
    const ctg_r1cs* rd = r->parts[d] ? r->parts[d] : r;     // devices without chains only read the shape
    Fr* abc = g->d_abc[d] ? g->d_abc[d] : g->d_hslice[d];   // non-null either way: the lane must not reserve its own 3 n buffer
    return prove_group_a(g->ctx[d], pk->parts[d], rd, z, r_m, s_m, mask[d], abc, g->ev_a[d]);
  }));
  // B: device 0 -- pull the other chains, pointwise step, last coset_ifft
  CTG_TRY(group_run_on(g, 0, [&]() { return prove_group_b(g, r->parts[0], r_m, s_m); }));
  // C: every device -- its slice of h, the H MSM, partial sums to device 0
  CTG_TRY(group_run(g, [&](int d) { return prove_group_c(g, d, pk->parts[d], r); }));
  // D: device 0 -- sum, assemble, copy out (the only host synchronisation of the call)
  CTG_TRY(group_run_on(g, 0, [&]() { return prove_group_d(g, r_m, s_m, out); }));
  if (tm) {
    memset(tm, 0, sizeof(*tm));
    tm->total_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
  }
  return CTG_OK;
}

// ---- sharded base tables and MSM ----------------------------------------------------------------------------------------
static void part_range(uint64_t n, int d, int parts, uint64_t* lo, uint64_t* cnt) {
  const uint64_t a = (uint64_t)((unsigned __int128)n * d / parts), b = (uint64_t)((unsigned __int128)n * (d + 1) / parts);
  *lo = a; *cnt = b - a;
}

// make(d, lo, cnt, &part): creates device d's part through the single-device entry points
int32_t group_bases_new(ctg_ctx* ctx, int group, uint64_t n, const std::function<int32_t(int, uint64_t, uint64_t, ctg_bases**)>& make, ctg_bases** out) {
  Group* g = ctx->group;
  ctg_bases* h = new ctg_bases();
  h->parts.assign(g->n, nullptr);
  h->part_lo.assign(g->n + 1, 0);
  for (int d = 0; d < g->n; d++) { uint64_t lo, cnt; part_range(n, d, g->n, &lo, &cnt); h->part_lo[d] = lo; h->part_lo[d + 1] = lo + cnt; }
  int32_t rc = group_run(g, [&](int d) { return make(d, h->part_lo[d], h->part_lo[d + 1] - h->part_lo[d], &h->parts[d]); });
  if (rc != CTG_OK) {
    for (int d = 0; d < g->n; d++) { ctg_bases* p = h->parts[d]; ctg_ctx* c = g->ctx[d]; if (p) group_run_on(g, d, [p, c]() -> int32_t { ctg_bases_free(c, p); return CTG_OK; }); }
    delete h;
    return rc;
  }
  h->b = h->parts[0]->b;
  h->b.pts = nullptr;
  h->b.group = group;
  h->b.n = n;
  *out = h;
  return CTG_OK;
}
void group_bases_free(ctg_ctx* ctx, ctg_bases* h) {
  Group* g = ctx->group;
  for (int d = 0; d < (int)h->parts.size(); d++) {
    ctg_bases* p = h->parts[d];
    ctg_ctx* c = g->ctx[d];
    if (p) group_run_on(g, d, [p, c]() -> int32_t { ctg_bases_free(c, p); return CTG_OK; });
  }
  delete h;
}

// scalars: host pointer (from_host) or a pointer into device 0's memory; result: affine to host (out_affine), and / or the
// projective sum left on device 0 (d_out_xyzz)
int32_t group_msm(ctg_ctx* ctx, const ctg_bases* bases, int group, const void* scalars, bool from_host, uint64_t n, uint32_t flags, void* out_affine,
                  void* d_out_xyzz) {
  Group* g = ctx->group;
  if ((int)bases->parts.size() != g->n) return fail(ctx, CTG_ERR_STATE, "msm: on a multi-device context the base table must have been created through the same context");
  if (bases->b.group != group) return fail(ctx, CTG_ERR_INVALID, "msm: base table is for the other group");
  if (n > bases->b.n) return fail(ctx, CTG_ERR_INVALID, "msm: more scalars than bases");
  if (n && !scalars) return fail(ctx, CTG_ERR_INVALID, "msm: null scalars");
  const bool mont = (flags & CTG_SCALARS_MONT) != 0;
  const size_t xb = group == 1 ? sizeof(G1XYZZ) : sizeof(G2XYZZ), ab = group == 1 ? sizeof(G1Affine) : sizeof(G2Affine);
  char* parts0 = (char*)g->d_msm_parts0;
  // the caller's stream on device 0 orders the read of device-resident scalars
  if (!from_host) CTG_TRY(group_run_on(g, 0, [&]() -> int32_t { CTG_CUDA(ctx, cudaEventRecord(g->ev_msm_go, ctx->stream)); return CTG_OK; }));
  CTG_TRY(group_run(g, [&](int d) -> int32_t {
    ctg_ctx* c = g->ctx[d];
    const uint64_t lo = bases->part_lo[d], hi = bases->part_lo[d + 1] < n ? bases->part_lo[d + 1] : n;
    const uint64_t cnt = hi > lo ? hi - lo : 0;
    cudaStream_t st = c->stream;
    const Fr* sc = nullptr;
    if (cnt) {
      if (!from_host && d == 0) {
        sc = (const Fr*)scalars + lo;
      } else {
        if (cnt * sizeof(Fr) > g->scalars_bytes[d]) {
          CTG_CUDA(c, cudaStreamSynchronize(st));
          if (g->d_scalars[d]) cudaFree(g->d_scalars[d]);
          g->d_scalars[d] = nullptr; g->scalars_bytes[d] = 0;
          if (cudaMalloc(&g->d_scalars[d], cnt * sizeof(Fr)) != cudaSuccess) { cudaGetLastError(); return fail(c, CTG_ERR_OOM, "msm: out of device memory for the scalars"); }
          g->scalars_bytes[d] = cnt * sizeof(Fr);
        }
        if (from_host) CTG_CUDA(c, cudaMemcpyAsync(g->d_scalars[d], (const Fr*)scalars + lo, cnt * sizeof(Fr), cudaMemcpyHostToDevice, st));
        else {
          CTG_CUDA(c, cudaStreamWaitEvent(st, g->ev_msm_go, 0));
          CTG_CUDA(c, cudaMemcpyPeerAsync(g->d_scalars[d], c->device, (const Fr*)scalars + lo, ctx->device, cnt * sizeof(Fr), st));
        }
        sc = (const Fr*)g->d_scalars[d];
      }
    }
    if (!c->ws[0].result) CTG_CUDA(c, cudaMalloc(&c->ws[0].result, sizeof(G2XYZZ)));
    void* d_part = d == 0 ? (void*)(parts0) : c->ws[0].result;
    const Bases& b = bases->parts[d]->b;
    if (group == 1) CTG_TRY(msm_g1_run(c, c->ws[0], b, (const uint32_t*)sc, cnt, mont, (G1XYZZ*)d_part, st, nullptr));
    else CTG_TRY(msm_g2_run(c, c->ws[0], b, (const uint32_t*)sc, cnt, mont, (G2XYZZ*)d_part, st, nullptr));
    if (d != 0) CTG_CUDA(c, cudaMemcpyPeerAsync(parts0 + (size_t)d * xb, ctx->device, d_part, c->device, xb, st));
    CTG_CUDA(c, cudaEventRecord(g->ev_msm[d], st));
    return CTG_OK;
  }));
  return group_run_on(g, 0, [&]() -> int32_t {
    cudaStream_t st = ctx->stream;
    for (int d = 1; d < g->n; d++) CTG_CUDA(ctx, cudaStreamWaitEvent(st, g->ev_msm[d], 0));
    char* d_sum = parts0 + (size_t)g->n * xb;
    char* d_aff = d_sum + sizeof(G2XYZZ);
    void* xy_out = d_out_xyzz ? d_out_xyzz : (void*)d_sum;
    if (group == 1) CTG_TRY(g1_combine_dev(ctx, (const G1XYZZ*)parts0, (uint32_t)g->n, out_affine ? (G1Affine*)d_aff : nullptr, (G1XYZZ*)xy_out, st));
    else CTG_TRY(g2_combine_dev(ctx, (const G2XYZZ*)parts0, (uint32_t)g->n, out_affine ? (G2Affine*)d_aff : nullptr, (G2XYZZ*)xy_out, st));
    if (out_affine) {
      CTG_CUDA(ctx, cudaMemcpyAsync(out_affine, d_aff, ab, cudaMemcpyDeviceToHost, st));
      CTG_CUDA(ctx, cudaStreamSynchronize(st));
    }
    return CTG_OK;
  });
}

}  // namespace ctg
