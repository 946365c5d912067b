// Internal host-side declarations shared by the translation units of libctg_groth16.so.
// Nothing here crosses the C ABI (include/ctg_groth16.h does).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/ctg_groth16.h"
#include "curve.cuh"
#include "fp.cuh"

namespace ctg {

struct NttDomain {
  int log_n = 0;
  Fr* tw_fwd = nullptr;   // w^k,  k < n/2
  Fr* tw_inv = nullptr;   // w^-k, k < n/2
  Fr* gpow = nullptr;     // g^i            (coset_fft premultiplier)
  Fr* gpow_n = nullptr;   // g^i / n        (ifft followed by coset_fft, fused)
  Fr* gipow_n = nullptr;  // g^-i / n       (coset_ifft postmultiplier)
  Fr n_inv;               // 1/n (Montgomery)
  Fr zinv;                // 1/(g^n - 1)
};

// Per-MSM scratch.  One instance per concurrently running MSM.
struct MsmWorkspace {
  uint32_t* counts = nullptr;   // [nb + 1] bucket sizes, then exclusive starts
  uint32_t* cursor = nullptr;   // [nb]
  uint32_t* entries = nullptr;  // [n * W] sorted (point | sign << 31)
  void* buckets = nullptr;      // [nb] XYZZ
  void* part_pt = nullptr;      // [2 * T] XYZZ partial sums of buckets split across threads
  uint32_t* part_gb = nullptr;  // [2 * T]
  void* chunk = nullptr;        // bucket-reduction partials
  void* result = nullptr;       // 1 XYZZ
  uint32_t* tile_sums = nullptr;  // scan scratch
  cudaEvent_t ev_sorted = nullptr;  // recorded after the digit sort: another MSM over the same scalars and plan may reuse counts / entries
  // batched-affine tree (msm_acc.cuh): compact point arrays of the odd / even levels and their bucket offsets
  void* ba_pts[2] = {nullptr, nullptr};
  uint32_t* ba_starts[2] = {nullptr, nullptr};   // [nb + 1] each
  size_t cap_ba_bytes[2] = {0, 0}, cap_ba_nb = 0;
  size_t cap_nb = 0, cap_entries = 0, cap_bucket_bytes = 0, cap_part_bytes = 0, cap_chunk_bytes = 0;
  int part_threads = 0;
};

struct MsmPlan {
  int c;        // window bits
  int nwin;     // ceil(256 / c)
  int copies;   // precomputed multiples per base (1 = none)
  int nsets;    // bucket sets = ceil(nwin / copies)
  uint32_t nbuckets_per_set;  // 2^(c-1)
};

struct Bases {  // device-resident base table (a proving-key query or a standalone upload)
  int group;        // 1 = G1, 2 = G2
  uint64_t n;       // points per copy
  MsmPlan plan;
  void* pts;        // [copies][n] affine
  uint64_t stride = 0;  // distance between copies in points; 0 = n (a point-range view of a larger table sets it to the parent's n)
};

}  // namespace ctg

namespace ctg {
// One in-flight proof: 5 streams (0 = witness map + H MSM, 1..4 = L, A, B1, B2 MSMs), one MSM workspace per stream,
// private scratch.  Lane 0's stream 0 is the context stream.  Batched proving round-robins over the lanes.
struct ProveLane {
  bool ready = false;
  cudaStream_t st[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  MsmWorkspace ws[5];
  void* scratch = nullptr; size_t scratch_bytes = 0;
  cudaEvent_t ev_z = nullptr, ev_done[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  // the witness map and the H MSM are the critical path of a proof (the other four MSMs only need z): they run on a
  // high-priority stream so that their blocks are scheduled ahead of the queued MSM blocks
  cudaStream_t st_hi = nullptr;
  cudaEvent_t ev_zfull = nullptr, ev_wm = nullptr;
  cudaEvent_t ev_t[16];   // stage timers
};
static constexpr int kMaxLanes = 4;
struct LaneBufs { Fr* d_z; Fr* d_abc; G1XYZZ* d_g1; G2XYZZ* d_g2; ctg_proof* d_proof; G1XYZZ* d_early; };
}  // namespace ctg

#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>

namespace ctg {
// One host thread per device of a multi-device context: it binds the device once and then runs the jobs the
// orchestrating (caller's) thread hands it, so that the per-device enqueue work of a sharded call proceeds in parallel.
struct DeviceWorker {
  int device = 0;
  std::thread th;
  std::mutex mu;
  std::condition_variable cv;
  std::function<int32_t()> job;
  bool has_job = false, done = false, stop = false;
  int32_t rc = 0;
};
static constexpr int kMaxDevices = 8;
static constexpr size_t kPartialBytes = 4 * sizeof(G1XYZZ) + sizeof(G2XYZZ);   // ctg_proof_partial
// The devices of a multi-device context (ctg_ctx_create with n_devices > 1), one process, peer access over NVLink.
// ctx[0] is the primary context (the handle the caller holds); ctx[1..] are ordinary single-device contexts it owns.
struct Group {
  int n = 0;
  ctg_ctx* ctx[kMaxDevices] = {};
  DeviceWorker* worker[kMaxDevices] = {};
  // sharded proof state (multi.cu / prove.cu prove_group_*)
  Fr* d_abc[kMaxDevices] = {};        // 3 n Fr on device 0 and on every chain owner
  Fr* d_hslice[kMaxDevices] = {};     // this device's slice of h (devices other than 0)
  uint64_t abc_n = 0;                 // domain size the buffers above were sized for
  void* d_parts0 = nullptr;           // device 0: n * kPartialBytes (+ proof)
  void* h_proof = nullptr;            // pinned staging of the proof
  cudaEvent_t ev_a[kMaxDevices] = {}, ev_part[kMaxDevices] = {}, ev_h = nullptr, ev_t0 = nullptr, ev_t1 = nullptr;
  // sharded stand-alone MSM
  void* d_msm_parts0 = nullptr;       // device 0: n * sizeof(G2XYZZ) partial sums + result
  void* d_scalars[kMaxDevices] = {}; size_t scalars_bytes[kMaxDevices] = {};
  cudaEvent_t ev_msm[kMaxDevices] = {}, ev_msm_go = nullptr;
  int chain_owner[3] = {0, 0, 0};
};
// runs fn(d) on every device's worker thread (device d current) and waits; returns the first non-OK code
int32_t group_run(Group* g, const std::function<int32_t(int)>& fn);
int32_t group_run_on(Group* g, int d, const std::function<int32_t()>& fn);
}  // namespace ctg

struct ctg_ctx {
  int device = 0;
  ctg::Group* group = nullptr;   // non-null on the primary context of a multi-device context (ctg_ctx_create with n_devices > 1)
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  bool own_stream = true;
  std::string last_error;
  uint64_t launches = 0;
  int ba_levels = -1;   // batched-affine tree levels per MSM: -1 = chosen per call (msm.cuh ba_num_levels), 0..6 forced
  std::map<int, ctg::NttDomain> domains;
  ctg::MsmWorkspace ws[2];
  // split MSM (abi.cu msm_split): point-range parts of one large MSM on their own streams, so that one part's scalar upload,
  // bucket sort and latency-bound reduction overlap another part's accumulate kernel
  static constexpr int kSplit = 4;
  cudaStream_t split_st[kSplit] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t split_done[kSplit] = {nullptr, nullptr, nullptr, nullptr}, split_start = nullptr;
  ctg::MsmWorkspace split_ws[kSplit];
  void* split_part = nullptr;   // kSplit XYZZ<Fq2> partial sums
  int msm_split = -1;           // parts per large MSM: -1 = automatic, 1 = never split
  ctg::ProveLane lanes[ctg::kMaxLanes];
  void* pinned_proofs = nullptr; size_t pinned_proofs_bytes = 0;
  void* scratch = nullptr; size_t scratch_bytes = 0;     // general device scratch (staging)
  void* pinned = nullptr; size_t pinned_bytes = 0;       // pinned host staging
  cudaEvent_t ev[16];
  bool events_ready = false;
};

struct ctg_bases {
  ctg::Bases b;
  // multi-device context: point-range parts, one per device (b is then unused except for group / n / plan of part 0)
  std::vector<ctg_bases*> parts;
  std::vector<uint64_t> part_lo;   // first point of each part
};

struct ctg_pk {
  uint64_t m, ni, nw, n;
  int log_n;
  // point-range shard (multi-GPU): this key holds positions [lo, lo + cnt) of each extended query
  int shard_index = 0, shard_count = 1;
  uint64_t ext_lo = 0, ext_cnt = 0;  // of a_ext / b1_ext / b2_ext (m + 3 positions)
  uint64_t l_lo = 0, l_cnt = 0;      // of l (num_witness positions)
  uint64_t h_lo = 0, h_cnt = 0;      // of h_ext (n positions, bit-reversed order)
  ctg::Bases a_ext, b1_ext, b2_ext, l, h_ext;
  // small whole keys (domain <= 2^15, library-chosen windows): a second copy of the five tables laid out for the throughput-optimal
  // window of ctg_prove_batch (wider windows, more precomputed multiples; a few tens of MB) -- ctg_prove keeps the latency-optimal one
  ctg_pk* batch_twin = nullptr;
  std::vector<ctg_pk*> parts;        // multi-device context: one shard per device (the members above are then unused)
};
struct ctg_r1cs {
  uint64_t ni, nw, nc, n;
  int log_n;
  uint64_t* row_ptr[3];
  uint32_t* col[3];
  ctg::Fr* val[3];
  uint64_t nnz[3];
  std::vector<ctg_r1cs*> parts;      // multi-device context: a replica per device that computes witness-map chains (nullptr elsewhere)
};

// ark_groth16::ProvingKey (+ VerifyingKey) resident in HBM in arkworks' order: produced by key generation (setup.cu) or by
// reading a circuit.zkey image (codec.cu)
struct ctg_params {
  uint64_t ni = 0, nw = 0, nc = 0, m = 0, n = 0;
  ctg::G1Affine *a_query = nullptr, *b_g1_query = nullptr, *h_query = nullptr, *l_query = nullptr, *gamma_abc_g1 = nullptr;
  ctg::G2Affine* b_g2_query = nullptr;
  ctg::G1Affine* singles_g1 = nullptr;  // alpha_g1, beta_g1, delta_g1
  ctg::G2Affine* singles_g2 = nullptr;  // beta_g2, gamma_g2, delta_g2
};

namespace ctg {

#define CTG_CUDA(ctx, call)                                                                  \
  do {                                                                                       \
    cudaError_t e__ = (call);                                                                \
    if (e__ != cudaSuccess) {                                                                \
      (ctx)->last_error = std::string(#call) + ": " + cudaGetErrorString(e__);               \
      return CTG_ERR_CUDA;                                                                   \
    }                                                                                        \
  } while (0)
#define CTG_TRY(expr)            \
  do {                           \
    int32_t rc__ = (expr);       \
    if (rc__ != CTG_OK) return rc__; \
  } while (0)
#define CTG_KERNEL_CHECK(ctx)                                                                \
  do {                                                                                       \
    (ctx)->launches++;                                                                       \
    cudaError_t e__ = cudaGetLastError();                                                    \
    if (e__ != cudaSuccess) {                                                                \
      (ctx)->last_error = std::string("kernel launch: ") + cudaGetErrorString(e__);          \
      return CTG_ERR_CUDA;                                                                   \
    }                                                                                        \
  } while (0)

int32_t fail(ctg_ctx* ctx, int32_t code, const std::string& msg);
// prove.cu pieces the multi-device orchestration (multi.cu) drives
int32_t lane_init(ctg_ctx* ctx, int idx);
int32_t lane_bufs(ctg_ctx* ctx, ProveLane& L, uint64_t m, uint64_t n, Fr* d_abc_ext, LaneBufs* out);
int32_t prove_group_a(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r, const ctg_fr* z, const Fr& r_m, const Fr& s_m, uint32_t chain_mask, Fr* d_abc,
                      cudaEvent_t ev_done);
int32_t prove_group_b(Group* g, const ctg_r1cs* r0, const Fr& r_m, const Fr& s_m);
int32_t prove_group_c(Group* g, int d, const ctg_pk* pk, const ctg_r1cs* r);
int32_t prove_group_d(Group* g, const Fr& r_m, const Fr& s_m, ctg_proof* out);
int32_t ensure_scratch(ctg_ctx* ctx, size_t bytes);
int32_t ensure_pinned(ctg_ctx* ctx, size_t bytes);

// ntt.cu
int32_t ntt_get_domain(ctg_ctx* ctx, int log_n, NttDomain** out);
// natural order in, natural order out (ark-poly Radix2EvaluationDomain::{fft,ifft,coset_fft,coset_ifft}_in_place)
int32_t ntt_run(ctg_ctx* ctx, Fr* d_data, int log_n, int inverse, int coset, cudaStream_t st);
// witness-map building blocks: no bit-reversal passes (see ntt.cu)
// batch: that many transforms of 2^log_n elements stored back to back, one launch per pass for all of them
int32_t ntt_ifft_then_coset_fft(ctg_ctx* ctx, Fr* d_data, int log_n, cudaStream_t st, uint32_t batch = 1);      // natural in, natural out
int32_t ntt_coset_ifft_bitrev(ctg_ctx* ctx, Fr* d_data, int log_n, cudaStream_t st, uint32_t batch = 1);        // natural in, bit-reversed out

int32_t fr_pow_table(ctg_ctx* ctx, Fr* d_out, uint64_t n, const Fr& base, const Fr& scale, cudaStream_t st);  // out[i] = scale * base^i
Fr fr_root_of_unity(int log_n);   // generator of the 2^log_n domain (ark-poly Radix2EvaluationDomain::group_gen), Montgomery
Fr fr_from_u64(uint64_t x);

// setup.cu
void params_free(ctg_params* p);
int32_t params_alloc(ctg_ctx* ctx, uint64_t ni, uint64_t nw, uint64_t nc, uint64_t n, ctg_params** out);   // device arrays sized, contents undefined

// msm_g1.cu / msm_g2.cu
MsmPlan msm_make_plan(uint64_t n, int copies_wanted, int c_override, size_t point_bytes, bool batch_throughput = false);
// batch / vec_stride / sorted: see MsmHost::run (msm.cuh)
int32_t msm_g1_run(ctg_ctx* ctx, MsmWorkspace& ws, const Bases& bases, const uint32_t* d_scalars, uint64_t n, bool scalars_mont,
                   G1XYZZ* d_out, cudaStream_t st, float* stage_ms, uint32_t batch = 1, uint64_t vec_stride = 0, const MsmWorkspace* sorted = nullptr);
int32_t msm_g2_run(ctg_ctx* ctx, MsmWorkspace& ws, const Bases& bases, const uint32_t* d_scalars, uint64_t n, bool scalars_mont,
                   G2XYZZ* d_out, cudaStream_t st, float* stage_ms, uint32_t batch = 1, uint64_t vec_stride = 0, const MsmWorkspace* sorted = nullptr);
int32_t g1_precompute(ctg_ctx* ctx, Bases& b, cudaStream_t st);
int32_t g2_precompute(ctg_ctx* ctx, Bases& b, cudaStream_t st);
int32_t g1_to_affine_dev(ctg_ctx* ctx, const G1XYZZ* d_in, G1Affine* d_out, int count, cudaStream_t st);
int32_t g2_to_affine_dev(ctg_ctx* ctx, const G2XYZZ* d_in, G2Affine* d_out, int count, cudaStream_t st);
int32_t g1_progression_dev(ctg_ctx* ctx, const G1Affine& g, const Fr& a, const Fr& b, uint64_t n, G1Affine* d_out, cudaStream_t st);
int32_t g2_progression_dev(ctg_ctx* ctx, const G2Affine& g, const Fr& a, const Fr& b, uint64_t n, G2Affine* d_out, cudaStream_t st);
int32_t g1_combine_dev(ctg_ctx* ctx, const G1XYZZ* d_in, uint32_t k, G1Affine* d_out_affine, G1XYZZ* d_out_xyzz, cudaStream_t st);
int32_t g2_combine_dev(ctg_ctx* ctx, const G2XYZZ* d_in, uint32_t k, G2Affine* d_out_affine, G2XYZZ* d_out_xyzz, cudaStream_t st);
void msm_ws_free(MsmWorkspace& ws);

}  // namespace ctg
