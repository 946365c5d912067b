/* ctg_groth16.h -- C ABI of libctg_groth16.so, a B200-native (sm_100a) Groth16 proving
 * backend for BLS12-381.  Drop-in boundary for ONE path of nulltea/contangle-zkcp:
 *     Groth16::<Bls12_381>::prove(&pk, circuit, rng)
 * as called at lib/src/zk/encryption.rs:76, lib/src/zk/verifiable_encryption.rs:92,
 * lib/src/zk/sample_entries.rs:86 and lib/src/zk/property.rs:133, i.e.
 * ark_groth16::create_random_proof / create_proof (ark-groth16 ^0.3.0, not in the tree;
 * declared at circuits-ark/Cargo.toml:17).  Constraint synthesis stays on the host
 * (reference language: Rust); everything from "R1CS matrices + full assignment" to
 * "affine (A, B, C)" runs on the GPU.  INTEGRATION.md shows the Rust `extern "C"` block
 * and the safe wrapper a maintainer would add; contangle-zkcp_b200/host/ctg_groth16.hpp
 * mirrors the ark-groth16 API shape in C++ on top of this header.
 *
 * Conventions
 *  - Plain C, no exceptions cross the boundary.  Every function returns int32_t:
 *    CTG_OK (0) or a negative CTG_ERR_*; ctg_last_error(ctx) gives the message.
 *  - Opaque handles.  A context is bound to one CUDA device (one process per GPU);
 *    handles are thread-compatible: one in-flight call per context.
 *  - Field elements are little-endian Montgomery limb images with R = 2^256 (Fr) and
 *    R = 2^384 (Fq): byte-identical to ark-ff 0.3.0 Fp256 / Fp384 ([u64; 4] / [u64; 6]).
 *    Affine points are x || y; G2 coordinates are c0 || c1.  The point at infinity is the
 *    all-zero image ((0,0) is not on either curve).  ark-ec's GroupAffine is not repr(C);
 *    the Rust shim repacks.
 *  - "host" pointers are caller-owned host memory, copied during the call.  "_dev"
 *    entry points take device pointers (e.g. torch tensors' data_ptr()).
 *  - There is no CPU fallback: without a usable CUDA device ctg_ctx_create fails.
 */
#ifndef CTG_GROTH16_H
#define CTG_GROTH16_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CTG_OK 0
#define CTG_ERR_INVALID (-1)
#define CTG_ERR_CUDA (-2)
#define CTG_ERR_OOM (-3)
#define CTG_ERR_STATE (-4)

/* scalar encodings accepted by the MSM entry points */
#define CTG_SCALARS_CANONICAL 0u /* ark-ff BigInteger256 (into_repr) images, as VariableBaseMSM takes them */
#define CTG_SCALARS_MONT 1u      /* ark-ff Fr memory images (Montgomery) */
#define CTG_MSM_TIMED 0x100u      /* record per-stage CUDA-event times (adds one stream sync); read with ctg_last_timings */

typedef struct ctg_ctx ctg_ctx;
typedef struct ctg_bases ctg_bases; /* device-resident base table (one proving-key query) */
typedef struct ctg_pk ctg_pk;       /* device-resident proving key (replaces the host ProvingKey cache,
                                       lib/src/zk/encryption.rs:22,31-32; lib/src/utils.rs:104-110) */
typedef struct ctg_r1cs ctg_r1cs;   /* device-resident R1CS matrices (ark-relations to_matrices()) */

typedef struct { uint32_t v[8]; } ctg_fr;
typedef struct { uint32_t v[12]; } ctg_fq;
typedef struct { ctg_fq x, y; } ctg_g1_affine;                     /* 96 B  */
typedef struct { ctg_fq x_c0, x_c1, y_c0, y_c1; } ctg_g2_affine;   /* 192 B */
typedef struct { ctg_fq x, y, zz, zzz; } ctg_g1_xyzz;              /* 192 B, x = X/ZZ, y = Y/ZZZ; ZZ = 0 is infinity */
typedef struct { ctg_fq c[8]; } ctg_g2_xyzz;                       /* 384 B */

/* ark_groth16::Proof<Bls12_381> { a: G1Affine, b: G2Affine, c: G1Affine } */
typedef struct { ctg_g1_affine a; ctg_g2_affine b; ctg_g1_affine c; } ctg_proof;

/* one GPU's share of a point-range sharded proof: partial sums of the five MSMs (see ctg_prove_partial) */
typedef struct { ctg_g1_xyzz a, b1, l, h; ctg_g2_xyzz b2; } ctg_proof_partial;               /* 1152 B */

/* ark_groth16::ProvingKey<Bls12_381> as pointers + lengths. */
typedef struct {
  const ctg_g1_affine* alpha_g1; /* vk.alpha_g1 */
  const ctg_g1_affine* beta_g1;
  const ctg_g1_affine* delta_g1;
  const ctg_g2_affine* beta_g2;  /* vk.beta_g2 */
  const ctg_g2_affine* delta_g2; /* vk.delta_g2 */
  const ctg_g1_affine* a_query;    uint64_t a_len;    /* m = num_instance + num_witness */
  const ctg_g1_affine* b_g1_query; uint64_t b_g1_len; /* m */
  const ctg_g2_affine* b_g2_query; uint64_t b_g2_len; /* m */
  const ctg_g1_affine* h_query;    uint64_t h_len;    /* n - 1 */
  const ctg_g1_affine* l_query;    uint64_t l_len;    /* num_witness */
} ctg_pk_view;

/* One R1CS matrix in CSR form.  Columns index z = instance (z[0] = 1) ++ witness. */
typedef struct {
  const uint64_t* row_ptr; /* num_constraints + 1 */
  const uint32_t* col;     /* nnz */
  const ctg_fr* val;       /* nnz, Montgomery */
} ctg_csr;

/* Per-stage device times (ms) of the last ctg_prove / ctg_msm_* call, CUDA events on the
 * context's stream.  Stage names follow arkworks' start_timer! spans. */
typedef struct {
  float h2d_ms;          /* assignment upload */
  float witness_map_ms;  /* "R1CS to QAP witness map": SpMV + 7 NTT + pointwise */
  float msm_h_ms, msm_l_ms, msm_a_ms, msm_b_g1_ms, msm_b_g2_ms;
  float finalize_ms;     /* s*A + r*B1 + ..., to_affine, D2H */
  float total_ms;
  /* inside the last MSM: */
  float msm_sort_ms;       /* digit recode + bucket counting sort */
  float msm_accumulate_ms; /* bucket accumulation (dominant kernel) */
  float msm_reduce_ms;     /* bucket-sum reduction + window combine */
} ctg_timings;

/* ---- context ---------------------------------------------------------------------- */
/* n_devices == 1: the context is bound to that GPU.
 * n_devices in 2..8 (one process, several GPUs of one NVLink / NVSwitch box; peer access between all of them is required):
 * device_ids[0] is the PRIMARY device -- every entry point without a sharded form (NTT, key generation, codecs, Poseidon, ...)
 * runs there exactly as on a one-device context -- and the calls on the prover path shard over all devices by themselves:
 *   ctg_pk_upload      keeps a point-range shard of every query on each device (1/N of the key per GPU)
 *   ctg_r1cs_upload    replicates the matrices on the devices that compute witness-map chains
 *   ctg_prove[_batch]  the five MSMs by point range, the three NTT chains on three devices, one exchange of the chains and of the
 *                      slices of h with cudaMemcpyPeerAsync over NVLink, 1152 B of partial sums per device summed on the primary;
 *                      no host round trip between the phases; the proof is bit-identical to the one-device proof
 *   ctg_bases_upload_* / ctg_bases_progression_* / ctg_msm_g1 / ctg_msm_g2 [_dev]   the same for a stand-alone MSM (scalars of
 *                      the _dev forms live on the primary device; so does the projective result)
 * A device id may be listed more than once (its shards then share that GPU: no speed-up, but a one-GPU machine can run this path).
 * Handles made by a multi-device context work with that context only (CTG_ERR_STATE otherwise); the shard / partial entry
 * points (ctg_pk_upload_shard, ctg_prove_partial, ...) are the one-process-per-GPU form of the same scheme and refuse them. */
int32_t ctg_ctx_create(const int* device_ids, int n_devices, ctg_ctx** out);
int32_t ctg_ctx_device_count(ctg_ctx* ctx); /* 1, or the n_devices of a multi-device context */
void ctg_ctx_destroy(ctg_ctx* ctx);
const char* ctg_last_error(ctg_ctx* ctx);
/* Run all work of this context on a caller-owned cudaStream_t (e.g. torch's current stream). */
int32_t ctg_ctx_set_stream(ctg_ctx* ctx, void* cuda_stream);
int32_t ctg_sync(ctg_ctx* ctx);
/* Bucket accumulation schedule of the MSMs: `levels` pairwise batched-affine tree levels (5M + 1S per addition, one shared
 * binary-GCD inversion per 64 additions) before the XYZZ accumulation (8M + 2S).  -1 (default) = chosen per call from the
 * size; 0 = XYZZ only; 1..6 forced (tests, sweeps).  The result is the same group element either way. */
int32_t ctg_ctx_set_ba_levels(ctg_ctx* ctx, int levels);
/* Large stand-alone MSMs (ctg_msm_*) are cut into `parts` point ranges that run as complete MSMs on their own streams, so that
 * one part's scalar upload, bucket sort and reduction overlap another part's accumulate kernel; the partial sums are added on
 * the device.  -1 (default) = automatic (2 for host scalars from 2^21 points up, else 1), 1 = never, up to 4. */
int32_t ctg_ctx_set_msm_split(ctg_ctx* ctx, int parts);
uint64_t ctg_launch_count(ctg_ctx* ctx); /* kernels launched by this context so far */
int32_t ctg_last_timings(ctg_ctx* ctx, ctg_timings* out);

/* ---- device memory helpers (for callers that do not own a CUDA allocator) ----------- */
int32_t ctg_dev_alloc(ctg_ctx* ctx, size_t bytes, void** d_out);
int32_t ctg_dev_free(ctg_ctx* ctx, void* d_ptr);
int32_t ctg_h2d(ctg_ctx* ctx, void* d_dst, const void* h_src, size_t bytes);
int32_t ctg_d2h(ctg_ctx* ctx, void* h_dst, const void* d_src, size_t bytes);
/* Page-locked host memory for assignments and scalars: host -> device copies from it run at full PCIe rate and asynchronously
 * (the Rust shim packs z = instance ++ witness straight into such a buffer).  Pageable buffers are accepted everywhere, slower. */
int32_t ctg_host_alloc(ctg_ctx* ctx, size_t bytes, void** h_out);
int32_t ctg_host_free(ctg_ctx* ctx, void* h_ptr);

/* ---- field kernels (replaces ark-ff Fp256/Fp384 mul_assign/add_assign/sub_assign; used by
 *      tests and the ncu integer-pipe captures) --------------------------------------- */
#define CTG_FIELD_FR 0
#define CTG_FIELD_FQ 1
#define CTG_FIELD_FQ2 2
#define CTG_OP_MUL 0
#define CTG_OP_ADD 1
#define CTG_OP_SUB 2
#define CTG_OP_INV 3 /* out[i] = a[i]^-1 (b ignored, may equal a); 0 -> 0.  The binary-GCD inversion the batched-affine MSM uses */
/* CTG_FIELD_FQ2 only -- the lane-paired Fq2 arithmetic of the G2 bucket-reduction tail (one component per lane of a lane pair),
 * exposed so that it can be checked element by element: a * b, a^2, a * b - b * a^2, and [a == 0] (as the element 1 or 0) */
#define CTG_OP_PAIRED_MUL 4
#define CTG_OP_PAIRED_SQR 5
#define CTG_OP_PAIRED_MUL2SUB 6
#define CTG_OP_PAIRED_IS_ZERO 7
/* out[i] = a[i] (op) b[i]; host buffers of n elements (32 / 48 / 96 bytes each) */
int32_t ctg_field_op_batch(ctg_ctx* ctx, int field, int op, const void* a, const void* b, void* out, size_t n);
/* ctg_field_op_batch for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO counterpart in the reference.  Same arguments and results as
 * ctg_field_op_batch (host buffers; op = CTG_OP_MUL / ADD / SUB only), but operands and result are staged in a fixed 3 MiB
 * file-scope device array instead of the context's growable scratch, and the call performs no cudaMalloc / cudaFree:
 *   limit   3 * n * element_bytes <= 3 MiB, i.e. n <= 32 768 (Fr), 21 845 (Fq), 10 922 (Fq2); beyond it CTG_ERR_INVALID, no fallback
 *   cost    3 MiB of HBM on every GPU that loads the library, until the process exits
 *   sharing every call on a device uses the SAME array, which keeps the previous call's contents (the result region is cleared at
 *           the start of each call): calls that overlap on one GPU -- two contexts, two streams, two threads -- race; serialise them */
int32_t ctg_field_op_batch_noalloc(ctg_ctx* ctx, int field, int op, const void* a, const void* b, void* out, size_t n);
int32_t ctg_fr_mul_batch(ctg_ctx* ctx, const ctg_fr* a, const ctg_fr* b, ctg_fr* out, size_t n);
int32_t ctg_fq_mul_batch(ctg_ctx* ctx, const ctg_fq* a, const ctg_fq* b, ctg_fq* out, size_t n);
/* Measured integer-pipe peaks of this GPU: IMAD (32-bit) instructions/s over all SMs, and the
 * sustained rate of dependent Fq Montgomery products (300 MAC32 each). */
int32_t ctg_imad_peak_probe(ctg_ctx* ctx, double* imad_per_s, double* fq_mul_per_s);

/* ---- NTT (replaces ark-poly 0.3.0 Radix2EvaluationDomain::{fft,ifft,coset_fft,coset_ifft}_in_place
 *      over Fr; natural order in, natural order out; ifft scales by 1/n; coset shift g = 7) */
int32_t ctg_ntt_fr(ctg_ctx* ctx, ctg_fr* host_data, uint32_t log_n, int inverse, int coset);
int32_t ctg_ntt_fr_dev(ctg_ctx* ctx, void* d_data, uint32_t log_n, int inverse, int coset);

/* ---- base tables and MSM (replaces ark-ec 0.3.0 VariableBaseMSM::multi_scalar_mul) ---- */
/* host_points may also be a device pointer (unified addressing; e.g. from ctg_params_pk_view).
 * copies >= 1: keep 2^(c*G*j) * P for j < copies in HBM so that only G = ceil(W/copies) bucket
 * sets are reduced (0 = let the library choose); c_bits = 0 lets the library choose. */
int32_t ctg_bases_upload_g1(ctg_ctx* ctx, const ctg_g1_affine* host_points, uint64_t n, int copies, int c_bits, ctg_bases** out);
int32_t ctg_bases_upload_g2(ctg_ctx* ctx, const ctg_g2_affine* host_points, uint64_t n, int copies, int c_bits, ctg_bases** out);
/* synthetic tables generated on the device: P_i = (a + i*b) * g  (distinct subgroup points with known dlogs) */
int32_t ctg_bases_progression_g1(ctg_ctx* ctx, const ctg_g1_affine* g, const ctg_fr* a, const ctg_fr* b, uint64_t n, int copies, int c_bits, ctg_bases** out);
int32_t ctg_bases_progression_g2(ctg_ctx* ctx, const ctg_g2_affine* g, const ctg_fr* a, const ctg_fr* b, uint64_t n, int copies, int c_bits, ctg_bases** out);
int32_t ctg_bases_download(ctg_ctx* ctx, const ctg_bases* bases, uint64_t first, uint64_t count, void* host_points);
void ctg_bases_free(ctg_ctx* ctx, ctg_bases* bases);
uint64_t ctg_bases_len(const ctg_bases* bases);
/* the MSM schedule chosen for this table: window bits c, windows ceil(256/c), precomputed copies kept, bucket sets */
typedef struct { int32_t window_bits, windows, copies, bucket_sets; uint32_t buckets_per_set; } ctg_msm_plan;
int32_t ctg_bases_plan(const ctg_bases* bases, ctg_msm_plan* out);

/* sum_{i<n} scalars[i] * bases[i], n <= len(bases); affine result in host memory */
int32_t ctg_msm_g1(ctg_ctx* ctx, const ctg_bases* bases, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, ctg_g1_affine* out);
int32_t ctg_msm_g2(ctg_ctx* ctx, const ctg_bases* bases, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, ctg_g2_affine* out);
/* scalars already in HBM; d_out_xyzz (device, 192 / 384 B) receives the projective sum; no host sync */
int32_t ctg_msm_g1_dev(ctg_ctx* ctx, const ctg_bases* bases, const void* d_scalars, uint64_t n, uint32_t flags, void* d_out_xyzz);
int32_t ctg_msm_g2_dev(ctg_ctx* ctx, const ctg_bases* bases, const void* d_scalars, uint64_t n, uint32_t flags, void* d_out_xyzz);
/* multi-GPU: each rank runs the MSM over its point range and returns the projective partial sum;
 * the partials are gathered (NCCL all_gather of 192 B per rank) and combined on one rank. */
int32_t ctg_msm_g1_partial(ctg_ctx* ctx, const ctg_bases* bases, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, ctg_g1_xyzz* out);
int32_t ctg_msm_g2_partial(ctg_ctx* ctx, const ctg_bases* bases, const ctg_fr* host_scalars, uint64_t n, uint32_t flags, ctg_g2_xyzz* out);
int32_t ctg_g1_combine(ctg_ctx* ctx, const ctg_g1_xyzz* host_partials, uint32_t k, ctg_g1_affine* out);
int32_t ctg_g2_combine(ctg_ctx* ctx, const ctg_g2_xyzz* host_partials, uint32_t k, ctg_g2_affine* out);
int32_t ctg_g1_combine_dev(ctg_ctx* ctx, const void* d_partials, uint32_t k, void* d_out_affine);
int32_t ctg_g2_combine_dev(ctg_ctx* ctx, const void* d_partials, uint32_t k, void* d_out_affine);

/* ---- verifier side of the same MSM ---------------------------------------------
 * ark_groth16::prepare_inputs(pvk, public_inputs): gamma_abc_g1[0] + sum_i public_inputs[i] * gamma_abc_g1[i + 1], the only
 * size-dependent step of Groth16::verify (lib/src/zk/encryption.rs:152, lib/src/zk/sample_entries.rs:126: n + 2 public inputs
 * for an n-chunk ciphertext, circuits-ark/src/encryption.rs:139-152).  gamma_abc_g1: a base table of exactly n_inputs + 1 points
 * (CTG_ERR_INVALID otherwise, arkworks' SynthesisError::MalformedVerifyingKey); public_inputs: Montgomery Fr images. */
int32_t ctg_prepare_inputs(ctg_ctx* ctx, const ctg_bases* gamma_abc_g1, const ctg_fr* public_inputs, uint64_t n_inputs, ctg_g1_affine* out);

/* ---- proving key / R1CS caches and the prover ---------------------------------------- */
/* copies / c_bits as for ctg_bases_upload_* (applied to every query). */
int32_t ctg_pk_upload(ctg_ctx* ctx, const ctg_pk_view* host_pk, uint64_t num_instance, uint64_t num_constraints, int copies, int c_bits, ctg_pk** out);
/* Multi-GPU: shard `shard_index` of `shard_count` keeps only its point range of every query (one process / context per
 * GPU; host_pk is the full key on every rank).  A sharded key works with ctg_prove_partial only. */
int32_t ctg_pk_upload_shard(ctg_ctx* ctx, const ctg_pk_view* host_pk, uint64_t num_instance, uint64_t num_constraints, int copies, int c_bits,
                            int shard_index, int shard_count, ctg_pk** out);
/* Weighted shards: keep positions [total * lo_num / den, total * hi_num / den) of every query.  The ranks' (lo_num, hi_num) must
 * tile [0, den); ranks that also compute NTT chains of the witness map can be given fewer points (ctg/sharding.py: shard_bounds).
 * ctg_pk_upload_shard(index, count) is the special case (index, index + 1, count). */
int32_t ctg_pk_upload_range(ctg_ctx* ctx, const ctg_pk_view* host_pk, uint64_t num_instance, uint64_t num_constraints, int copies, int c_bits,
                            uint64_t lo_num, uint64_t hi_num, uint64_t den, ctg_pk** out);
void ctg_pk_free(ctg_ctx* ctx, ctg_pk* pk);
int32_t ctg_r1cs_upload(ctg_ctx* ctx, uint64_t num_instance, uint64_t num_witness, uint64_t num_constraints, const ctg_csr* a,
                        const ctg_csr* b, const ctg_csr* c, ctg_r1cs** out);
void ctg_r1cs_free(ctg_ctx* ctx, ctg_r1cs* r1cs);

/* ark_groth16 R1CStoQAP::witness_map: h_out receives n = next_pow2(num_constraints + num_instance)
 * coefficients in natural order (host). */
int32_t ctg_witness_map(ctg_ctx* ctx, const ctg_r1cs* r1cs, const ctg_fr* z, ctg_fr* h_out);

/* ark_groth16::create_proof(circuit, pk, r, s) after synthesis: z = instance ++ witness
 * (m elements, z[0] = 1, Montgomery), r and s drawn by the caller (Fr::rand, r first). */
int32_t ctg_prove(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r1cs, const ctg_fr* z, const ctg_fr* r, const ctg_fr* s,
                  ctg_proof* out, ctg_timings* timings /* may be NULL */);
/* n_proofs independent assignments (z_batch: n_proofs * m), rs_batch: n_proofs * 2 (r, s); the proofs are pipelined
 * over several stream sets so that one proof's latency-bound tails overlap the next one's kernels */
int32_t ctg_prove_batch(ctg_ctx* ctx, const ctg_pk* pk, const ctg_r1cs* r1cs, uint32_t n_proofs, const ctg_fr* z_batch,
                        const ctg_fr* rs_batch, ctg_proof* out);

/* Sharded proof: every rank computes the witness map (replicated, cheap) and the five MSMs over ITS point range, giving
 * 1152 bytes of partial sums; the partials are gathered (NCCL all_gather over NVLink) and one rank sums them and
 * finishes the proof: C = s*A + r*B1 + L + H, affine (A, B, C).  Replaces the same create_proof call as ctg_prove. */
int32_t ctg_prove_partial(ctg_ctx* ctx, const ctg_pk* pk_shard, const ctg_r1cs* r1cs, const ctg_fr* z, const ctg_fr* r, const ctg_fr* s,
                          ctg_proof_partial* out, ctg_timings* timings /* may be NULL */);
int32_t ctg_prove_combine(ctg_ctx* ctx, const ctg_proof_partial* partials, uint32_t k, const ctg_fr* r, const ctg_fr* s, ctg_proof* out);
/* The same with the witness map distributed over the ranks (the seven NTTs of R1CStoQAP::witness_map form three independent
 * chains a, b, c -- SpMV, ifft, coset_fft -- joined by the pointwise step and the last coset_ifft).
 *   begin : uploads z, starts this shard's A / B1 / B2 / L MSMs, and computes the chains in chain_mask (bit 0 = a, 1 = b, 2 = c)
 *           into d_abc[k * n .. (k + 1) * n) -- a DEVICE buffer of 3 * n Fr owned by the caller; returns when they are complete;
 *   (the caller broadcasts each chain from its owner to every rank: NCCL over NVLink, n * 32 bytes per chain; the MSMs
 *    started by begin keep the SMs busy meanwhile)
 *   finish: (a b - c)/Z, coset_ifft, this shard's H MSM; writes the 1152 bytes of partial sums.
 * begin(chain_mask = 7) + finish with no exchange equals ctg_prove_partial.  One begin/finish pair in flight per context. */
int32_t ctg_prove_partial_begin(ctg_ctx* ctx, const ctg_pk* pk_shard, const ctg_r1cs* r1cs, const ctg_fr* z, const ctg_fr* r, const ctg_fr* s,
                                uint32_t chain_mask, void* d_abc);
int32_t ctg_prove_partial_finish(ctg_ctx* ctx, const ctg_pk* pk_shard, const ctg_r1cs* r1cs, const ctg_fr* r, const ctg_fr* s, void* d_abc,
                                 ctg_proof_partial* out);

/* ---- key generation ---------------------------------------------------------------
 * ark_groth16::generate_parameters(circuit, alpha, beta, gamma, delta, g1_generator, g2_generator, rng) after synthesis
 * (the body of Groth16::setup / generate_random_parameters: lib/src/zk/encryption.rs:169, sample_entries.rs:141,
 * property.rs:192).  The caller draws the trapdoor -- alpha, beta, gamma, delta, the two generators, then
 * tau = domain.sample_element_outside_domain(rng), in arkworks' order -- and passes it in (Montgomery Fr images, affine
 * generators); the R1CS is the same to_matrices() triple ctg_r1cs_upload takes.  The key stays in HBM:
 * ctg_params_pk_view hands DEVICE pointers that ctg_pk_upload / ctg_pk_upload_shard / ctg_bases_upload_g1 accept as they
 * accept host pointers; ctg_params_download copies any part to caller-owned host buffers (for serialisation).
 * Errors: CTG_ERR_INVALID if gamma or delta is 0 or tau lies in the evaluation domain (arkworks panics on both). */
typedef struct { ctg_fr alpha, beta, gamma, delta, tau; ctg_g1_affine g1; ctg_g2_affine g2; } ctg_trapdoor;
typedef struct ctg_params ctg_params; /* ark_groth16::ProvingKey (with its VerifyingKey) resident in HBM */
typedef struct {                      /* caller-allocated host buffers; NULL members are skipped */
  ctg_g1_affine* alpha_g1; ctg_g1_affine* beta_g1; ctg_g1_affine* delta_g1;
  ctg_g2_affine* beta_g2; ctg_g2_affine* gamma_g2; ctg_g2_affine* delta_g2;
  ctg_g1_affine* gamma_abc_g1; /* num_instance */
  ctg_g1_affine* a_query;      /* m */
  ctg_g1_affine* b_g1_query;   /* m */
  ctg_g2_affine* b_g2_query;   /* m */
  ctg_g1_affine* h_query;      /* n - 1 */
  ctg_g1_affine* l_query;      /* num_witness */
} ctg_params_out;
int32_t ctg_generate_parameters(ctg_ctx* ctx, uint64_t num_instance, uint64_t num_witness, uint64_t num_constraints, const ctg_csr* a,
                                const ctg_csr* b, const ctg_csr* c, const ctg_trapdoor* trapdoor, ctg_params** out);
int32_t ctg_params_pk_view(ctg_ctx* ctx, const ctg_params* params, ctg_pk_view* out, const ctg_g1_affine** gamma_abc_g1 /* may be NULL */,
                           uint64_t* gamma_abc_len /* may be NULL */);
int32_t ctg_params_download(ctg_ctx* ctx, const ctg_params* params, const ctg_params_out* dst);
void ctg_params_free(ctg_ctx* ctx, ctg_params* params);

/* ---- arkworks canonical wire formats (ark-serialize 0.3 conventions, see csrc/codec.cu) --------------
 * circuit.zkey = ProvingKey::serialize_unchecked (lib/src/utils.rs:85-110): decoded ON THE DEVICE straight into the packed
 * Montgomery layout (a ctg_params, then ctg_params_pk_view -> ctg_pk_upload), or encoded from it.  Passing out = NULL to a
 * serialize call only reports the size in *written.  verification.key = VerifyingKey::serialize (compressed,
 * lib/src/utils.rs:94).  Reading rejects non-canonical field elements and invalid flag bits with CTG_ERR_INVALID
 * (arkworks: SerializationError::InvalidData); like deserialize_unchecked it does not test curve or subgroup membership. */
int32_t ctg_params_deserialize_unchecked(ctg_ctx* ctx, const uint8_t* key_image, uint64_t len, ctg_params** out);
int32_t ctg_params_serialize_unchecked(ctg_ctx* ctx, const ctg_params* params, uint8_t* out, uint64_t capacity, uint64_t* written);
int32_t ctg_params_serialize_vk(ctg_ctx* ctx, const ctg_params* params, uint8_t* out, uint64_t capacity, uint64_t* written);
/* Proof::serialize / Proof::deserialize (compressed, 48 + 96 + 48 = 192 bytes: ark_to_bytes / ark_from_bytes,
 * circuits-ark/src/utils.rs:12-22; lib/src/zk/encryption.rs:78-84,135-141).  Host-only, no context needed.
 * check_subgroup != 0 adds arkworks' r * P == O test of the checked deserialize. */
#define CTG_PROOF_BYTES 192
int32_t ctg_proof_serialize(const ctg_proof* proof, uint8_t* out192);
int32_t ctg_proof_deserialize(const uint8_t* in192, int check_subgroup, ctg_proof* out);

/* ---- bulk Poseidon over Fr: sponge hashes and the ciphertext Merkle tree ----------------------
 * Replaces, for many items at once, the native (out-of-circuit) Poseidon work upstream of the prover:
 *   circuits-ark/src/sample_entries.rs:127-134  build_merkle_tree = MerkleTree::new(params, params, leaves) with one field
 *       element per leaf (leaf digest = poseidon::CRH([e]), inner digest = TwoToOneCRH(l, r) = CRH([l, r]));
 *   circuits-ark/src/encryption.rs:154-172      PoseidonSponge::new(params).absorb(..).squeeze_field_elements(1).
 * Sponge as ark-sponge ^0.3.0 PoseidonSponge: width 3 (rate 2, capacity 1, rate cells first), full_rounds/2 full rounds,
 * partial_rounds partial rounds (S-box on cell 0), full_rounds/2 full rounds; every round adds ark[round][0..3), applies
 * x^alpha and multiplies by mds (new[i] = sum_j mds[i][j] * state[j]).  Parameters are the caller's PoseidonParameters
 * (circuits-ark/src/poseidon.rs:13-41): mds = 3 x 3 row-major, ark = (full + partial) x 3, Montgomery Fr images.
 * Errors: CTG_ERR_INVALID for an odd or zero full_rounds, full + partial > 128, alpha < 2, k = 0, or a leaf count that is
 * not a power of two >= 2 (arkworks' MerkleTree::new asserts the same). */
typedef struct ctg_poseidon ctg_poseidon;
int32_t ctg_poseidon_new(ctg_ctx* ctx, uint32_t full_rounds, uint32_t partial_rounds, uint32_t alpha, const ctg_fr* mds, const ctg_fr* ark,
                         ctg_poseidon** out);
void ctg_poseidon_free(ctg_ctx* ctx, ctg_poseidon* p);
/* out[i] = the first squeezed element after absorbing in[i*k .. i*k + k) into a fresh sponge (k >= 1; k = 1 is the leaf CRH,
 * k = 2 the two-to-one CRH, k = 2 with an affine point's (x, y) the ElGamal pad of encrypt()).  _dev: device pointers. */
int32_t ctg_poseidon_hash_batch(ctg_ctx* ctx, const ctg_poseidon* p, const ctg_fr* in, uint64_t n, uint32_t k, ctg_fr* out);
int32_t ctg_poseidon_hash_batch_dev(ctg_ctx* ctx, const ctg_poseidon* p, const void* d_in, uint64_t n, uint32_t k, void* d_out);
/* Whole tree over n = 2^d leaves: leaf_digests[n] (may be NULL in the host variant) and inner[n - 1] in heap order -- root at 0,
 * children of node i at 2i + 1 and 2i + 2, the last level's children being leaf digests 2j and 2j + 1 (the order of arkworks'
 * MerkleTree::non_leaf_nodes / leaf_nodes, from which generate_proof reads its Path). */
int32_t ctg_poseidon_merkle_tree(ctg_ctx* ctx, const ctg_poseidon* p, const ctg_fr* leaves, uint64_t n, ctg_fr* leaf_digests, ctg_fr* inner);
int32_t ctg_poseidon_merkle_tree_dev(ctg_ctx* ctx, const ctg_poseidon* p, const void* d_leaves, uint64_t n, void* d_leaf_digests, void* d_inner);

/* ---- bulk scalar multiplication on a twisted Edwards curve over Fr (JubJub): the group part of hashed ElGamal ----------
 * out[i] = scalars[i] * points[n_points == 1 ? 0 : i] on -x^2 + y^2 = 1 + d x^2 y^2 (a = -1; d a non-square, e.g. JubJub's
 * -10240/10241), i.e. for many randomness values at once the c1 = G.mul(r), p_r = pk.mul(r), into_affine() of encrypt()
 * (circuits-ark/src/encryption.rs:160-165) and the pk = G.mul(sk) of keygen (:126-137).  points / out: affine (x, y) pairs of
 * Montgomery Fr images (ark_ed_on_bls12_381::EdwardsAffine field by field); scalars: 256-bit little-endian integers, 8 words
 * each (the scalar field's into_repr(); any 256-bit value is accepted); d: Montgomery.  The identity is (0, 1).
 * n_points must be 1 (one shared base) or n.  The ElGamal pad is then ctg_poseidon_hash_batch over the p_r pairs with k = 2 and
 * c2 = pad + m (ctg_field_op_batch).  Points are not checked to be on the curve (arkworks does not check in mul either).
 * Timing: variable-time in the scalar, like ark-ec's `mul` (bit-dependent branches / digit-indexed table); the library zeroes
 * the scratch that held the scalars before returning. */
int32_t ctg_te_mul_batch(ctg_ctx* ctx, const ctg_fr* d, const ctg_fr* points, uint64_t n_points, const uint32_t* scalars, uint64_t n, ctg_fr* out);
int32_t ctg_te_mul_batch_dev(ctg_ctx* ctx, const ctg_fr* d, const void* d_points, uint64_t n_points, const void* d_scalars, uint64_t n, void* d_out);

#ifdef __cplusplus
}
#endif
#endif /* CTG_GROTH16_H */
