//! Raw binding of include/ctg_groth16.h.  One `extern "C"` item per declaration in the header; layouts are
//! `#[repr(C)]` copies of the header's structs.  Nothing here is safe; see the `ctg-groth16` crate.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

pub const CTG_OK: i32 = 0;
pub const CTG_ERR_INVALID: i32 = -1;
pub const CTG_ERR_CUDA: i32 = -2;
pub const CTG_ERR_OOM: i32 = -3;
pub const CTG_ERR_STATE: i32 = -4;
pub const CTG_SCALARS_CANONICAL: u32 = 0;
pub const CTG_SCALARS_MONT: u32 = 1;
pub const CTG_MSM_TIMED: u32 = 0x100;

#[repr(C)] pub struct ctg_ctx { _p: [u8; 0] }
#[repr(C)] pub struct ctg_bases { _p: [u8; 0] }
#[repr(C)] pub struct ctg_pk { _p: [u8; 0] }
#[repr(C)] pub struct ctg_r1cs { _p: [u8; 0] }
#[repr(C)] pub struct ctg_params { _p: [u8; 0] }
#[repr(C)] pub struct ctg_poseidon { _p: [u8; 0] }

/// ark-ff Fp256 memory image: [u64; 4] little-endian Montgomery limbs == [u32; 8]
#[repr(C)] #[derive(Clone, Copy, Default)] pub struct ctg_fr { pub v: [u32; 8] }
/// ark-ff Fp384 memory image
#[repr(C)] #[derive(Clone, Copy)] pub struct ctg_fq { pub v: [u32; 12] }
#[repr(C)] #[derive(Clone, Copy)] pub struct ctg_g1_affine { pub x: ctg_fq, pub y: ctg_fq }
#[repr(C)] #[derive(Clone, Copy)] pub struct ctg_g2_affine { pub x_c0: ctg_fq, pub x_c1: ctg_fq, pub y_c0: ctg_fq, pub y_c1: ctg_fq }
#[repr(C)] #[derive(Clone, Copy)] pub struct ctg_g1_xyzz { pub x: ctg_fq, pub y: ctg_fq, pub zz: ctg_fq, pub zzz: ctg_fq }
#[repr(C)] #[derive(Clone, Copy)] pub struct ctg_g2_xyzz { pub c: [ctg_fq; 8] }
#[repr(C)] #[derive(Clone, Copy)] pub struct ctg_proof { pub a: ctg_g1_affine, pub b: ctg_g2_affine, pub c: ctg_g1_affine }

#[repr(C)]
pub struct ctg_pk_view {
    pub alpha_g1: *const ctg_g1_affine, pub beta_g1: *const ctg_g1_affine, pub delta_g1: *const ctg_g1_affine,
    pub beta_g2: *const ctg_g2_affine, pub delta_g2: *const ctg_g2_affine,
    pub a_query: *const ctg_g1_affine, pub a_len: u64,
    pub b_g1_query: *const ctg_g1_affine, pub b_g1_len: u64,
    pub b_g2_query: *const ctg_g2_affine, pub b_g2_len: u64,
    pub h_query: *const ctg_g1_affine, pub h_len: u64,
    pub l_query: *const ctg_g1_affine, pub l_len: u64,
}
/// trapdoor of ark_groth16::generate_parameters, drawn by the caller in arkworks' order
#[repr(C)] #[derive(Clone, Copy)]
pub struct ctg_trapdoor { pub alpha: ctg_fr, pub beta: ctg_fr, pub gamma: ctg_fr, pub delta: ctg_fr, pub tau: ctg_fr, pub g1: ctg_g1_affine, pub g2: ctg_g2_affine }
#[repr(C)]
pub struct ctg_params_out {
    pub alpha_g1: *mut ctg_g1_affine, pub beta_g1: *mut ctg_g1_affine, pub delta_g1: *mut ctg_g1_affine,
    pub beta_g2: *mut ctg_g2_affine, pub gamma_g2: *mut ctg_g2_affine, pub delta_g2: *mut ctg_g2_affine,
    pub gamma_abc_g1: *mut ctg_g1_affine, pub a_query: *mut ctg_g1_affine, pub b_g1_query: *mut ctg_g1_affine,
    pub b_g2_query: *mut ctg_g2_affine, pub h_query: *mut ctg_g1_affine, pub l_query: *mut ctg_g1_affine,
}
#[repr(C)] #[derive(Clone, Copy)] pub struct ctg_proof_partial { pub a: ctg_g1_xyzz, pub b1: ctg_g1_xyzz, pub l: ctg_g1_xyzz, pub h: ctg_g1_xyzz, pub b2: ctg_g2_xyzz }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct ctg_msm_plan { pub window_bits: i32, pub windows: i32, pub copies: i32, pub bucket_sets: i32, pub buckets_per_set: u32 }
#[repr(C)] pub struct ctg_csr { pub row_ptr: *const u64, pub col: *const u32, pub val: *const ctg_fr }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct ctg_timings {
    pub h2d_ms: f32, pub witness_map_ms: f32, pub msm_h_ms: f32, pub msm_l_ms: f32, pub msm_a_ms: f32, pub msm_b_g1_ms: f32,
    pub msm_b_g2_ms: f32, pub finalize_ms: f32, pub total_ms: f32, pub msm_sort_ms: f32, pub msm_accumulate_ms: f32, pub msm_reduce_ms: f32,
}

extern "C" {
    pub fn ctg_ctx_create(device_ids: *const c_int, n_devices: c_int, out: *mut *mut ctg_ctx) -> i32;
    pub fn ctg_ctx_device_count(ctx: *mut ctg_ctx) -> i32;
    pub fn ctg_ctx_destroy(ctx: *mut ctg_ctx);
    pub fn ctg_last_error(ctx: *mut ctg_ctx) -> *const c_char;
    pub fn ctg_ctx_set_stream(ctx: *mut ctg_ctx, cuda_stream: *mut c_void) -> i32;
    pub fn ctg_sync(ctx: *mut ctg_ctx) -> i32;
    pub fn ctg_launch_count(ctx: *mut ctg_ctx) -> u64;
    pub fn ctg_last_timings(ctx: *mut ctg_ctx, out: *mut ctg_timings) -> i32;
    /// No counterpart in the reference: ctg_field_op_batch without any allocation, staged in a fixed 3 MiB device array.
    pub fn ctg_field_op_batch_noalloc(ctx: *mut ctg_ctx, field: c_int, op: c_int, a: *const c_void, b: *const c_void, out: *mut c_void, n: usize) -> i32;
    pub fn ctg_ntt_fr(ctx: *mut ctg_ctx, host_data: *mut ctg_fr, log_n: u32, inverse: c_int, coset: c_int) -> i32;
    pub fn ctg_bases_upload_g1(ctx: *mut ctg_ctx, pts: *const ctg_g1_affine, n: u64, copies: c_int, c_bits: c_int, out: *mut *mut ctg_bases) -> i32;
    pub fn ctg_bases_upload_g2(ctx: *mut ctg_ctx, pts: *const ctg_g2_affine, n: u64, copies: c_int, c_bits: c_int, out: *mut *mut ctg_bases) -> i32;
    pub fn ctg_bases_free(ctx: *mut ctg_ctx, bases: *mut ctg_bases);
    pub fn ctg_msm_g1(ctx: *mut ctg_ctx, bases: *const ctg_bases, scalars: *const ctg_fr, n: u64, flags: u32, out: *mut ctg_g1_affine) -> i32;
    pub fn ctg_msm_g2(ctx: *mut ctg_ctx, bases: *const ctg_bases, scalars: *const ctg_fr, n: u64, flags: u32, out: *mut ctg_g2_affine) -> i32;
    pub fn ctg_pk_upload(ctx: *mut ctg_ctx, pk: *const ctg_pk_view, num_instance: u64, num_constraints: u64, copies: c_int, c_bits: c_int, out: *mut *mut ctg_pk) -> i32;
    pub fn ctg_pk_free(ctx: *mut ctg_ctx, pk: *mut ctg_pk);
    pub fn ctg_r1cs_upload(ctx: *mut ctg_ctx, num_instance: u64, num_witness: u64, num_constraints: u64,
                           a: *const ctg_csr, b: *const ctg_csr, c: *const ctg_csr, out: *mut *mut ctg_r1cs) -> i32;
    pub fn ctg_r1cs_free(ctx: *mut ctg_ctx, r1cs: *mut ctg_r1cs);
    pub fn ctg_witness_map(ctx: *mut ctg_ctx, r1cs: *const ctg_r1cs, z: *const ctg_fr, h_out: *mut ctg_fr) -> i32;
    pub fn ctg_prove(ctx: *mut ctg_ctx, pk: *const ctg_pk, r1cs: *const ctg_r1cs, z: *const ctg_fr, r: *const ctg_fr, s: *const ctg_fr,
                     out: *mut ctg_proof, timings: *mut ctg_timings) -> i32;
    pub fn ctg_prove_batch(ctx: *mut ctg_ctx, pk: *const ctg_pk, r1cs: *const ctg_r1cs, n_proofs: u32, z_batch: *const ctg_fr,
                           rs_batch: *const ctg_fr, out: *mut ctg_proof) -> i32;
    // multi-GPU: point-range key shards, partial sums, distributed witness map
    pub fn ctg_pk_upload_shard(ctx: *mut ctg_ctx, pk: *const ctg_pk_view, num_instance: u64, num_constraints: u64, copies: c_int, c_bits: c_int,
                               shard_index: c_int, shard_count: c_int, out: *mut *mut ctg_pk) -> i32;
    pub fn ctg_prove_partial(ctx: *mut ctg_ctx, pk: *const ctg_pk, r1cs: *const ctg_r1cs, z: *const ctg_fr, r: *const ctg_fr, s: *const ctg_fr,
                             out: *mut ctg_proof_partial, timings: *mut ctg_timings) -> i32;
    pub fn ctg_prove_partial_begin(ctx: *mut ctg_ctx, pk: *const ctg_pk, r1cs: *const ctg_r1cs, z: *const ctg_fr, r: *const ctg_fr, s: *const ctg_fr,
                                   chain_mask: u32, d_abc: *mut c_void) -> i32;
    pub fn ctg_prove_partial_finish(ctx: *mut ctg_ctx, pk: *const ctg_pk, r1cs: *const ctg_r1cs, r: *const ctg_fr, s: *const ctg_fr,
                                    d_abc: *mut c_void, out: *mut ctg_proof_partial) -> i32;
    pub fn ctg_prove_combine(ctx: *mut ctg_ctx, partials: *const ctg_proof_partial, k: u32, r: *const ctg_fr, s: *const ctg_fr, out: *mut ctg_proof) -> i32;
    // verifier-side aggregation, key generation, wire formats
    pub fn ctg_prepare_inputs(ctx: *mut ctg_ctx, gamma_abc_g1: *const ctg_bases, public_inputs: *const ctg_fr, n_inputs: u64, out: *mut ctg_g1_affine) -> i32;
    pub fn ctg_generate_parameters(ctx: *mut ctg_ctx, num_instance: u64, num_witness: u64, num_constraints: u64, a: *const ctg_csr, b: *const ctg_csr,
                                   c: *const ctg_csr, trapdoor: *const ctg_trapdoor, out: *mut *mut ctg_params) -> i32;
    pub fn ctg_params_pk_view(ctx: *mut ctg_ctx, params: *const ctg_params, out: *mut ctg_pk_view, gamma_abc_g1: *mut *const ctg_g1_affine,
                              gamma_abc_len: *mut u64) -> i32;
    pub fn ctg_params_download(ctx: *mut ctg_ctx, params: *const ctg_params, dst: *const ctg_params_out) -> i32;
    pub fn ctg_params_free(ctx: *mut ctg_ctx, params: *mut ctg_params);
    pub fn ctg_params_deserialize_unchecked(ctx: *mut ctg_ctx, key_image: *const u8, len: u64, out: *mut *mut ctg_params) -> i32;
    pub fn ctg_params_serialize_unchecked(ctx: *mut ctg_ctx, params: *const ctg_params, out: *mut u8, capacity: u64, written: *mut u64) -> i32;
    pub fn ctg_params_serialize_vk(ctx: *mut ctg_ctx, params: *const ctg_params, out: *mut u8, capacity: u64, written: *mut u64) -> i32;
    pub fn ctg_proof_serialize(proof: *const ctg_proof, out192: *mut u8) -> i32;
    pub fn ctg_proof_deserialize(in192: *const u8, check_subgroup: c_int, out: *mut ctg_proof) -> i32;
    pub fn ctg_host_alloc(ctx: *mut ctg_ctx, bytes: usize, h_out: *mut *mut c_void) -> i32;
    pub fn ctg_host_free(ctx: *mut ctg_ctx, h_ptr: *mut c_void) -> i32;
    pub fn ctg_pk_upload_range(ctx: *mut ctg_ctx, pk: *const ctg_pk_view, num_instance: u64, num_constraints: u64, copies: c_int, c_bits: c_int,
                               lo_num: u64, hi_num: u64, den: u64, out: *mut *mut ctg_pk) -> i32;
    // bulk Poseidon / ciphertext Merkle tree (circuits-ark/src/sample_entries.rs:127-134, encryption.rs:154-172)
    pub fn ctg_poseidon_new(ctx: *mut ctg_ctx, full_rounds: u32, partial_rounds: u32, alpha: u32, mds: *const ctg_fr, ark: *const ctg_fr,
                            out: *mut *mut ctg_poseidon) -> i32;
    pub fn ctg_poseidon_free(ctx: *mut ctg_ctx, p: *mut ctg_poseidon);
    pub fn ctg_poseidon_hash_batch(ctx: *mut ctg_ctx, p: *const ctg_poseidon, input: *const ctg_fr, n: u64, k: u32, out: *mut ctg_fr) -> i32;
    pub fn ctg_poseidon_hash_batch_dev(ctx: *mut ctg_ctx, p: *const ctg_poseidon, d_in: *const c_void, n: u64, k: u32, d_out: *mut c_void) -> i32;
    pub fn ctg_poseidon_merkle_tree(ctx: *mut ctg_ctx, p: *const ctg_poseidon, leaves: *const ctg_fr, n: u64, leaf_digests: *mut ctg_fr,
                                    inner: *mut ctg_fr) -> i32;
    pub fn ctg_poseidon_merkle_tree_dev(ctx: *mut ctg_ctx, p: *const ctg_poseidon, d_leaves: *const c_void, n: u64, d_leaf_digests: *mut c_void,
                                        d_inner: *mut c_void) -> i32;
    // bulk JubJub scalar multiplication: the group part of encrypt() / keygen() (circuits-ark/src/encryption.rs:126-137,160-165)
    pub fn ctg_te_mul_batch(ctx: *mut ctg_ctx, d: *const ctg_fr, points: *const ctg_fr, n_points: u64, scalars: *const u32, n: u64, out: *mut ctg_fr) -> i32;
    pub fn ctg_te_mul_batch_dev(ctx: *mut ctg_ctx, d: *const ctg_fr, d_points: *const c_void, n_points: u64, d_scalars: *const c_void, n: u64,
                                d_out: *mut c_void) -> i32;
    // primitives, device-pointer variants, tuning knobs and harness helpers (tests / benches / ncu captures)
    pub fn ctg_ctx_set_ba_levels(ctx: *mut ctg_ctx, levels: c_int) -> i32;
    pub fn ctg_ctx_set_msm_split(ctx: *mut ctg_ctx, parts: c_int) -> i32;
    pub fn ctg_dev_alloc(ctx: *mut ctg_ctx, bytes: usize, d_out: *mut *mut c_void) -> i32;
    pub fn ctg_dev_free(ctx: *mut ctg_ctx, d_ptr: *mut c_void) -> i32;
    pub fn ctg_h2d(ctx: *mut ctg_ctx, d_dst: *mut c_void, h_src: *const c_void, bytes: usize) -> i32;
    pub fn ctg_d2h(ctx: *mut ctg_ctx, h_dst: *mut c_void, d_src: *const c_void, bytes: usize) -> i32;
    pub fn ctg_field_op_batch(ctx: *mut ctg_ctx, field: c_int, op: c_int, a: *const c_void, b: *const c_void, out: *mut c_void, n: usize) -> i32;
    pub fn ctg_fr_mul_batch(ctx: *mut ctg_ctx, a: *const ctg_fr, b: *const ctg_fr, out: *mut ctg_fr, n: usize) -> i32;
    pub fn ctg_fq_mul_batch(ctx: *mut ctg_ctx, a: *const ctg_fq, b: *const ctg_fq, out: *mut ctg_fq, n: usize) -> i32;
    pub fn ctg_imad_peak_probe(ctx: *mut ctg_ctx, imad_per_s: *mut f64, fq_mul_per_s: *mut f64) -> i32;
    pub fn ctg_ntt_fr_dev(ctx: *mut ctg_ctx, d_data: *mut c_void, log_n: u32, inverse: c_int, coset: c_int) -> i32;
    pub fn ctg_bases_progression_g1(ctx: *mut ctg_ctx, g: *const ctg_g1_affine, a: *const ctg_fr, b: *const ctg_fr, n: u64, copies: c_int, c_bits: c_int, out: *mut *mut ctg_bases) -> i32;
    pub fn ctg_bases_progression_g2(ctx: *mut ctg_ctx, g: *const ctg_g2_affine, a: *const ctg_fr, b: *const ctg_fr, n: u64, copies: c_int, c_bits: c_int, out: *mut *mut ctg_bases) -> i32;
    pub fn ctg_bases_download(ctx: *mut ctg_ctx, bases: *const ctg_bases, first: u64, count: u64, host_points: *mut c_void) -> i32;
    pub fn ctg_bases_len(bases: *const ctg_bases) -> u64;
    pub fn ctg_bases_plan(bases: *const ctg_bases, out: *mut ctg_msm_plan) -> i32;
    pub fn ctg_msm_g1_dev(ctx: *mut ctg_ctx, bases: *const ctg_bases, d_scalars: *const c_void, n: u64, flags: u32, d_out_xyzz: *mut c_void) -> i32;
    pub fn ctg_msm_g2_dev(ctx: *mut ctg_ctx, bases: *const ctg_bases, d_scalars: *const c_void, n: u64, flags: u32, d_out_xyzz: *mut c_void) -> i32;
    pub fn ctg_msm_g1_partial(ctx: *mut ctg_ctx, bases: *const ctg_bases, scalars: *const ctg_fr, n: u64, flags: u32, out: *mut ctg_g1_xyzz) -> i32;
    pub fn ctg_msm_g2_partial(ctx: *mut ctg_ctx, bases: *const ctg_bases, scalars: *const ctg_fr, n: u64, flags: u32, out: *mut ctg_g2_xyzz) -> i32;
    pub fn ctg_g1_combine(ctx: *mut ctg_ctx, partials: *const ctg_g1_xyzz, k: u32, out: *mut ctg_g1_affine) -> i32;
    pub fn ctg_g2_combine(ctx: *mut ctg_ctx, partials: *const ctg_g2_xyzz, k: u32, out: *mut ctg_g2_affine) -> i32;
    pub fn ctg_g1_combine_dev(ctx: *mut ctg_ctx, d_partials: *const c_void, k: u32, d_out_affine: *mut c_void) -> i32;
    pub fn ctg_g2_combine_dev(ctx: *mut ctg_ctx, d_partials: *const c_void, k: u32, d_out_affine: *mut c_void) -> i32;
}
