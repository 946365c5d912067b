//! `ctg-groth16`: the ark-groth16 prove API backed by libctg_groth16.so (B200, sm_100a).
//! SOURCE ONLY -- never compiled on the build box (no Rust toolchain there); see ../README.md.
//!
//! Replaces, for `E = Bls12_381` only:
//!   ark_groth16::create_random_proof(circuit, &pk, rng)      -> ctg_groth16::create_random_proof(circuit, &dpk, rng)
//!   ark_groth16::create_proof(circuit, &pk, r, s)            -> ctg_groth16::create_proof(circuit, &dpk, r, s)
//!   <Groth16<Bls12_381> as SNARK<Fr>>::prove(&pk, c, rng)    -> <Groth16Gpu as SNARK<Fr>>::prove(&dpk, c, rng)
//! Synthesis runs on the CPU exactly as in ark-groth16 0.3.0 prover.rs (OptimizationGoal::Constraints,
//! finalize, to_matrices); r then s are drawn with Fr::rand; only the arithmetic after synthesis is offloaded.
mod pack;
pub mod witness;

use ark_bls12_381::{Bls12_381, Fr};
use ark_groth16::{Proof, ProvingKey, VerifyingKey};
use ark_relations::r1cs::{ConstraintSynthesizer, ConstraintSystem, OptimizationGoal, SynthesisError};
use ark_std::{rand::{CryptoRng, RngCore}, UniformRand};
use ctg_groth16_sys as sys;
use std::{ptr, sync::Mutex};

/// One CUDA device.  Calls are serialized per context (the C ABI allows one in-flight call per context).
pub struct Context { raw: Mutex<*mut sys::ctg_ctx> }
unsafe impl Send for Context {}
unsafe impl Sync for Context {}

impl Context {
    pub fn new(device: i32) -> Result<Self, SynthesisError> {
        let mut raw = ptr::null_mut();
        let rc = unsafe { sys::ctg_ctx_create(&device, 1, &mut raw) };
        if rc != sys::CTG_OK { return Err(SynthesisError::Unsatisfiable) } // no device: there is no CPU fallback
        Ok(Self { raw: Mutex::new(raw) })
    }
}
impl Drop for Context { fn drop(&mut self) { unsafe { sys::ctg_ctx_destroy(*self.raw.lock().unwrap()) } } }

/// Device-resident replacement of `Option<ProvingKey<Bls12_381>>` (reference: lib/src/zk/encryption.rs:22,31-32).
pub struct DeviceProvingKey<'c> {
    ctx: &'c Context,
    pk: *mut sys::ctg_pk,
    /// R1CS matrices are per-circuit constants: uploaded on the first proof, then reused.
    r1cs: Mutex<Option<(*mut sys::ctg_r1cs, usize, usize, usize)>>,
    pub vk: VerifyingKey<Bls12_381>,
    num_instance: usize,
    num_constraints: usize,
}
unsafe impl<'c> Send for DeviceProvingKey<'c> {}
unsafe impl<'c> Sync for DeviceProvingKey<'c> {}

impl<'c> DeviceProvingKey<'c> {
    /// `num_instance` counts the constant 1; both counts come from one Setup-mode synthesis of the circuit
    /// (or from the ConstraintSystem used by `Groth16::setup`).
    pub fn upload(ctx: &'c Context, pk: &ProvingKey<Bls12_381>, num_instance: usize, num_constraints: usize) -> Result<Self, SynthesisError> {
        let packed = pack::PackedPk::from(pk); // GroupAffine is not repr(C): repack to x || y, infinity = all-zero
        let view = packed.view();
        let mut raw = ptr::null_mut();
        let g = ctx.raw.lock().unwrap();
        let rc = unsafe { sys::ctg_pk_upload(*g, &view, num_instance as u64, num_constraints as u64, 0, 0, &mut raw) };
        if rc != sys::CTG_OK { return Err(SynthesisError::MalformedVerifyingKey) }
        Ok(Self { ctx, pk: raw, r1cs: Mutex::new(None), vk: pk.vk.clone(), num_instance, num_constraints })
    }
}
impl<'c> DeviceProvingKey<'c> {
    /// `read_proving_key` (reference: lib/src/utils.rs:104-110) without the host-side ProvingKey: the circuit.zkey image
    /// (ProvingKey::serialize_unchecked) is decoded on the device.  `vk` comes from verification.key as before.
    pub fn from_zkey_bytes(ctx: &'c Context, image: &[u8], vk: VerifyingKey<Bls12_381>, num_constraints: usize) -> Result<Self, SynthesisError> {
        let g = ctx.raw.lock().unwrap();
        let (mut params, mut raw) = (ptr::null_mut(), ptr::null_mut());
        if unsafe { sys::ctg_params_deserialize_unchecked(*g, image.as_ptr(), image.len() as u64, &mut params) } != sys::CTG_OK {
            return Err(SynthesisError::MalformedVerifyingKey)
        }
        let num_instance = vk.gamma_abc_g1.len();
        let mut view = std::mem::MaybeUninit::<sys::ctg_pk_view>::uninit();
        let rc = unsafe {
            sys::ctg_params_pk_view(*g, params, view.as_mut_ptr(), ptr::null_mut(), ptr::null_mut());
            let rc = sys::ctg_pk_upload(*g, view.as_ptr(), num_instance as u64, num_constraints as u64, 0, 0, &mut raw); // device to device
            sys::ctg_params_free(*g, params);
            rc
        };
        if rc != sys::CTG_OK { return Err(SynthesisError::MalformedVerifyingKey) }
        Ok(Self { ctx, pk: raw, r1cs: Mutex::new(None), vk, num_instance, num_constraints })
    }
}

/// ark_groth16::generate_random_parameters with the arithmetic on the device (ctg_generate_parameters).  The trapdoor is
/// drawn here in ark-groth16 0.3.0's order so the RNG stream matches generator.rs: alpha, beta, gamma, delta,
/// G1Projective::rand, G2Projective::rand, then tau = domain.sample_element_outside_domain(rng) after Setup-mode synthesis.
/// Returns the host ProvingKey (downloaded once, for `write_circuit_artifacts`) -- or keep the handle and call
/// `ctg_params_serialize_unchecked` to write circuit.zkey without materialising it on the host.
pub fn generate_random_parameters<C, R>(ctx: &Context, circuit: C, rng: &mut R) -> Result<ProvingKey<Bls12_381>, SynthesisError>
where C: ConstraintSynthesizer<Fr>, R: RngCore + CryptoRng {
    use ark_bls12_381::{G1Projective, G2Projective};
    use ark_ec::ProjectiveCurve;
    use ark_poly::{EvaluationDomain, GeneralEvaluationDomain};
    let (alpha, beta, gamma, delta) = (Fr::rand(rng), Fr::rand(rng), Fr::rand(rng), Fr::rand(rng));
    let (g1, g2) = (G1Projective::rand(rng).into_affine(), G2Projective::rand(rng).into_affine());
    let cs = ConstraintSystem::new_ref();
    cs.set_optimization_goal(OptimizationGoal::Constraints);
    cs.set_mode(ark_relations::r1cs::SynthesisMode::Setup);
    circuit.generate_constraints(cs.clone())?;
    cs.finalize();
    let (ni, nw, nc) = (cs.num_instance_variables(), cs.num_witness_variables(), cs.num_constraints());
    let domain = GeneralEvaluationDomain::<Fr>::new(nc + ni).ok_or(SynthesisError::PolynomialDegreeTooLarge)?;
    let tau = domain.sample_element_outside_domain(rng);
    let m = cs.to_matrices().ok_or(SynthesisError::AssignmentMissing)?;
    let (a, b, c) = (pack::Csr::from_rows(&m.a), pack::Csr::from_rows(&m.b), pack::Csr::from_rows(&m.c));
    let td = sys::ctg_trapdoor { alpha: pack::fr_image(&alpha), beta: pack::fr_image(&beta), gamma: pack::fr_image(&gamma),
                                 delta: pack::fr_image(&delta), tau: pack::fr_image(&tau), g1: pack::g1_image(&g1), g2: pack::g2_image(&g2) };
    let g = ctx.raw.lock().unwrap();
    let mut params = ptr::null_mut();
    if unsafe { sys::ctg_generate_parameters(*g, ni as u64, nw as u64, nc as u64, &a.view(), &b.view(), &c.view(), &td, &mut params) } != sys::CTG_OK {
        return Err(SynthesisError::Unsatisfiable)
    }
    let pk = pack::download_pk(*g, params, ni, nw, domain.size());   // ctg_params_download into Vec<G1Affine> / Vec<G2Affine>
    unsafe { sys::ctg_params_free(*g, params) };
    pk
}

impl<'c> Drop for DeviceProvingKey<'c> {
    fn drop(&mut self) {
        let g = self.ctx.raw.lock().unwrap();
        unsafe {
            if let Some((r, ..)) = self.r1cs.lock().unwrap().take() { sys::ctg_r1cs_free(*g, r) }
            sys::ctg_pk_free(*g, self.pk)
        }
    }
}

/// ark_groth16::create_random_proof: r first, then s.
pub fn create_random_proof<C, R>(circuit: C, pk: &DeviceProvingKey<'_>, rng: &mut R) -> Result<Proof<Bls12_381>, SynthesisError>
where C: ConstraintSynthesizer<Fr>, R: RngCore + CryptoRng {
    let r = Fr::rand(rng);
    let s = Fr::rand(rng);
    create_proof(circuit, pk, r, s)
}

/// ark_groth16::create_proof after the same CPU synthesis arkworks performs.
pub fn create_proof<C: ConstraintSynthesizer<Fr>>(circuit: C, pk: &DeviceProvingKey<'_>, r: Fr, s: Fr) -> Result<Proof<Bls12_381>, SynthesisError> {
    let cs = ConstraintSystem::new_ref();
    cs.set_optimization_goal(OptimizationGoal::Constraints);
    circuit.generate_constraints(cs.clone())?;
    cs.finalize();
    let (ni, nw, nc) = (cs.num_instance_variables(), cs.num_witness_variables(), cs.num_constraints());
    if ni != pk.num_instance || nc != pk.num_constraints { return Err(SynthesisError::MalformedVerifyingKey) }
    let g = pk.ctx.raw.lock().unwrap();
    let mut slot = pk.r1cs.lock().unwrap();
    if slot.is_none() {
        let m = cs.to_matrices().ok_or(SynthesisError::AssignmentMissing)?;
        let (a, b, c) = (pack::Csr::from_rows(&m.a), pack::Csr::from_rows(&m.b), pack::Csr::from_rows(&m.c));
        let mut raw = ptr::null_mut();
        let rc = unsafe { sys::ctg_r1cs_upload(*g, ni as u64, nw as u64, nc as u64, &a.view(), &b.view(), &c.view(), &mut raw) };
        if rc != sys::CTG_OK { return Err(SynthesisError::Unsatisfiable) }
        *slot = Some((raw, ni, nw, nc));
    }
    let (r1cs, ..) = slot.unwrap();
    let inner = cs.borrow().ok_or(SynthesisError::MissingCS)?;
    // z = instance ++ witness; Fp256<..> is #[repr(transparent)]-like over BigInteger256([u64; 4]): Montgomery image
    let mut z: Vec<sys::ctg_fr> = Vec::with_capacity(ni + nw);
    z.extend(inner.instance_assignment.iter().map(pack::fr_image));
    z.extend(inner.witness_assignment.iter().map(pack::fr_image));
    let (rr, ss) = (pack::fr_image(&r), pack::fr_image(&s));
    let mut out = std::mem::MaybeUninit::<sys::ctg_proof>::uninit();
    let rc = unsafe { sys::ctg_prove(*g, pk.pk, r1cs, z.as_ptr(), &rr, &ss, out.as_mut_ptr(), ptr::null_mut()) };
    if rc != sys::CTG_OK { return Err(SynthesisError::Unsatisfiable) }
    Ok(pack::proof_from(unsafe { &out.assume_init() }))
}

/// `SNARK` front: the reference's call sites change `Groth16::<Bls12_381>` to `Groth16Gpu`.
pub struct Groth16Gpu;
impl Groth16Gpu {
    pub fn prove<C: ConstraintSynthesizer<Fr>, R: RngCore + CryptoRng>(pk: &DeviceProvingKey<'_>, circuit: C, rng: &mut R)
        -> Result<Proof<Bls12_381>, SynthesisError> { create_random_proof(circuit, pk, rng) }
    /// Verification and setup are unchanged: delegate to ark_groth16 with `pk.vk`.
    pub fn verify(vk: &VerifyingKey<Bls12_381>, public: &[Fr], proof: &Proof<Bls12_381>) -> Result<bool, SynthesisError> {
        ark_groth16::verify_proof(&ark_groth16::prepare_verifying_key(vk), proof, public)
    }
}
