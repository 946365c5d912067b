//! `ctg-groth16`: the ark-groth16 prove API backed by libctg_groth16.so (B200, sm_100a).
//! SOURCE ONLY -- never compiled on the build box (no Rust toolchain there); see ../README.md.
//!
//! Replaces, for `E = Bls12_381` only:
//!   ark_groth16::create_random_proof(circuit, &pk, rng)      -> ctg_groth16::create_random_proof(circuit, &dpk, rng)
//!   ark_groth16::create_proof(circuit, &pk, r, s)            -> ctg_groth16::create_proof(circuit, &dpk, r, s)
//!   <Groth16<Bls12_381> as SNARK<Fr>>::prove(&pk, c, rng)    -> <Groth16Gpu as SNARK<Fr>>::prove(&dpk, c, rng)
//! Synthesis runs on the CPU exactly as in ark-groth16 0.3.0 prover.rs (OptimizationGoal::Constraints,
//! finalize, to_matrices); r then s are drawn with Fr::rand; only the arithmetic after synthesis is offloaded.
//!
//! One process may drive several GPUs: `Context::with_devices(&[0, 1, .., 7])` makes a multi-device context
//! (ctg_ctx_create with n_devices > 1); every call below is unchanged on it -- the key is sharded by point range at upload, a
//! proof's MSMs and NTT chains are spread over the devices and exchanged over NVLink inside the library.
mod pack;
pub mod witness;

use ark_bls12_381::{Bls12_381, Fr};
use ark_groth16::{PreparedVerifyingKey, Proof, ProvingKey, VerifyingKey};
use ark_relations::r1cs::{ConstraintSynthesizer, ConstraintSystem, ConstraintSystemRef, OptimizationGoal, SynthesisError};
use ark_snark::SNARK;
use ark_std::{rand::{CryptoRng, RngCore}, UniformRand};
use ctg_groth16_sys as sys;
use std::{ffi::CStr, fmt, ptr, sync::{Arc, Mutex, OnceLock}};

/// What a call into the library can report.  `SNARK::Error` of `Groth16Gpu`; every C-ABI status keeps its identity and the
/// library's message (`ctg_last_error`) instead of collapsing into one `SynthesisError` variant.
#[derive(Debug)]
pub enum Error {
    /// Raised by the circuit's own `generate_constraints`, or a size mismatch between circuit and key.
    Synthesis(SynthesisError),
    /// CTG_ERR_INVALID: an argument the library rejected (sizes, a malformed CSR, a key of another context, ...).
    InvalidArgument(String),
    /// CTG_ERR_CUDA: a CUDA call or kernel launch failed; there is no CPU fallback.
    Device(String),
    /// CTG_ERR_OOM: device or pinned host memory exhausted.
    OutOfMemory(String),
    /// CTG_ERR_STATE: the call does not fit the handle's state (a shard handed to ctg_prove, a closed context, ...).
    State(String),
    Other(Box<dyn std::error::Error>),
}
impl fmt::Display for Error {
    fn fmt(&self, f: &mut fmt::Formatter<'_>) -> fmt::Result {
        match self {
            Error::Synthesis(e) => write!(f, "synthesis: {e}"),
            Error::InvalidArgument(m) => write!(f, "ctg: invalid argument: {m}"),
            Error::Device(m) => write!(f, "ctg: CUDA failure: {m}"),
            Error::OutOfMemory(m) => write!(f, "ctg: out of memory: {m}"),
            Error::State(m) => write!(f, "ctg: wrong state: {m}"),
            Error::Other(e) => write!(f, "{e}"),
        }
    }
}
impl std::error::Error for Error {}
impl From<SynthesisError> for Error { fn from(e: SynthesisError) -> Self { Error::Synthesis(e) } }
impl From<Box<dyn std::error::Error>> for Error { fn from(e: Box<dyn std::error::Error>) -> Self { Error::Other(e) } }

/// Status code -> `Error`, with the library's message for this context.
fn check(ctx: *mut sys::ctg_ctx, rc: i32) -> Result<(), Error> {
    if rc == sys::CTG_OK { return Ok(()) }
    let msg = if ctx.is_null() { String::new() } else {
        let p = unsafe { sys::ctg_last_error(ctx) };
        if p.is_null() { String::new() } else { unsafe { CStr::from_ptr(p) }.to_string_lossy().into_owned() }
    };
    Err(match rc {
        sys::CTG_ERR_INVALID => Error::InvalidArgument(msg),
        sys::CTG_ERR_CUDA => Error::Device(msg),
        sys::CTG_ERR_OOM => Error::OutOfMemory(msg),
        _ => Error::State(msg),
    })
}

/// One CUDA device, or several GPUs of one NVLink box.  Calls are serialized per context (the C ABI allows one in-flight
/// call per context).
pub struct Context { raw: Mutex<*mut sys::ctg_ctx>, devices: Vec<i32> }
unsafe impl Send for Context {}
unsafe impl Sync for Context {}

impl Context {
    pub fn new(device: i32) -> Result<Arc<Self>, Error> { Self::with_devices(&[device]) }
    /// `devices[0]` is the primary device; with more than one device the prover-path calls shard by themselves.
    pub fn with_devices(devices: &[i32]) -> Result<Arc<Self>, Error> {
        let mut raw = ptr::null_mut();
        let rc = unsafe { sys::ctg_ctx_create(devices.as_ptr(), devices.len() as i32, &mut raw) };
        if rc != sys::CTG_OK { return Err(Error::Device(format!("ctg_ctx_create({devices:?}) failed with {rc}: no usable sm_100 device (there is no CPU fallback)"))) }
        Ok(Arc::new(Self { raw: Mutex::new(raw), devices: devices.to_vec() }))
    }
    /// The context `SNARK::circuit_specific_setup` uses (its signature has no room for one): device `CTG_DEVICE` or 0.
    pub fn global() -> Result<Arc<Self>, Error> {
        static G: OnceLock<Arc<Context>> = OnceLock::new();
        if let Some(c) = G.get() { return Ok(c.clone()) }
        let dev = std::env::var("CTG_DEVICE").ok().and_then(|v| v.parse().ok()).unwrap_or(0);
        let c = Self::new(dev)?;
        Ok(G.get_or_init(|| c).clone())
    }
    pub fn devices(&self) -> &[i32] { &self.devices }
}
impl Drop for Context { fn drop(&mut self) { unsafe { sys::ctg_ctx_destroy(*self.raw.lock().unwrap()) } } }

/// Page-locked staging for the assignment (ctg_host_alloc): host -> device copies from it run asynchronously at full PCIe rate.
struct PinnedZ { ptr: *mut sys::ctg_fr, cap: usize }
impl PinnedZ {
    fn ensure(&mut self, ctx: *mut sys::ctg_ctx, elems: usize) -> Result<(), Error> {
        if elems <= self.cap { return Ok(()) }
        if !self.ptr.is_null() { unsafe { sys::ctg_host_free(ctx, self.ptr as *mut _) }; self.ptr = ptr::null_mut(); self.cap = 0; }
        let mut p = ptr::null_mut();
        check(ctx, unsafe { sys::ctg_host_alloc(ctx, elems * std::mem::size_of::<sys::ctg_fr>(), &mut p) })?;
        self.ptr = p as *mut sys::ctg_fr;
        self.cap = elems;
        Ok(())
    }
}

struct KeyInner {
    ctx: Arc<Context>,
    pk: *mut sys::ctg_pk,
    /// R1CS matrices are per-circuit constants: uploaded on the first proof, then reused.
    r1cs: Mutex<Option<(*mut sys::ctg_r1cs, usize, usize, usize)>>,
    staging: Mutex<PinnedZ>,
    vk: VerifyingKey<Bls12_381>,
    num_instance: usize,
    num_constraints: usize,
}
unsafe impl Send for KeyInner {}
unsafe impl Sync for KeyInner {}

/// Device-resident replacement of `Option<ProvingKey<Bls12_381>>` (reference: lib/src/zk/encryption.rs:22,31-32).  Cheap to
/// clone (`SNARK::ProvingKey: Clone`): the clones share one key cache in HBM.
#[derive(Clone)]
pub struct DeviceProvingKey(Arc<KeyInner>);

impl DeviceProvingKey {
    /// `num_instance` counts the constant 1; both counts come from one Setup-mode synthesis of the circuit
    /// (or from the ConstraintSystem used by `Groth16::setup`).
    pub fn upload(ctx: &Arc<Context>, pk: &ProvingKey<Bls12_381>, num_instance: usize, num_constraints: usize) -> Result<Self, Error> {
        let packed = pack::PackedPk::from(pk); // GroupAffine is not repr(C): repack to x || y, infinity = all-zero
        let view = packed.view();
        let mut raw = ptr::null_mut();
        let g = ctx.raw.lock().unwrap();
        check(*g, unsafe { sys::ctg_pk_upload(*g, &view, num_instance as u64, num_constraints as u64, 0, 0, &mut raw) })?;
        Ok(Self::wrap(ctx, raw, pk.vk.clone(), num_instance, num_constraints))
    }
    fn wrap(ctx: &Arc<Context>, pk: *mut sys::ctg_pk, vk: VerifyingKey<Bls12_381>, num_instance: usize, num_constraints: usize) -> Self {
        DeviceProvingKey(Arc::new(KeyInner { ctx: ctx.clone(), pk, r1cs: Mutex::new(None), staging: Mutex::new(PinnedZ { ptr: ptr::null_mut(), cap: 0 }),
                                             vk, num_instance, num_constraints }))
    }
    /// `read_proving_key` (reference: lib/src/utils.rs:104-110) without the host-side ProvingKey: the circuit.zkey image
    /// (ProvingKey::serialize_unchecked) is decoded on the device.  `vk` comes from verification.key as before.
    pub fn from_zkey_bytes(ctx: &Arc<Context>, image: &[u8], vk: VerifyingKey<Bls12_381>, num_constraints: usize) -> Result<Self, Error> {
        let g = ctx.raw.lock().unwrap();
        let (mut params, mut raw) = (ptr::null_mut(), ptr::null_mut());
        check(*g, unsafe { sys::ctg_params_deserialize_unchecked(*g, image.as_ptr(), image.len() as u64, &mut params) })?;
        let num_instance = vk.gamma_abc_g1.len();
        let mut view = std::mem::MaybeUninit::<sys::ctg_pk_view>::uninit();
        let rc = unsafe {
            sys::ctg_params_pk_view(*g, params, view.as_mut_ptr(), ptr::null_mut(), ptr::null_mut());
            let rc = sys::ctg_pk_upload(*g, view.as_ptr(), num_instance as u64, num_constraints as u64, 0, 0, &mut raw); // device to device
            sys::ctg_params_free(*g, params);
            rc
        };
        check(*g, rc)?;
        Ok(Self::wrap(ctx, raw, vk, num_instance, num_constraints))
    }
    pub fn vk(&self) -> &VerifyingKey<Bls12_381> { &self.0.vk }
    pub fn context(&self) -> &Arc<Context> { &self.0.ctx }
}

impl Drop for KeyInner {
    fn drop(&mut self) {
        let g = self.ctx.raw.lock().unwrap();
        unsafe {
            if let Some((r, ..)) = self.r1cs.lock().unwrap().take() { sys::ctg_r1cs_free(*g, r) }
            let st = self.staging.lock().unwrap();
            if !st.ptr.is_null() { sys::ctg_host_free(*g, st.ptr as *mut _); }
            sys::ctg_pk_free(*g, self.pk)
        }
    }
}

/// ark_groth16::generate_random_parameters with the arithmetic on the device (ctg_generate_parameters).  The trapdoor is
/// drawn here in ark-groth16 0.3.0's order so the RNG stream matches generator.rs: alpha, beta, gamma, delta,
/// G1Projective::rand, G2Projective::rand, then tau = domain.sample_element_outside_domain(rng) after Setup-mode synthesis.
/// Returns the host ProvingKey (downloaded once, for `write_circuit_artifacts`) -- or keep the handle and call
/// `ctg_params_serialize_unchecked` to write circuit.zkey without materialising it on the host.
pub fn generate_random_parameters<C, R>(ctx: &Arc<Context>, circuit: C, rng: &mut R) -> Result<ProvingKey<Bls12_381>, Error>
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
    check(*g, unsafe { sys::ctg_generate_parameters(*g, ni as u64, nw as u64, nc as u64, &a.view(), &b.view(), &c.view(), &td, &mut params) })?;
    let pk = pack::download_pk(*g, params, ni, nw, domain.size());   // ctg_params_download into Vec<G1Affine> / Vec<G2Affine>
    unsafe { sys::ctg_params_free(*g, params) };
    Ok(pk?)
}


/// CPU synthesis exactly as ark-groth16 0.3.0 prover.rs performs it.
fn synthesize<C: ConstraintSynthesizer<Fr>>(circuit: C) -> Result<ConstraintSystemRef<Fr>, Error> {
    let cs = ConstraintSystem::new_ref();
    cs.set_optimization_goal(OptimizationGoal::Constraints);
    circuit.generate_constraints(cs.clone())?;
    cs.finalize();
    Ok(cs)
}

impl DeviceProvingKey {
    /// The circuit's matrices, uploaded once per key (ctg_r1cs_upload caches them in HBM).
    fn r1cs_for(&self, g: *mut sys::ctg_ctx, cs: &ConstraintSystemRef<Fr>) -> Result<*mut sys::ctg_r1cs, Error> {
        let (ni, nw, nc) = (cs.num_instance_variables(), cs.num_witness_variables(), cs.num_constraints());
        // the key fixes the instance size and the evaluation domain (h_query has domain - 1 elements); any constraint count that maps to
        // the same radix-2 domain is the same key shape
        let domain = |c: usize| (c + ni).next_power_of_two();
        if ni != self.0.num_instance || domain(nc) != domain(self.0.num_constraints) { return Err(Error::Synthesis(SynthesisError::MalformedVerifyingKey)) }
        let mut slot = self.0.r1cs.lock().unwrap();
        if let Some((raw, _, w, _)) = *slot {
            if w != nw { return Err(Error::Synthesis(SynthesisError::MalformedVerifyingKey)) }
            return Ok(raw)
        }
        let m = cs.to_matrices().ok_or(SynthesisError::AssignmentMissing)?;
        let (a, b, c) = (pack::Csr::from_rows(&m.a), pack::Csr::from_rows(&m.b), pack::Csr::from_rows(&m.c));
        let mut raw = ptr::null_mut();
        check(g, unsafe { sys::ctg_r1cs_upload(g, ni as u64, nw as u64, nc as u64, &a.view(), &b.view(), &c.view(), &mut raw) })?;
        *slot = Some((raw, ni, nw, nc));
        Ok(raw)
    }
}

/// ark_groth16::create_random_proof: r first, then s.
pub fn create_random_proof<C, R>(circuit: C, pk: &DeviceProvingKey, rng: &mut R) -> Result<Proof<Bls12_381>, Error>
where C: ConstraintSynthesizer<Fr>, R: RngCore + CryptoRng {
    let r = Fr::rand(rng);
    let s = Fr::rand(rng);
    create_proof(circuit, pk, r, s)
}

/// ark_groth16::create_proof after the same CPU synthesis arkworks performs.
pub fn create_proof<C: ConstraintSynthesizer<Fr>>(circuit: C, pk: &DeviceProvingKey, r: Fr, s: Fr) -> Result<Proof<Bls12_381>, Error> {
    let cs = synthesize(circuit)?;
    let g = pk.0.ctx.raw.lock().unwrap();
    let r1cs = pk.r1cs_for(*g, &cs)?;
    let inner = cs.borrow().ok_or(SynthesisError::MissingCS)?;
    // z = instance ++ witness, packed straight into page-locked memory; Fp256 is a BigInteger256([u64; 4]) Montgomery image
    let m = inner.instance_assignment.len() + inner.witness_assignment.len();
    let mut st = pk.0.staging.lock().unwrap();
    st.ensure(*g, m)?;
    let z = unsafe { std::slice::from_raw_parts_mut(st.ptr, m) };
    for (dst, x) in z.iter_mut().zip(inner.instance_assignment.iter().chain(inner.witness_assignment.iter())) { *dst = pack::fr_image(x) }
    let (rr, ss) = (pack::fr_image(&r), pack::fr_image(&s));
    let mut out = std::mem::MaybeUninit::<sys::ctg_proof>::uninit();
    check(*g, unsafe { sys::ctg_prove(*g, pk.0.pk, r1cs, st.ptr, &rr, &ss, out.as_mut_ptr(), ptr::null_mut()) })?;
    Ok(pack::proof_from(unsafe { &out.assume_init() }))
}

/// Many proofs of ONE circuit under one key (the reference's production shape: a key-encryption proof per buyer request,
/// lib/src/seller.rs:136-159, and the per-block proofs of docs/protocol.md:56): synthesis per circuit on the CPU, then one
/// `ctg_prove_batch` -- a single digit sort / bucket accumulation / reduction per MSM for a whole chunk of proofs.
/// (r, s) are drawn per proof in order, r first, as `create_random_proof` does.
pub fn create_random_proofs<C, R>(circuits: Vec<C>, pk: &DeviceProvingKey, rng: &mut R) -> Result<Vec<Proof<Bls12_381>>, Error>
where C: ConstraintSynthesizer<Fr>, R: RngCore + CryptoRng {
    let n = circuits.len();
    if n == 0 { return Ok(Vec::new()) }
    let systems = circuits.into_iter().map(synthesize).collect::<Result<Vec<_>, _>>()?;
    let g = pk.0.ctx.raw.lock().unwrap();
    let r1cs = pk.r1cs_for(*g, &systems[0])?;
    let m = systems[0].num_instance_variables() + systems[0].num_witness_variables();
    let mut st = pk.0.staging.lock().unwrap();
    st.ensure(*g, n * m)?;
    let z = unsafe { std::slice::from_raw_parts_mut(st.ptr, n * m) };
    let mut rs = Vec::with_capacity(2 * n);
    for (i, cs) in systems.iter().enumerate() {
        let inner = cs.borrow().ok_or(SynthesisError::MissingCS)?;
        if inner.instance_assignment.len() + inner.witness_assignment.len() != m { return Err(Error::Synthesis(SynthesisError::MalformedVerifyingKey)) }
        for (dst, x) in z[i * m..(i + 1) * m].iter_mut().zip(inner.instance_assignment.iter().chain(inner.witness_assignment.iter())) { *dst = pack::fr_image(x) }
        let r = Fr::rand(rng);
        let s = Fr::rand(rng);
        rs.push(pack::fr_image(&r));
        rs.push(pack::fr_image(&s));
    }
    let mut out = vec![std::mem::MaybeUninit::<sys::ctg_proof>::uninit(); n];
    check(*g, unsafe { sys::ctg_prove_batch(*g, pk.0.pk, r1cs, n as u32, st.ptr, rs.as_ptr(), out.as_mut_ptr() as *mut sys::ctg_proof) })?;
    Ok(out.iter().map(|p| pack::proof_from(unsafe { &*p.as_ptr() })).collect())
}

/// `SNARK` front: the reference's call sites change `Groth16::<Bls12_381>` to `Groth16Gpu` and `ProvingKey` to `DeviceProvingKey`.
pub struct Groth16Gpu;
impl SNARK<Fr> for Groth16Gpu {
    type ProvingKey = DeviceProvingKey;
    type VerifyingKey = VerifyingKey<Bls12_381>;
    type Proof = Proof<Bls12_381>;
    type ProcessedVerifyingKey = PreparedVerifyingKey<Bls12_381>;
    type Error = Error;

    /// `Groth16::setup` (reference: circuits-ark/src/encryption.rs:352, lib/src/utils.rs:87-101): key generation on the global
    /// context's device, the key then goes to that device's prover cache.
    fn circuit_specific_setup<C: ConstraintSynthesizer<Fr>, R: RngCore + CryptoRng>(circuit: C, rng: &mut R)
        -> Result<(Self::ProvingKey, Self::VerifyingKey), Self::Error> {
        let ctx = Context::global()?;
        let pk = generate_random_parameters(&ctx, circuit, rng)?;
        // num_instance = |gamma_abc_g1|; num_constraints + num_instance <= |h_query| + 1 = the domain size: the domain is all the
        // prover needs from the constraint count, so the largest count that maps to the same domain stands in for it
        let num_instance = pk.vk.gamma_abc_g1.len();
        let num_constraints = pk.h_query.len() + 1 - num_instance;
        let dpk = DeviceProvingKey::upload(&ctx, &pk, num_instance, num_constraints)?;
        Ok((dpk, pk.vk))
    }
    fn prove<C: ConstraintSynthesizer<Fr>, R: RngCore + CryptoRng>(pk: &Self::ProvingKey, circuit: C, rng: &mut R) -> Result<Self::Proof, Self::Error> {
        create_random_proof(circuit, pk, rng)
    }
    /// Verification is unchanged: ark_groth16 on the host (one proof is four pairings; nothing to offload).
    fn process_vk(vk: &Self::VerifyingKey) -> Result<Self::ProcessedVerifyingKey, Self::Error> { Ok(ark_groth16::prepare_verifying_key(vk)) }
    fn verify_with_processed_vk(pvk: &Self::ProcessedVerifyingKey, public: &[Fr], proof: &Self::Proof) -> Result<bool, Self::Error> {
        Ok(ark_groth16::verify_proof(pvk, proof, public)?)
    }
}
