//! Witness-side native crypto in bulk (SOURCE ONLY, uncompiled here): the Poseidon Merkle tree over a ciphertext and the
//! hashed-ElGamal encryption of many messages, on the device.  Replaces, for large inputs,
//!   circuits-ark/src/sample_entries.rs:127-134  SampleEntries::build_merkle_tree
//!   circuits-ark/src/encryption.rs:154-172      EncryptCircuit::encrypt (one call per message there)
//! Values are repacked through `pack::fr_image` like everything else in this crate (ark-ff's Fp256 is not repr(C)).
use crate::{pack::fr_image, Context};
use ark_bls12_381::Fr;
use ark_ed_on_bls12_381::{EdwardsAffine, EdwardsParameters, Fr as JubJubScalar};
use ark_ec::models::TEModelParameters;
use ark_ff::{BigInteger256, Fp256, PrimeField};
use ark_relations::r1cs::SynthesisError;
use ark_sponge::poseidon::PoseidonParameters;
use ctg_groth16_sys as sys;
use std::ptr;

fn fr_from(v: &sys::ctg_fr) -> Fr {
    let mut l = [0u64; 4];
    for i in 0..4 { l[i] = v.v[2 * i] as u64 | (v.v[2 * i + 1] as u64) << 32; }
    Fp256::new(BigInteger256(l)) // `new` takes the Montgomery image as is
}
const ZR: sys::ctg_fr = sys::ctg_fr { v: [0; 8] };

/// `PoseidonParameters<Fr>` (rate 2, capacity 1: what `PoseidonParameters::new` of ark-sponge 0.3 builds) on the device.
pub struct PoseidonDevice<'c> { ctx: &'c Context, raw: *mut sys::ctg_poseidon }
unsafe impl<'c> Send for PoseidonDevice<'c> {}
unsafe impl<'c> Sync for PoseidonDevice<'c> {}

/// `MerkleTree { leaf_nodes, non_leaf_nodes }` of ark-crypto-primitives, same order (non-leaf nodes in heap order, root first).
pub struct MerkleNodes { pub leaf_nodes: Vec<Fr>, pub non_leaf_nodes: Vec<Fr> }
impl MerkleNodes {
    pub fn root(&self) -> Fr { self.non_leaf_nodes[0] }
    /// (leaf_sibling_hash, auth_path root side first) of `MerkleTree::generate_proof(index)`.
    pub fn path(&self, index: usize) -> (Fr, Vec<Fr>) {
        let n = self.leaf_nodes.len();
        let sibling = self.leaf_nodes[index ^ 1];
        let mut node = (n - 1 + index - 1) / 2; // heap index of the leaf's parent
        let mut path = Vec::new();
        while node != 0 {
            path.push(self.non_leaf_nodes[if node % 2 == 1 { node + 1 } else { node - 1 }]);
            node = (node - 1) / 2;
        }
        path.reverse();
        (sibling, path)
    }
}

impl<'c> PoseidonDevice<'c> {
    pub fn new(ctx: &'c Context, params: &PoseidonParameters<Fr>) -> Result<Self, SynthesisError> {
        let mds: Vec<sys::ctg_fr> = params.mds.iter().flatten().map(fr_image).collect();
        let ark: Vec<sys::ctg_fr> = params.ark.iter().flatten().map(fr_image).collect();
        if mds.len() != 9 || ark.len() != 3 * (params.full_rounds + params.partial_rounds) as usize { return Err(SynthesisError::Unsatisfiable) }
        let mut raw = ptr::null_mut();
        let g = ctx.raw.lock().unwrap();
        let rc = unsafe { sys::ctg_poseidon_new(*g, params.full_rounds, params.partial_rounds, params.alpha as u32, mds.as_ptr(), ark.as_ptr(), &mut raw) };
        if rc != sys::CTG_OK { return Err(SynthesisError::Unsatisfiable) }
        Ok(Self { ctx, raw })
    }

    /// `poseidon::CRH::evaluate` for every row of `k` elements (`k = 2`: `TwoToOneCRH::evaluate`).
    pub fn hash_batch(&self, inputs: &[Fr], k: usize) -> Result<Vec<Fr>, SynthesisError> {
        if k == 0 || inputs.len() % k != 0 { return Err(SynthesisError::Unsatisfiable) }
        let n = inputs.len() / k;
        let packed: Vec<sys::ctg_fr> = inputs.iter().map(fr_image).collect();
        let mut out = vec![ZR; n];
        let g = self.ctx.raw.lock().unwrap();
        let rc = unsafe { sys::ctg_poseidon_hash_batch(*g, self.raw, packed.as_ptr(), n as u64, k as u32, out.as_mut_ptr()) };
        if rc != sys::CTG_OK { return Err(SynthesisError::Unsatisfiable) }
        Ok(out.iter().map(fr_from).collect())
    }

    /// `MerkleTree::new(params, params, leaves)` with one field element per leaf (a power of two >= 2 of them).
    pub fn merkle_tree(&self, leaves: &[Fr]) -> Result<MerkleNodes, SynthesisError> {
        let n = leaves.len();
        let packed: Vec<sys::ctg_fr> = leaves.iter().map(fr_image).collect();
        let mut leaf_nodes = vec![ZR; n];
        let mut inner = vec![ZR; n.saturating_sub(1)];
        let g = self.ctx.raw.lock().unwrap();
        let rc = unsafe { sys::ctg_poseidon_merkle_tree(*g, self.raw, packed.as_ptr(), n as u64, leaf_nodes.as_mut_ptr(), inner.as_mut_ptr()) };
        if rc != sys::CTG_OK { return Err(SynthesisError::Unsatisfiable) }
        Ok(MerkleNodes { leaf_nodes: leaf_nodes.iter().map(fr_from).collect(), non_leaf_nodes: inner.iter().map(fr_from).collect() })
    }

    /// `EncryptCircuit::encrypt(pk, msg, r, params)` for every (r, msg) pair under one public key:
    /// c1 = G * r, pad = sponge(pk * r), c2 = pad + m for every chunk m of the message.
    pub fn encrypt_batch(&self, generator: &EdwardsAffine, pk: &EdwardsAffine, randomness: &[JubJubScalar], messages: &[Vec<Fr>])
        -> Result<Vec<(EdwardsAffine, Vec<Fr>)>, SynthesisError> {
        let n = randomness.len();
        if messages.len() != n { return Err(SynthesisError::Unsatisfiable) }
        let d = fr_image(&<EdwardsParameters as TEModelParameters>::COEFF_D);
        let ks: Vec<[u64; 4]> = randomness.iter().map(|r| r.into_repr().0).collect(); // little-endian limbs = 8 u32 words each
        let base = |p: &EdwardsAffine| [fr_image(&p.x), fr_image(&p.y)];
        let (gen_xy, pk_xy) = (base(generator), base(pk));
        let mut c1 = vec![ZR; 2 * n];
        let mut shared = vec![ZR; 2 * n];
        let mut pad = vec![ZR; n];
        let g = self.ctx.raw.lock().unwrap();
        let rc = unsafe {
            let mut rc = sys::ctg_te_mul_batch(*g, &d, gen_xy.as_ptr(), 1, ks.as_ptr().cast(), n as u64, c1.as_mut_ptr());
            if rc == sys::CTG_OK { rc = sys::ctg_te_mul_batch(*g, &d, pk_xy.as_ptr(), 1, ks.as_ptr().cast(), n as u64, shared.as_mut_ptr()); }
            if rc == sys::CTG_OK { rc = sys::ctg_poseidon_hash_batch(*g, self.raw, shared.as_ptr(), n as u64, 2, pad.as_mut_ptr()); }
            rc
        };
        if rc != sys::CTG_OK { return Err(SynthesisError::Unsatisfiable) }
        Ok((0..n).map(|i| {
            let dh = fr_from(&pad[i]);
            (EdwardsAffine::new(fr_from(&c1[2 * i]), fr_from(&c1[2 * i + 1])), messages[i].iter().map(|m| dh + m).collect())
        }).collect())
    }
}
impl<'c> Drop for PoseidonDevice<'c> {
    fn drop(&mut self) { unsafe { sys::ctg_poseidon_free(*self.ctx.raw.lock().unwrap(), self.raw) } }
}
