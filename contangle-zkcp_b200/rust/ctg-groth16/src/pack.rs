//! Repacking between arkworks in-memory types and the wire layout of include/ctg_groth16.h.
use ark_bls12_381::{Bls12_381, Fq, Fq2, Fr, G1Affine, G2Affine};
use ark_ff::{Fp256, Fp384, Zero};
use ark_groth16::{Proof, ProvingKey};
use ctg_groth16_sys as sys;

/// Montgomery image of an Fr: ark-ff 0.3 `Fp256<P>(pub BigInteger256, PhantomData)`, `.0 .0 : [u64; 4]`.
pub fn fr_image(x: &Fr) -> sys::ctg_fr {
    let l = (x.0).0;
    let mut v = [0u32; 8];
    for i in 0..4 { v[2 * i] = l[i] as u32; v[2 * i + 1] = (l[i] >> 32) as u32; }
    sys::ctg_fr { v }
}
fn fq_image(x: &Fq) -> sys::ctg_fq {
    let l = (x.0).0;
    let mut v = [0u32; 12];
    for i in 0..6 { v[2 * i] = l[i] as u32; v[2 * i + 1] = (l[i] >> 32) as u32; }
    sys::ctg_fq { v }
}
fn fq_from(v: &sys::ctg_fq) -> Fq {
    let mut l = [0u64; 6];
    for i in 0..6 { l[i] = v.v[2 * i] as u64 | (v.v[2 * i + 1] as u64) << 32; }
    Fp384::new(ark_ff::BigInteger384(l)) // `new` takes the Montgomery image as is
}
const ZQ: sys::ctg_fq = sys::ctg_fq { v: [0; 12] };
pub fn g1_image(p: &G1Affine) -> sys::ctg_g1_affine {
    if p.infinity { sys::ctg_g1_affine { x: ZQ, y: ZQ } } else { sys::ctg_g1_affine { x: fq_image(&p.x), y: fq_image(&p.y) } }
}
pub fn g2_image(p: &G2Affine) -> sys::ctg_g2_affine {
    if p.infinity { sys::ctg_g2_affine { x_c0: ZQ, x_c1: ZQ, y_c0: ZQ, y_c1: ZQ } }
    else { sys::ctg_g2_affine { x_c0: fq_image(&p.x.c0), x_c1: fq_image(&p.x.c1), y_c0: fq_image(&p.y.c0), y_c1: fq_image(&p.y.c1) } }
}
fn g1_from(p: &sys::ctg_g1_affine) -> G1Affine {
    let (x, y) = (fq_from(&p.x), fq_from(&p.y));
    if x.is_zero() && y.is_zero() { G1Affine::zero() } else { G1Affine::new(x, y, false) }
}
fn g2_from(p: &sys::ctg_g2_affine) -> G2Affine {
    let (x, y) = (Fq2::new(fq_from(&p.x_c0), fq_from(&p.x_c1)), Fq2::new(fq_from(&p.y_c0), fq_from(&p.y_c1)));
    if x.is_zero() && y.is_zero() { G2Affine::zero() } else { G2Affine::new(x, y, false) }
}
pub fn proof_from(p: &sys::ctg_proof) -> Proof<Bls12_381> { Proof { a: g1_from(&p.a), b: g2_from(&p.b), c: g1_from(&p.c) } }

pub struct PackedPk {
    singles1: [sys::ctg_g1_affine; 3], singles2: [sys::ctg_g2_affine; 2],
    a: Vec<sys::ctg_g1_affine>, b1: Vec<sys::ctg_g1_affine>, b2: Vec<sys::ctg_g2_affine>, h: Vec<sys::ctg_g1_affine>, l: Vec<sys::ctg_g1_affine>,
}
impl From<&ProvingKey<Bls12_381>> for PackedPk {
    fn from(pk: &ProvingKey<Bls12_381>) -> Self {
        Self {
            singles1: [g1_image(&pk.vk.alpha_g1), g1_image(&pk.beta_g1), g1_image(&pk.delta_g1)],
            singles2: [g2_image(&pk.vk.beta_g2), g2_image(&pk.vk.delta_g2)],
            a: pk.a_query.iter().map(g1_image).collect(), b1: pk.b_g1_query.iter().map(g1_image).collect(),
            b2: pk.b_g2_query.iter().map(g2_image).collect(), h: pk.h_query.iter().map(g1_image).collect(),
            l: pk.l_query.iter().map(g1_image).collect(),
        }
    }
}
impl PackedPk {
    pub fn view(&self) -> sys::ctg_pk_view {
        sys::ctg_pk_view {
            alpha_g1: &self.singles1[0], beta_g1: &self.singles1[1], delta_g1: &self.singles1[2],
            beta_g2: &self.singles2[0], delta_g2: &self.singles2[1],
            a_query: self.a.as_ptr(), a_len: self.a.len() as u64, b_g1_query: self.b1.as_ptr(), b_g1_len: self.b1.len() as u64,
            b_g2_query: self.b2.as_ptr(), b_g2_len: self.b2.len() as u64, h_query: self.h.as_ptr(), h_len: self.h.len() as u64,
            l_query: self.l.as_ptr(), l_len: self.l.len() as u64,
        }
    }
}

/// ark-relations `Matrix<F> = Vec<Vec<(F, usize)>>` -> CSR
pub struct Csr { row_ptr: Vec<u64>, col: Vec<u32>, val: Vec<sys::ctg_fr> }
impl Csr {
    pub fn from_rows(rows: &[Vec<(Fr, usize)>]) -> Self {
        let mut c = Csr { row_ptr: vec![0], col: vec![], val: vec![] };
        for row in rows {
            for (coeff, idx) in row { c.col.push(*idx as u32); c.val.push(fr_image(coeff)); }
            c.row_ptr.push(c.col.len() as u64);
        }
        c
    }
    pub fn view(&self) -> sys::ctg_csr { sys::ctg_csr { row_ptr: self.row_ptr.as_ptr(), col: self.col.as_ptr(), val: self.val.as_ptr() } }
}
#[allow(dead_code)] fn _assert_layout(_: Fp256<ark_bls12_381::FrParameters>) {}

/// ctg_params_download -> ark_groth16::ProvingKey (used by `generate_random_parameters`)
pub fn download_pk(ctx: *mut sys::ctg_ctx, params: *const sys::ctg_params, ni: usize, nw: usize, n: usize)
    -> Result<ProvingKey<Bls12_381>, ark_relations::r1cs::SynthesisError> {
    let m = ni + nw;
    let z1 = sys::ctg_g1_affine { x: ZQ, y: ZQ };
    let z2 = sys::ctg_g2_affine { x_c0: ZQ, x_c1: ZQ, y_c0: ZQ, y_c1: ZQ };
    let (mut s1, mut s2) = ([z1; 3], [z2; 3]);
    let (mut abc, mut a, mut b1, mut b2, mut h, mut l) = (vec![z1; ni], vec![z1; m], vec![z1; m], vec![z2; m], vec![z1; n - 1], vec![z1; nw]);
    let dst = sys::ctg_params_out {
        alpha_g1: &mut s1[0], beta_g1: &mut s1[1], delta_g1: &mut s1[2], beta_g2: &mut s2[0], gamma_g2: &mut s2[1], delta_g2: &mut s2[2],
        gamma_abc_g1: abc.as_mut_ptr(), a_query: a.as_mut_ptr(), b_g1_query: b1.as_mut_ptr(), b_g2_query: b2.as_mut_ptr(),
        h_query: h.as_mut_ptr(), l_query: l.as_mut_ptr(),
    };
    if unsafe { sys::ctg_params_download(ctx, params, &dst) } != sys::CTG_OK { return Err(ark_relations::r1cs::SynthesisError::Unsatisfiable) }
    let vk = ark_groth16::VerifyingKey { alpha_g1: g1_from(&s1[0]), beta_g2: g2_from(&s2[0]), gamma_g2: g2_from(&s2[1]), delta_g2: g2_from(&s2[2]),
                                         gamma_abc_g1: abc.iter().map(g1_from).collect() };
    Ok(ProvingKey { vk, beta_g1: g1_from(&s1[1]), delta_g1: g1_from(&s1[2]), a_query: a.iter().map(g1_from).collect(),
                    b_g1_query: b1.iter().map(g1_from).collect(), b_g2_query: b2.iter().map(g2_from).collect(),
                    h_query: h.iter().map(g1_from).collect(), l_query: l.iter().map(g1_from).collect() })
}
