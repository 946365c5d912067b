// ctg_groth16.hpp -- C++17 host-side mirror of the ark-groth16 0.3.0 prove API over the C ABI of
// libctg_groth16.so (include/ctg_groth16.h).  The reference's host language is Rust (crates under ../rust,
// source only: no Rust toolchain on the build box); this header is the compiled-language stand-in with the same
// names, argument meaning and error behaviour:
//
//   ark_groth16::ProvingKey<Bls12_381>           -> ctg::ProvingKey            (host vectors, wire layout)
//   (new) device residency of the key            -> ctg::DeviceProvingKey      (lib/src/zk/encryption.rs:22,31-32)
//   ark_relations ConstraintMatrices             -> ctg::ConstraintMatrices    (rows of (coeff, column))
//   ark_groth16::create_proof(circuit, pk, r, s) -> ctg::create_proof(dpk, cs, r, s)
//   ark_groth16::create_random_proof(.., rng)    -> ctg::create_random_proof(dpk, cs, rng)   (r first, then s)
//   ark_groth16::Proof<Bls12_381>                -> ctg::Proof { a, b, c }
//   ark_groth16::generate_parameters(..)         -> ctg::generate_parameters(ctx, cs, trapdoor) -> ctg::DeviceParams
//   read_proving_key (lib/src/utils.rs:104-110)  -> ctg::read_proving_key(ctx, circuit.zkey bytes)  -> ctg::DeviceParams
//   ark_to_bytes(proof) / ark_from_bytes         -> ctg::proof_to_bytes / ctg::proof_from_bytes   (192 compressed bytes)
//   SynthesisError                               -> ctg::Error (std::runtime_error with the C ABI message)
//
// "cs" is the synthesized constraint system (matrices + full assignment): synthesis itself is circuit code and
// stays with the caller, as it stays in Rust in the reference (circuits-ark/src/encryption.rs:303-317).
// Header-only; link with -lctg_groth16.  No CPU fallback: Context's constructor throws without an sm_100 device.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/ctg_groth16.h"

namespace ctg {

struct Error : std::runtime_error { int32_t code; Error(int32_t c, const std::string& m) : std::runtime_error(m), code(c) {} };

class Context {
 public:
  explicit Context(int device = 0) {
    int32_t rc = ctg_ctx_create(&device, 1, &ctx_);
    if (rc != CTG_OK) throw Error(rc, "ctg_ctx_create failed: no usable sm_100 CUDA device (there is no CPU fallback)");
  }
  // several GPUs of one NVLink box in one process (ctg_ctx_create with n_devices > 1): devices[0] is the primary device; key
  // upload, prove and the stand-alone MSMs shard over all of them by themselves, every other call runs on the primary
  explicit Context(const std::vector<int>& devices) {
    int32_t rc = ctg_ctx_create(devices.data(), (int)devices.size(), &ctx_);
    if (rc != CTG_OK) throw Error(rc, "ctg_ctx_create failed: no usable sm_100 CUDA devices with peer access (there is no CPU fallback)");
  }
  int device_count() const { return ctg_ctx_device_count(ctx_); }
  ~Context() { ctg_ctx_destroy(ctx_); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  ctg_ctx* raw() const { return ctx_; }
  void check(int32_t rc) const { if (rc != CTG_OK) throw Error(rc, ctg_last_error(ctx_)); }

 private:
  ctg_ctx* ctx_ = nullptr;
};

// ark_groth16::ProvingKey<Bls12_381>; points in the wire layout (Montgomery, x || y, all-zero = infinity)
struct VerifyingKey {
  ctg_g1_affine alpha_g1{};
  ctg_g2_affine beta_g2{}, gamma_g2{}, delta_g2{};
  std::vector<ctg_g1_affine> gamma_abc_g1;
};
struct ProvingKey {
  VerifyingKey vk;
  ctg_g1_affine beta_g1{}, delta_g1{};
  std::vector<ctg_g1_affine> a_query, b_g1_query, h_query, l_query;
  std::vector<ctg_g2_affine> b_g2_query;
};
struct Proof { ctg_g1_affine a; ctg_g2_affine b; ctg_g1_affine c; };

// ark_relations::r1cs::ConstraintMatrices + the assignment of one synthesis run
using Row = std::vector<std::pair<ctg_fr, size_t>>;  // (coefficient (Montgomery), column); columns: instance first, then witness
struct ConstraintSystem {
  size_t num_instance_variables = 1;  // includes the constant 1
  size_t num_witness_variables = 0;
  std::vector<Row> a, b, c;           // num_constraints rows each
  std::vector<ctg_fr> instance_assignment, witness_assignment;
  size_t num_constraints() const { return a.size(); }
};

// CSR views of a ConstraintSystem's three matrices (storage lives in the object)
struct FlatMatrices {
  struct Flat { std::vector<uint64_t> rp; std::vector<uint32_t> col; std::vector<ctg_fr> val; ctg_csr view; };
  Flat f[3];
  explicit FlatMatrices(const ConstraintSystem& cs) {
    const std::vector<Row>* mats[3] = {&cs.a, &cs.b, &cs.c};
    for (int k = 0; k < 3; k++) {
      f[k].rp.push_back(0);
      for (const Row& row : *mats[k]) {
        for (const auto& e : row) { f[k].val.push_back(e.first); f[k].col.push_back((uint32_t)e.second); }
        f[k].rp.push_back(f[k].col.size());
      }
      f[k].view = ctg_csr{f[k].rp.data(), f[k].col.data(), f[k].val.data()};
    }
  }
};

// A Groth16 key resident in HBM in arkworks' order (ctg_params): the output of key generation or of reading circuit.zkey.
class DeviceParams {
 public:
  DeviceParams(const Context& ctx, ctg_params* p) : ctx_(ctx), p_(p) {}
  ~DeviceParams() { ctg_params_free(ctx_.raw(), p_); }
  DeviceParams(const DeviceParams&) = delete;
  DeviceParams& operator=(const DeviceParams&) = delete;
  const Context& ctx() const { return ctx_; }
  const ctg_params* raw() const { return p_; }
  ctg_pk_view view(uint64_t* num_instance = nullptr) const {
    ctg_pk_view v{};
    ctx_.check(ctg_params_pk_view(ctx_.raw(), p_, &v, nullptr, num_instance));
    return v;
  }
  // ProvingKey::serialize_unchecked (circuit.zkey) and VerifyingKey::serialize (verification.key): write_circuit_artifacts
  std::vector<uint8_t> serialize_unchecked() const { return ser(ctg_params_serialize_unchecked); }
  std::vector<uint8_t> serialize_vk() const { return ser(ctg_params_serialize_vk); }

 private:
  template <class Fn> std::vector<uint8_t> ser(Fn fn) const {
    uint64_t n = 0;
    ctx_.check(fn(ctx_.raw(), p_, nullptr, 0, &n));
    std::vector<uint8_t> out(n);
    ctx_.check(fn(ctx_.raw(), p_, out.data(), n, &n));
    return out;
  }
  const Context& ctx_;
  ctg_params* p_;
};

// ark_groth16::generate_parameters after Setup-mode synthesis; the trapdoor is drawn by the caller in arkworks' order
inline std::unique_ptr<DeviceParams> generate_parameters(const Context& ctx, const ConstraintSystem& cs, const ctg_trapdoor& trapdoor) {
  FlatMatrices m(cs);
  ctg_params* p = nullptr;
  ctx.check(ctg_generate_parameters(ctx.raw(), cs.num_instance_variables, cs.num_witness_variables, cs.num_constraints(), &m.f[0].view, &m.f[1].view,
                                    &m.f[2].view, &trapdoor, &p));
  return std::make_unique<DeviceParams>(ctx, p);
}
// read_proving_key: ProvingKey::deserialize_unchecked of a circuit.zkey image, decoded on the device
inline std::unique_ptr<DeviceParams> read_proving_key(const Context& ctx, const std::vector<uint8_t>& image) {
  ctg_params* p = nullptr;
  ctx.check(ctg_params_deserialize_unchecked(ctx.raw(), image.data(), image.size(), &p));
  return std::make_unique<DeviceParams>(ctx, p);
}

class DeviceProvingKey {
 public:
  // from a key already in HBM (device-to-device)
  DeviceProvingKey(const DeviceParams& params, size_t num_constraints, int copies = 0, int c_bits = 0) : ctx_(params.ctx()), nc_(num_constraints) {
    uint64_t ni = 0;
    ctg_pk_view v = params.view(&ni);
    ni_ = ni;
    ctx_.check(ctg_pk_upload(ctx_.raw(), &v, ni, num_constraints, copies, c_bits, &pk_));
  }
  // num_instance / num_constraints: shape of the circuit the key was generated for
  DeviceProvingKey(const Context& ctx, const ProvingKey& pk, size_t num_instance, size_t num_constraints, int copies = 0, int c_bits = 0)
      : ctx_(ctx), ni_(num_instance), nc_(num_constraints) {
    ctg_pk_view v{};
    v.alpha_g1 = &pk.vk.alpha_g1; v.beta_g1 = &pk.beta_g1; v.delta_g1 = &pk.delta_g1;
    v.beta_g2 = &pk.vk.beta_g2; v.delta_g2 = &pk.vk.delta_g2;
    v.a_query = pk.a_query.data(); v.a_len = pk.a_query.size();
    v.b_g1_query = pk.b_g1_query.data(); v.b_g1_len = pk.b_g1_query.size();
    v.b_g2_query = pk.b_g2_query.data(); v.b_g2_len = pk.b_g2_query.size();
    v.h_query = pk.h_query.data(); v.h_len = pk.h_query.size();
    v.l_query = pk.l_query.data(); v.l_len = pk.l_query.size();
    ctx_.check(ctg_pk_upload(ctx_.raw(), &v, num_instance, num_constraints, copies, c_bits, &pk_));
  }
  ~DeviceProvingKey() {
    if (r1cs_) ctg_r1cs_free(ctx_.raw(), r1cs_);
    ctg_pk_free(ctx_.raw(), pk_);
  }
  DeviceProvingKey(const DeviceProvingKey&) = delete;
  DeviceProvingKey& operator=(const DeviceProvingKey&) = delete;

  // R1CS matrices are per-circuit constants: uploaded with the first proof, reused afterwards.
  const ctg_r1cs* r1cs_for(const ConstraintSystem& cs) {
    if (cs.num_instance_variables != ni_ || cs.num_constraints() != nc_) throw Error(CTG_ERR_INVALID, "constraint system does not match the proving key");
    if (r1cs_) return r1cs_;
    FlatMatrices m(cs);
    auto& f = m.f;
    ctx_.check(ctg_r1cs_upload(ctx_.raw(), cs.num_instance_variables, cs.num_witness_variables, cs.num_constraints(), &f[0].view, &f[1].view,
                               &f[2].view, &r1cs_));
    return r1cs_;
  }
  const Context& ctx() const { return ctx_; }
  const ctg_pk* raw() const { return pk_; }

 private:
  const Context& ctx_;
  ctg_pk* pk_ = nullptr;
  ctg_r1cs* r1cs_ = nullptr;
  size_t ni_, nc_;
};

// ark_groth16::create_proof(circuit, pk, r, s) after synthesis
inline Proof create_proof(DeviceProvingKey& pk, const ConstraintSystem& cs, const ctg_fr& r, const ctg_fr& s, ctg_timings* timings = nullptr) {
  const ctg_r1cs* r1cs = pk.r1cs_for(cs);
  if (cs.instance_assignment.size() != cs.num_instance_variables || cs.witness_assignment.size() != cs.num_witness_variables)
    throw Error(CTG_ERR_INVALID, "AssignmentMissing");  // SynthesisError::AssignmentMissing
  std::vector<ctg_fr> z;
  z.reserve(cs.instance_assignment.size() + cs.witness_assignment.size());
  z.insert(z.end(), cs.instance_assignment.begin(), cs.instance_assignment.end());
  z.insert(z.end(), cs.witness_assignment.begin(), cs.witness_assignment.end());
  ctg_proof out;
  pk.ctx().check(ctg_prove(pk.ctx().raw(), pk.raw(), r1cs, z.data(), &r, &s, &out, timings));
  return Proof{out.a, out.b, out.c};
}

// Many proofs of one circuit under one key (ctg_prove_batch: one digit sort / accumulation / reduction per MSM for a chunk of
// proofs).  systems[i] carries the i-th assignment; rs = r_0, s_0, r_1, s_1, ...
inline std::vector<Proof> create_proofs(DeviceProvingKey& pk, const std::vector<const ConstraintSystem*>& systems, const std::vector<ctg_fr>& rs) {
  if (systems.empty()) return {};
  if (rs.size() != 2 * systems.size()) throw Error(CTG_ERR_INVALID, "create_proofs: two blinding scalars per proof");
  const ctg_r1cs* r1cs = pk.r1cs_for(*systems[0]);
  const size_t m = systems[0]->instance_assignment.size() + systems[0]->witness_assignment.size();
  std::vector<ctg_fr> z;
  z.reserve(m * systems.size());
  for (const ConstraintSystem* cs : systems) {
    if (cs->instance_assignment.size() + cs->witness_assignment.size() != m) throw Error(CTG_ERR_INVALID, "create_proofs: assignments of different circuits");
    z.insert(z.end(), cs->instance_assignment.begin(), cs->instance_assignment.end());
    z.insert(z.end(), cs->witness_assignment.begin(), cs->witness_assignment.end());
  }
  std::vector<ctg_proof> raw(systems.size());
  pk.ctx().check(ctg_prove_batch(pk.ctx().raw(), pk.raw(), r1cs, (uint32_t)systems.size(), z.data(), rs.data(), raw.data()));
  std::vector<Proof> out;
  out.reserve(raw.size());
  for (const ctg_proof& p : raw) out.push_back(Proof{p.a, p.b, p.c});
  return out;
}

// ark_groth16::create_random_proof: `rng()` must return one uniformly random Fr as a Montgomery image, as ark-ff's
// Fr::rand does (rejection sampling of the image itself); it is called for r first, then for s.
template <class Rng> inline Proof create_random_proof(DeviceProvingKey& pk, const ConstraintSystem& cs, Rng&& rng) {
  const ctg_fr r = rng();
  const ctg_fr s = rng();
  return create_proof(pk, cs, r, s);
}

// ark_to_bytes(proof) / ark_from_bytes::<Proof>: arkworks' compressed canonical form, 48 + 96 + 48 bytes
inline std::vector<uint8_t> proof_to_bytes(const Proof& p) {
  ctg_proof raw{p.a, p.b, p.c};
  std::vector<uint8_t> out(CTG_PROOF_BYTES);
  if (ctg_proof_serialize(&raw, out.data()) != CTG_OK) throw Error(CTG_ERR_INVALID, "proof_to_bytes");
  return out;
}
inline Proof proof_from_bytes(const std::vector<uint8_t>& bytes, bool check_subgroup = true) {
  ctg_proof raw;
  if (bytes.size() != CTG_PROOF_BYTES || ctg_proof_deserialize(bytes.data(), check_subgroup, &raw) != CTG_OK)
    throw Error(CTG_ERR_INVALID, "SerializationError::InvalidData");
  return Proof{raw.a, raw.b, raw.c};
}

// ---- bulk Poseidon / ciphertext Merkle tree (circuits-ark/src/sample_entries.rs:127-134, encryption.rs:154-172) ----
// ark_sponge::poseidon::PoseidonParameters (rate 2, capacity 1) resident on the device.
struct MerkleTree {               // ark_crypto_primitives::merkle_tree::MerkleTree node arrays, in its order
  std::vector<ctg_fr> leaf_nodes;      // CRH([leaf])
  std::vector<ctg_fr> non_leaf_nodes;  // heap order, root first
  const ctg_fr& root() const { return non_leaf_nodes[0]; }
};
class Poseidon {
 public:
  // mds: 9 elements row-major; ark: (full_rounds + partial_rounds) * 3 elements; Montgomery images
  Poseidon(const Context& ctx, uint32_t full_rounds, uint32_t partial_rounds, uint32_t alpha, const std::vector<ctg_fr>& mds, const std::vector<ctg_fr>& ark)
      : ctx_(ctx) {
    if (mds.size() != 9 || ark.size() != (size_t)3 * (full_rounds + partial_rounds)) throw Error(CTG_ERR_INVALID, "poseidon: parameter shape");
    ctx_.check(ctg_poseidon_new(ctx_.raw(), full_rounds, partial_rounds, alpha, mds.data(), ark.data(), &p_));
  }
  ~Poseidon() { ctg_poseidon_free(ctx_.raw(), p_); }
  Poseidon(const Poseidon&) = delete;
  Poseidon& operator=(const Poseidon&) = delete;
  // CRH::evaluate for n inputs of k elements each (k = 2: TwoToOneCRH::evaluate)
  std::vector<ctg_fr> hash_batch(const std::vector<ctg_fr>& inputs, uint32_t k) const {
    if (k == 0 || inputs.size() % k) throw Error(CTG_ERR_INVALID, "poseidon: inputs must be n * k elements");
    std::vector<ctg_fr> out(inputs.size() / k);
    ctx_.check(ctg_poseidon_hash_batch(ctx_.raw(), p_, inputs.data(), out.size(), k, out.data()));
    return out;
  }
  // MerkleTree::new(params, params, leaves) with one field element per leaf
  MerkleTree merkle_tree(const std::vector<ctg_fr>& leaves) const {
    MerkleTree t;
    t.leaf_nodes.resize(leaves.size());
    t.non_leaf_nodes.resize(leaves.empty() ? 0 : leaves.size() - 1);
    ctx_.check(ctg_poseidon_merkle_tree(ctx_.raw(), p_, leaves.data(), leaves.size(), t.leaf_nodes.data(), t.non_leaf_nodes.data()));
    return t;
  }

 private:
  const Context& ctx_;
  ctg_poseidon* p_ = nullptr;
};

// bulk JubJub scalar multiplication (c1 = G.mul(r), p_r = pk.mul(r) of encrypt(), circuits-ark/src/encryption.rs:160-165):
// points = one shared (x, y) pair or one per scalar; scalars = 8 words each (into_repr()); returns n (x, y) pairs
inline std::vector<ctg_fr> te_mul_batch(const Context& ctx, const ctg_fr& d, const std::vector<ctg_fr>& points_xy, const std::vector<uint32_t>& scalars) {
  if (scalars.size() % 8 || points_xy.size() % 2) throw Error(CTG_ERR_INVALID, "te_mul: shapes");
  const uint64_t n = scalars.size() / 8;
  std::vector<ctg_fr> out(2 * n);
  ctx.check(ctg_te_mul_batch(ctx.raw(), &d, points_xy.data(), points_xy.size() / 2, scalars.data(), n, out.data()));
  return out;
}

}  // namespace ctg
