// Exercises contangle-zkcp_b200/host/ctg_groth16.hpp (the compiled-language mirror of the ark-groth16 API) end to end, the
// way the reference's seller process uses arkworks: setup -> write circuit.zkey / verification.key -> (restart) read the key
// -> prove -> serialize the proof (lib/src/zk/encryption.rs:156-176, 27-40, 65-85).
// usage: roundtrip <input.bin> <out_dir>     input.bin is written by tests/test_host_cpp.py (format below)
//   u64 ni, nw, nc | 3 x { u64 nnz, row_ptr[nc + 1] u64, col[nnz] u32, val[nnz] Fr } | z[ni + nw] Fr | ctg_trapdoor | r | s
// writes <out_dir>/circuit.zkey, verification.key, proof.bin (192 B).  Exit code 0 on success, 3 when no device is usable.
#include <cstdio>
#include <fstream>
#include <iostream>

#include "../../contangle-zkcp_b200/host/ctg_groth16.hpp"

static std::vector<uint8_t> slurp(const std::string& path) {
  std::ifstream f(path, std::ios::binary);
  return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}
static void spit(const std::string& path, const std::vector<uint8_t>& data) {
  std::ofstream f(path, std::ios::binary);
  f.write((const char*)data.data(), (std::streamsize)data.size());
}

int main(int argc, char** argv) {
  if (argc != 3) { std::fprintf(stderr, "usage: roundtrip <input.bin> <out_dir>\n"); return 2; }
  const std::vector<uint8_t> in = slurp(argv[1]);
  const std::string out_dir = argv[2];
  size_t off = 0;
  auto take = [&](void* dst, size_t bytes) { if (off + bytes > in.size()) throw std::runtime_error("short input"); std::memcpy(dst, in.data() + off, bytes); off += bytes; };
  try {
    uint64_t ni, nw, nc;
    take(&ni, 8); take(&nw, 8); take(&nc, 8);
    ctg::ConstraintSystem cs;
    cs.num_instance_variables = ni;
    cs.num_witness_variables = nw;
    std::vector<ctg::Row>* mats[3] = {&cs.a, &cs.b, &cs.c};
    for (int k = 0; k < 3; k++) {
      uint64_t nnz;
      take(&nnz, 8);
      std::vector<uint64_t> rp(nc + 1);
      std::vector<uint32_t> col(nnz);
      std::vector<ctg_fr> val(nnz);
      take(rp.data(), rp.size() * 8); take(col.data(), nnz * 4); take(val.data(), nnz * sizeof(ctg_fr));
      mats[k]->resize(nc);
      for (uint64_t i = 0; i < nc; i++)
        for (uint64_t e = rp[i]; e < rp[i + 1]; e++) (*mats[k])[i].push_back({val[e], col[e]});
    }
    cs.instance_assignment.resize(ni);
    cs.witness_assignment.resize(nw);
    take(cs.instance_assignment.data(), ni * sizeof(ctg_fr));
    take(cs.witness_assignment.data(), nw * sizeof(ctg_fr));
    ctg_trapdoor td;
    ctg_fr r, s;
    take(&td, sizeof(td)); take(&r, sizeof(r)); take(&s, sizeof(s));

    ctg::Context ctx(0);
    {  // "compile": Groth16::setup + write_circuit_artifacts
      auto params = ctg::generate_parameters(ctx, cs, td);
      spit(out_dir + "/circuit.zkey", params->serialize_unchecked());
      spit(out_dir + "/verification.key", params->serialize_vk());
    }
    // "sell": read_proving_key at start-up, then prove
    auto params = ctg::read_proving_key(ctx, slurp(out_dir + "/circuit.zkey"));
    ctg::DeviceProvingKey dpk(*params, nc);
    params.reset();   // the prover's cache holds its own copy
    int calls = 0;
    ctg::Proof proof = ctg::create_random_proof(dpk, cs, [&]() { return calls++ == 0 ? r : s; });   // r first, then s
    const std::vector<uint8_t> bytes = ctg::proof_to_bytes(proof);
    ctg::Proof back = ctg::proof_from_bytes(bytes);
    if (std::memcmp(&back, &proof, sizeof(proof)) != 0) { std::fprintf(stderr, "proof bytes do not round-trip\n"); return 1; }
    spit(out_dir + "/proof.bin", bytes);
    // a second proof reuses the cached R1CS and must be identical for the same (r, s)
    calls = 0;
    ctg::Proof again = ctg::create_random_proof(dpk, cs, [&]() { return calls++ == 0 ? r : s; });
    if (std::memcmp(&again, &proof, sizeof(proof)) != 0) { std::fprintf(stderr, "second proof differs\n"); return 1; }
    // error behaviour: a constraint system of another shape is refused (arkworks asserts on the lengths)
    ctg::ConstraintSystem other = cs;
    other.a.pop_back(); other.b.pop_back(); other.c.pop_back();
    try { ctg::create_proof(dpk, other, r, s); std::fprintf(stderr, "shape mismatch accepted\n"); return 1; } catch (const ctg::Error&) {}
    // batch entry point: three proofs of the same assignment with three (r, s) pairs; the one with the (r, s) above must be that proof
    {
      ctg_fr r2 = s, s2 = r;
      std::vector<const ctg::ConstraintSystem*> systems = {&cs, &cs, &cs};
      std::vector<ctg::Proof> batch = ctg::create_proofs(dpk, systems, {r2, s2, r, s, s, s});
      if (batch.size() != 3 || std::memcmp(&batch[1], &proof, sizeof(proof)) != 0) { std::fprintf(stderr, "batched proof differs\n"); return 1; }
      ctg::Proof one = ctg::create_proof(dpk, cs, r2, s2);
      if (std::memcmp(&batch[0], &one, sizeof(one)) != 0) { std::fprintf(stderr, "batched proof 0 differs\n"); return 1; }
    }
    // multi-device context (the device listed twice: the sharded code path on a one-GPU box): same key image, same proof
    {
      ctg::Context multi(std::vector<int>{0, 0});
      if (multi.device_count() != 2) { std::fprintf(stderr, "device_count\n"); return 1; }
      auto mparams = ctg::read_proving_key(multi, slurp(out_dir + "/circuit.zkey"));
      ctg::DeviceProvingKey mpk(*mparams, nc);
      mparams.reset();
      ctg::Proof sharded = ctg::create_proof(mpk, cs, r, s);
      if (std::memcmp(&sharded, &proof, sizeof(proof)) != 0) { std::fprintf(stderr, "sharded proof differs\n"); return 1; }
    }
    // bulk Poseidon: the Merkle tree over 8 leaves equals the tree rebuilt with the bulk hashes (values themselves are
    // checked against the big-int restatement in tests/test_poseidon.py); r, s and small constants serve as field elements
    {
      std::vector<ctg_fr> mds(9), ark(3 * 39), leaves(8);
      for (size_t i = 0; i < mds.size(); i++) { mds[i] = r; mds[i].v[0] ^= (uint32_t)(i + 1); mds[i].v[7] &= 0x3fffffffu; }
      for (size_t i = 0; i < ark.size(); i++) { ark[i] = s; ark[i].v[1] ^= (uint32_t)(i * 2654435761u); ark[i].v[7] &= 0x3fffffffu; }
      for (size_t i = 0; i < leaves.size(); i++) { leaves[i] = r; leaves[i].v[2] ^= (uint32_t)(i + 7); leaves[i].v[7] &= 0x3fffffffu; }
      ctg::Poseidon pos(ctx, 8, 31, 17, mds, ark);
      ctg::MerkleTree tree = pos.merkle_tree(leaves);
      std::vector<ctg_fr> level = pos.hash_batch(leaves, 1);
      if (std::memcmp(level.data(), tree.leaf_nodes.data(), 8 * sizeof(ctg_fr)) != 0) { std::fprintf(stderr, "leaf digests differ\n"); return 1; }
      for (size_t width = 4; width >= 1; width /= 2) {
        level = pos.hash_batch(level, 2);
        if (std::memcmp(level.data(), &tree.non_leaf_nodes[width - 1], width * sizeof(ctg_fr)) != 0) { std::fprintf(stderr, "tree level differs\n"); return 1; }
      }
      if (std::memcmp(&tree.root(), level.data(), sizeof(ctg_fr)) != 0) { std::fprintf(stderr, "root differs\n"); return 1; }
      try { pos.merkle_tree(std::vector<ctg_fr>(leaves.begin(), leaves.begin() + 3)); std::fprintf(stderr, "3 leaves accepted\n"); return 1; } catch (const ctg::Error&) {}
    }
    std::printf("ok %zu bytes of key, proof %zu bytes\n", slurp(out_dir + "/circuit.zkey").size(), bytes.size());
    return 0;
  } catch (const ctg::Error& e) {
    std::fprintf(stderr, "ctg::Error %d: %s\n", e.code, e.what());
    return e.code == CTG_ERR_CUDA || std::string(e.what()).find("no usable") != std::string::npos ? 3 : 1;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
}
