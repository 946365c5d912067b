#!/usr/bin/env python
"""Generates tests/golden/poseidon_reference_vectors.json: sponge digests, Merkle trees, a path and hashed-ElGamal ciphertexts
computed by the big-int restatement (oracle/py) UNDER THE REFERENCE'S OWN Poseidon parameters (tests/golden/reference_constants.json,
extracted from circuits-ark/src/parameters.rs:12-229 by make_reference_constants.py) and the reference-held public key ALICE_PK
(circuits-ark/src/utils.rs:115).  The constants and the key are reference bytes; the digests are still OUR restatement of
ark-sponge 0.3.0 applied to them (the sponge layout stays unpinned until a Rust-side known-answer test exists).

    python tests/golden/make_golden_reference_vectors.py       # deterministic; needs no /root/reference
"""
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.py import bls12_381 as B  # noqa: E402
from oracle.py import jubjub as J  # noqa: E402
from oracle.py import poseidon as P  # noqa: E402

H = lambda x: format(x, "x")


def reference_parameters():
    g = json.load(open(os.path.join(HERE, "reference_constants.json")))
    return P.PoseidonParameters(g["full_rounds"], g["partial_rounds"], g["alpha"], [[int(v, 16) for v in row] for row in g["mds_hex"]],
                                [[int(v, 16) for v in row] for row in g["ark_hex"]]), g


def main():
    rnd = random.Random(20261105)
    R = B.R_MOD
    p, g = reference_parameters()
    hashes = []
    for k in (1, 2, 3, 5):
        for row in ([0] * k, [R - 1] * k, [rnd.randrange(R) for _ in range(k)], [rnd.randrange(256) for _ in range(k)]):
            hashes.append({"inputs": [H(v) for v in row], "digest": H(P.crh(p, row))})
    trees = []
    for n in (2, 8, 32):
        leaves = [rnd.randrange(R) for _ in range(n)]
        t = P.MerkleTree(p, leaves)
        sib, path, idx = t.generate_proof(n - 3 if n > 2 else 1)
        trees.append({"leaves": [H(v) for v in leaves], "leaf_nodes": [H(v) for v in t.leaf_nodes], "non_leaf_nodes": [H(v) for v in t.non_leaf_nodes],
                      "path": {"leaf_index": idx, "leaf_sibling_hash": H(sib), "auth_path": [H(v) for v in path]}})
    pt = lambda q: [H(q[0]), H(q[1])]
    alice_pk = J.point_decompress(bytes.fromhex(g["ALICE_PK"]))
    alice_sk = J.scalar_from_bytes(bytes.fromhex(g["ALICE_SK"]))
    assert alice_pk is not None and alice_sk is not None
    ciphertexts = []
    for chunks in (1, 3):
        r = rnd.randrange(J.SUBGROUP_ORDER)
        msg = [rnd.randrange(256) for _ in range(chunks)]       # plaintext bytes, as bytes_to_plaintext_chunks_direct feeds them
        c1, c2 = J.encrypt(p, alice_pk, msg, r)
        ciphertexts.append({"r": H(r), "msg": [H(m) for m in msg], "c1": pt(c1), "c2": [H(c) for c in c2]})
    out = {"field": "BLS12-381 Fr", "parameters": "tests/golden/reference_constants.json (circuits-ark/src/parameters.rs:12-229)",
           "hashes": hashes, "trees": trees,
           "alice_pk": pt(alice_pk), "alice_sk": H(alice_sk), "alice_sk_times_generator": pt(J.mul(J.GEN, alice_sk)),
           "ciphertexts_to_alice_pk": ciphertexts}
    path = os.path.join(HERE, "poseidon_reference_vectors.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=0, sort_keys=True)
        f.write("\n")
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
