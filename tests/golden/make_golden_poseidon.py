#!/usr/bin/env python
"""Generates tests/golden/poseidon_bls12_381_fr.json with the big-int restatement oracle/py/poseidon.py: frozen parameters,
sponge digests, Merkle trees, JubJub multiples and hashed-ElGamal ciphertexts that the CUDA kernels (and the restatement itself, after any edit) are checked against.

The reference holds no hash value for this path and cannot be run here (no Rust; ark-sponge / ark-crypto-primitives are not
in the tree), so the fixture freezes OUR restatement of ark-sponge 0.3.0's conventions: parity stays unpinned.  The constants
are seeded random field elements in the reference's shape (8 + 31 rounds, alpha 17) and two other shapes; they are stored
in the fixture, so it is self-contained.

    python tests/golden/make_golden_poseidon.py        # rewrites the JSON next to this script (deterministic)
"""
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.py import bls12_381 as B  # noqa: E402
from oracle.py import jubjub as J  # noqa: E402
from oracle.py import poseidon as P  # noqa: E402

H = lambda x: format(x, "x")


def main():
    rnd = random.Random(20261104)
    R = B.R_MOD
    cases = []
    for seed, full, partial, alpha in ((1, 8, 31, 17), (2, 8, 56, 5), (3, 4, 0, 3)):
        p = P.random_parameters(seed, full, partial, alpha)
        hashes = []
        for k in (1, 2, 3, 5):
            for row in ([0] * k, [R - 1] * k, [rnd.randrange(R) for _ in range(k)], [rnd.randrange(R) for _ in range(k)]):
                hashes.append({"inputs": [H(v) for v in row], "digest": H(P.crh(p, row))})
        trees = []
        for n in (2, 8):
            leaves = [rnd.randrange(R) for _ in range(n)]
            t = P.MerkleTree(p, leaves)
            sib, path, idx = t.generate_proof(n - 3 if n > 2 else 1)
            trees.append({"leaves": [H(v) for v in leaves], "leaf_nodes": [H(v) for v in t.leaf_nodes],
                          "non_leaf_nodes": [H(v) for v in t.non_leaf_nodes],
                          "path": {"leaf_index": idx, "leaf_sibling_hash": H(sib), "auth_path": [H(v) for v in path]}})
        cases.append({"full_rounds": full, "partial_rounds": partial, "alpha": alpha, "mds": [[H(v) for v in row] for row in p.mds],
                      "ark": [[H(v) for v in row] for row in p.ark], "hashes": hashes, "trees": trees})
    # hashed ElGamal on JubJub under the first parameter set (oracle/py/jubjub.py): key pair, scalar multiples, ciphertexts
    p0 = P.random_parameters(1, 8, 31, 17)
    sk = rnd.randrange(J.SUBGROUP_ORDER)
    pk = J.keygen(sk)
    pt = lambda q: [H(q[0]), H(q[1])]
    multiples = [{"k": H(k), "point": pt(J.mul(J.GEN, k))} for k in
                 (0, 1, 2, J.SUBGROUP_ORDER - 1, J.SUBGROUP_ORDER, (1 << 256) - 1, rnd.randrange(1 << 256), rnd.randrange(J.SUBGROUP_ORDER))]
    ciphertexts = []
    for chunks in (1, 4):
        r = rnd.randrange(J.SUBGROUP_ORDER)
        msg = [rnd.randrange(R) for _ in range(chunks)]
        c1, c2 = J.encrypt(p0, pk, msg, r)
        ciphertexts.append({"r": H(r), "msg": [H(m) for m in msg], "c1": pt(c1), "c2": [H(c) for c in c2]})
    elgamal = {"d": H(J.D), "generator": pt(J.GEN), "subgroup_order": H(J.SUBGROUP_ORDER), "sk": H(sk), "pk": pt(pk),
               "poseidon_case": 0, "multiples_of_generator": multiples, "ciphertexts": ciphertexts}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "poseidon_bls12_381_fr.json")
    with open(path, "w") as f:
        json.dump({"field": "BLS12-381 Fr", "modulus": H(R), "cases": cases, "elgamal": elgamal}, f, indent=0, sort_keys=True)
        f.write("\n")
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
