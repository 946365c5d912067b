"""The PoPRP circuit of the reference (SampleEntries, circuits-ark/src/sample_entries.rs:136-232) as built by ctg/circuits.py.

CPU: the R1CS is satisfied by an honest witness made with the big-int oracle (oracle/py: JubJub hashed ElGamal + Poseidon Merkle tree),
has the instance layout [1, sample_entry, merkle_root] and the size SURVEY Appendix B estimates, and is NOT satisfied by a wrong
sample, a wrong leaf index or a wrong secret key.
GPU (the f4 -> prover chain): ctg_elgamal_encrypt_batch -> ctg_poseidon_merkle_tree -> Merkle path by index
arithmetic -> witness -> ctg_generate_parameters -> ctg_prove, every stage against the oracle and the proof checked by the big-int
pairing verifier against the public inputs (sample, root) the GPU itself produced.
"""
import random

import numpy as np
import pytest

import ctg
from ctg import circuits as CI
from ctg import synthetic as S
from oracle import ark
from oracle.py import bls12_381 as B
from oracle.py import jubjub as J
from oracle.py import poseidon as P
from tests import util

rnd = random.Random(4242)


def _reference_poseidon():
    fr, pr, alpha, mds_w, ark_w = ctg.poseidon_reference_params()
    pc = CI.PoseidonConstants.from_words(fr, pr, alpha, mds_w, ark_w)
    return pc, P.PoseidonParameters(fr, pr, alpha, pc.mds, pc.ark), (fr, pr, alpha, mds_w, ark_w)


def _satisfied(r1cs, z):
    vals = {k: S.fr_from_mont_words(r1cs[k]["val"]) for k in "abc"}
    for i in range(r1cs["num_constraints"]):
        d = []
        for k in "abc":
            lo, hi = int(r1cs[k]["row_ptr"][i]), int(r1cs[k]["row_ptr"][i + 1])
            d.append(sum(vals[k][j] * z[int(r1cs[k]["col"][j])] for j in range(lo, hi)) % B.R_MOD)
        if d[0] * d[1] % B.R_MOD != d[2]:
            return False
    return True


def _statement(params, n_leaves, idx, seed):
    r = random.Random(seed)
    sk = r.randrange(J.SUBGROUP_ORDER)
    msgs = [r.randrange(256) for _ in range(n_leaves)]
    c1, c2 = J.encrypt(params, J.keygen(sk), msgs, r.randrange(J.SUBGROUP_ORDER))
    tree = P.MerkleTree(params, c2)
    sib, path, _ = tree.generate_proof(idx)
    return sk, msgs, c1, c2, tree, sib, path


@pytest.mark.parametrize("n_leaves,idx", [(2, 1), (16, 11), (64, 0)])
def test_sample_entries_circuit_is_satisfied_and_sized(n_leaves, idx):
    pc, params, _ = _reference_poseidon()
    sk, msgs, c1, c2, tree, sib, path = _statement(params, n_leaves, idx, 7 + n_leaves)
    cs = CI.sample_entries_circuit(pc, c1, c2[idx], sib, path, idx, sk, msgs[idx], tree.root())
    r1cs, z = cs.r1cs(), cs.assignment()
    assert r1cs["num_instance"] == 3 and z[:3] == [1, msgs[idx], tree.root()]
    depth = n_leaves.bit_length() - 1
    # one scalar multiplication (256 x (add + select + double)), one sponge, depth + 1 Poseidon CRH, the selectors
    assert r1cs["num_constraints"] == 3 + 275 + depth * (2 + 275) + 1 + 256 + 256 * (7 + 2) + 255 * 7 + 275 + 1
    assert 4500 + 250 * depth < r1cs["num_constraints"] < 5500 + 300 * depth        # ~5 k + 300 log2 n
    assert _satisfied(r1cs, z)
    # the assignment-only mode (further witnesses of the same circuit) produces the same z
    assert CI.sample_entries_circuit(pc, c1, c2[idx], sib, path, idx, sk, msgs[idx], tree.root(), record=False).assignment() == z


def test_sample_entries_circuit_rejects_false_statements():
    pc, params, _ = _reference_poseidon()
    idx = 5
    sk, msgs, c1, c2, tree, sib, path = _statement(params, 8, idx, 99)
    for kw in ({"sample_entry": msgs[idx] + 1}, {"leaf_index": idx ^ 1}, {"sk": sk + 1}, {"merkle_root": tree.root() + 1}):
        args = dict(c1=c1, leaf=c2[idx], leaf_sibling_hash=sib, auth_path=path, leaf_index=idx, sk=sk, sample_entry=msgs[idx], merkle_root=tree.root())
        args.update(kw)
        cs = CI.sample_entries_circuit(pc, **args)
        assert not _satisfied(cs.r1cs(), cs.assignment()), kw


@pytest.mark.parametrize("n_msg,n_slots", [(1, 1), (3, 5)])
def test_encrypt_circuit_is_satisfied_sized_and_rejects(n_msg, n_slots):
    """The reference's Proof-of-Encryption circuit (EncryptCircuit, circuits-ark/src/encryption.rs:219-317) as built by ctg/circuits.py:
    instance [1, c1.x, c1.y, c2[0..n)], satisfied by an honest encryption made with the big-int oracle (empty slots included), not by a
    tampered ciphertext, another randomness or another plaintext."""
    pc, params, _ = _reference_poseidon()
    r = random.Random(50 + n_slots)
    sk, rr = r.randrange(J.SUBGROUP_ORDER), r.randrange(J.SUBGROUP_ORDER)
    pk = J.keygen(sk)
    msgs = [r.randrange(1, 256) for _ in range(n_msg)]
    c1, c2 = J.encrypt(params, pk, msgs, rr)
    cs = CI.encrypt_circuit(pc, pk, rr, msgs, c1, c2, n_slots=n_slots)
    r1cs, z = cs.r1cs(), cs.assignment()
    assert r1cs["num_instance"] == 3 + n_slots and z[:3 + n_msg] == [1, c1[0], c1[1]] + c2 and z[3 + n_msg:3 + n_slots] == [0] * (n_slots - n_msg)
    # 256 bit checks, the on-curve check, a variable-base and a fixed-base scalar multiplication, one sponge, c1 == c1', 3 per slot
    assert r1cs["num_constraints"] == 256 + 3 + (256 * 9 + 255 * 7) + 256 * 5 + 275 + 2 + 3 * n_slots
    assert _satisfied(r1cs, z)
    assert CI.encrypt_circuit(pc, pk, rr, msgs, c1, c2, n_slots=n_slots, record=False).assignment() == z
    bad = list(c2)
    bad[0] = (bad[0] + 1) % B.R_MOD
    for kw in ({"c2": bad}, {"r": rr + 1}, {"msg": [msgs[0] + 1] + msgs[1:]}, {"c1": J.add(c1, J.GEN)}):
        args = dict(pk=pk, r=rr, msg=msgs, c1=c1, c2=c2)
        args.update(kw)
        cs2 = CI.encrypt_circuit(pc, n_slots=n_slots, **args)
        assert not _satisfied(cs2.r1cs(), cs2.assignment()), list(kw)


@pytest.mark.gpu
def test_encrypt_then_prove_encryption_on_gpu():
    """BASELINE.json configs[0] with the reference's statement instead of a shape: encrypt on the device, prove the PoE circuit on the
    device (one lone proof and a batch), verify with the big-int pairing check against the ciphertext as public input."""
    _, params, words = _reference_poseidon()
    ctx = ctg.Context(0)
    r1cs, zb, publics = CI.encrypt_workload(ctx, words, 2, 3, seed=77)
    assert S.domain_size(r1cs) == 8192 and r1cs["num_instance"] == 5
    td_words, _ = S.trapdoor()
    dp = ctx.generate_parameters(r1cs, td_words)
    key = dp.download()
    dpk = dp.to_device_pk()
    dp.free()
    dr = ctx.upload_r1cs(r1cs)
    rs = ark.fr_mont_from_ints([rnd.randrange(B.R_MOD) for _ in range(6)]).reshape(3, 2, 8)
    lone = ctx.prove(dpk, dr, zb[0], rs[0][0], rs[0][1])
    assert all(np.array_equal(g, w) for g, w in zip(lone, ark.create_proof(key, r1cs, zb[0], rs[0][0], rs[0][1])))
    out = ctx.prove_batch(dpk, dr, zb, rs)
    assert np.array_equal(out[0], np.concatenate(lone))
    for k in range(3):
        pk_ = (out[k][:24], out[k][24:72], out[k][72:])
        assert util.groth16_verify_words(key, pk_, publics[k])
    wrong = list(publics[1])
    wrong[3] = (wrong[3] + 1) % B.R_MOD                                                  # another ciphertext element
    assert not util.groth16_verify_words(key, (out[1][:24], out[1][24:72], out[1][72:]), wrong)
    dpk.free(); dr.free(); ctx.close()


@pytest.mark.gpu
def test_encrypt_merkle_path_prove_chain_on_gpu():
    pc, params, words = _reference_poseidon()
    ctx = ctg.Context(0)
    pos = ctx.poseidon(*words)
    n_leaves, idx = 32, 19
    sk = rnd.randrange(J.SUBGROUP_ORDER)
    rr = rnd.randrange(J.SUBGROUP_ORDER)
    msgs = [rnd.randrange(256) for _ in range(n_leaves)]
    d_w = ark.fr_mont_from_ints([J.D])[0]
    gen_w = ark.fr_mont_from_ints(list(J.GEN))
    pk = J.keygen(sk)
    # 1. bulk hashed ElGamal on the device (one randomness, n_leaves plaintext elements), against the oracle
    c1_w, c2_w = ctx.elgamal_encrypt_batch(pos, d_w, gen_w, ark.fr_mont_from_ints(list(pk)), ark.ints_to_words([rr], 8),
                                           ark.fr_mont_from_ints(msgs).reshape(1, n_leaves, 8))
    c1 = tuple(ark.fr_ints_from_mont(c1_w[0]))
    c2 = ark.fr_ints_from_mont(c2_w[0])
    want_c1, want_c2 = J.encrypt(params, pk, msgs, rr)
    assert c1 == want_c1 and c2 == want_c2
    # 2. Merkle tree of the ciphertext on the device, path by index arithmetic
    tree = pos.merkle_tree(c2_w[0])
    otree = P.MerkleTree(params, c2)
    root = ark.fr_ints_from_mont(tree.root().reshape(1, 8))[0]
    assert root == otree.root()
    sib_w, path_w, _ = tree.generate_proof(idx)
    sib = ark.fr_ints_from_mont(sib_w.reshape(1, 8))[0]
    path = ark.fr_ints_from_mont(np.stack(path_w)) if path_w else []
    assert (sib, path, idx) == otree.generate_proof(idx)
    assert P.verify_path(params, root, c2[idx], (sib, path, idx))
    # 3. the PoPRP statement about it: circuit + witness, key generated on the device, proof
    cs = CI.sample_entries_circuit(pc, c1, c2[idx], sib, path, idx, sk, msgs[idx], root)
    r1cs, z = cs.r1cs(), cs.assignment()
    zm = ark.fr_mont_from_ints(z)
    td_words, td = S.trapdoor()
    dp = ctx.generate_parameters(r1cs, td_words)
    key = dp.download()
    dpk = dp.to_device_pk()
    dp.free()
    dr = ctx.upload_r1cs(r1cs)
    r, s = ark.fr_mont_from_ints([rnd.randrange(B.R_MOD)])[0], ark.fr_mont_from_ints([rnd.randrange(B.R_MOD)])[0]
    proof = ctx.prove(dpk, dr, zm, r, s)
    want = ark.create_proof(key, r1cs, zm, r, s)                       # the arkworks algorithm on the host, same key
    assert all(np.array_equal(g, w) for g, w in zip(proof, want))
    # 4. the verifier accepts it for the public inputs (sample, root) and rejects it for another sample
    assert util.groth16_verify_words(key, proof, [msgs[idx], root])
    assert not util.groth16_verify_words(key, proof, [(msgs[idx] + 1) % B.R_MOD, root])
    # a batch of two more samples of the same ciphertext under the same key (assignment-only synthesis)
    zs, rs = [], []
    for j in (0, 31):
        sj, pj, _ = tree.generate_proof(j)
        cj = CI.sample_entries_circuit(pc, c1, c2[j], ark.fr_ints_from_mont(sj.reshape(1, 8))[0], ark.fr_ints_from_mont(np.stack(pj)), j, sk, msgs[j], root, record=False)
        zs.append(ark.fr_mont_from_ints(cj.assignment()))
        rs.append(np.stack([r, s]))
    out = ctx.prove_batch(dpk, dr, np.stack(zs), np.stack(rs))
    for k, j in enumerate((0, 31)):
        pj = (out[k][:24], out[k][24:72], out[k][72:])
        assert util.groth16_verify_words(key, pj, [msgs[j], root])
    dpk.free(); dr.free(); pos.free(); ctx.close()
