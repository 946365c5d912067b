"""Bulk Poseidon / ciphertext Merkle tree: csrc/poseidon.cu through the C ABI against the big-int
duplex-sponge restatement oracle/py/poseidon.py (parity unpinned, see its header).  The CPU part pins the oracle's own
structure: the sponge state machine against a hand-unrolled permutation, tree against paths."""
import json
import os
import random

import numpy as np
import pytest

import ctg
from oracle import ark
from oracle.py import bls12_381 as B
from oracle.py import poseidon as P

rnd = random.Random(77001)
R = B.R_MOD


def _params_words(p):
    mds = ark.fr_mont_from_ints([v for row in p.mds for v in row]).reshape(3, 3, 8)
    rc = ark.fr_mont_from_ints([v for row in p.ark for v in row]).reshape(len(p.ark), 3, 8)
    return mds, rc


def _ints(words):
    return ark.fr_ints_from_mont(np.ascontiguousarray(words).reshape(-1, 8))


# ----------------------------------------------------------------------------- CPU: the oracle itself
def _permute_unrolled(p, st):
    """The permutation written out a second way (explicit round kinds, column-sum MDS)."""
    st = list(st)
    kinds = ["F"] * (p.full_rounds // 2) + ["P"] * p.partial_rounds + ["F"] * (p.full_rounds // 2)
    for r, kind in enumerate(kinds):
        st = [(st[i] + p.ark[r][i]) % R for i in range(3)]
        for i in range(3 if kind == "F" else 1):
            st[i] = pow(st[i], p.alpha, R)
        new = [0, 0, 0]
        for j in range(3):
            for i in range(3):
                new[i] = (new[i] + p.mds[i][j] * st[j]) % R
        st = new
    return st


def test_oracle_sponge_against_unrolled_permutation():
    p = P.random_parameters(1)
    a, b, c, d, e = (rnd.randrange(R) for _ in range(5))
    assert P.crh(p, [a]) == _permute_unrolled(p, [a, 0, 0])[0]
    assert P.crh(p, [a, b]) == _permute_unrolled(p, [a, b, 0])[0]
    s1 = _permute_unrolled(p, [a, b, 0])
    assert P.crh(p, [a, b, c]) == _permute_unrolled(p, [(s1[0] + c) % R, s1[1], s1[2]])[0]
    s2 = _permute_unrolled(p, [(s1[0] + c) % R, (s1[1] + d) % R, s1[2]])
    assert P.crh(p, [a, b, c, d, e]) == _permute_unrolled(p, [(s2[0] + e) % R, s2[1], s2[2]])[0]


def test_oracle_sponge_modes():
    p = P.random_parameters(2, full_rounds=4, partial_rounds=3, alpha=5)
    xs = [rnd.randrange(R) for _ in range(5)]
    # absorbing in pieces equals absorbing at once
    s, t = P.PoseidonSponge(p), P.PoseidonSponge(p)
    s.absorb(xs)
    t.absorb(xs[:1]); t.absorb(xs[1:2]); t.absorb(xs[2:])
    assert s.squeeze(3) == t.squeeze(3)
    # squeezing in pieces equals squeezing at once; a squeeze of the whole rate then one more permutes once more
    s, t = P.PoseidonSponge(p), P.PoseidonSponge(p)
    s.absorb(xs); t.absorb(xs)
    whole = s.squeeze(5)
    assert t.squeeze(2) + t.squeeze(1) + t.squeeze(2) == whole
    # absorb after squeeze starts a new block
    s = P.PoseidonSponge(p)
    s.absorb(xs[:1])
    first = s.squeeze(1)[0]
    s.absorb(xs[1:2])
    s1 = _permute_unrolled(p, [xs[0], 0, 0])
    assert first == s1[0]
    s1 = _permute_unrolled(p, s1)
    s1[0] = (s1[0] + xs[1]) % R
    assert s.squeeze(1)[0] == _permute_unrolled(p, s1)[0]


@pytest.mark.parametrize("n", [2, 4, 16])
def test_oracle_merkle_paths_verify(n):
    p = P.random_parameters(3, full_rounds=4, partial_rounds=5, alpha=5)
    leaves = [rnd.randrange(R) for _ in range(n)]
    tree = P.MerkleTree(p, leaves)
    for i in range(n):
        proof = tree.generate_proof(i)
        assert len(proof[1]) == n.bit_length() - 2
        assert P.verify_path(p, tree.root(), leaves[i], proof)
        assert not P.verify_path(p, tree.root(), (leaves[i] + 1) % R, proof)
    with pytest.raises(AssertionError):
        P.MerkleTree(p, leaves[:n - 1] if n > 2 else leaves[:1])


@pytest.mark.parametrize("shape", [(8, 31, 17), (4, 3, 5), (2, 5, 2), (4, 0, 3)])
def test_oracle_c_matches_oracle_py(shape):
    """oracle/c (machine words, threads: the timed CPU baseline and the checker at large sizes) against the big-int restatement."""
    p = P.random_parameters(40 + shape[2], *shape)
    mds, rc = _params_words(p)
    for k in (1, 2, 3, 5):
        rows = [[rnd.randrange(R) for _ in range(k)] for _ in range(12)] + [[0] * k, [R - 1] * k]
        words = ark.fr_mont_from_ints([v for row in rows for v in row]).reshape(len(rows), k, 8)
        assert _ints(ark.poseidon_hash_batch(*shape, mds, rc, words)) == [P.crh(p, row) for row in rows]
    leaves = [rnd.randrange(R) for _ in range(32)]
    leaf_nodes, inner = ark.poseidon_merkle_tree(*shape, mds, rc, ark.fr_mont_from_ints(leaves))
    want = P.MerkleTree(p, leaves)
    assert _ints(leaf_nodes) == want.leaf_nodes and _ints(inner) == want.non_leaf_nodes


# ----------------------------------------------------------------------------- golden fixture (tests/golden/make_golden_poseidon.py)
I = lambda h: int(h, 16)


def _golden():
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "poseidon_bls12_381_fr.json")) as f:
        return json.load(f)


def _golden_params(case):
    return P.PoseidonParameters(case["full_rounds"], case["partial_rounds"], case["alpha"], [[I(v) for v in row] for row in case["mds"]],
                                [[I(v) for v in row] for row in case["ark"]])


def test_golden_generator_is_deterministic(tmp_path, monkeypatch):
    """Re-running tests/golden/make_golden_poseidon.py reproduces the committed fixture byte for byte."""
    import importlib.util
    here = os.path.dirname(os.path.abspath(__file__))
    spec = importlib.util.spec_from_file_location("make_golden_poseidon", os.path.join(here, "golden", "make_golden_poseidon.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    real_join = os.path.join
    target = str(tmp_path / "poseidon_bls12_381_fr.json")
    monkeypatch.setattr(mod.os.path, "join", lambda *a: target if a[-1] == "poseidon_bls12_381_fr.json" else real_join(*a))
    mod.main()
    with open(target, "rb") as f, open(real_join(here, "golden", "poseidon_bls12_381_fr.json"), "rb") as g:
        assert f.read() == g.read()


@pytest.mark.parametrize("ci", [0, 1, 2])
def test_oracle_reproduces_golden(ci):
    case = _golden()["cases"][ci]
    p = _golden_params(case)
    for h in case["hashes"]:
        assert P.crh(p, [I(v) for v in h["inputs"]]) == I(h["digest"])
    for t in case["trees"]:
        tree = P.MerkleTree(p, [I(v) for v in t["leaves"]])
        assert tree.leaf_nodes == [I(v) for v in t["leaf_nodes"]] and tree.non_leaf_nodes == [I(v) for v in t["non_leaf_nodes"]]
        sib, path, idx = tree.generate_proof(t["path"]["leaf_index"])
        assert sib == I(t["path"]["leaf_sibling_hash"]) and path == [I(v) for v in t["path"]["auth_path"]]
        assert P.verify_path(p, tree.root(), I(t["leaves"][idx]), (sib, path, idx))


# ----------------------------------------------------------------------------- GPU parity through the C ABI
@pytest.fixture(scope="module")
def ctx():
    c = ctg.Context(0)
    yield c
    c.close()


EDGE = [0, 1, 2, R - 1, R - 2, (R - 1) // 2, (1 << 255) % R, (1 << 64) - 1]


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(8, 31, 17), (8, 31, 5), (4, 0, 3), (2, 57, 2), (8, 56, 5)])
def test_hash_batch_matches_oracle(ctx, shape):
    full, partial, alpha = shape
    p = P.random_parameters(100 + alpha + partial, full, partial, alpha)
    pos = ctx.poseidon(full, partial, alpha, *_params_words(p))
    for k in (1, 2, 3, 4, 7):
        rows = [[EDGE[(i + j) % len(EDGE)] for j in range(k)] for i in range(len(EDGE))]
        rows += [[rnd.randrange(R) for _ in range(k)] for _ in range(150)]
        words = ark.fr_mont_from_ints([v for row in rows for v in row]).reshape(len(rows), k, 8)
        got = _ints(pos.hash_batch(words))
        assert got == [P.crh(p, row) for row in rows], (shape, k)
    assert pos.hash_batch(np.zeros((0, 2, 8), dtype=np.uint32)).shape == (0, 8)
    pos.free()


@pytest.mark.gpu
@pytest.mark.parametrize("ci", [0, 1, 2])
def test_gpu_reproduces_golden(ctx, ci):
    case = _golden()["cases"][ci]
    p = _golden_params(case)
    pos = ctx.poseidon(p.full_rounds, p.partial_rounds, p.alpha, *_params_words(p))
    for h in case["hashes"]:
        words = ark.fr_mont_from_ints([I(v) for v in h["inputs"]]).reshape(1, len(h["inputs"]), 8)
        assert _ints(pos.hash_batch(words)) == [I(h["digest"])]
    for t in case["trees"]:
        tree = pos.merkle_tree(ark.fr_mont_from_ints([I(v) for v in t["leaves"]]))
        assert _ints(tree.leaf_nodes) == [I(v) for v in t["leaf_nodes"]]
        assert _ints(tree.non_leaf_nodes) == [I(v) for v in t["non_leaf_nodes"]]
        sib, path, idx = tree.generate_proof(t["path"]["leaf_index"])
        assert _ints(sib) == [I(t["path"]["leaf_sibling_hash"])]
        assert _ints(np.array(path).reshape(-1, 8)) == [I(v) for v in t["path"]["auth_path"]]
    pos.free()


@pytest.mark.gpu
@pytest.mark.parametrize("n", [2, 4, 8, 256, 1024])
def test_merkle_tree_matches_oracle(ctx, n):
    p = P.random_parameters(5)
    pos = ctx.poseidon(p.full_rounds, p.partial_rounds, p.alpha, *_params_words(p))
    leaves = [rnd.randrange(R) for _ in range(n)]
    tree = pos.merkle_tree(ark.fr_mont_from_ints(leaves))
    want = P.MerkleTree(p, leaves)
    assert _ints(tree.leaf_nodes) == want.leaf_nodes
    assert _ints(tree.non_leaf_nodes) == want.non_leaf_nodes
    for i in (0, 1, n // 2, n - 1):
        sib, path, idx = tree.generate_proof(i)
        wsib, wpath, _ = want.generate_proof(i)
        assert _ints(sib) == [wsib] and _ints(np.array(path).reshape(-1, 8)) == wpath and idx == i
        assert P.verify_path(p, _ints(tree.root())[0], leaves[i], (wsib, wpath, i))
    pos.free()


@pytest.mark.gpu
def test_merkle_tree_2_16_leaves_node_by_node(ctx):
    """Both kernels (one thread per hash on the wide levels, four lanes per hash on the narrow ones) against oracle/c on every node."""
    n = 1 << 16
    p = P.random_parameters(8)
    mds, rc = _params_words(p)
    pos = ctx.poseidon(p.full_rounds, p.partial_rounds, p.alpha, mds, rc)
    leaves = ark.random_fr_mont(n, 10)
    tree = pos.merkle_tree(leaves)
    leaf_nodes, inner = ark.poseidon_merkle_tree(p.full_rounds, p.partial_rounds, p.alpha, mds, rc, leaves)
    assert np.array_equal(tree.leaf_nodes, leaf_nodes) and np.array_equal(tree.non_leaf_nodes, inner)
    pos.free()


@pytest.mark.gpu
def test_merkle_tree_full_size_properties(ctx):
    """2^20 leaves (the ciphertext of a BASELINE-sized PoE): paths from the device tree verify under the oracle's Path::verify,
    a changed leaf changes the root, and the tree equals the one rebuilt level by level with the bulk two-to-one hash."""
    n = 1 << 20
    p = P.random_parameters(6)
    pos = ctx.poseidon(p.full_rounds, p.partial_rounds, p.alpha, *_params_words(p))
    leaves = ark.random_fr_mont(n, 9)
    tree = pos.merkle_tree(leaves)
    level = pos.hash_batch(leaves.reshape(n, 1, 8))
    assert np.array_equal(level, tree.leaf_nodes)
    width = n // 2
    while width >= 1:
        level = pos.hash_batch(level.reshape(width, 2, 8))
        assert np.array_equal(level, tree.non_leaf_nodes[width - 1:2 * width - 1])
        width //= 2
    root = _ints(tree.root())[0]
    for i in (0, 12345, n - 1):
        sib, path, _ = tree.generate_proof(i)
        proof = (_ints(sib)[0], _ints(np.array(path).reshape(-1, 8)), i)
        assert P.verify_path(p, root, _ints(leaves[i])[0], proof)
    changed = leaves.copy()
    changed[777] = ark.fr_mont_from_ints([123456789])[0]
    assert not np.array_equal(pos.merkle_tree(changed).root(), tree.root())
    pos.free()


@pytest.mark.gpu
def test_poseidon_rejects_bad_arguments(ctx):
    p = P.random_parameters(7)
    mds, rc = _params_words(p)
    with pytest.raises(ctg.CtgError):
        ctx.poseidon(7, 32, 17, mds, rc)            # odd number of full rounds
    with pytest.raises(ctg.CtgError):
        ctx.poseidon(8, 31, 1, mds, rc)             # alpha < 2
    pos = ctx.poseidon(8, 31, 17, mds, rc)
    for n in (0, 1, 3, 6):
        with pytest.raises(ctg.CtgError):
            pos.merkle_tree(np.zeros((n, 8), dtype=np.uint32))
    with pytest.raises(ctg.CtgError):
        pos.hash_batch(np.zeros((4, 0, 8), dtype=np.uint32))
    pos.free()
