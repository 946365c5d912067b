"""The only bytes the reference holds on the witness-side crypto path, as fixtures:

  * the Poseidon table P1 of circuits-ark/src/parameters.rs:12-229, parsed into BLS12-381 Fr the way Fr::from_str does
    (circuits-ark/src/poseidon.rs:18-33): tests/golden/reference_constants.json and the shipped loader
    ctg.poseidon_reference_params();
  * the JubJub key strings ALICE_PK / ALICE_SK of circuits-ark/src/utils.rs:114-115, which the reference's own tests
    (:117-127) decode with ark_from_bytes -- a reference-held pin for the curve coefficient d, the prime subgroup order and the
    compressed twisted-Edwards encoding (sign flag polarity included).

CPU part: loader, oracle and (when /root/reference is present, i.e. in the build container) a fresh extraction agree; the
oracle reproduces the frozen vectors computed under these parameters.  GPU part: the CUDA sponge / tree / ElGamal kernels
under the reference's parameters against those vectors, and r * ALICE_PK = O on the device.
"""
import importlib.util
import json
import os
import random

import numpy as np
import pytest

import ctg
from oracle import ark
from oracle.py import bls12_381 as B
from oracle.py import jubjub as J
from oracle.py import poseidon as P

HERE = os.path.dirname(os.path.abspath(__file__))
R = B.R_MOD
I = lambda h: int(h, 16)
rnd = random.Random(5150)


def _const():
    with open(os.path.join(HERE, "golden", "reference_constants.json")) as f:
        return json.load(f)


def _vectors():
    with open(os.path.join(HERE, "golden", "poseidon_reference_vectors.json")) as f:
        return json.load(f)


def _params():
    g = _const()
    return P.PoseidonParameters(g["full_rounds"], g["partial_rounds"], g["alpha"], [[I(v) for v in row] for row in g["mds_hex"]],
                                [[I(v) for v in row] for row in g["ark_hex"]])


def _ints(words):
    return ark.fr_ints_from_mont(np.ascontiguousarray(words).reshape(-1, 8))


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(HERE, "golden", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


# ----------------------------------------------------------------------------- CPU
def test_loader_matches_frozen_table():
    """ctg.poseidon_reference_params() (decimal strings reduced digit by digit, as from_str) == the frozen reduced table."""
    g = _const()
    full, partial, alpha, mds, rc = ctg.poseidon_reference_params()
    assert (full, partial, alpha) == (g["full_rounds"], g["partial_rounds"], g["alpha"]) == (8, 31, 17)
    assert mds.shape == (3, 3, 8) and rc.shape == (39, 3, 8)
    assert _ints(mds) == [I(v) for row in g["mds_hex"] for v in row]
    assert _ints(rc) == [I(v) for row in g["ark_hex"] for v in row]


def test_table_exceeds_modulus_and_is_reduced():
    """Every decimal of the table is a 370+-bit number (the table was generated for BLS12-377's base field), so from_str's
    in-field accumulation matters: digit-by-digit reduction == int(dec) mod r, and the plain integer is never a valid Fr."""
    with open(os.path.join(os.path.dirname(ctg.__file__), "data", "poseidon_bls12_381_rate2.json")) as f:
        d = json.load(f)
    decs = [v for row in d["mds"] + d["ark"] for v in row]
    assert len(decs) == 9 + 39 * 3
    assert all(int(v) >= R for v in decs)
    assert [ctg._fr_from_str(v) for v in decs] == [int(v) % R for v in decs]
    for bad in ("", "01", "12a", "-3"):
        with pytest.raises(ctg.CtgError):
            ctg._fr_from_str(bad)


@pytest.mark.skipif(not os.path.exists("/root/reference/circuits-ark/src/parameters.rs"), reason="the reference tree exists only in the build container")
def test_fresh_extraction_matches_committed_files():
    mod = _load("make_reference_constants")
    ark_t, mds_t, scal, keys = mod.extract()
    g = _const()
    assert [[format(mod.from_str(v), "x") for v in row] for row in ark_t] == g["ark_hex"]
    assert [[format(mod.from_str(v), "x") for v in row] for row in mds_t] == g["mds_hex"]
    assert keys == {"ALICE_SK": g["ALICE_SK"], "ALICE_PK": g["ALICE_PK"]}
    with open(os.path.join(os.path.dirname(ctg.__file__), "data", "poseidon_bls12_381_rate2.json")) as f:
        d = json.load(f)
    assert d["ark"] == ark_t and d["mds"] == mds_t and all(d[k] == v for k, v in scal.items())


def test_vector_generator_is_deterministic(tmp_path, monkeypatch):
    mod = _load("make_golden_reference_vectors")
    real_join = os.path.join
    target = str(tmp_path / "v.json")
    monkeypatch.setattr(mod.os.path, "join", lambda *a: target if a[-1] == "poseidon_reference_vectors.json" else real_join(*a))
    mod.main()
    with open(target, "rb") as f, open(real_join(HERE, "golden", "poseidon_reference_vectors.json"), "rb") as g:
        assert f.read() == g.read()


def test_alice_keys_decode_as_the_reference_tests_require():
    """circuits-ark/src/utils.rs:117-127: ark_from_bytes::<JubJub>(ALICE_PK) and ark_from_bytes::<Fr>(ALICE_SK) succeed.
    For the oracle's restatement of ark-ec's compressed twisted-Edwards form that means: x is canonical, (1 + x^2)/(1 - d x^2) is
    a square (pins d), the root picked by the sign flag lies in the prime-order subgroup (pins the subgroup order AND the flag
    polarity: the other root is outside the subgroup), and re-encoding gives the same 32 bytes."""
    g = _const()
    raw = bytes.fromhex(g["ALICE_PK"])
    pk = J.point_decompress(raw)
    assert pk is not None and J.on_curve(pk)
    assert J.mul(pk, J.SUBGROUP_ORDER) == J.IDENTITY and pk != J.IDENTITY
    assert J.point_compress(pk) == raw
    flipped = bytearray(raw); flipped[31] ^= 0x80
    assert J.point_decompress(bytes(flipped)) is None                       # arkworks: InvalidData (not in the subgroup)
    other = J.point_decompress(bytes(flipped), check_subgroup=False)
    assert other == (pk[0], R - pk[1]) and J.on_curve(other)
    sk = J.scalar_from_bytes(bytes.fromhex(g["ALICE_SK"]))
    assert sk is not None and sk < R and sk < J.SUBGROUP_ORDER
    # with another d the x coordinate is (overwhelmingly likely) not on the curve or not in the subgroup
    assert J.point_decompress(raw, d=(J.D + 1) % R) is None
    # the identity and codec round trips on other points
    assert J.point_decompress(J.point_compress(J.IDENTITY)) == J.IDENTITY
    for _ in range(5):
        q = J.mul(J.GEN, rnd.randrange(J.SUBGROUP_ORDER))
        assert J.point_decompress(J.point_compress(q)) == q
    v = _vectors()
    assert [I(c) for c in v["alice_pk"]] == list(pk) and I(v["alice_sk"]) == sk


def test_oracle_reproduces_reference_parameter_vectors():
    p, v = _params(), _vectors()
    for h in v["hashes"]:
        assert P.crh(p, [I(x) for x in h["inputs"]]) == I(h["digest"])
    for t in v["trees"]:
        tree = P.MerkleTree(p, [I(x) for x in t["leaves"]])
        assert tree.leaf_nodes == [I(x) for x in t["leaf_nodes"]] and tree.non_leaf_nodes == [I(x) for x in t["non_leaf_nodes"]]
        sib, path, idx = tree.generate_proof(t["path"]["leaf_index"])
        assert sib == I(t["path"]["leaf_sibling_hash"]) and path == [I(x) for x in t["path"]["auth_path"]]
        assert P.verify_path(p, tree.root(), I(t["leaves"][idx]), (sib, path, idx))
    pk = tuple(I(c) for c in v["alice_pk"])
    for c in v["ciphertexts_to_alice_pk"]:
        c1, c2 = J.encrypt(p, pk, [I(m) for m in c["msg"]], I(c["r"]))
        assert list(c1) == [I(x) for x in c["c1"]] and c2 == [I(x) for x in c["c2"]]


def test_oracle_c_under_reference_parameters():
    """oracle/c (the timed CPU leg of bench.py's Poseidon rows) under the reference's table."""
    p, v = _params(), _vectors()
    full, partial, alpha, mds, rc = ctg.poseidon_reference_params()
    for h in v["hashes"]:
        words = ark.fr_mont_from_ints([I(x) for x in h["inputs"]]).reshape(1, len(h["inputs"]), 8)
        assert _ints(ark.poseidon_hash_batch(full, partial, alpha, mds, rc, words)) == [I(h["digest"])]
    t = v["trees"][-1]
    leaf_nodes, inner = ark.poseidon_merkle_tree(full, partial, alpha, mds, rc, ark.fr_mont_from_ints([I(x) for x in t["leaves"]]))
    assert _ints(leaf_nodes) == [I(x) for x in t["leaf_nodes"]] and _ints(inner) == [I(x) for x in t["non_leaf_nodes"]]


# ----------------------------------------------------------------------------- GPU
@pytest.fixture(scope="module")
def ctx():
    c = ctg.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
def test_gpu_sponge_and_tree_under_reference_parameters(ctx):
    v = _vectors()
    pos = ctx.poseidon(*ctg.poseidon_reference_params())
    for h in v["hashes"]:
        words = ark.fr_mont_from_ints([I(x) for x in h["inputs"]]).reshape(1, len(h["inputs"]), 8)
        assert _ints(pos.hash_batch(words)) == [I(h["digest"])]
    for t in v["trees"]:
        tree = pos.merkle_tree(ark.fr_mont_from_ints([I(x) for x in t["leaves"]]))
        assert _ints(tree.leaf_nodes) == [I(x) for x in t["leaf_nodes"]]
        assert _ints(tree.non_leaf_nodes) == [I(x) for x in t["non_leaf_nodes"]]
        sib, path, idx = tree.generate_proof(t["path"]["leaf_index"])
        assert _ints(sib) == [I(t["path"]["leaf_sibling_hash"])]
        assert _ints(np.array(path).reshape(-1, 8)) == [I(x) for x in t["path"]["auth_path"]]
    # a wide batch (one thread per hash) and a 2^12-leaf tree (both kernels) against oracle/c under the same table
    full, partial, alpha, mds, rc = ctg.poseidon_reference_params()
    rows = ark.random_fr_mont(3000 * 2, 31).reshape(3000, 2, 8)
    assert np.array_equal(pos.hash_batch(rows), ark.poseidon_hash_batch(full, partial, alpha, mds, rc, rows))
    leaves = ark.random_fr_mont(1 << 12, 32)
    tree = pos.merkle_tree(leaves)
    leaf_nodes, inner = ark.poseidon_merkle_tree(full, partial, alpha, mds, rc, leaves)
    assert np.array_equal(tree.leaf_nodes, leaf_nodes) and np.array_equal(tree.non_leaf_nodes, inner)
    pos.free()


@pytest.mark.gpu
def test_gpu_alice_pk_is_in_the_prime_subgroup_and_encrypts(ctx):
    """Device JubJub arithmetic on the reference-held key: order * ALICE_PK = (0, 1), 8 * order * (the other root) = (0, 1) but
    order * (the other root) != (0, 1); hashed-ElGamal ciphertexts to ALICE_PK under the reference's Poseidon table equal the
    frozen vectors."""
    v = _vectors()
    d = ark.fr_mont_from_ints([J.D])[0]
    pk = tuple(I(c) for c in v["alice_pk"])
    other = (pk[0], R - pk[1])
    pts = ark.fr_mont_from_ints([pk[0], pk[1], other[0], other[1], other[0], other[1]]).reshape(3, 2, 8)
    ks = ark.ints_to_words([J.SUBGROUP_ORDER, J.SUBGROUP_ORDER, 8 * J.SUBGROUP_ORDER], 8)
    out = ctx.te_mul_batch(d, pts, ks)
    got = [tuple(_ints(o)) for o in out]
    assert got[0] == J.IDENTITY and got[1] != J.IDENTITY and got[2] == J.IDENTITY
    pos = ctx.poseidon(*ctg.poseidon_reference_params())
    gen = ark.fr_mont_from_ints([J.GEN_X, J.GEN_Y]).reshape(2, 8)
    pkw = ark.fr_mont_from_ints(list(pk)).reshape(2, 8)
    for c in v["ciphertexts_to_alice_pk"]:
        msg = ark.fr_mont_from_ints([I(m) for m in c["msg"]]).reshape(1, -1, 8)
        c1, c2 = ctx.elgamal_encrypt_batch(pos, d, gen, pkw, ark.ints_to_words([I(c["r"])], 8), msg)
        assert _ints(c1[0]) == [I(x) for x in c["c1"]] and _ints(c2[0]) == [I(x) for x in c["c2"]]
    # sk * G on the device equals the oracle's (the reference never asserts ALICE_PK = ALICE_SK * G, and it is not)
    skg = ctx.te_mul_batch(d, gen.reshape(1, 2, 8), ark.ints_to_words([I(v["alice_sk"])], 8))
    assert _ints(skg[0]) == [I(x) for x in v["alice_sk_times_generator"]]
    pos.free()
