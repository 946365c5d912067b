"""Bulk JubJub scalar multiplication and hashed-ElGamal encryption: csrc/edwards.cu (+ poseidon.cu) through
the C ABI against the big-int restatement oracle/py/jubjub.py (affine unified law; the kernel uses extended coordinates)."""
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

rnd = random.Random(991)
R = B.R_MOD


def _pts(points):
    return ark.fr_mont_from_ints([c for p in points for c in p]).reshape(len(points), 2, 8)


def _scalars(ks):
    return ark.ints_to_words(ks, 8)


def _pts_back(words):
    v = ark.fr_ints_from_mont(np.ascontiguousarray(words).reshape(-1, 8))
    return [(v[2 * i], v[2 * i + 1]) for i in range(len(v) // 2)]


# ----------------------------------------------------------------------------- CPU: the oracle itself
def test_oracle_jubjub_constants_and_group_law():
    J.selfcheck()
    a, b = rnd.randrange(J.SUBGROUP_ORDER), rnd.randrange(J.SUBGROUP_ORDER)
    pa, pb = J.mul(J.GEN, a), J.mul(J.GEN, b)
    assert J.on_curve(pa) and J.add(pa, pb) == J.mul(J.GEN, (a + b) % J.SUBGROUP_ORDER)
    assert J.mul(pa, b) == J.mul(pb, a) == J.mul(J.GEN, a * b % J.SUBGROUP_ORDER)
    assert J.add(pa, J.IDENTITY) == pa and J.add(pa, (R - pa[0], pa[1])) == J.IDENTITY
    assert J.mul(pa, 0) == J.IDENTITY
    found = [J.point_from_x(x) for x in range(2, 40)]
    assert sum(p is not None for p in found) > 5 and all(J.on_curve(p) for p in found if p)


def test_oracle_elgamal_round_trip():
    p = P.random_parameters(21)
    sk = rnd.randrange(J.SUBGROUP_ORDER)
    pk = J.keygen(sk)
    msg = [rnd.randrange(R) for _ in range(5)]
    cipher = J.encrypt(p, pk, msg, rnd.randrange(J.SUBGROUP_ORDER))
    assert J.decrypt(p, cipher, sk) == msg
    assert J.decrypt(p, cipher, (sk + 1) % J.SUBGROUP_ORDER) != msg


# ----------------------------------------------------------------------------- golden fixture (tests/golden/make_golden_poseidon.py)
I = lambda h: int(h, 16)


def _golden():
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "poseidon_bls12_381_fr.json")) as f:
        g = json.load(f)
    case = g["cases"][g["elgamal"]["poseidon_case"]]
    params = P.PoseidonParameters(case["full_rounds"], case["partial_rounds"], case["alpha"], [[I(v) for v in row] for row in case["mds"]],
                                  [[I(v) for v in row] for row in case["ark"]])
    return g["elgamal"], params


def _pt(h):
    return (I(h[0]), I(h[1]))


def test_oracle_reproduces_elgamal_golden():
    e, params = _golden()
    assert I(e["d"]) == J.D and _pt(e["generator"]) == J.GEN and I(e["subgroup_order"]) == J.SUBGROUP_ORDER
    assert J.keygen(I(e["sk"])) == _pt(e["pk"])
    for m in e["multiples_of_generator"]:
        assert J.mul(J.GEN, I(m["k"])) == _pt(m["point"])
    for c in e["ciphertexts"]:
        c1, c2 = J.encrypt(params, _pt(e["pk"]), [I(v) for v in c["msg"]], I(c["r"]))
        assert c1 == _pt(c["c1"]) and c2 == [I(v) for v in c["c2"]]
        assert J.decrypt(params, (c1, c2), I(e["sk"])) == [I(v) for v in c["msg"]]


# ----------------------------------------------------------------------------- GPU parity through the C ABI
@pytest.fixture(scope="module")
def ctx():
    c = ctg.Context(0)
    yield c
    c.close()


D_WORDS = None


def _d():
    return ark.fr_mont_from_ints([J.D])[0]


@pytest.mark.gpu
def test_te_mul_shared_base_matches_oracle(ctx):
    ks = [0, 1, 2, 3, 8, J.SUBGROUP_ORDER - 1, J.SUBGROUP_ORDER, J.SUBGROUP_ORDER + 1, (1 << 256) - 1, 1 << 255, (1 << 32) - 1, 1 << 32]
    ks += [rnd.randrange(1 << 256) for _ in range(40)] + [rnd.randrange(J.SUBGROUP_ORDER) for _ in range(40)]
    got = _pts_back(ctx.te_mul_batch(_d(), _pts([J.GEN]), _scalars(ks)))
    assert got == [J.mul(J.GEN, k) for k in ks]
    assert got[0] == J.IDENTITY and got[6] == J.IDENTITY


@pytest.mark.gpu
def test_te_mul_per_item_bases_and_small_order_points(ctx):
    bases = [J.mul(J.GEN, rnd.randrange(1, J.SUBGROUP_ORDER)) for _ in range(30)]
    bases += [J.IDENTITY, (0, R - 1)]                                    # identity and the point of order 2
    bases += [p for p in (J.point_from_x(x) for x in range(2, 30)) if p]  # points outside the prime subgroup too
    ks = [rnd.randrange(1 << 253) for _ in bases]
    got = _pts_back(ctx.te_mul_batch(_d(), _pts(bases), _scalars(ks)))
    assert got == [J.mul(b, k) for b, k in zip(bases, ks)]
    assert ctx.te_mul_batch(_d(), _pts([J.GEN]), np.zeros((0, 8), dtype=np.uint32)).shape == (0, 2, 8)
    with pytest.raises(ctg.CtgError):
        ctx.te_mul_batch(_d(), _pts(bases[:3]), _scalars(ks[:5]))     # neither one base nor one per scalar


@pytest.mark.gpu
def test_te_mul_window_table_path_equals_double_and_add(ctx):
    """From 4096 scalars on one shared base the library switches to a window table (no doublings): same points as the plain
    double-and-add kernel (forced by passing the base once per scalar), on edge scalars and bases outside the prime subgroup."""
    edge = [0, 1, 255, 256, 257, (1 << 256) - 1, 1 << 255, J.SUBGROUP_ORDER, J.SUBGROUP_ORDER - 1, 0xFF << 248, 0xFF00FF << 120]
    n = 5000
    ks = edge + [rnd.randrange(1 << 256) for _ in range(n - len(edge))]
    odd_base = next(p for p in (J.point_from_x(x) for x in range(2, 30)) if p and J.mul(p, J.SUBGROUP_ORDER) != J.IDENTITY)
    for base in (J.GEN, odd_base, J.IDENTITY):
        shared = ctx.te_mul_batch(_d(), _pts([base]), _scalars(ks))
        per_item = ctx.te_mul_batch(_d(), np.ascontiguousarray(np.broadcast_to(_pts([base]), (n, 2, 8))), _scalars(ks))
        assert np.array_equal(shared, per_item)
        got = _pts_back(shared[:len(edge) + 4])
        assert got == [J.mul(base, k) for k in ks[:len(edge) + 4]]


@pytest.mark.gpu
def test_te_mul_other_curve_coefficient(ctx):
    """d is the caller's: another non-square coefficient gives that curve's group law."""
    d = next(v for v in range(5, 200) if pow(v, (R - 1) // 2, R) == R - 1)
    base = next(p for p in (J.point_from_x(x, d) for x in range(2, 60)) if p)
    ks = [rnd.randrange(1 << 256) for _ in range(20)]
    got = _pts_back(ctx.te_mul_batch(ark.fr_mont_from_ints([d])[0], _pts([base]), _scalars(ks)))
    assert got == [J.mul(base, k, d) for k in ks]


@pytest.mark.gpu
def test_elgamal_encrypt_batch_matches_oracle_and_decrypts(ctx):
    p = P.random_parameters(31)
    mds = ark.fr_mont_from_ints([v for row in p.mds for v in row]).reshape(3, 3, 8)
    rc = ark.fr_mont_from_ints([v for row in p.ark for v in row]).reshape(len(p.ark), 3, 8)
    pos = ctx.poseidon(p.full_rounds, p.partial_rounds, p.alpha, mds, rc)
    sk = rnd.randrange(J.SUBGROUP_ORDER)
    pk = J.keygen(sk)
    assert _pts_back(ctx.te_mul_batch(_d(), _pts([J.GEN]), _scalars([sk]))) == [pk]          # keygen
    n, chunks = 24, 6
    rs = [rnd.randrange(J.SUBGROUP_ORDER) for _ in range(n)]
    msgs = [[rnd.randrange(R) for _ in range(chunks)] for _ in range(n)]
    c1, c2 = ctx.elgamal_encrypt_batch(pos, _d(), _pts([J.GEN])[0], _pts([pk])[0], _scalars(rs),
                                       ark.fr_mont_from_ints([m for row in msgs for m in row]).reshape(n, chunks, 8))
    c1 = _pts_back(c1)
    c2 = ark.fr_ints_from_mont(c2.reshape(-1, 8))
    for i in range(n):
        want = J.encrypt(p, pk, msgs[i], rs[i])
        got = (c1[i], c2[i * chunks:(i + 1) * chunks])
        assert got == want
        assert J.decrypt(p, got, sk) == msgs[i]
    pos.free()


@pytest.mark.gpu
def test_gpu_reproduces_elgamal_golden(ctx):
    e, p = _golden()
    ks = [I(m["k"]) for m in e["multiples_of_generator"]] + [I(e["sk"])]
    got = _pts_back(ctx.te_mul_batch(_d(), _pts([_pt(e["generator"])]), _scalars(ks)))
    assert got == [_pt(m["point"]) for m in e["multiples_of_generator"]] + [_pt(e["pk"])]
    mds = ark.fr_mont_from_ints([v for row in p.mds for v in row]).reshape(3, 3, 8)
    rc = ark.fr_mont_from_ints([v for row in p.ark for v in row]).reshape(len(p.ark), 3, 8)
    pos = ctx.poseidon(p.full_rounds, p.partial_rounds, p.alpha, mds, rc)
    for c in e["ciphertexts"]:
        msg = ark.fr_mont_from_ints([I(v) for v in c["msg"]]).reshape(1, len(c["msg"]), 8)
        c1, c2 = ctx.elgamal_encrypt_batch(pos, _d(), _pts([_pt(e["generator"])])[0], _pts([_pt(e["pk"])])[0], _scalars([I(c["r"])]), msg)
        assert _pts_back(c1) == [_pt(c["c1"])]
        assert ark.fr_ints_from_mont(c2.reshape(-1, 8)) == [I(v) for v in c["c2"]]
    pos.free()


@pytest.mark.gpu
def test_te_mul_full_size_properties(ctx):
    """2^18 scalars: linearity against a second launch -- (k + 1) * G == k * G + G through the oracle's addition on a sample, and
    k * (s * G) == (k * s mod order) * G for every item."""
    n = 1 << 18
    ks = np.frombuffer(np.random.default_rng(5).bytes(32 * n), dtype=np.uint32).reshape(n, 8).copy()
    ks[:, 7] &= 0x0FFFFFFF                                      # < 2^252 < subgroup order * 2: plain integers below
    g = _pts([J.GEN])
    kg = ctx.te_mul_batch(_d(), g, ks)
    s = rnd.randrange(2, J.SUBGROUP_ORDER)
    sg = ctx.te_mul_batch(_d(), g, _scalars([s]))
    k_sg = ctx.te_mul_batch(_d(), sg, ks)
    ints = ark.words_to_ints(ks)
    ks2 = _scalars([k * s % J.SUBGROUP_ORDER for k in ints])
    assert np.array_equal(k_sg, ctx.te_mul_batch(_d(), g, ks2))
    for i in (0, 1, n // 3, n - 1):
        assert _pts_back(kg[i:i + 1])[0] == J.mul(J.GEN, ints[i])
