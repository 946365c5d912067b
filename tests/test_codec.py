"""arkworks canonical wire formats: csrc/codec.cu through the C ABI against the big-int restatement
oracle/py/serialize.py.  The proof codec is host-only and runs in the CPU suite; the key-image codec converts on the device."""
import random

import numpy as np
import pytest

import ctg
from ctg import synthetic as S
from oracle import ark
from oracle.py import bls12_381 as B
from oracle.py import groth16 as G
from oracle.py import serialize as SER
from tests import util

rnd = random.Random(20261104)


def _proof_words(a, b, c):
    return ark.g1_from_py([a])[0], ark.g2_from_py([b])[0], ark.g1_from_py([c])[0]


def _rand_proof():
    return (B.G1.mul(B.G1_GEN, rnd.randrange(1, B.R_MOD)), B.G2.mul(B.G2_GEN, rnd.randrange(1, B.R_MOD)),
            B.G1.mul(B.G1_GEN, rnd.randrange(1, B.R_MOD)))


def test_proof_bytes_match_oracle_and_round_trip():
    seen_flags = set()
    for _ in range(12):
        pr = _rand_proof()
        want = SER.proof_compressed(pr)
        got = ctg.proof_to_bytes(_proof_words(*pr))
        assert got == want
        seen_flags.add((got[47] >> 7, got[143] >> 7))
        back = ctg.proof_from_bytes(got)
        assert util.proof_to_py(back) == pr == SER.proof_from_compressed(want)
    assert len(seen_flags) > 1  # both y orders occurred


def test_proof_bytes_identity_and_negation():
    pr = (None, B.G2.mul(B.G2_GEN, 5), None)
    data = ctg.proof_to_bytes(_proof_words(*pr))
    assert data == SER.proof_compressed(pr)
    assert data[47] == 0x40 and data[191] == 0x40 and not any(data[:47])
    assert util.proof_to_py(ctg.proof_from_bytes(data)) == pr
    # P and -P differ exactly in the 0x80 flag
    P = B.G1.mul(B.G1_GEN, 77)
    Q = B.G2.mul(B.G2_GEN, 78)
    d1 = ctg.proof_to_bytes(_proof_words(P, Q, P))
    d2 = ctg.proof_to_bytes(_proof_words(B.G1.neg(P), B.G2.neg(Q), P))
    assert d1[:47] == d2[:47] and d1[47] ^ d2[47] == 0x80 and d1[143] ^ d2[143] == 0x80 and d1[144:] == d2[144:]


def test_proof_from_bytes_rejects_invalid_encodings():
    good = bytearray(SER.proof_compressed(_rand_proof()))
    # both flag bits
    bad = bytearray(good); bad[47] |= 0xC0
    with pytest.raises(ctg.CtgError):
        ctg.proof_from_bytes(bad)
    # x >= p (all ones below the flag bits)
    bad = bytearray(good); bad[0:48] = b"\xff" * 47 + b"\x1f"
    with pytest.raises(ctg.CtgError):
        ctg.proof_from_bytes(bad)
    # x with no point on the curve
    x = 1
    while SER._sqrt_fq((x ** 3 + 4) % B.P_MOD) is not None:
        x += 1
    bad = bytearray(good); bad[0:48] = x.to_bytes(48, "little")
    with pytest.raises(ctg.CtgError):
        ctg.proof_from_bytes(bad)
    # on the curve but outside the r-order subgroup: accepted unchecked, rejected by the checked read
    x = 1
    while True:
        y = SER._sqrt_fq((x ** 3 + 4) % B.P_MOD)
        if y is not None and B.G1.mul((x, y), B.R_MOD) is not None:
            break
        x += 1
    bad = bytearray(good); bad[0:48] = x.to_bytes(48, "little")
    assert ctg.proof_from_bytes(bad, check_subgroup=False)[0] is not None
    with pytest.raises(ctg.CtgError):
        ctg.proof_from_bytes(bad)
    with pytest.raises(ctg.CtgError):
        ctg.proof_from_bytes(bytes(good)[:100])


# ----------------------------------------------------------------------------- key images (device codec)
@pytest.fixture(scope="module")
def ctx():
    c = ctg.Context(0)
    yield c
    c.close()


def _trapdoor():
    return G.Trapdoor(*[rnd.randrange(1, B.R_MOD) for _ in range(5)],
                      B.G1.mul(B.G1_GEN, rnd.randrange(1, B.R_MOD)), B.G2.mul(B.G2_GEN, rnd.randrange(1, B.R_MOD)))


@pytest.mark.gpu
def test_key_image_matches_oracle_bytes(ctx):
    cs, _ = G.mul_chain(4, rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD))
    td = _trapdoor()
    pk = G.generate_parameters(cs, td)
    # force identity points into a query, as real keys have them (variables absent from B)
    pk.b_g1_query[3] = None; pk.b_g2_query[3] = None; pk.b_g2_query[0] = None
    image = SER.proving_key_unchecked(pk)
    dp = ctx.load_proving_key(image)
    got, want = dp.download(), ark.pk_from_py(pk)
    for k, w in want.items():
        assert np.array_equal(got[k].reshape(w.shape), w), k
    assert dp.serialize_unchecked() == image
    assert dp.serialize_vk() == SER.verifying_key_compressed(pk)
    assert (dp.ni, dp.nw, dp.n) == (cs.num_instance, cs.num_witness, cs.domain_size)
    dp.free()


@pytest.mark.gpu
def test_generated_key_round_trips_through_its_image_and_proves(ctx):
    """setup on the device -> circuit.zkey bytes -> load (as the seller process does at start-up, lib/src/zk/encryption.rs:31-32)
    -> prove -> compressed proof bytes -> checked read -> pairing equation accepts."""
    r1cs, zm, z = S.poe_shape(1, gates=200, bits=32)
    td = _trapdoor()
    dp = ctx.generate_parameters(r1cs, util.trapdoor_words(td))
    image = dp.serialize_unchecked()
    key = dp.download()
    want_pk = ark.g1_to_py(key["a_query"])
    assert SER.point_uncompressed(want_pk[0], 1) == image[96 + 3 * 192 + 8 + 96 * r1cs["num_instance"] + 2 * 96 + 8:][:96]
    dp2 = ctx.load_proving_key(image)
    got = dp2.download()
    for k in key:
        assert np.array_equal(got[k], key[k]), k
    dpk, dr = dp2.to_device_pk(), ctx.upload_r1cs(r1cs)
    r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    fm = lambda v: ark.fr_mont_from_ints([v])[0]
    proof = ctx.prove(dpk, dr, zm, fm(r), fm(s))
    data = ctg.proof_to_bytes(proof)
    assert data == SER.proof_compressed(util.proof_to_py(proof))
    back = ctg.proof_from_bytes(data)
    assert util.groth16_verify_words(key, back, z[1:r1cs["num_instance"]])
    dpk.free(); dr.free(); dp.free(); dp2.free()


@pytest.mark.gpu
def test_key_image_rejects_malformed(ctx):
    cs, _ = G.mul_chain(3, 5, 7)
    image = SER.proving_key_unchecked(G.generate_parameters(cs, _trapdoor()))
    for bad in (image[:-1], image + b"\0", image[:500], b""):
        with pytest.raises(ctg.CtgError):
            ctx.load_proving_key(bad)
    bad = bytearray(image); bad[0:48] = b"\xff" * 48          # alpha_g1.x >= p
    with pytest.raises(ctg.CtgError):
        ctx.load_proving_key(bytes(bad))
    bad = bytearray(image); bad[95] |= 0xC0                    # both flags on alpha_g1.y
    with pytest.raises(ctg.CtgError):
        ctx.load_proving_key(bytes(bad))
    bad = bytearray(image); off = 96 + 3 * 192; bad[off:off + 8] = (10 ** 12).to_bytes(8, "little")   # absurd gamma_abc length
    with pytest.raises(ctg.CtgError):
        ctx.load_proving_key(bytes(bad))
