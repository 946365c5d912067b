"""GPU parity tests (run on the B200 box): every call goes through the C ABI of libctg_groth16.so
(contangle-zkcp_b200/ctg) and is compared bit-for-bit with the oracle (oracle/c C++ restatement of
arkworks 0.3.0, itself pinned against oracle/py and the trapdoor closed form in test_oracle.py)."""
import random

import numpy as np
import pytest

import ctg
from ctg import synthetic as S
from oracle import ark
from oracle.py import bls12_381 as B
from oracle.py import groth16 as G
from tests import util

pytestmark = pytest.mark.gpu
rnd = random.Random(424242)


@pytest.fixture(scope="module")
def ctx():
    c = ctg.Context(0)
    yield c
    c.close()


def fm(v):
    return ark.fr_mont_from_ints([v])[0]


# ----------------------------------------------------------------------------- fields
def _edge(mod, n):
    vals = [0, 1, 2, mod - 1, mod - 2, (1 << 255) % mod, (mod - 1) // 2, 0xFFFFFFFF, 1 << 32, (1 << 64) - 1]
    return vals + [rnd.randrange(mod) for _ in range(n)]


@pytest.mark.parametrize("field,w,mod", [(ctg.FIELD_FR, 8, B.R_MOD), (ctg.FIELD_FQ, 12, B.P_MOD)])
def test_field_ops(ctx, field, w, mod):
    av = _edge(mod, 5000)
    bv = list(reversed(_edge(mod, 5000)))
    a, b = ark.ints_to_words(av, w), ark.ints_to_words(bv, w)
    name = "fr" if field == ctg.FIELD_FR else "fq"
    for op, opn in ((ctg.OP_MUL, "mul"), (ctg.OP_ADD, "add"), (ctg.OP_SUB, "sub")):
        got = ctx.field_op(field, op, a, b)
        assert np.array_equal(got, getattr(ark, f"{name}_{opn}")(a, b)), opn
    assert np.array_equal(ctx.fr_mul(a[:, :8], b[:, :8]) if w == 8 else ctx.fq_mul(a, b), getattr(ark, f"{name}_mul")(a, b))


@pytest.mark.parametrize("field,w,mod,R", [(ctg.FIELD_FR, 8, B.R_MOD, B.FR_R), (ctg.FIELD_FQ, 12, B.P_MOD, B.FQ_R)])
def test_field_inverse_bingcd(ctx, field, w, mod, R):
    vals = _edge(mod, 1000) + [rnd.randrange(1 << k) % mod for k in range(1, w * 32 - 2, 3)]
    a = ark.ints_to_words(vals, w)
    got = ark.words_to_ints(ctx.field_op(field, ctg.OP_INV, a, a))
    rinv = pow(R, -1, mod)
    assert got == [0 if v == 0 else pow(v * rinv % mod, -1, mod) * R % mod for v in vals]


def test_fq2_mul(ctx):
    n = 3000
    a = ark.ints_to_words([rnd.randrange(B.P_MOD) for _ in range(2 * n)], 12).reshape(n, 24)
    b = ark.ints_to_words([rnd.randrange(B.P_MOD) for _ in range(2 * n)], 12).reshape(n, 24)
    assert np.array_equal(ctx.field_op(ctg.FIELD_FQ2, ctg.OP_MUL, a, b), ark.fq2_mul(a, b))


def test_fq2_lane_paired_arithmetic(ctx):
    """csrc/msm_pair.cuh: an Fq2 element on an even / odd lane pair (the G2 bucket-reduction tail).  Products, squares, the fused
    a b - c d and the pair's agreement on is_zero, element by element against the oracle -- with operands whose components are
    0, 1, p - 1 or random, an odd element count (a warp with an idle tail) and every zero / non-zero mix of the two components."""
    edge = [0, 1, B.P_MOD - 1, 2, (B.P_MOD - 1) // 2]
    vals = [(x, y) for x in edge for y in edge] + [(rnd.randrange(B.P_MOD), rnd.randrange(B.P_MOD)) for _ in range(2974)]
    vals2 = vals[::-1]
    n = len(vals)
    assert n % 2 == 1
    a = ark.ints_to_words([c for v in vals for c in v], 12).reshape(n, 24)
    b = ark.ints_to_words([c for v in vals2 for c in v], 12).reshape(n, 24)
    ab = ark.fq2_mul(a, b)
    aa = ark.fq2_mul(a, a)
    assert np.array_equal(ctx.field_op(ctg.FIELD_FQ2, ctg.OP_PAIRED_MUL, a, b), ab)
    assert np.array_equal(ctx.field_op(ctg.FIELD_FQ2, ctg.OP_PAIRED_SQR, a, b), aa)
    baa = ark.fq2_mul(b, aa)
    want = ark.fq_sub(ab.reshape(2 * n, 12), baa.reshape(2 * n, 12)).reshape(n, 24)      # component-wise in the oracle
    assert np.array_equal(ctx.field_op(ctg.FIELD_FQ2, ctg.OP_PAIRED_MUL2SUB, a, b), want)
    isz = ctx.field_op(ctg.FIELD_FQ2, ctg.OP_PAIRED_IS_ZERO, a, b)
    one = ark.ints_to_words([B.FQ_R % B.P_MOD if hasattr(B, "FQ_R") else pow(2, 384, B.P_MOD), 0], 12).reshape(24)
    for i, (x, y) in enumerate(vals):
        assert np.array_equal(isz[i], one if (x == 0 and y == 0) else np.zeros(24, dtype=np.uint32)), (i, x == 0, y == 0)


def test_empty_inputs(ctx):
    e = np.zeros((0, 8), dtype=np.uint32)
    assert ctx.field_op(ctg.FIELD_FR, ctg.OP_MUL, e, e).shape == (0, 8)


# ----------------------------------------------------------------------------- NTT
@pytest.mark.parametrize("log_n", [0, 1, 2, 3, 7, 10, 11, 12, 13, 16, 19, 20])
def test_ntt_all_modes(ctx, log_n):
    x = ark.random_fr_mont(1 << log_n, 1000 + log_n)
    for inverse, coset in ((False, False), (True, False), (False, True), (True, True)):
        got = ctx.ntt(x, inverse=inverse, coset=coset)
        assert np.array_equal(got, ark.ntt_fr(x, inverse=inverse, coset=coset)), (log_n, inverse, coset)


@pytest.mark.parametrize("log_n", [22])
def test_ntt_roundtrip_large(ctx, log_n):
    # size-independent properties at a BASELINE size: inverse(forward(x)) == x, coset likewise
    x = ark.random_fr_mont(1 << log_n, 77)
    assert np.array_equal(ctx.ntt(ctx.ntt(x), inverse=True), x)
    assert np.array_equal(ctx.ntt(ctx.ntt(x, coset=True), inverse=True, coset=True), x)


def test_ntt_rejects_bad_length(ctx):
    with pytest.raises(ctg.CtgError):
        ctx.ntt(np.zeros((3, 8), dtype=np.uint32))


# ----------------------------------------------------------------------------- bases
def test_progression_bases_match_oracle(ctx):
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    n = 1000
    t = ctx.bases_progression_g1(ark.g1_generator(), fm(a), fm(b), n)
    assert np.array_equal(t.download(), ark.g1_progression(ark.g1_generator(), fm(a), fm(b), n))
    t.free()
    t2 = ctx.bases_progression_g2(ark.g2_generator(), fm(a), fm(b), 100)
    assert np.array_equal(t2.download(), ark.g2_progression(ark.g2_generator(), fm(a), fm(b), 100))
    t2.free()
    assert np.array_equal(S.g1_gen_words(), ark.g1_generator()) and np.array_equal(S.g2_gen_words(), ark.g2_generator())


def test_precomputed_copies_are_shifted_multiples(ctx):
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    n = 50
    pts = ark.g1_progression(ark.g1_generator(), fm(a), fm(b), n)
    pts[3] = 0
    t = ctx.bases_g1(pts, copies=4, c_bits=8)   # 32 windows -> 8 sets, copy j = 2^(8*8*j) P
    all_pts = t.download(0, 4 * n).reshape(4, n, 24)
    assert np.array_equal(all_pts[0], pts)
    for j in (1, 3):
        for i in (0, 3, 17):
            want = ark.g1_mul(pts[i], ark.ints_to_words([1 << (64 * j)], 8)[0])
            assert np.array_equal(all_pts[j][i], want)
    t.free()


# ----------------------------------------------------------------------------- MSM
def _g1_bases(n, seed, with_identity=True):
    r = random.Random(seed)
    pts = ark.g1_progression(ark.g1_generator(), fm(r.randrange(B.R_MOD)), fm(r.randrange(B.R_MOD)), n)
    if with_identity and n > 5:
        pts[5] = 0
    return pts


@pytest.mark.parametrize("n", [1, 2, 31, 33, 1000, 1 << 12, (1 << 14) + 7])
@pytest.mark.parametrize("kind", ["uniform", "poe"])
def test_msm_g1_vs_oracle(ctx, n, kind):
    pts = _g1_bases(n, n)
    sc = ark.random_scalars_canon(n, n + 1, kind)
    t = ctx.bases_g1(pts)
    assert np.array_equal(ctx.msm(t, sc), ark.msm_g1(pts, sc))
    # Montgomery-encoded scalars give the same point
    assert np.array_equal(ctx.msm(t, ark.fr_to_mont(sc), mont=True), ark.msm_g1(pts, sc))
    t.free()


@pytest.mark.parametrize("kind", ["equal", "rminus1", "zeros", "ones", "duplicates"])
def test_msm_g1_adversarial(ctx, kind):
    n = 3000
    pts = _g1_bases(n, 99)
    if kind == "zeros":
        sc = np.zeros((n, 8), dtype=np.uint32)
    elif kind == "ones":
        sc = np.zeros((n, 8), dtype=np.uint32); sc[:, 0] = 1
    elif kind == "duplicates":
        pts[:] = pts[0]                      # every bucket add hits the doubling path
        pts[1::2, 12:] = ark.g1_from_py([B.G1.neg(ark.g1_to_py(pts[:1])[0])])[0][12:]   # alternate P, -P
        sc = ark.random_scalars_canon(n, 5, "equal")
    else:
        sc = ark.random_scalars_canon(n, 5, kind)
    t = ctx.bases_g1(pts)
    assert np.array_equal(ctx.msm(t, sc), ark.msm_g1(pts, sc)), kind
    t.free()


@pytest.mark.parametrize("c_bits,copies", [(2, 1), (4, 1), (8, 1), (8, 4), (11, 2), (13, 20), (16, 1), (16, 16), (16, 3), (17, 16), (18, 15), (19, 5), (20, 13),
                                           (21, 13)])
def test_msm_g1_window_and_precompute_choices(ctx, c_bits, copies):
    n = 5000
    pts = _g1_bases(n, 7)
    sc = ark.random_scalars_canon(n, 8, "uniform")
    sc[:50] = ark.random_scalars_canon(50, 9, "rminus1")
    t = ctx.bases_g1(pts, copies=copies, c_bits=c_bits)
    assert np.array_equal(ctx.msm(t, sc), ark.msm_g1(pts, sc)), (c_bits, copies)
    t.free()


# ---- batched-affine tree levels (msm_acc.cuh stage 4a): forced on small inputs, every pair kind ----
@pytest.mark.parametrize("levels", [1, 2, 3, 6])
@pytest.mark.parametrize("kind", ["uniform", "poe", "equal", "duplicates", "cancel", "inf", "tiny"])
def test_msm_g1_batched_affine_levels(ctx, levels, kind):
    n = 2500 if kind != "tiny" else 3
    pts = _g1_bases(n, 41)
    sc = ark.random_scalars_canon(n, 43, kind if kind in ("uniform", "poe", "equal") else "uniform")
    if kind == "duplicates":      # equal points with equal scalars share every bucket: tangent (doubling) pairs at every level
        pts[:] = pts[0]
        sc[:] = sc[0]
    elif kind == "cancel":        # P and -P with the same scalar sit next to each other in every bucket: chord through infinity
        neg = ark.g1_from_py([B.G1.neg(P) for P in ark.g1_to_py(pts[0::2])])
        pts[1::2] = neg[: pts[1::2].shape[0]]
        sc[1::2] = sc[0::2][: sc[1::2].shape[0]]
        sc[-1] = sc[0]            # one survivor so that the total is not the identity
    elif kind == "inf":
        pts[::3] = 0
    ctx.set_ba_levels(levels)
    try:
        for c_bits in (4, 9):     # 8 / 256 buckets per window: deep and shallow trees
            t = ctx.bases_g1(pts, copies=1, c_bits=c_bits)
            assert np.array_equal(ctx.msm(t, sc), ark.msm_g1(pts, sc)), (kind, levels, c_bits)
            t.free()
    finally:
        ctx.set_ba_levels(-1)


@pytest.mark.parametrize("levels", [1, 3])
@pytest.mark.parametrize("kind", ["uniform", "duplicates", "cancel"])
def test_msm_g2_batched_affine_levels(ctx, levels, kind):
    n = 700
    pts = ark.g2_progression(ark.g2_generator(), fm(rnd.randrange(B.R_MOD)), fm(rnd.randrange(B.R_MOD)), n)
    sc = ark.random_scalars_canon(n, 47, "uniform")
    pts[5] = 0
    if kind == "duplicates":
        pts[:] = pts[0]
        sc[:] = sc[0]
    elif kind == "cancel":
        neg = ark.g2_from_py([B.G2.neg(P) for P in ark.g2_to_py(pts[0::2])])
        pts[1::2] = neg
        sc[1::2] = sc[0::2]
        sc[-1] = sc[0]
    ctx.set_ba_levels(levels)
    try:
        t = ctx.bases_g2(pts, copies=2, c_bits=5)
        assert np.array_equal(ctx.msm(t, sc), ark.msm_g2(pts, sc)), (kind, levels)
        t.free()
    finally:
        ctx.set_ba_levels(-1)


@pytest.mark.parametrize("n_inputs", [0, 1, 5, 2050])
def test_prepare_inputs_matches_arkworks_loop(ctx, n_inputs):
    """ark_groth16::prepare_inputs: g_ic = gamma_abc[0]; g_ic += gamma_abc[i + 1] * x_i -- against that loop in the oracle."""
    gamma = _g1_bases(n_inputs + 1, 1234)
    if n_inputs > 3:
        gamma[2] = 0
    xs = ark.random_scalars_canon(n_inputs, 77, "poe") if n_inputs else np.zeros((0, 8), dtype=np.uint32)
    t = ctx.bases_g1(gamma)
    got = ctx.prepare_inputs(t, ark.fr_to_mont(xs) if n_inputs else xs)
    want = gamma[0] if n_inputs == 0 else ark.g1_add(gamma[0], ark.msm_g1(gamma[1:], xs))
    assert np.array_equal(got, want)
    with pytest.raises(ctg.CtgError):
        ctx.prepare_inputs(t, np.zeros((n_inputs + 1, 8), dtype=np.uint32))     # malformed verifying key
    t.free()


@pytest.mark.parametrize("parts", [2, 3, 4])
def test_msm_split_parts_match(ctx, parts):
    """Point-range parts on separate streams + on-device combine give the same point (G1 with precomputed copies, G2)."""
    n = 5003
    pts = _g1_bases(n, 17)
    sc = ark.random_scalars_canon(n, 18, "poe")
    want = ark.msm_g1(pts, sc)
    p2 = ark.g2_progression(ark.g2_generator(), fm(3), fm(5), 600)
    sc2 = ark.random_scalars_canon(600, 19, "uniform")
    want2 = ark.msm_g2(p2, sc2)
    ctx.set_msm_split(parts)
    try:
        for copies, c_bits in ((1, 0), (4, 8)):
            t = ctx.bases_g1(pts, copies=copies, c_bits=c_bits)
            assert np.array_equal(ctx.msm(t, sc), want)
            assert np.array_equal(ctx.msm(t, sc[:777]), ark.msm_g1(pts[:777], sc[:777]))
            t.free()
        t2 = ctx.bases_g2(p2, copies=2, c_bits=6)
        assert np.array_equal(ctx.msm(t2, sc2), want2)
        t2.free()
    finally:
        ctx.set_msm_split(-1)


def test_msm_g1_prefix_and_errors(ctx):
    n = 2000
    pts = _g1_bases(n, 3)
    sc = ark.random_scalars_canon(n, 4)
    t = ctx.bases_g1(pts)
    assert np.array_equal(ctx.msm(t, sc[:777]), ark.msm_g1(pts[:777], sc[:777]))   # min(len) semantics
    assert not ctx.msm(t, sc[:0]).any()                                              # empty -> identity
    with pytest.raises(ctg.CtgError):
        ctx.msm(t, np.zeros((n + 1, 8), dtype=np.uint32))
    t2 = ctx.bases_g2(ark.g2_progression(ark.g2_generator(), fm(1), fm(1), 4))
    with pytest.raises(ctg.CtgError):
        load_fn = ctg.load().ctg_msm_g1
        ctx._check(load_fn(ctx.h, t2.handle, None, 0, 0, None))
    t.free(); t2.free()


@pytest.mark.parametrize("n,kind", [(1, "uniform"), (40, "poe"), (1500, "uniform"), (4096, "poe")])
def test_msm_g2_vs_oracle(ctx, n, kind):
    pts = ark.g2_progression(ark.g2_generator(), fm(rnd.randrange(B.R_MOD)), fm(rnd.randrange(B.R_MOD)), n)
    if n > 5:
        pts[2] = 0
    sc = ark.random_scalars_canon(n, n + 11, kind)
    t = ctx.bases_g2(pts)
    assert np.array_equal(ctx.msm(t, sc), ark.msm_g2(pts, sc))
    t.free()
    t = ctx.bases_g2(pts, copies=3, c_bits=9)
    assert np.array_equal(ctx.msm(t, sc), ark.msm_g2(pts, sc))
    t.free()


@pytest.mark.parametrize("log_n,kind", [(16, "uniform"), (18, "poe"), (20, "uniform")])
def test_msm_g1_dlog_closed_form_large(ctx, log_n, kind):
    """Bases generated on the device with known discrete logs: the MSM must equal (sum s_i (a + i b)) G."""
    n = 1 << log_n
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    t = ctx.bases_progression_g1(ark.g1_generator(), fm(a), fm(b), n, copies=2)
    sc = ark.random_scalars_canon(n, log_n, kind)
    got = ark.g1_to_py(ctx.msm(t, sc))[0]
    s_int = ark.words_to_ints(sc)
    dl = (a * sum(s_int) + b * sum(i * s for i, s in enumerate(s_int))) % B.R_MOD
    assert got == B.G1.mul(B.G1_GEN, dl)
    t.free()


def test_msm_partials_combine(ctx):
    """Point-range sharding as the multi-GPU path does it: partial sums per range, combined on one device."""
    n = 6000
    pts = _g1_bases(n, 21)
    sc = ark.random_scalars_canon(n, 22)
    parts = []
    for lo, hi in ((0, 1500), (1500, 3000), (3000, 4500), (4500, 6000)):
        t = ctx.bases_g1(pts[lo:hi])
        parts.append(ctx.msm_partial(t, sc[lo:hi]))
        t.free()
    assert np.array_equal(ctx.combine(np.stack(parts), group=1), ark.msm_g1(pts, sc))


# ----------------------------------------------------------------------------- witness map and prover
def _trapdoor():
    return G.Trapdoor(*[rnd.randrange(1, B.R_MOD) for _ in range(5)],
                      B.G1.mul(B.G1_GEN, rnd.randrange(1, B.R_MOD)), B.G2.mul(B.G2_GEN, rnd.randrange(1, B.R_MOD)))


@pytest.mark.parametrize("log_n", [2, 3, 6, 10, 12, 14])
def test_witness_map_vs_oracle(ctx, log_n):
    r1cs, zm, z = S.mul_chain(log_n)
    d = ctx.upload_r1cs(r1cs)
    assert np.array_equal(ctx.witness_map(d, zm), ark.witness_map(r1cs, zm))
    d.free()


def test_witness_map_poe_shape(ctx):
    r1cs, zm, z = S.poe_shape(100)
    d = ctx.upload_r1cs(r1cs)
    h = ctx.witness_map(d, zm)
    assert np.array_equal(h, ark.witness_map(r1cs, zm))
    assert not h[-1].any()     # deg h <= n - 2
    d.free()


@pytest.mark.parametrize("log_n,r_zero,copies", [(3, False, 1), (5, True, 1), (8, False, 1), (10, False, 4)])
def test_prove_trapdoor_key_vs_oracle_and_closed_form(ctx, log_n, r_zero, copies):
    cs, z = G.mul_chain(log_n, rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD))
    td = _trapdoor()
    pk = util.trapdoor_pk_arrays(cs, td)
    r1 = ark.r1cs_from_py(cs)
    r = 0 if r_zero else rnd.randrange(B.R_MOD)
    s = rnd.randrange(B.R_MOD)
    zm = ark.fr_mont_from_ints(z)
    dpk = ctx.upload_pk(pk, cs.num_instance, cs.num_constraints, copies=copies)
    dr = ctx.upload_r1cs(r1)
    got = ctx.prove(dpk, dr, zm, fm(r), fm(s))
    want = ark.create_proof(pk, r1, zm, fm(r), fm(s))
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    if not r_zero:
        assert util.proof_to_py(got) == G.closed_form_proof(cs, td, z, r, s)
    dpk.free(); dr.free()


def test_prove_poe_shape_vs_oracle(ctx):
    r1cs, zm, z = S.poe_shape(100)
    pk, dl = S.progression_pk(ctx, r1cs)
    r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    dpk = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"])
    dr = ctx.upload_r1cs(r1cs)
    got, tm = ctx.prove(dpk, dr, zm, fm(r), fm(s), timed=True)
    want = ark.create_proof(pk, r1cs, zm, fm(r), fm(s))
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    assert tm["total_ms"] > 0
    dpk.free(); dr.free()


@pytest.mark.parametrize("log_n", [16])
def test_prove_progression_key_dlog_closed_form(ctx, log_n):
    """Full-size pin that needs no MSM / NTT oracle for the group part: the key's discrete logs are known."""
    r1cs, zm, z = S.mul_chain(log_n)
    pk, dl = S.progression_pk(ctx, r1cs)
    r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    dpk = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"], copies=2)
    dr = ctx.upload_r1cs(r1cs)
    got = ctx.prove(dpk, dr, zm, fm(r), fm(s))
    h = ark.fr_ints_from_mont(ark.witness_map(r1cs, zm))
    assert util.proof_to_py(got) == util.dlog_expected_proof(dl, r1cs, z, h, r, s)
    dpk.free(); dr.free()


def test_prove_batch_matches_single_proofs(ctx):
    """ctg_prove_batch pipelines independent proofs over several stream sets; each must equal its own ctg_prove / oracle proof."""
    r1cs, zm0, z0 = S.poe_shape(3, gates=300, bits=40)
    pk, _ = S.progression_pk(ctx, r1cs)
    dpk = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"])
    dr = ctx.upload_r1cs(r1cs)
    nproofs = 7
    zs, rss = [], []
    for i in range(nproofs):
        _, zm, _ = S.poe_shape(3, seed=1000 + i, gates=300, bits=40)     # same shape, different witness
        zs.append(zm)
        rss.append(np.stack([fm(rnd.randrange(B.R_MOD)), fm(rnd.randrange(B.R_MOD))]))
    out = ctx.prove_batch(dpk, dr, np.stack(zs), np.stack(rss))
    for i in range(nproofs):
        single = ctx.prove(dpk, dr, zs[i], rss[i][0], rss[i][1])
        assert np.array_equal(out[i], np.concatenate(single)), i
    for i in (0, nproofs - 1):
        want = ark.create_proof(pk, r1cs, zs[i], rss[i][0], rss[i][1])
        assert np.array_equal(out[i], np.concatenate(want)), i
    dpk.free(); dr.free()


@pytest.mark.parametrize("shape,nproofs", [("mulchain1", 5), ("mulchain5", 300), ("mulchain12", 9), ("poe1", 40)])
def test_prove_batch_fused_chunks(ctx, shape, nproofs):
    """The fused batch pipeline (one sort / accumulate / reduce per MSM for a whole chunk of proofs; batched SpMV and NTT passes):
    independent witnesses AND independent (r, s) per proof, more proofs than one chunk holds (300 > 256), the smallest domain (2),
    a two-pass NTT domain (2^12) and the PoE shape with its skewed witness; every proof equals ctg_prove's, samples equal the oracle's."""
    def make(seed):
        if shape.startswith("mulchain"):
            return S.mul_chain(int(shape[8:]), seed=seed)
        return S.poe_shape(1, seed=seed, gates=600, bits=256)
    r1cs, _, _ = make(1)
    pk, _ = S.progression_pk(ctx, r1cs)
    dpk = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"])
    dr = ctx.upload_r1cs(r1cs)
    zs = np.stack([make(1 + (i % 11))[1] for i in range(nproofs)])
    rss = ark.fr_mont_from_ints([rnd.randrange(B.R_MOD) for _ in range(2 * nproofs)]).reshape(nproofs, 2, 8)
    rss[min(3, nproofs - 1), 0] = 0                      # r = 0: arkworks skips B1 (the product r * B1 vanishes either way)
    out = ctx.prove_batch(dpk, dr, zs, rss)
    step = max(1, nproofs // 25)
    for i in list(range(0, nproofs, step)) + [nproofs - 1]:
        single = ctx.prove(dpk, dr, zs[i], rss[i][0], rss[i][1])
        assert np.array_equal(out[i], np.concatenate(single)), i
    for i in (0, min(3, nproofs - 1), nproofs - 1):
        want = ark.create_proof(pk, r1cs, zs[i], rss[i][0], rss[i][1])
        assert np.array_equal(out[i], np.concatenate(want)), i
    dpk.free(); dr.free()


def test_poe_shape_witness_batch_satisfies_and_proves(ctx):
    """ctg.synthetic.poe_shape_witnesses: independent assignments of the PoE shape built with the device's Fr product.  Each must
    satisfy the R1CS (host big-int check), differ from the others, and prove in a batch to what ctg_prove / the oracle give."""
    r1cs, _, _ = S.poe_shape(4, gates=120, bits=24)
    zb = S.poe_shape_witnesses(ctx, 4, 6, seed=77, gates=120, bits=24)
    assert zb.shape == (6, r1cs["num_instance"] + r1cs["num_witness"], 8)
    for i in range(6):
        assert S.r1cs_is_satisfied(r1cs, ark.fr_ints_from_mont(zb[i])), i
    assert len({zb[i].tobytes() for i in range(6)}) == 6
    pk, _ = S.progression_pk(ctx, r1cs)
    dpk = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"])
    dr = ctx.upload_r1cs(r1cs)
    rss = ark.fr_mont_from_ints([rnd.randrange(B.R_MOD) for _ in range(12)]).reshape(6, 2, 8)
    out = ctx.prove_batch(dpk, dr, zb, rss)
    for i in range(6):
        assert np.array_equal(out[i], np.concatenate(ctx.prove(dpk, dr, zb[i], rss[i][0], rss[i][1]))), i
    assert np.array_equal(out[5], np.concatenate(ark.create_proof(pk, r1cs, zb[5], rss[5][0], rss[5][1])))
    dpk.free(); dr.free()


def test_prove_batch_1024_independent_poe_proofs(ctx):
    """BASELINE.json configs[4] at its stated batch: 1024 independent PoE-shaped assignments (domain 2^13) under one
    device-resident generated key; sampled proofs equal ctg_prove's, two equal the oracle's, and the batch stays under
    20 kernel launches per proof."""
    r1cs, _, _ = S.poe_shape(1)
    zb = S.poe_shape_witnesses(ctx, 1, 1024, seed=5)
    dp = ctx.generate_parameters(r1cs, S.trapdoor()[0])
    key_host = dp.download()
    dpk = dp.to_device_pk()
    dp.free()
    dr = ctx.upload_r1cs(r1cs)
    rss = ark.fr_mont_from_ints([rnd.randrange(B.R_MOD) for _ in range(2048)]).reshape(1024, 2, 8)
    l0 = ctx.launches
    out = ctx.prove_batch(dpk, dr, zb, rss)
    assert (ctx.launches - l0) / 1024 <= 20
    for i in list(range(0, 1024, 93)) + [1023]:
        assert np.array_equal(out[i], np.concatenate(ctx.prove(dpk, dr, zb[i], rss[i][0], rss[i][1]))), i
    for i in (0, 1023):
        assert np.array_equal(out[i], np.concatenate(ark.create_proof(key_host, r1cs, zb[i], rss[i][0], rss[i][1]))), i
    dpk.free(); dr.free()


@pytest.mark.parametrize("log_n,shards", [(4, 2), (9, 3), (12, 8)])
def test_prove_sharded_partials_combine(ctx, log_n, shards):
    """The multi-GPU path on one GPU: every shard's ctg_prove_partial, then ctg_prove_combine, equals ctg_prove."""
    r1cs, zm, z = S.mul_chain(log_n)
    pk, _ = S.progression_pk(ctx, r1cs)
    r, s = fm(rnd.randrange(B.R_MOD)), fm(rnd.randrange(B.R_MOD))
    dr = ctx.upload_r1cs(r1cs)
    full = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"])
    want = ctx.prove(full, dr, zm, r, s)
    full.free()
    parts = []
    for k in range(shards):
        sh = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"], shard=(k, shards))
        with pytest.raises(ctg.CtgError):
            ctx.prove(sh, dr, zm, r, s)          # a shard cannot produce a whole proof
        parts.append(ctx.prove_partial(sh, dr, zm, r, s))
        sh.free()
    got = ctx.prove_combine(np.stack(parts), r, s)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    if log_n <= 9:
        for g, w in zip(got, ark.create_proof(pk, r1cs, zm, r, s)):
            assert np.array_equal(g, w)
    dr.free()


def test_prove_rejects_mismatched_shapes(ctx):
    r1a, zma, _ = S.mul_chain(4)
    r1b, zmb, _ = S.mul_chain(5)
    pk, _ = S.progression_pk(ctx, r1a)
    dpk = ctx.upload_pk(pk, r1a["num_instance"], r1a["num_constraints"])
    drb = ctx.upload_r1cs(r1b)
    with pytest.raises(ctg.CtgError):
        ctx.prove(dpk, drb, zmb, fm(1), fm(2))
    bad = dict(pk); bad["h_query"] = pk["h_query"][:-1]
    with pytest.raises(ctg.CtgError):
        ctx.upload_pk(bad, r1a["num_instance"], r1a["num_constraints"])
    dpk.free(); drb.free()
    # malformed CSR structure is refused at upload (the SpMV kernel trusts row_ptr): non-zero start, decreasing offsets
    for mutate in (lambda rp: rp.__setitem__(0, 1), lambda rp: rp.__setitem__(3, int(rp[4]) + 1)):
        bad_r = dict(r1b); bad_r["a"] = dict(r1b["a"]); bad_r["a"]["row_ptr"] = r1b["a"]["row_ptr"].copy()
        mutate(bad_r["a"]["row_ptr"])
        with pytest.raises(ctg.CtgError):
            ctx.upload_r1cs(bad_r)
        with pytest.raises(ctg.CtgError):
            ctx.generate_parameters(bad_r, S.trapdoor()[0])


# ----------------------------------------------------------------------------- key generation
KEY_MEMBERS = ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "gamma_g2", "delta_g2", "gamma_abc_g1",
               "a_query", "b_g1_query", "b_g2_query", "h_query", "l_query")


def test_generate_parameters_vs_bigint_oracle(ctx):
    """Every member of the key against oracle/py generate_parameters (independent big-int restatement)."""
    cs, _ = G.mul_chain(3, rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD))
    td = _trapdoor()
    want = ark.pk_from_py(G.generate_parameters(cs, td))
    dp = ctx.generate_parameters(ark.r1cs_from_py(cs), util.trapdoor_words(td))
    got = dp.download()
    for k in KEY_MEMBERS:
        assert np.array_equal(got[k].reshape(want[k].shape), want[k]), k
    dp.free()


@pytest.mark.parametrize("shape", ["mulchain10", "mulchain14", "poe100"])
def test_generate_parameters_vs_oracle(ctx, shape):
    """Against the arkworks-algorithm C++ restatement (Lagrange coefficients + fixed-base multiplication); poe100 has the
    heavy constant column (hundreds of rows touch z[0]) and 103 instance columns; mulchain14 takes the 16-bit-window path."""
    if shape.startswith("mulchain"):
        r1cs, _, _ = S.mul_chain(int(shape[8:]))
    else:
        r1cs, _, _ = S.poe_shape(100)
    td = _trapdoor()
    want = util.trapdoor_pk_arrays(util.cs_from_arrays(r1cs), td)
    dp = ctx.generate_parameters(r1cs, util.trapdoor_words(td))
    got = dp.download()
    for k in KEY_MEMBERS:
        assert np.array_equal(got[k].reshape(want[k].shape), want[k]), k
    dp.free()


@pytest.mark.parametrize("shape", ["mulchain6", "poe1"])
def test_setup_prove_verify_round_trip(ctx, shape):
    """The reference's own test shape (circuits-ark/src/encryption.rs:403-411): setup -> prove -> verify accepts, and rejects a
    wrong public input.  Setup and prove run on the GPU (key never leaves HBM between them); prepare_inputs runs on the GPU
    and is compared with the oracle's; the pairing equation is the big-int oracle (oracle/py/pairing.py)."""
    if shape.startswith("mulchain"):
        r1cs, zm, z = S.mul_chain(int(shape[8:]))
    else:
        r1cs, zm, z = S.poe_shape(1, gates=300, bits=64)
    assert S.r1cs_is_satisfied(r1cs, z)
    td = _trapdoor()
    dp = ctx.generate_parameters(r1cs, util.trapdoor_words(td))
    dpk = dp.to_device_pk()
    dr = ctx.upload_r1cs(r1cs)
    r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    proof = ctx.prove(dpk, dr, zm, fm(r), fm(s))
    key = dp.download(("alpha_g1", "beta_g2", "gamma_g2", "delta_g2", "gamma_abc_g1"))
    ni = r1cs["num_instance"]
    public = z[1:ni]
    assert util.groth16_verify_words(key, proof, public)
    bad = list(public); bad[-1] = (bad[-1] + 1) % B.R_MOD
    assert not util.groth16_verify_words(key, proof, bad)
    # the proof also equals the trapdoor closed form
    assert util.proof_to_py(proof) == G.closed_form_proof(util.cs_from_arrays(r1cs), td, z, r, s)
    # verifier-side aggregation on the device from the device-resident gamma_abc_g1
    from oracle.py import pairing as PR
    gb = dp.gamma_abc_bases()
    acc = ctx.prepare_inputs(gb, ark.fr_mont_from_ints(public))
    assert ark.g1_to_py(acc)[0] == PR.prepare_inputs(ark.g1_to_py(key["gamma_abc_g1"]), public)
    gb.free(); dpk.free(); dr.free(); dp.free()


def test_generate_parameters_sharded_key_from_device(ctx):
    """Point-range shards built device-to-device from the generated key give the same proof as the whole key."""
    r1cs, zm, z = S.mul_chain(9)
    td = _trapdoor()
    dp = ctx.generate_parameters(r1cs, util.trapdoor_words(td))
    dr = ctx.upload_r1cs(r1cs)
    r, s = fm(rnd.randrange(B.R_MOD)), fm(rnd.randrange(B.R_MOD))
    whole = dp.to_device_pk()
    want = ctx.prove(whole, dr, zm, r, s)
    parts = []
    for i in range(3):
        sh = dp.to_device_pk(shard=(i, 3))
        parts.append(ctx.prove_partial(sh, dr, zm, r, s))
        sh.free()
    got = ctx.prove_combine(np.stack(parts), r, s)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    whole.free(); dr.free(); dp.free()


def test_generate_parameters_rejects_bad_trapdoor(ctx):
    r1cs, _, _ = S.mul_chain(4)
    td = _trapdoor()
    for field_name, value in (("delta", 0), ("gamma", 0), ("tau", 1), ("tau", B.fr_root_of_unity(4))):
        w = util.trapdoor_words(td)
        w[field_name] = fm(value)
        with pytest.raises(ctg.CtgError):
            ctx.generate_parameters(r1cs, w)
    bad = dict(r1cs)
    bad["a"] = dict(r1cs["a"]); bad["a"]["col"] = r1cs["a"]["col"].copy(); bad["a"]["col"][0] = 10**6
    with pytest.raises(ctg.CtgError):
        ctx.generate_parameters(bad, util.trapdoor_words(td))


# ----------------------------------------------------------------------------- distributed witness map
@pytest.mark.parametrize("log_n,shards", [(6, 1), (11, 3)])
def test_prove_partial_begin_finish_with_chain_masks(ctx, log_n, shards):
    """ctg_prove_partial_begin / _finish: (1) all chains on this rank; (2) chains computed in two separate begins (masks 0b001
    and 0b110, what two ranks would do) and stitched together.  Partial sums are projective (their representation depends on
    the order atomics filled the buckets in), so the comparison is on the combined, affine proof."""
    import torch
    r1cs, zm, _ = S.mul_chain(log_n)
    pk, _ = S.progression_pk(ctx, r1cs)
    dr = ctx.upload_r1cs(r1cs)
    r, s = fm(rnd.randrange(B.R_MOD)), fm(rnd.randrange(B.R_MOD))
    full = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"])
    want = ctx.prove(full, dr, zm, r, s)
    full.free()
    dev = torch.device("cuda", 0)
    abc = torch.zeros((3, dr.n, 8), dtype=torch.int32, device=dev)
    whole, stitched, chainless = [], [], []
    for index in range(shards):
        dpk = ctx.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"], shard=(index, shards))
        abc.zero_(); torch.cuda.synchronize()
        ctx.prove_partial_begin(dpk, dr, zm, r, s, 7, abc.data_ptr())
        whole.append(ctx.prove_partial_finish(dpk, dr, r, s, abc.data_ptr()))
        # rank "0": chain a only
        abc.zero_(); torch.cuda.synchronize()
        ctx.prove_partial_begin(dpk, dr, zm, r, s, 0b001, abc.data_ptr())
        a_chain = abc[0].clone()
        assert bool(a_chain.any()) and not bool(abc[1:].any())
        torch.cuda.synchronize()
        ctx.prove_partial_finish(dpk, dr, r, s, abc.data_ptr())   # drains the MSMs begin started; result unused
        # rank "1": chains b and c, then the "broadcast" of a
        abc.zero_(); torch.cuda.synchronize()
        ctx.prove_partial_begin(dpk, dr, zm, r, s, 0b110, abc.data_ptr())
        assert not bool(abc[0].any()) and bool(abc[1].any()) and bool(abc[2].any())
        abc[0].copy_(a_chain)
        torch.cuda.synchronize()
        all_chains = abc.clone()
        stitched.append(ctx.prove_partial_finish(dpk, dr, r, s, abc.data_ptr()))
        # rank "3+": no chain of its own (and only its slice of z is uploaded); everything arrives by broadcast
        abc.zero_(); torch.cuda.synchronize()
        ctx.prove_partial_begin(dpk, dr, zm, r, s, 0, abc.data_ptr())
        assert not bool(abc.any())
        abc.copy_(all_chains)
        torch.cuda.synchronize()
        chainless.append(ctx.prove_partial_finish(dpk, dr, r, s, abc.data_ptr()))
        with pytest.raises(ctg.CtgError):
            ctx.prove_partial_begin(dpk, dr, zm, r, s, 8, abc.data_ptr())
        dpk.free()
    for parts in (whole, stitched, chainless):
        got = ctx.prove_combine(np.stack(parts), r, s)
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
    dr.free()


@pytest.mark.parametrize("log_n", [20])
def test_prove_full_size_generated_key_closed_form(ctx, log_n):
    """BASELINE.json configs[2] size (2^20-constraint synthetic R1CS): key GENERATED on the device, proof on the device, and
    the result pinned by the trapdoor closed form -- three scalar multiplications whose exponents come from O(nnz) big-int
    arithmetic (tests/util.py: closed_form_proof_arrays), with no MSM / NTT / key oracle in between."""
    r1cs, zm, z = S.mul_chain(log_n)
    td_words, td = S.trapdoor(seed=0x5EED + log_n)
    dp = ctx.generate_parameters(r1cs, td_words)
    dpk = dp.to_device_pk()
    dp.free()
    dr = ctx.upload_r1cs(r1cs)
    r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    got = ctx.prove(dpk, dr, zm, fm(r), fm(s))
    dpk.free(); dr.free()
    assert util.proof_to_py(got) == util.closed_form_proof_arrays(r1cs, td, z, r, s)


def test_setup_prove_verify_circuit_without_witnesses(ctx):
    """Edge case: every variable is public (num_witness = 0), so l_query is empty and the L MSM has no points; one constraint
    row is empty in C.  Setup and prove on the device against oracle/py, then the pairing equation."""
    from oracle.py import pairing as PR
    cs = G.R1CS(num_instance=3, num_witness=0)           # z = (1, x, y):  x * x = y,  (x - 3) * 0 = 0
    cs.a = [[(1, 1)], [(1, 1), (B.R_MOD - 3, 0)]]
    cs.b = [[(1, 1)], []]
    cs.c = [[(1, 2)], []]
    z = [1, 3, 9]
    assert cs.is_satisfied(z)
    td = _trapdoor()
    r1 = ark.r1cs_from_py(cs)
    want_pk = G.generate_parameters(cs, td)
    dp = ctx.generate_parameters(r1, util.trapdoor_words(td))
    got = dp.download()
    for k, w in ark.pk_from_py(want_pk).items():
        assert np.array_equal(got[k].reshape(w.shape), w), k
    assert got["l_query"].shape == (0, 24)
    dpk, dr = dp.to_device_pk(), ctx.upload_r1cs(r1)
    r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    proof = util.proof_to_py(ctx.prove(dpk, dr, ark.fr_mont_from_ints(z), fm(r), fm(s)))
    assert proof == G.create_proof(cs, want_pk, z, r, s)
    assert PR.groth16_verify(want_pk.alpha_g1, want_pk.beta_g2, want_pk.gamma_g2, want_pk.delta_g2, want_pk.gamma_abc_g1, proof, z[1:])
    image = dp.serialize_unchecked()
    dp2 = ctx.load_proving_key(image)
    assert dp2.serialize_unchecked() == image and (dp2.ni, dp2.nw) == (3, 0)
    dp2.free(); dpk.free(); dr.free(); dp.free()


@pytest.mark.parametrize("shape", ["n1-empty", "n2-one-constraint", "n4-ragged"])
def test_setup_prove_tiny_domains(ctx, shape):
    """Smallest domains (n = 1, 2, 4): empty h_query (n - 1 = 0), no constraints at all, and a domain that the constraint count
    does not fill (zero padding rows).  Key and proof against oracle/py."""
    if shape == "n1-empty":
        cs, z = G.R1CS(num_instance=1, num_witness=0), [1]
    elif shape == "n2-one-constraint":
        cs, z = G.R1CS(num_instance=1, num_witness=1), [1, 1]
        cs.a, cs.b, cs.c = [[(1, 1)]], [[(1, 1)]], [[(1, 1)]]          # w * w = w
    else:
        cs, z = G.R1CS(num_instance=2, num_witness=1), [1, 25, 5]      # n = next_pow2(1 + 2) = 4, one padding row
        cs.a, cs.b, cs.c = [[(1, 2)]], [[(1, 2)]], [[(1, 1)]]          # w * w = x
    assert cs.is_satisfied(z)
    td = _trapdoor()
    r1 = ark.r1cs_from_py(cs)
    want_pk = G.generate_parameters(cs, td)
    dp = ctx.generate_parameters(r1, util.trapdoor_words(td))
    got = dp.download()
    for k, w in ark.pk_from_py(want_pk).items():
        assert np.array_equal(got[k].reshape(w.shape), w), k
    dpk, dr = dp.to_device_pk(), ctx.upload_r1cs(r1)
    r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    proof = util.proof_to_py(ctx.prove(dpk, dr, ark.fr_mont_from_ints(z), fm(r), fm(s)))
    assert proof == G.create_proof(cs, want_pk, z, r, s)
    dpk.free(); dr.free(); dp.free()


@pytest.mark.parametrize("world", [2, 5])
def test_prove_weighted_shards_combine(ctx, world):
    """ctg_pk_upload_range with the weighted bounds the multi-GPU path uses (NTT-chain owners keep fewer points): the shards'
    partial sums combine to the whole-key proof; malformed ranges are refused."""
    from ctg import sharding
    r1cs, zm, _ = S.mul_chain(10)
    td_words, _ = S.trapdoor(seed=77)
    dp = ctx.generate_parameters(r1cs, td_words)
    dr = ctx.upload_r1cs(r1cs)
    r, s = fm(rnd.randrange(B.R_MOD)), fm(rnd.randrange(B.R_MOD))
    whole = dp.to_device_pk(shard=(0, 1, 1))          # the full range is a whole key: ctg_prove accepts it
    want = ctx.prove(whole, dr, zm, r, s)
    whole.free()
    parts = []
    for rank in range(world):
        b = sharding.shard_bounds(rank, world, True, 0.033)
        sh = dp.to_device_pk(shard=b)
        parts.append(ctx.prove_partial(sh, dr, zm, r, s))
        if rank == 0:
            with pytest.raises(ctg.CtgError):
                ctx.prove(sh, dr, zm, r, s)            # a partial key cannot produce a whole proof
        sh.free()
    got = ctx.prove_combine(np.stack(parts), r, s)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    for bad in ((3, 2, 8), (0, 9, 8), (0, 1, 0)):
        with pytest.raises(ctg.CtgError):
            dp.to_device_pk(shard=bad)
    dr.free(); dp.free()
