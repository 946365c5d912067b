"""Multi-device contexts (ctg_ctx_create with n_devices > 1; csrc/multi.cu): one process, the prover-path calls shard over the
devices by themselves and must return exactly what a one-device context returns.

The driver's GPU test box has ONE GPU, so every case runs with repeated device ids (the shards then share the GPU: same code
path -- worker threads, per-shard contexts, cross-stream events, cudaMemcpyPeerAsync -- without the speed-up); when the box has
more GPUs the same cases also run over distinct devices.
"""
import random

import numpy as np
import pytest

import ctg
from ctg import synthetic as S
from oracle import ark
from oracle.py import bls12_381 as B
from tests import util

pytestmark = pytest.mark.gpu
rnd = random.Random(808)


def fm(v):
    return ark.fr_mont_from_ints([v])[0]


def _device_sets():
    import torch
    have = torch.cuda.device_count()
    sets = [[0, 0], [0, 0, 0], [0, 0, 0, 0, 0]]
    if have >= 2:
        sets.append([0, 1])
    if have >= 4:
        sets.append([0, 1, 2, 3])
    if have >= 8:
        sets.append(list(range(8)))
    return sets


@pytest.fixture(scope="module")
def one():
    c = ctg.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module", params=_device_sets(), ids=lambda s: "dev" + "".join(map(str, s)))
def multi(request):
    c = ctg.Context(request.param)
    yield c
    c.close()


def test_device_count_and_primary_device_calls(multi):
    assert ctg.load().ctg_ctx_device_count(multi.h) == len(multi.devices)
    x = ark.random_fr_mont(1 << 9, 3)
    assert np.array_equal(multi.ntt(x, coset=True), ark.ntt_fr(x, coset=True))        # unsharded calls run on the primary device


@pytest.mark.parametrize("n,kind", [(1, "uniform"), (7, "poe"), (5003, "poe"), (1 << 14, "uniform")])
def test_sharded_msm_matches_oracle(multi, n, kind):
    r = random.Random(n)
    pts = ark.g1_progression(ark.g1_generator(), fm(r.randrange(B.R_MOD)), fm(r.randrange(B.R_MOD)), n)
    if n > 5:
        pts[5] = 0
    sc = ark.random_scalars_canon(n, n + 1, kind)
    t = multi.bases_g1(pts)
    assert len(t) == n
    assert np.array_equal(t.download(), pts)
    want = ark.msm_g1(pts, sc)
    assert np.array_equal(multi.msm(t, sc), want)
    assert np.array_equal(multi.msm(t, ark.fr_to_mont(sc), mont=True), want)
    k = max(1, n // 3)
    assert np.array_equal(multi.msm(t, sc[:k]), ark.msm_g1(pts[:k], sc[:k]))           # a prefix: trailing devices get nothing
    assert not multi.msm(t, sc[:0]).any()
    t.free()
    if n <= 5003:
        p2 = ark.g2_progression(ark.g2_generator(), fm(3 + n), fm(5), min(n, 700))
        s2 = ark.random_scalars_canon(p2.shape[0], 19, "uniform")
        t2 = multi.bases_g2(p2)
        assert np.array_equal(multi.msm(t2, s2), ark.msm_g2(p2, s2))
        t2.free()


def test_sharded_msm_device_scalars_and_progression(multi):
    """Scalars resident on the primary device, bases generated per device from a + lo b: the dlog closed form at 2^18."""
    import torch
    n = 1 << 18
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    t = multi.bases_progression_g1(ark.g1_generator(), fm(a), fm(b), n)
    first = t.download(0, 3)
    assert np.array_equal(first, ark.g1_progression(ark.g1_generator(), fm(a), fm(b), 3))
    last = t.download(n - 2, 2)
    assert np.array_equal(last, ark.g1_progression(ark.g1_generator(), fm((a + (n - 2) * b) % B.R_MOD), fm(b), 2))
    sc = np.random.default_rng(5).integers(0, 1 << 32, size=(n, 8), dtype=np.uint64).astype(np.uint32)
    sc[:, 7] &= 0x3FFFFFFF
    s0, s1 = util.weighted_sums(sc)
    want = B.G1.mul(B.G1_GEN, (a * s0 + b * s1) % B.R_MOD)
    dev = torch.device("cuda", multi.devices[0])
    d_sc = torch.from_numpy(sc.view(np.int32)).to(dev)
    d_out = torch.zeros(48, dtype=torch.int32, device=dev)
    d_aff = torch.zeros(24, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    multi.msm_dev(t, d_sc.data_ptr(), n, d_out.data_ptr())
    multi.combine_dev(d_out.data_ptr(), 1, d_aff.data_ptr())
    multi.sync()
    assert ark.g1_to_py(d_aff.cpu().numpy().view(np.uint32))[0] == want
    assert ark.g1_to_py(multi.msm(t, sc))[0] == want
    t.free()


@pytest.mark.parametrize("shape", ["mulchain4", "mulchain9", "mulchain13", "poe100"])
def test_sharded_prove_equals_one_device_proof(one, multi, shape):
    if shape.startswith("mulchain"):
        r1cs, zm, z = S.mul_chain(int(shape[8:]))
    else:
        r1cs, zm, z = S.poe_shape(100)
    pk, _ = S.progression_pk(one, r1cs)
    r, s = fm(rnd.randrange(B.R_MOD)), fm(rnd.randrange(B.R_MOD))
    dpk1 = one.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"])
    dr1 = one.upload_r1cs(r1cs)
    want = one.prove(dpk1, dr1, zm, r, s)
    dpk = multi.upload_pk(pk, r1cs["num_instance"], r1cs["num_constraints"])
    dr = multi.upload_r1cs(r1cs)
    for _ in range(2):                                   # twice: buffers and events are reused from the second call on
        got, tm = multi.prove(dpk, dr, zm, r, s, timed=True)
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
        assert tm["total_ms"] > 0
    if shape in ("mulchain4", "mulchain9"):
        for g, w in zip(got, ark.create_proof(pk, r1cs, zm, r, s)):
            assert np.array_equal(g, w)
    assert np.array_equal(multi.witness_map(dr, zm), one.witness_map(dr1, zm))
    # a batch through the sharded key: each proof equals its one-device proof
    zs = np.stack([zm, zm])
    rs = np.stack([np.stack([r, s]), np.stack([s, r])])
    out = multi.prove_batch(dpk, dr, zs, rs)
    assert np.array_equal(out[0], np.concatenate(want))
    assert np.array_equal(out[1], np.concatenate(one.prove(dpk1, dr1, zm, s, r)))
    # handles do not cross contexts
    with pytest.raises(ctg.CtgError):
        one.prove(dpk, dr1, zm, r, s)
    with pytest.raises(ctg.CtgError):
        multi.prove(dpk1, dr1, zm, r, s)
    with pytest.raises(ctg.CtgError):
        multi.prove_partial(dpk, dr, zm, r, s)
    dpk.free(); dr.free(); dpk1.free(); dr1.free()


def test_sharded_prove_generated_key_2_16_closed_form(multi):
    """Key generated on the primary device, sharded device-to-device, proof over all devices, pinned by the trapdoor closed form."""
    log_n = 16
    r1cs, zm, z = S.mul_chain(log_n)
    td_words, td = S.trapdoor(seed=0x5EED + log_n)
    dp = multi.generate_parameters(r1cs, td_words)
    dpk = dp.to_device_pk()
    dp.free()
    dr = multi.upload_r1cs(r1cs)
    r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    zp = multi.pinned_like(zm)
    got = multi.prove(dpk, dr, zp, fm(r), fm(s))
    multi.pinned_free(zp)
    dpk.free(); dr.free()
    assert util.proof_to_py(got) == util.closed_form_proof_arrays(r1cs, td, z, r, s)
