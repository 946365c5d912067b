"""GPU parity at the sizes BASELINE.json quotes the metric on (2^22, 2^24; G2 at 2^20, 2^22), through the C ABI, against checks
that do not need a CPU MSM / NTT of that size ("Next round" item 1):

  * MSMs over progression bases P_i = (a + i b) G generated on the device: the result must be (sum_i s_i (a + i b)) G -- one
    scalar multiplication of the generator by a number obtained with O(n) machine arithmetic (tests/util.py weighted_sums);
  * NTTs against something that is not an NTT: Horner evaluation (oracle/c ark_fr_poly_eval, pinned against big-int evaluation in
    tests/test_oracle.py) of the input at w^j / 7 w^j for 16 random j per mode, plus oracle/c bit-exact at 2^22;
  * a whole 2^22 Groth16 proof (key generated on the device) against the trapdoor closed form (tests/util.py
    closed_form_proof_arrays: O(nnz) big-int work, no MSM / NTT / key oracle in between).
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
rnd = random.Random(20261104)


@pytest.fixture(scope="module")
def ctx():
    c = ctg.Context(0)
    yield c
    c.close()


def fm(v):
    return ark.fr_mont_from_ints([v])[0]


def _scalars(n, seed, kind):
    """uniform: 254-bit canonical scalars; poe: the witness mix of SURVEY 8(d) (zeros, ones, bytes, full width)."""
    g = np.random.default_rng(seed)
    sc = g.integers(0, 1 << 32, size=(n, 8), dtype=np.uint64).astype(np.uint32)
    sc[:, 7] &= 0x3FFFFFFF
    if kind == "poe":
        sel = g.random(n)
        small = sel < 0.80
        sc[small, 1:] = 0
        sc[sel < 0.40, 0] = 0
        sc[(sel >= 0.40) & (sel < 0.65), 0] = 1
        sc[(sel >= 0.65) & small, 0] &= 0xFF
    return sc


# ----------------------------------------------------------------------------- MSM
@pytest.mark.parametrize("log_n,kind", [(22, "uniform"), (22, "poe"), (24, "uniform")])
def test_msm_g1_dlog_closed_form_baseline_sizes(ctx, log_n, kind):
    """The headline workload of bench.py (2^22 uniform scalars, default plan) and 2^24, host-scalar and device-scalar entry points."""
    import torch
    n = 1 << log_n
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    t = ctx.bases_progression_g1(ark.g1_generator(), fm(a), fm(b), n)
    sc = _scalars(n, log_n, kind)
    s0, s1 = util.weighted_sums(sc)
    want = B.G1.mul(B.G1_GEN, (a * s0 + b * s1) % B.R_MOD)
    assert ark.g1_to_py(ctx.msm(t, sc))[0] == want                      # ctg_msm_g1: pageable host scalars, split schedule
    dev = torch.device("cuda", 0)
    d_sc = torch.from_numpy(sc.view(np.int32)).to(dev)
    d_out = torch.zeros(48, dtype=torch.int32, device=dev)
    d_aff = torch.zeros(24, dtype=torch.int32, device=dev)
    ctx.msm_dev(t, d_sc.data_ptr(), n, d_out.data_ptr())                 # ctg_msm_g1_dev: what bench.py's `value` times
    ctx.combine_dev(d_out.data_ptr(), 1, d_aff.data_ptr())
    ctx.sync()
    torch.cuda.synchronize()
    assert ark.g1_to_py(d_aff.cpu().numpy().view(np.uint32))[0] == want
    # a strict prefix (min(len) semantics of VariableBaseMSM) at full size
    k = n - 12345
    p0, p1 = util.weighted_sums(sc[:k])
    assert ark.g1_to_py(ctx.msm(t, sc[:k]))[0] == B.G1.mul(B.G1_GEN, (a * p0 + b * p1) % B.R_MOD)
    t.free()


@pytest.mark.parametrize("log_n,kind", [(20, "uniform"), (20, "poe"), (22, "uniform")])
def test_msm_g2_dlog_closed_form_baseline_sizes(ctx, log_n, kind):
    n = 1 << log_n
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    t = ctx.bases_progression_g2(ark.g2_generator(), fm(a), fm(b), n)
    sc = _scalars(n, 100 + log_n, kind)
    s0, s1 = util.weighted_sums(sc)
    assert ark.g2_to_py(ctx.msm(t, sc))[0] == B.G2.mul(B.G2_GEN, (a * s0 + b * s1) % B.R_MOD)
    t.free()


# ----------------------------------------------------------------------------- NTT
@pytest.mark.parametrize("log_n", [22, 24])
def test_ntt_against_horner_evaluation_baseline_sizes(ctx, log_n):
    """fft(x)[j] = x(w^j), coset_fft(x)[j] = x(7 w^j); ifft(x) = the coefficients c with c(w^j) = x[j]; coset_ifft likewise at 7 w^j."""
    n = 1 << log_n
    x = _scalars(n, 7 + log_n, "uniform")            # any 254-bit pattern is a valid Montgomery image
    w = B.fr_root_of_unity(log_n)
    js = sorted({0, 1, n - 1, n // 2, n // 2 + 1} | {rnd.randrange(n) for _ in range(11)})
    pts = ark.fr_mont_from_ints([pow(w, j, B.R_MOD) for j in js])
    cpts = ark.fr_mont_from_ints([7 * pow(w, j, B.R_MOD) % B.R_MOD for j in js])
    want_f, want_c = ark.fr_poly_eval(x, pts), ark.fr_poly_eval(x, cpts)
    fwd = ctx.ntt(x)
    assert np.array_equal(fwd[js], want_f)
    cos = ctx.ntt(x, coset=True)
    assert np.array_equal(cos[js], want_c)
    inv = ctx.ntt(x, inverse=True)
    assert np.array_equal(ark.fr_poly_eval(inv, pts), x[js])
    cinv = ctx.ntt(x, inverse=True, coset=True)
    assert np.array_equal(ark.fr_poly_eval(cinv, cpts), x[js])
    # and the transforms invert each other on the whole vector
    assert np.array_equal(ctx.ntt(fwd, inverse=True), x)
    assert np.array_equal(ctx.ntt(cos, inverse=True, coset=True), x)
    if log_n == 22:
        assert np.array_equal(fwd, ark.ntt_fr(x))                           # oracle/c (ark-poly's algorithm), every element
        assert np.array_equal(cinv, ark.ntt_fr(x, inverse=True, coset=True))


# ----------------------------------------------------------------------------- prove
def test_prove_2_22_generated_key_closed_form(ctx):
    """BASELINE.json configs[3] size on one GPU: setup and prove on the device, pinned by the trapdoor closed form."""
    log_n = 22
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
