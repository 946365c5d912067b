"""CPU tests: the two oracles (oracle/py big-int, oracle/c arkworks-algorithm C++) pin each
other and the Groth16 trapdoor closed form.  The reference has no vectors for this path
(SURVEY 8(c)); these are the pins we manufacture."""
import random

import numpy as np
import pytest

from oracle import ark
from oracle.py import bls12_381 as B
from oracle.py import groth16 as G

rnd = random.Random(20261104)


def test_constants_selfcheck():
    assert B.selfcheck()
    assert ark.g1_on_curve(ark.g1_generator()) and ark.g2_on_curve(ark.g2_generator())


def test_field_ops_c_vs_py():
    n = 64
    a = [rnd.randrange(B.R_MOD) for _ in range(n)] + [0, 1, B.R_MOD - 1]
    b = [rnd.randrange(B.R_MOD) for _ in range(n)] + [B.R_MOD - 1, B.R_MOD - 1, B.R_MOD - 1]
    am, bm = ark.fr_mont_from_ints(a), ark.fr_mont_from_ints(b)
    # the Montgomery image is x * 2^256 mod r (ark-ff Fp256)
    assert ark.words_to_ints(am) == [x * B.FR_R % B.R_MOD for x in a]
    assert ark.fr_ints_from_mont(ark.fr_mul(am, bm)) == [x * y % B.R_MOD for x, y in zip(a, b)]
    assert ark.fr_ints_from_mont(ark.fr_add(am, bm)) == [(x + y) % B.R_MOD for x, y in zip(a, b)]
    assert ark.fr_ints_from_mont(ark.fr_sub(am, bm)) == [(x - y) % B.R_MOD for x, y in zip(a, b)]
    a = [rnd.randrange(B.P_MOD) for _ in range(n)] + [0, 1, B.P_MOD - 1]
    b = [rnd.randrange(B.P_MOD) for _ in range(n)] + [B.P_MOD - 1] * 3
    am, bm = ark.fq_to_mont(ark.ints_to_words(a, 12)), ark.fq_to_mont(ark.ints_to_words(b, 12))
    assert ark.words_to_ints(am) == [x * B.FQ_R % B.P_MOD for x in a]
    assert ark.words_to_ints(ark.fq_from_mont(ark.fq_mul(am, bm))) == [x * y % B.P_MOD for x, y in zip(a, b)]
    assert ark.words_to_ints(ark.fq_from_mont(ark.fq_sub(am, bm))) == [(x - y) % B.P_MOD for x, y in zip(a, b)]


@pytest.mark.parametrize("log_n", [0, 1, 2, 5, 10])
@pytest.mark.parametrize("inverse,coset", [(False, False), (True, False), (False, True), (True, True)])
def test_ntt_c_vs_py(log_n, inverse, coset):
    n = 1 << log_n
    x = [rnd.randrange(B.R_MOD) for _ in range(n)]
    want = B.ntt(x, inverse=inverse, coset=coset)
    got = ark.fr_ints_from_mont(ark.ntt_fr(ark.fr_mont_from_ints(x), inverse=inverse, coset=coset))
    assert got == want


def test_ntt_is_polynomial_evaluation():
    n = 32
    x = [rnd.randrange(B.R_MOD) for _ in range(n)]
    w = B.fr_root_of_unity(5)
    X = ark.fr_ints_from_mont(ark.ntt_fr(ark.fr_mont_from_ints(x)))
    Xc = ark.fr_ints_from_mont(ark.ntt_fr(ark.fr_mont_from_ints(x), coset=True))
    for j in (0, 1, 7, 31):
        assert X[j] == B.poly_eval(x, pow(w, j, B.R_MOD))
        assert Xc[j] == B.poly_eval(x, 7 * pow(w, j, B.R_MOD) % B.R_MOD)


def test_group_ops_c_vs_py():
    k = rnd.randrange(B.R_MOD)
    kc = ark.ints_to_words([k], 8)[0]
    assert ark.g1_to_py(ark.g1_mul(ark.g1_generator(), kc))[0] == B.G1.mul(B.G1_GEN, k)
    assert ark.g2_to_py(ark.g2_mul(ark.g2_generator(), kc))[0] == B.G2.mul(B.G2_GEN, k)
    rm1 = ark.ints_to_words([B.R_MOD - 1], 8)[0]
    P = ark.g1_mul(ark.g1_generator(), rm1)
    assert ark.g1_to_py(ark.g1_add(P, ark.g1_generator()))[0] is None  # (r-1)G + G = O


@pytest.mark.parametrize("kind", ["uniform", "poe", "equal", "rminus1"])
def test_msm_g1_c_vs_py_and_dlog(kind):
    n = 96
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    bases = ark.g1_progression(ark.g1_generator(), ark.fr_mont_from_ints([a])[0], ark.fr_mont_from_ints([b])[0], n)
    bases[5] = 0  # an identity base, as in b_g1_query
    sc = ark.random_scalars_canon(n, 7, kind)
    got = ark.g1_to_py(ark.msm_g1(bases, sc))[0]
    s_int = ark.words_to_ints(sc)
    dl = sum(s * (a + i * b) for i, s in enumerate(s_int) if i != 5) % B.R_MOD
    assert got == B.G1.mul(B.G1_GEN, dl)
    if kind in ("uniform", "poe"):
        assert got == B.G1.msm(ark.g1_to_py(bases), s_int)


def test_msm_g2_c_vs_py():
    n = 40
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    bases = ark.g2_progression(ark.g2_generator(), ark.fr_mont_from_ints([a])[0], ark.fr_mont_from_ints([b])[0], n)
    sc = ark.random_scalars_canon(n, 9, "poe")
    got = ark.g2_to_py(ark.msm_g2(bases, sc))[0]
    s_int = ark.words_to_ints(sc)
    assert got == B.G2.mul(B.G2_GEN, sum(s * (a + i * b) for i, s in enumerate(s_int)) % B.R_MOD)
    assert got == B.G2.msm_naive(ark.g2_to_py(bases), s_int)


def test_fixed_base_matches_scalar_mul():
    ks = [0, 1, B.R_MOD - 1] + [rnd.randrange(B.R_MOD) for _ in range(5)]
    out = ark.g1_to_py(ark.g1_fixed_base(ark.g1_generator(), ark.fr_mont_from_ints(ks)))
    assert out == [B.G1.mul(B.G1_GEN, k) for k in ks]
    out2 = ark.g2_to_py(ark.g2_fixed_base(ark.g2_generator(), ark.fr_mont_from_ints(ks[:4])))
    assert out2 == [B.G2.mul(B.G2_GEN, k) for k in ks[:4]]


def _trapdoor():
    return G.Trapdoor(*[rnd.randrange(1, B.R_MOD) for _ in range(5)],
                      B.G1.mul(B.G1_GEN, rnd.randrange(1, B.R_MOD)), B.G2.mul(B.G2_GEN, rnd.randrange(1, B.R_MOD)))


@pytest.mark.parametrize("log_n,r_zero", [(3, False), (5, False), (4, True)])
def test_groth16_py_c_closed_form(log_n, r_zero):
    cs, z = G.mul_chain(log_n, rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD))
    assert cs.is_satisfied(z)
    td = _trapdoor()
    pk = G.generate_parameters(cs, td)
    r = 0 if r_zero else rnd.randrange(B.R_MOD)
    s = rnd.randrange(B.R_MOD)
    want = G.create_proof(cs, pk, z, r, s)
    if not r_zero:
        assert want == G.closed_form_proof(cs, td, z, r, s)
    # witness map, C++ vs python
    r1 = ark.r1cs_from_py(cs)
    zm = ark.fr_mont_from_ints(z)
    assert ark.fr_ints_from_mont(ark.witness_map(r1, zm)) == G.witness_map(cs, z)
    a, b, c = ark.create_proof(ark.pk_from_py(pk), r1, zm, ark.fr_mont_from_ints([r])[0], ark.fr_mont_from_ints([s])[0])
    assert (ark.g1_to_py(a)[0], ark.g2_to_py(b)[0], ark.g1_to_py(c)[0]) == want


def test_lagrange_coeffs_c_vs_py():
    tau = rnd.randrange(B.R_MOD)
    got = ark.fr_ints_from_mont(ark.lagrange_coeffs(4, ark.fr_mont_from_ints([tau])[0]))
    assert got == G.lagrange_coeffs(16, tau)


# ----------------------------------------------------------------------------- pairing / Groth16 verification (oracle/py/pairing.py)
def test_pairing_bilinear_nondegenerate_order_r():
    from oracle.py import pairing as PR
    e = PR.pairing(B.G1_GEN, B.G2_GEN)
    assert e != PR.F12_ONE
    assert PR.f12_pow(e, B.R_MOD) == PR.F12_ONE
    a, b = 0x1234567, 0x7654321ABC
    assert PR.pairing(B.G1.mul(B.G1_GEN, a), B.G2.mul(B.G2_GEN, b)) == PR.f12_pow(e, a * b % B.R_MOD)
    assert PR.pairing(None, B.G2_GEN) == PR.F12_ONE
    x = (1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12)
    assert PR.f12_mul(x, PR.f12_inv(x)) == PR.F12_ONE


def test_oracle_setup_prove_verify_accepts_and_rejects():
    """The reference's own test shape (circuits-ark/src/encryption.rs:403-411) on the oracle alone: the restated
    generate_parameters + create_proof produce a proof the Groth16 pairing equation accepts, so the oracle's conventions
    (query offsets, instance rows, h_query/l_query scaling, r/s blinding) form a sound Groth16 system."""
    from oracle.py import pairing as PR
    from tests import util
    cs, z = G.mul_chain(3, 5, 7)
    td = G.Trapdoor(alpha=11111, beta=22222, gamma=33333, delta=44444, tau=55555, g1=B.G1.mul(B.G1_GEN, 97), g2=B.G2.mul(B.G2_GEN, 89))
    pk = G.generate_parameters(cs, td)
    pr = G.create_proof(cs, pk, z, 123456789, 987654321)
    pub = z[1:cs.num_instance]
    assert PR.groth16_verify(pk.alpha_g1, pk.beta_g2, pk.gamma_g2, pk.delta_g2, pk.gamma_abc_g1, pr, pub)
    assert not PR.groth16_verify(pk.alpha_g1, pk.beta_g2, pk.gamma_g2, pk.delta_g2, pk.gamma_abc_g1, pr, [(pub[0] + 1) % B.R_MOD])
    assert not PR.groth16_verify(pk.alpha_g1, pk.beta_g2, pk.gamma_g2, pk.delta_g2, pk.gamma_abc_g1, (pr[0], pr[1], pr[0]), pub)
    # the C++ restatement's key generator agrees member for member (incl. gamma_g2 / gamma_abc_g1)
    want = ark.pk_from_py(pk)
    got = util.trapdoor_pk_arrays(cs, td)
    for k, w in want.items():
        assert np.array_equal(np.asarray(got[k]).reshape(w.shape), w), k


def test_closed_form_proof_arrays_matches_bigint_oracle():
    """tests/util.py's O(nnz) closed form (used at 2^20 on the GPU) against oracle/py's on a small circuit."""
    from tests import util
    cs, z = G.mul_chain(4, rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD))
    td = G.Trapdoor(*[rnd.randrange(1, B.R_MOD) for _ in range(5)], B.G1_GEN, B.G2_GEN)
    r, s = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    got = util.closed_form_proof_arrays(ark.r1cs_from_py(cs), dict(alpha=td.alpha, beta=td.beta, delta=td.delta, tau=td.tau), z, r, s)
    assert got == G.closed_form_proof(cs, td, z, r, s)


def test_dedicated_squaring_equals_product():
    """oracle/c squares with ark-ff's square_in_place schedule (off-diagonal once, doubled); it must equal a * a on edge and random values."""
    rnd2 = random.Random(99)
    for mod, w, sqr, mul in ((B.R_MOD, 8, ark.fr_sqr, ark.fr_mul), (B.P_MOD, 12, ark.fq_sqr, ark.fq_mul)):
        vals = [0, 1, 2, mod - 1, mod - 2, (mod - 1) // 2, (1 << (32 * w - 1)) % mod, (1 << 64) - 1] + [rnd2.randrange(mod) for _ in range(2000)]
        a = ark.ints_to_words(vals, w)
        assert np.array_equal(sqr(a), mul(a, a))
        R = 1 << (32 * w)
        assert ark.words_to_ints(sqr(a)) == [v * v * pow(R, -1, mod) % mod for v in vals]


@pytest.mark.parametrize("log_n", [0, 1, 5, 12])
def test_poly_eval_is_the_transform_at_domain_points(log_n):
    """ark_fr_poly_eval (Horner) against big-int evaluation, and against the FFT at w^j and 7 w^j: the check the large GPU NTT tests rely on."""
    n = 1 << log_n
    x = ark.random_fr_mont(n, 5 + log_n)
    xi = ark.fr_ints_from_mont(x)
    w = B.fr_root_of_unity(log_n)
    js = sorted({0, n - 1, n // 2, 3 % n})
    pts = [pow(w, j, B.R_MOD) for j in js] + [7 * pow(w, j, B.R_MOD) % B.R_MOD for j in js] + [12345]
    got = ark.fr_ints_from_mont(ark.fr_poly_eval(x, ark.fr_mont_from_ints(pts)))
    assert got == [sum(c * pow(p, i, B.R_MOD) for i, c in enumerate(xi)) % B.R_MOD for p in pts]
    fwd, cos = ark.fr_ints_from_mont(ark.ntt_fr(x)), ark.fr_ints_from_mont(ark.ntt_fr(x, coset=True))
    assert got[:len(js)] == [fwd[j] for j in js] and got[len(js):2 * len(js)] == [cos[j] for j in js]
