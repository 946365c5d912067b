"""CPU tests of the DEVICE field/curve headers compiled for the host (csrc/ptx.cuh emulates
the PTX carry chain): the exact mont_mul / XYZZ code the kernels run, against the oracle."""
import ctypes as C
import os
import random
import subprocess

import numpy as np
import pytest

from oracle import ark
from oracle.py import bls12_381 as B

HERE = os.path.dirname(os.path.abspath(__file__))
rnd = random.Random(99)


@pytest.fixture(scope="module")
def hm():
    subprocess.check_call(["make", "-C", os.path.join(HERE, "host_math")], stdout=subprocess.DEVNULL)
    return C.CDLL(os.path.join(HERE, "host_math", "libhost_math.so"))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _edge(mod, n):
    vals = [0, 1, 2, mod - 1, mod - 2, (1 << 255) % mod, (mod - 1) // 2, 0xFFFFFFFF, 1 << 32, (1 << 64) - 1]
    vals += [rnd.randrange(mod) for _ in range(n)]
    return vals


@pytest.mark.parametrize("field", ["fr", "fq"])
def test_field_ops_match_oracle(hm, field):
    mod, w = (B.R_MOD, 8) if field == "fr" else (B.P_MOD, 12)
    av = _edge(mod, 500)
    bv = list(reversed(_edge(mod, 500)))
    # Montgomery images straight from ints (x*R mod p) -- any value < p is a valid image
    a = ark.ints_to_words(av, w)
    b = ark.ints_to_words(bv, w)
    for op in ("mul", "add", "sub"):
        out = np.zeros_like(a)
        getattr(hm, f"hm_{field}_{op}")(_p(a), _p(b), _p(out), C.c_uint64(len(av)))
        want = getattr(ark, f"{field}_{op}")(a, b)
        assert np.array_equal(out, want), op


def test_fr_mul_special_form_reduction_rows(hm):
    """Fp<FrParams>::redc_row exploits r = ... 2^64 - 2^32 + 1 (m = -E[0], the two low limbs of m * r without a multiplier): operands
    whose limbs are drawn from {0, 1, 2^31, 2^32 - 1} drive the row through m = 0, m = 1, m = 2^32 - 1 and the carry out of limb 1."""
    limbs = [0, 1, 0x80000000, 0xFFFFFFFF]
    vals = [sum(rnd.choice(limbs) << (32 * i) for i in range(8)) % B.R_MOD for _ in range(4000)]
    a = ark.ints_to_words(vals, 8)
    b = ark.ints_to_words(vals[::-1], 8)
    out = np.zeros_like(a)
    hm.hm_fr_mul(_p(a), _p(b), _p(out), C.c_uint64(len(vals)))
    assert np.array_equal(out, ark.fr_mul(a, b))
    one = ark.ints_to_words([1] * len(vals), 8)              # from_mont: rows where only E[0] of the first row is non-zero
    hm.hm_fr_mul(_p(a), _p(one), _p(out), C.c_uint64(len(vals)))
    assert np.array_equal(out, ark.fr_mul(a, one))


@pytest.mark.parametrize("field", ["fr", "fq"])
def test_dedicated_squaring_equals_product(hm, field):
    """Fp::sqr (the symmetric half of the partial products left out, rows multiplied by the limbs of 2a) against a * a from the
    oracle: edge values, limbs drawn from {0, 1, 2^31, 2^32 - 1} (every carry of the doubling and of the skipped positions), random."""
    mod, w = (B.R_MOD, 8) if field == "fr" else (B.P_MOD, 12)
    limbs = [0, 1, 0x80000000, 0xFFFFFFFF, 0x7FFFFFFF]
    vals = _edge(mod, 2000) + [sum(rnd.choice(limbs) << (32 * i) for i in range(w)) % mod for _ in range(3000)]
    a = ark.ints_to_words(vals, w)
    out = np.zeros_like(a)
    getattr(hm, f"hm_{field}_sqr")(_p(a), _p(out), C.c_uint64(len(vals)))
    assert np.array_equal(out, getattr(ark, f"{field}_mul")(a, a))


def test_fr_mont_conversions(hm):
    vals = _edge(B.R_MOD, 50)
    a = ark.ints_to_words(vals, 8)
    out = np.zeros_like(a)
    hm.hm_fr_to_mont(_p(a), _p(out), C.c_uint64(len(vals)))
    assert np.array_equal(out, ark.fr_to_mont(a))
    back = np.zeros_like(a)
    hm.hm_fr_from_mont(_p(out), _p(back), C.c_uint64(len(vals)))
    assert np.array_equal(back, a)


def test_fq2_and_inverse(hm):
    n = 40
    a = ark.ints_to_words([rnd.randrange(B.P_MOD) for _ in range(2 * n)], 12).reshape(n, 24)
    b = ark.ints_to_words([rnd.randrange(B.P_MOD) for _ in range(2 * n)], 12).reshape(n, 24)
    out = np.zeros_like(a)
    hm.hm_fq2_mul(_p(a), _p(b), _p(out), C.c_uint64(n))
    assert np.array_equal(out, ark.fq2_mul(a, b))
    hm.hm_fq2_sqr(_p(a), _p(out), C.c_uint64(n))
    assert np.array_equal(out, ark.fq2_mul(a, a))
    x = a.reshape(-1, 12)[:8].copy()
    inv = np.zeros_like(x)
    hm.hm_fq_inv(_p(x), _p(inv), C.c_uint64(8))
    one = ark.fq_to_mont(ark.ints_to_words([1] * 8, 12))
    assert np.array_equal(ark.fq_mul(x, inv), one)


@pytest.mark.parametrize("field", ["fr", "fq"])
def test_bingcd_inverse_matches_pow(hm, field):
    """inv_bingcd (Pornin's binary GCD, the inversion of the batched-affine MSM) against x^(p-2) in Python."""
    mod, w, R = (B.R_MOD, 8, B.FR_R) if field == "fr" else (B.P_MOD, 12, B.FQ_R)
    vals = _edge(mod, 400) + [(1 << 62) - 1, 1 << 61, 3] + [rnd.randrange(1 << k) % mod for k in range(1, w * 32 - 2, 5)]
    a = ark.ints_to_words(vals, w)
    out = np.zeros_like(a)
    getattr(hm, f"hm_{field}_inv_bingcd")(_p(a), _p(out), C.c_uint64(len(vals)))
    rinv = pow(R, -1, mod)
    want = [0 if v == 0 else pow(v * rinv % mod, -1, mod) * R % mod for v in vals]   # image of the inverse of the value a is the image of
    assert ark.words_to_ints(out) == want


def test_fq2_bingcd_inverse(hm):
    n = 16
    a = ark.ints_to_words([rnd.randrange(B.P_MOD) for _ in range(2 * n)], 12).reshape(n, 24)
    a[3] = 0
    inv = np.zeros_like(a)
    hm.hm_fq2_inv_bingcd(_p(a), _p(inv), C.c_uint64(n))
    one = np.zeros_like(a)
    one[:, :12] = ark.fq_to_mont(ark.ints_to_words([1], 12))[0]
    one[3] = 0
    assert np.array_equal(ark.fq2_mul(a, inv), one)


def test_xyzz_sum_with_special_cases(hm):
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    n = 24
    fm = lambda v: ark.fr_mont_from_ints([v])[0]
    pts = ark.g1_progression(ark.g1_generator(), fm(a), fm(b), n)
    dl = [(a + i * b) % B.R_MOD for i in range(n)]
    # duplicates (forces the doubling path), an inverse pair (P + -P), identity points
    pts[1] = pts[0]; dl[1] = dl[0]
    pts[7] = 0; dl[7] = 0
    neg = np.zeros(n, dtype=np.uint8)
    neg[3] = 1; dl[3] = -dl[3]
    pts[10] = pts[9]; neg[10] = 1; dl[10] = -dl[9]
    out = np.zeros(24, dtype=np.uint32)
    hm.hm_g1_sum(_p(pts), _p(neg), C.c_uint64(n), _p(out))
    assert ark.g1_to_py(out)[0] == B.G1.mul(B.G1_GEN, sum(dl) % B.R_MOD)
    # cancel to infinity: P, P, -P, -P
    four = np.stack([pts[0]] * 4)
    hm.hm_g1_sum(_p(four), _p(np.array([0, 0, 1, 1], dtype=np.uint8)), C.c_uint64(4), _p(out))
    assert not out.any()
    # G2
    p2 = ark.g2_progression(ark.g2_generator(), fm(a), fm(b), 8)
    p2[2] = p2[1]
    out2 = np.zeros(48, dtype=np.uint32)
    hm.hm_g2_sum(_p(p2), _p(np.zeros(8, dtype=np.uint8)), C.c_uint64(8), _p(out2))
    d2 = [(a + i * b) for i in range(8)]; d2[2] = d2[1]
    assert ark.g2_to_py(out2)[0] == B.G2.mul(B.G2_GEN, sum(d2) % B.R_MOD)


def test_xyzz_full_add_tree_and_scalar_mul(hm):
    a, b = rnd.randrange(B.R_MOD), rnd.randrange(B.R_MOD)
    fm = lambda v: ark.fr_mont_from_ints([v])[0]
    pts = ark.g1_progression(ark.g1_generator(), fm(a), fm(b), 16)
    pts[4] = pts[5]  # equal operands in the first tree level -> xyzz_add doubling path
    pts[8] = 0
    out = np.zeros(24, dtype=np.uint32)
    hm.hm_g1_tree(_p(pts), C.c_uint64(16), _p(out))
    dl = [(a + i * b) for i in range(16)]; dl[4] = dl[5]; dl[8] = 0
    assert ark.g1_to_py(out)[0] == B.G1.mul(B.G1_GEN, sum(dl) % B.R_MOD)
    for k in (0, 1, 2, B.R_MOD - 1, rnd.randrange(B.R_MOD)):
        kc = ark.ints_to_words([k], 8)[0]
        hm.hm_g1_mul(_p(ark.g1_generator()), _p(kc), _p(out))
        assert ark.g1_to_py(out)[0] == B.G1.mul(B.G1_GEN, k)
    out2 = np.zeros(48, dtype=np.uint32)
    k = rnd.randrange(B.R_MOD)
    hm.hm_g2_mul(_p(ark.g2_generator()), _p(ark.ints_to_words([k], 8)[0]), _p(out2))
    assert ark.g2_to_py(out2)[0] == B.G2.mul(B.G2_GEN, k)


@pytest.mark.parametrize("c", [3, 8, 11, 13, 15, 16, 17, 20, 23])
def test_signed_digits_recompose(hm, c):
    nwin = (256 + c - 1) // c
    for k in [0, 1, (1 << (c - 1)), (1 << (c - 1)) + 1, B.R_MOD - 1, (1 << 255) - 1] + [rnd.randrange(B.R_MOD) for _ in range(50)]:
        dig = np.zeros(nwin, dtype=np.int32)
        hm.hm_signed_digits(_p(ark.ints_to_words([k], 8)[0]), C.c_int(c), C.c_int(nwin), _p(dig))
        assert all(abs(int(d)) <= 1 << (c - 1) for d in dig)
        assert sum(int(d) << (c * w) for w, d in enumerate(dig)) == k


def test_lazy_sums_of_products(hm):
    """Fq::mul2_add / mul2_sub / Fq2::mul2_sub (one Montgomery reduction for several products) incl. extreme operands."""
    P = B.P_MOD
    Rinv = pow(1 << 384, -1, P)
    edge = [0, 1, P - 1, P - 2, (P - 1) // 2, (1 << 380) % P]
    quads = [(a, b, c, d) for a in edge for b in (0, P - 1, 1) for c in (P - 1, 0, 2) for d in (P - 1, 0, 1)]
    quads += [tuple(rnd.randrange(P) for _ in range(4)) for _ in range(300)]
    arr = [ark.ints_to_words([q[k] for q in quads], 12) for k in range(4)]
    n = len(quads)
    for name, sign in (("hm_fq_mul2_add", 1), ("hm_fq_mul2_sub", -1)):
        out = np.zeros_like(arr[0])
        getattr(hm, name)(*[_p(x) for x in arr], _p(out), C.c_uint64(n))
        want = [((a * b + sign * c * d) * Rinv) % P for a, b, c, d in quads]   # Montgomery images in, Montgomery image out
        assert ark.words_to_ints(out) == want, name
    # Fq2: a*b - c*d with a = (a0, a1) etc.
    m = n // 2 * 2
    f2 = [x[:m].reshape(m // 2, 24) for x in arr]
    out = np.zeros_like(f2[0])
    hm.hm_fq2_mul2_sub(*[_p(np.ascontiguousarray(x)) for x in f2], _p(out), C.c_uint64(m // 2))
    got = ark.words_to_ints(out.reshape(m, 12))
    for i in range(m // 2):
        (a0, b0, c0, d0), (a1, b1, c1, d1) = quads[2 * i], quads[2 * i + 1]
        w0 = ((a0 * b0 - a1 * b1) - (c0 * d0 - c1 * d1)) * Rinv % P
        w1 = ((a0 * b1 + a1 * b0) - (c0 * d1 + c1 * d0)) * Rinv % P
        assert (got[2 * i], got[2 * i + 1]) == (w0, w1), i
