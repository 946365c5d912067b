"""TEST INFRASTRUCTURE ONLY -- slow big-int ate pairing on BLS12-381 and the Groth16 verification equation.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.
PARITY UNPINNED at byte level (see oracle/py/bls12_381.py).  What this file adds is an INDEPENDENT acceptance
check in the reference's own sense: the reference's Groth16 tests assert exactly `Groth16::verify(...) == true`
(circuits-ark/src/encryption.rs:403-411, circom_wrapper.rs:206-224; product call sites lib/src/zk/encryption.rs:152,
sample_entries.rs:126).  ark-groth16 0.3.0 verifier.rs::verify_proof checks
    e(A, B) * e(prepared_inputs, -gamma) * e(C, -delta) == e(alpha, beta)
with prepared_inputs = gamma_abc_g1[0] + sum_i x_i * gamma_abc_g1[i + 1].  A key produced by the CUDA key
generator and a proof produced by the CUDA prover that pass this equation are correct Groth16 objects whatever
else might differ; the pairing itself is pinned by bilinearity / order tests (tests/test_oracle.py).

Construction (textbook, not ark-ec's optimised tower): Fq12 = Fq[w]/(w^12 - 2 w^6 + 2) (u = w^6 - 1, u^2 = -1,
xi = 1 + u = w^6), sextic untwist psi(x', y') = (x' w^-2, y' w^-3) from E'(Fq2): y^2 = x^3 + 4 xi to
E(Fq12): y^2 = x^3 + 4, Miller loop over |z| = 0xd201000000010000 with slopes computed on the twist, final
exponentiation by (p^12 - 1)/r as one big-integer power.  The sign of z is ignored (consistently), which
replaces e by its inverse: still bilinear and non-degenerate, and the verification equation is unaffected.
"""
from __future__ import annotations

from .bls12_381 import BLS_Z, P_MOD, R_MOD, Fq2Ops, G1, G2

F2 = Fq2Ops
ATE_LOOP = -BLS_Z
FINAL_EXP = (P_MOD**12 - 1) // R_MOD
assert (P_MOD**12 - 1) % R_MOD == 0

F12_ONE = (1,) + (0,) * 11


def f12_mul(a, b):
    t = [0] * 23
    for i, ai in enumerate(a):
        if ai:
            for j, bj in enumerate(b):
                t[i + j] += ai * bj
    for k in range(22, 11, -1):  # w^12 = 2 w^6 - 2
        c = t[k]
        if c:
            t[k - 6] += 2 * c
            t[k - 12] -= 2 * c
    return tuple(x % P_MOD for x in t[:12])


def f12_pow(a, e):
    acc, base = F12_ONE, a
    while e:
        if e & 1:
            acc = f12_mul(acc, base)
        base = f12_mul(base, base)
        e >>= 1
    return acc


def f12_inv(a):
    """Solve (mult-by-a) x = 1 by Gaussian elimination over Fq (12 x 12)."""
    cols = []
    wj = F12_ONE
    w = (0, 1) + (0,) * 10
    for _ in range(12):
        cols.append(f12_mul(a, wj))
        wj = f12_mul(wj, w)
    m = [[cols[j][i] for j in range(12)] + [1 if i == 0 else 0] for i in range(12)]
    for c in range(12):
        piv = next(r for r in range(c, 12) if m[r][c])
        m[c], m[piv] = m[piv], m[c]
        inv = pow(m[c][c], P_MOD - 2, P_MOD)
        m[c] = [x * inv % P_MOD for x in m[c]]
        for r in range(12):
            if r != c and m[r][c]:
                f = m[r][c]
                m[r] = [(x - f * y) % P_MOD for x, y in zip(m[r], m[c])]
    return tuple(m[i][12] for i in range(12))


def f12_from_fq2(c):
    """c0 + c1 u with u = w^6 - 1."""
    return ((c[0] - c[1]) % P_MOD, 0, 0, 0, 0, 0, c[1] % P_MOD, 0, 0, 0, 0, 0)


W = (0, 1) + (0,) * 10
W_INV = f12_inv(W)
W_INV3 = f12_mul(W_INV, f12_mul(W_INV, W_INV))
assert f12_mul(W, W_INV) == F12_ONE


def untwist(Q):
    """psi: E'(Fq2) -> E(Fq12)."""
    w2 = f12_mul(W_INV, W_INV)
    return f12_mul(f12_from_fq2(Q[0]), w2), f12_mul(f12_from_fq2(Q[1]), W_INV3)


def _line(T, lam, P):
    """Line through psi(T) with slope lam * w^-1, evaluated at P = (xP, yP) in E(Fq):
    yP - lam xP w^-1 + (lam x_T - y_T) w^-3."""
    xP, yP = P
    a = f12_mul(f12_from_fq2(F2.mul(lam, (xP, 0))), W_INV)
    b = f12_mul(f12_from_fq2(F2.sub(F2.mul(lam, T[0]), T[1])), W_INV3)
    out = [(-x + y) % P_MOD for x, y in zip(a, b)]
    out[0] = (out[0] + yP) % P_MOD
    return tuple(out)


def miller_loop(P, Q):
    """f_{|z|, Q}(P) without the final exponentiation.  P in G1 (affine ints), Q in G2 (affine Fq2 pairs); None = infinity."""
    if P is None or Q is None:
        return F12_ONE
    f, T = F12_ONE, Q
    for bit in bin(ATE_LOOP)[3:]:
        lam = F2.mul(F2.mul(F2.small(3), F2.sqr(T[0])), F2.inv(F2.mul(F2.small(2), T[1])))
        f = f12_mul(f12_mul(f, f), _line(T, lam, P))
        x3 = F2.sub(F2.sqr(lam), F2.mul(F2.small(2), T[0]))
        T = (x3, F2.sub(F2.mul(lam, F2.sub(T[0], x3)), T[1]))
        if bit == "1":
            lam = F2.mul(F2.sub(Q[1], T[1]), F2.inv(F2.sub(Q[0], T[0])))
            f = f12_mul(f, _line(T, lam, P))
            x3 = F2.sub(F2.sub(F2.sqr(lam), T[0]), Q[0])
            T = (x3, F2.sub(F2.mul(lam, F2.sub(T[0], x3)), T[1]))
    return f


def final_exponentiation(f):
    return f12_pow(f, FINAL_EXP)


def pairing(P, Q):
    return final_exponentiation(miller_loop(P, Q))


def g1_neg(P):
    return None if P is None else (P[0], (-P[1]) % P_MOD)


def prepare_inputs(gamma_abc_g1, public_inputs):
    """ark-groth16 0.3.0 verifier.rs::prepare_inputs; public_inputs excludes the constant 1."""
    if len(public_inputs) + 1 != len(gamma_abc_g1):
        raise ValueError("MalformedVerifyingKey")
    acc = G1.to_jac(gamma_abc_g1[0])
    for x, b in zip(public_inputs, gamma_abc_g1[1:]):
        acc = G1.jadd(acc, G1.jmul(G1.to_jac(b), x % R_MOD))
    return G1.to_affine(acc)


def groth16_verify(alpha_g1, beta_g2, gamma_g2, delta_g2, gamma_abc_g1, proof, public_inputs):
    """ark-groth16 0.3.0 verifier.rs::verify_proof as one product of Miller loops."""
    A, B, Cc = proof
    acc = prepare_inputs(gamma_abc_g1, public_inputs)
    f = miller_loop(A, B)
    f = f12_mul(f, miller_loop(g1_neg(acc), gamma_g2))
    f = f12_mul(f, miller_loop(g1_neg(Cc), delta_g2))
    f = f12_mul(f, miller_loop(g1_neg(alpha_g1), beta_g2))
    return final_exponentiation(f) == F12_ONE
