"""TEST INFRASTRUCTURE ONLY -- big-int restatement of ark-groth16 0.3.0 (generator.rs,
r1cs_to_qap.rs, prover.rs) for BLS12-381.  PARITY UNPINNED at byte level (see
bls12_381.py header): the reference's own Groth16 tests only assert accept
(circuits-ark/src/encryption.rs:403-411, circom_wrapper.rs:206-224).

Reference call sites that execute this path: lib/src/zk/encryption.rs:76,
verifiable_encryption.rs:92, sample_entries.rs:86, property.rs:133 (Groth16::prove),
lib/src/zk/encryption.rs:169 (Groth16::setup).

The pin this module provides on top of the restatement is `closed_form_proof`: with a
known trapdoor the proof is three scalar multiplications of the generators, independent
of any MSM / NTT code.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Tuple

from .bls12_381 import (G1, G2, R_MOD, FR_GENERATOR, fr_inv, fr_root_of_unity, ntt)

Row = List[Tuple[int, int]]  # (coeff in [0, r), column); columns: instance first, then witness


@dataclass
class R1CS:
    """ark-relations ConstraintSystem::to_matrices() shape."""
    num_instance: int  # includes the constant 1 at column 0
    num_witness: int
    a: List[Row] = field(default_factory=list)
    b: List[Row] = field(default_factory=list)
    c: List[Row] = field(default_factory=list)

    @property
    def num_constraints(self):
        return len(self.a)

    @property
    def num_vars(self):
        return self.num_instance + self.num_witness

    @property
    def domain_size(self):
        k = self.num_constraints + self.num_instance
        n = 1
        while n < k:
            n *= 2
        return n

    def is_satisfied(self, z):
        dot = lambda row: sum(cf * z[i] for cf, i in row) % R_MOD
        return all(dot(ra) * dot(rb) % R_MOD == dot(rc) for ra, rb, rc in zip(self.a, self.b, self.c))


def mul_chain(log_n: int, x0: int, x1: int):
    """MulChain(k) of SURVEY 8(d): instance {1, x_out}; x_i * x_{i+1} = x_{i+2};
    num_constraints + num_instance = 2^k exactly."""
    nc = (1 << log_n) - 2
    xs = [x0 % R_MOD, x1 % R_MOD]
    for i in range(nc):
        xs.append(xs[i] * xs[i + 1] % R_MOD)
    # xs has nc + 2 values; the last one is the public output
    cs = R1CS(num_instance=2, num_witness=nc + 1)
    col = lambda i: 1 if i == nc + 1 else 2 + i
    for i in range(nc):
        cs.a.append([(1, col(i))])
        cs.b.append([(1, col(i + 1))])
        cs.c.append([(1, col(i + 2))])
    z = [1, xs[nc + 1]] + xs[: nc + 1]
    return cs, z


def lagrange_coeffs(n: int, tau: int):
    """ark-poly EvaluationDomain::evaluate_all_lagrange_coefficients (tau outside domain)."""
    log_n = n.bit_length() - 1
    w = fr_root_of_unity(log_n)
    zt = (pow(tau, n, R_MOD) - 1) % R_MOD
    assert zt != 0
    out, wi = [], 1
    ninv = fr_inv(n)
    for _ in range(n):
        out.append(zt * ninv % R_MOD * wi % R_MOD * fr_inv((tau - wi) % R_MOD) % R_MOD)
        wi = wi * w % R_MOD
    return out


@dataclass
class Trapdoor:
    alpha: int
    beta: int
    gamma: int
    delta: int
    tau: int
    g1: tuple  # affine generator of G1 actually used (arkworks draws it at random)
    g2: tuple


@dataclass
class ProvingKey:
    """ark-groth16 0.3.0 ProvingKey / VerifyingKey field-for-field."""
    alpha_g1: tuple
    beta_g1: tuple
    beta_g2: tuple
    gamma_g2: tuple
    delta_g1: tuple
    delta_g2: tuple
    gamma_abc_g1: list
    a_query: list
    b_g1_query: list
    b_g2_query: list
    h_query: list
    l_query: list


def qap_at_tau(cs: R1CS, tau: int):
    n = cs.domain_size
    u = lagrange_coeffs(n, tau)
    m = cs.num_vars
    a, b, c = [0] * m, [0] * m, [0] * m
    for i in range(cs.num_constraints):
        for cf, j in cs.a[i]:
            a[j] = (a[j] + u[i] * cf) % R_MOD
        for cf, j in cs.b[i]:
            b[j] = (b[j] + u[i] * cf) % R_MOD
        for cf, j in cs.c[i]:
            c[j] = (c[j] + u[i] * cf) % R_MOD
    for j in range(cs.num_instance):  # input-consistency rows
        a[j] = (a[j] + u[cs.num_constraints + j]) % R_MOD
    zt = (pow(tau, n, R_MOD) - 1) % R_MOD
    return a, b, c, zt, n


def generate_parameters(cs: R1CS, td: Trapdoor) -> ProvingKey:
    """ark-groth16 0.3.0 generator.rs::generate_parameters."""
    a, b, c, zt, n = qap_at_tau(cs, td.tau)
    ni = cs.num_instance
    dinv, ginv = fr_inv(td.delta), fr_inv(td.gamma)
    j1, j2 = G1.to_jac(td.g1), G2.to_jac(td.g2)
    m1 = lambda k: G1.to_affine(G1.jmul(j1, k % R_MOD))
    m2 = lambda k: G2.to_affine(G2.jmul(j2, k % R_MOD))
    abc = [(td.beta * a[i] + td.alpha * b[i] + c[i]) % R_MOD for i in range(cs.num_vars)]
    h, t = [], zt * dinv % R_MOD
    for _ in range(n - 1):
        h.append(m1(t))
        t = t * td.tau % R_MOD
    return ProvingKey(
        alpha_g1=m1(td.alpha), beta_g1=m1(td.beta), beta_g2=m2(td.beta), gamma_g2=m2(td.gamma),
        delta_g1=m1(td.delta), delta_g2=m2(td.delta),
        gamma_abc_g1=[m1(abc[i] * ginv) for i in range(ni)],
        a_query=[m1(x) for x in a], b_g1_query=[m1(x) for x in b], b_g2_query=[m2(x) for x in b],
        h_query=h, l_query=[m1(abc[i] * dinv) for i in range(ni, cs.num_vars)],
    )


def witness_map(cs: R1CS, z):
    """ark-groth16 0.3.0 r1cs_to_qap.rs::witness_map: returns n coefficients."""
    n, nc, ni = cs.domain_size, cs.num_constraints, cs.num_instance
    dot = lambda row: sum(cf * z[i] for cf, i in row) % R_MOD
    a = [dot(r) for r in cs.a] + [0] * (n - nc)
    b = [dot(r) for r in cs.b] + [0] * (n - nc)
    c = [dot(r) for r in cs.c] + [0] * (n - nc)
    for j in range(ni):
        a[nc + j] = z[j]
    a = ntt(ntt(a, inverse=True), coset=True)
    b = ntt(ntt(b, inverse=True), coset=True)
    c = ntt(ntt(c, inverse=True), coset=True)
    zinv = fr_inv((pow(FR_GENERATOR, n, R_MOD) - 1) % R_MOD)
    ab = [((x * y - w) % R_MOD) * zinv % R_MOD for x, y, w in zip(a, b, c)]
    return ntt(ab, inverse=True, coset=True)


def create_proof(cs: R1CS, pk: ProvingKey, z, r: int, s: int, msm1=None, msm2=None):
    """ark-groth16 0.3.0 prover.rs::create_proof.  Returns affine (A, B, C)."""
    msm1 = msm1 or G1.msm
    msm2 = msm2 or G2.msm
    ni = cs.num_instance
    h = witness_map(cs, z)
    h_acc = msm1(pk.h_query, h)          # min(len) = n - 1 pairs
    aux = z[ni:]
    l_acc = msm1(pk.l_query, aux)
    z1 = z[1:]
    J = G1.to_jac
    g_a = G1.jmul(J(pk.delta_g1), r)
    for P in (pk.a_query[0], msm1(pk.a_query[1:], z1), pk.alpha_g1):
        g_a = G1.jadd(g_a, J(P))
    if r == 0:
        g1_b = J(None)
    else:
        g1_b = G1.jmul(J(pk.delta_g1), s)
        for P in (pk.b_g1_query[0], msm1(pk.b_g1_query[1:], z1), pk.beta_g1):
            g1_b = G1.jadd(g1_b, J(P))
    g2_b = G2.jmul(G2.to_jac(pk.delta_g2), s)
    for P in (pk.b_g2_query[0], msm2(pk.b_g2_query[1:], z1), pk.beta_g2):
        g2_b = G2.jadd(g2_b, G2.to_jac(P))
    g_c = G1.jadd(G1.jmul(g_a, s), G1.jmul(g1_b, r))
    g_c = G1.jadd(g_c, G1.jneg(G1.jmul(J(pk.delta_g1), r * s % R_MOD)))
    g_c = G1.jadd(g_c, J(l_acc))
    g_c = G1.jadd(g_c, J(h_acc))
    return G1.to_affine(g_a), G2.to_affine(g2_b), G1.to_affine(g_c)


def closed_form_proof(cs: R1CS, td: Trapdoor, z, r: int, s: int):
    """The proof as three scalar multiplications (SURVEY 8(c) pin 4).  h(tau) is
    obtained by evaluating the QAP identity at tau, not through any NTT:
    h(tau) = (A(tau) B(tau) - C(tau)) / t(tau)."""
    a, b, c, zt, n = qap_at_tau(cs, td.tau)
    ni = cs.num_instance
    az = sum(x * y for x, y in zip(a, z)) % R_MOD
    bz = sum(x * y for x, y in zip(b, z)) % R_MOD
    cz = sum(x * y for x, y in zip(c, z)) % R_MOD
    h_tau = (az * bz - cz) * fr_inv(zt) % R_MOD
    ea = (td.alpha + az + r * td.delta) % R_MOD
    eb = (td.beta + bz + s * td.delta) % R_MOD
    laux = sum((td.beta * a[i] + td.alpha * b[i] + c[i]) * z[i] for i in range(ni, cs.num_vars)) % R_MOD
    ec = ((laux + h_tau * zt) * fr_inv(td.delta) + s * ea + r * eb - r * s % R_MOD * td.delta) % R_MOD
    return G1.mul(td.g1, ea), G2.mul(td.g2, eb), G1.mul(td.g1, ec)
