"""Test helpers: trapdoor keys built with the C++ oracle's fixed-base multiplier, and the
discrete-log closed form for progression keys (ctg.synthetic.progression_pk)."""
import numpy as np

from oracle import ark
from oracle.py import bls12_381 as B
from oracle.py import groth16 as G


def trapdoor_pk_arrays(cs, td):
    """ark-groth16 generate_parameters (oracle/py/groth16.py) with the scalar multiplications done by
    oracle/c (fast enough for 2^12-2^14 domains).  cs: oracle.py R1CS; td: Trapdoor (ints, affine gens)."""
    n = cs.domain_size
    log_n = n.bit_length() - 1
    u = ark.fr_ints_from_mont(ark.lagrange_coeffs(log_n, ark.fr_mont_from_ints([td.tau])[0]))
    m, ni = cs.num_vars, cs.num_instance
    a, b, c = [0] * m, [0] * m, [0] * m
    for i in range(cs.num_constraints):
        for cf, j in cs.a[i]:
            a[j] = (a[j] + u[i] * cf) % B.R_MOD
        for cf, j in cs.b[i]:
            b[j] = (b[j] + u[i] * cf) % B.R_MOD
        for cf, j in cs.c[i]:
            c[j] = (c[j] + u[i] * cf) % B.R_MOD
    for j in range(ni):
        a[j] = (a[j] + u[cs.num_constraints + j]) % B.R_MOD
    zt = (pow(td.tau, n, B.R_MOD) - 1) % B.R_MOD
    dinv = B.fr_inv(td.delta)
    g1, g2 = ark.g1_from_py([td.g1])[0], ark.g2_from_py([td.g2])[0]
    fb1 = lambda ks: ark.g1_fixed_base(g1, ark.fr_mont_from_ints([k % B.R_MOD for k in ks]))
    fb2 = lambda ks: ark.g2_fixed_base(g2, ark.fr_mont_from_ints([k % B.R_MOD for k in ks]))
    abc = [(td.beta * a[i] + td.alpha * b[i] + c[i]) % B.R_MOD for i in range(m)]
    hs, t = [], zt * dinv % B.R_MOD
    for _ in range(n - 1):
        hs.append(t)
        t = t * td.tau % B.R_MOD
    singles = fb1([td.alpha, td.beta, td.delta])
    singles2 = fb2([td.beta, td.delta, td.gamma])
    ginv = B.fr_inv(td.gamma)
    return {"alpha_g1": singles[0], "beta_g1": singles[1], "delta_g1": singles[2], "beta_g2": singles2[0], "delta_g2": singles2[1],
            "gamma_g2": singles2[2], "gamma_abc_g1": fb1([abc[i] * ginv for i in range(ni)]),
            "a_query": fb1(a), "b_g1_query": fb1(b), "b_g2_query": fb2(b), "h_query": fb1(hs),
            "l_query": fb1([abc[i] * dinv for i in range(ni, m)]) if m > ni else np.zeros((0, 24), dtype=np.uint32)}


def dlog_expected_proof(dl, r1cs, z, h, r, s):
    """Expected (A, B, C) for a progression key (dl from ctg.synthetic.progression_pk): every base has a
    known discrete log, so each proof element is ONE scalar multiplication of the generator.
    z, h: ints (h natural order, n entries); r, s ints."""
    R = B.R_MOD
    ni = r1cs["num_instance"]
    m = len(z)
    lin = lambda key, vals: sum(v * (dl[key][0] + i * dl[key][1]) for i, v in enumerate(vals)) % R
    ea = (dl["alpha"] + r * dl["delta"] + lin("a", z)) % R
    eb = (dl["beta"] + s * dl["delta"] + lin("b", z)) % R
    ec = (s * ea + r * eb - r * s % R * dl["delta"] + lin("l", z[ni:]) + lin("h", h[:len(h) - 1])) % R
    return B.G1.mul(B.G1_GEN, ea), B.G2.mul(B.G2_GEN, eb), B.G1.mul(B.G1_GEN, ec)


def proof_to_py(proof):
    a, b, c = proof
    return ark.g1_to_py(a)[0], ark.g2_to_py(b)[0], ark.g1_to_py(c)[0]


def trapdoor_words(td):
    """oracle.py Trapdoor -> the word arrays ctg.Context.generate_parameters takes (ctg_trapdoor)."""
    d = {k: ark.fr_mont_from_ints([getattr(td, k)])[0] for k in ("alpha", "beta", "gamma", "delta", "tau")}
    d["g1"] = ark.g1_from_py([td.g1])[0]
    d["g2"] = ark.g2_from_py([td.g2])[0]
    return d


def groth16_verify_words(key, proof, public_inputs):
    """Pairing check (oracle/py/pairing.py) of a proof given as word arrays against a key dict of word arrays
    holding alpha_g1, beta_g2, gamma_g2, delta_g2, gamma_abc_g1.  public_inputs: ints, without the constant 1."""
    from oracle.py import pairing as PR
    a, b, c = proof_to_py(proof)
    return PR.groth16_verify(ark.g1_to_py(key["alpha_g1"])[0], ark.g2_to_py(key["beta_g2"])[0], ark.g2_to_py(key["gamma_g2"])[0],
                             ark.g2_to_py(key["delta_g2"])[0], ark.g1_to_py(key["gamma_abc_g1"]), (a, b, c), public_inputs)


def cs_from_arrays(r1cs):
    """Array-dict R1CS (ctg.synthetic / ark.r1cs_from_py layout) -> oracle.py.groth16.R1CS with integer coefficients."""
    cs = G.R1CS(num_instance=r1cs["num_instance"], num_witness=r1cs["num_witness"])
    for name in "abc":
        m = r1cs[name]
        vals = ark.fr_ints_from_mont(m["val"]) if len(m["col"]) else []
        rows = []
        for i in range(r1cs["num_constraints"]):
            lo, hi = int(m["row_ptr"][i]), int(m["row_ptr"][i + 1])
            rows.append([(vals[k], int(m["col"][k])) for k in range(lo, hi)])
        setattr(cs, name, rows)
    return cs


def closed_form_proof_arrays(r1cs, td, z, r, s, g1=B.G1_GEN, g2=B.G2_GEN):
    """The Groth16 proof as three scalar multiplications of the generators (SURVEY 8(c) pin 4) for an array-dict R1CS of any
    size: O(nnz) big-int work, Lagrange coefficients from oracle/c.  td: dict of ints alpha, beta, delta, tau (gamma unused).
    h(tau) t(tau) comes from the QAP identity A(tau) B(tau) - C(tau), not from any NTT."""
    R = B.R_MOD
    ni, nc = r1cs["num_instance"], r1cs["num_constraints"]
    n = 1
    while n < ni + nc:
        n *= 2
    L = ark.fr_ints_from_mont(ark.lagrange_coeffs(n.bit_length() - 1, ark.fr_mont_from_ints([td["tau"]])[0]))
    tot, inst = [], []
    for name in "abc":
        m = r1cs[name]
        vals = ark.fr_ints_from_mont(m["val"]) if len(m["col"]) else []
        cols, rp = m["col"].tolist(), m["row_ptr"].tolist()
        total, icol = 0, [0] * ni
        for i in range(nc):
            Li = L[i]
            for k in range(rp[i], rp[i + 1]):
                c = cols[k]
                t = vals[k] * Li
                total += t * z[c]
                if c < ni:
                    icol[c] += t
        if name == "a":
            for j in range(ni):
                total += L[nc + j] * z[j]
                icol[j] += L[nc + j]
        tot.append(total % R)
        inst.append([x % R for x in icol])
    az, bz, cz = tot
    inst_part = sum((td["beta"] * inst[0][j] + td["alpha"] * inst[1][j] + inst[2][j]) * z[j] for j in range(ni)) % R
    laux = (td["beta"] * az + td["alpha"] * bz + cz - inst_part) % R
    ht = (az * bz - cz) % R          # h(tau) * t(tau)
    ea = (td["alpha"] + az + r * td["delta"]) % R
    eb = (td["beta"] + bz + s * td["delta"]) % R
    ec = ((laux + ht) * B.fr_inv(td["delta"]) + s * ea + r * eb - r * s % R * td["delta"]) % R
    return B.G1.mul(g1, ea), B.G2.mul(g2, eb), B.G1.mul(g1, ec)


from ctg.synthetic import weighted_sums  # noqa: E402,F401  (lives beside the generators: bench.py uses it at N > 1)
