"""Loader for tests/golden/groth16_bls12_381.json (made by tests/golden/make_golden.py from oracle/py)."""
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
I = lambda s: int(s, 16)


def load():
    with open(os.path.join(HERE, "golden", "groth16_bls12_381.json")) as f:
        return json.load(f)


def g1(p):
    return None if p is None else (I(p[0]), I(p[1]))


def g2(p):
    return None if p is None else ((I(p[0][0]), I(p[0][1])), (I(p[1][0]), I(p[1][1])))


def r1cs_and_pk(case):
    """-> (oracle.py R1CS, array-dict r1cs, array-dict pk, z ints, r, s, h ints, expected proof as py tuples)."""
    from oracle import ark
    from oracle.py import groth16 as G
    cs = G.R1CS(num_instance=case["num_instance"], num_witness=case["num_witness"])
    for name in "abc":
        setattr(cs, name, [[(I(cf), j) for cf, j in row] for row in case[name + "_rows"]])
    pkj = case["pk"]
    pk = {}
    for k in ("alpha_g1", "beta_g1", "delta_g1"):
        pk[k] = ark.g1_from_py([g1(pkj[k])])[0]
    for k in ("beta_g2", "delta_g2"):
        pk[k] = ark.g2_from_py([g2(pkj[k])])[0]
    for k in ("a_query", "b_g1_query", "h_query", "l_query"):
        pk[k] = ark.g1_from_py([g1(p) for p in pkj[k]])
    pk["b_g2_query"] = ark.g2_from_py([g2(p) for p in pkj["b_g2_query"]])
    want = (g1(case["proof"]["a"]), g2(case["proof"]["b"]), g1(case["proof"]["c"]))
    return cs, ark.r1cs_from_py(cs), pk, [I(v) for v in case["z"]], I(case["r"]), I(case["s"]), [I(v) for v in case["h"]], want
