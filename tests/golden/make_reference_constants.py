#!/usr/bin/env python
"""Extracts the only bytes the reference holds on the witness-side crypto path into fixtures:

  /root/reference/circuits-ark/src/parameters.rs:12-229   P1 = bls12377_rate2_constraints: ark (39 x 3), mds (3 x 3) as
        decimal strings, rate 2, alpha 17, full_rounds 8, partial_rounds 31 -- parsed by get_poseidon_params
        (circuits-ark/src/poseidon.rs:13-41) with C::BaseField::from_str, C = JubJub, i.e. into BLS12-381 Fr.  The 126
        decimals are 377-bit numbers; ark-ff 0.3 FromStr accumulates res = res * 10 + digit IN THE FIELD, which is
        reduction mod r.
  /root/reference/circuits-ark/src/utils.rs:114-115       ALICE_SK / ALICE_PK hex strings (tests :117-127 decode them with
        ark_from_bytes as Fr and as a JubJub point).

Writes contangle-zkcp_b200/ctg/data/poseidon_bls12_381_rate2.json (shipped: ctg.poseidon_reference_params() loads it; the
decimal strings are kept exactly as the reference holds them and reduced at load time the way from_str does) and
tests/golden/reference_constants.json (the key hex strings + the reduced table as hex, so that loader and oracle are checked
against frozen values).  Needs /root/reference (this container only); tests/test_reference_constants.py re-runs the
extraction when the reference is present and compares.

    python tests/golden/make_reference_constants.py
"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference/circuits-ark/src"
R_MOD = 0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001


def from_str(dec: str) -> int:
    """ark-ff 0.3 FromStr for Fp: digit by digit in the field (== int(dec) mod r; written out as the crate does it)."""
    if not dec or not dec.isdigit() or (dec[0] == "0" and len(dec) > 1):
        raise ValueError("ark-ff from_str rejects this string")
    acc = 0
    for ch in dec:
        acc = (acc * 10 + int(ch)) % R_MOD
    return acc


def extract():
    src = open(os.path.join(REF, "parameters.rs")).read()
    body = src[src.index("pub static ref P1"):]

    def table(name):
        start = body.index('"%s" => array![' % name)
        depth, i = 0, start + len('"%s" => ' % name)
        j = i
        while True:
            if body.startswith("array![", j):
                depth += 1
                j += 7
                continue
            if body[j] == "]":
                depth -= 1
                if depth == 0:
                    break
            j += 1
        rows = re.findall(r"array!\[\s*((?:\"\d+\"\s*,?\s*)+)\]", body[i:j + 1])
        return [re.findall(r"\"(\d+)\"", r) for r in rows]

    ark, mds = table("ark"), table("mds")
    scal = {k: int(re.search(r'"%s"\s*=>\s*(\d+)' % k, body).group(1)) for k in ("rate", "alpha", "full_rounds", "partial_rounds")}
    assert len(mds) == 3 and all(len(r) == 3 for r in mds)
    assert len(ark) == scal["full_rounds"] + scal["partial_rounds"] and all(len(r) == 3 for r in ark)
    utils = open(os.path.join(REF, "utils.rs")).read()
    keys = {k: re.search(r'const %s: &str = "([0-9a-f]+)";' % k, utils).group(1) for k in ("ALICE_SK", "ALICE_PK")}
    return ark, mds, scal, keys


def main():
    ark, mds, scal, keys = extract()
    shipped = {"source": "nulltea/contangle-zkcp circuits-ark/src/parameters.rs:12-229 (P1, bls12377_rate2_constraints), parsed into BLS12-381 Fr by from_str (circuits-ark/src/poseidon.rs:18-33)",
               **scal, "mds": mds, "ark": ark}
    p1 = os.path.join(ROOT, "contangle-zkcp_b200", "ctg", "data", "poseidon_bls12_381_rate2.json")
    with open(p1, "w") as f:
        json.dump(shipped, f, indent=0)
        f.write("\n")
    golden = {"source": "circuits-ark/src/parameters.rs:12-229 reduced mod r as Fr::from_str does; circuits-ark/src/utils.rs:114-115",
              **scal, "mds_hex": [[format(from_str(v), "x") for v in row] for row in mds],
              "ark_hex": [[format(from_str(v), "x") for v in row] for row in ark], **keys}
    p2 = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_constants.json")
    with open(p2, "w") as f:
        json.dump(golden, f, indent=0, sort_keys=True)
        f.write("\n")
    print("wrote", p1, os.path.getsize(p1), "and", p2, os.path.getsize(p2))


if __name__ == "__main__":
    main()
