"""R1CS of the reference's Proof-of-Property circuit (SampleEntries, circuits-ark/src/sample_entries.rs:136-232), built gate by
gate in Python so that the prover can be driven with the REAL statement instead of a shape:

    instance   [1, sample_entry, merkle_root]                      (new_input order of generate_constraints, :221-232)
    witness    ciphertext (c1 point, c2 = the sampled leaf), Merkle path, secret key bits
    relations  verify_membership (:136-166): CRH(leaf) folded with the sibling hashes up to merkle_root
               compare_decrypted (:168-195): s = sk * c1 (scalar_mul_le over the 256 bits of sk), dh = sponge(s.x, s.y),
               c2 - dh == sample_entry

Synthesis is the host's job (it stays in Rust in the reference: r1cs-std gadgets); this module is the workload generator the
tests and bench.py use above the C ABI.  The gate-level decomposition is ours (arkworks' gadget library is not on this box), the
statement and the instance layout are the reference's: ~4.9 k + 277 log2(n_leaves) constraints, which matches SURVEY
Appendix B's estimate (5 k + 300 log2 n).  Poseidon: width 3 (rate 2, capacity 1), x^alpha S-box, the sponge layout of
ark-sponge 0.3 (rate cells first, squeeze state[0]) -- the same convention as csrc/poseidon.cu.
"""
from __future__ import annotations

import numpy as np

from .synthetic import R_MOD, fr_mont

ONE = ("i", 0)
JUBJUB_A = R_MOD - 1
JUBJUB_D = (-10240 * pow(10241, -1, R_MOD)) % R_MOD


class ConstraintSystem:
    """Rank-1 constraints <A, z> * <B, z> = <C, z> over z = instance ++ witness; a linear combination is a list of
    (coefficient, variable), a variable is ("i" | "w", index).  record=False computes the assignment only (the matrices of a
    circuit do not depend on the witness, so further witnesses of the same circuit skip the bookkeeping)."""

    def __init__(self, record=True):
        self.inst, self.wit, self.rows, self.record = [1], [], ([], [], []), record

    def new_input(self, v):
        self.inst.append(v % R_MOD)
        return [(1, ("i", len(self.inst) - 1))]

    def new_witness(self, v):
        self.wit.append(v % R_MOD)
        return [(1, ("w", len(self.wit) - 1))]

    def value(self, lc):
        return sum(c * (self.inst[k] if kind == "i" else self.wit[k]) for c, (kind, k) in lc) % R_MOD

    def enforce(self, a, b, c):
        if self.record:
            for rows, lc in zip(self.rows, (a, b, c)):
                rows.append(lc)

    # ---- linear-combination helpers
    @staticmethod
    def add(a, b):
        return a + b

    @staticmethod
    def scale(a, k):
        return [(c * k % R_MOD, v) for c, v in a]

    @classmethod
    def sub(cls, a, b):
        return a + cls.scale(b, R_MOD - 1)

    @staticmethod
    def const(k):
        return [(k % R_MOD, ONE)]

    def mul(self, a, b):
        """A fresh witness w with a * b = w."""
        w = self.new_witness(self.value(a) * self.value(b))
        self.enforce(a, b, w)
        return w

    def boolean(self, bit):
        b = self.new_witness(bit & 1)
        self.enforce(b, self.sub(self.const(1), b), [])
        return b

    def select(self, bit, if_one, if_zero):
        """bit ? if_one : if_zero  =  if_zero + bit * (if_one - if_zero)."""
        return self.add(if_zero, self.mul(bit, self.sub(if_one, if_zero)))

    def enforce_equal(self, a, b):
        self.enforce(self.sub(a, b), self.const(1), [])

    @staticmethod
    def compact(lc):
        acc = {}
        for c, v in lc:
            acc[v] = (acc.get(v, 0) + c) % R_MOD
        return [(c, v) for v, c in acc.items() if c]

    # ---- export
    def r1cs(self):
        """(r1cs dict for Context.upload_r1cs / generate_parameters).  Variables: instance first, then witness."""
        ni = len(self.inst)
        out = {"num_instance": ni, "num_witness": len(self.wit), "num_constraints": len(self.rows[0])}
        for name, rows in zip("abc", self.rows):
            rp, col, val = [0], [], []
            for lc in rows:
                for c, (kind, k) in sorted(self.compact(lc), key=lambda t: (t[1][0] != "i", t[1][1])):
                    col.append(k if kind == "i" else ni + k)
                    val.append(c)
                rp.append(len(col))
            out[name] = {"row_ptr": np.array(rp, dtype=np.uint64), "col": np.array(col, dtype=np.uint32),
                         "val": fr_mont(val).reshape(-1, 8) if val else np.zeros((0, 8), dtype=np.uint32)}
        return out

    def assignment(self):
        return self.inst + self.wit


# ---- Poseidon ---------------------------------------------------------------------------------------------------------------
class PoseidonConstants:
    """full_rounds, partial_rounds, alpha and the mds / ark tables as canonical integers."""

    def __init__(self, full_rounds, partial_rounds, alpha, mds, ark):
        self.full_rounds, self.partial_rounds, self.alpha, self.mds, self.ark = full_rounds, partial_rounds, alpha, mds, ark

    @classmethod
    def from_words(cls, full_rounds, partial_rounds, alpha, mds_words, ark_words):
        """From the Montgomery-word arrays of ctg.poseidon_reference_params() / Context.poseidon()."""
        from .synthetic import fr_from_mont_words
        mds = fr_from_mont_words(np.asarray(mds_words).reshape(-1, 8))
        ark = fr_from_mont_words(np.asarray(ark_words).reshape(-1, 8))
        return cls(full_rounds, partial_rounds, alpha, [mds[3 * i:3 * i + 3] for i in range(3)], [ark[3 * i:3 * i + 3] for i in range(len(ark) // 3)])


def _sbox(cs, x, alpha):
    """x^alpha by square-and-multiply: 5 constraints for alpha = 17."""
    acc, base, e = None, x, alpha
    while e:
        if e & 1:
            acc = base if acc is None else cs.mul(acc, base)
        e >>= 1
        if e:
            base = cs.mul(base, base)
    return acc


def poseidon_permute(cs, pc, state):
    """PoseidonSponge::permute on three linear combinations: add round constants, S-box (all cells in the full rounds, cell 0
    in the partial ones), MDS.  The linear layers cost no constraints; only the S-boxes do (55 * 5 = 275 for 8 + 31 rounds)."""
    half = pc.full_rounds // 2
    for r in range(pc.full_rounds + pc.partial_rounds):
        state = [cs.compact(cs.add(s, cs.const(c))) for s, c in zip(state, pc.ark[r])]
        if r < half or r >= half + pc.partial_rounds:
            state = [_sbox(cs, s, pc.alpha) for s in state]
        else:
            state[0] = _sbox(cs, state[0], pc.alpha)
        state = [cs.compact([t for m, s in zip(row, state) for t in cs.scale(s, m)]) for row in pc.mds]
    return state


def poseidon_crh(cs, pc, inputs):
    """CRH::evaluate / TwoToOneCRH::evaluate / sponge.absorb(inputs).squeeze(1) for at most two inputs (one rate block)."""
    assert 1 <= len(inputs) <= 2
    state = [list(x) for x in inputs] + [[] for _ in range(3 - len(inputs))]
    return poseidon_permute(cs, pc, state)[0]


# ---- JubJub -----------------------------------------------------------------------------------------------------------------
def te_add(cs, p, q):
    """Complete twisted-Edwards addition (a = -1) on affine coordinates: 7 constraints.
    x3 (1 + d x1 x2 y1 y2) = x1 y2 + y1 x2,   y3 (1 - d x1 x2 y1 y2) = y1 y2 + x1 x2."""
    (x1, y1), (x2, y2) = p, q
    a, b, c, dd = cs.mul(x1, y2), cs.mul(y1, x2), cs.mul(x1, x2), cs.mul(y1, y2)
    t = cs.mul(cs.scale(c, JUBJUB_D), dd)
    num_x, num_y = cs.add(a, b), cs.add(dd, c)
    den_x, den_y = cs.add(cs.const(1), t), cs.sub(cs.const(1), t)
    x3 = cs.new_witness(cs.value(num_x) * pow(cs.value(den_x), -1, R_MOD))
    y3 = cs.new_witness(cs.value(num_y) * pow(cs.value(den_y), -1, R_MOD))
    cs.enforce(x3, den_x, num_x)
    cs.enforce(y3, den_y, num_y)
    return x3, y3


def te_scalar_mul_le(cs, point, bits):
    """scalar_mul_le: acc = O; for each bit, least significant first: acc = bit ? acc + base : acc; base = 2 base."""
    acc = (cs.const(0), cs.const(1))
    base = point
    for i, bit in enumerate(bits):
        s = te_add(cs, acc, base)
        acc = (cs.select(bit, s[0], acc[0]), cs.select(bit, s[1], acc[1]))
        if i + 1 < len(bits):
            base = te_add(cs, base, base)
    return acc


# ---- the circuit --------------------------------------------------------------------------------------------------------------
def sample_entries_circuit(pc, c1, leaf, leaf_sibling_hash, auth_path, leaf_index, sk, sample_entry, merkle_root, record=True, sk_bits=256):
    """SampleEntries::generate_constraints.  c1: (x, y) of the ciphertext's group element; leaf = c2[leaf_index];
    (leaf_sibling_hash, auth_path root side first, leaf_index) = MerkleTree::generate_proof(leaf_index); all canonical integers.
    Returns the ConstraintSystem (r1cs() / assignment())."""
    cs = ConstraintSystem(record)
    sample = cs.new_input(sample_entry)                                   # :226
    c1v = (cs.new_witness(c1[0]), cs.new_witness(c1[1]))                  # ciphertext_var, Witness mode (:197-212)
    leafv = cs.new_witness(leaf)
    # on-curve check of the witness point: -x^2 + y^2 = 1 + d x^2 y^2
    xx, yy = cs.mul(c1v[0], c1v[0]), cs.mul(c1v[1], c1v[1])
    cs.enforce(cs.scale(xx, JUBJUB_D), yy, cs.sub(cs.sub(yy, xx), cs.const(1)))
    # verify_membership (:136-166)
    sib = cs.new_witness(leaf_sibling_hash)
    path = [cs.new_witness(h) for h in auth_path]
    root = cs.new_input(merkle_root)
    cur = poseidon_crh(cs, pc, [leafv])
    idx = leaf_index
    for sibling in [sib] + path[::-1]:                                   # leaf level first, root-side entries last
        b = cs.boolean(idx & 1)                                           # 1: the node is a right child
        left = cs.select(b, sibling, cur)
        right = cs.sub(cs.add(cur, sibling), left)
        cur = poseidon_crh(cs, pc, [left, right])
        idx >>= 1
    cs.enforce_equal(cur, root)
    # compare_decrypted (:168-195)
    bits = [cs.boolean((sk >> i) & 1) for i in range(sk_bits)]
    s = te_scalar_mul_le(cs, c1v, bits)
    dh = poseidon_crh(cs, pc, [s[0], s[1]])
    cs.enforce_equal(cs.sub(leafv, dh), sample)
    return cs
