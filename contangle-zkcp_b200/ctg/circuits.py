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
JUBJUB_GEN = (8076246640662884909881801758704306714034609987455869804520522091855516602923,
              13262374693698910701929044844600465831413122818447359594527400194675274060458)
JUBJUB_ORDER = 6554484396890773809930967563523245729705921265872317281365359162392183254199


class LC:
    """A linear combination with its value: terms = [(coefficient, variable)], variable = ("i" | "w", index); terms is None
    when the constraint system only computes the assignment."""
    __slots__ = ("t", "v")

    def __init__(self, terms, value):
        self.t, self.v = terms, value % R_MOD


class ConstraintSystem:
    """Rank-1 constraints <A, z> * <B, z> = <C, z> over z = instance ++ witness.  record=False computes the assignment only
    (the matrices of a circuit do not depend on the witness, so further witnesses of the same circuit skip the bookkeeping)."""

    def __init__(self, record=True):
        self.inst, self.wit, self.rows, self.record = [1], [], ([], [], []), record

    def new_input(self, v):
        self.inst.append(v % R_MOD)
        return LC([(1, ("i", len(self.inst) - 1))] if self.record else None, v)

    def new_witness(self, v):
        self.wit.append(v % R_MOD)
        return LC([(1, ("w", len(self.wit) - 1))] if self.record else None, v)

    def value(self, lc):
        return lc.v

    def enforce(self, a, b, c):
        if self.record:
            for rows, lc in zip(self.rows, (a, b, c)):
                rows.append(lc.t)

    # ---- linear-combination helpers
    def add(self, a, b):
        return LC(a.t + b.t if self.record else None, a.v + b.v)

    def scale(self, a, k):
        return LC([(c * k % R_MOD, v) for c, v in a.t] if self.record else None, a.v * k)

    def sub(self, a, b):
        return self.add(a, self.scale(b, R_MOD - 1))

    def const(self, k):
        return LC([(k % R_MOD, ONE)] if self.record and k % R_MOD else ([] if self.record else None), k)

    def zero(self):
        return LC([] if self.record else None, 0)

    def mul(self, a, b):
        """A fresh witness w with a * b = w."""
        w = self.new_witness(a.v * b.v)
        self.enforce(a, b, w)
        return w

    def boolean(self, bit):
        b = self.new_witness(bit & 1)
        self.enforce(b, self.sub(self.const(1), b), self.zero())
        return b

    def select(self, bit, if_one, if_zero):
        """bit ? if_one : if_zero  =  if_zero + bit * (if_one - if_zero)."""
        return self.add(if_zero, self.mul(bit, self.sub(if_one, if_zero)))

    def enforce_equal(self, a, b):
        self.enforce(self.sub(a, b), self.const(1), self.zero())

    def compact(self, lc):
        if not self.record:
            return lc
        acc = {}
        for c, v in lc.t:
            acc[v] = (acc.get(v, 0) + c) % R_MOD
        return LC([(c, v) for v, c in acc.items() if c], lc.v)

    def linear(self, coeffs, lcs):
        """sum_k coeffs[k] * lcs[k] (one row of the MDS layer)."""
        out = self.zero()
        for k, x in zip(coeffs, lcs):
            out = self.add(out, self.scale(x, k))
        return self.compact(out)

    # ---- export
    def r1cs(self):
        """(r1cs dict for Context.upload_r1cs / generate_parameters).  Variables: instance first, then witness."""
        ni = len(self.inst)
        out = {"num_instance": ni, "num_witness": len(self.wit), "num_constraints": len(self.rows[0])}
        for name, rows in zip("abc", self.rows):
            rp, col, val = [0], [], []
            for terms in rows:
                acc = {}
                for c, v in terms:
                    acc[v] = (acc.get(v, 0) + c) % R_MOD
                for (kind, k), c in sorted(acc.items(), key=lambda t: (t[0][0] != "i", t[0][1])):
                    if c:
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
        state = [cs.add(s, cs.const(c)) for s, c in zip(state, pc.ark[r])]
        if r < half or r >= half + pc.partial_rounds:
            state = [_sbox(cs, s, pc.alpha) for s in state]
        else:
            state[0] = _sbox(cs, state[0], pc.alpha)
        state = [cs.linear(row, state) for row in pc.mds]
    return state


def poseidon_crh(cs, pc, inputs):
    """CRH::evaluate / TwoToOneCRH::evaluate / sponge.absorb(inputs).squeeze(1) for at most two inputs (one rate block)."""
    assert 1 <= len(inputs) <= 2
    state = list(inputs) + [cs.zero() for _ in range(3 - len(inputs))]
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
    x3 = cs.new_witness(num_x.v * pow(den_x.v, -1, R_MOD))
    y3 = cs.new_witness(num_y.v * pow(den_y.v, -1, R_MOD))
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


# ---- Proof of Encryption (EncryptCircuit, circuits-ark/src/encryption.rs:219-317) -----------------------------------------------------
def _te_add_ints(p, q):
    """The same affine group law on integers (constants of the fixed-base multiplication; witness-side arithmetic)."""
    (x1, y1), (x2, y2) = p, q
    t = JUBJUB_D * x1 % R_MOD * x2 % R_MOD * y1 % R_MOD * y2 % R_MOD
    return ((x1 * y2 + y1 * x2) * pow(1 + t, -1, R_MOD) % R_MOD, (y1 * y2 + x1 * x2) * pow(1 - t, -1, R_MOD) % R_MOD)


def te_add_const(cs, p, q_ints):
    """p + Q for a CONSTANT point Q: the four cross products are linear in p, so 3 constraints (t, x3, y3) instead of 7."""
    (x1, y1), (x2, y2) = p, q_ints
    a, b, c, dd = cs.scale(x1, y2), cs.scale(y1, x2), cs.scale(x1, x2), cs.scale(y1, y2)
    t = cs.mul(cs.scale(c, JUBJUB_D), dd)
    num_x, num_y = cs.add(a, b), cs.add(dd, c)
    den_x, den_y = cs.add(cs.const(1), t), cs.sub(cs.const(1), t)
    x3 = cs.new_witness(num_x.v * pow(den_x.v, -1, R_MOD))
    y3 = cs.new_witness(num_y.v * pow(den_y.v, -1, R_MOD))
    cs.enforce(x3, den_x, num_x)
    cs.enforce(y3, den_y, num_y)
    return x3, y3


def te_fixed_base_mul_le(cs, base_ints, bits):
    """scalar_mul_le on a constant point (the generator, `CV::new_constant`, :228): the multiples 2^i G are constants."""
    acc = (cs.const(0), cs.const(1))
    base = base_ints
    for i, bit in enumerate(bits):
        s = te_add_const(cs, acc, base)
        acc = (cs.select(bit, s[0], acc[0]), cs.select(bit, s[1], acc[1]))
        if i + 1 < len(bits):
            base = _te_add_ints(base, base)
    return acc


def encrypt_circuit(pc, pk, r, msg, c1, c2, n_slots=None, record=True, r_bits=256):
    """EncryptCircuit::generate_constraints (:289-317) + verify_encryption (:219-262): plaintext slots as witnesses, then the
    ciphertext as INPUT -- instance [1, c1.x, c1.y, c2[0 .. n)] -- then: the 256 bits of r, pk as a witness point, s = r pk,
    c1' = r G, dh = sponge(s.x, s.y), c1' == c1, and per slot (dh + m_i - c2_i) * [c2_i != 0] = 0 with the is_zero pair.
    pk, c1: (x, y) integers; r: integer; msg, c2: integers (slots beyond len(msg) are empty: zero plaintext, zero ciphertext)."""
    n = len(c2) if n_slots is None else n_slots
    assert len(msg) <= n and len(c2) <= n
    cs = ConstraintSystem(record)
    m_vars = [cs.new_witness(msg[i] if i < len(msg) else 0) for i in range(n)]              # :293-300
    c1v = (cs.new_input(c1[0]), cs.new_input(c1[1]))                                          # ciphertext_var, Input mode (:264-287)
    c2v = [cs.new_input(c2[i] if i < len(c2) else 0) for i in range(n)]
    bits = [cs.boolean((r >> i) & 1) for i in range(r_bits)]                                  # :231-235
    pkv = (cs.new_witness(pk[0]), cs.new_witness(pk[1]))                                      # :237
    xx, yy = cs.mul(pkv[0], pkv[0]), cs.mul(pkv[1], pkv[1])                                   # on-curve check of the witness point
    cs.enforce(cs.scale(xx, JUBJUB_D), yy, cs.sub(cs.sub(yy, xx), cs.const(1)))
    s = te_scalar_mul_le(cs, pkv, bits)                                                       # :240
    c1calc = te_fixed_base_mul_le(cs, JUBJUB_GEN, bits)                                       # :243
    dh = poseidon_crh(cs, pc, [s[0], s[1]])                                                   # :245-249
    cs.enforce_equal(c1calc[0], c1v[0])                                                       # :251
    cs.enforce_equal(c1calc[1], c1v[1])
    for mv, exp in zip(m_vars, c2v):                                                          # :253-262
        z = cs.new_witness(1 if exp.v == 0 else 0)                                            # exp.is_zero()
        inv = cs.new_witness(pow(exp.v, -1, R_MOD) if exp.v else 0)
        cs.enforce(exp, inv, cs.sub(cs.const(1), z))
        cs.enforce(exp, z, cs.zero())
        cs.enforce(cs.sub(cs.add(dh, mv), exp), cs.sub(cs.const(1), z), cs.zero())            # conditional_enforce_equal
    return cs


# ---- the workload above the C ABI: encrypt -> Merkle tree -> sampled-entry statements ---------------------------------------


def sample_entries_workload(ctx, poseidon_words, n_leaves, count, seed=1):
    """What the seller does before proving (lib/src/seller.rs:81-127): encrypt a dataset of n_leaves byte-valued elements with
    the device's bulk hashed ElGamal, build the Poseidon Merkle tree of the ciphertext on the device, and state `count`
    sampled-entry proofs about it (leaf indices spread over the tree).  poseidon_words = (full, partial, alpha, mds words,
    ark words), e.g. ctg.poseidon_reference_params().  Returns (r1cs, z (count, m, 8) Montgomery words, public inputs
    [(sample, root)] as integers)."""
    import random
    from .synthetic import fr_from_mont_words, ints_to_words
    rnd = random.Random(seed)
    pc = PoseidonConstants.from_words(*poseidon_words)
    pos = ctx.poseidon(*poseidon_words)
    sk, rr = rnd.randrange(1, JUBJUB_ORDER), rnd.randrange(1, JUBJUB_ORDER)
    msgs = [rnd.randrange(256) for _ in range(n_leaves)]
    d_w, gen_w = fr_mont([JUBJUB_D])[0], fr_mont(list(JUBJUB_GEN))
    pk_w = ctx.te_mul_batch(d_w, gen_w.reshape(1, 2, 8), ints_to_words([sk], 8))[0]          # keygen: pk = G * sk
    c1_w, c2_w = ctx.elgamal_encrypt_batch(pos, d_w, gen_w, pk_w, ints_to_words([rr], 8), fr_mont(msgs).reshape(1, n_leaves, 8))
    tree = pos.merkle_tree(c2_w[0])
    pos.free()
    c1 = tuple(fr_from_mont_words(c1_w[0]))
    c2 = fr_from_mont_words(c2_w[0])
    root = fr_from_mont_words(tree.root().reshape(1, 8))[0]
    r1cs, zs, publics = None, [], []
    for k in range(count):
        idx = (k * 2654435761 + seed) % n_leaves
        sib_w, path_w, _ = tree.generate_proof(idx)
        sib = fr_from_mont_words(sib_w.reshape(1, 8))[0]
        path = fr_from_mont_words(np.stack(path_w)) if path_w else []
        cs = sample_entries_circuit(pc, c1, c2[idx], sib, path, idx, sk, msgs[idx], root, record=r1cs is None)
        if r1cs is None:
            r1cs = cs.r1cs()
        zs.append(cs.assignment())
        publics.append((msgs[idx], root))
    m = len(zs[0])
    z = fr_mont([v for row in zs for v in row]).reshape(count, m, 8)
    return r1cs, z, publics


def encrypt_workload(ctx, poseidon_words, n_data, count, seed=1):
    """What the seller does per buyer request / per block (lib/src/zk/encryption.rs:65-85): `count` independent hashed-ElGamal
    encryptions of n_data byte-valued elements under one public key -- on the device (ctg_te_mul_batch + ctg_poseidon_hash_batch) --
    and the Proof-of-Encryption statement of each.  Returns (r1cs, z (count, m, 8) Montgomery words, public inputs per proof
    [c1.x, c1.y, c2...] as integers)."""
    import random
    from .synthetic import fr_from_mont_words, ints_to_words
    rnd = random.Random(seed)
    pc = PoseidonConstants.from_words(*poseidon_words)
    pos = ctx.poseidon(*poseidon_words)
    sk = rnd.randrange(1, JUBJUB_ORDER)
    d_w, gen_w = fr_mont([JUBJUB_D])[0], fr_mont(list(JUBJUB_GEN))
    pk_w = ctx.te_mul_batch(d_w, gen_w.reshape(1, 2, 8), ints_to_words([sk], 8))[0]
    pk = tuple(fr_from_mont_words(pk_w))
    rs = [rnd.randrange(1, JUBJUB_ORDER) for _ in range(count)]
    msgs = [[rnd.randrange(1, 256) for _ in range(n_data)] for _ in range(count)]
    c1_w, c2_w = ctx.elgamal_encrypt_batch(pos, d_w, gen_w, pk_w, ints_to_words(rs, 8), fr_mont([v for m in msgs for v in m]).reshape(count, n_data, 8))
    pos.free()
    c1s = fr_from_mont_words(c1_w.reshape(-1, 8))
    c2s = fr_from_mont_words(c2_w.reshape(-1, 8))
    r1cs, zs, publics = None, [], []
    for k in range(count):
        c1 = (c1s[2 * k], c1s[2 * k + 1])
        c2 = c2s[k * n_data:(k + 1) * n_data]
        cs = encrypt_circuit(pc, pk, rs[k], msgs[k], c1, c2, record=r1cs is None)
        if r1cs is None:
            r1cs = cs.r1cs()
        zs.append(cs.assignment())
        publics.append([c1[0], c1[1]] + list(c2))
    z = fr_mont([v for row in zs for v in row]).reshape(count, len(zs[0]), 8)
    return r1cs, z, publics

