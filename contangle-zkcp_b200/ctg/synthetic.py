"""Seeded synthetic inputs of the shapes SURVEY 8(d) names: MulChain(k) R1CS instances,
PoE-shaped R1CS instances, and proving keys whose queries are arithmetic progressions
P_i = (a + i*b) * G generated ON THE DEVICE (ctg_bases_progression_*).  Such keys are not
trapdoor-consistent, which the prover does not care about (a proof is a deterministic
function of key, assignment, r, s) -- and every base has a known discrete log, so tests can
pin whole proofs at 2^20+ with O(n) field arithmetic (tests/util.py: dlog_expected_proof).

Host-side helper for tests/bench only (plain Python integers; nothing here is on the hot path).
"""
from __future__ import annotations

import numpy as np

R_MOD = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001
P_MOD = 0x1A0111EA397FE69A4B1BA7B6434BACD764774B84F38512BF6730D2A0F6B0F6241EABFFFEB153FFFFB9FEFFFFFFFFAAAB
FR_R, FQ_R = 1 << 256, 1 << 384
G1_GEN = (0x17F1D3A73197D7942695638C4FA9AC0FC3688C4F9774B905A14E3A3F171BAC586C55E83FF97A1AEFFB3AF00ADB22C6BB,
          0x08B3F481E3AAA0F1A09E30ED741D8AE4FCF5E095D5D00AF600DB18CB2C04B3EDD03CC744A2888AE40CAA232946C5E7E1)
G2_GEN = ((0x024AA2B2F08F0A91260805272DC51051C6E47AD4FA403B02B4510B647AE3D1770BAC0326A805BBEFD48056C8C121BDB8,
           0x13E02B6052719F607DACD3A088274F65596BD0D09920B61AB5DA61BBDC7F5049334CF11213945D57E5AC7D055D042B7E),
          (0x0CE5D527727D6E118CC9CDC6DA2E351AADFD9BAA8CBDD3A76D429A695160D12C923AC9CC3BACA289E193548608B82801,
           0x0606C4A02EA734CC32ACD2B02BC28B99CB3E287E85A763AF267492AB572E99AB3F370D275CEC1DA1AAA9075FF05F79BE))


def ints_to_words(vals, words):
    buf = b"".join(int(v).to_bytes(4 * words, "little") for v in vals)
    return np.frombuffer(buf, dtype=np.uint32).reshape(len(vals), words).copy()


def words_to_ints(arr):
    arr = np.ascontiguousarray(arr, dtype=np.uint32)
    w = arr.shape[-1]
    flat = arr.reshape(-1, w)
    raw = flat.tobytes()
    return [int.from_bytes(raw[i * 4 * w:(i + 1) * 4 * w], "little") for i in range(flat.shape[0])]


def fr_mont(vals):
    """canonical ints -> (n, 8) Montgomery images (x * 2^256 mod r)"""
    return ints_to_words([v % R_MOD * FR_R % R_MOD for v in vals], 8)


def fr_from_mont_words(arr):
    rinv = pow(FR_R, -1, R_MOD)
    return [v * rinv % R_MOD for v in words_to_ints(arr)]


def g1_gen_words():
    return ints_to_words([c * FQ_R % P_MOD for c in G1_GEN], 12).reshape(24)


def g2_gen_words():
    return ints_to_words([c * FQ_R % P_MOD for c in (G2_GEN[0][0], G2_GEN[0][1], G2_GEN[1][0], G2_GEN[1][1])], 12).reshape(48)


class SplitMix64:
    """Documented tiny PRNG for seeds (seed 0xC0FFEE + size)."""

    def __init__(self, seed):
        self.s = seed & 0xFFFFFFFFFFFFFFFF

    def next(self):
        self.s = (self.s + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        return z ^ (z >> 31)

    def fr(self):
        return (self.next() | self.next() << 64 | self.next() << 128 | self.next() << 192) % R_MOD


def mul_chain(log_n, seed=0xC0FFEE):
    """MulChain(k): instance {1, x_out}; x_i * x_{i+1} = x_{i+2}; num_constraints + num_instance = 2^k.
    Returns (r1cs dict, z_mont (m, 8), z ints)."""
    rng = SplitMix64(seed + log_n)
    nc = (1 << log_n) - 2
    xs = [rng.fr(), rng.fr()]
    for i in range(nc):
        xs.append(xs[i] * xs[i + 1] % R_MOD)
    z = [1, xs[nc + 1]] + xs[:nc + 1]
    one = fr_mont([1])
    idx = np.arange(nc, dtype=np.int64)
    col = lambda k: np.where(idx + k == nc + 1, 1, 2 + idx + k).astype(np.uint32)
    rp = np.arange(nc + 1, dtype=np.uint64)
    val = np.repeat(one, nc, axis=0)
    r1cs = {"num_instance": 2, "num_witness": nc + 1, "num_constraints": nc,
            "a": {"row_ptr": rp, "col": col(0), "val": val},
            "b": {"row_ptr": rp, "col": col(1), "val": val},
            "c": {"row_ptr": rp, "col": col(2), "val": val}}
    return r1cs, fr_mont(z), z


def poe_shape(n_data, seed=0xC0FFEE, gates=7500, bits=512):
    """PoEShape(n_data): an R1CS with the variable / constraint counts and the witness value mix of the
    reference's Proof-of-Encryption circuit (circuits-ark/src/encryption.rs:219-317):
    instance = {1, c1.x, c1.y, c2[0..n_data)}; `bits` booleanity checks, `gates` full-width multiplication
    gates (stand-ins for the two JubJub scalar multiplications and the Poseidon sponge), then 3 constraints
    per data element over byte-valued plaintext (is_zero pair + the ciphertext link c2 = m + mask).
    Satisfied by construction.  Returns (r1cs dict, z_mont, z ints)."""
    rng = SplitMix64(seed ^ (n_data * 0x9E37 + 1))
    ni = 3 + n_data
    wit = []
    rows = ([], [], [])  # A, B, C rows: lists of (coeff, ('i'|'w', index))
    NEG1 = R_MOD - 1

    def add(a, b, c):
        rows[0].append(a); rows[1].append(b); rows[2].append(c)

    for _ in range(bits):  # b * (1 - b) = 0
        wit.append(rng.next() & 1)
        j = len(wit) - 1
        add([(1, ("w", j))], [(1, ("i", 0)), (NEG1, ("w", j))], [])
    wit += [rng.fr(), rng.fr()]
    for _ in range(gates):  # w_{j-2} * w_{j-1} = w_j
        j = len(wit)
        wit.append(wit[j - 2] * wit[j - 1] % R_MOD)
        add([(1, ("w", j - 2))], [(1, ("w", j - 1))], [(1, ("w", j))])
    mask_j = len(wit) - 1
    mask = wit[mask_j]
    c2 = []
    for k in range(n_data):
        m = rng.next() & 0xFF
        inv = pow(m, -1, R_MOD) if m else 0
        zf = 0 if m else 1
        j = len(wit)
        wit += [m, inv, zf]
        c2.append((m + mask) % R_MOD)
        add([(1, ("w", j))], [(1, ("w", j + 1))], [(1, ("i", 0)), (NEG1, ("w", j + 2))])   # m * inv = 1 - z
        add([(1, ("w", j))], [(1, ("w", j + 2))], [])                                        # m * z = 0
        add([(1, ("i", 3 + k)), (NEG1, ("w", mask_j))], [(1, ("i", 0))], [(1, ("w", j))])   # (c2 - mask) * 1 = m
    z = [1, rng.fr(), rng.fr()] + c2 + wit
    one_m, neg_m = fr_mont([1])[0], fr_mont([NEG1])[0]
    r1cs = {"num_instance": ni, "num_witness": len(wit), "num_constraints": len(rows[0])}
    for name, rr in zip("abc", rows):
        rp, col, val = [0], [], []
        for row in rr:
            for cf, (kind, idx) in row:
                col.append(idx if kind == "i" else ni + idx)
                val.append(one_m if cf == 1 else neg_m)
            rp.append(len(col))
        r1cs[name] = {"row_ptr": np.array(rp, dtype=np.uint64), "col": np.array(col, dtype=np.uint32),
                      "val": np.array(val, dtype=np.uint32).reshape(-1, 8)}
    return r1cs, fr_mont(z), z


def poe_shape_witnesses(ctx, n_data, count, seed=0xC0FFEE, gates=7500, bits=512):
    """`count` INDEPENDENT satisfying assignments of PoEShape(n_data) (the matrices of poe_shape do not depend on the seed),
    as (count, m, 8) Montgomery words -- one per buyer request / ciphertext block (lib/src/seller.rs:136-159,
    docs/protocol.md:56).  The multiplication chain is advanced for all assignments at once with the device's Fr product
    (Montgomery images multiply to Montgomery images), everything else is numpy table look-ups."""
    FR = 0   # CTG_FIELD_FR
    ADD = 1  # CTG_OP_ADD
    g = np.random.default_rng(seed ^ (n_data * 0x9E37 + 1))
    ni = 3 + n_data
    nw = bits + 2 + gates + 3 * n_data
    z = np.zeros((count, ni + nw, 8), dtype=np.uint32)
    tab = fr_mont(list(range(256)))                                        # bytes
    inv = fr_mont([pow(v, -1, R_MOD) if v else 0 for v in range(256)])
    z[:, 0] = tab[1]

    def rand_fr(k):
        w = g.integers(0, 1 << 32, size=(k, 8), dtype=np.uint64).astype(np.uint32)
        w[:, 7] &= 0x3FFFFFFF                                              # < 2^254 < r: a valid Montgomery image of SOME element
        return w
    z[:, 1], z[:, 2] = rand_fr(count), rand_fr(count)
    w0 = ni
    z[:, w0:w0 + bits] = tab[g.integers(0, 2, size=(count, bits))]
    j = w0 + bits
    a, b = rand_fr(count), rand_fr(count)
    z[:, j], z[:, j + 1] = a, b
    j += 2
    for _ in range(gates):
        a, b = b, ctx.fr_mul(a, b)
        z[:, j] = b
        j += 1
    mask = b
    if n_data:
        m = g.integers(0, 256, size=(count, n_data))
        z[:, j + 0:j + 3 * n_data:3] = tab[m]
        z[:, j + 1:j + 3 * n_data:3] = inv[m]
        z[:, j + 2:j + 3 * n_data:3] = tab[(m == 0).astype(np.int64)]
        c2 = ctx.field_op(FR, ADD, tab[m].reshape(-1, 8), np.repeat(mask, n_data, axis=0))
        z[:, 3:3 + n_data] = c2.reshape(count, n_data, 8)
    return z


def r1cs_is_satisfied(r1cs, z):
    """Host check with Python ints (small instances only)."""
    one_m = tuple(fr_mont([1])[0])

    def dots(m):
        out = []
        vals = [1 if tuple(v) == one_m else R_MOD - 1 for v in m["val"]]
        for i in range(r1cs["num_constraints"]):
            lo, hi = int(m["row_ptr"][i]), int(m["row_ptr"][i + 1])
            out.append(sum(vals[k] * z[int(m["col"][k])] for k in range(lo, hi)) % R_MOD)
        return out
    a, b, c = dots(r1cs["a"]), dots(r1cs["b"]), dots(r1cs["c"])
    return all(x * y % R_MOD == w for x, y, w in zip(a, b, c))


def domain_size(r1cs):
    n = 1
    while n < r1cs["num_instance"] + r1cs["num_constraints"]:
        n *= 2
    return n


def progression_pk(ctx, r1cs, seed=0xC0FFEE):
    """A proving key of the right shape whose queries are progressions generated on the device.
    Returns (pk dict of host arrays, dlogs dict) where dlogs holds the scalars behind every element:
    single elements as ints, queries as (start, step)."""
    rng = SplitMix64(seed ^ 0xA5A5)
    ni, nw = r1cs["num_instance"], r1cs["num_witness"]
    m, n = ni + nw, domain_size(r1cs)
    g1, g2 = g1_gen_words(), g2_gen_words()
    dl = {k: rng.fr() for k in ("alpha", "beta", "delta")}
    dl.update({k: (rng.fr(), rng.fr()) for k in ("a", "b", "h", "l")})
    pk = {}

    def single(group, k):
        b = (ctx.bases_progression_g1 if group == 1 else ctx.bases_progression_g2)(g1 if group == 1 else g2, fr_mont([k])[0], fr_mont([0])[0], 1, copies=1)
        out = b.download()[0]
        b.free()
        return out

    def query(group, key, count):
        a, b = dl[key]
        if count == 0:
            return np.zeros((0, 24 if group == 1 else 48), dtype=np.uint32)
        t = (ctx.bases_progression_g1 if group == 1 else ctx.bases_progression_g2)(g1 if group == 1 else g2, fr_mont([a])[0], fr_mont([b])[0], count, copies=1)
        out = t.download()
        t.free()
        return out

    pk["alpha_g1"] = single(1, dl["alpha"]); pk["beta_g1"] = single(1, dl["beta"]); pk["delta_g1"] = single(1, dl["delta"])
    pk["beta_g2"] = single(2, dl["beta"]); pk["delta_g2"] = single(2, dl["delta"])
    pk["a_query"] = query(1, "a", m); pk["b_g1_query"] = query(1, "b", m); pk["b_g2_query"] = query(2, "b", m)
    pk["h_query"] = query(1, "h", n - 1); pk["l_query"] = query(1, "l", nw)
    return pk, dl


def trapdoor(seed=0xC0FFEE):
    """A seeded Groth16 trapdoor for Context.generate_parameters: Montgomery words alpha, beta, gamma, delta, tau and the
    standard generators (arkworks draws random generators; any subgroup generator gives a valid key).  Returns
    (words dict, ints dict)."""
    rng = SplitMix64(seed ^ 0x7D00)
    ints = {k: rng.fr() or 1 for k in ("alpha", "beta", "gamma", "delta", "tau")}
    words = {k: fr_mont([v])[0] for k, v in ints.items()}
    words["g1"], words["g2"] = g1_gen_words(), g2_gen_words()
    return words, ints


def weighted_sums(sc_words, first_index=0):
    """(sum_i s_i, sum_i (first_index + i) * s_i) as Python integers for an (n, 8) uint32 array of 256-bit scalars -- numpy on
    16-bit half-limbs in chunks of 2^20 (every partial dot product stays below 2^63), so that the discrete-log closed form of
    an MSM over progression bases costs O(n) machine arithmetic even at 2^24 points."""
    sc = np.ascontiguousarray(sc_words, dtype=np.uint32)
    n = sc.shape[0]
    halves = sc.view(np.uint16).reshape(n, 16).astype(np.uint64)
    plain = [int(x) for x in halves.sum(axis=0, dtype=np.uint64)] if n < (1 << 47) else None
    weighted = [0] * 16
    CH = 1 << 20
    for lo in range(0, n, CH):
        hi = min(n, lo + CH)
        idx = np.arange(lo, hi, dtype=np.uint64)          # local index < 2^24 per chunk offset: (idx - lo) < 2^20
        loc = idx - np.uint64(lo)
        blk = halves[lo:hi]
        dots = loc @ blk                                   # < 2^20 * 2^20 * 2^16 = 2^56
        sums = blk.sum(axis=0, dtype=np.uint64)
        for k in range(16):
            weighted[k] += int(dots[k]) + (first_index + lo) * int(sums[k])
    s0 = sum(v << (16 * k) for k, v in enumerate(plain))
    s1 = sum(v << (16 * k) for k, v in enumerate(weighted))
    return s0, s1
