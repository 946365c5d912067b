"""TEST INFRASTRUCTURE ONLY -- big-int restatement of arkworks' canonical serialisation for BLS12-381 points,
Groth16 keys and proofs.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.

PARITY UNPINNED at byte level: ark-serialize / ark-ec ^0.3.0 are crates.io dependencies absent from
/root/reference (circuits-ark/Cargo.toml:10-25); the reference holds no serialized Groth16 key or proof to
compare with (its only byte goldens, circuits-ark/src/utils.rs:114-115, are JubJub keys on a twisted-Edwards
model).  Conventions restated, from ark-ec 0.3.0 models/short_weierstrass_jacobian.rs
(CanonicalSerialize for GroupAffine) and ark-serialize 0.3.0 flags.rs (SWFlags):
  Fq = 48 bytes LE canonical; Fq2 = c0 || c1; flags in the top two bits of the last byte (0x80: y > -y,
  0x40: infinity); compressed = x | flags; uncompressed = x || (y | flags), identity = (0, 1 | 0x40);
  Vec<T> = u64-LE length || items; structs field by field.
Reference call sites: lib/src/utils.rs:85-118 (circuit.zkey via serialize_unchecked / deserialize_unchecked,
verification.key via ark_to_bytes), circuits-ark/src/utils.rs:12-22 (ark_to_bytes / ark_from_bytes on proofs).
"""
from __future__ import annotations

import struct

from .bls12_381 import P_MOD

FLAG_LARGER, FLAG_INF = 0x80, 0x40


def _fq(v):
    return int(v).to_bytes(48, "little")


def _coords(P, group):
    if group == 1:
        return [P[0]], [P[1]]
    return [P[0][0], P[0][1]], [P[1][0], P[1][1]]


def _y_is_larger(ycs):
    """y > -y in arkworks' order: Fq by integer value; Fq2 compares c1 first, then c0."""
    for c in reversed(ycs):
        if c != 0:
            return c > P_MOD - c
    return False


def point_compressed(P, group):
    nx = 1 if group == 1 else 2
    if P is None:
        out = bytearray(48 * nx)
        out[-1] |= FLAG_INF
        return bytes(out)
    xs, ys = _coords(P, group)
    out = bytearray(b"".join(_fq(c) for c in xs))
    if _y_is_larger(ys):
        out[-1] |= FLAG_LARGER
    return bytes(out)


def point_uncompressed(P, group):
    nx = 1 if group == 1 else 2
    if P is None:
        out = bytearray(96 * nx)
        out[48 * nx] = 1  # y = one
        out[-1] |= FLAG_INF
        return bytes(out)
    xs, ys = _coords(P, group)
    return b"".join(_fq(c) for c in xs + ys)


def _vec(points, group, fn):
    return struct.pack("<Q", len(points)) + b"".join(fn(P, group) for P in points)


def proving_key_unchecked(pk):
    """ProvingKey::serialize_unchecked; pk: oracle.py.groth16.ProvingKey (affine tuples / None)."""
    u = point_uncompressed
    return b"".join([
        u(pk.alpha_g1, 1), u(pk.beta_g2, 2), u(pk.gamma_g2, 2), u(pk.delta_g2, 2), _vec(pk.gamma_abc_g1, 1, u),
        u(pk.beta_g1, 1), u(pk.delta_g1, 1), _vec(pk.a_query, 1, u), _vec(pk.b_g1_query, 1, u), _vec(pk.b_g2_query, 2, u),
        _vec(pk.h_query, 1, u), _vec(pk.l_query, 1, u)])


def verifying_key_compressed(pk):
    c = point_compressed
    return b"".join([c(pk.alpha_g1, 1), c(pk.beta_g2, 2), c(pk.gamma_g2, 2), c(pk.delta_g2, 2), _vec(pk.gamma_abc_g1, 1, c)])


def proof_compressed(proof):
    a, b, c = proof
    return point_compressed(a, 1) + point_compressed(b, 2) + point_compressed(c, 1)


# ---- reading (independent of the writer above: decompression by square root) ----------------------------
def _sqrt_fq(a):
    r = pow(a, (P_MOD + 1) // 4, P_MOD)
    return r if r * r % P_MOD == a % P_MOD else None


def _sqrt_fq2(a):
    """Complex method for p = 3 mod 4 (Adj, Rodriguez-Henriquez, Alg. 9) -- a different route from the norm-based one
    in csrc/codec.cu."""
    from .bls12_381 import Fq2Ops as F
    if a == (0, 0):
        return (0, 0)

    def fpow(x, e):
        acc = F.one
        while e:
            if e & 1:
                acc = F.mul(acc, x)
            x = F.sqr(x)
            e >>= 1
        return acc
    a1 = fpow(a, (P_MOD - 3) // 4)
    alpha = F.mul(F.sqr(a1), a)
    a0 = F.mul(fpow(alpha, P_MOD), alpha)
    if a0 == (P_MOD - 1, 0):
        return None
    x0 = F.mul(a1, a)
    if alpha == (P_MOD - 1, 0):
        x = F.mul((0, 1), x0)
    else:
        b = fpow(F.add(F.one, alpha), (P_MOD - 1) // 2)
        x = F.mul(b, x0)
    return x if F.sqr(x) == a else None


def point_from_compressed(data, group):
    """GroupAffine::deserialize without the subgroup check; raises ValueError on invalid data."""
    nx = 1 if group == 1 else 2
    if len(data) != 48 * nx:
        raise ValueError("length")
    flags = data[-1] & 0xC0
    body = bytearray(data)
    body[-1] &= 0x3F
    xs = [int.from_bytes(body[48 * k:48 * k + 48], "little") for k in range(nx)]
    if any(x >= P_MOD for x in xs) or flags == 0xC0:
        raise ValueError("InvalidData")
    if flags & FLAG_INF:
        return None
    if group == 1:
        y = _sqrt_fq((xs[0] ** 3 + 4) % P_MOD)
        if y is None:
            raise ValueError("not on curve")
        if _y_is_larger([y]) != bool(flags & FLAG_LARGER):
            y = (-y) % P_MOD
        return (xs[0], y)
    from .bls12_381 import Fq2Ops as F
    x = (xs[0], xs[1])
    y = _sqrt_fq2(F.add(F.mul(F.sqr(x), x), (4, 4)))
    if y is None:
        raise ValueError("not on curve")
    if _y_is_larger([y[0], y[1]]) != bool(flags & FLAG_LARGER):
        y = F.neg(y)
    return (x, y)


def proof_from_compressed(data):
    if len(data) != 192:
        raise ValueError("length")
    return point_from_compressed(data[:48], 1), point_from_compressed(data[48:144], 2), point_from_compressed(data[144:], 1)
