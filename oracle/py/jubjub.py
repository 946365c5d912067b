"""TEST INFRASTRUCTURE ONLY -- big-int JubJub (the twisted Edwards curve over BLS12-381 Fr the reference encrypts on) and the
reference's hashed-ElGamal encrypt / decrypt restated on it.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg may import this.

PARITY UNPINNED: ark-ed-on-bls12-381 / ark-ec ^0.3.0 are crates.io dependencies absent from /root/reference.  Restated:
  curve  a x^2 + y^2 = 1 + d x^2 y^2 with a = -1, d = -(10240/10241) mod r (ark-ed-on-bls12-381 curves/mod.rs), cofactor 8,
         prime subgroup order SUBGROUP_ORDER, generator (GEN_X, GEN_Y) = EdwardsParameters::AFFINE_GENERATOR_COEFFS;
         the constants are checked below (on the curve, order exactly SUBGROUP_ORDER)
  group law: the unified affine law (used here because the CUDA kernel uses extended coordinates -- two different routes)
  reference call sites: circuits-ark/src/encryption.rs:126-137 keygen (pk = G * sk), :154-172 encrypt
         (c1 = G * r, p_r = pk * r, dh = PoseidonSponge.absorb(p_r affine = [x, y]).squeeze(1), c2[i] = dh + m[i]),
         :174-192 decrypt (c2[i] - sponge(c1 * sk)).
"""
from __future__ import annotations

from . import poseidon as P
from .bls12_381 import R_MOD

A = R_MOD - 1
D = (-10240 * pow(10241, -1, R_MOD)) % R_MOD
SUBGROUP_ORDER = 6554484396890773809930967563523245729705921265872317281365359162392183254199
GEN_X = 8076246640662884909881801758704306714034609987455869804520522091855516602923
GEN_Y = 13262374693698910701929044844600465831413122818447359594527400194675274060458
GEN = (GEN_X, GEN_Y)
IDENTITY = (0, 1)


def on_curve(pt, d=D):
    x, y = pt
    return (A * x * x + y * y) % R_MOD == (1 + d * x * x % R_MOD * y * y) % R_MOD


def add(p, q, d=D):
    x1, y1 = p
    x2, y2 = q
    t = d * x1 * x2 % R_MOD * y1 * y2 % R_MOD
    x3 = (x1 * y2 + y1 * x2) * pow(1 + t, -1, R_MOD) % R_MOD
    y3 = (y1 * y2 - A * x1 * x2) * pow(1 - t, -1, R_MOD) % R_MOD
    return (x3, y3)


def mul(p, k, d=D):
    """ark-ec mul: double-and-add over the bits of k, most significant first."""
    acc = IDENTITY
    for bit in bin(k)[2:] if k else "":
        acc = add(acc, acc, d)
        if bit == "1":
            acc = add(acc, p, d)
    return acc


def point_from_x(x, d=D):
    """A curve point with this x, or None: y^2 = (1 + x^2) / (1 - d x^2); square root by Tonelli-Shanks (r - 1 = 2^32 * odd)."""
    num, den = (1 + x * x) % R_MOD, (1 - d * x * x) % R_MOD
    if den == 0:
        return None
    y2 = num * pow(den, -1, R_MOD) % R_MOD
    if pow(y2, (R_MOD - 1) // 2, R_MOD) != 1:
        return None if y2 else (x, 0)
    q, s = R_MOD - 1, 0
    while q % 2 == 0:
        q //= 2
        s += 1
    z = 7                                   # 7 generates Fr* (ark-bls12-381 Fr GENERATOR), so it is a non-residue
    m, c, t, r = s, pow(z, q, R_MOD), pow(y2, q, R_MOD), pow(y2, (q + 1) // 2, R_MOD)
    while t != 1:
        i, t2 = 0, t
        while t2 != 1:
            t2 = t2 * t2 % R_MOD
            i += 1
        b = pow(c, 1 << (m - i - 1), R_MOD)
        m, c = i, b * b % R_MOD
        t, r = t * c % R_MOD, r * b % R_MOD
    assert r * r % R_MOD == y2
    return (x, r)


def keygen(sk):
    return mul(GEN, sk)


def encrypt(params, pk, msg, r):
    """(c1, c2) of circuits-ark/src/encryption.rs:154-172."""
    c1 = mul(GEN, r)
    p_r = mul(pk, r)
    sponge = P.PoseidonSponge(params)
    sponge.absorb([p_r[0], p_r[1]])
    dh = sponge.squeeze(1)[0]
    return c1, [(dh + m) % R_MOD for m in msg]


def decrypt(params, cipher, sk):
    c1, c2 = cipher
    s = mul(c1, sk)
    sponge = P.PoseidonSponge(params)
    sponge.absorb([s[0], s[1]])
    dh = sponge.squeeze(1)[0]
    return [(c - dh) % R_MOD for c in c2]


def selfcheck():
    assert on_curve(GEN) and mul(GEN, SUBGROUP_ORDER) == IDENTITY and mul(GEN, 1) == GEN
    assert pow(D, (R_MOD - 1) // 2, R_MOD) == R_MOD - 1, "d must be a non-square (complete addition law)"


# ---- ark-serialize 0.3 canonical form of a twisted Edwards affine point (ark-ec 0.3.0 twisted_edwards_extended.rs, restated):
#   serialize   : the identity is x = 0 with no flag; otherwise x (32 bytes little-endian canonical) with EdwardsFlags in the top
#                 bit of the last byte: 0x80 = Positive (y > -y as integers), 0 = Negative (the default flag).  PINNED by the
#                 reference: ALICE_PK has the bit set, and only the root with y > -y lies in the prime-order subgroup, which the
#                 reference's test_public_key_decode (circuits-ark/src/utils.rs:117-121) requires by unwrapping the decode
#   deserialize : x == 0 -> identity; else get_point_from_x(x, greatest = positive) (y^2 = (1 - a x^2) / (1 - d x^2), the larger
#                 root when greatest), InvalidData if no root or if the point is outside the prime-order subgroup
# Reference call sites: ark_from_bytes::<JubJub> / ark_to_bytes on the ElGamal keys (circuits-ark/src/utils.rs:12-22, :117-121);
# the only key bytes the reference holds are ALICE_PK / ALICE_SK (circuits-ark/src/utils.rs:114-115).
def point_compress(pt):
    x, y = pt
    if pt == IDENTITY:
        return (0).to_bytes(32, "little")
    raw = bytearray(x.to_bytes(32, "little"))
    if y > (R_MOD - y) % R_MOD:
        raw[31] |= 0x80
    return bytes(raw)


def point_decompress(data, d=D, check_subgroup=True):
    """-> affine point, or None where arkworks returns SerializationError::InvalidData."""
    if len(data) != 32:
        return None
    negative = not bool(data[31] & 0x80)
    x = int.from_bytes(bytes(data[:31]) + bytes([data[31] & 0x7F]), "little")
    if x >= R_MOD:
        return None
    if x == 0:
        return IDENTITY
    p = point_from_x(x, d)
    if p is None:
        return None
    y = p[1]
    ny = (R_MOD - y) % R_MOD
    big, small = (y, ny) if y > ny else (ny, y)
    p = (x, small if negative else big)
    if check_subgroup and mul(p, SUBGROUP_ORDER, d) != IDENTITY:
        return None
    return p


def scalar_from_bytes(data):
    """ark_from_bytes::<Fr> of 32 little-endian bytes: the integer, or None if it is not below the modulus (the reference decodes
    ALICE_SK as ark_bls12_381::Fr, circuits-ark/src/utils.rs:123-127)."""
    v = int.from_bytes(data, "little")
    return v if v < R_MOD else None
