"""TEST INFRASTRUCTURE ONLY -- slow big-int oracle for BLS12-381 (Fr, Fq, Fq2, G1, G2).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.
PARITY UNPINNED at byte level: the reference (/root/reference, nulltea/contangle-zkcp)
holds no golden proof / MSM / FFT vectors (its Groth16 tests only assert accept,
circuits-ark/src/encryption.rs:403-411) and its arithmetic lives in crates.io
dependencies that are absent from the tree: ark-ff / ark-ec / ark-poly / ark-groth16 /
ark-bls12-381, all "^0.3.0" (circuits-ark/Cargo.toml:10-25, Cargo.lock git-ignored).
This file restates the PUBLISHED algorithms/conventions of those crates with Python
integers; it is deliberately independent of oracle/c (different language, different
limb model) so the two can pin each other.  Pins used instead of reference vectors:
  * the standard BLS12-381 constants (generators, r, p, BLS parameter z) self-checked
    below (on-curve, r*G = O, r = z^4 - z^2 + 1, p = (z-1)^2 r / 3 + z);
  * algebraic identities and the Groth16 trapdoor closed form (oracle/py/groth16.py).

Call sites in the reference that reach this arithmetic: lib/src/lib.rs:21 (Bls12_381
engine), lib/src/zk/encryption.rs:76 (Groth16::prove).
"""
from __future__ import annotations

# ---------------------------------------------------------------- parameters
# ark-bls12-381 0.3.0 fields/{fr,fq}.rs
R_MOD = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001
P_MOD = 0x1A0111EA397FE69A4B1BA7B6434BACD764774B84F38512BF6730D2A0F6B0F6241EABFFFEB153FFFFB9FEFFFFFFFFAAAB
BLS_Z = -0xD201000000010000
assert R_MOD == BLS_Z**4 - BLS_Z**2 + 1
assert P_MOD == ((BLS_Z - 1) ** 2 * R_MOD) // 3 + BLS_Z

FR_LIMBS32, FQ_LIMBS32 = 8, 12
FR_R = 1 << 256  # Montgomery radix, ark-ff Fp256 (4 x u64)
FQ_R = 1 << 384  # ark-ff Fp384 (6 x u64)
FR_GENERATOR = 7  # Fr::multiplicative_generator(), also the coset shift
FR_TWO_ADICITY = 32
FR_TWO_ADIC_ROOT = pow(FR_GENERATOR, (R_MOD - 1) >> FR_TWO_ADICITY, R_MOD)
assert FR_TWO_ADIC_ROOT == 10238227357739495823651030575849232062558860180284477541189508159991286009131

G1_B = 4
G1_GEN = (
    0x17F1D3A73197D7942695638C4FA9AC0FC3688C4F9774B905A14E3A3F171BAC586C55E83FF97A1AEFFB3AF00ADB22C6BB,
    0x08B3F481E3AAA0F1A09E30ED741D8AE4FCF5E095D5D00AF600DB18CB2C04B3EDD03CC744A2888AE40CAA232946C5E7E1,
)
G2_B = (4, 4)
G2_GEN = (
    (0x024AA2B2F08F0A91260805272DC51051C6E47AD4FA403B02B4510B647AE3D1770BAC0326A805BBEFD48056C8C121BDB8,
     0x13E02B6052719F607DACD3A088274F65596BD0D09920B61AB5DA61BBDC7F5049334CF11213945D57E5AC7D055D042B7E),
    (0x0CE5D527727D6E118CC9CDC6DA2E351AADFD9BAA8CBDD3A76D429A695160D12C923AC9CC3BACA289E193548608B82801,
     0x0606C4A02EA734CC32ACD2B02BC28B99CB3E287E85A763AF267492AB572E99AB3F370D275CEC1DA1AAA9075FF05F79BE),
)


# ---------------------------------------------------------------- field ops
class FqOps:
    """Fq as Python ints in [0, p)."""
    zero, one = 0, 1

    @staticmethod
    def add(a, b): return (a + b) % P_MOD
    @staticmethod
    def sub(a, b): return (a - b) % P_MOD
    @staticmethod
    def mul(a, b): return (a * b) % P_MOD
    @staticmethod
    def sqr(a): return (a * a) % P_MOD
    @staticmethod
    def neg(a): return (-a) % P_MOD
    @staticmethod
    def inv(a): return pow(a, P_MOD - 2, P_MOD)
    @staticmethod
    def is_zero(a): return a == 0
    @staticmethod
    def small(k): return k % P_MOD


class Fq2Ops:
    """Fq2 = Fq[u]/(u^2+1), elements are (c0, c1). ark-ff fp2.rs / ark-bls12-381 fq2.rs."""
    zero, one = (0, 0), (1, 0)

    @staticmethod
    def add(a, b): return ((a[0] + b[0]) % P_MOD, (a[1] + b[1]) % P_MOD)
    @staticmethod
    def sub(a, b): return ((a[0] - b[0]) % P_MOD, (a[1] - b[1]) % P_MOD)
    @staticmethod
    def mul(a, b):
        return ((a[0] * b[0] - a[1] * b[1]) % P_MOD, (a[0] * b[1] + a[1] * b[0]) % P_MOD)
    @staticmethod
    def sqr(a):
        return ((a[0] + a[1]) * (a[0] - a[1]) % P_MOD, 2 * a[0] * a[1] % P_MOD)
    @staticmethod
    def neg(a): return ((-a[0]) % P_MOD, (-a[1]) % P_MOD)
    @staticmethod
    def inv(a):
        n = pow((a[0] * a[0] + a[1] * a[1]) % P_MOD, P_MOD - 2, P_MOD)
        return (a[0] * n % P_MOD, (-a[1]) * n % P_MOD)
    @staticmethod
    def is_zero(a): return a[0] == 0 and a[1] == 0
    @staticmethod
    def small(k): return (k % P_MOD, 0)


# ---------------------------------------------------------------- curves
class Curve:
    """Short Weierstrass y^2 = x^3 + b over field F (a = 0).

    Affine points are (x, y) or None (infinity); Jacobian points are (X, Y, Z) with
    x = X/Z^2, y = Y/Z^3 (ark-ec models/short_weierstrass_jacobian.rs).  Only the
    affine value of any result is observable, so formula choice is free.
    """

    def __init__(self, F, b, gen, name):
        self.F, self.b, self.gen, self.name = F, b, gen, name

    def on_curve(self, P):
        if P is None:
            return True
        F = self.F
        x, y = P
        return F.sqr(y) == F.add(F.mul(F.sqr(x), x), self.b)

    # --- Jacobian arithmetic
    def to_jac(self, P):
        F = self.F
        return (F.one, F.one, F.zero) if P is None else (P[0], P[1], F.one)

    def to_affine(self, J):
        F = self.F
        X, Y, Z = J
        if F.is_zero(Z):
            return None
        zi = F.inv(Z)
        zi2 = F.sqr(zi)
        return (F.mul(X, zi2), F.mul(Y, F.mul(zi2, zi)))

    def jdbl(self, J):
        F = self.F
        X, Y, Z = J
        if F.is_zero(Z):
            return J
        A = F.sqr(X); B = F.sqr(Y); C = F.sqr(B)
        t = F.sub(F.sub(F.sqr(F.add(X, B)), A), C)
        D = F.add(t, t)
        E = F.add(F.add(A, A), A)
        Fv = F.sqr(E)
        X3 = F.sub(Fv, F.add(D, D))
        C8 = F.add(C, C); C8 = F.add(C8, C8); C8 = F.add(C8, C8)
        Y3 = F.sub(F.mul(E, F.sub(D, X3)), C8)
        YZ = F.mul(Y, Z)
        return (X3, Y3, F.add(YZ, YZ))

    def jadd(self, J1, J2):
        F = self.F
        X1, Y1, Z1 = J1
        X2, Y2, Z2 = J2
        if F.is_zero(Z1):
            return J2
        if F.is_zero(Z2):
            return J1
        Z1Z1 = F.sqr(Z1); Z2Z2 = F.sqr(Z2)
        U1 = F.mul(X1, Z2Z2); U2 = F.mul(X2, Z1Z1)
        S1 = F.mul(F.mul(Y1, Z2), Z2Z2); S2 = F.mul(F.mul(Y2, Z1), Z1Z1)
        if U1 == U2:
            if S1 == S2:
                return self.jdbl(J1)
            return (F.one, F.one, F.zero)
        H = F.sub(U2, U1); Rr = F.sub(S2, S1)
        HH = F.sqr(H); HHH = F.mul(H, HH); V = F.mul(U1, HH)
        X3 = F.sub(F.sub(F.sqr(Rr), HHH), F.add(V, V))
        Y3 = F.sub(F.mul(Rr, F.sub(V, X3)), F.mul(S1, HHH))
        Z3 = F.mul(F.mul(Z1, Z2), H)
        return (X3, Y3, Z3)

    def jneg(self, J):
        return (J[0], self.F.neg(J[1]), J[2])

    def jmul(self, J, k):
        F = self.F
        acc = (F.one, F.one, F.zero)
        if k < 0:
            J, k = self.jneg(J), -k
        for bit in bin(k)[2:] if k else "":
            acc = self.jdbl(acc)
            if bit == "1":
                acc = self.jadd(acc, J)
        return acc

    # --- affine conveniences
    def add(self, P, Q):
        return self.to_affine(self.jadd(self.to_jac(P), self.to_jac(Q)))

    def neg(self, P):
        return None if P is None else (P[0], self.F.neg(P[1]))

    def mul(self, P, k):
        return self.to_affine(self.jmul(self.to_jac(P), k))

    def batch_to_affine(self, Js):
        """Montgomery batch inversion of the Z coordinates."""
        F = self.F
        pref, acc = [], F.one
        for J in Js:
            pref.append(acc)
            if not F.is_zero(J[2]):
                acc = F.mul(acc, J[2])
        inv = F.inv(acc)
        out = [None] * len(Js)
        for i in range(len(Js) - 1, -1, -1):
            X, Y, Z = Js[i]
            if F.is_zero(Z):
                continue
            zi = F.mul(inv, pref[i])
            inv = F.mul(inv, Z)
            zi2 = F.sqr(zi)
            out[i] = (F.mul(X, zi2), F.mul(Y, F.mul(zi2, zi)))
        return out

    def msm_naive(self, bases, scalars):
        """sum s_i * P_i, double-and-add per term (independent of any bucket method)."""
        acc = self.to_jac(None)
        for P, s in zip(bases, scalars):
            if P is None or s == 0:
                continue
            acc = self.jadd(acc, self.jmul(self.to_jac(P), s))
        return self.to_affine(acc)

    def msm(self, bases, scalars, c=None):
        """ark-ec 0.3.0 msm/variable_base.rs: unsigned-window Pippenger, zero scalars
        skipped, scalar == 1 added directly in window 0, running-sum bucket reduction,
        windows combined high -> low with c doublings."""
        n = min(len(bases), len(scalars))
        if c is None:
            c = 3 if n < 32 else ark_ln_without_floats(n) + 2
        nbits = 255
        zero = self.to_jac(None)
        window_sums = []
        for w_start in range(0, nbits, c):
            res = zero
            buckets = [zero] * ((1 << c) - 1)
            for i in range(n):
                s = scalars[i]
                if s == 0:
                    continue
                if s == 1:
                    if w_start == 0:
                        res = self.jadd(res, self.to_jac(bases[i]))
                    continue
                d = (s >> w_start) & ((1 << c) - 1)
                if d:
                    buckets[d - 1] = self.jadd(buckets[d - 1], self.to_jac(bases[i]))
            running = zero
            for bkt in reversed(buckets):
                running = self.jadd(running, bkt)
                res = self.jadd(res, running)
            window_sums.append(res)
        total = window_sums[-1]
        for ws in reversed(window_sums[:-1]):
            for _ in range(c):
                total = self.jdbl(total)
            total = self.jadd(total, ws)
        return self.to_affine(total)


def ark_ln_without_floats(a: int) -> int:
    """ark-ec 0.3.0 msm::ln_without_floats: (log2(a) * 69 / 100) with ark_std::log2 =
    ceil(log2) for non powers of two."""
    lg = (a - 1).bit_length() if a > 1 else 0
    return lg * 69 // 100


G1 = Curve(FqOps, G1_B, G1_GEN, "G1")
G2 = Curve(Fq2Ops, G2_B, G2_GEN, "G2")


# ---------------------------------------------------------------- Fr helpers and NTT
def fr_inv(a): return pow(a, R_MOD - 2, R_MOD)


def fr_root_of_unity(log_n: int) -> int:
    """ark-ff FftField::get_root_of_unity(n) for n = 2^log_n (ark-poly radix2 domain)."""
    assert 0 <= log_n <= FR_TWO_ADICITY
    return pow(FR_TWO_ADIC_ROOT, 1 << (FR_TWO_ADICITY - log_n), R_MOD)


def bit_reverse(i: int, bits: int) -> int:
    return int(format(i, "0%db" % bits)[::-1], 2) if bits else 0


def ntt(vals, inverse=False, coset=False):
    """ark-poly 0.3.0 Radix2EvaluationDomain::{fft,ifft,coset_fft,coset_ifft}_in_place.
    Natural order in, natural order out; ifft scales by 1/n; coset shift g = 7:
    coset_fft(x) = fft(x_i g^i); coset_ifft(y)_i = ifft(y)_i g^-i."""
    n = len(vals)
    log_n = n.bit_length() - 1
    assert 1 << log_n == n
    a = list(vals)
    if coset and not inverse:
        g, acc = FR_GENERATOR, 1
        for i in range(n):
            a[i] = a[i] * acc % R_MOD
            acc = acc * g % R_MOD
    w = fr_root_of_unity(log_n)
    if inverse:
        w = fr_inv(w)
    a = [a[bit_reverse(i, log_n)] for i in range(n)]
    m = 1
    while m < n:
        wm = pow(w, n // (2 * m), R_MOD)
        for k in range(0, n, 2 * m):
            t = 1
            for j in range(m):
                u, v = a[k + j], a[k + j + m] * t % R_MOD
                a[k + j], a[k + j + m] = (u + v) % R_MOD, (u - v) % R_MOD
                t = t * wm % R_MOD
        m *= 2
    if inverse:
        ninv = fr_inv(n)
        a = [x * ninv % R_MOD for x in a]
        if coset:
            gi, acc = fr_inv(FR_GENERATOR), 1
            for i in range(n):
                a[i] = a[i] * acc % R_MOD
                acc = acc * gi % R_MOD
    return a


def poly_eval(coeffs, x):
    acc = 0
    for c in reversed(coeffs):
        acc = (acc * x + c) % R_MOD
    return acc


# ---------------------------------------------------------------- wire layout helpers
def to_mont(x: int, mod: int, R: int) -> int:
    return x * R % mod


def from_mont(x: int, mod: int, R: int) -> int:
    return x * pow(R, -1, mod) % mod


def int_to_limbs(x: int, n32: int):
    return [(x >> (32 * i)) & 0xFFFFFFFF for i in range(n32)]


def limbs_to_int(limbs) -> int:
    acc = 0
    for i, l in enumerate(limbs):
        acc |= int(l) << (32 * i)
    return acc


def selfcheck():
    assert G1.on_curve(G1_GEN) and G2.on_curve(G2_GEN)
    assert G1.mul(G1_GEN, R_MOD) is None and G2.mul(G2_GEN, R_MOD) is None
    assert pow(FR_TWO_ADIC_ROOT, 1 << 32, R_MOD) == 1 and pow(FR_TWO_ADIC_ROOT, 1 << 31, R_MOD) != 1
    return True
