"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/c/libark_oracle.so (the
arkworks-0.3.0 algorithm restatement, see oracle/c/ark_oracle.cpp header) plus numpy
helpers for the shared wire layout.  PARITY UNPINNED at byte level (no reference
vectors exist; see oracle/py/bls12_381.py header).  Only tests/, smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "c", "libark_oracle.so")
_lib = None

FR_W, FQ_W = 8, 12          # 32-bit words per element
G1_W, G2_W = 24, 48         # words per affine point


def build(force=False):
    src = os.path.join(_HERE, "c", "ark_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", os.path.join(_HERE, "c")], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.ark_get_threads.restype = C.c_int
        _lib.ark_g1_on_curve.restype = C.c_int
        _lib.ark_g2_on_curve.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, words=None):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    if words is not None:
        assert a.shape[-1] == words, (a.shape, words)
    return a


def set_threads(t):
    lib().ark_set_threads(int(t))


def threads():
    return lib().ark_get_threads()


# ------------------------------------------------------------ int <-> limb arrays
def ints_to_words(vals, words):
    buf = b"".join(int(v).to_bytes(4 * words, "little") for v in vals)
    return np.frombuffer(buf, dtype=np.uint32).reshape(len(vals), words).copy()


def words_to_ints(arr):
    arr = np.ascontiguousarray(arr, dtype=np.uint32)
    w = arr.shape[-1]
    raw = arr.reshape(-1, w).tobytes()
    return [int.from_bytes(raw[i * 4 * w:(i + 1) * 4 * w], "little") for i in range(arr.reshape(-1, w).shape[0])]


def fr_to_mont(canon):
    canon = _c(canon, FR_W); out = np.empty_like(canon)
    lib().ark_fr_to_mont(_p(canon), _p(out), C.c_uint64(canon.shape[0])); return out


def fr_from_mont(mont):
    mont = _c(mont, FR_W); out = np.empty_like(mont)
    lib().ark_fr_from_mont(_p(mont), _p(out), C.c_uint64(mont.shape[0])); return out


def fq_to_mont(canon):
    canon = _c(canon, FQ_W); out = np.empty_like(canon)
    lib().ark_fq_to_mont(_p(canon), _p(out), C.c_uint64(canon.shape[0])); return out


def fq_from_mont(mont):
    mont = _c(mont, FQ_W); out = np.empty_like(mont)
    lib().ark_fq_from_mont(_p(mont), _p(out), C.c_uint64(mont.shape[0])); return out


def fr_mont_from_ints(vals):
    return fr_to_mont(ints_to_words(vals, FR_W))


def fr_ints_from_mont(arr):
    return words_to_ints(fr_from_mont(arr))


def g1_from_py(points):
    """[(x, y) | None] (canonical ints) -> (n, 24) uint32 Montgomery affine images."""
    flat = []
    for P in points:
        flat += [0, 0] if P is None else [P[0], P[1]]
    m = fq_to_mont(ints_to_words(flat, FQ_W))
    return m.reshape(len(points), G1_W)


def g1_to_py(arr):
    arr = _c(arr).reshape(-1, G1_W)
    vals = words_to_ints(fq_from_mont(arr.reshape(-1, FQ_W)))
    out = []
    for i in range(arr.shape[0]):
        x, y = vals[2 * i], vals[2 * i + 1]
        out.append(None if x == 0 and y == 0 else (x, y))
    return out


def g2_from_py(points):
    flat = []
    for P in points:
        flat += [0, 0, 0, 0] if P is None else [P[0][0], P[0][1], P[1][0], P[1][1]]
    m = fq_to_mont(ints_to_words(flat, FQ_W))
    return m.reshape(len(points), G2_W)


def g2_to_py(arr):
    arr = _c(arr).reshape(-1, G2_W)
    vals = words_to_ints(fq_from_mont(arr.reshape(-1, FQ_W)))
    out = []
    for i in range(arr.shape[0]):
        v = vals[4 * i:4 * i + 4]
        out.append(None if not any(v) else ((v[0], v[1]), (v[2], v[3])))
    return out


# ------------------------------------------------------------ seeded synthetic inputs
def random_fr_mont(n, seed):
    """n pseudo-random Fr elements as Montgomery images, (n, 8) uint32.  Any value < r
    is a valid image; 256 random bits are reduced through to_mont."""
    rng = np.random.Generator(np.random.PCG64(seed))
    raw = rng.integers(0, 1 << 32, size=(n, FR_W), dtype=np.uint64).astype(np.uint32)
    return fr_to_mont(raw)


def random_scalars_canon(n, seed, kind="uniform"):
    """Canonical scalars (BigInteger256 images) for MSM, (n, 8) uint32.
    kind: uniform | poe (40% zero, 25% one, 15% bytes, 20% uniform) | equal | rminus1."""
    rng = np.random.Generator(np.random.PCG64(seed))
    uni = fr_from_mont(random_fr_mont(n, seed ^ 0x5EED))
    if kind == "uniform":
        return uni
    if kind == "poe":
        sel = rng.random(n)
        out = uni.copy()
        z = sel < 0.40
        o = (sel >= 0.40) & (sel < 0.65)
        b = (sel >= 0.65) & (sel < 0.80)
        out[z] = 0
        out[o] = 0; out[o, 0] = 1
        out[b] = 0; out[b, 0] = rng.integers(0, 256, size=int(b.sum()), dtype=np.uint64).astype(np.uint32)
        return out
    if kind == "equal":
        return np.repeat(uni[:1], n, axis=0)
    if kind == "rminus1":
        from .py.bls12_381 import R_MOD
        return np.repeat(ints_to_words([R_MOD - 1], FR_W), n, axis=0)
    raise ValueError(kind)


# ------------------------------------------------------------ oracle entry points
def fr_mul(a, b):
    a, b = _c(a, FR_W), _c(b, FR_W); out = np.empty_like(a)
    lib().ark_fr_mul_batch(_p(a), _p(b), _p(out), C.c_uint64(a.shape[0])); return out


def fq_mul(a, b):
    a, b = _c(a, FQ_W), _c(b, FQ_W); out = np.empty_like(a)
    lib().ark_fq_mul_batch(_p(a), _p(b), _p(out), C.c_uint64(a.shape[0])); return out


def _binop(name, a, b, w):
    a, b = _c(a, w), _c(b, w); out = np.empty_like(a)
    getattr(lib(), name)(_p(a), _p(b), _p(out), C.c_uint64(a.shape[0])); return out


def fr_add(a, b): return _binop("ark_fr_add_batch", a, b, FR_W)
def fr_sub(a, b): return _binop("ark_fr_sub_batch", a, b, FR_W)
def fq_add(a, b): return _binop("ark_fq_add_batch", a, b, FQ_W)
def fq_sub(a, b): return _binop("ark_fq_sub_batch", a, b, FQ_W)
def fq2_mul(a, b): return _binop("ark_fq2_mul_batch", a, b, 2 * FQ_W)


def fr_sqr(a):
    a = _c(a, FR_W); out = np.empty_like(a)
    lib().ark_fr_sqr_batch(_p(a), _p(out), C.c_uint64(a.shape[0])); return out


def fq_sqr(a):
    a = _c(a, FQ_W); out = np.empty_like(a)
    lib().ark_fq_sqr_batch(_p(a), _p(out), C.c_uint64(a.shape[0])); return out


def fr_poly_eval(coeffs_mont, points_mont):
    """sum_i coeffs[i] * x^i at each x of points (Horner, one host thread per point); Montgomery in and out."""
    coeffs_mont, points_mont = _c(coeffs_mont, FR_W), _c(points_mont, FR_W)
    out = np.empty_like(points_mont)
    lib().ark_fr_poly_eval(_p(coeffs_mont), C.c_uint64(coeffs_mont.shape[0]), _p(points_mont), C.c_uint64(points_mont.shape[0]), _p(out))
    return out


def ntt_fr(data, inverse=False, coset=False):
    data = _c(data, FR_W).copy()
    n = data.shape[0]
    log_n = n.bit_length() - 1
    assert 1 << log_n == n
    lib().ark_ntt_fr(_p(data), C.c_int(log_n), C.c_int(int(inverse)), C.c_int(int(coset)))
    return data


def msm_g1(bases, scalars_canon):
    bases, scalars_canon = _c(bases, G1_W), _c(scalars_canon, FR_W)
    n = min(bases.shape[0], scalars_canon.shape[0])
    out = np.zeros(G1_W, dtype=np.uint32)
    lib().ark_msm_g1(_p(bases), _p(scalars_canon), C.c_uint64(n), _p(out)); return out


def msm_g2(bases, scalars_canon):
    bases, scalars_canon = _c(bases, G2_W), _c(scalars_canon, FR_W)
    n = min(bases.shape[0], scalars_canon.shape[0])
    out = np.zeros(G2_W, dtype=np.uint32)
    lib().ark_msm_g2(_p(bases), _p(scalars_canon), C.c_uint64(n), _p(out)); return out


def g1_mul(p, k_canon):
    p, k_canon = _c(p).reshape(G1_W), _c(k_canon).reshape(FR_W); out = np.zeros(G1_W, dtype=np.uint32)
    lib().ark_g1_mul(_p(p), _p(k_canon), _p(out)); return out


def g2_mul(p, k_canon):
    p, k_canon = _c(p).reshape(G2_W), _c(k_canon).reshape(FR_W); out = np.zeros(G2_W, dtype=np.uint32)
    lib().ark_g2_mul(_p(p), _p(k_canon), _p(out)); return out


def g1_add(p, q):
    p, q = _c(p).reshape(G1_W), _c(q).reshape(G1_W); out = np.zeros(G1_W, dtype=np.uint32)
    lib().ark_g1_add(_p(p), _p(q), _p(out)); return out


def g2_add(p, q):
    p, q = _c(p).reshape(G2_W), _c(q).reshape(G2_W); out = np.zeros(G2_W, dtype=np.uint32)
    lib().ark_g2_add(_p(p), _p(q), _p(out)); return out


def g1_on_curve(p): return bool(lib().ark_g1_on_curve(_p(_c(p).reshape(G1_W))))
def g2_on_curve(p): return bool(lib().ark_g2_on_curve(_p(_c(p).reshape(G2_W))))


def g1_generator():
    from .py.bls12_381 import G1_GEN
    return g1_from_py([G1_GEN])[0]


def g2_generator():
    from .py.bls12_381 import G2_GEN
    return g2_from_py([G2_GEN])[0]


def g1_fixed_base(g, k_mont):
    g, k_mont = _c(g).reshape(G1_W), _c(k_mont, FR_W); out = np.zeros((k_mont.shape[0], G1_W), dtype=np.uint32)
    lib().ark_g1_fixed_base(_p(g), _p(k_mont), C.c_uint64(k_mont.shape[0]), _p(out)); return out


def g2_fixed_base(g, k_mont):
    g, k_mont = _c(g).reshape(G2_W), _c(k_mont, FR_W); out = np.zeros((k_mont.shape[0], G2_W), dtype=np.uint32)
    lib().ark_g2_fixed_base(_p(g), _p(k_mont), C.c_uint64(k_mont.shape[0]), _p(out)); return out


def g1_progression(g, a_mont, b_mont, n):
    """out[i] = (a + i*b) * g -- distinct subgroup points with known discrete logs."""
    g = _c(g).reshape(G1_W); out = np.zeros((n, G1_W), dtype=np.uint32)
    lib().ark_g1_progression(_p(g), _p(_c(a_mont).reshape(FR_W)), _p(_c(b_mont).reshape(FR_W)), C.c_uint64(n), _p(out)); return out


def g2_progression(g, a_mont, b_mont, n):
    g = _c(g).reshape(G2_W); out = np.zeros((n, G2_W), dtype=np.uint32)
    lib().ark_g2_progression(_p(g), _p(_c(a_mont).reshape(FR_W)), _p(_c(b_mont).reshape(FR_W)), C.c_uint64(n), _p(out)); return out


def lagrange_coeffs(log_n, tau_mont):
    out = np.zeros((1 << log_n, FR_W), dtype=np.uint32)
    lib().ark_lagrange_coeffs(C.c_int(log_n), _p(_c(tau_mont).reshape(FR_W)), _p(out)); return out


class PkView(C.Structure):
    """Mirror of ctg_pk_view (include/ctg_groth16.h)."""
    _fields_ = [("alpha_g1", C.c_void_p), ("beta_g1", C.c_void_p), ("delta_g1", C.c_void_p),
                ("beta_g2", C.c_void_p), ("delta_g2", C.c_void_p),
                ("a_query", C.c_void_p), ("a_len", C.c_uint64),
                ("b_g1_query", C.c_void_p), ("b_g1_len", C.c_uint64),
                ("b_g2_query", C.c_void_p), ("b_g2_len", C.c_uint64),
                ("h_query", C.c_void_p), ("h_len", C.c_uint64),
                ("l_query", C.c_void_p), ("l_len", C.c_uint64)]


def pk_view(pk: dict):
    """pk: dict of numpy arrays (see ctg.layout.ProvingKeyArrays); returns (view, keepalive)."""
    v = PkView()
    keep = {}
    for k in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2"):
        keep[k] = _c(pk[k]); setattr(v, k, keep[k].ctypes.data)
    for k in ("a_query", "b_g1_query", "b_g2_query", "h_query", "l_query"):
        keep[k] = _c(pk[k]); setattr(v, k, keep[k].ctypes.data)
        setattr(v, k.replace("_query", "_len"), keep[k].shape[0])
    return v, keep


def _csr_args(m):
    rp = np.ascontiguousarray(m["row_ptr"], dtype=np.uint64)
    col = np.ascontiguousarray(m["col"], dtype=np.uint32)
    val = _c(m["val"]).reshape(-1, FR_W)
    return (rp, col, val)


def witness_map(r1cs: dict, z_mont):
    """r1cs: {num_instance, num_witness, num_constraints, a/b/c: {row_ptr u64, col u32, val Fr mont}}"""
    ni, nc = r1cs["num_instance"], r1cs["num_constraints"]
    n = 1
    while n < ni + nc:
        n *= 2
    h = np.zeros((n, FR_W), dtype=np.uint32)
    z = _c(z_mont, FR_W)
    args = []
    keep = []
    for k in "abc":
        t = _csr_args(r1cs[k]); keep.append(t); args += [_p(x) for x in t]
    lib().ark_witness_map(C.c_uint64(ni), C.c_uint64(nc), *args, _p(z), _p(h))
    return h


def create_proof(pk: dict, r1cs: dict, z_mont, r_mont, s_mont):
    """ark_groth16::create_proof restated; returns (A (24,), B (48,), C (24,)) Montgomery affine."""
    v, keep = pk_view(pk)
    z = _c(z_mont, FR_W)
    args = []
    keep2 = []
    for k in "abc":
        t = _csr_args(r1cs[k]); keep2.append(t); args += [_p(x) for x in t]
    a = np.zeros(G1_W, dtype=np.uint32); b = np.zeros(G2_W, dtype=np.uint32); c = np.zeros(G1_W, dtype=np.uint32)
    lib().ark_create_proof(C.byref(v), C.c_uint64(r1cs["num_instance"]), C.c_uint64(r1cs["num_witness"]),
                           C.c_uint64(r1cs["num_constraints"]), *args, _p(z),
                           _p(_c(r_mont).reshape(FR_W)), _p(_c(s_mont).reshape(FR_W)), _p(a), _p(b), _p(c))
    return a, b, c


# ------------------------------------------------------------ bridges from oracle/py objects
def r1cs_from_py(cs):
    """oracle.py.groth16.R1CS -> array dict (CSR with Montgomery coefficients)."""
    out = {"num_instance": cs.num_instance, "num_witness": cs.num_witness, "num_constraints": cs.num_constraints}
    for name, rows in (("a", cs.a), ("b", cs.b), ("c", cs.c)):
        rp, col, val = [0], [], []
        for row in rows:
            for cf, j in row:
                col.append(j); val.append(cf)
            rp.append(len(col))
        out[name] = {"row_ptr": np.array(rp, dtype=np.uint64), "col": np.array(col, dtype=np.uint32),
                     "val": fr_mont_from_ints(val) if val else np.zeros((0, FR_W), dtype=np.uint32)}
    return out


def pk_from_py(pk):
    """oracle.py.groth16.ProvingKey -> array dict in the wire layout."""
    d = {}
    for k in ("alpha_g1", "beta_g1", "delta_g1"):
        d[k] = g1_from_py([getattr(pk, k)])[0]
    for k in ("beta_g2", "delta_g2", "gamma_g2"):
        d[k] = g2_from_py([getattr(pk, k)])[0]
    for k in ("a_query", "b_g1_query", "h_query", "l_query", "gamma_abc_g1"):
        d[k] = g1_from_py(getattr(pk, k))
    d["b_g2_query"] = g2_from_py(pk.b_g2_query)
    return d


# ------------------------------------------------------------ Poseidon / Merkle tree (oracle/py/poseidon.py holds the conventions)
def poseidon_hash_batch(full_rounds, partial_rounds, alpha, mds, rc, inputs):
    """inputs (n, k, 8) Montgomery -> (n, 8): CRH of each row (fresh sponge, absorb k, squeeze 1)."""
    mds, rc, inputs = _c(mds, FR_W), _c(rc, FR_W), _c(inputs, FR_W)
    n, k = inputs.shape[0], inputs.shape[1]
    out = np.empty((n, FR_W), dtype=np.uint32)
    lib().ark_poseidon_hash_batch(int(full_rounds), int(partial_rounds), C.c_uint64(alpha), _p(mds), _p(rc), _p(inputs), C.c_uint64(n), C.c_uint64(k), _p(out))
    return out


def poseidon_merkle_tree(full_rounds, partial_rounds, alpha, mds, rc, leaves):
    """leaves (n, 8) -> (leaf_nodes (n, 8), non_leaf_nodes (n - 1, 8) in heap order)."""
    mds, rc, leaves = _c(mds, FR_W), _c(rc, FR_W), _c(leaves, FR_W)
    n = leaves.shape[0]
    assert n >= 2 and n & (n - 1) == 0
    leaf_nodes = np.empty((n, FR_W), dtype=np.uint32)
    inner = np.empty((n - 1, FR_W), dtype=np.uint32)
    lib().ark_poseidon_merkle_tree(int(full_rounds), int(partial_rounds), C.c_uint64(alpha), _p(mds), _p(rc), _p(leaves), C.c_uint64(n), _p(leaf_nodes), _p(inner))
    return leaf_nodes, inner
