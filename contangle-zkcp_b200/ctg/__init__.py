"""ctypes binding of libctg_groth16.so (include/ctg_groth16.h) -- the harness tests/ and
bench.py drive the C ABI through; a thin mirror of what the Rust `ctg-groth16-sys` crate
binds.  There is no CPU path here: importing works anywhere (so the CPU
test-suite can check that the library loads and exports its symbols), but every compute
call needs a CUDA device and raises CtgError otherwise.

Arrays are numpy uint32 in the wire layout of the header: Fr (n, 8), Fq (n, 12),
G1 affine (n, 24), G2 affine (n, 48), Montgomery limbs, all-zero = point at infinity.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CTG_LIB") or os.path.join(os.path.dirname(_HERE), "libctg_groth16.so")   # CTG_LIB: alternative builds in kernel sweeps

FR_W, FQ_W, G1_W, G2_W = 8, 12, 24, 48
PARTIAL_W = 4 * 2 * G1_W + 2 * G2_W   # ctg_proof_partial: 4 G1 XYZZ + 1 G2 XYZZ = 288 words (1152 B)
SCALARS_CANONICAL, SCALARS_MONT, MSM_TIMED = 0, 1, 0x100
FIELD_FR, FIELD_FQ, FIELD_FQ2 = 0, 1, 2
OP_MUL, OP_ADD, OP_SUB, OP_INV = 0, 1, 2, 3
OP_PAIRED_MUL, OP_PAIRED_SQR, OP_PAIRED_MUL2SUB, OP_PAIRED_IS_ZERO = 4, 5, 6, 7   # FIELD_FQ2 only: the lane-paired arithmetic of csrc/msm_pair.cuh

# every symbol include/ctg_groth16.h declares (checked by tests/test_abi_cpu.py)
EXPORTS = [
    "ctg_ctx_create", "ctg_ctx_device_count", "ctg_ctx_destroy", "ctg_last_error", "ctg_ctx_set_stream", "ctg_ctx_set_ba_levels", "ctg_ctx_set_msm_split", "ctg_sync", "ctg_launch_count",
    "ctg_last_timings", "ctg_dev_alloc", "ctg_dev_free", "ctg_h2d", "ctg_d2h", "ctg_field_op_batch", "ctg_field_op_batch_noalloc",
    "ctg_fr_mul_batch", "ctg_fq_mul_batch", "ctg_imad_peak_probe", "ctg_ntt_fr", "ctg_ntt_fr_dev",
    "ctg_bases_upload_g1", "ctg_bases_upload_g2", "ctg_bases_progression_g1", "ctg_bases_progression_g2",
    "ctg_bases_download", "ctg_bases_free", "ctg_bases_len", "ctg_bases_plan", "ctg_msm_g1", "ctg_msm_g2", "ctg_msm_g1_dev",
    "ctg_msm_g2_dev", "ctg_msm_g1_partial", "ctg_msm_g2_partial", "ctg_g1_combine", "ctg_g2_combine",
    "ctg_g1_combine_dev", "ctg_g2_combine_dev", "ctg_pk_upload", "ctg_pk_free", "ctg_r1cs_upload",
    "ctg_r1cs_free", "ctg_witness_map", "ctg_prove", "ctg_prove_batch", "ctg_pk_upload_shard", "ctg_prove_partial",
    "ctg_prove_combine", "ctg_prepare_inputs", "ctg_host_alloc", "ctg_host_free",
    "ctg_generate_parameters", "ctg_params_pk_view", "ctg_params_download", "ctg_params_free",
    "ctg_params_deserialize_unchecked", "ctg_params_serialize_unchecked", "ctg_params_serialize_vk",
    "ctg_proof_serialize", "ctg_proof_deserialize", "ctg_prove_partial_begin", "ctg_prove_partial_finish",
    "ctg_pk_upload_range",
    "ctg_poseidon_new", "ctg_poseidon_free", "ctg_poseidon_hash_batch", "ctg_poseidon_hash_batch_dev",
    "ctg_poseidon_merkle_tree", "ctg_poseidon_merkle_tree_dev", "ctg_te_mul_batch", "ctg_te_mul_batch_dev",
]


class CtgError(RuntimeError):
    pass


class PkView(C.Structure):
    _fields_ = [("alpha_g1", C.c_void_p), ("beta_g1", C.c_void_p), ("delta_g1", C.c_void_p),
                ("beta_g2", C.c_void_p), ("delta_g2", C.c_void_p),
                ("a_query", C.c_void_p), ("a_len", C.c_uint64),
                ("b_g1_query", C.c_void_p), ("b_g1_len", C.c_uint64),
                ("b_g2_query", C.c_void_p), ("b_g2_len", C.c_uint64),
                ("h_query", C.c_void_p), ("h_len", C.c_uint64),
                ("l_query", C.c_void_p), ("l_len", C.c_uint64)]


class Trapdoor(C.Structure):
    _fields_ = [(k, C.c_uint32 * FR_W) for k in ("alpha", "beta", "gamma", "delta", "tau")] + [("g1", C.c_uint32 * G1_W), ("g2", C.c_uint32 * G2_W)]


PARAMS_OUT_FIELDS = ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "gamma_g2", "delta_g2", "gamma_abc_g1",
                     "a_query", "b_g1_query", "b_g2_query", "h_query", "l_query")


class ParamsOut(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in PARAMS_OUT_FIELDS]


class Csr(C.Structure):
    _fields_ = [("row_ptr", C.c_void_p), ("col", C.c_void_p), ("val", C.c_void_p)]


class MsmPlan(C.Structure):
    _fields_ = [("window_bits", C.c_int32), ("windows", C.c_int32), ("copies", C.c_int32), ("bucket_sets", C.c_int32),
                ("buckets_per_set", C.c_uint32)]


class Timings(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("h2d_ms", "witness_map_ms", "msm_h_ms", "msm_l_ms", "msm_a_ms", "msm_b_g1_ms",
                                         "msm_b_g2_ms", "finalize_ms", "total_ms", "msm_sort_ms", "msm_accumulate_ms",
                                         "msm_reduce_ms")]

    def as_dict(self):
        return {k: float(getattr(self, k)) for k, _ in self._fields_}


_lib = None


def load():
    """Load the CUDA extension; fail loudly if it was not built (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CtgError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(make -C contangle-zkcp_b200/csrc). There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        lib.ctg_last_error.restype = C.c_char_p
        lib.ctg_launch_count.restype = C.c_uint64
        lib.ctg_bases_len.restype = C.c_uint64
        for name in ("ctg_ctx_destroy", "ctg_bases_free", "ctg_pk_free", "ctg_r1cs_free", "ctg_params_free", "ctg_poseidon_free"):
            getattr(lib, name).restype = None
        _lib = lib
    return _lib


def _arr(a, words=None):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    if words is not None and (a.ndim == 0 or a.shape[-1] != words):
        raise CtgError(f"expected trailing dimension {words}, got shape {a.shape}")
    return a


def _ptr(a):
    return C.c_void_p(a.ctypes.data)


class Bases:
    def __init__(self, ctx, handle, group):
        self.ctx, self.handle, self.group = ctx, handle, group

    def __len__(self):
        return int(load().ctg_bases_len(self.handle))

    def plan(self):
        p = MsmPlan()
        self.ctx._check(load().ctg_bases_plan(self.handle, C.byref(p)))
        return {k: int(getattr(p, k)) for k, _ in MsmPlan._fields_}

    def download(self, first=0, count=None):
        count = len(self) - first if count is None else count
        out = np.zeros((count, G1_W if self.group == 1 else G2_W), dtype=np.uint32)
        self.ctx._check(load().ctg_bases_download(self.ctx.h, self.handle, C.c_uint64(first), C.c_uint64(count), _ptr(out)))
        return out

    def free(self):
        if self.handle:
            load().ctg_bases_free(self.ctx.h, self.handle)
            self.handle = None


class Context:
    """One CUDA device (ctg_ctx), or -- device = a list of ids -- a multi-device context: one process driving several GPUs, the
    prover-path calls (upload_pk, upload_r1cs, prove, bases_*, msm*) shard over them by themselves, everything else runs on
    the first device.  `stream`: a raw cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream) for the (first) device."""

    def __init__(self, device=0, stream=None):
        lib = load()
        self.h = C.c_void_p()
        devs = [int(d) for d in device] if isinstance(device, (list, tuple)) else [int(device)]
        self.devices = devs
        ids = (C.c_int * len(devs))(*devs)
        rc = lib.ctg_ctx_create(ids, len(devs), C.byref(self.h))
        if rc != 0:
            raise CtgError(f"ctg_ctx_create failed with {rc}: no usable sm_100 CUDA device (there is no CPU fallback)")
        if stream is not None:
            self._check(lib.ctg_ctx_set_stream(self.h, C.c_void_p(stream)))

    def _check(self, rc):
        if rc != 0:
            raise CtgError(f"ctg error {rc}: {load().ctg_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            load().ctg_ctx_destroy(self.h)
            self.h = None

    def sync(self):
        self._check(load().ctg_sync(self.h))

    @property
    def launches(self):
        return int(load().ctg_launch_count(self.h))

    def last_timings(self):
        t = Timings()
        self._check(load().ctg_last_timings(self.h, C.byref(t)))
        return t.as_dict()

    # ---- fields
    def set_ba_levels(self, levels=-1):
        """MSM bucket accumulation: number of batched-affine tree levels (-1 = automatic, 0 = XYZZ only)."""
        self._check(load().ctg_ctx_set_ba_levels(self.h, int(levels)))

    def set_msm_split(self, parts=-1):
        """Point-range parts per stand-alone MSM (-1 = automatic, 1 = never split)."""
        self._check(load().ctg_ctx_set_msm_split(self.h, int(parts)))

    def field_op(self, field, op, a, b):
        w = {FIELD_FR: FR_W, FIELD_FQ: FQ_W, FIELD_FQ2: 2 * FQ_W}[field]
        a, b = _arr(a, w), _arr(b, w)
        out = np.empty_like(a)
        self._check(load().ctg_field_op_batch(self.h, field, op, _ptr(a), _ptr(b), _ptr(out), C.c_size_t(a.shape[0])))
        return out

    def field_op_noalloc(self, field, op, a, b):
        """ctg_field_op_batch_noalloc: the same element-wise kernel staged in the library's fixed 3 MiB device array (no cudaMalloc);
        raises CtgError beyond 32 768 Fr / 21 845 Fq / 10 922 Fq2 elements.  Calls that overlap on one GPU race on that array."""
        w = {FIELD_FR: FR_W, FIELD_FQ: FQ_W, FIELD_FQ2: 2 * FQ_W}[field]
        a, b = _arr(a, w), _arr(b, w)
        out = np.empty_like(a)
        self._check(load().ctg_field_op_batch_noalloc(self.h, field, op, _ptr(a), _ptr(b), _ptr(out), C.c_size_t(a.shape[0])))
        return out

    def fr_mul(self, a, b):
        a, b = _arr(a, FR_W), _arr(b, FR_W)
        out = np.empty_like(a)
        self._check(load().ctg_fr_mul_batch(self.h, _ptr(a), _ptr(b), _ptr(out), C.c_size_t(a.shape[0])))
        return out

    def fq_mul(self, a, b):
        a, b = _arr(a, FQ_W), _arr(b, FQ_W)
        out = np.empty_like(a)
        self._check(load().ctg_fq_mul_batch(self.h, _ptr(a), _ptr(b), _ptr(out), C.c_size_t(a.shape[0])))
        return out

    def imad_peak_probe(self):
        imad, fq = C.c_double(), C.c_double()
        self._check(load().ctg_imad_peak_probe(self.h, C.byref(imad), C.byref(fq)))
        return imad.value, fq.value

    # ---- NTT
    def ntt(self, data, inverse=False, coset=False):
        data = _arr(data, FR_W).copy()
        n = data.shape[0]
        log_n = n.bit_length() - 1
        if n == 0 or (1 << log_n) != n:
            raise CtgError("ntt: length must be a power of two")
        self._check(load().ctg_ntt_fr(self.h, _ptr(data), C.c_uint32(log_n), int(inverse), int(coset)))
        return data

    def ntt_dev(self, d_ptr, log_n, inverse=False, coset=False):
        self._check(load().ctg_ntt_fr_dev(self.h, C.c_void_p(d_ptr), C.c_uint32(log_n), int(inverse), int(coset)))

    # ---- bulk Poseidon (witness-side crypto upstream of the prover)
    def poseidon(self, full_rounds, partial_rounds, alpha, mds, ark):
        """ark_sponge::poseidon::PoseidonParameters::new(full_rounds, partial_rounds, alpha, mds, ark) on the device;
        mds (3, 3, 8) and ark (rounds, 3, 8): Montgomery Fr words."""
        mds = np.ascontiguousarray(mds, dtype=np.uint32)
        ark = np.ascontiguousarray(ark, dtype=np.uint32)
        if mds.shape != (3, 3, FR_W) or ark.shape != (int(full_rounds) + int(partial_rounds), 3, FR_W):
            raise CtgError("poseidon: mds must be (3, 3, 8) and ark (full + partial, 3, 8) Fr words")
        h = C.c_void_p()
        self._check(load().ctg_poseidon_new(self.h, C.c_uint32(full_rounds), C.c_uint32(partial_rounds), C.c_uint32(alpha),
                                            _ptr(mds), _ptr(ark), C.byref(h)))
        return Poseidon(self, h)

    def te_mul_batch(self, d, points, scalars):
        """scalars[i] * points[i] (or * points[0] when one point is given) on -x^2 + y^2 = 1 + d x^2 y^2 over Fr (JubJub):
        d (8,) Montgomery words, points (1 | n, 2, 8) affine Montgomery, scalars (n, 8) 256-bit integers -> (n, 2, 8)."""
        d = np.ascontiguousarray(d, dtype=np.uint32).reshape(FR_W)
        points = np.ascontiguousarray(points, dtype=np.uint32)
        scalars = _arr(scalars, FR_W)
        if points.ndim != 3 or points.shape[1:] != (2, FR_W):
            raise CtgError("te_mul: points must be (1 | n, 2, 8) Fr words")
        n = scalars.shape[0]
        out = np.empty((n, 2, FR_W), dtype=np.uint32)
        self._check(load().ctg_te_mul_batch(self.h, _ptr(d), _ptr(points), C.c_uint64(points.shape[0]), _ptr(scalars), C.c_uint64(n), _ptr(out)))
        return out

    def te_mul_batch_dev(self, d, d_points, n_points, d_scalars, n, d_out):
        d = np.ascontiguousarray(d, dtype=np.uint32).reshape(FR_W)
        self._check(load().ctg_te_mul_batch_dev(self.h, _ptr(d), C.c_void_p(d_points), C.c_uint64(n_points), C.c_void_p(d_scalars),
                                                C.c_uint64(n), C.c_void_p(d_out)))

    def elgamal_encrypt_batch(self, poseidon, d, generator, pk, randomness, messages):
        """EncryptCircuit::encrypt (circuits-ark/src/encryption.rs:154-172) for n (randomness, message) pairs under one public
        key: c1[i] = G * r[i]; pad[i] = sponge(pk * r[i]); c2[i][j] = pad[i] + messages[i][j].  generator, pk: (2, 8) affine;
        randomness (n, 8) integers; messages (n, chunks, 8) Montgomery -> (c1 (n, 2, 8), c2 (n, chunks, 8))."""
        messages = np.ascontiguousarray(messages, dtype=np.uint32)
        n, chunks = messages.shape[0], messages.shape[1]
        c1 = self.te_mul_batch(d, np.asarray(generator, dtype=np.uint32).reshape(1, 2, FR_W), randomness)
        shared = self.te_mul_batch(d, np.asarray(pk, dtype=np.uint32).reshape(1, 2, FR_W), randomness)
        pad = poseidon.hash_batch(shared)
        c2 = self.field_op(FIELD_FR, OP_ADD, messages.reshape(n * chunks, FR_W), np.repeat(pad, chunks, axis=0))
        return c1, c2.reshape(n, chunks, FR_W)

    # ---- bases / MSM
    def bases_g1(self, points, copies=0, c_bits=0):
        points = _arr(points, G1_W)
        h = C.c_void_p()
        self._check(load().ctg_bases_upload_g1(self.h, _ptr(points), C.c_uint64(points.shape[0]), copies, c_bits, C.byref(h)))
        return Bases(self, h, 1)

    def bases_g2(self, points, copies=0, c_bits=0):
        points = _arr(points, G2_W)
        h = C.c_void_p()
        self._check(load().ctg_bases_upload_g2(self.h, _ptr(points), C.c_uint64(points.shape[0]), copies, c_bits, C.byref(h)))
        return Bases(self, h, 2)

    def bases_progression_g1(self, g, a_mont, b_mont, n, copies=0, c_bits=0):
        g, a_mont, b_mont = _arr(g).reshape(G1_W), _arr(a_mont).reshape(FR_W), _arr(b_mont).reshape(FR_W)
        h = C.c_void_p()
        self._check(load().ctg_bases_progression_g1(self.h, _ptr(g), _ptr(a_mont), _ptr(b_mont), C.c_uint64(n), copies, c_bits, C.byref(h)))
        return Bases(self, h, 1)

    def bases_progression_g2(self, g, a_mont, b_mont, n, copies=0, c_bits=0):
        g, a_mont, b_mont = _arr(g).reshape(G2_W), _arr(a_mont).reshape(FR_W), _arr(b_mont).reshape(FR_W)
        h = C.c_void_p()
        self._check(load().ctg_bases_progression_g2(self.h, _ptr(g), _ptr(a_mont), _ptr(b_mont), C.c_uint64(n), copies, c_bits, C.byref(h)))
        return Bases(self, h, 2)

    def msm(self, bases: Bases, scalars, mont=False, n=None, timed=False):
        """sum_i scalars[i] * bases[i] -> affine point (24 or 48 words).  Scalars are host memory."""
        scalars = _arr(scalars, FR_W)
        n = scalars.shape[0] if n is None else n
        flags = (SCALARS_MONT if mont else 0) | (MSM_TIMED if timed else 0)
        out = np.zeros(G1_W if bases.group == 1 else G2_W, dtype=np.uint32)
        fn = load().ctg_msm_g1 if bases.group == 1 else load().ctg_msm_g2
        self._check(fn(self.h, bases.handle, _ptr(scalars), C.c_uint64(n), C.c_uint32(flags), _ptr(out)))
        return out

    def msm_dev(self, bases: Bases, d_scalars, n, d_out_xyzz, mont=False, timed=False):
        flags = (SCALARS_MONT if mont else 0) | (MSM_TIMED if timed else 0)
        fn = load().ctg_msm_g1_dev if bases.group == 1 else load().ctg_msm_g2_dev
        self._check(fn(self.h, bases.handle, C.c_void_p(d_scalars), C.c_uint64(n), C.c_uint32(flags), C.c_void_p(d_out_xyzz)))

    def pinned_like(self, arr):
        """Copy of `arr` (uint32 words) in page-locked host memory (ctg_host_alloc); keep the returned array alive while in use.
        Free with ctx.pinned_free(a)."""
        arr = np.ascontiguousarray(arr, dtype=np.uint32)
        p = C.c_void_p()
        self._check(load().ctg_host_alloc(self.h, C.c_size_t(arr.nbytes), C.byref(p)))
        buf = (C.c_uint32 * arr.size).from_address(p.value)
        out = np.frombuffer(buf, dtype=np.uint32).reshape(arr.shape)
        out[...] = arr
        self._pinned = getattr(self, "_pinned", {})
        self._pinned[out.ctypes.data] = p.value
        return out

    def pinned_free(self, a):
        p = getattr(self, "_pinned", {}).pop(a.ctypes.data, None)
        if p is not None:
            self._check(load().ctg_host_free(self.h, C.c_void_p(p)))

    def prepare_inputs(self, gamma_abc: Bases, public_inputs_mont):
        """ark_groth16::prepare_inputs: gamma_abc_g1[0] + sum_i x_i * gamma_abc_g1[i + 1] (affine, (24,) uint32)."""
        x = _arr(public_inputs_mont, FR_W) if len(public_inputs_mont) else np.zeros((0, FR_W), dtype=np.uint32)
        out = np.zeros(G1_W, dtype=np.uint32)
        self._check(load().ctg_prepare_inputs(self.h, gamma_abc.handle, _ptr(x), C.c_uint64(x.shape[0]), _ptr(out)))
        return out

    def msm_partial(self, bases: Bases, scalars, mont=False):
        scalars = _arr(scalars, FR_W)
        out = np.zeros(2 * (G1_W if bases.group == 1 else G2_W), dtype=np.uint32)
        fn = load().ctg_msm_g1_partial if bases.group == 1 else load().ctg_msm_g2_partial
        self._check(fn(self.h, bases.handle, _ptr(scalars), C.c_uint64(scalars.shape[0]), C.c_uint32(SCALARS_MONT if mont else 0), _ptr(out)))
        return out

    def combine(self, partials, group=1):
        w = 2 * (G1_W if group == 1 else G2_W)
        partials = _arr(partials, w)
        out = np.zeros(G1_W if group == 1 else G2_W, dtype=np.uint32)
        fn = load().ctg_g1_combine if group == 1 else load().ctg_g2_combine
        self._check(fn(self.h, _ptr(partials), C.c_uint32(partials.shape[0]), _ptr(out)))
        return out

    def combine_dev(self, d_partials, k, d_out_affine, group=1):
        fn = load().ctg_g1_combine_dev if group == 1 else load().ctg_g2_combine_dev
        self._check(fn(self.h, C.c_void_p(d_partials), C.c_uint32(k), C.c_void_p(d_out_affine)))

    # ---- proving key / R1CS / prover
    def upload_pk(self, pk: dict, num_instance, num_constraints, copies=0, c_bits=0, shard=None):
        """pk: dict of arrays alpha_g1, beta_g1, delta_g1, beta_g2, delta_g2, a_query, b_g1_query, b_g2_query, h_query, l_query.
        shard=(index, count): keep only that point range of every query (multi-GPU; use with prove_partial)."""
        v = PkView()
        keep = []
        for k, w in (("alpha_g1", G1_W), ("beta_g1", G1_W), ("delta_g1", G1_W), ("beta_g2", G2_W), ("delta_g2", G2_W)):
            a = _arr(pk[k]).reshape(w); keep.append(a); setattr(v, k, a.ctypes.data)
        for k, w in (("a_query", G1_W), ("b_g1_query", G1_W), ("b_g2_query", G2_W), ("h_query", G1_W), ("l_query", G1_W)):
            a = _arr(pk[k]).reshape(-1, w); keep.append(a); setattr(v, k, a.ctypes.data)
            setattr(v, k.replace("_query", "_len"), a.shape[0])
        return DevicePk(self, self._pk_upload(v, num_instance, num_constraints, copies, c_bits, shard), pk["a_query"].shape[0])

    def _pk_upload(self, v, num_instance, num_constraints, copies, c_bits, shard):
        """shard: None (whole key), (index, count) equal point-range shards, or (lo_num, hi_num, den) a weighted range."""
        h = C.c_void_p()
        if shard is None:
            self._check(load().ctg_pk_upload(self.h, C.byref(v), C.c_uint64(num_instance), C.c_uint64(num_constraints), copies, c_bits, C.byref(h)))
        elif len(shard) == 2:
            self._check(load().ctg_pk_upload_shard(self.h, C.byref(v), C.c_uint64(num_instance), C.c_uint64(num_constraints), copies, c_bits,
                                                   int(shard[0]), int(shard[1]), C.byref(h)))
        else:
            self._check(load().ctg_pk_upload_range(self.h, C.byref(v), C.c_uint64(num_instance), C.c_uint64(num_constraints), copies, c_bits,
                                                   C.c_uint64(int(shard[0])), C.c_uint64(int(shard[1])), C.c_uint64(int(shard[2])), C.byref(h)))
        return h

    @staticmethod
    def _csr_triplet(r1cs):
        mats, keep = [], []
        for k in "abc":
            m = r1cs[k]
            rp = np.ascontiguousarray(m["row_ptr"], dtype=np.uint64)
            col = np.ascontiguousarray(m["col"], dtype=np.uint32)
            val = _arr(m["val"]).reshape(-1, FR_W)
            keep += [rp, col, val]
            mats.append(Csr(rp.ctypes.data, col.ctypes.data, val.ctypes.data))
        return mats, keep

    def generate_parameters(self, r1cs: dict, trapdoor: dict):
        """ark_groth16::generate_parameters on the device.  trapdoor: Montgomery Fr word arrays alpha, beta, gamma, delta, tau
        and affine generators g1 (24 words), g2 (48 words).  Returns DeviceParams (key resident in HBM)."""
        mats, keep = self._csr_triplet(r1cs)
        td = Trapdoor()
        for k, w in (("alpha", FR_W), ("beta", FR_W), ("gamma", FR_W), ("delta", FR_W), ("tau", FR_W), ("g1", G1_W), ("g2", G2_W)):
            a = _arr(trapdoor[k]).reshape(w)
            getattr(td, k)[:] = [int(x) for x in a]
        h = C.c_void_p()
        self._check(load().ctg_generate_parameters(self.h, C.c_uint64(r1cs["num_instance"]), C.c_uint64(r1cs["num_witness"]),
                                                   C.c_uint64(r1cs["num_constraints"]), C.byref(mats[0]), C.byref(mats[1]), C.byref(mats[2]),
                                                   C.byref(td), C.byref(h)))
        n = 1
        while n < r1cs["num_instance"] + r1cs["num_constraints"]:
            n *= 2
        return DeviceParams(self, h, r1cs["num_instance"], r1cs["num_witness"], r1cs["num_constraints"], n)

    def load_proving_key(self, key_image: bytes):
        """ProvingKey::deserialize_unchecked of a circuit.zkey image (lib/src/utils.rs:104-110), decoded on the device."""
        buf = np.frombuffer(key_image, dtype=np.uint8)
        h = C.c_void_p()
        self._check(load().ctg_params_deserialize_unchecked(self.h, C.c_void_p(buf.ctypes.data), C.c_uint64(buf.size), C.byref(h)))
        v, gl = PkView(), C.c_uint64()
        self._check(load().ctg_params_pk_view(self.h, h, C.byref(v), None, C.byref(gl)))
        ni, m, n = int(gl.value), int(v.a_len), int(v.h_len) + 1
        return DeviceParams(self, h, ni, m - ni, n - ni, n)

    def upload_r1cs(self, r1cs: dict):
        mats, keep = self._csr_triplet(r1cs)
        h = C.c_void_p()
        self._check(load().ctg_r1cs_upload(self.h, C.c_uint64(r1cs["num_instance"]), C.c_uint64(r1cs["num_witness"]),
                                           C.c_uint64(r1cs["num_constraints"]), C.byref(mats[0]), C.byref(mats[1]), C.byref(mats[2]), C.byref(h)))
        n = 1
        while n < r1cs["num_instance"] + r1cs["num_constraints"]:
            n *= 2
        return DeviceR1cs(self, h, r1cs["num_instance"] + r1cs["num_witness"], n)

    def witness_map(self, r1cs: "DeviceR1cs", z):
        z = _arr(z, FR_W)
        if z.shape[0] != r1cs.m:
            raise CtgError("witness_map: assignment length differs from num_instance + num_witness")
        h = np.zeros((r1cs.n, FR_W), dtype=np.uint32)
        self._check(load().ctg_witness_map(self.h, r1cs.handle, _ptr(z), _ptr(h)))
        return h

    def prove(self, pk: "DevicePk", r1cs: "DeviceR1cs", z, r, s, timed=False):
        """ark_groth16::create_proof after synthesis.  Returns (A (24,), B (48,), C (24,))[, timings]."""
        z = _arr(z, FR_W)
        if z.shape[0] != r1cs.m or pk.m != r1cs.m:
            raise CtgError("prove: assignment / key / R1CS sizes differ")
        r, s = _arr(r).reshape(FR_W), _arr(s).reshape(FR_W)
        out = np.zeros(G1_W + G2_W + G1_W, dtype=np.uint32)
        t = Timings()
        self._check(load().ctg_prove(self.h, pk.handle, r1cs.handle, _ptr(z), _ptr(r), _ptr(s), _ptr(out), C.byref(t) if timed else None))
        proof = (out[:G1_W].copy(), out[G1_W:G1_W + G2_W].copy(), out[G1_W + G2_W:].copy())
        return (proof, t.as_dict()) if timed else proof

    def prove_batch(self, pk: "DevicePk", r1cs: "DeviceR1cs", z_batch, rs_batch):
        """n independent proofs with one key: z_batch (n, m, 8), rs_batch (n, 2, 8).  Returns (n, 96) words: A | B | C."""
        z_batch = _arr(z_batch, FR_W)
        rs_batch = _arr(rs_batch, FR_W)
        n = z_batch.shape[0]
        if z_batch.ndim != 3 or z_batch.shape[1] != r1cs.m or rs_batch.shape != (n, 2, FR_W) or pk.m != r1cs.m:
            raise CtgError("prove_batch: shapes must be z (n, m, 8), rs (n, 2, 8)")
        out = np.zeros((n, G1_W + G2_W + G1_W), dtype=np.uint32)
        self._check(load().ctg_prove_batch(self.h, pk.handle, r1cs.handle, C.c_uint32(n), _ptr(z_batch), _ptr(rs_batch), _ptr(out)))
        return out

    def prove_partial(self, pk_shard: "DevicePk", r1cs: "DeviceR1cs", z, r, s, timed=False):
        """This GPU's share of a sharded proof: (288,) words of partial sums (ctg_proof_partial)."""
        z = _arr(z, FR_W)
        if z.shape[0] != r1cs.m or pk_shard.m != r1cs.m:
            raise CtgError("prove_partial: assignment / key / R1CS sizes differ")
        r, s = _arr(r).reshape(FR_W), _arr(s).reshape(FR_W)
        out = np.zeros(PARTIAL_W, dtype=np.uint32)
        t = Timings()
        self._check(load().ctg_prove_partial(self.h, pk_shard.handle, r1cs.handle, _ptr(z), _ptr(r), _ptr(s), _ptr(out), C.byref(t) if timed else None))
        return (out, t.as_dict()) if timed else out

    def prove_partial_begin(self, pk_shard: "DevicePk", r1cs: "DeviceR1cs", z, r, s, chain_mask, d_abc_ptr):
        """Phase 1 of a sharded proof with a distributed witness map: computes the chains in chain_mask (bit 0 = a, 1 = b,
        2 = c) into the caller's device buffer (3 * n Fr at d_abc_ptr) and starts the assignment MSMs."""
        z = _arr(z, FR_W)
        if z.shape[0] != r1cs.m or pk_shard.m != r1cs.m:
            raise CtgError("prove_partial_begin: assignment / key / R1CS sizes differ")
        r, s = _arr(r).reshape(FR_W), _arr(s).reshape(FR_W)
        self._check(load().ctg_prove_partial_begin(self.h, pk_shard.handle, r1cs.handle, _ptr(z), _ptr(r), _ptr(s), C.c_uint32(chain_mask),
                                                   C.c_void_p(d_abc_ptr)))

    def prove_partial_finish(self, pk_shard: "DevicePk", r1cs: "DeviceR1cs", r, s, d_abc_ptr):
        """Phase 2: pointwise step, coset_ifft and the H MSM; returns the (288,) words of partial sums."""
        r, s = _arr(r).reshape(FR_W), _arr(s).reshape(FR_W)
        out = np.zeros(PARTIAL_W, dtype=np.uint32)
        self._check(load().ctg_prove_partial_finish(self.h, pk_shard.handle, r1cs.handle, _ptr(r), _ptr(s), C.c_void_p(d_abc_ptr), _ptr(out)))
        return out

    def prove_combine(self, partials, r, s):
        """Sum the shards' partials (k, 288) and assemble the proof (A, B, C)."""
        partials = _arr(partials, PARTIAL_W).reshape(-1, PARTIAL_W)
        r, s = _arr(r).reshape(FR_W), _arr(s).reshape(FR_W)
        out = np.zeros(G1_W + G2_W + G1_W, dtype=np.uint32)
        self._check(load().ctg_prove_combine(self.h, _ptr(partials), C.c_uint32(partials.shape[0]), _ptr(r), _ptr(s), _ptr(out)))
        return out[:G1_W].copy(), out[G1_W:G1_W + G2_W].copy(), out[G1_W + G2_W:].copy()

    def prove_ptr(self, pk, r1cs, z_ptr, r, s, out):
        """Same with a raw host pointer for z (e.g. a pinned torch tensor) and a preallocated output (96 words)."""
        self._check(load().ctg_prove(self.h, pk.handle, r1cs.handle, C.c_void_p(z_ptr), _ptr(r), _ptr(s), _ptr(out), None))


class DevicePk:
    def __init__(self, ctx, handle, m):
        self.ctx, self.handle, self.m = ctx, handle, m

    def free(self):
        if self.handle:
            load().ctg_pk_free(self.ctx.h, self.handle)
            self.handle = None


class DeviceR1cs:
    def __init__(self, ctx, handle, m, n):
        self.ctx, self.handle, self.m, self.n = ctx, handle, m, n

    def free(self):
        if self.handle:
            load().ctg_r1cs_free(self.ctx.h, self.handle)
            self.handle = None


class Poseidon:
    """Device-resident PoseidonParameters (rate 2, capacity 1): bulk CRH / sponge hashes and the ciphertext Merkle tree
    (circuits-ark/src/sample_entries.rs:127-134, encryption.rs:154-172)."""

    def __init__(self, ctx, handle):
        self.ctx, self.handle = ctx, handle

    def free(self):
        if self.handle:
            load().ctg_poseidon_free(self.ctx.h, self.handle)
            self.handle = None

    def hash_batch(self, inputs):
        """inputs (n, k, 8) -> (n, 8): first squeezed element of a fresh sponge that absorbed the k elements of each row
        (k = 1: leaf CRH, k = 2: TwoToOneCRH / the ElGamal pad of an affine point)."""
        inputs = np.ascontiguousarray(inputs, dtype=np.uint32)
        if inputs.ndim != 3 or inputs.shape[2] != FR_W:
            raise CtgError("poseidon: inputs must be (n, k, 8) Fr words")
        n, k = inputs.shape[0], inputs.shape[1]
        out = np.empty((n, FR_W), dtype=np.uint32)
        self.ctx._check(load().ctg_poseidon_hash_batch(self.ctx.h, self.handle, _ptr(inputs), C.c_uint64(n), C.c_uint32(k), _ptr(out)))
        return out

    def hash_batch_dev(self, d_in, n, k, d_out):
        self.ctx._check(load().ctg_poseidon_hash_batch_dev(self.ctx.h, self.handle, C.c_void_p(d_in), C.c_uint64(n), C.c_uint32(k), C.c_void_p(d_out)))

    def merkle_tree(self, leaves):
        """MerkleTree::new(params, params, leaves) with one field element per leaf -> MerkleTree (leaf_nodes, non_leaf_nodes)."""
        leaves = _arr(leaves, FR_W)
        n = leaves.shape[0]
        digests = np.empty((n, FR_W), dtype=np.uint32)
        inner = np.empty((max(n - 1, 0), FR_W), dtype=np.uint32)
        self.ctx._check(load().ctg_poseidon_merkle_tree(self.ctx.h, self.handle, _ptr(leaves), C.c_uint64(n), _ptr(digests), _ptr(inner)))
        return MerkleTree(digests, inner)

    def merkle_tree_dev(self, d_leaves, n, d_leaf_digests, d_inner):
        self.ctx._check(load().ctg_poseidon_merkle_tree_dev(self.ctx.h, self.handle, C.c_void_p(d_leaves), C.c_uint64(n),
                                                            C.c_void_p(d_leaf_digests), C.c_void_p(d_inner)))


class MerkleTree:
    """Nodes of ark_crypto_primitives::merkle_tree::MerkleTree in its own order: leaf_nodes (n, 8) and non_leaf_nodes
    (n - 1, 8) in heap order, root first."""

    def __init__(self, leaf_nodes, non_leaf_nodes):
        self.leaf_nodes, self.non_leaf_nodes = leaf_nodes, non_leaf_nodes

    def root(self):
        return self.non_leaf_nodes[0]

    def generate_proof(self, index):
        """MerkleTree::generate_proof: (leaf_sibling_hash, auth_path root side first, leaf_index)."""
        n = self.leaf_nodes.shape[0]
        if not 0 <= index < n:
            raise CtgError("merkle tree: leaf index out of range")
        sibling = self.leaf_nodes[index ^ 1]
        node = (n - 1 + index - 1) // 2
        path = []
        while node != 0:
            path.append(self.non_leaf_nodes[node + 1 if node % 2 == 1 else node - 1])
            node = (node - 1) // 2
        return sibling, path[::-1], index


class DeviceParams:
    """Output of Context.generate_parameters: ark_groth16::ProvingKey (+ VerifyingKey) in HBM."""

    def __init__(self, ctx, handle, ni, nw, nc, n):
        self.ctx, self.handle, self.ni, self.nw, self.nc, self.n = ctx, handle, ni, nw, nc, n

    def _view(self):
        v, gp, gl = PkView(), C.c_void_p(), C.c_uint64()
        self.ctx._check(load().ctg_params_pk_view(self.ctx.h, self.handle, C.byref(v), C.byref(gp), C.byref(gl)))
        return v, gp, int(gl.value)

    def to_device_pk(self, copies=0, c_bits=0, shard=None):
        """Device-to-device: build the prover's key cache (ctg_pk) straight from the generated key."""
        v, _, _ = self._view()
        return DevicePk(self.ctx, self.ctx._pk_upload(v, self.ni, self.nc, copies, c_bits, shard), self.ni + self.nw)

    def gamma_abc_bases(self, copies=1, c_bits=0):
        """vk.gamma_abc_g1 as a base table for Context.prepare_inputs (device-to-device)."""
        _, gp, gl = self._view()
        h = C.c_void_p()
        self.ctx._check(load().ctg_bases_upload_g1(self.ctx.h, gp, C.c_uint64(gl), copies, c_bits, C.byref(h)))
        return Bases(self.ctx, h, 1)

    def download(self, fields=PARAMS_OUT_FIELDS):
        """Host copies (dict of uint32 word arrays) of the requested key members."""
        m, shapes = self.ni + self.nw, {}
        for k in ("alpha_g1", "beta_g1", "delta_g1"):
            shapes[k] = (G1_W,)
        for k in ("beta_g2", "gamma_g2", "delta_g2"):
            shapes[k] = (G2_W,)
        shapes.update(gamma_abc_g1=(self.ni, G1_W), a_query=(m, G1_W), b_g1_query=(m, G1_W), b_g2_query=(m, G2_W),
                      h_query=(self.n - 1, G1_W), l_query=(self.nw, G1_W))
        out, dst = {}, ParamsOut()
        for k in fields:
            out[k] = np.zeros(shapes[k], dtype=np.uint32)
            setattr(dst, k, out[k].ctypes.data)
        self.ctx._check(load().ctg_params_download(self.ctx.h, self.handle, C.byref(dst)))
        return out

    def _serialize(self, fn):
        n = C.c_uint64()
        self.ctx._check(fn(self.ctx.h, self.handle, None, C.c_uint64(0), C.byref(n)))
        out = np.zeros(int(n.value), dtype=np.uint8)
        self.ctx._check(fn(self.ctx.h, self.handle, C.c_void_p(out.ctypes.data), C.c_uint64(out.size), C.byref(n)))
        return out.tobytes()

    def serialize_unchecked(self):
        """circuit.zkey image: ProvingKey::serialize_unchecked (lib/src/utils.rs:91)."""
        return self._serialize(load().ctg_params_serialize_unchecked)

    def serialize_vk(self):
        """verification.key image: VerifyingKey::serialize, compressed (lib/src/utils.rs:94)."""
        return self._serialize(load().ctg_params_serialize_vk)

    def free(self):
        if self.handle:
            load().ctg_params_free(self.ctx.h, self.handle)
            self.handle = None


def proof_to_bytes(proof):
    """ark_to_bytes(proof): 192 compressed bytes.  proof = (a, b, c) word arrays.  Host-only."""
    a, b, c = proof
    buf = np.concatenate([_arr(a).reshape(G1_W), _arr(b).reshape(G2_W), _arr(c).reshape(G1_W)])
    out = np.zeros(192, dtype=np.uint8)
    if load().ctg_proof_serialize(_ptr(buf), C.c_void_p(out.ctypes.data)) != 0:
        raise CtgError("proof_serialize failed")
    return out.tobytes()


def proof_from_bytes(data, check_subgroup=True):
    """ark_from_bytes::<Proof>: decompress (square roots) and, as arkworks' checked deserialize, test the subgroup."""
    raw = np.frombuffer(bytes(data), dtype=np.uint8)
    if raw.size != 192:
        raise CtgError("proof_from_bytes: expected 192 bytes")
    out = np.zeros(G1_W + G2_W + G1_W, dtype=np.uint32)
    if load().ctg_proof_deserialize(C.c_void_p(raw.ctypes.data), int(bool(check_subgroup)), _ptr(out)) != 0:
        raise CtgError("proof_from_bytes: invalid encoding (arkworks: SerializationError::InvalidData)")
    return out[:G1_W].copy(), out[G1_W:G1_W + G2_W].copy(), out[G1_W + G2_W:].copy()


# ---- the reference's own Poseidon parameter set -----------------------------------------------------------
R_MOD_FR = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001


def _fr_from_str(dec):
    """ark-ff 0.3 FromStr for a prime field: the decimal digits are accumulated IN the field (res = res * 10 + digit), which
    reduces numbers larger than the modulus -- the reference's table holds 377-bit decimals (circuits-ark/src/poseidon.rs:18-33)."""
    if not dec or not dec.isdigit() or (dec[0] == "0" and len(dec) > 1):
        raise CtgError("from_str: not a decimal field element")
    acc = 0
    for ch in dec:
        acc = (acc * 10 + ord(ch) - 48) % R_MOD_FR
    return acc


def _fr_mont_words(vals):
    buf = b"".join(((v % R_MOD_FR) * (1 << 256) % R_MOD_FR).to_bytes(32, "little") for v in vals)
    return np.frombuffer(buf, dtype=np.uint32).reshape(len(vals), FR_W).copy()


def poseidon_reference_params():
    """get_poseidon_params::<JubJub>(2) of the reference (circuits-ark/src/poseidon.rs:13-41): the P1 table of
    circuits-ark/src/parameters.rs:12-229 (rate 2, alpha 17, 8 full + 31 partial rounds) parsed into BLS12-381 Fr as
    Fr::from_str does.  Returns (full_rounds, partial_rounds, alpha, mds (3, 3, 8), ark (39, 3, 8)) -- Montgomery words,
    ready for Context.poseidon(*poseidon_reference_params())."""
    import json
    with open(os.path.join(_HERE, "data", "poseidon_bls12_381_rate2.json")) as f:
        d = json.load(f)
    if d["rate"] != 2:
        raise CtgError("poseidon_reference_params: the device sponge is rate 2 / capacity 1")
    mds = _fr_mont_words([_fr_from_str(v) for row in d["mds"] for v in row]).reshape(3, 3, FR_W)
    ark = _fr_mont_words([_fr_from_str(v) for row in d["ark"] for v in row]).reshape(len(d["ark"]), 3, FR_W)
    return int(d["full_rounds"]), int(d["partial_rounds"]), int(d["alpha"]), mds, ark
