"""ctg_field_op_batch_noalloc (csrc/noalloc.cu): the element-wise field kernel staged in a fixed 3 MiB file-scope device array, for
callers that must not call cudaMalloc.  No counterpart in the reference, so the
checker here is the existing entry point ctg_field_op_batch -- itself checked against the oracle in tests/test_gpu_parity.py --
and the comparison is bit-exact (integer arithmetic), through the C ABI, on two calls in a row that share the array.

Collected last on purpose; passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

import ctg

pytestmark = pytest.mark.gpu

LIMIT = {ctg.FIELD_FR: 32768, ctg.FIELD_FQ: 21845, ctg.FIELD_FQ2: 10922}
WORDS = {ctg.FIELD_FR: 8, ctg.FIELD_FQ: 12, ctg.FIELD_FQ2: 24}
# top words below these keep every component under the modulus (r = 0x73ed..., p = 0x1a01...)
TOP = {ctg.FIELD_FR: 0x73000000, ctg.FIELD_FQ: 0x1A000000, ctg.FIELD_FQ2: 0x1A000000}


def _rand(field, n, seed):
    g = np.random.default_rng(seed)
    w = WORDS[field]
    a = g.integers(0, 1 << 32, size=(n, w), dtype=np.uint64).astype(np.uint32)
    for top in range(11 if field != ctg.FIELD_FR else 7, w, 12):
        a[:, top] %= TOP[field]
    return a


@pytest.fixture()
def ctx():
    c = ctg.Context(0)
    yield c
    c.close()


def test_noalloc_entry_point_equals_field_op_batch_on_two_calls_in_a_row(ctx):
    # first call: the largest size the array holds; second call right after it: a smaller batch of other operands and another
    # op, so that anything left over from the first call (operands or results beyond the new n) would show
    for field in (ctg.FIELD_FR, ctg.FIELD_FQ, ctg.FIELD_FQ2):
        n1, n2 = LIMIT[field], 1000 + field
        a1, b1 = _rand(field, n1, 10 + field), _rand(field, n1, 20 + field)
        a2, b2 = _rand(field, n2, 30 + field), _rand(field, n2, 40 + field)
        got1 = ctx.field_op_noalloc(field, ctg.OP_MUL, a1, b1)
        got2 = ctx.field_op_noalloc(field, ctg.OP_SUB, a2, b2)
        assert np.array_equal(got1, ctx.field_op(field, ctg.OP_MUL, a1, b1)), field
        assert np.array_equal(got2, ctx.field_op(field, ctg.OP_SUB, a2, b2)), field
        # and once more with the first operands: the array holds the second call's data now
        assert np.array_equal(ctx.field_op_noalloc(field, ctg.OP_ADD, a1, b1), ctx.field_op(field, ctg.OP_ADD, a1, b1)), field


def test_noalloc_entry_point_rejects_what_does_not_fit(ctx):
    lib = ctg.load()
    for field in (ctg.FIELD_FR, ctg.FIELD_FQ, ctg.FIELD_FQ2):
        n = LIMIT[field] + 1
        a = np.zeros((n, WORDS[field]), dtype=np.uint32)
        out = np.zeros_like(a)
        rc = lib.ctg_field_op_batch_noalloc(ctx.h, field, ctg.OP_ADD, a.ctypes.data_as(C.c_void_p), a.ctypes.data_as(C.c_void_p),
                                            out.ctypes.data_as(C.c_void_p), C.c_size_t(n))
        assert rc == -1                                               # CTG_ERR_INVALID
        assert b"exceed the fixed 3 MiB scratch" in lib.ctg_last_error(ctx.h)
    with pytest.raises(ctg.CtgError):                                 # only mul / add / sub
        ctx.field_op_noalloc(ctg.FIELD_FR, ctg.OP_INV, np.zeros((4, 8), dtype=np.uint32), np.zeros((4, 8), dtype=np.uint32))
    assert ctx.field_op_noalloc(ctg.FIELD_FR, ctg.OP_MUL, np.zeros((0, 8), dtype=np.uint32), np.zeros((0, 8), dtype=np.uint32)).shape == (0, 8)
