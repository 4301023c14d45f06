"""The allocation-free decode entry point, called through the C ABI.

dx_decoder_decode_noalloc exists for hosts that fix all of their device memory at start-up (a Spark
executor sharing the GPU, a latency-bound ingest loop) and therefore cannot let a per-batch call reach
cudaMalloc/cudaFree, which synchronise the whole device.  It must produce exactly what the ordinary
entry point (dx_decoder_decode) produces — bit for bit, the tolerance used by every decode test —
with caller-owned outputs, also on a second call that reuses the shared in-library scratch array right
after the first.  This file sorts last on purpose: a run that stops at the first failure has then
already exercised every other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

import dxo
from adversarial import all_records, to_ndjson

pytestmark = pytest.mark.gpu

SCHEMA2 = dxo.iot_schema_json(2)


def bind(ctx):
    fn = ctx.L.dx_decoder_decode_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.POINTER(C.c_void_p), C.c_int, C.c_void_p,
                   C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    return fn


def decoder_dictionary(ctx, dec, column):
    """code -> bytes for a string leaf of a decoder that only ever ran the allocation-free call."""
    L = ctx.L
    L.dx_decoder_dict_size.restype, L.dx_decoder_dict_size.argtypes = C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int64)]
    L.dx_decoder_dict_string.restype = C.c_int
    L.dx_decoder_dict_string.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    n = C.c_int64()
    ctx.check(L.dx_decoder_dict_size(dec.h, column, C.byref(n)))
    out = []
    for code in range(n.value):
        p, ln = C.c_void_p(), C.c_int64()
        ctx.check(L.dx_decoder_dict_string(dec.h, column, code, C.byref(p), C.byref(ln)))
        out.append(C.string_at(p.value, ln.value) if ln.value else b"")
    return out


def decode_both_ways(dx, ctx, ref_dec, dec, fn, data, n_events):
    """-> None; asserts the caller-owned outputs equal the ordinary entry point's table: numbers and validity bit for bit,
    strings byte for byte through each decoder's own dictionary.  `dec` never sees the ordinary entry point, so every
    dictionary entry it holds was inserted by the allocation-free call itself."""
    buf = np.frombuffer(data, dtype=np.uint8)
    src = ctx.alloc(len(data) + 64)
    ctx.h2d(src, buf)
    ref = ref_dec.decode_device(src, len(data), n_events=n_events)      # existing entry point, on its own decoder
    names = ref.columns
    widths = [4 if ref.column_type(i) == 3 else 8 for i in range(len(names))]
    cols = [ctx.alloc(max(1, n_events) * w) for w in widths]             # outputs owned by the caller
    mask_w = 1 if len(names) <= 7 else 4
    mask = ctx.alloc(max(1, n_events) * mask_w)
    arr = (C.c_void_p * len(cols))(*cols)
    rows, nulls = C.c_int64(-1), C.c_int64(-1)
    ctx.check(fn(dec.h, C.c_void_p(src), len(data), n_events, arr, len(cols), C.c_void_p(mask), C.byref(rows), C.byref(nulls)))
    assert rows.value == n_events
    got_mask = ctx.d2h(mask, n_events * mask_w)
    got_mask = got_mask if mask_w == 1 else got_mask.view(np.uint32)
    n_null_rows = 0
    for i, name in enumerate(names):
        raw = ctx.d2h(cols[i], n_events * widths[i])
        got = raw.view(np.int32 if widths[i] == 4 else np.int64)
        exp = np.zeros(n_events, dtype=got.dtype)
        exp_valid = np.zeros(n_events, dtype=np.uint8)
        ctx.check(ctx.L.dx_table_column_to_host(ref.h, i, exp.ctypes.data_as(C.c_void_p), exp_valid.ctypes.data_as(C.c_void_p)))
        assert (((got_mask >> i) & 1).astype(np.uint8) == exp_valid).all(), name
        if widths[i] == 4:
            # the two decoders number their strings independently: compare the text, and the NULL marker (-1) as it is
            mine, theirs = decoder_dictionary(ctx, dec, i), ref.dictionary(i)
            ok = exp_valid.astype(bool)
            assert (got[~ok] == exp[~ok]).all(), name
            assert [mine[c] for c in got[ok]] == [theirs[c] for c in exp[ok]], name
        else:
            assert (got == exp).all(), name                               # bit-exact, including the zeroed NULL slots
    top = 7 if mask_w == 1 else 31
    n_null_rows = int((((got_mask >> top) & 1) == 0).sum())
    assert nulls.value == n_null_rows == dec.stats()["null_rows"] == ref_dec.stats()["null_rows"]
    for p in cols + [mask, src]:
        ctx.free(p)
    ref.release()


def test_noalloc_decode_matches_ordinary_entry_point_twice_in_a_row(dx, ctx):
    fn = bind(ctx)
    ref_dec = dx.Decoder(ctx, SCHEMA2)      # only ever runs dx_decoder_decode
    dec = dx.Decoder(ctx, SCHEMA2)          # only ever runs dx_decoder_decode_noalloc: its dictionary inserts happen there
    # first call: clean generated events spanning many tiles
    p = dxo.GenParams(20260104, 5000, 500, 1_700_000_040_000, 60_000, 60_000, 2, 10, 1, 0)
    n1 = 120_000
    decode_both_ways(dx, ctx, ref_dec, dec, fn, dxo.gen_events(p, 0, n1), n1)
    # second call straight after, smaller and messy: stale counters / tile states from call 1 must not leak in
    recs = all_records(4000, seed=11)
    decode_both_ways(dx, ctx, ref_dec, dec, fn, to_ndjson(recs), len(recs))
    # and a third, larger again
    decode_both_ways(dx, ctx, ref_dec, dec, fn, dxo.gen_events(p, n1, 40_000), 40_000)
    dec.close()
    ref_dec.close()


def test_noalloc_decode_rejects_what_it_cannot_hold(dx, ctx):
    fn = bind(ctx)
    dec = dx.Decoder(ctx, SCHEMA2)
    n_cols = 7
    cols = [ctx.alloc(64) for _ in range(n_cols)]
    mask = ctx.alloc(64)
    arr = (C.c_void_p * n_cols)(*cols)
    src = ctx.alloc(64)
    too_big = 65536 * 32768 + 1                       # one byte past the fixed scratch's reach; rejected before any access
    rc = fn(dec.h, C.c_void_p(src), too_big, 1, arr, n_cols, C.c_void_p(mask), None, None)
    assert rc != 0
    msg = ctx.L.dx_ctx_last_error(ctx.h).decode()
    assert "65536 tiles" in msg and "split the batch" in msg
    assert fn(dec.h, C.c_void_p(src), 10, 1, arr, n_cols - 1, C.c_void_p(mask), None, None) != 0      # wrong column count
    assert fn(dec.h, C.c_void_p(src), 10, -1, arr, n_cols, C.c_void_p(mask), None, None) != 0          # n_events must be stated
    for p in cols + [mask, src]:
        ctx.free(p)
    dec.close()
