"""Thin object wrappers over the C ABI: Context, Decoder, Table.  numpy only; torch is optional
plumbing used by bench.py, never required here."""
import ctypes as C

import numpy as np

from . import _native as N
from ._native import DxError, GenParams

TYPE_NAMES = {1: "long", 2: "double", 3: "string", 4: "boolean", 5: "timestamp", 6: "flags"}


class Context:
    def __init__(self, device=0):
        self.L = N.lib()
        h = C.c_void_p()
        rc = self.L.dx_ctx_create(device, C.byref(h))
        if rc:
            raise DxError(f"dx_ctx_create(device={device}) failed with status {rc} "
                          "(needs an sm_100 GPU; this library has no other code path)")
        self.h = h
        self.device = device

    def check(self, rc):
        if rc:
            raise DxError(self.L.dx_ctx_last_error(self.h).decode("utf-8", "replace"))

    def close(self):
        if self.h:
            self.L.dx_ctx_destroy(self.h)
            self.h = None

    def set_timing(self, on):
        self.check(self.L.dx_ctx_set_timing(self.h, 1 if on else 0))

    def take_stage_ms(self, stage):
        t, n = C.c_double(), C.c_int64()
        self.check(self.L.dx_ctx_take_stage_ms(self.h, stage.encode(), C.byref(t), C.byref(n)))
        return t.value, n.value

    def take_launches(self):
        return int(self.L.dx_ctx_take_launches(self.h))

    def take_counter(self, name):
        fn = self.L.dx_ctx_take_counter
        fn.restype, fn.argtypes = C.c_int64, [C.c_void_p, C.c_char_p]
        return int(fn(self.h, name.encode()))

    def synchronize(self):
        self.check(self.L.dx_ctx_synchronize(self.h))

    def stream(self):
        return self.L.dx_ctx_stream(self.h)

    def alloc(self, nbytes):
        p = C.c_void_p()
        self.check(self.L.dx_dev_alloc(self.h, nbytes, C.byref(p)))
        return p.value

    def free(self, ptr):
        self.L.dx_dev_free(self.h, C.c_void_p(ptr))

    def h2d(self, dst, arr):
        arr = np.ascontiguousarray(arr)
        self.check(self.L.dx_copy_h2d(self.h, C.c_void_p(dst), arr.ctypes.data_as(C.c_void_p), arr.nbytes))

    def d2h(self, src, nbytes):
        out = np.empty(nbytes, dtype=np.uint8)
        if nbytes:
            self.check(self.L.dx_copy_d2h(self.h, out.ctypes.data_as(C.c_void_p), C.c_void_p(src), nbytes))
        return out

    def gunzip(self, members):
        """gzip members (list of bytes, one per blob file) -> (device pointer, n_bytes) of their contents back to back, inflated on
        the device (dx_gunzip_device); free the pointer with free()."""
        fn = self.L.dx_gunzip_device
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
        blob = np.frombuffer(b"".join(members) + b"\0", dtype=np.uint8)
        off = np.zeros(len(members) + 1, dtype=np.int64)
        off[1:] = np.cumsum([len(m) for m in members])
        p, n = C.c_void_p(), C.c_int64()
        self.check(fn(self.h, blob.ctypes.data_as(C.c_void_p), off.ctypes.data_as(C.c_void_p), len(members), 0, C.byref(p), C.byref(n)))
        return p.value, n.value

    def gen_events(self, params: GenParams, first, n, want_offsets=False):
        """Synthetic events on the device -> (device_ptr, n_bytes, offsets_ptr|None)."""
        b, nb, off = C.c_void_p(), C.c_int64(), C.c_void_p()
        self.check(self.L.dx_gen_events(self.h, C.byref(params), first, n, C.byref(b), C.byref(nb),
                                        C.byref(off) if want_offsets else None))
        return b.value, nb.value, (off.value if want_offsets else None)


class Table:
    def __init__(self, ctx, handle):
        self.ctx, self.h = ctx, handle
        self.L = ctx.L

    def release(self):
        if self.h and self.ctx.h:      # handles die with their context: never touch one after dx_ctx_destroy
            self.L.dx_table_release(self.h)
        self.h = None

    __del__ = release

    @property
    def num_rows(self):
        return int(self.L.dx_table_num_rows(self.h))

    @property
    def columns(self):
        n = self.L.dx_table_num_columns(self.h)
        return [self.L.dx_table_column_name(self.h, i).decode() for i in range(n)]

    def column_type(self, col):
        return self.L.dx_table_column_type(self.h, self._idx(col))

    def nesting(self, col):
        """(key path, container kinds) of a MAP/STRUCT leaf column; ([], "") for an ordinary column."""
        i = self._idx(col)
        kinds = C.c_char_p()
        n = self.L.dx_table_column_nesting(self.h, i, C.byref(kinds))
        return [self.L.dx_table_column_path(self.h, i, k).decode() for k in range(n)], (kinds.value or b"").decode()

    def _idx(self, col):
        if isinstance(col, int):
            return col
        i = self.L.dx_table_find_column(self.h, col.encode())
        if i < 0:
            raise KeyError(col)
        return i

    def column(self, col):
        """-> (values ndarray, valid bool ndarray).  Strings come back as object arrays of bytes."""
        i = self._idx(col)
        t = self.L.dx_table_column_type(self.h, i)
        n = self.num_rows
        valid = np.zeros(n, dtype=np.uint8)
        if t == 6:
            w = self.L.dx_table_flag_words(self.h, i)
            data = np.zeros((w, n), dtype=np.uint32)   # word-major on the device: flags[w][row]
        else:
            data = np.zeros(n, dtype={1: np.int64, 2: np.float64, 3: np.int32, 4: np.int64, 5: np.int64}[t])
        self.ctx.check(self.L.dx_table_column_to_host(self.h, i, data.ctypes.data_as(C.c_void_p),
                                                      valid.ctypes.data_as(C.c_void_p)))
        valid = valid.astype(bool)
        if t == 6:
            return np.ascontiguousarray(data.T), valid      # -> (n_rows, words)
        if t == 3:
            d = self.dictionary(i)
            out = np.empty(n, dtype=object)
            for r in range(n):
                out[r] = d[data[r]] if valid[r] else None
            return out, valid
        return data, valid

    def codes(self, col):
        i = self._idx(col)
        n = self.num_rows
        data = np.zeros(n, dtype=np.int32)
        valid = np.zeros(n, dtype=np.uint8)
        self.ctx.check(self.L.dx_table_column_to_host(self.h, i, data.ctypes.data_as(C.c_void_p),
                                                      valid.ctypes.data_as(C.c_void_p)))
        return data, valid.astype(bool)

    def dictionary(self, col):
        i = self._idx(col)
        n = C.c_int64()
        self.ctx.check(self.L.dx_table_dict_size(self.h, i, C.byref(n)))
        out = []
        for code in range(n.value):
            p, ln = C.c_void_p(), C.c_int64()
            self.ctx.check(self.L.dx_table_dict_string(self.h, i, code, C.byref(p), C.byref(ln)))
            out.append(C.string_at(p.value, ln.value) if ln.value else b"")
        return out

    def to_json(self, device=False):
        """The sink hand-off (OutputManager.sink, sink/OutputManager.scala:103-118): to_json(struct(output columns)) per row,
        serialised on the device.  -> list of bytes, one JSON object per row; device=True -> (bytes_ptr, offsets_ptr, n_rows,
        n_bytes, release) without the copy to the host."""
        L = self.L
        L.dx_table_to_json.restype, L.dx_table_to_json.argtypes = C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]
        L.dx_json_num_rows.restype, L.dx_json_num_rows.argtypes = C.c_int64, [C.c_void_p]
        L.dx_json_num_bytes.restype, L.dx_json_num_bytes.argtypes = C.c_int64, [C.c_void_p]
        L.dx_json_device.restype, L.dx_json_device.argtypes = C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.dx_json_to_host.restype, L.dx_json_to_host.argtypes = C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]
        L.dx_json_release.restype, L.dx_json_release.argtypes = None, [C.c_void_p]
        j = C.c_void_p()
        self.ctx.check(L.dx_table_to_json(self.h, C.byref(j)))
        n, nb = L.dx_json_num_rows(j), L.dx_json_num_bytes(j)
        if device:
            b, o = C.c_void_p(), C.c_void_p()
            self.ctx.check(L.dx_json_device(j, C.byref(b), C.byref(o)))
            return b.value, o.value, n, nb, (lambda: L.dx_json_release(j))
        data = np.empty(max(nb, 1), dtype=np.uint8)
        off = np.zeros(n + 1, dtype=np.int64)
        try:
            self.ctx.check(L.dx_json_to_host(j, data.ctypes.data_as(C.c_void_p), off.ctypes.data_as(C.c_void_p)))
        finally:
            L.dx_json_release(j)
        raw = data.tobytes()
        return [raw[off[r]:off[r + 1] - 1] for r in range(n)]

    def filter_flags(self, expr):
        """One sink's filter column (OutputManager.scala:53-61): bool per row, TRUE only where the expression is TRUE."""
        L = self.L
        L.dx_table_filter_flags.restype, L.dx_table_filter_flags.argtypes = C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p]
        out = np.zeros(max(self.num_rows, 1), dtype=np.uint8)
        self.ctx.check(L.dx_table_filter_flags(self.h, expr.encode("utf-8"), out.ctypes.data_as(C.c_void_p)))
        return out[: self.num_rows].astype(bool)

    def device_view(self, col):
        i = self._idx(col)
        d, m, w, b = C.c_void_p(), C.c_void_p(), C.c_int(), C.c_int()
        self.ctx.check(self.L.dx_table_column_device(self.h, i, C.byref(d), C.byref(m), C.byref(w), C.byref(b)))
        return d.value, m.value, w.value, b.value


class Decoder:
    """K1.  Mirrors from_json(Raw, rawBlockSchema) (CommonProcessorFactory.scala:93)."""

    def __init__(self, ctx, schema_json: str):
        self.ctx = ctx
        self.L = ctx.L
        h = C.c_void_p()
        rc = self.L.dx_decoder_create(ctx.h, schema_json.encode("utf-8"), C.byref(h))
        ctx.check(rc)
        self.h = h

    def close(self):
        if self.h:
            self.L.dx_decoder_destroy(self.h)
            self.h = None

    def decode_host(self, data: bytes, offsets=None, n_events=-1):
        buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        t = C.c_void_p()
        off_p = None
        if offsets is not None:
            offsets = np.ascontiguousarray(offsets, dtype=np.int64)
            off_p = offsets.ctypes.data_as(C.c_void_p)
            n_events = len(offsets) - 1
        ptr = buf.ctypes.data_as(C.c_void_p) if buf.size else None
        self.ctx.check(self.L.dx_decoder_decode(self.h, ptr, buf.size, off_p, n_events, 0, C.byref(t)))
        return Table(self.ctx, t)

    def decode_device(self, ptr, n_bytes, offsets_ptr=None, n_events=-1):
        t = C.c_void_p()
        self.ctx.check(self.L.dx_decoder_decode(self.h, C.c_void_p(ptr), n_bytes,
                                                C.c_void_p(offsets_ptr) if offsets_ptr else None,
                                                n_events, 1, C.byref(t)))
        return Table(self.ctx, t)

    def stats(self):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        self.L.dx_decoder_stats(self.h, C.byref(a), C.byref(b), C.byref(c))
        return {"rows": a.value, "null_rows": b.value, "fast_rows": c.value}
