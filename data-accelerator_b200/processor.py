"""Host-side mirror of the reference's processor interface, over the C ABI.

The reference builds its processor with
    CommonProcessorFactory.createProcessor(config: UnifiedConfig): CommonProcessor
(datax/processor/CommonProcessorFactory.scala:50) and drives it per micro-batch through
    processJson / processEventData / ... -> processDataset(data, batchTime, batchInterval,
                                                           outputPartitionTime, targets, namespace)
(CommonProcessorFactory.scala:333-338, 405-461) which returns the metrics Map[String, Double].
There is no JVM in this image, so the Scala shim cannot be compiled here; this
module is the same surface in Python — same names, argument meaning and error behaviour — and
is what the parity tests drive.  All work happens in libdxgpu.so; nothing here computes.
"""
import ctypes as C
import os

import numpy as np

from . import _native as N
from ._native import DxError
from .device import Context, Table

vp, i64, i32, cp = C.c_void_p, C.c_int64, C.c_int, C.c_char_p
P = C.POINTER
_SIG = {
    "dx_flow_create": (i32, [vp, cp, cp, P(cp), i32, cp, P(vp)]),
    "dx_flow_destroy": (None, [vp]),
    "dx_flow_process_batch": (i32, [vp, cp, vp, i64, vp, i64, i32, i64, i64, i64, i64, P(vp)]),
    "dx_result_metrics": (i32, [vp, P(P(cp)), P(P(C.c_double)), P(i32)]),
    "dx_result_num_tables": (i32, [vp]),
    "dx_result_table_name": (cp, [vp, i32]),
    "dx_result_table": (i32, [vp, cp, P(vp)]),
    "dx_result_release": (None, [vp]),
    "dx_flow_state_export": (i32, [vp, cp, P(vp), P(i32)]),
    "dx_table_rule_count": (i32, [vp, i32]),
    "dx_table_rule_json": (i32, [vp, i32, i32, P(vp), P(i64)]),
    "dx_sql_canonical": (i32, [cp, C.c_char_p, i64]),
    "dx_transform_split": (i32, [cp, C.c_char_p, i64]),
    "dx_string_to_timestamp": (i32, [cp, P(i64)]),
    "dx_duration_seconds": (i64, [cp]),
}
_bound = False


def _lib():
    global _bound
    L = N.lib()
    if not _bound:
        for name, (res, args) in _SIG.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _bound = True
    return L


# ---- host-side artefact parsers (no GPU) ---------------------------------------------------------
def sql_canonical(sql: str) -> str:
    buf = C.create_string_buffer(1 << 20)
    rc = _lib().dx_sql_canonical(sql.encode("utf-8"), buf, len(buf))
    if rc:
        raise DxError(buf.value.decode("utf-8", "replace"))
    return buf.value.decode("utf-8")


def transform_split(text: str):
    """TransformSQLParser.parse -> (commands [(name|None, type, text)], viewReferenceCount dict)."""
    buf = C.create_string_buffer(4 << 20)
    rc = _lib().dx_transform_split(text.encode("utf-8"), buf, len(buf))
    if rc:
        raise DxError(buf.value.decode("utf-8", "replace"))
    cmds, refs = [], {}
    for line in buf.value.decode("utf-8").split("\n"):
        if not line:
            continue
        a, b, c = line.split("\t", 2)
        if a == "#ref":
            refs[b] = int(c)
        else:
            cmds.append((a or None, b, c))
    return cmds, refs


def string_to_timestamp(s: str):
    """stringToTimestamp built-in -> epoch microseconds (UTC session) or None."""
    us = C.c_int64()
    return us.value if _lib().dx_string_to_timestamp(s.encode("utf-8"), C.byref(us)) else None


def duration_seconds(s: str) -> int:
    return int(_lib().dx_duration_seconds(s.encode("utf-8")))


# ---- the processor ------------------------------------------------------------------------------
class BatchResult:
    def __init__(self, ctx, handle):
        self.ctx, self.h, self.L = ctx, handle, ctx.L

    @property
    def metrics(self):
        keys, vals, n = P(cp)(), P(C.c_double)(), C.c_int()
        self.ctx.check(self.L.dx_result_metrics(self.h, C.byref(keys), C.byref(vals), C.byref(n)))
        return {keys[i].decode(): vals[i] for i in range(n.value)}

    @property
    def table_names(self):
        return [self.L.dx_result_table_name(self.h, i).decode() for i in range(self.L.dx_result_num_tables(self.h))]

    def table(self, name) -> Table:
        t = C.c_void_p()
        self.ctx.check(self.L.dx_result_table(self.h, name.encode(), C.byref(t)))
        return Table(self.ctx, t)

    def release(self):
        if self.h and self.ctx.h:      # handles die with their context: never touch one after dx_ctx_destroy
            self.L.dx_result_release(self.h)
        self.h = None

    __del__ = release


def rules_of(table: Table, col="Rules"):
    """Static rule table behind a FLAGS column -> list of dicts (the MAP(...) constants)."""
    import json
    L = _lib()
    i = table._idx(col)
    out = []
    for r in range(L.dx_table_rule_count(table.h, i)):
        p, ln = C.c_void_p(), C.c_int64()
        table.ctx.check(L.dx_table_rule_json(table.h, i, r, C.byref(p), C.byref(ln)))
        out.append(json.loads(C.string_at(p.value, ln.value).decode("utf-8")))
    return out


def rule_hits(table: Table, col="Rules"):
    """Rows for which each rule of a FLAGS column fired -> int64 ndarray, one entry per rule."""
    L = _lib()
    i = table._idx(col)
    n = L.dx_table_rule_count(table.h, i)
    out = np.zeros(max(n, 1), dtype=np.int64)
    fn = L.dx_table_rule_hits
    fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    table.ctx.check(fn(table.h, i, out.ctypes.data_as(C.c_void_p), n))
    return out[:n]


class CommonProcessor:
    """What createProcessor returns (CommonProcessor case class, datax/processor/CommonProcessor.scala:17-31),
    reduced to the ingest adapters this ABI serves: JSON strings, in host or device memory."""

    def __init__(self, ctx: Context, conf_text, schema_json, projection_steps, transform_text):
        self.ctx = ctx
        self.L = _lib()
        steps = (cp * len(projection_steps))(*[s.encode("utf-8") for s in projection_steps])
        h = C.c_void_p()
        rc = self.L.dx_flow_create(ctx.h, conf_text.encode("utf-8"), schema_json.encode("utf-8"), steps,
                                   len(projection_steps), transform_text.encode("utf-8"), C.byref(h))
        ctx.check(rc)
        self.h = h

    def close(self):
        if self.h and self.ctx.h:
            self.L.dx_flow_destroy(self.h)
        self.h = None

    def process_json(self, data, batch_time_ms, batch_interval_ms, output_partition_time_ms=None, namespace="",
                     now_ms=None, offsets=None, n_events=-1) -> BatchResult:
        """processJson (CPF:458-461): data = NDJSON bytes (or bytes + offsets) in HOST memory."""
        buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        off_p = None
        if offsets is not None:
            offsets = np.ascontiguousarray(offsets, dtype=np.int64)
            off_p = offsets.ctypes.data_as(C.c_void_p)
            n_events = len(offsets) - 1
        r = C.c_void_p()
        ptr = buf.ctypes.data_as(C.c_void_p) if buf.size else None
        self.ctx.check(self.L.dx_flow_process_batch(
            self.h, namespace.encode(), ptr, buf.size, off_p, n_events, 0, batch_time_ms, batch_interval_ms,
            output_partition_time_ms if output_partition_time_ms is not None else batch_time_ms,
            now_ms if now_ms is not None else batch_time_ms, C.byref(r)))
        return BatchResult(self.ctx, r)

    def process_device(self, ptr, n_bytes, batch_time_ms, batch_interval_ms, namespace="", now_ms=None, n_events=-1,
                       offsets_ptr=None) -> BatchResult:
        """Same, with the batch already resident in device memory."""
        r = C.c_void_p()
        self.ctx.check(self.L.dx_flow_process_batch(
            self.h, namespace.encode(), C.c_void_p(ptr), n_bytes, C.c_void_p(offsets_ptr) if offsets_ptr else None,
            n_events, 1, batch_time_ms, batch_interval_ms, batch_time_ms,
            now_ms if now_ms is not None else batch_time_ms, C.byref(r)))
        return BatchResult(self.ctx, r)

    def set_batch_info(self, input_path=None, output_folders=None, output_file_name=None, file_time_ms=None, rule_index_prefix=None,
                       target=None, properties=None, system_properties=None, executor_id=None, reset=False):
        """The per-batch constants behind __DataX_FileInfo / DataXProperties / Properties / SystemProperties (dx_flow_set_batch_info):
        what a blob or EventHub adapter knows about the partition it hands over next.  reset=True: back to FileInternal()."""
        fn = self.L.dx_flow_set_batch_info
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_void_p]
        if reset:
            self.ctx.check(fn(self.h, None))
            return
        PP = C.POINTER(C.c_char_p)

        class Info(C.Structure):
            _fields_ = [("input_path", C.c_char_p), ("of_k", PP), ("of_v", PP), ("n_of", C.c_int), ("output_file_name", C.c_char_p),
                        ("file_time_ms", C.c_int64), ("has_file_time", C.c_int), ("rule_index_prefix", C.c_char_p), ("target", C.c_char_p),
                        ("p_k", PP), ("p_v", PP), ("n_p", C.c_int), ("sp_k", PP), ("sp_v", PP), ("n_sp", C.c_int), ("executor_id", C.c_char_p)]
        enc = lambda x: None if x is None else x.encode("utf-8")

        def arr(d):
            d = d or {}
            k = (C.c_char_p * max(1, len(d)))(*[x.encode("utf-8") for x in d.keys()])
            v = (C.c_char_p * max(1, len(d)))(*[x.encode("utf-8") for x in d.values()])
            return k, v, len(d)
        ofk, ofv, nof = arr(output_folders)
        pk, pv, np_ = arr(properties)
        spk, spv, nsp = arr(system_properties)
        info = Info(enc(input_path), ofk, ofv, nof if output_folders is not None else -1, enc(output_file_name),
                    file_time_ms or 0, 0 if file_time_ms is None else 1, enc(rule_index_prefix), enc(target), pk, pv, np_, spk, spv, nsp, enc(executor_id))
        self.ctx.check(fn(self.h, C.byref(info)))

    def set_event_property(self, key, values, system=False):
        """Per-event values of one Properties (system=True: SystemProperties) key for the NEXT batch (dx_flow_set_event_property):
        values[i] is bytes / str, or None where event i does not carry the key."""
        fn = self.L.dx_flow_set_event_property
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        enc = [b"" if v is None else (v.encode("utf-8") if isinstance(v, str) else bytes(v)) for v in values]
        off = np.zeros(len(enc) + 1, dtype=np.int64)
        off[1:] = np.cumsum([len(e) for e in enc])
        blob = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8)
        valid = np.array([0 if v is None else 1 for v in values], dtype=np.uint8)
        self.ctx.check(fn(self.h, 1 if system else 0, key.encode("utf-8"), blob.ctypes.data_as(C.c_void_p), off.ctypes.data_as(C.c_void_p),
                          valid.ctypes.data_as(C.c_void_p) if len(values) else None, len(values)))

    def stage_json(self, host_ptr, n_bytes):
        """Starts the host->device copy of the NEXT batch (page-locked host memory at host_ptr) and returns a handle for
        process_staged; the copy overlaps whatever the processor is doing meanwhile."""
        fn = self.L.dx_flow_stage_batch
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]
        h = C.c_void_p()
        self.ctx.check(fn(self.h, C.c_void_p(host_ptr), n_bytes, C.byref(h)))
        return h

    def process_staged(self, staged, batch_time_ms, batch_interval_ms, namespace="", now_ms=None, n_events=-1) -> BatchResult:
        """processJson on a batch whose copy was started with stage_json (the handle is consumed)."""
        fn = self.L.dx_flow_process_staged
        fn.restype = C.c_int
        fn.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.POINTER(C.c_void_p)]
        r = C.c_void_p()
        self.ctx.check(fn(self.h, namespace.encode(), staged, n_events, batch_time_ms, batch_interval_ms, batch_time_ms,
                          now_ms if now_ms is not None else batch_time_ms, C.byref(r)))
        return BatchResult(self.ctx, r)

    def add_refdata(self, name, csv_text, delimiter=",", header=True):
        """Reference data view (ReferenceDataHandler.scala:42-60): CSV text, all columns strings."""
        fn = self.L.dx_flow_add_refdata
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char, C.c_int]
        self.ctx.check(fn(self.h, name.encode(), csv_text.encode("utf-8"), delimiter.encode()[:1], 1 if header else 0))

    def register_udf(self, name, func):
        """A String -> String user function (UDF1), e.g. datax.sample.udf.UdfHelloWorld.  func(str) -> str | None.
        Called by the library once per distinct argument value."""
        proto = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_char), C.c_int64, C.POINTER(C.c_char), C.c_int64, C.POINTER(C.c_int64))

        def thunk(user, arg, arg_len, out, out_cap, out_len):
            try:
                r = func(None if arg_len < 0 else C.string_at(arg, arg_len).decode("utf-8"))
                if r is None:
                    out_len[0] = -1
                    return 0
                b = r.encode("utf-8")
                if len(b) > out_cap:
                    return 1
                C.memmove(out, b, len(b))
                out_len[0] = len(b)
                return 0
            except Exception:
                return 1
        cb = proto(thunk)
        self._udf_keepalive = getattr(self, "_udf_keepalive", []) + [cb]
        fn = self.L.dx_flow_register_udf
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_char_p, proto, C.c_void_p]
        self.ctx.check(fn(self.h, name.encode(), cb, None))

    def register_udfn(self, name, n_args, func, refresh_each_batch=False):
        """A user function of 0 to 3 arguments returning a string (DynamicUDF.Generator0..3 / Java UDF0..UDF3): func(*str) -> str | None.
        Called once per distinct combination of argument values; every batch anew with refresh_each_batch."""
        proto = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.POINTER(C.POINTER(C.c_char)), C.POINTER(C.c_int64), C.POINTER(C.c_char), C.c_int64, C.POINTER(C.c_int64))

        def thunk(user, n, args, lens, out, out_cap, out_len):
            try:
                r = func(*[C.string_at(args[k], lens[k]).decode("utf-8") for k in range(n)])
                if r is None:
                    out_len[0] = -1
                    return 0
                b = r.encode("utf-8")
                if len(b) > out_cap:
                    return 1
                C.memmove(out, b, len(b))
                out_len[0] = len(b)
                return 0
            except Exception:
                return 1
        cb = proto(thunk)
        self._udf_keepalive = getattr(self, "_udf_keepalive", []) + [cb]
        fn = self.L.dx_flow_register_udfn
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_char_p, C.c_int, proto, C.c_void_p, C.c_int]
        self.ctx.check(fn(self.h, name.encode(), n_args, cb, None, 1 if refresh_each_batch else 0))

    def set_interval_hook(self, func):
        """IntervalUpdateHandler.onInterval: func(batch_time_ms) at the start of every batch (CPF:351-353)."""
        proto = C.CFUNCTYPE(None, C.c_void_p, C.c_int64)
        cb = proto(lambda user, t: func(t))
        self._udf_keepalive = getattr(self, "_udf_keepalive", []) + [cb]
        fn = self.L.dx_flow_set_interval_hook
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, proto, C.c_void_p]
        self.ctx.check(fn(self.h, cb, None))

    def set_normalizer(self, func):
        """StringNormalizer (extension/StringNormalizer.scala:10-12): func(str) -> str applied to every record before from_json,
        on the host, over batches passed in host memory."""
        proto = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_char), C.c_int64, C.POINTER(C.c_char), C.c_int64, C.POINTER(C.c_int64))

        def thunk(user, inp, n_in, out, out_cap, n_out):
            try:
                lines = C.string_at(inp, n_in).split(b"\n")
                res = b"\n".join(func(l.decode("utf-8", "replace")).encode("utf-8") if l else l for l in lines)
                if len(res) > out_cap:
                    return 1
                C.memmove(out, res, len(res))
                n_out[0] = len(res)
                return 0
            except Exception:
                return 1
        cb = proto(thunk)
        self._udf_keepalive = getattr(self, "_udf_keepalive", []) + [cb]
        fn = self.L.dx_flow_set_normalizer
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, proto, C.c_void_p]
        self.ctx.check(fn(self.h, cb, None))

    def import_state(self, name, columns, active="A"):
        """Restart (StateTableHandler.scala:70-96): `columns` = [(values ndarray | list of bytes/None for strings, valid
        bool ndarray | None)] in the order of the table's declared schema, as read back from the persisted copy."""
        n = len(columns[0][0]) if columns else 0
        keep, data, valid, sbytes = [], [], [], []
        for vals, ok in columns:
            if isinstance(vals, np.ndarray) and vals.dtype != object:
                a = np.ascontiguousarray(vals.astype(np.float64) if vals.dtype.kind == "f" else vals.astype(np.int64))
                keep.append(a); data.append(a.ctypes.data); sbytes.append(None)
            else:
                off, blob = np.zeros(n + 1, dtype=np.int64), bytearray()
                for r, v in enumerate(vals):
                    if v is not None and (ok is None or ok[r]):
                        blob += v
                    off[r + 1] = len(blob)
                b = np.frombuffer(bytes(blob) or b"\0", dtype=np.uint8)
                keep += [off, b]; data.append(off.ctypes.data); sbytes.append(b.ctypes.data)
            if ok is None:
                valid.append(None)
            else:
                m = np.ascontiguousarray(np.asarray(ok).astype(np.uint8))
                keep.append(m); valid.append(m.ctypes.data)
        k = len(columns)
        fn = self.L.dx_flow_state_import
        fn.restype = C.c_int
        fn.argtypes = [C.c_void_p, C.c_char_p, C.c_int64, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int]
        arr = lambda xs: (C.c_void_p * k)(*[C.c_void_p(x) if x else None for x in xs])
        self.ctx.check(fn(self.h, name.encode(), n, k, arr(data), arr(valid), arr(sbytes), 1 if active == "B" else 0))

    def save_state(self, name, location):
        """The table's active copy as Parquet under <location>/<A|B>/ plus <location>/metadata.info (dx_flow_state_save: what
        StateTableHandler's INSERT OVERWRITE + flip + persist leave behind, StateTableHandler.scala:99-125)."""
        fn = self.L.dx_flow_state_save
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p]
        self.ctx.check(fn(self.h, name.encode(), str(location).encode()))

    def load_state(self, name, location):
        """Restart from <location> (dx_flow_state_load): metadata.info says which copy is active; its part files — written by
        save_state or by Spark 2.4 — become the flow's table."""
        fn = self.L.dx_flow_state_load
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p]
        self.ctx.check(fn(self.h, name.encode(), str(location).encode()))

    def state_table(self, name):
        t, b = C.c_void_p(), C.c_int()
        self.ctx.check(self.L.dx_flow_state_export(self.h, name.encode(), C.byref(t), C.byref(b)))
        return Table(self.ctx, t), ("B" if b.value else "A")


class CommonProcessorFactory:
    """CommonProcessorFactory.createProcessor (CPF:50).  `config` carries what UnifiedConfig.dict points
    at: the .conf text plus the contents of the files it names (schema, projection steps, transform)."""

    @staticmethod
    def createProcessor(ctx: Context, conf_text: str, schema_json: str, projection_steps, transform_text: str):
        return CommonProcessor(ctx, conf_text, schema_json, list(projection_steps), transform_text)

    @staticmethod
    def createProcessorFromFiles(ctx: Context, conf_path: str, read=lambda p: open(p, encoding="utf-8-sig").read()):
        """Loads the artefacts the way the reference does: the .conf names the other files
        (datax.job.input.default.blobschemafile, .process.projection (';'-separated), .process.transform)."""
        conf_text = read(conf_path)
        d = {}
        for line in conf_text.splitlines():
            line = line.strip().lstrip("﻿")
            if line and not line.startswith("#") and "=" in line:
                k, v = line.split("=", 1)
                d[k.strip()] = v.strip()
        base = os.path.dirname(conf_path)
        rel = lambda p: p if os.path.isabs(p) else os.path.join(base, p)
        schema = read(rel(d["datax.job.input.default.blobschemafile"]))
        steps = [read(rel(p)) for p in d.get("datax.job.process.projection", "").split(";") if p]
        transform = read(rel(d["datax.job.process.transform"])) if d.get("datax.job.process.transform") else ""
        proc = CommonProcessor(ctx, conf_text, schema, steps, transform)
        # reference data views (ReferenceDataHandler.scala:25-60)
        pre = "datax.job.input.default.referencedata."
        for name in sorted({k[len(pre):].split(".")[0] for k in d if k.startswith(pre)}):
            path = d.get(pre + name + ".path")
            if path and d.get(pre + name + ".format", "csv").lower() == "csv":
                proc.add_refdata(name, read(rel(path)), d.get(pre + name + ".delimiter", ","),
                                 d.get(pre + name + ".header", "true").lower() == "true")
        return proc
