"""CPU ORACLE — Python side (test infrastructure, NOT product code).

Wraps oracle/libdxo.so (dxo_core.c) and restates the reference's host-side control flow in plain
Python/numpy.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs import this module; the product path (data-accelerator_b200/) never does.

PARITY UNPINNED at the Spark boundary (see dxo_core.c header and DESIGN.md §2); pinned pieces:
  TransformSQLParser.parse / replaceTableNames  <- datax/sql/TransformSqlParser.scala:27-104,
      KAT datax/test/sql/TransformSQLParserTests.scala:11-21
  string_to_timestamp                           <- datax/utility/ConcurrentDateFormat.scala:38-62,
      KAT datax/utility/test/ConcurrentDateFormatTests.scala:13-15
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
last_limited = 0

LONG, DOUBLE, STRING, BOOL, STRUCT = 1, 2, 3, 4, 5
_TYPE = {"long": LONG, "integer": LONG, "int": LONG, "short": LONG, "byte": LONG, "double": DOUBLE,
         "float": DOUBLE, "string": STRING, "boolean": BOOL}


class _Field(C.Structure):
    _fields_ = [("name", C.c_char_p), ("name_len", C.c_int32), ("type", C.c_int32), ("target", C.c_int32)]


class _Struct(C.Structure):
    _fields_ = [("first_field", C.c_int32), ("n_fields", C.c_int32), ("first_leaf", C.c_int32),
                ("n_leaves", C.c_int32)]


class GenParams(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_keys", C.c_int64), ("n_homes", C.c_int64), ("base_ms", C.c_int64),
                ("interval_ms", C.c_int64), ("watermark_ms", C.c_int64), ("shape", C.c_int32),
                ("late_permille", C.c_int32), ("future_permille", C.c_int32), ("pad", C.c_int32)]


class Agg(C.Structure):
    _fields_ = [("count_star", C.c_int64), ("count_v", C.c_int64), ("sum_i", C.c_int64), ("sum_d", C.c_double),
                ("min_i", C.c_int64), ("max_i", C.c_int64), ("min_d", C.c_double), ("max_d", C.c_double)]


def _lib_name():
    """libdxo.<tag>.so, tag = hash of this host's CPU model + flags: the oracle is built -march=native, so a copy built on
    another machine (the build container vs the GPU box) must not be loaded — it is rebuilt here instead (2 s of gcc)."""
    import hashlib
    key = ""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith(("model name", "flags")):
                key += line
            if line.startswith("flags"):
                break
    except OSError:
        pass
    return "libdxo.%s.so" % hashlib.md5(key.encode()).hexdigest()[:8]


def build():
    tmp = "libdxo.tmp%d.so" % os.getpid()       # built aside and renamed, so a concurrent loader never sees a partial file
    subprocess.check_call(["make", "-s", "-C", _HERE, "OUT=" + tmp])
    os.replace(os.path.join(_HERE, tmp), os.path.join(_HERE, _lib_name()))


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, _lib_name())
        src_time = max(os.path.getmtime(os.path.join(_HERE, f)) for f in ("dxo_core.c", "dxo_rules.c"))
        if not os.path.exists(path) or os.path.getmtime(path) < src_time:
            build()
        _LIB = C.CDLL(path)
        _LIB.dxo_decode.restype = C.c_int64
        _LIB.dxo_decode_mt.restype = C.c_int64
        _LIB.dxo_decode_mt_h.restype = C.c_int64
        _LIB.dxo_split.restype = C.c_int64
        _LIB.dxo_gen_events.restype = C.c_int64
        _LIB.dxo_group_agg.restype = C.c_int64
    return _LIB


class Schema:
    """Flattened view of a Spark StructType JSON (SchemaFile.loadInputSchema, SchemaFile.scala:22-34)."""

    def __init__(self, schema_json: str):
        root = json.loads(schema_json.lstrip("﻿"))
        self.structs, self.fields, self.leaf_names, self.leaf_types = [], [], [], []
        self._keep = []
        self._add(root, "")
        self.c_structs = (_Struct * len(self.structs))(*self.structs)
        self.c_fields = (_Field * len(self.fields))(*self.fields)

    def _add(self, st, prefix):
        sidx = len(self.structs)
        self.structs.append(None)
        flds = st["fields"]
        first = len(self.fields)
        self.fields.extend([None] * len(flds))
        first_leaf = len(self.leaf_names)
        for i, f in enumerate(flds):
            name = f["name"].encode("utf-8")
            self._keep.append(name)
            ty = f["type"]
            if isinstance(ty, str):
                if ty not in _TYPE:
                    raise ValueError(f"unsupported field type {ty}")
                t = _TYPE[ty]
                self.fields[first + i] = _Field(name, len(name), t, len(self.leaf_names))
                self.leaf_names.append(prefix + f["name"])
                self.leaf_types.append(t)
            elif isinstance(ty, dict) and ty.get("type") == "struct":
                child = self._add(ty, prefix + f["name"] + ".")
                self.fields[first + i] = _Field(name, len(name), STRUCT, child)
            else:
                raise ValueError("unsupported field type (array/map)")
        self.structs[sidx] = _Struct(first, len(flds), first_leaf, len(self.leaf_names) - first_leaf)
        return sidx


def split_ndjson(data: bytes):
    """NDJSON framing -> (starts, ends) int64 arrays."""
    buf = np.frombuffer(data, dtype=np.uint8)
    cap = int((buf == 10).sum()) + 1
    starts = np.zeros(cap, dtype=np.int64)
    ends = np.zeros(cap, dtype=np.int64)
    n = lib().dxo_split(buf.ctypes.data_as(C.c_void_p), C.c_int64(buf.size), starts.ctypes.data_as(C.c_void_p),
                        ends.ctypes.data_as(C.c_void_p), C.c_int64(cap))
    return starts[:n].copy(), ends[:n].copy()


def decode(schema: Schema, data: bytes, offsets=None, threads=0):
    """from_json over a batch.  Returns dict leaf -> (values, valid) plus '__rowok'.
    Strings come back as object arrays of bytes (None when NULL)."""
    buf = np.frombuffer(data, dtype=np.uint8)
    if offsets is None:
        starts, ends = split_ndjson(data)
    else:
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        starts, ends = offsets[:-1].copy(), offsets[1:].copy()
    n = len(starts)
    nl = len(schema.leaf_names)
    vals = np.zeros((nl, max(n, 1)), dtype=np.int64)
    lens = np.zeros((nl, max(n, 1)), dtype=np.int32)
    valid = np.zeros((nl, max(n, 1)), dtype=np.uint8)
    rowok = np.zeros(max(n, 1), dtype=np.uint8)
    heap = np.zeros(buf.size + 16, dtype=np.uint8)
    used = C.c_int64(0)
    vp = C.c_void_p
    if threads and threads > 0:
        # threaded baseline decode: string leaves come back as the FNV-1a hash of their bytes (see str_hash)
        is_str = np.array([1 if t == STRING else 0 for t in schema.leaf_types], dtype=np.uint8)
        rc = lib().dxo_decode_mt_h(buf.ctypes.data_as(vp), starts.ctypes.data_as(vp), ends.ctypes.data_as(vp),
                                   C.c_int64(n), schema.c_structs, schema.c_fields, C.c_int32(nl),
                                   vals.ctypes.data_as(vp), lens.ctypes.data_as(vp), valid.ctypes.data_as(vp),
                                   rowok.ctypes.data_as(vp), C.c_int(threads), is_str.ctypes.data_as(vp))
    else:
        rc = lib().dxo_decode(buf.ctypes.data_as(vp), starts.ctypes.data_as(vp), ends.ctypes.data_as(vp),
                              C.c_int64(n), schema.c_structs, schema.c_fields, C.c_int32(nl),
                              vals.ctypes.data_as(vp), lens.ctypes.data_as(vp), valid.ctypes.data_as(vp),
                              rowok.ctypes.data_as(vp), heap.ctypes.data_as(vp), C.c_int64(heap.size), C.byref(used))
    if rc < 0:
        raise RuntimeError(f"oracle limitation hit (code {rc}): -2 heap")
    global last_limited
    last_limited = int((rowok[:n] == 2).sum())       # rows made NULL by a limit of the restatement (nested value under a string field)
    out = {"__rowok": rowok[:n] == 1}
    for k, name in enumerate(schema.leaf_names):
        v = valid[k, :n].astype(bool)
        t = schema.leaf_types[k]
        if t == DOUBLE:
            out[name] = (vals[k, :n].view(np.float64).copy(), v)
        elif t == STRING and not threads:
            hb = heap.tobytes()
            s = np.empty(n, dtype=object)
            for r in range(n):
                s[r] = hb[vals[k, r]: vals[k, r] + lens[k, r]] if v[r] else None
            out[name] = (s, v)
        else:
            out[name] = (vals[k, :n].copy(), v)
    return out


def str_hash(b: bytes) -> int:
    """FNV-1a 64 of a byte string as a signed int64: the string identity the threaded baseline decoder reports."""
    h = 0xcbf29ce484222325
    for c in b:
        h = ((h ^ c) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
    return h - (1 << 64) if h >= (1 << 63) else h


TYPES = [b"DoorLock", b"WindowLock", b"Heating", b"GarageDoorLock"]


def gen_events(p: GenParams, first, n):
    cap = n * 200 + 64
    out = np.zeros(cap, dtype=np.uint8)
    used = lib().dxo_gen_events(C.byref(p), C.c_int64(first), C.c_int64(n), out.ctypes.data_as(C.c_void_p),
                                C.c_int64(cap))
    assert used >= 0
    return out[:used].tobytes()


def gen_fields(p: GenParams, first, n):
    a = {k: np.zeros(n, dtype=np.int64) for k in ("deviceId", "eventTime", "homeId", "status")}
    ty = np.zeros(n, dtype=np.int32)
    temp = np.zeros(n, dtype=np.float64)
    hum = np.zeros(n, dtype=np.float64)
    vp = C.c_void_p
    lib().dxo_gen_fields(C.byref(p), C.c_int64(first), C.c_int64(n), a["deviceId"].ctypes.data_as(vp),
                         ty.ctypes.data_as(vp), a["eventTime"].ctypes.data_as(vp), a["homeId"].ctypes.data_as(vp),
                         a["status"].ctypes.data_as(vp), temp.ctypes.data_as(vp), hum.ctypes.data_as(vp))
    a["deviceType"] = ty
    a["temperature"] = temp
    a["humidity"] = hum
    return a


def iot_schema_json(shape=2, event_time_type="long"):
    """Spark StructType JSON of the home-automation event (HomeAutomationLocal.json:288), plus the
    two double fields of shape S2 (BasicLocal.json:185)."""
    f = [{"name": "deviceId", "type": "long", "nullable": True, "metadata": {}},
         {"name": "deviceType", "type": "string", "nullable": True, "metadata": {}},
         {"name": "eventTime", "type": event_time_type, "nullable": True, "metadata": {}},
         {"name": "homeId", "type": "long", "nullable": True, "metadata": {}},
         {"name": "status", "type": "long", "nullable": True, "metadata": {}}]
    if shape == 2:
        f += [{"name": "temperature", "type": "double", "nullable": True, "metadata": {}},
              {"name": "humidity", "type": "double", "nullable": True, "metadata": {}}]
    return json.dumps({"type": "struct", "fields": [
        {"name": "deviceDetails", "type": {"type": "struct", "fields": f}, "nullable": True, "metadata": {}}]})


# --------------------------------------------------------------------------------------------------
# threaded rule evaluation (dxo_rules.c) — CPU baseline; checked against dxo_flow's numpy evaluator
# --------------------------------------------------------------------------------------------------
class _ROp(C.Structure):
    _fields_ = [("op", C.c_int32), ("arg", C.c_int32), ("i", C.c_int64), ("f", C.c_double)]


_R = dict(LOAD_I=1, LOAD_F=2, CONST_I=3, CONST_F=4, CONST_NULL=5, I2F_TOP=6, I2F_SECOND=7, CMP_I=8, CMP_F=9, AND=10, OR=11, NOT=12,
          ISNULL=13, ISNOTNULL=14)
_CMP = {"=": 0, "==": 0, "<>": 1, "!=": 1, "<": 2, "<=": 3, ">": 4, ">=": 5}


def compile_rules(conds, columns, str_code):
    """conds: condition trees as dxo_flow.P produces them; columns: [(dotted name, 'long'|'double'|'string')];
    str_code(column index, literal) -> the integer a string column holds for that literal.  -> (ops, rule_start)."""
    names = [n.lower() for n, _ in columns]
    ops, starts = [], [0]

    def emit(op, arg=0, i=0, f=0.0):
        ops.append((_R[op], arg, i, f))

    def walk(e):
        k = e[0]
        if k == "col":
            ci = names.index(".".join(e[1]).lower())
            ty = columns[ci][1]
            emit("LOAD_F" if ty == "double" else "LOAD_I", ci)
            return ty, ci
        if k == "int":
            emit("CONST_I", i=int(e[1])); return "long", None
        if k == "dec":
            emit("CONST_F", f=float(e[1])); return "double", None
        if k == "null":
            emit("CONST_NULL"); return "null", None
        if k == "not":
            walk(e[1]); emit("NOT"); return "bool", None
        if k == "isnull":
            walk(e[1]); emit("ISNOTNULL" if e[2] else "ISNULL"); return "bool", None
        if k == "bin" and e[1] in ("AND", "OR"):
            walk(e[2]); walk(e[3]); emit(e[1]); return "bool", None
        if k == "bin" and e[1] in _CMP:
            a, b = e[2], e[3]
            if a[0] == "str" or b[0] == "str":
                col, lit = (b, a) if a[0] == "str" else (a, b)
                if col[0] != "col" or e[1] not in ("=", "<>", "==", "!="):
                    raise ValueError("string comparison shape outside the baseline evaluator")
                ty, ci = walk(col)
                emit("CONST_I", i=int(str_code(ci, lit[1])))
                emit("CMP_I", _CMP[e[1]]); return "bool", None
            ta, _ = walk(a)
            tb, _ = walk(b)
            if "double" in (ta, tb):
                if ta != "double":
                    emit("I2F_SECOND")
                if tb != "double":
                    emit("I2F_TOP")
                emit("CMP_F", _CMP[e[1]])
            else:
                emit("CMP_I", _CMP[e[1]])
            return "bool", None
        raise ValueError("expression outside the baseline evaluator: %r" % (e,))

    for c in conds:
        walk(c)
        starts.append(len(ops))
    arr = (_ROp * max(1, len(ops)))()
    for j, (op, arg, i, f) in enumerate(ops):
        arr[j].op, arr[j].arg, arr[j].i, arr[j].f = op, arg, i, f
    return arr, np.asarray(starts, dtype=np.int32)


def eval_rules_mt(program, column_arrays, valid_arrays, n_rows, n_rules, threads=0):
    """-> uint32 flags (n_rows, ceil(n_rules/32)), bit r of a row set iff rule r is TRUE."""
    ops, starts = program
    words = (n_rules + 31) // 32
    flags = np.empty((max(n_rows, 1), words), dtype=np.uint32)
    k = len(column_arrays)
    keep = [np.ascontiguousarray(a) for a in column_arrays]
    vkeep = [None if v is None else np.ascontiguousarray(v.astype(np.uint8)) for v in valid_arrays]
    cols = (C.c_void_p * k)(*[a.ctypes.data for a in keep])
    valid = (C.c_void_p * k)(*[None if v is None else v.ctypes.data for v in vkeep])
    fn = lib().dxo_eval_rules
    fn.restype = C.c_int
    rc = fn(ops, starts.ctypes.data_as(C.c_void_p), C.c_int(n_rules), cols, valid, C.c_int64(n_rows), flags.ctypes.data_as(C.c_void_p),
            C.c_int(words), C.c_int(threads or (os.cpu_count() or 1)))
    if rc:
        raise ValueError("dxo_eval_rules: malformed program")
    return flags[:n_rows]
