"""The state tables' on-disk format (dx_parquet_write / dx_parquet_read) against pyarrow: what this library writes
must read back in pyarrow exactly (it stands in for Spark's reader: same Parquet format), and what a Spark-2.4-style writer produces
(snappy pages, dictionary encoding, INT96 timestamps, several row groups and part files) must read back here exactly.  No GPU."""
import ctypes as C
import os

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

import dxgpu as dx



def lib():
    from importlib import import_module
    N = import_module("data-accelerator_b200._native")
    Lb = N.lib()
    Lb.dx_parquet_write.restype = C.c_int
    Lb.dx_parquet_write.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int), C.c_int64, C.POINTER(C.c_void_p),
                                    C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_char_p, C.c_int64]
    Lb.dx_parquet_read.restype, Lb.dx_parquet_read.argtypes = C.c_int, [C.c_char_p, C.POINTER(C.c_void_p), C.c_char_p, C.c_int64]
    Lb.dx_parquet_num_rows.restype, Lb.dx_parquet_num_rows.argtypes = C.c_int64, [C.c_void_p]
    Lb.dx_parquet_num_columns.restype, Lb.dx_parquet_num_columns.argtypes = C.c_int, [C.c_void_p]
    Lb.dx_parquet_column_name.restype, Lb.dx_parquet_column_name.argtypes = C.c_char_p, [C.c_void_p, C.c_int]
    Lb.dx_parquet_column_type.restype, Lb.dx_parquet_column_type.argtypes = C.c_int, [C.c_void_p, C.c_int]
    Lb.dx_parquet_column.restype = C.c_int
    Lb.dx_parquet_column.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    Lb.dx_parquet_release.restype, Lb.dx_parquet_release.argtypes = None, [C.c_void_p]
    return Lb


def sample(n, seed=3):
    rnd = np.random.default_rng(seed)
    ids = rnd.integers(-10**12, 10**12, n)
    vals = rnd.normal(0, 1e6, n)
    vals[::17] = np.nan
    flags = rnd.integers(0, 2, n)
    ts = rnd.integers(-10**15, 2 * 10**15, n)           # microseconds, before and after the epoch
    words = [("w%d" % (i % 37)).encode() if i % 5 else ("é\n\"%d" % i).encode() for i in range(n)]
    valid = {k: rnd.integers(0, 10, n) > 0 for k in ("id", "v", "f", "t", "s")}
    return ids, vals, flags, ts, words, valid


def write_ours(Lb, path, ids, vals, flags, ts, words, valid):
    n = len(ids)
    names = [b"deviceId", b"Reading", b"Flag", b"EventTime", b"deviceType"]
    types = [1, 2, 4, 5, 3]
    off = np.zeros(n + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(w) for w in words])
    blob = np.frombuffer(b"".join(words) + b"\0", dtype=np.uint8)
    cols = [ids.astype(np.int64), vals.astype(np.float64), flags.astype(np.int64), ts.astype(np.int64), off]
    vs = [valid[k].astype(np.uint8) for k in ("id", "v", "f", "t", "s")]
    data = (C.c_void_p * 5)(*[c.ctypes.data for c in cols])
    vptr = (C.c_void_p * 5)(*[v.ctypes.data for v in vs])
    sptr = (C.c_void_p * 5)(None, None, None, None, blob.ctypes.data)
    err = C.create_string_buffer(512)
    rc = Lb.dx_parquet_write(path.encode(), 5, (C.c_char_p * 5)(*names), (C.c_int * 5)(*types), n, data, vptr, sptr, err, 512)
    assert rc == 0, err.value


def read_ours(Lb, path):
    h = C.c_void_p()
    err = C.create_string_buffer(512)
    rc = Lb.dx_parquet_read(path.encode(), C.byref(h), err, 512)
    assert rc == 0, err.value
    n, out = Lb.dx_parquet_num_rows(h), {}
    for c in range(Lb.dx_parquet_num_columns(h)):
        name, ty = Lb.dx_parquet_column_name(h, c).decode(), Lb.dx_parquet_column_type(h, c)
        d, v, s = C.c_void_p(), C.c_void_p(), C.c_void_p()
        assert Lb.dx_parquet_column(h, c, C.byref(d), C.byref(v), C.byref(s)) == 0
        valid = np.ctypeslib.as_array(C.cast(v, C.POINTER(C.c_uint8)), (n,)).astype(bool).copy() if n else np.zeros(0, bool)
        if ty == 3:
            off = np.ctypeslib.as_array(C.cast(d, C.POINTER(C.c_int64)), (n + 1,)).copy()
            raw = C.string_at(s.value, int(off[-1])) if off[-1] else b""
            vals = [raw[off[i]:off[i + 1]] if valid[i] else None for i in range(n)]
        else:
            arr = np.ctypeslib.as_array(C.cast(d, C.POINTER(C.c_int64)), (n,)).copy() if n else np.zeros(0, np.int64)
            vals = [(arr[i:i + 1].view(np.float64)[0] if ty == 2 else int(arr[i])) if valid[i] else None for i in range(n)]
        out[name] = (ty, vals)
    Lb.dx_parquet_release(h)
    return n, out


def same(a, b):
    if a is None or b is None:
        return a is None and b is None
    if isinstance(a, float) and isinstance(b, float):
        return (np.isnan(a) and np.isnan(b)) or a == b
    return a == b


def test_what_we_write_pyarrow_reads_back_exactly(tmp_path):
    Lb = lib()
    ids, vals, flags, ts, words, valid = sample(5000)
    path = str(tmp_path / "A" / "part-00000.parquet")
    write_ours(Lb, path, ids, vals, flags, ts, words, valid)
    t = pq.read_table(path)
    assert t.column_names == ["deviceId", "Reading", "Flag", "EventTime", "deviceType"] and t.num_rows == 5000
    got = t.to_pydict()
    epoch = np.datetime64(0, "us")
    for i in range(5000):
        assert same(got["deviceId"][i], int(ids[i]) if valid["id"][i] else None)
        assert same(got["Reading"][i], float(vals[i]) if valid["v"][i] else None)
        assert same(got["Flag"][i], bool(flags[i]) if valid["f"][i] else None)
        assert same(got["deviceType"][i], words[i].decode() if valid["s"][i] else None)
    tcol = t.column("EventTime").cast(pa.timestamp("us")).cast(pa.int64()).to_pylist()       # INT96 comes back as a timestamp
    assert all(same(tcol[i], int(ts[i]) if valid["t"][i] else None) for i in range(5000))
    # and our own reader gives the same
    n, ours = read_ours(Lb, path)
    assert n == 5000 and [same(x, int(ids[i]) if valid["id"][i] else None) for i, x in enumerate(ours["deviceId"][1])].count(False) == 0
    assert ours["EventTime"][0] == 5 and all(same(x, int(ts[i]) if valid["t"][i] else None) for i, x in enumerate(ours["EventTime"][1]))
    assert all(same(x, words[i] if valid["s"][i] else None) for i, x in enumerate(ours["deviceType"][1]))


@pytest.mark.parametrize("compression,dictionary,row_group", [("snappy", True, 700), ("none", False, 100000), ("snappy", ["deviceType"], 1500)])
def test_spark_style_files_read_back_exactly(tmp_path, compression, dictionary, row_group):
    Lb = lib()
    ids, vals, flags, ts, words, valid = sample(6000, seed=11)

    def arr(x, ok, ty):
        return pa.array([x[i] if ok[i] else None for i in range(len(x))], type=ty)
    ts_ok = np.where((ts > -6 * 10**13) & (ts < 4 * 10**15), ts, 0)               # INT96 via pyarrow needs nanosecond-representable values
    table = pa.table({"deviceId": arr(ids.tolist(), valid["id"], pa.int64()), "Reading": arr(vals.tolist(), valid["v"], pa.float64()),
                      "Flag": arr([bool(f) for f in flags], valid["f"], pa.bool_()),
                      "EventTime": arr(ts_ok.tolist(), valid["t"], pa.timestamp("us")),
                      "deviceType": arr([w.decode() for w in words], valid["s"], pa.string())})
    d = tmp_path / "B"
    os.makedirs(d)
    half = 3500
    for k, (a, b) in enumerate([(0, half), (half, 6000)]):          # two part files, as a two-task Spark job leaves them
        pq.write_table(table.slice(a, b - a), str(d / ("part-%05d-c000.snappy.parquet" % k)), compression=compression, use_dictionary=dictionary,
                       row_group_size=row_group, use_deprecated_int96_timestamps=True, flavor="spark", data_page_version="1.0")
    open(d / "_SUCCESS", "w").close()                                # Hadoop's marker files are skipped
    open(d / ".part-00000.crc", "wb").write(b"crc")
    n, ours = read_ours(Lb, str(d))
    assert n == 6000 and list(ours) == ["deviceId", "Reading", "Flag", "EventTime", "deviceType"]
    assert [t for t, _ in ours.values()] == [1, 2, 4, 5, 3]
    assert all(same(x, int(ids[i]) if valid["id"][i] else None) for i, x in enumerate(ours["deviceId"][1]))
    assert all(same(x, float(vals[i]) if valid["v"][i] else None) for i, x in enumerate(ours["Reading"][1]))
    assert all(same(x, int(flags[i]) if valid["f"][i] else None) for i, x in enumerate(ours["Flag"][1]))
    assert all(same(x, int(ts_ok[i]) if valid["t"][i] else None) for i, x in enumerate(ours["EventTime"][1]))
    assert all(same(x, words[i] if valid["s"][i] else None) for i, x in enumerate(ours["deviceType"][1]))


def test_bad_files_are_errors(tmp_path):
    Lb = lib()
    p = tmp_path / "x.parquet"
    p.write_bytes(b"PAR1 this is not parquet PAR1")
    h, err = C.c_void_p(), C.create_string_buffer(256)
    assert Lb.dx_parquet_read(str(p).encode(), C.byref(h), err, 256) != 0 and err.value
    assert Lb.dx_parquet_read(str(tmp_path / "missing").encode(), C.byref(h), err, 256) != 0
    t = pa.table({"a": pa.array([[1, 2], [3]], type=pa.list_(pa.int64()))})
    pq.write_table(t, str(tmp_path / "nested.parquet"))
    assert Lb.dx_parquet_read(str(tmp_path / "nested.parquet").encode(), C.byref(h), err, 256) != 0 and b"flat" in err.value
