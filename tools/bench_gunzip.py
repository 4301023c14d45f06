#!/usr/bin/env python3
"""K7 timing: gzip members (blob files of S2 JSON lines) inflated on the device, one warp per member.
  python tools/bench_gunzip.py [members] [events_per_member] [level]"""
import gzip, io, json, os, sys, time, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def run(dx, ctx, members=512, events=20_000, level=6, reps=5):
    import ctypes as C
    plain = []
    for k in range(8):      # eight distinct files, repeated: the decoder does not care, the host prep stays short
        p = dx.GenParams(20260200 + k, 1_000_000, 100_000, 1_700_000_040_000, 60_000, 60_000, 2, 10, 1, 0)
        ptr, nb, _ = ctx.gen_events(p, k * events, events)          # the library's own generator; compressed on the host below
        plain.append(ctx.d2h(ptr, nb).tobytes())
        ctx.free(ptr)
    gz = []
    for k in range(8):
        buf = io.BytesIO()
        with gzip.GzipFile(mode="wb", fileobj=buf, compresslevel=level, mtime=0) as f:
            f.write(plain[k])
        gz.append(buf.getvalue())
    mem = [gz[k % 8] for k in range(members)]
    blob = np.frombuffer(b"".join(mem), dtype=np.uint8)
    off = np.zeros(members + 1, dtype=np.int64); off[1:] = np.cumsum([len(m) for m in mem])
    dptr = ctx.alloc(blob.size + 64); ctx.h2d(dptr, blob)
    fn = ctx.L.dx_gunzip_device
    fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    ctx.set_timing(True)
    n_out = 0
    for i in range(2 + reps):
        if i == 2:
            ctx.take_stage_ms("k7_inflate")
        p, n = C.c_void_p(), C.c_int64()
        ctx.check(fn(ctx.h, C.c_void_p(dptr), off.ctypes.data_as(C.c_void_p), members, 1, C.byref(p), C.byref(n)))
        n_out = n.value
        ctx.free(p.value)
    ms, cnt = ctx.take_stage_ms("k7_inflate")
    ctx.set_timing(False)
    ctx.free(dptr)
    ms /= max(cnt, 1)
    t0 = time.perf_counter(); zlib.decompress(gz[0], 31); cpu = time.perf_counter() - t0
    return {"members": members, "compressed_bytes": int(blob.size), "json_bytes": int(n_out), "ratio": round(n_out / blob.size, 2), "level": level,
            "k7_inflate_ms": round(ms, 3), "out_gbs": round(n_out / ms / 1e6, 2), "in_gbs": round(blob.size / ms / 1e6, 2),
            "alg_gbs": round((n_out + blob.size) / ms / 1e6, 2), "events_per_s": round(members * events / ms * 1e3, 1),
            "cpu_zlib_one_thread_out_gbs": round(len(plain[0]) / cpu / 1e9, 3)}


if __name__ == "__main__":
    import dxgpu as dx
    ctx = dx.Context(0)
    a = [int(x) for x in sys.argv[1:]]
    print(json.dumps(run(dx, ctx, *(a + [512, 20_000, 6][len(a):]))))
