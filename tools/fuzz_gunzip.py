"""Robustness fuzz of K7 (test infrastructure, run by hand on the GPU box): gzip members with one to three damaged bytes, a cut, or a
forged ISIZE, each inflated next to intact members; the call must fail with a message, or succeed with exactly the bytes zlib's raw
inflate gives for the same stream — and the intact members of a later call must still inflate.  python tools/fuzz_gunzip.py [seconds] [seed]"""
import gzip
import io
import os
import random
import sys
import time
import zlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import dxo                                       # noqa: E402
import dxgpu as dx                               # noqa: E402


def member(data, level):
    buf = io.BytesIO()
    with gzip.GzipFile(filename="", mode="wb", fileobj=buf, compresslevel=level, mtime=0) as f:
        f.write(data)
    return buf.getvalue()


def raw_inflate(m):
    """what the deflate stream of member m decodes to, or None (header without optional fields: 10 bytes)"""
    try:
        d = zlib.decompressobj(-15)
        out = d.decompress(m[10:-8]) + d.flush()
        return out if d.eof else None
    except zlib.error:
        return None


def main():
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 25
    rng = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    ctx = dx.Context(0)
    plain = []
    for k in range(6):
        p = dxo.GenParams(20260200 + k, 5000, 500, 1_700_000_040_000, 60_000, 60_000, 2, 10, 1, 0)
        plain.append(dxo.gen_events(p, 0, 150 + 40 * k))
    plain += [bytes(rng.randrange(256) for _ in range(3000)), b"ab" * 4000, b""]
    good = [member(d, lv) for d in plain for lv in (0, 1, 6, 9)]
    want_good = [d for d in plain for lv in (0, 1, 6, 9)]
    t0, calls, failed, same, differ = time.time(), 0, 0, 0, 0
    while time.time() - t0 < seconds:
        k = rng.randrange(len(good))
        m = bytearray(good[k])
        how = rng.randrange(5)
        if how == 0:
            for _ in range(rng.randrange(1, 4)):
                m[rng.randrange(10, len(m))] = rng.randrange(256)
        elif how == 1:
            m[rng.randrange(10, len(m))] ^= 1 << rng.randrange(8)
        elif how == 2:
            cut = rng.randrange(10, len(m) - 8); m = m[:cut] + m[-8:]
        elif how == 3:
            m[-4:] = rng.randrange(0, 1 << 20).to_bytes(4, "little")            # forged ISIZE
        else:
            a = rng.randrange(10, len(m) - 8); m[a:a] = bytes(rng.randrange(256) for _ in range(rng.randrange(1, 9)))
        m = bytes(m)
        j = rng.randrange(len(good))
        calls += 1
        try:
            ptr, n = ctx.gunzip([good[j], m, good[j]])
        except dx.DxError:
            failed += 1
            continue
        got = ctx.d2h(ptr, n).tobytes() if n else b""
        ctx.free(ptr)
        ref = raw_inflate(m)
        isize = int.from_bytes(m[-4:], "little")
        if ref is not None and len(ref) == isize and got == want_good[j] + ref + want_good[j]:
            same += 1
        else:
            differ += 1
            print("DIFFERS: how", how, "member", k, "len", len(m), "isize", isize, "zlib", None if ref is None else len(ref), "got", n, flush=True)
    ptr, n = ctx.gunzip(good)
    ok = ctx.d2h(ptr, n).tobytes() == b"".join(want_good)
    ctx.free(ptr)
    print("calls %d: refused %d, inflated and equal to zlib %d, inflated but different %d; intact members afterwards: %s" % (calls, failed, same, differ, "ok" if ok else "WRONG"))
    ctx.close()
    return 0 if differ == 0 and ok else 1


if __name__ == "__main__":
    sys.exit(main())
