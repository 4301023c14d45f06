"""K7: gzip members inflated on the device (dx_gunzip_device, the blob input's compressed files) against zlib."""
import gzip
import io
import os
import zlib

import numpy as np
import pytest

import dxo

pytestmark = pytest.mark.gpu


def member(data, level=6, name=None, mtime=0):
    buf = io.BytesIO()
    with gzip.GzipFile(filename=name or "", mode="wb", fileobj=buf, compresslevel=level, mtime=mtime) as f:
        f.write(data)
    return buf.getvalue()


def events(n, seed=1):
    p = dxo.GenParams(20260120 + seed, 5000, 500, 1_700_000_040_000, 60_000, 60_000, 2, 10, 1, 0)
    return dxo.gen_events(p, 0, n)


def test_members_of_every_kind_inflate_to_what_zlib_gives(dx, ctx):
    rnd = np.random.default_rng(5)
    plain = [events(20_000, 1), events(3, 2), b"", b"x", events(60_000, 3), rnd.integers(0, 256, 300_000, dtype=np.uint8).tobytes(),
             b"ab" * 70_000, bytes(range(256)) * 40, events(5_000, 4)]
    members = [member(plain[0], 6), member(plain[1], 9, name="part-00000.json"), member(plain[2], 6), member(plain[3], 1),
               member(plain[4], 1), member(plain[5], 6), member(plain[6], 9), member(plain[7], 0), member(plain[8], 4, name="blob name.json", mtime=1700000000)]
    # a member made of fixed-Huffman blocks only (zlib strategy Z_FIXED) and one with many small dynamic blocks (full flushes)
    co = zlib.compressobj(6, zlib.DEFLATED, 31, 8, zlib.Z_FIXED)
    fixed_src = events(2_000, 5)
    members.append(co.compress(fixed_src) + co.flush())
    plain.append(fixed_src)
    co = zlib.compressobj(6, zlib.DEFLATED, 31)
    chunks = events(30_000, 6)
    out = b""
    for k in range(0, len(chunks), 50_000):
        out += co.compress(chunks[k:k + 50_000]) + co.flush(zlib.Z_FULL_FLUSH)
    members.append(out + co.flush())
    plain.append(chunks)
    ptr, n = ctx.gunzip(members)
    want = b"".join(plain)
    assert n == len(want)
    got = ctx.d2h(ptr, n).tobytes()
    assert got == want
    ctx.free(ptr)


def test_inflated_blob_files_feed_the_decoder(dx, ctx):
    files = [events(40_000, k) for k in range(6)]
    ptr, n = ctx.gunzip([member(f, 6) for f in files])
    dec = dx.Decoder(ctx, dxo.iot_schema_json(2))
    t = dec.decode_device(ptr, n, n_events=6 * 40_000)
    assert t.num_rows == 240_000 and dec.stats()["null_rows"] == 0
    ids, ok = t.column("deviceDetails.deviceId")
    ref = dxo.decode(dxo.Schema(dxo.iot_schema_json(2)), b"".join(files))
    assert ok.all() and (ids == ref["deviceDetails.deviceId"][0]).all()
    t.release(); dec.close(); ctx.free(ptr)


def test_damaged_members_fail_the_call(dx, ctx):
    good = member(events(5_000, 9), 6)
    with pytest.raises(dx.DxError):
        ctx.gunzip([b"PK\x03\x04" + good[4:]])                       # not gzip
    broken = bytearray(good)
    broken[len(broken) // 2] ^= 0x55                                 # deflate data damaged: bad code, bad distance or wrong length
    broken[len(broken) // 2 + 1] ^= 0xAA
    with pytest.raises(dx.DxError):
        ctx.gunzip([bytes(broken)])
    trunc = good[:-8][: len(good) // 2] + good[-8:]                  # truncated stream with an intact trailer
    with pytest.raises(dx.DxError):
        ctx.gunzip([trunc])
    ptr, n = ctx.gunzip([good])                                      # and the context is still usable
    assert n == len(events(5_000, 9))
    ctx.free(ptr)
