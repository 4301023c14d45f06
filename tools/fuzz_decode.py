"""Differential fuzz of K1 against the CPU oracle (test infrastructure, run by hand on the GPU box): records of the S2 shape with one
to three byte-level edits each (structure characters, escapes, digits, exponents, whitespace, control and high bytes, cuts,
duplicated spans), decoded by both; every row's validity and value must agree.  python tools/fuzz_decode.py [records] [rounds] [seed]"""
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import dxo                                       # noqa: E402
import dxgpu as dx                               # noqa: E402
from adversarial import GOOD, random_number      # noqa: E402

ALPHA = [b'"', b"\\", b"{", b"}", b"[", b"]", b":", b",", b" ", b"\t", b"\r", b"0", b"9", b"-", b"+", b".", b"e", b"E", b"n", b"t", b"f",
         b"null", b"true", b"false", b"\\u0041", b"\\u00e9", b"\\ud83d\\ude00", b"\\n", b"\\/", b"\\x", b"\x01", b"\x1f", b"\x7f", b"\xc3\xa9",
         b"\xe2\x82\xac", b"\xff", b"\xc3", b"1e5", b"1E-3", b"0.0", b"-0", b"00", b"1.", b".5", b"Infinity", b"NaN", b"9223372036854775808",
         b"-9223372036854775809", b"1" * 25, b'"a":1,', b'{"a":[]}', b"[1,2]", b'""']


def mutate(rng, rec):
    b = bytearray(rec)
    for _ in range(rng.randrange(1, 4)):
        k, pos = rng.randrange(6), rng.randrange(len(b) + 1)
        tok = rng.choice(ALPHA)
        if k == 0:
            b[pos:pos] = tok                                     # insert
        elif k == 1:
            b[pos:pos + len(tok)] = tok                          # overwrite
        elif k == 2:
            del b[pos:pos + rng.randrange(1, 6)]                 # delete
        elif k == 3:
            b = b[:pos]                                          # cut
        elif k == 4:
            e = rng.randrange(len(b) + 1); b[pos:pos] = b[min(pos, e):max(pos, e)][:40]      # duplicate a span
        else:
            if len(b): b[rng.randrange(len(b))] = rng.randrange(1, 256)
        if not b:
            b = bytearray(b"{}")
    return bytes(b).replace(b"\n", b" ")


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    seed0 = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    schema_json = dxo.iot_schema_json(2)
    S = dxo.Schema(schema_json)
    ctx = dx.Context(0)
    dec = dx.Decoder(ctx, schema_json)
    types = ["DoorLock", "WindowLock", "Heating", "GarageDoorLock", "Sp ace", "quo\\\"te", "ünï", "", "x" * 40]
    bad_total = 0
    for rd in range(rounds):
        rng = random.Random(seed0 * 1000 + rd)
        recs = []
        for _ in range(n):
            rec = (GOOD % (rng.randrange(1, 10**6), rng.choice(types), 1600000000000 + rng.randrange(10**9), rng.randrange(1, 10**5),
                           rng.randrange(2), random_number(rng), random_number(rng))).encode("utf-8")
            recs.append(mutate(rng, rec) if rng.randrange(10) else rec)
        data = b"\n".join(recs) + b"\n"
        t = dec.decode_host(data)
        ora = dxo.decode(S, data)
        assert t.num_rows == len(ora["__rowok"]) == n, (t.num_rows, len(ora["__rowok"]), n)
        bad = np.zeros(n, bool)
        for k, name in enumerate(S.leaf_names):
            ev, eok = ora[name]
            gv, gok = t.column(name)
            bad |= gok != eok
            both = gok & eok
            ty = S.leaf_types[k]
            if ty == dxo.DOUBLE:
                a, b = gv.view(np.int64), ev.view(np.int64)
                bad |= both & (a != b) & ~(np.isnan(gv) & np.isnan(ev))
            elif ty == dxo.STRING:
                for i in np.nonzero(both)[0]:
                    if gv[i] != ev[i]:
                        bad[i] = True
            else:
                bad |= both & (gv != ev)
        idx = np.nonzero(bad)[0]
        bad_total += len(idx)
        print("round %d: %d records, %d null rows, %d mismatches" % (rd, n, int((~ora["__rowok"]).sum()), len(idx)), flush=True)
        for i in idx[:15]:
            print("  record %d: %r" % (i, recs[i]))
            for k, name in enumerate(S.leaf_names):
                ev, eok = ora[name]; gv, gok = t.column(name)
                if gok[i] != eok[i] or (eok[i] and gv[i] != ev[i] and not (S.leaf_types[k] == dxo.DOUBLE and np.isnan(gv[i]) and np.isnan(ev[i]))):
                    print("    %s: gpu %r (%s) oracle %r (%s)" % (name, gv[i], gok[i], ev[i], eok[i]))
    dec.close(); ctx.close()
    print("total mismatches:", bad_total)
    return 1 if bad_total else 0


if __name__ == "__main__":
    sys.exit(main())
