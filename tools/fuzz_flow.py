"""Robustness fuzz of the flow compiler (test infrastructure, run by hand on the GPU box): the sample flow's transform and projection
with token-level edits (dropped, swapped, duplicated and foreign tokens); every variant must either be refused by createProcessor
with a message or run a 120-event batch — never crash.  The text under test is written to gpurun_out/fuzz_flow_last.txt before each
attempt.  python tools/fuzz_flow.py [seconds] [seed]"""
import json
import os
import random
import re
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dxgpu as dx                               # noqa: E402

C1 = os.path.join(ROOT, "tests", "golden", "c1_homeautomation")
FOREIGN = ["(", ")", ",", "*", "=", "<>", ">", "AND", "OR", "NOT", "NULL", "IN", "BETWEEN", "CASE", "WHEN", "THEN", "ELSE", "END", "AS", "FROM", "WHERE",
           "GROUP", "BY", "JOIN", "ON", "LEFT", "UNION", "ALL", "SELECT", "DISTINCT", "HAVING", "COUNT", "SUM", "MIN", "MAX", "AVG", "MAP", "STRUCT",
           "Array", "filterNull", "IF", "CAST", "timestamp", "string", "long", "double", "1", "0", "-1", "1.5", "'x'", "''", "`a.b`", "Raw", "Raw.*",
           "deviceDetails.deviceId", "deviceDetails.status", "eventTimeStamp", "DataXProcessedInput", "DataXProcessedInput_5minutes", "myDevicesRefData",
           "whoOpened", "current_timestamp()", "unix_timestamp()", "DATE_TRUNC", "CONCAT", "TIMEWINDOW('5 minutes')", "--DataXQuery--", "\n", ".", "/", "%", "||"]


def mutate(rng, text):
    toks = re.findall(r"--DataXQuery--|[A-Za-z_][A-Za-z_0-9.]*|\d+\.?\d*|'[^']*'|`[^`]*`|\s+|.", text)
    for _ in range(rng.randrange(1, 4)):
        k, i = rng.randrange(5), rng.randrange(len(toks))
        if k == 0:
            del toks[i]
        elif k == 1:
            toks.insert(i, rng.choice(FOREIGN) + " ")
        elif k == 2:
            toks[i] = rng.choice(FOREIGN)
        elif k == 3:
            j = rng.randrange(len(toks)); toks[i], toks[j] = toks[j], toks[i]
        else:
            j = rng.randrange(len(toks)); toks[i:i] = toks[j:j + rng.randrange(1, 12)]
    return "".join(toks)


def main():
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 60
    rng = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    rd = lambda f: open(os.path.join(C1, f), encoding="utf-8-sig").read()
    conf, schema, proj, tr, devices = rd("HomeAutomationLocal.conf"), rd("inputschema.json"), rd("projection.txt"), rd("HomeAutomationLocal-combined.txt"), rd("devices.csv")
    ctx = dx.Context(0)
    now = 1_700_000_041_500
    ev = ("\n".join(json.dumps({"deviceDetails": {"deviceId": rng.choice([1, 2, 3, 4, 5, 6]), "deviceType": rng.choice(["DoorLock", "WindowLock", "Heating"]),
                                                  "eventTime": now - rng.randrange(1000), "homeId": rng.choice([32, 150, 25, 81]), "status": rng.choice([0, 1])}},
                               separators=(",", ":")) for _ in range(120)) + "\n").encode()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    last = os.path.join(ROOT, "gpurun_out", "fuzz_flow_last.txt")
    t0, n, created, ran, msgs = time.time(), 0, 0, 0, {}
    while time.time() - t0 < seconds:
        which = rng.randrange(4)
        p, t = (mutate(rng, proj), tr) if which == 0 else (proj, mutate(rng, tr))
        with open(last, "w") as f:
            f.write("==== projection ====\n" + p + "\n==== transform ====\n" + t + "\n"); f.flush(); os.fsync(f.fileno())
        n += 1
        try:
            g = dx.CommonProcessor(ctx, conf, schema, [p], t)
        except dx.DxError as e:
            key = re.sub(r"'[^']*'|\d+", "#", str(e))[:60]; msgs[key] = msgs.get(key, 0) + 1
            continue
        created += 1
        try:
            g.add_refdata("myDevicesRefData", devices, ",", True)
            g.register_udf("whoOpened", lambda s: "hello, " + str(s))
            r = g.process_json(ev, 1_700_000_040_000, 60_000, now_ms=now)
            for name in r.table_names:
                r.table(name).num_rows
            r.release(); ran += 1
        except dx.DxError as e:
            key = "run: " + re.sub(r"'[^']*'|\d+", "#", str(e))[:60]; msgs[key] = msgs.get(key, 0) + 1
        g.close()
    os.remove(last)
    print("variants %d, created %d, ran a batch %d" % (n, created, ran))
    for k, v in sorted(msgs.items(), key=lambda kv: -kv[1])[:25]:
        print("%6d  %s" % (v, k))
    ctx.close()


if __name__ == "__main__":
    main()
