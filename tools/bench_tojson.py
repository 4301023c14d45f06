#!/usr/bin/env python3
"""K6 timing: rows of a result table -> NDJSON on the device (dx_table_to_json).
  python tools/bench_tojson.py [events] [alerts|flat]
flat: the seven decoded S2 columns + the projected timestamp of every event (the shape of a pass-through output);
alerts: the rows of an alert output (id, kind, timestamp, a MAP of mixed types)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

FLAT = "--DataXQuery--\nT1 = SELECT deviceDetails.deviceId AS deviceId, deviceDetails.deviceType AS deviceType, eventTimeStamp, deviceDetails.homeId AS homeId, deviceDetails.status AS status, deviceDetails.temperature AS temperature, deviceDetails.humidity AS humidity FROM DataXProcessedInput\n"
ALERTS = "--DataXQuery--\nT1 = SELECT deviceDetails.deviceId AS id, deviceDetails.deviceType AS kind, eventTimeStamp, MAP('deviceId', deviceDetails.deviceId, 'type', deviceDetails.deviceType, 'temperature', deviceDetails.temperature) AS pivots FROM DataXProcessedInput WHERE deviceDetails.temperature > 50\n"


T0, INTERVAL_MS = 1_700_000_040_000, 60_000


def run(dxm, ctx, n=5_000_000, shape="flat", reps=5):
    conf = dxm.flowgen.conf_text("tojson", outputs=["T1"])
    proj = "CAST(Raw.deviceDetails.eventTime / 1000 AS timestamp) AS eventTimeStamp\nRaw.*\n"
    g = dxm.CommonProcessorFactory.createProcessor(ctx, conf, dxm.flowgen.iot_schema_json(2), [proj], FLAT if shape == "flat" else ALERTS)
    gp = dxm.GenParams(20260102, 1_000_000, 100_000, T0, INTERVAL_MS, 60_000, 2, 10, 1, 0)      # the C2 event recipe (S2, 1 M keys)
    ptr, nb, _ = ctx.gen_events(gp, 0, n)
    res = g.process_device(ptr, nb, T0, INTERVAL_MS, n_events=n)
    t = res.table("T1")
    rows = t.num_rows
    col_bytes = sum(4 if t.column_type(i) == 3 else 8 for i in range(len(t.columns)))
    ctx.set_timing(True)
    out = {}
    for i in range(2 + reps):
        if i == 2:
            for k in ("k6_json_len", "k6_json_scan", "k6_json_write"):
                ctx.take_stage_ms(k)
        b, o, nr, nbytes, rel = t.to_json(device=True)
        rel()
    tot = 0.0
    for k in ("k6_json_len", "k6_json_scan", "k6_json_write"):
        ms, cnt = ctx.take_stage_ms(k)
        out[k + "_ms"] = round(ms / max(cnt, 1), 4)
        tot += ms / max(cnt, 1)
    ctx.set_timing(False)
    alg = rows * (2 * col_bytes + 1 + 12) + nbytes          # both passes read the columns; 4 B length + 8 B offset; the JSON bytes once
    out.update({"shape": shape, "rows": rows, "json_bytes": nbytes, "bytes_per_row": round(nbytes / max(rows, 1), 1), "ms": round(tot, 4),
                "rows_per_s": round(rows / tot * 1e3, 1), "alg_bytes": alg, "alg_gbs": round(alg / tot / 1e6, 1)})
    res.release(); g.close()
    return out


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
    shape = sys.argv[2] if len(sys.argv) > 2 else "flat"
    import dxgpu as dx
    ctx = dx.Context(0)
    print(json.dumps(run(dx, ctx, n, shape)))
