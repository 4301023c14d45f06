#!/usr/bin/env python3
"""Auxiliary measurements for BASELINE.json configs[2] (C3: tumbling + sliding window aggregates,
1 M device keys) and configs[3] (C4: state-table merge over consecutive micro-batches).
bench.py stays on the headline configuration (C2); the lines printed here are kept under profiles/.

  python tools/bench_windows.py --workload c3 --batches 100 --events-per-batch 5000000
  python tools/bench_windows.py --workload c4 --batches 100 --events-per-batch 5000000
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

T0, I = 1_700_000_040_000, 60_000
PROJ = "CAST(Raw.deviceDetails.eventTime / 1000 AS timestamp) AS eventTimeStamp\nRaw.*\n"
AGGS = ("COUNT(*) AS Count, SUM(deviceDetails.status) AS SumStatus, MIN(deviceDetails.status) AS MinStatus, "
        "MAX(deviceDetails.status) AS MaxStatus, AVG(deviceDetails.status) AS AvgStatus")


def c3_flow(dx):
    tr = ""
    for name, table in (("Tumbling", "DataXProcessedInput"), ("Sliding5m", "DataXProcessedInput_5minutes"), ("Sliding1h", "DataXProcessedInput_1hour")):
        tr += "--DataXQuery--\n%s = SELECT deviceDetails.deviceId, deviceDetails.homeId, %s FROM %s GROUP BY deviceDetails.deviceId, deviceDetails.homeId\n" % (name, AGGS, table)
    conf = dx.flowgen.conf_text("c3windows", timestamp_column="eventTimeStamp", watermark="60 second",
                                windows=[("DataXProcessedInput_5minutes", "5 minutes"), ("DataXProcessedInput_1hour", "1 hour")])
    return conf, tr


def c4_flow(dx):
    tr = """--DataXQuery--
PerKey = SELECT deviceDetails.deviceId, deviceDetails.homeId, MAX(eventTimeStamp) AS WindowEnd, COUNT(*) AS Count
    FROM DataXProcessedInput GROUP BY deviceDetails.deviceId, deviceDetails.homeId
--DataXQuery--
acc_new = SELECT deviceId, homeId, WindowEnd, Count FROM PerKey
    UNION ALL SELECT deviceId, homeId, WindowEnd, Count FROM acc WHERE unix_timestamp() - to_unix_timestamp(WindowEnd, 'yyyy-MM-dd') < 3600
--DataXQuery--
acc = SELECT deviceId, homeId, WindowEnd, Count FROM acc_new
--DataXQuery--
merged = SELECT deviceId, homeId, Count FROM PerKey UNION ALL SELECT deviceId, homeId, Count FROM totals
--DataXQuery--
totals = SELECT deviceId, homeId, SUM(Count) AS Count FROM merged GROUP BY deviceId, homeId
"""
    conf = dx.flowgen.conf_text("c4state", state_tables=[("acc", "deviceId long, homeId long, WindowEnd Timestamp, Count long"),
                                                         ("totals", "deviceId long, homeId long, Count long")])
    return conf, tr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--batches", type=int, default=100)
    ap.add_argument("--events-per-batch", type=int, default=5_000_000)
    ap.add_argument("--keys", type=int, default=1_000_000)
    ap.add_argument("--no-smem", action="store_true", help="pane partials through the global-memory hash table (gpu.panesmem=false)")
    ap.add_argument("--no-panes", action="store_true", help="rescan every cached batch per window (gpu.panepartials=false)")
    args = ap.parse_args()
    import torch
    import dxgpu as dx
    import dxo
    ctx = dx.Context(0)
    conf, tr = c3_flow(dx) if args.workload == "c3" else c4_flow(dx)
    if args.no_smem:
        conf += "datax.job.process.gpu.panesmem=false\n"
    if args.no_panes:
        conf += "datax.job.process.gpu.panepartials=false\n"
    flow = dx.CommonProcessorFactory.createProcessor(ctx, conf, dxo.iot_schema_json(1), [PROJ], tr)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", 0))
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    n = args.events_per_batch
    ctx.set_timing(True)          # per-kernel CUDA events (adds a host sync per kernel; the wall total below includes it)
    gen_ms = 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    proc_ms = 0.0
    rows_out = {}
    t_wall = time.perf_counter()
    for b in range(args.batches):
        T = T0 + b * I
        gp = dx.GenParams(20260103, args.keys, max(1, args.keys // 10), T, I, 60_000, 1, 10, 1, 0)
        ptr, nb, _ = ctx.gen_events(gp, b * n, n)
        e0.record(stream)
        r = flow.process_device(ptr, nb, T, I, n_events=n, now_ms=T + 5)
        e1.record(stream)
        e1.synchronize()
        proc_ms += e0.elapsed_time(e1)
        if b == args.batches - 1:
            for name in r.table_names:
                rows_out[name] = r.table(name).num_rows
        r.release()
        ctx.free(ptr)
    wall = time.perf_counter() - t_wall
    k3_ms, k3_n = ctx.take_stage_ms("k3_hash_agg")
    k1_ms, _ = ctx.take_stage_ms("k1_decode")
    kg_ms, kg_n = ctx.take_stage_ms("k_gather")
    k3_rows = ctx.take_counter("k3_rows")
    g_bytes = ctx.take_counter("gather_bytes")
    host_us = {k: ctx.take_counter("host_us:" + k) for k in ["project", "route", "k3p_build_scatter", "k3p_build_total_to_agg"] + ["stmt:" + nm for nm in rows_out]}
    total_events = n * args.batches
    row_bytes = 33 if args.workload == "c3" else 8 * 4          # keys + ts + value (+validity) per visited row
    out = {"metric": "events/sec per micro-batch (%s)" % ("window aggregates C3" if args.workload == "c3" else "state-table merge C4"),
           "value": round(total_events / (proc_ms / 1e3), 1), "unit": "events/s", "batches": args.batches, "events_per_batch": n,
           "keys": args.keys, "ms_per_batch": round(proc_ms / args.batches, 3),
           "k3_hash_agg": {"launches": k3_n, "total_ms": round(k3_ms, 2), "rows_visited": k3_rows,
                           "alg_bytes_per_row": row_bytes, "gbs": round(k3_rows * row_bytes / k3_ms / 1e6, 2) if k3_n else None,
                           "frac_of_measured_hbm": round(k3_rows * row_bytes / k3_ms / 1e6 / peak, 4) if k3_n else None,
                           "rows_per_s": round(k3_rows / k3_ms * 1e3, 1) if k3_n else None},
           "k_gather": {"launches": kg_n, "total_ms": round(kg_ms, 2), "bytes": g_bytes, "gbs": round(g_bytes / max(kg_ms, 1e-9) / 1e6, 2),
                        "frac_of_measured_hbm": round(g_bytes / max(kg_ms, 1e-9) / 1e6 / peak, 4)},
           "k1_decode_total_ms": round(k1_ms, 2), "k3p": {k: ctx.take_counter(k) for k in ("k3p_partitions", "k3p_repartitions", "k3p_output_regrows", "k3_partial_rows_built")},
           "k3p_aggregate_ms": ctx.take_stage_ms("k3p_aggregate"), "k3p_scatter_ms": ctx.take_stage_ms("k3p_scatter"), "host_ms_per_batch": {k: round(v / 1e3 / args.batches, 3) for k, v in host_us.items()}, "wall_s": round(wall, 2), "result_rows_last_batch": rows_out,
           "note": "per-kernel timing enabled (host sync per kernel); process time by CUDA events around each batch on the library stream"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
