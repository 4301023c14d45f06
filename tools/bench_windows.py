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


def run(dx, ctx, workload="c3", batches=100, events_per_batch=5_000_000, keys=1_000_000, no_smem=False, no_panes=False, peak=None, device=0):
    """Runs one workload on an existing context and returns the result dict (bench.py puts it under `stages`)."""
    import torch
    conf, tr = c3_flow(dx) if workload == "c3" else c4_flow(dx)
    if no_smem:
        conf += "datax.job.process.gpu.panesmem=false\n"
    if no_panes:
        conf += "datax.job.process.gpu.panepartials=false\n"
    flow = dx.CommonProcessorFactory.createProcessor(ctx, conf, dx.flowgen.iot_schema_json(1), [PROJ], tr)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", device))
    if peak is None:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    n = events_per_batch
    names = ["k1_decode", "k2_eval_column", "k3_hash_agg", "k3p_scatter", "k3p_aggregate", "k_gather"]

    def loop(first, count, timing):
        """-> (CUDA-event ms over the batches, rows of the last batch's tables)"""
        ctx.set_timing(timing)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ms, rows_out = 0.0, {}
        for b in range(first, first + count):
            T = T0 + b * I
            gp = dx.GenParams(20260103, keys, max(1, keys // 10), T, I, 60_000, 1, 10, 1, 0)
            ptr, nb, _ = ctx.gen_events(gp, b * n, n)
            e0.record(stream)
            r = flow.process_device(ptr, nb, T, I, n_events=n, now_ms=T + 5)
            e1.record(stream)
            e1.synchronize()
            ms += e0.elapsed_time(e1)
            if b == first + count - 1:
                for name in r.table_names:
                    rows_out[name] = r.table(name).num_rows
            r.release()
            ctx.free(ptr)
        ctx.set_timing(False)
        return ms, rows_out

    # the headline of the stage: all batches without per-kernel timing (no host sync per kernel)
    for k in names:
        ctx.take_stage_ms(k)
    for k in ("k3_rows", "gather_bytes"):
        ctx.take_counter(k)
    t_wall = time.perf_counter()
    proc_ms, rows_out = loop(0, batches, False)
    wall = time.perf_counter() - t_wall
    k3_rows = ctx.take_counter("k3_rows")
    g_bytes = ctx.take_counter("gather_bytes")
    host_us = {k: ctx.take_counter("host_us:" + k) for k in ["project", "route"] + ["stmt:" + nm for nm in rows_out]}
    # per-kernel times: a few more batches in steady state with the per-kernel CUDA events on
    probe = max(2, min(8, batches // 10))
    for k in ("k3_rows", "gather_bytes"):
        ctx.take_counter(k)
    loop(batches, probe, True)
    st = {k: ctx.take_stage_ms(k) for k in names}
    p_k3_rows, p_g_bytes = ctx.take_counter("k3_rows"), ctx.take_counter("gather_bytes")
    row_bytes = 33 if workload == "c3" else 8 * 4          # keys + ts + value (+validity) per visited row
    kern = {}
    for k, (ms, cnt) in st.items():
        if cnt:
            kern[k] = {"ms_per_batch": round(ms / probe, 3), "launches_per_batch": round(cnt / probe, 1)}
    agg_ms = sum(st[k][0] for k in ("k3_hash_agg", "k3p_scatter", "k3p_aggregate"))
    if agg_ms > 0:
        kern["aggregation"] = {"rows_visited_per_batch": p_k3_rows // probe, "alg_bytes_per_row": row_bytes,
                               "gbs": round(p_k3_rows * row_bytes / agg_ms / 1e6, 2), "frac": round(p_k3_rows * row_bytes / agg_ms / 1e6 / peak, 4),
                               "what": "k3p_scatter + k3p_aggregate + k3_hash_agg time against the rows they visit"}
    if st["k_gather"][1]:
        kern["k_gather"].update({"bytes_per_batch": p_g_bytes // probe, "gbs": round(p_g_bytes / st["k_gather"][0] / 1e6, 2),
                                 "frac": round(p_g_bytes / st["k_gather"][0] / 1e6 / peak, 4)})
    if st["k1_decode"][1]:
        b1 = (115.5 + 37) * n * probe
        kern["k1_decode"].update({"gbs": round(b1 / st["k1_decode"][0] / 1e6, 2), "frac": round(b1 / st["k1_decode"][0] / 1e6 / peak, 4)})
    total_events = n * batches
    out = {"what": ("C3: tumbling + 5-minute + 1-hour COUNT/SUM/MIN/MAX/AVG(status) per (deviceId, homeId)" if workload == "c3" else
                    "C4: per-key tumbling counts merged into an append-with-retention accumulator and a SUM(Count) upsert table") +
                   ", %d x %d S1 events, %d keys" % (batches, n, keys),
           "value": round(total_events / (proc_ms / 1e3), 1), "unit": "events/s", "batches": batches, "events_per_batch": n,
           "keys": keys, "ms_per_batch": round(proc_ms / batches, 3), "kernel_ms_per_batch": round(sum(v[0] for v in st.values()) / probe, 3),
           "kernels": kern, "host_ms_per_batch": {k: round(v / 1e3 / batches, 3) for k, v in host_us.items()},
           "wall_s": round(wall, 2), "result_rows_last_batch": rows_out, "rows_visited_by_aggregation": k3_rows, "gather_bytes": g_bytes,
           "k3p": {k: ctx.take_counter(k) for k in ("k3p_partitions", "k3p_repartitions", "k3p_output_regrows", "k3_partial_rows_built")}}
    flow.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--batches", type=int, default=100)
    ap.add_argument("--events-per-batch", type=int, default=5_000_000)
    ap.add_argument("--keys", type=int, default=1_000_000)
    ap.add_argument("--no-smem", action="store_true", help="pane partials through the global-memory hash table (gpu.panesmem=false)")
    ap.add_argument("--no-panes", action="store_true", help="rescan every cached batch per window (gpu.panepartials=false)")
    args = ap.parse_args()
    import dxgpu as dx
    ctx = dx.Context(0)
    out = run(dx, ctx, args.workload, args.batches, args.events_per_batch, args.keys, args.no_smem, args.no_panes)
    out["metric"] = "events/sec per micro-batch (%s)" % ("window aggregates C3" if args.workload == "c3" else "state-table merge C4")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
