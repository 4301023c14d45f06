import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import dxgpu as dx, dxo
import bench_windows as bw
ctx = dx.Context(0)
conf, tr = bw.c3_flow(dx)
flow = dx.CommonProcessorFactory.createProcessor(ctx, conf, dxo.iot_schema_json(1), [bw.PROJ], tr)
n = 5_000_000
for b in range(8):
    T = bw.T0 + b * bw.I
    gp = dx.GenParams(20260103, 1_000_000, 100_000, T, bw.I, 60_000, 1, 10, 1, 0)
    ptr, nb, _ = ctx.gen_events(gp, b * n, n)
    ctx.set_timing(True)
    r = flow.process_device(ptr, nb, T, bw.I, n_events=n, now_ms=T + 5)
    ctx.set_timing(False)
    ms, k = ctx.take_stage_ms("k3_hash_agg")
    rows = ctx.take_counter("k3_rows")
    ev, ke = ctx.take_stage_ms("k2_eval_flags")
    print(f"batch {b}: k3 {ms:9.3f} ms in {k} launches, rows {rows}, eval_flags {ev:.3f} ms/{ke}; metrics {r.metrics}; out rows {[ (t, r.table(t).num_rows) for t in r.table_names]}")
    r.release(); ctx.free(ptr)
