import sys, os, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import dxgpu as dx, bench_windows as bw
ctx = dx.Context(0)
conf, tr = bw.c3_flow(dx)
flow = dx.CommonProcessorFactory.createProcessor(ctx, conf, dx.flowgen.iot_schema_json(1), [bw.PROJ], tr)
n = 5_000_000
keys = 1_000_000
T0, I = bw.T0, bw.I
tt = []
NAMES = ['k1_decode','k2_eval_column','k3_hash_agg','k3p_scatter','k3p_aggregate','k_gather','k2_eval_flags']
phase = {}
for b in range(80):
    T = T0 + b * I
    gp = dx.GenParams(20260103, keys, keys // 10, T, I, 60_000, 1, 10, 1, 0)
    ptr, nb, _ = ctx.gen_events(gp, b * n, n)
    if b in (30, 70):
        ctx.set_timing(True)
        for k in NAMES: ctx.take_stage_ms(k)
        for k in ('k3_rows','gather_bytes','k3_partial_rows_built'): ctx.take_counter(k)
    if b in (35, 75):
        phase[b] = ({k: tuple(round(x, 3) for x in ctx.take_stage_ms(k)) for k in NAMES}, {k: ctx.take_counter(k) for k in ('k3_rows','gather_bytes','k3_partial_rows_built')})
        ctx.set_timing(False)
    ctx.synchronize()
    t0 = time.perf_counter()
    r = flow.process_device(ptr, nb, T, I, n_events=n, now_ms=T + 5)
    ctx.synchronize()
    t1 = time.perf_counter()
    tt.append((t1 - t0) * 1e3)
    r.release(); ctx.free(ptr)
    if b == 69:
        for k in ["project", "route", "stmt:Tumbling", "stmt:Sliding5m", "stmt:Sliding1h"]: ctx.take_counter("host_us:" + k)
print("wall ms per batch:", [round(x, 1) for x in tt])
print({k: ctx.take_counter("host_us:" + k) / 10 / 1e3 for k in ["project", "route", "stmt:Tumbling", "stmt:Sliding5m", "stmt:Sliding1h"]})
print(phase)
