import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import bench, dxgpu as dx, torch, time
ctx = dx.Context(0)
conf, schema, proj, transform = bench.artefacts(dx)
flow = dx.CommonProcessorFactory.createProcessor(ctx, conf, schema, proj, transform)
gp = bench.gen_params(dx.GenParams, 0)
n = 25_000_000
ptr, nb, _ = ctx.gen_events(gp, 0, n)
for i in range(3):
    r = flow.process_device(ptr, nb, bench.BATCH_MS, bench.INTERVAL_MS, n_events=n); r.release()
for k in ("project", "route", "stmt:Rules"): ctx.take_counter("host_us:" + k)
ctx.set_timing(True)
t0 = time.perf_counter()
for i in range(5):
    r = flow.process_device(ptr, nb, bench.BATCH_MS, bench.INTERVAL_MS, n_events=n); r.release()
ctx.synchronize()
t1 = time.perf_counter()
print("wall ms/batch", (t1 - t0) / 5 * 1e3)
print({k: ctx.take_counter("host_us:" + k) / 5 / 1e3 for k in ("project", "route", "stmt:Rules")})
print("k1", ctx.take_stage_ms("k1_decode"), "k2", ctx.take_stage_ms("k2_rules_bucket"))
