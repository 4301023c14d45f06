import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import dxgpu as dx, dxo
ctx = dx.Context(0)
PROJ = "CAST(Raw.deviceDetails.eventTime / 1000 AS timestamp) AS eventTimeStamp\nRaw.*\n"
T0, I = 1_700_000_040_000, 60_000
n = 5_000_000
gp = dx.GenParams(20260103, 1_000_000, 100_000, T0, I, 60_000, 1, 10, 1, 0)
ptr, nb, _ = ctx.gen_events(gp, 0, n)
variants = {
 "count": "COUNT(*) AS Count",
 "count+sum": "COUNT(*) AS Count, SUM(deviceDetails.status) AS S",
 "count+min": "COUNT(*) AS Count, MIN(deviceDetails.status) AS Mn",
 "count+max": "COUNT(*) AS Count, MAX(deviceDetails.status) AS Mx",
 "count+avg": "COUNT(*) AS Count, AVG(deviceDetails.status) AS Av",
 "count+maxts": "COUNT(*) AS Count, MAX(eventTimeStamp) AS Last",
 "all": "COUNT(*) AS Count, SUM(deviceDetails.status) AS S, MIN(deviceDetails.status) AS Mn, MAX(deviceDetails.status) AS Mx, AVG(deviceDetails.status) AS Av",
}
for name, aggs in variants.items():
    tr = "--DataXQuery--\nQ = SELECT deviceDetails.deviceId, deviceDetails.homeId, %s FROM DataXProcessedInput GROUP BY deviceDetails.deviceId, deviceDetails.homeId\n" % aggs
    flow = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("p"), dxo.iot_schema_json(1), [PROJ], tr)
    for rep in range(3):
        ctx.set_timing(True)
        r = flow.process_device(ptr, nb, T0, I, n_events=n)
        ctx.set_timing(False)
        ms, k = ctx.take_stage_ms("k3_hash_agg")
        rows = r.table("Q").num_rows
        r.release()
    print(f"{name:12s} k3 {ms:8.3f} ms launches {k} groups {rows}")
    flow.close()
