"""Full-size checks (BASELINE.json sizes) through properties that do not need a row-by-row CPU pass:
decode checksums against the generator's CPU twin, per-rule hit counts from two independent rule kernels,
additivity of the hit counts over a split batch, and windows from pane partials against the full rescan."""
import numpy as np
import pytest

import dxo
from flowcmp import gpu_rows, same_metrics

pytestmark = pytest.mark.gpu
N = 25_000_000                     # one C2 micro-batch
T0, I = 1_700_000_040_000, 60_000
ARGS = (20260102, 1_000_000, 100_000, T0, I, 60_000, 2, 10, 1, 0)


def test_c2_micro_batch_decode_checksums_and_rule_counts(dx, ctx):
    ptr, nb, _ = ctx.gen_events(dx.GenParams(*ARGS), 0, N)
    truth = dxo.gen_fields(dxo.GenParams(*ARGS), 0, N)            # the generator's CPU twin: field values without any parsing
    rules = dx.flowgen.make_rules(200, seed=20260102, n_keys=1_000_000, n_homes=100_000)
    tr = dx.flowgen.process_rules_sql(rules)
    schema = dxo.iot_schema_json(2)
    fast = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("full"), schema, ["Raw.*\n"], tr)
    generic = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("full") + "datax.job.process.gpu.genericrules=true\n",
                                                        schema, ["Raw.*\n"], tr)
    r = fast.process_device(ptr, nb, T0, I, n_events=N)
    assert r.metrics["Input_Normalized_Events_Count"] == N
    t = r.table("Rules")
    assert t.num_rows == N
    # decode: every column's wrapping sum (bit patterns for doubles) equals the twin's; rows are in input order
    for name, key in (("deviceDetails.deviceId", "deviceId"), ("deviceDetails.eventTime", "eventTime"), ("deviceDetails.homeId", "homeId"),
                      ("deviceDetails.status", "status"), ("deviceDetails.temperature", "temperature"), ("deviceDetails.humidity", "humidity")):
        v, ok = t.column(name)
        assert ok.all()
        assert v.view(np.int64).sum() == truth[key].view(np.int64).sum(), name
        assert (v.view(np.int64)[::997] == truth[key].view(np.int64)[::997]).all(), name
    hits = dx.rule_hits(t, "Rules")
    assert hits.shape == (200,) and hits.max() <= N and (hits > 0).sum() > 150
    # the bytecode interpreter is an independent implementation of the same rules
    r2 = generic.process_device(ptr, nb, T0, I, n_events=N)
    assert (dx.rule_hits(r2.table("Rules"), "Rules") == hits).all()
    r2.release()
    # additivity: the two halves of the batch (cut at a record boundary) give hit counts that add up
    head = bytes(ctx.d2h(ptr + nb // 2, 400))
    cut = nb // 2 + head.index(b"\n") + 1
    ha = fast.process_device(ptr, cut, T0, I)
    hb = fast.process_device(ptr + cut, nb - cut, T0, I)
    ta, tb = ha.table("Rules"), hb.table("Rules")
    assert ta.num_rows + tb.num_rows == N
    assert (dx.rule_hits(ta, "Rules") + dx.rule_hits(tb, "Rules") == hits).all()
    for x in (ha, hb, r):
        x.release()
    fast.close(); generic.close()
    ctx.free(ptr)


def test_c3_windows_pane_partials_equal_rescan_at_full_key_count(dx, ctx):
    tr = ""
    for name, table in (("Tumbling", "DataXProcessedInput"), ("Sliding5m", "DataXProcessedInput_5minutes")):
        tr += ("--DataXQuery--\n%s = SELECT deviceDetails.deviceId, deviceDetails.homeId, COUNT(*) AS Count, SUM(deviceDetails.status) AS S, "
               "MIN(deviceDetails.status) AS Mn, MAX(deviceDetails.status) AS Mx FROM %s GROUP BY deviceDetails.deviceId, deviceDetails.homeId\n" % (name, table))
    conf = dx.flowgen.conf_text("c3", timestamp_column="eventTimeStamp", watermark="60 second", windows=[("DataXProcessedInput_5minutes", "5 minutes")])
    proj = ["CAST(Raw.deviceDetails.eventTime / 1000 AS timestamp) AS eventTimeStamp\nRaw.*\n"]
    schema = dxo.iot_schema_json(1)
    panes = dx.CommonProcessorFactory.createProcessor(ctx, conf, schema, proj, tr)
    rescan = dx.CommonProcessorFactory.createProcessor(ctx, conf + "datax.job.process.gpu.panepartials=false\n", schema, proj, tr)
    n = 5_000_000
    for b in range(7):
        T = T0 + b * I
        ptr, nb, _ = ctx.gen_events(dx.GenParams(20260103, 1_000_000, 100_000, T, I, 60_000, 1, 10, 1, 0), b * n, n)
        ra = panes.process_device(ptr, nb, T, I, n_events=n, now_ms=T + 5)
        rb = rescan.process_device(ptr, nb, T, I, n_events=n, now_ms=T + 5)
        assert same_metrics(ra.metrics, rb.metrics)
        for name in ("Tumbling", "Sliding5m"):
            ta, tb = ra.table(name), rb.table(name)
            assert ta.num_rows == tb.num_rows
            ca = [ta.column(i)[0] for i in range(6)]
            cb = [tb.column(i)[0] for i in range(6)]
            # every event of the window is counted exactly once, whichever way the window was computed
            assert ca[2].sum() == cb[2].sum() and ca[3].sum() == cb[3].sum()
            if b in (3, 6):     # full comparison of the ~1 M groups on two of the batches
                oa = np.lexsort((ca[1], ca[0])); ob = np.lexsort((cb[1], cb[0]))
                for x, y in zip(ca, cb):
                    assert (x[oa] == y[ob]).all(), (b, name)
        ra.release(); rb.release()
        ctx.free(ptr)
    assert ctx.take_counter("k3_partial_rows_built") > 0
    panes.close(); rescan.close()
