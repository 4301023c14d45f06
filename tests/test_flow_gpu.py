"""Flow-level parity: libdxgpu's processor (C ABI) vs the CPU oracle's restatement of
CommonProcessorFactory.processDataset on the same batches."""
import json

import numpy as np
import pytest

import dxo
import dxo_flow
from adversarial import all_records, to_ndjson
from flowcmp import assert_same_rows, same_metrics

pytestmark = pytest.mark.gpu

SCHEMA2 = dxo.iot_schema_json(2)
SCHEMA1 = dxo.iot_schema_json(1)
PROJ_TS = "CAST(Raw.deviceDetails.eventTime / 1000 AS timestamp) AS eventTimeStamp\nRaw.*\n"
T0 = 1_700_000_040_000      # a multiple of 60 s
I = 60_000


def both(dx, ctx, conf, schema, proj, transform):
    g = dx.CommonProcessorFactory.createProcessor(ctx, conf, schema, proj, transform)
    o = dxo_flow.Flow(conf, schema, proj, transform)
    return g, o


def gen(dx, n, first=0, shape=2, n_keys=1000, base=T0, seed=20260103, late=10, future=1, wm=60_000):
    p = dxo.GenParams(seed, n_keys, max(1, n_keys // 10), base, I, wm, shape, late, future, 0)
    return dxo.gen_events(p, first, n)


def check_batch(dx, g, o, data, batch_ms, now_ms=None, tables=None):
    rg = g.process_json(data, batch_ms, I, now_ms=now_ms)
    ro, mo = o.process(data, batch_ms, I, now_ms=now_ms)
    assert same_metrics(rg.metrics, mo), (rg.metrics, mo)
    for name in (tables or rg.table_names):
        assert_same_rows(rg.table(name), ro[name], dx)
    rg.release()


def test_rules_flags_match_oracle(dx, ctx):
    rules = dx.flowgen.make_rules(200, n_keys=1000, n_homes=100)
    tr = dx.flowgen.process_rules_sql(rules)
    conf = dx.flowgen.conf_text("c2", outputs=["Rules"])
    g, o = both(dx, ctx, conf, SCHEMA2, ["Raw.*\n"], tr)
    data = gen(dx, 50_000)
    rg = g.process_json(data, T0, I)
    ro, mo = o.process(data, T0, I)
    t = rg.table("Rules")
    assert dx.rules_of(t, "Rules") == ro["Rules"].rule_table
    assert_same_rows(t, ro["Rules"], dx, ordered=True)
    flags, _ = t.column("Rules")
    assert flags.shape == (50_000, 7)
    # per-rule hit counts computed on the device == column sums of the flag bits
    want = np.unpackbits(flags.view(np.uint8), axis=1, bitorder="little")[:, :200].sum(axis=0)
    assert (dx.rule_hits(t, "Rules") == want).all()
    # every rule fires at least once on 50k events except the very selective ones; none fires always
    fired = np.unpackbits(flags.view(np.uint8), axis=1, bitorder="little")[:, :200].sum(axis=0)
    assert (fired > 0).sum() > 150 and (fired < 50_000).all()
    g.close()


def test_rules_on_adversarial_events_three_valued_logic(dx, ctx):
    rules = [{"ruleId": "R%d" % i, "ruleDescription": "d", "severity": "Critical", "tagName": "Tag", "tag": "t", "condition": c}
             for i, c in enumerate([
                 "deviceDetails.temperature > 50", "deviceDetails.temperature <= 50 OR deviceDetails.humidity > 1e300",
                 "NOT (deviceDetails.temperature > 50)", "deviceDetails.status = 1 AND deviceDetails.deviceType = 'Heating'",
                 "deviceDetails.deviceType <> 'Heating'", "deviceDetails.deviceType LIKE '%Lock'", "deviceDetails.deviceType NOT LIKE 'Door%'",
                 "deviceDetails.deviceType LIKE '_eating'", "deviceDetails.deviceId IN (1, 2, 3, 5, 8)", "deviceDetails.homeId BETWEEN 10 AND 5000",
                 "deviceDetails.status IS NULL", "deviceDetails.status IS NOT NULL AND deviceDetails.temperature IS NULL",
                 "deviceDetails.temperature = deviceDetails.humidity", "deviceDetails.deviceId + deviceDetails.homeId * 2 > 1000000",
                 "deviceDetails.deviceId / 0 > 1", "deviceDetails.temperature / deviceDetails.humidity >= 1.5",
                 "deviceDetails.deviceId % 7 = 3", "deviceDetails.temperature > deviceDetails.deviceId",
                 "IF(deviceDetails.status = 1, deviceDetails.temperature, deviceDetails.humidity) < 20",
                 "deviceDetails.status = 1 OR NULL", "deviceDetails.eventTime >= 1600000000000 AND deviceDetails.eventTime < 1600000500000",
                 "-deviceDetails.temperature < -99.5", "deviceDetails.homeId <> 150 OR deviceDetails.deviceId = 3"])]
    tr = dx.flowgen.process_rules_sql(rules)
    g, o = both(dx, ctx, dx.flowgen.conf_text("adv"), SCHEMA2, ["Raw.*\n"], tr)
    data = to_ndjson(all_records(3000, seed=21))
    rg = g.process_json(data, T0, I)
    ro, _ = o.process(data, T0, I)
    assert_same_rows(rg.table("Rules"), ro["Rules"], dx, ordered=True)
    g.close()


def test_filter_project_and_alert_statements(dx, ctx):
    tr = """--DataXQuery--
sa1_1_1 = SELECT *, 'R100' AS ruleId, 'DoorLock Close' AS ruleDescription, 'Critical' AS severity, 'CLOSE' AS Tag
    FROM DataXProcessedInput WHERE deviceDetails.deviceType = 'DoorLock' AND deviceDetails.status = 1
--DataXQuery--
sa2_1_1 = SELECT * FROM sa1_1_1
--DataXQuery--
CLOSEAlert = SELECT DISTINCT DATE_TRUNC('second', current_timestamp()) AS EventTime, 'CLOSEAlert' AS MetricName, 0 as Metric,
    'iotsample' AS Product, 'DoorLock Close' AS Pivot1 FROM sa1_1_1
--DataXQuery--
HeaterStateFiltered = SELECT eventTimeStamp, deviceDetails.deviceId, deviceDetails.deviceType, deviceDetails.homeId, deviceDetails.status
    FROM DataXProcessedInput WHERE deviceDetails.homeId = 15 AND deviceDetails.deviceType = 'Heating'
    GROUP BY eventTimeStamp, deviceId, deviceType, homeId, status
--DataXQuery--
HeaterStateOneIsOn = SELECT DISTINCT DATE_TRUNC('second', current_timestamp()) AS EventTime, 'HeaterStateOneIsOn' AS MetricName,
    status AS Metric, 'iotsample' AS Product, '' AS Pivot1 FROM HeaterStateFiltered GROUP BY EventTime, MetricName, Metric, Product, Pivot1
--DataXQuery--
Derived = SELECT deviceDetails.deviceId AS id, deviceDetails.temperature * 1.8 + 32 AS fahrenheit, hour(eventTimeStamp) AS h,
    deviceDetails.status = 1 AS isOn, CAST(deviceDetails.humidity AS long) AS hum FROM DataXProcessedInput WHERE deviceDetails.temperature > 90
"""
    g, o = both(dx, ctx, dx.flowgen.conf_text("alerts"), SCHEMA2, [PROJ_TS], tr)
    data = gen(dx, 40_000, n_keys=200)
    check_batch(dx, g, o, data, T0, now_ms=T0 + 1234)
    check_batch(dx, g, o, b"", T0 + I, now_ms=T0 + I)      # empty micro-batch
    g.close()


AGG_TR = """--DataXQuery--
PerKey = SELECT deviceDetails.deviceId, deviceDetails.homeId, COUNT(*) AS Count, SUM(deviceDetails.status) AS SumStatus,
    MIN(deviceDetails.status) AS MinStatus, MAX(deviceDetails.status) AS MaxStatus, AVG(deviceDetails.status) AS AvgStatus,
    SUM(deviceDetails.temperature) AS SumT, MIN(deviceDetails.temperature) AS MinT, MAX(deviceDetails.temperature) AS MaxT,
    AVG(deviceDetails.temperature) AS AvgT, MAX(eventTimeStamp) AS LastSeen, COUNT(deviceDetails.humidity) AS NHum
    FROM DataXProcessedInput GROUP BY deviceDetails.deviceId, deviceDetails.homeId
--DataXQuery--
PerType = SELECT deviceDetails.deviceType, COUNT(*) AS Count, AVG(deviceDetails.temperature) AS Temperature_AVG FROM DataXProcessedInput GROUP BY deviceDetails.deviceType
--DataXQuery--
Total = SELECT COUNT(*) AS Count, MAX(deviceDetails.temperature) AS MaxT FROM DataXProcessedInput WHERE deviceDetails.status = 1
--DataXQuery--
Nothing = SELECT COUNT(*) AS Count, MIN(deviceDetails.homeId) AS m FROM DataXProcessedInput WHERE deviceDetails.status = 7
--DataXQuery--
ar2 = SELECT *, IF(Temperature_AVG > 50 AND Count > 10, MAP('ruleId', 'R3', 'ruleDescription', 'hot on average', 'severity', 'Critical', 'Tag', 'HotAgg'), NULL) AS RuleObject FROM PerType
"""


def test_group_by_aggregates(dx, ctx):
    g, o = both(dx, ctx, dx.flowgen.conf_text("agg"), SCHEMA2, [PROJ_TS], AGG_TR)
    check_batch(dx, g, o, gen(dx, 60_000, n_keys=3000), T0)
    # NULL keys / NULL values / malformed rows take part as Spark says
    check_batch(dx, g, o, to_ndjson(all_records(3000, seed=31)), T0 + I)
    g.close()


WIN_TR = """--DataXQuery--
Tumbling = SELECT deviceDetails.deviceId, deviceDetails.homeId, COUNT(*) AS Count, SUM(deviceDetails.status) AS S, MIN(deviceDetails.status) AS Mn,
    MAX(deviceDetails.status) AS Mx, AVG(deviceDetails.status) AS Av FROM DataXProcessedInput GROUP BY deviceDetails.deviceId, deviceDetails.homeId
--DataXQuery--
Sliding5 = SELECT deviceDetails.deviceId, deviceDetails.homeId, COUNT(*) AS Count, SUM(deviceDetails.status) AS S, MAX(eventTimeStamp) AS Last
    FROM DataXProcessedInput_5minutes GROUP BY deviceDetails.deviceId, deviceDetails.homeId
--DataXQuery--
Sliding2 = SELECT COUNT(*) AS Count FROM DataXProcessedInput_2minutes
--DataXQuery--
WholeWindow = SELECT COUNT(*) AS Count, MIN(eventTimeStamp) AS First FROM DataXProcessedInput_Window
--DataXQuery--
ThisBatch = SELECT COUNT(*) AS Count FROM DataXProcessedInput_Batch
"""


@pytest.mark.parametrize("wm", ["60 second", "0 second", "90 second"])
def test_time_windows_with_reference_quirks(dx, ctx, wm):
    conf = dx.flowgen.conf_text("win", timestamp_column="eventTimeStamp", watermark=wm,
                                windows=[("DataXProcessedInput_5minutes", "5 minutes"), ("DataXProcessedInput_2minutes", "2 minutes")])
    g, o = both(dx, ctx, conf, SCHEMA1, [PROJ_TS], WIN_TR)
    wm_ms = dxo_flow.parse_duration(wm) * 1000
    for b in range(12):
        T = T0 + b * I
        data = gen(dx, 4000, first=b * 4000, shape=1, n_keys=300, base=T, late=30, future=20, wm=wm_ms)
        check_batch(dx, g, o, data, T)
    g.close()


def test_pane_partials_match_rescan(dx, ctx):
    """Window aggregates served from per-batch pane partials (the default when the window ends fall on interval
    boundaries) against the full rescan of every cached batch (gpu.panepartials=false) and against the oracle:
    integer and double aggregates, a string group key, NULL keys and values, late and future events."""
    tr = """--DataXQuery--
ByKey = SELECT deviceDetails.deviceId, deviceDetails.deviceType, COUNT(*) AS Count, COUNT(deviceDetails.status) AS N, SUM(deviceDetails.status) AS S,
    MIN(deviceDetails.temperature) AS Tmin, MAX(deviceDetails.temperature) AS Tmax, AVG(deviceDetails.humidity) AS Havg, SUM(deviceDetails.temperature) AS Tsum,
    MAX(eventTimeStamp) AS Last FROM DataXProcessedInput_5minutes GROUP BY deviceDetails.deviceId, deviceDetails.deviceType
--DataXQuery--
ByHome = SELECT deviceDetails.homeId, MIN(deviceDetails.eventTime) AS FirstMs, COUNT(*) AS Count FROM DataXProcessedInput_3minutes GROUP BY deviceDetails.homeId
--DataXQuery--
Total = SELECT COUNT(*) AS Count, AVG(deviceDetails.temperature) AS Tavg FROM DataXProcessedInput_5minutes
"""
    windows = [("DataXProcessedInput_5minutes", "5 minutes"), ("DataXProcessedInput_3minutes", "3 minutes")]
    conf = dx.flowgen.conf_text("pane", timestamp_column="eventTimeStamp", watermark="60 second", windows=windows)
    g, o = both(dx, ctx, conf, SCHEMA2, [PROJ_TS], tr)
    g_rescan = dx.CommonProcessorFactory.createProcessor(ctx, conf + "datax.job.process.gpu.panepartials=false\n", SCHEMA2, [PROJ_TS], tr)
    g_global = dx.CommonProcessorFactory.createProcessor(ctx, conf + "datax.job.process.gpu.panesmem=false\n", SCHEMA2, [PROJ_TS], tr)
    # one partition cannot hold these groups in its shared table: the overflow -> finer partitions -> rebuild path
    g_coarse = dx.CommonProcessorFactory.createProcessor(ctx, conf + "datax.job.process.gpu.panepartitions=1\n", SCHEMA2, [PROJ_TS], tr)
    ctx.take_counter("k3p_repartitions")
    ctx.take_counter("k3_partial_rows_built")
    for b in range(9):
        T = T0 + b * I
        if b == 4:
            data = to_ndjson(all_records(1500, seed=33))          # a batch of adversarial events: NULL timestamps, keys and values
        else:
            data = gen(dx, 3000, first=b * 3000, shape=2, n_keys=600, base=T, late=30, future=20)
        ro, mo = o.process(data, T, I)
        for proc in (g, g_rescan, g_global, g_coarse):
            rg = proc.process_json(data, T, I)
            assert same_metrics(rg.metrics, mo)
            for name in ("ByKey", "ByHome", "Total"):
                assert_same_rows(rg.table(name), ro[name], dx)
            rg.release()
    assert ctx.take_counter("k3_partial_rows_built") > 0           # the partial path really ran
    assert ctx.take_counter("k3p_repartitions") > 0
    g.close(); g_rescan.close(); g_global.close(); g_coarse.close()


def test_pane_partials_shared_between_statements_and_skewed_keys(dx, ctx):
    """Two statements over the same keys share one partial per cached batch when the accumulators of one cover the
    other's (whichever comes first in the file), and a hot key that outgrows its partition's fixed slice in the
    single-pass scatter takes the exact two-pass form: results equal the oracle's either way."""
    def flow(order):
        a = "Narrow = SELECT deviceDetails.deviceId, COUNT(*) AS Count, SUM(deviceDetails.status) AS S FROM DataXProcessedInput_5minutes GROUP BY deviceDetails.deviceId\n"
        b = ("Wide = SELECT deviceDetails.deviceId, COUNT(*) AS Count, SUM(deviceDetails.status) AS S, MIN(deviceDetails.status) AS Mn, MAX(deviceDetails.status) AS Mx, "
             "AVG(deviceDetails.status) AS Av FROM DataXProcessedInput_3minutes GROUP BY deviceDetails.deviceId\n")
        return "".join("--DataXQuery--\n" + q for q in ((a, b) if order == 0 else (b, a)))
    windows = [("DataXProcessedInput_5minutes", "5 minutes"), ("DataXProcessedInput_3minutes", "3 minutes")]
    conf = dx.flowgen.conf_text("share", timestamp_column="eventTimeStamp", watermark="60 second", windows=windows)
    for order in (0, 1):
        g, o = both(dx, ctx, conf, SCHEMA1, [PROJ_TS], flow(order))
        ctx.take_counter("k3_partial_rows_built"); ctx.take_counter("k3p_two_pass_fallbacks")
        built = []
        for b in range(6):
            T = T0 + b * I
            if b == 3:      # 9 of 10 events carry one device id: its partition gets far more than mean + 8 sigma records
                recs = [json.dumps({"deviceDetails": {"deviceId": 7 if k % 10 else 1000 + k, "deviceType": "DoorLock", "eventTime": T - 1000 - (k % 50000),
                                                      "homeId": 3, "status": k % 2}}, separators=(",", ":")) for k in range(60_000)]
                data = ("\n".join(recs) + "\n").encode()
            else:
                data = gen(dx, 3000, first=b * 3000, shape=1, n_keys=500, base=T, late=30, future=20)
            check_batch(dx, g, o, data, T)
            built.append(ctx.take_counter("k3_partial_rows_built"))
        assert ctx.take_counter("k3p_two_pass_fallbacks") >= 1
        # from the third batch on (windows span several cached batches, the wider partial exists) one partial per batch serves both statements
        if order == 1:
            assert built[4] > 0 and built[5] > 0
        g.close()


def test_state_table_append_with_retention_and_upsert(dx, ctx):
    tr = """--DataXQuery--
PerKey = SELECT deviceDetails.deviceId, deviceDetails.homeId, MAX(eventTimeStamp) AS WindowEnd, COUNT(*) AS Count
    FROM DataXProcessedInput GROUP BY deviceDetails.deviceId, deviceDetails.homeId
--DataXQuery--
acc_new = SELECT deviceId, homeId, WindowEnd, Count FROM PerKey
    UNION ALL SELECT deviceId, homeId, WindowEnd, Count FROM acc WHERE unix_timestamp() - to_unix_timestamp(WindowEnd, 'yyyy-MM-dd') < 200
--DataXQuery--
acc = SELECT deviceId, homeId, WindowEnd, Count FROM acc_new
--DataXQuery--
merged = SELECT deviceId, homeId, Count FROM PerKey UNION ALL SELECT deviceId, homeId, Count FROM totals
--DataXQuery--
totals = SELECT deviceId, homeId, SUM(Count) AS Count FROM merged GROUP BY deviceId, homeId
"""
    conf = dx.flowgen.conf_text("state", state_tables=[("acc", "deviceId long, homeId long, WindowEnd Timestamp, Count long"),
                                                       ("totals", "deviceId long, homeId long, Count long")])
    g, o = both(dx, ctx, conf, SCHEMA1, [PROJ_TS], tr)
    for b in range(8):
        T = T0 + b * I
        check_batch(dx, g, o, gen(dx, 3000, first=b * 3000, shape=1, n_keys=200, base=T), T, now_ms=T + 5)
        t, which = g.state_table("totals")
        assert which == o.state["totals"]["active"]
        assert_same_rows(t, o.state["totals"]["tables"][o.state["totals"]["active"]], dx)
    g.close()


def test_state_export_import_restart(dx, ctx):
    """A processor restarted from the exported state table (dx_flow_state_export -> host -> dx_flow_state_import,
    what the JVM does through the Parquet A/B copies) continues exactly like one that never stopped; the table has a
    string column, so the restored dictionary and the decoder's meet in the UNION ALL."""
    tr = """--DataXQuery--
acc_new = SELECT deviceDetails.deviceId, deviceDetails.deviceType, eventTimeStamp AS EventTime, deviceDetails.status AS Reading
    FROM DataXProcessedInput WHERE deviceDetails.status = 1
    UNION ALL SELECT deviceId, deviceType, EventTime, Reading FROM acc WHERE unix_timestamp() - to_unix_timestamp(EventTime, 'yyyy-MM-dd') < 150
--DataXQuery--
acc = SELECT deviceId, deviceType, EventTime, Reading FROM acc_new
--DataXQuery--
perType = SELECT deviceType, COUNT(*) AS Cnt, MAX(EventTime) AS Last FROM acc GROUP BY deviceType
"""
    conf = dx.flowgen.conf_text("restart", state_tables=[("acc", "deviceId long, deviceType string, EventTime Timestamp, Reading long")])
    g, o = both(dx, ctx, conf, SCHEMA1, [PROJ_TS], tr)
    batch = lambda b: gen(dx, 2000, first=b * 2000, shape=1, n_keys=100, base=T0 + b * I)
    for b in range(3):
        check_batch(dx, g, o, batch(b), T0 + b * I, now_ms=T0 + b * I + 5)
    t, which = g.state_table("acc")
    saved = [t.column(i) for i in range(len(t.columns))]
    assert t.num_rows > 0 and any(v is None for v in saved[1][0]) is False
    g.close()
    g2 = dx.CommonProcessorFactory.createProcessor(ctx, conf, SCHEMA1, [PROJ_TS], tr)
    g2.import_state("acc", saved, active=which)
    t2, which2 = g2.state_table("acc")
    assert which2 == which
    assert_same_rows(t2, o.state["acc"]["tables"][which], dx)
    for b in range(3, 7):
        check_batch(dx, g2, o, batch(b), T0 + b * I, now_ms=T0 + b * I + 5)
        t2, which2 = g2.state_table("acc")
        assert which2 == o.state["acc"]["active"]
        assert_same_rows(t2, o.state["acc"]["tables"][which2], dx)
    g2.close()


def test_timestamp_to_string_cast_and_union_widening(dx, ctx):
    """CAST(timestamp AS STRING) (yyyy-MM-dd HH:mm:ss[.fraction without trailing zeros], UTC) and the string promotion Spark
    applies when a UNION ALL meets a string column and a timestamp column."""
    tr = """--DataXQuery--
A = SELECT CAST(eventTimeStamp AS STRING) AS s, CONCAT('at ', eventTimeStamp) AS c, deviceDetails.deviceId FROM DataXProcessedInput
--DataXQuery--
U = SELECT deviceDetails.deviceType AS x, deviceDetails.deviceId FROM DataXProcessedInput WHERE deviceDetails.status = 1
    UNION ALL SELECT eventTimeStamp AS x, deviceDetails.deviceId FROM DataXProcessedInput WHERE deviceDetails.status = 0
--DataXQuery--
D = SELECT x, COUNT(*) AS n FROM U GROUP BY x
"""
    g, o = both(dx, ctx, dx.flowgen.conf_text("tsstr"), SCHEMA1, [PROJ_TS], tr)
    check_batch(dx, g, o, gen(dx, 4000, shape=1, n_keys=50), T0)
    check_batch(dx, g, o, to_ndjson(all_records(1500, seed=5)), T0 + I)
    g.close()


def test_left_outer_join(dx, ctx):
    """LEFT [OUTER] JOIN: unmatched left rows survive once with NULL right columns (reference data that does not
    know every device; a derived table that lacks some keys), NULL keys never match."""
    tr = """--DataXQuery--
Named = SELECT deviceDetails.deviceId, deviceDetails.status, names.deviceName, names.floor FROM DataXProcessedInput
    LEFT JOIN names ON deviceDetails.deviceId = names.deviceId
--DataXQuery--
PerHome = SELECT deviceDetails.homeId AS h, COUNT(*) AS n FROM DataXProcessedInput WHERE deviceDetails.status = 1 GROUP BY deviceDetails.homeId
--DataXQuery--
Both = SELECT a.deviceId, a.deviceName, b.n FROM Named a LEFT OUTER JOIN PerHome b ON a.deviceId = b.h WHERE a.status = 0
--DataXQuery--
Cnt = SELECT deviceName, COUNT(*) AS c, COUNT(floor) AS f FROM Named GROUP BY deviceName
"""
    csv = "deviceId,deviceName,floor\n1,door,0\n3,window,1\n5,heater,\n8,garage,2\n"
    g, o = both(dx, ctx, dx.flowgen.conf_text("lj"), SCHEMA1, ["Raw.*\n"], tr)
    g.add_refdata("names", csv)
    o.add_refdata("names", csv)
    check_batch(dx, g, o, gen(dx, 3000, shape=1, n_keys=12), T0)
    check_batch(dx, g, o, to_ndjson(all_records(1500, seed=21)), T0 + I)
    g.close()


def test_having(dx, ctx):
    """HAVING over aggregates that are and are not in the select list, a select alias and a group key."""
    tr = """--DataXQuery--
Busy = SELECT deviceDetails.homeId, COUNT(*) AS n, AVG(deviceDetails.temperature) AS t FROM DataXProcessedInput
    GROUP BY deviceDetails.homeId HAVING COUNT(*) > 20 AND MAX(deviceDetails.humidity) > 90 AND n < 100000 AND deviceDetails.homeId <> 3
--DataXQuery--
Hot = SELECT deviceDetails.deviceType, MIN(deviceDetails.temperature) AS lo FROM DataXProcessedInput
    WHERE deviceDetails.status = 1 GROUP BY deviceDetails.deviceType HAVING AVG(deviceDetails.temperature) >= 50 OR SUM(deviceDetails.status) IS NULL
"""
    g, o = both(dx, ctx, dx.flowgen.conf_text("having"), SCHEMA2, ["Raw.*\n"], tr)
    check_batch(dx, g, o, gen(dx, 6000, n_keys=400), T0)
    check_batch(dx, g, o, to_ndjson(all_records(2000, seed=21)), T0 + I)
    g.close()


def test_case_when(dx, ctx):
    """CASE WHEN / simple CASE (Spark's CaseWhen: first TRUE arm, a NULL condition is not taken, no ELSE -> NULL)."""
    tr = """--DataXQuery--
A = SELECT deviceDetails.deviceId, CASE WHEN deviceDetails.temperature > 50 THEN 1 WHEN deviceDetails.humidity > 50 THEN 2 END AS band,
    CASE deviceDetails.status WHEN 1 THEN deviceDetails.temperature ELSE deviceDetails.humidity END AS v
    FROM DataXProcessedInput WHERE CASE WHEN deviceDetails.homeId > 50 THEN deviceDetails.status = 1 ELSE deviceDetails.status = 0 END
--DataXQuery--
B = SELECT band, COUNT(*) AS n, MAX(v) AS mx FROM A GROUP BY band
"""
    g, o = both(dx, ctx, dx.flowgen.conf_text("case"), SCHEMA2, ["Raw.*\n"], tr)
    check_batch(dx, g, o, gen(dx, 5000, n_keys=200), T0)
    check_batch(dx, g, o, to_ndjson(all_records(2000, seed=21)), T0 + I)
    g.close()


def test_unsupported_sql_fails_at_create(dx, ctx):
    for bad in ["--DataXQuery--\nx = SELECT a FROM t ORDER BY a\n", "--DataXQuery--\nSET spark.x=1\n--DataXQuery--\n"]:
        with pytest.raises(dx.DxError):
            dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("bad"), SCHEMA1, ["Raw.*\n"], bad)
    g = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("bad2"), SCHEMA1, ["Raw.*\n"],
                                                 "--DataXQuery--\nx = SELECT nosuch FROM DataXProcessedInput\n")
    with pytest.raises(dx.DxError):
        g.process_json(b'{"deviceDetails":{"deviceId":1}}\n', T0, I)
    g.close()


def test_bucketed_and_generic_rule_kernels_agree(dx, ctx):
    """The bucketed bit-sliced K2 and the bytecode interpreter are two implementations of the same rules."""
    rules = dx.flowgen.make_rules(200, seed=5, n_keys=1000, n_homes=100)
    extra = ["deviceDetails.temperature > deviceDetails.humidity", "NOT (deviceDetails.status = 1 OR deviceDetails.temperature < 50)",
             "deviceDetails.deviceId IN (1, 2, 3) AND deviceDetails.homeId NOT IN (5, 6)", "deviceDetails.temperature NOT BETWEEN 20 AND 80",
             "deviceDetails.deviceType LIKE '%Lock' AND NOT (deviceDetails.deviceType = 'DoorLock')", "deviceDetails.status IS NULL OR deviceDetails.homeId > 50",
             "deviceDetails.homeId > 10.5 AND deviceDetails.homeId < 90",
             "(deviceDetails.status = 1 AND deviceDetails.homeId = 3) OR (deviceDetails.temperature > 99 AND deviceDetails.humidity < 1) OR deviceDetails.deviceId = 7 OR deviceDetails.deviceId = 8 OR deviceDetails.deviceId = 9"]
    for i, c in enumerate(extra):
        rules.append({"ruleId": "X%d" % i, "ruleDescription": "x", "severity": "Low", "tagName": "Tag", "tag": "x", "condition": c})
    tr = dx.flowgen.process_rules_sql(rules)
    conf = dx.flowgen.conf_text("k2")
    fast = dx.CommonProcessorFactory.createProcessor(ctx, conf, SCHEMA2, ["Raw.*\n"], tr)
    slow = dx.CommonProcessorFactory.createProcessor(ctx, conf + "datax.job.process.gpu.genericrules=true\n", SCHEMA2, ["Raw.*\n"], tr)
    o = dxo_flow.Flow(conf, SCHEMA2, ["Raw.*\n"], tr)
    for data in (gen(dx, 30_000), to_ndjson(all_records(3000, seed=21))):
        ctx.set_timing(True)
        a = fast.process_json(data, T0, I).table("Rules")
        assert ctx.take_stage_ms("k2_rules_bucket")[1] == 1 and ctx.take_stage_ms("k2_eval_flags")[1] == 0
        b = slow.process_json(data, T0, I).table("Rules")
        assert ctx.take_stage_ms("k2_eval_flags")[1] == 1
        ctx.set_timing(False)
        fa, _ = a.column("Rules")
        fb, _ = b.column("Rules")
        assert (fa == fb).all()
        ro, _ = o.process(data, T0, I)
        assert_same_rows(a, ro["Rules"], dx, ordered=True)
    fast.close()
    slow.close()


def test_string_to_timestamp_projection_on_reference_sample_events(dx, ctx):
    """The cloud IoT sample's projection (Local.Test/Resource/projection.txt:1-2) over the reference's own sample
    events (string eventTime 'MM/dd/yyyy HH:mm:ss'): stringToTimestamp is evaluated per dictionary entry."""
    import os
    data = open(os.path.join(os.path.dirname(__file__), "golden", "reference_sample_events.ndjson"), "rb").read()
    schema = dxo.iot_schema_json(1, event_time_type="string")
    proj = "stringToTimestamp(Raw.deviceDetails.eventTime) AS eventTimeStamp\nRaw.*\n"
    tr = """--DataXQuery--
ByHour = SELECT hour(eventTimeStamp) AS h, deviceDetails.deviceType, COUNT(*) AS Count, MAX(eventTimeStamp) AS Last FROM DataXProcessedInput GROUP BY hour(eventTimeStamp), deviceDetails.deviceType
--DataXQuery--
Named = SELECT CONCAT(deviceDetails.deviceType, '-', deviceDetails.deviceId) AS name, CAST(deviceDetails.homeId AS String) AS home, eventTimeStamp FROM DataXProcessedInput WHERE deviceDetails.status = 1
"""
    g, o = both(dx, ctx, dx.flowgen.conf_text("s2t"), schema, [proj], tr)
    extra = b'{"deviceDetails":{"deviceId":9,"deviceType":"Heating","eventTime":"not a time","homeId":1,"status":1}}\n' \
            b'{"deviceDetails":{"deviceId":9,"deviceType":null,"eventTime":"2018-08-10 22:55:03","homeId":null,"status":1}}\n' \
            b'{"deviceDetails":{"deviceId":9,"deviceType":"Heating","eventTime":"08/10/2018 22:55:3","homeId":1,"status":1}}\n'
    check_batch(dx, g, o, data + extra, T0)
    check_batch(dx, g, o, data[:2000] + extra, T0 + I)     # second batch: lookup tables extend, dictionaries persist
    g.close()


def test_staged_and_direct_rule_kernels_agree_on_ragged_sizes(dx, ctx):
    """K2 stages full row tiles in shared memory by bulk copies and reads the ragged last tile directly; the direct kernels
    (gpu.rulesstaging=false, also what row selections use) must give the same flags at every tile-boundary case, and so must
    the per-class kernel (gpu.rulesfast=false) the normal-form kernel replaces for rule sets like this one."""
    rules = dx.flowgen.make_rules(200, seed=11, n_keys=1000, n_homes=100)
    tr = dx.flowgen.process_rules_sql(rules)
    conf = dx.flowgen.conf_text("k2s")
    variants = [conf, conf + "datax.job.process.gpu.rulesstaging=false\n", conf + "datax.job.process.gpu.rulesfast=false\n",
                conf + "datax.job.process.gpu.rulesfast=false\ndatax.job.process.gpu.rulesstaging=false\n"]
    procs = [dx.CommonProcessorFactory.createProcessor(ctx, c, SCHEMA2, ["Raw.*\n"], tr) for c in variants]
    o = dxo_flow.Flow(conf, SCHEMA2, ["Raw.*\n"], tr)
    cases = [gen(dx, n, seed=20260200 + n) for n in (1, 255, 256, 257, 511, 513, 30_011)] + [to_ndjson(all_records(3000, seed=21))]
    for data in cases:
        ctx.take_launches()
        res = [p.process_json(data, T0, I) for p in procs]
        fa, _ = res[0].table("Rules").column("Rules")
        for r in res[1:]:
            fb, _ = r.table("Rules").column("Rules")
            assert fa.shape == fb.shape and (fa == fb).all()
        ro, _ = o.process(data, T0, I)
        assert_same_rows(res[0].table("Rules"), ro["Rules"], dx, ordered=True)
        for r in res:
            r.release()
    for p in procs:
        p.close()


def test_rule_widths_and_shapes_through_the_normal_form_kernel(dx, ctx):
    """Rule sets of every flag-word count from 1 to 8 (the normal-form kernel is instantiated per count), with every shape it
    takes — a, a AND b, a OR b, a AND (b OR c), (a OR b) AND c AND d, a AND b AND c, x IN (..) AND y, x BETWEEN .. AND y —
    plus constants at int32 / int64 edges and equal constants shared between rules; each against the per-class kernel and
    the oracle."""
    import random
    rng = random.Random(5)
    leaves = ["deviceDetails.temperature > %.2f", "deviceDetails.humidity <= %.2f", "deviceDetails.homeId = %d", "deviceDetails.status = %d",
              "deviceDetails.deviceId < %d", "deviceDetails.deviceType = '%s'", "deviceDetails.deviceId >= %d", "deviceDetails.temperature <= %d"]

    def leaf():
        k = rng.randrange(len(leaves))
        if k in (0, 1):
            return leaves[k] % rng.uniform(0, 100)
        if k == 5:
            return leaves[k] % rng.choice(["DoorLock", "WindowLock", "Heating", "GarageDoorLock", "Nope"])
        if k == 3:
            return leaves[k] % rng.randrange(3)
        if k == 7:
            return leaves[k] % rng.randrange(0, 100)
        return leaves[k] % rng.choice([1, 5, 50, 500, 2147483646, 2147483647, 2147483648, -2147483648, 4000000000, -1])
    shapes = ["%s", "%s AND %s", "%s OR %s", "%s AND (%s OR %s)", "(%s OR %s) AND %s AND %s", "%s AND %s AND %s",
              "deviceDetails.homeId IN (1, 2, 3) AND %s", "deviceDetails.temperature BETWEEN 20 AND 30.5 AND %s"]
    for n_rules in (7, 33, 64, 65, 100, 129, 161, 193, 225, 256, 40):
        rules = []
        n_shapes = len(shapes) if n_rules == 40 else 6        # the last set mixes in IN / BETWEEN rules: leaves of one rule in the same column
        for r in range(n_rules):
            sh = shapes[r % n_shapes]
            cond = sh % tuple(leaf() for _ in range(sh.count("%s")))
            rules.append({"ruleId": "R%d" % r, "ruleDescription": "d", "severity": "Low", "tagName": "Tag", "tag": "T", "condition": cond})
        tr = dx.flowgen.process_rules_sql(rules)
        conf = dx.flowgen.conf_text("k2w")
        g = dx.CommonProcessorFactory.createProcessor(ctx, conf, SCHEMA2, ["Raw.*\n"], tr)
        g2 = dx.CommonProcessorFactory.createProcessor(ctx, conf + "datax.job.process.gpu.rulesfast=false\n", SCHEMA2, ["Raw.*\n"], tr)
        o = dxo_flow.Flow(conf, SCHEMA2, ["Raw.*\n"], tr)
        for data in (gen(dx, 5000, seed=n_rules, n_keys=600), to_ndjson(all_records(1500, seed=n_rules))):
            ctx.take_counter("k2_rules_fast_launches")
            a, b = g.process_json(data, T0, I), g2.process_json(data, T0, I)
            assert ctx.take_counter("k2_rules_fast_launches") == 1
            fa, _ = a.table("Rules").column("Rules")
            fb, _ = b.table("Rules").column("Rules")
            assert (fa == fb).all(), n_rules
            ro, _ = o.process(data, T0, I)
            assert_same_rows(a.table("Rules"), ro["Rules"], dx, ordered=True)
            a.release(); b.release()
        g.close(); g2.close()


def test_staged_ingest_equals_process_json(dx, ctx):
    """dx_flow_stage_batch + dx_flow_process_staged (the copy of batch i+1 started before batch i is processed) give the
    tables of dx_flow_process_batch on host memory, over a windowed flow whose state carries from batch to batch."""
    import ctypes as C
    conf = dx.flowgen.conf_text("stg", timestamp_column="eventTimeStamp", watermark="60 second",
                                windows=[("DataXProcessedInput_3minutes", "3 minutes")])
    tr = """--DataXQuery--
K = SELECT deviceDetails.deviceId, COUNT(*) AS Count, MAX(deviceDetails.temperature) AS Tmax FROM DataXProcessedInput_3minutes GROUP BY deviceDetails.deviceId
"""
    g, o = both(dx, ctx, conf, SCHEMA2, [PROJ_TS], tr)
    L = ctx.L
    L.dx_host_alloc.restype, L.dx_host_alloc.argtypes = C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]
    L.dx_host_free.argtypes = [C.c_void_p, C.c_void_p]
    batches = [gen(dx, 4000, first=b * 4000, shape=2, n_keys=300, base=T0 + b * I) for b in range(5)]
    pinned = []
    for data in batches:
        hp = C.c_void_p()
        ctx.check(L.dx_host_alloc(ctx.h, len(data), C.byref(hp)))
        C.memmove(hp.value, data, len(data))
        pinned.append(hp)
    staged = g.stage_json(pinned[0].value, len(batches[0]))
    for b, data in enumerate(batches):
        nxt = g.stage_json(pinned[b + 1].value, len(batches[b + 1])) if b + 1 < len(batches) else None
        rg = g.process_staged(staged, T0 + b * I, I)
        ro, mo = o.process(data, T0 + b * I, I)
        assert same_metrics(rg.metrics, mo)
        assert_same_rows(rg.table("K"), ro["K"], dx)
        rg.release()
        staged = nxt
    # a handle that is never processed
    extra = g.stage_json(pinned[0].value, len(batches[0]))
    fn = L.dx_staged_release
    fn.restype, fn.argtypes = None, [C.c_void_p]
    fn(extra)
    for hp in pinned:
        L.dx_host_free(ctx.h, hp)
    g.close()


def test_transport_metadata_columns(dx, ctx):
    """__DataX_FileInfo beside every event and through every projection step and SELECT * (CPF:95-100, 405-421), DataXProperties
    (PropertiesHandler.scala:26-35, conf appendproperty.*) and the transport's Properties / SystemProperties maps as per-batch
    constants: the streaming default (all of FileInternal() NULL), then a batch a blob adapter described."""
    conf = dx.flowgen.conf_text("meta") + "datax.job.process.appendproperty.Region=westus2\n"
    proj = ["Raw.*\nDataXProperties.BatchTime AS bt\nDataXProperties.Region AS region\nProperties AS _Properties\nSystemProperties.offset AS off\n",
            "deviceDetails.deviceId AS id\nbt\nregion\n_Properties\noff\n"]
    tr = "--DataXQuery--\nAll = SELECT * FROM DataXProcessedInput WHERE id > 0\n--DataXQuery--\nPaths = SELECT DISTINCT __DataX_FileInfo.inputPath AS p, bt FROM DataXProcessedInput\n"
    g, o = both(dx, ctx, conf, SCHEMA2, proj, tr)
    data = gen(dx, 500, n_keys=50)
    rg = g.process_json(data, T0, I, now_ms=T0 + 1234)
    ro, mo = o.process(data, T0, I, now_ms=T0 + 1234)
    assert same_metrics(rg.metrics, mo)
    t = rg.table("All")
    assert t.columns == ["id", "bt", "region", "_Properties", "off", "__DataX_FileInfo.inputPath", "__DataX_FileInfo.outputFolders",
                         "__DataX_FileInfo.outputFileName", "__DataX_FileInfo.fileTime", "__DataX_FileInfo.ruleIndexPrefix", "__DataX_FileInfo.target"]
    assert t.nesting("__DataX_FileInfo.fileTime") == (["fileTime"], "s")
    for name in ("All", "Paths"):
        assert_same_rows(rg.table(name), ro[name], dx)
    assert set(t.column("bt")[0]) == {b"2023-11-14 22:14:00.0"} and set(t.column("region")[0]) == {b"westus2"}
    rg.release()
    info = dict(input_path="wasbs://in/2023/11/14/part-7.json", output_folders={"default": "out/2023/11/14", "alerts": "alerts/14"},
                output_file_name="part-7", file_time_ms=T0 - 5000, rule_index_prefix="r", target="t1",
                properties={"tenant": "contoso"}, system_properties={"offset": "42", "partition": "3"}, executor_id="7")
    g.set_batch_info(**info)
    o.set_batch_info(info)
    rg = g.process_json(data, T0 + I, I, now_ms=T0 + I + 5)
    ro, mo = o.process(data, T0 + I, I, now_ms=T0 + I + 5)
    t = rg.table("All")
    assert "__DataX_FileInfo.outputFolders.alerts" in t.columns and "_Properties.tenant" in t.columns
    assert t.nesting("__DataX_FileInfo.outputFolders.alerts") == (["outputFolders", "alerts"], "sm")
    assert set(t.column("off")[0]) == {b"42"} and set(t.column("__DataX_FileInfo.inputPath")[0]) == {info["input_path"].encode()}
    for name in ("All", "Paths"):
        assert_same_rows(rg.table(name), ro[name], dx)
    rg.release()
    g.close()


def test_user_function_arities_interval_hook_and_normalizer(dx, ctx):
    """The hook surface of datax-core's extension package: functions of 0, 2 and 3 arguments (DynamicUDF.Generator0..3), a
    function refreshed by onInterval whose answers change from batch to batch, and a StringNormalizer (RemoveInvalidChars of
    datax-udf-samples) applied before from_json — each against the oracle running the same Python callables."""
    state = {"greeting": "hi", "ticks": 0}

    def on_interval(batch_ms):
        state["ticks"] += 1
        state["greeting"] = "hi%d" % state["ticks"]
    udfs = {"SITE": (0, lambda: "plant-7"), "PAIR": (2, lambda a, b: a + "/" + b), "TRIPLE": (3, lambda a, b, c: None if b == "3" else "|".join([a, b, c])),
            "GREET": (1, lambda a: state["greeting"] + " " + a)}
    tr = """--DataXQuery--
U = SELECT deviceDetails.deviceId, site() AS s, pair(deviceDetails.deviceType, CAST(deviceDetails.homeId AS String)) AS p,
    triple(deviceDetails.deviceType, deviceDetails.status, deviceDetails.deviceType) AS t, greet(deviceDetails.deviceType) AS g FROM DataXProcessedInput
"""
    conf = dx.flowgen.conf_text("udfn")
    g, o = both(dx, ctx, conf, SCHEMA1, ["Raw.*\n"], tr)
    for name, (n, f) in udfs.items():
        g.register_udfn(name, n, f, refresh_each_batch=(name == "GREET"))
        dxo_flow.UDFS[name] = f
    g.set_interval_hook(on_interval)
    g.set_normalizer(lambda s: s.replace("\x01", ""))
    try:
        for b in range(3):
            data = gen(dx, 800, first=b * 800, shape=1, n_keys=40, base=T0 + b * I)
            dirty = data.replace(b'"deviceType":"', b'"deviceType":"\x01')       # an unescaped control character: the row would be NULL
            greeting_before = state["greeting"]
            rg = g.process_json(dirty, T0 + b * I, I)
            assert state["ticks"] == b + 1 and "Latency-Normalizer" in rg.metrics
            greeting_gpu = state["greeting"]
            state["greeting"] = greeting_gpu                                        # the oracle sees the refreshed state of this batch
            ro, mo = o.process(data, T0 + b * I, I)
            assert_same_rows(rg.table("U"), ro["U"], dx)
            vals = set(rg.table("U").column("g")[0])
            assert all(v.startswith(greeting_gpu.encode() + b" ") for v in vals) and greeting_before != greeting_gpu
            rg.release()
    finally:
        for name in udfs:
            dxo_flow.UDFS.pop(name, None)
    g.close()


def test_batch_of_only_late_events_leaves_the_windows_untouched(dx, ctx):
    """The late-event drop (CPF:161-167) is a predicate on the cached batch here, not a filtered copy: a batch in which EVERY
    event is older than the watermark must count zero kept events, add nothing to any window, and still age out like any
    other cached batch; a batch of events without any timestamp behaves the same."""
    conf = dx.flowgen.conf_text("late", timestamp_column="eventTimeStamp", watermark="60 second",
                                windows=[("DataXProcessedInput_3minutes", "3 minutes")])
    tr = """--DataXQuery--
W = SELECT deviceDetails.deviceId, COUNT(*) AS Count, SUM(deviceDetails.status) AS S, MIN(eventTimeStamp) AS First FROM DataXProcessedInput_3minutes GROUP BY deviceDetails.deviceId
--DataXQuery--
B = SELECT deviceDetails.homeId, COUNT(*) AS Count FROM DataXProcessedInput GROUP BY deviceDetails.homeId
--DataXQuery--
Rows = SELECT deviceDetails.deviceId, eventTimeStamp FROM DataXProcessedInput_3minutes WHERE deviceDetails.status = 1
"""
    g, o = both(dx, ctx, conf, SCHEMA1, [PROJ_TS], tr)
    for b in range(7):
        T = T0 + b * I
        if b == 2:      # events of an hour ago: all late
            data = gen(dx, 2500, first=b * 2500, shape=1, n_keys=200, base=T - 3_600_000, late=0, future=0)
        elif b == 4:    # no timestamps at all
            data = b"\n".join(b'{"deviceDetails":{"deviceId":%d,"deviceType":"Heating","homeId":%d,"status":1}}' % (i % 50, i % 7) for i in range(800)) + b"\n"
        else:
            data = gen(dx, 2500, first=b * 2500, shape=1, n_keys=200, base=T, late=20, future=10)
        rg = g.process_json(data, T, I)
        ro, mo = o.process(data, T, I)
        assert same_metrics(rg.metrics, mo), (b, rg.metrics, mo)
        if b in (2, 4):
            assert rg.metrics["Input_DataXProcessedInput_Events_Count"] == 0
        for name in rg.table_names:
            assert_same_rows(rg.table(name), ro[name], dx)
        rg.release()
    g.close()


def test_per_event_transport_properties(dx, ctx):
    """processEventData / processConsumerRecord put a Properties and a SystemProperties map beside every body (CPF:405-421,
    628-654).  Here the adapter hands over the per-event values of the keys the flow uses (dx_flow_set_event_property); they are
    string columns for that batch: selectable as a whole map, by key, groupable across the batches of a window, and NULL where an
    event does not carry the key.  A per-batch constant of the same key is replaced; the other constants stay."""
    import dxo_json
    proj = "CAST(Raw.deviceDetails.eventTime / 1000 AS timestamp) AS eventTimeStamp\nRaw.*\nProperties AS _Properties\nProperties.partition AS part\nSystemProperties.seq AS seq\n"
    tr = """--DataXQuery--
ByPart = SELECT part, COUNT(*) AS Count, COUNT(seq) AS WithSeq, MIN(deviceDetails.status) AS Mn FROM DataXProcessedInput_3minutes GROUP BY part
--DataXQuery--
P1 = SELECT deviceDetails.deviceId, _Properties.partition AS p, _Properties.source AS src, seq FROM DataXProcessedInput WHERE _Properties.partition = 'p1' AND deviceDetails.status = 1
--DataXQuery--
Maps = SELECT deviceDetails.deviceId, _Properties FROM DataXProcessedInput WHERE deviceDetails.deviceId < 40
"""
    conf = dx.flowgen.conf_text("props", timestamp_column="eventTimeStamp", watermark="60 second",
                                windows=[("DataXProcessedInput_3minutes", "3 minutes")], outputs=["Maps"])
    g, o = both(dx, ctx, conf, SCHEMA1, [proj], tr)
    n = 3000
    for b in range(5):
        T = T0 + b * I
        data = gen(dx, n, first=b * n, shape=1, n_keys=300, base=T, late=10, future=5)
        part = [None if (i + b) % 11 == 0 else "p%d" % ((i * 7 + b) % 5) for i in range(n)]
        seq = [None if i % 3 == 0 else str(b * n + i) for i in range(n)]            # high-cardinality: every value new
        props = {"partition": "constant-that-is-replaced", "source": "hub-%d" % b}
        g.set_batch_info(properties=props)
        o.set_batch_info({"properties": props})
        for f in (g, o):
            f.set_event_property("partition", part)
            f.set_event_property("seq", seq, system=True)
        rg = g.process_json(data, T, I)
        ro, mo = o.process(data, T, I)
        assert same_metrics(rg.metrics, mo)
        for name in rg.table_names:
            assert_same_rows(rg.table(name), ro[name], dx)
        rows = rg.table("Maps").to_json()
        assert rows == dxo_json.to_json_rows(ro["Maps"])
        if b >= 2:      # the base table lags by the watermark (CPF:225-227): the first batches' rows show up later
            assert rows and any(b'"partition":null' in r for r in rows) and all(b'"source":"hub-' in r for r in rows)
        rg.release()
    # the count must match the batch
    g.set_event_property("partition", ["x"] * 5)
    with pytest.raises(dx.DxError):
        g.process_json(gen(dx, 100, shape=1), T0 + 5 * I, I)
    g.close()


def test_restart_from_parquet_state_on_disk(dx, ctx, tmp_path):
    """Restart through the state table's on-disk form (StateTableHandler.scala:26-52, 70-125): the stopped processor leaves
    <location>/<A|B>/part-*.parquet + metadata.info (dx_flow_state_save), a new one loads them (dx_flow_state_load) and continues
    exactly like one that never stopped.  The files are plain Parquet: pyarrow (standing in for Spark's reader) sees the same
    rows, and a copy re-written by pyarrow the way Spark 2.4 writes (snappy, dictionary pages, INT96) loads just as well."""
    import pyarrow.parquet as pq
    tr = """--DataXQuery--
acc_new = SELECT deviceDetails.deviceId, deviceDetails.deviceType, eventTimeStamp AS EventTime, deviceDetails.status AS Reading
    FROM DataXProcessedInput WHERE deviceDetails.status = 1
    UNION ALL SELECT deviceId, deviceType, EventTime, Reading FROM acc WHERE unix_timestamp() - to_unix_timestamp(EventTime, 'yyyy-MM-dd') < 150
--DataXQuery--
acc = SELECT deviceId, deviceType, EventTime, Reading FROM acc_new
--DataXQuery--
perType = SELECT deviceType, COUNT(*) AS Cnt, MAX(EventTime) AS Last FROM acc GROUP BY deviceType
"""
    conf = dx.flowgen.conf_text("restart", state_tables=[("acc", "deviceId long, deviceType string, EventTime Timestamp, Reading long")])
    g, o = both(dx, ctx, conf, SCHEMA1, [PROJ_TS], tr)
    batch = lambda b: gen(dx, 2000, first=b * 2000, shape=1, n_keys=100, base=T0 + b * I)
    for b in range(3):
        check_batch(dx, g, o, batch(b), T0 + b * I, now_ms=T0 + b * I + 5)
    loc = tmp_path / "state" / "acc"
    g.save_state("acc", loc)
    _, which = g.state_table("acc")
    g.close()
    assert (loc / "metadata.info").read_text() == "active=%s\nstandby=%s" % (which, "A" if which == "B" else "B")
    seen = pq.read_table(str(loc / which))
    want = o.state["acc"]["tables"][which]
    assert seen.num_rows == want.n and seen.column_names == ["deviceId", "deviceType", "EventTime", "Reading"]
    assert sorted(seen.column("deviceId").to_pylist()) == sorted(int(x) for x in want.cols[0][2])
    # a Spark-style rewrite of the same rows in two part files
    for k, part in enumerate([seen.slice(0, seen.num_rows // 2), seen.slice(seen.num_rows // 2)]):
        pq.write_table(part, str(loc / which / ("part-%05d-spark.snappy.parquet" % k)), compression="snappy", use_dictionary=True,
                       use_deprecated_int96_timestamps=True, flavor="spark", data_page_version="1.0", row_group_size=300)
    (loc / which / "part-00000.parquet").unlink()
    g2 = dx.CommonProcessorFactory.createProcessor(ctx, conf, SCHEMA1, [PROJ_TS], tr)
    g2.load_state("acc", loc)
    t2, which2 = g2.state_table("acc")
    assert which2 == which
    assert_same_rows(t2, want, dx)
    for b in range(3, 6):
        check_batch(dx, g2, o, batch(b), T0 + b * I, now_ms=T0 + b * I + 5)
        t2, which2 = g2.state_table("acc")
        assert which2 == o.state["acc"]["active"]
        assert_same_rows(t2, o.state["acc"]["tables"][which2], dx)
    # a location nobody has written to: an empty table, active = A
    g3 = dx.CommonProcessorFactory.createProcessor(ctx, conf, SCHEMA1, [PROJ_TS], tr)
    g3.load_state("acc", tmp_path / "nowhere")
    t3, which3 = g3.state_table("acc")
    assert which3 == "A"
    g2.close(); g3.close()
