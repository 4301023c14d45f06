"""Flow-level oracle pinned against the reference's own known-answer tests, and cross-checked
against the product's host-side parsers on the reference's generated-SQL goldens."""
import glob
import os

import pytest

import dxo_flow

REF = "/root/reference"
KAT_SQL = "--DataXQuery--\niottestbatch5s = \nSELECT MIN(myTime) AS __receivedtime,\n      '00000000-0000-0000-0000-000000000000' AS __ruleid,\n\tIoTDeviceId AS __deviceid,\n        MAP('avg', AVG(temperature), 'max', MAX(temperature), 'min', MIN(temperature), 'count', COUNT(temperature)) AS temperature\nFROM DataXProcessedInput\nGROUP BY IoTDeviceId\n--DataXQuery--\niottestbatch5salert = \nSELECT 1 AS `doc.schemaversion`,\n\t'alarm' AS `doc.schema`,\n\t'open' AS status,\n\t'1Rule-1Device-NMessage' AS logic,\n\tunix_timestamp()*1000 AS created,\n\tunix_timestamp()*1000 AS modified,\n\t'Temperature > 80 degrees' AS `rule.description`,\n\t'Critical' AS `rule.severity`,\n\t__ruleid AS `rule.id`,\n\t__deviceid AS `device.id`,\n\tSTRUCT(__ruleid, __deviceid, temperature) AS __aggregates,\n   \t__receivedtime AS `device.msg.received`\nFROM iottestbatch5s\nWHERE temperature.avg>0"


def test_oracle_transform_parser_kat():
    # TransformSQLParserTests.scala:11-21
    cmds, refs = dxo_flow.parse_transform(KAT_SQL)
    assert [c[0] for c in cmds] == ["iottestbatch5s", "iottestbatch5salert"]
    assert cmds[0][2].startswith("SELECT MIN(myTime) AS __receivedtime, '00000000-0000-0000-0000-000000000000' AS __ruleid, IoTDeviceId AS __deviceid, MAP('avg'")
    assert cmds[1][2].endswith("__receivedtime AS `device.msg.received` FROM iottestbatch5s WHERE temperature.avg>0")
    assert refs == {"iottestbatch5s": 1, "iottestbatch5salert": 0}


def test_oracle_string_to_timestamp_kat():
    # ConcurrentDateFormatTests.scala:13-15: "08/10/2018 22:55:3" == Timestamp.valueOf("2018-08-10 22:55:03")
    assert dxo_flow.string_to_timestamp("08/10/2018 22:55:3") == dxo_flow.string_to_timestamp("2018-08-10 22:55:03") == 1533941703000000
    assert dxo_flow.string_to_timestamp("nope") is None


def test_product_parsers_agree_with_oracle(dx):
    for s in ["08/10/2018 22:55:3", "2018-08-10 22:55:03", "2018-8-1 2:5:3.25", "2018-08-10T22:55:03Z", "13/45/2018 25:61:61",
              "05/16/2017 20:38:22", "", "x", "2018-08-10", "2018-08-10 22:55", "2018-08-10 22:55:03.1234567891"]:
        assert dx.string_to_timestamp(s) == dxo_flow.string_to_timestamp(s), s
    for s in ["5 minutes", "60 second", "1 hour", "2 days", "90 s", "1500 ms"]:
        assert dx.duration_seconds(s) == dxo_flow.parse_duration(s), s


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree only exists in the build container")
def test_splitters_agree_on_reference_goldens(dx):
    files = glob.glob(REF + "/Services/DataX.Flow/DataX.Flow.CodegenRules.Tests/[Cc][Gg]en*.txt")
    files += glob.glob(REF + "/Services/DataX.Config/**/*combined*.txt", recursive=True)
    for f in files:
        text = open(f, encoding="utf-8-sig").read()
        a = dx.transform_split(text)
        b = dxo_flow.parse_transform(text)
        assert a[0] == b[0] and a[1] == b[1], f


def test_threaded_c_rule_evaluator_matches_numpy_oracle(dx):
    """oracle/dxo_rules.c (the CPU baseline's rule evaluation) against dxo_flow's numpy evaluator (the oracle of record),
    on generated events and on the adversarial set (NULLs, NaN, three-valued AND/OR/NOT)."""
    import numpy as np
    import dxo
    from adversarial import all_records, to_ndjson
    conds = ["deviceDetails.temperature > 50", "deviceDetails.temperature <= 50 OR deviceDetails.humidity > 1e300",
             "NOT (deviceDetails.temperature > 50)", "deviceDetails.status = 1 AND deviceDetails.deviceType = 'Heating'",
             "deviceDetails.deviceType <> 'Heating'", "deviceDetails.status IS NULL", "deviceDetails.status IS NOT NULL AND deviceDetails.temperature IS NULL",
             "deviceDetails.temperature = deviceDetails.humidity", "deviceDetails.temperature > deviceDetails.deviceId",
             "deviceDetails.status = 1 OR NULL", "deviceDetails.homeId <> 150 OR deviceDetails.deviceId = 3",
             "(deviceDetails.humidity < 30 OR deviceDetails.homeId >= 40) AND NOT deviceDetails.deviceType = 'DoorLock'"]
    rules = [{"ruleId": "R%d" % i, "ruleDescription": "d", "severity": "Critical", "tagName": "Tag", "tag": "t", "condition": c} for i, c in enumerate(conds)]
    rules += dx.flowgen.make_rules(60, seed=9, n_keys=500, n_homes=50)
    o = dxo_flow.Flow(dx.flowgen.conf_text("r"), dxo.iot_schema_json(2), ["Raw.*\n"], dx.flowgen.process_rules_sql(rules))
    p = dxo.GenParams(5, 500, 50, 1_700_000_040_000, 60_000, 60_000, 2, 10, 1, 0)
    for data in (dxo.gen_events(p, 0, 20_000), to_ndjson(all_records(3000, seed=21))):
        ro, _ = o.process(data, 1_700_000_040_000, 60_000)
        want = [c for c in ro["Rules"].cols if c[1] == "rules"][0][2]
        src = ro["Rules"]                        # SELECT *, <rules>: the input columns ride along
        cols, arrs, valids, dicts = [], [], [], {}
        for ci, (name, ty, v, ok) in enumerate(src.cols):
            if ty == "string":
                vocab = sorted({x for x, k in zip(v, ok) if k})
                dicts[ci] = {s: i for i, s in enumerate(vocab)}
                v = np.array([dicts[ci][x] if k else -1 for x, k in zip(v, ok)], dtype=np.int64)
            elif ty not in ("long", "double"):
                continue
            cols.append((name, ty)); arrs.append(v); valids.append(ok)
        cores, _ = dxo_flow.P(o.commands[0][2]).statement()
        rs = dxo_flow._rule_shape(cores[0]["items"][1][0])
        names = [n for n, _ in cols]
        str_code = lambda ci, lit: dicts[[i for i, c in enumerate(src.cols) if c[0] == names[ci]][0]].get(lit.encode(), -2)
        prog = dxo.compile_rules([c for c, _ in rs], cols, str_code)
        flags = dxo.eval_rules_mt(prog, arrs, valids, src.n, len(rs), threads=4)
        got = np.unpackbits(flags.view(np.uint8), axis=1, bitorder="little")[:, :len(rs)].astype(bool)
        assert np.array_equal(got, want)
