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
