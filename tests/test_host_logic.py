"""Host-side logic of libdxgpu.so that needs no GPU: the reference's own known-answer tests."""
import datetime
import glob
import os

import pytest

REF = "/root/reference"


def test_transform_sql_parser_kat(dx):
    # datax-host/src/test/scala/datax/test/sql/TransformSQLParserTests.scala:11-21 (input and expected, verbatim)
    sql = "--DataXQuery--\niottestbatch5s = \nSELECT MIN(myTime) AS __receivedtime,\n      '00000000-0000-0000-0000-000000000000' AS __ruleid,\n\tIoTDeviceId AS __deviceid,\n        MAP('avg', AVG(temperature), 'max', MAX(temperature), 'min', MIN(temperature), 'count', COUNT(temperature)) AS temperature\nFROM DataXProcessedInput\nGROUP BY IoTDeviceId\n--DataXQuery--\niottestbatch5salert = \nSELECT 1 AS `doc.schemaversion`,\n\t'alarm' AS `doc.schema`,\n\t'open' AS status,\n\t'1Rule-1Device-NMessage' AS logic,\n\tunix_timestamp()*1000 AS created,\n\tunix_timestamp()*1000 AS modified,\n\t'Temperature > 80 degrees' AS `rule.description`,\n\t'Critical' AS `rule.severity`,\n\t__ruleid AS `rule.id`,\n\t__deviceid AS `device.id`,\n\tSTRUCT(__ruleid, __deviceid, temperature) AS __aggregates,\n   \t__receivedtime AS `device.msg.received`\nFROM iottestbatch5s\nWHERE temperature.avg>0"
    cmds, refs = dx.transform_split(sql)
    assert cmds == [
        ("iottestbatch5s", "Query", "SELECT MIN(myTime) AS __receivedtime, '00000000-0000-0000-0000-000000000000' AS __ruleid, IoTDeviceId AS __deviceid, MAP('avg', AVG(temperature), 'max', MAX(temperature), 'min', MIN(temperature), 'count', COUNT(temperature)) AS temperature FROM DataXProcessedInput GROUP BY IoTDeviceId"),
        ("iottestbatch5salert", "Query", "SELECT 1 AS `doc.schemaversion`, 'alarm' AS `doc.schema`, 'open' AS status, '1Rule-1Device-NMessage' AS logic, unix_timestamp()*1000 AS created, unix_timestamp()*1000 AS modified, 'Temperature > 80 degrees' AS `rule.description`, 'Critical' AS `rule.severity`, __ruleid AS `rule.id`, __deviceid AS `device.id`, STRUCT(__ruleid, __deviceid, temperature) AS __aggregates, __receivedtime AS `device.msg.received` FROM iottestbatch5s WHERE temperature.avg>0"),
    ]
    assert refs == {"iottestbatch5s": 1, "iottestbatch5salert": 0}


def test_transform_parser_edge_cases(dx):
    cmds, refs = dx.transform_split("--DataXQuery-- \nx = SELECT 1 FROM t\n")      # trailing blank: not a separator, a comment
    assert cmds == [("x", "Query", "SELECT 1 FROM t")]
    with pytest.raises(dx.DxError):
        dx.transform_split("--DataXQuery--\na = SELECT 1 FROM t\n--DataXQuery--\na = SELECT 2 FROM t\n")
    cmds, _ = dx.transform_split("-- comment\n--DataXQuery--\n  -- another\nT1 = SELECT *\n  FROM x\n\n--DataXQuery--\nSET a=b\n--DataXQuery--\n")
    assert cmds[0] == ("T1", "Query", "SELECT * FROM x")
    # 'SET a=b' does not match ^\s*([a-zA-Z0-9_]+)\s*=(.*)$ (a word, then '='), so it is a Command
    assert cmds[1] == (None, "Command", "SET a=b")
    # a trailing statement without a name is dropped (TransformSqlParser.scala:88-92)
    cmds, _ = dx.transform_split("--DataXQuery--\nSELECT 1 FROM t\n")
    assert cmds == []


def ts(y, mo, d, h, mi, s, us=0):
    return int((datetime.datetime(y, mo, d, h, mi, s, tzinfo=datetime.timezone.utc).timestamp())) * 1_000_000 + us


def test_string_to_timestamp_kat(dx):
    # datax-utility/src/test/scala/datax/utility/test/ConcurrentDateFormatTests.scala:13-15
    assert dx.string_to_timestamp("08/10/2018 22:55:3") == ts(2018, 8, 10, 22, 55, 3)
    assert dx.string_to_timestamp("2018-08-10 22:55:03") == ts(2018, 8, 10, 22, 55, 3)           # Timestamp.valueOf
    assert dx.string_to_timestamp("2018-8-1 2:5:3.5") == ts(2018, 8, 1, 2, 5, 3, 500000)
    assert dx.string_to_timestamp("2018-08-10T22:55:03Z") == ts(2018, 8, 10, 22, 55, 3)          # quoted 'Z': literal
    assert dx.string_to_timestamp("05/16/2017 20:38:22") == ts(2017, 5, 16, 20, 38, 22)          # iotsample.json:52
    assert dx.string_to_timestamp("13/01/2018 00:00:00") == ts(2019, 1, 1, 0, 0, 0)              # lenient roll-over
    assert dx.string_to_timestamp("") is None and dx.string_to_timestamp("garbage") is None


def test_durations(dx):
    # values from the reference's own configs (jobConfig.conf:19-21, HomeAutomationLocal.json:309)
    assert dx.duration_seconds("5 minutes") == 300 and dx.duration_seconds("2 minutes") == 120
    assert dx.duration_seconds("60 second") == 60 and dx.duration_seconds("0 second") == 0
    assert dx.duration_seconds("1 hour") == 3600 and dx.duration_seconds("nonsense") == -1


def test_sql_subset_parser(dx):
    c = dx.sql_canonical
    assert c("SELECT a.b, `x.y` AS z, COUNT(*) AS Count FROM t WHERE a.b = 1 AND (c > 2.5 OR d = 'x') GROUP BY a.b") == \
        "SELECT a.b, x.y AS z, COUNT(*) AS Count FROM t WHERE ((a.b = 1) AND ((c > 2.5) OR (d = 'x'))) GROUP BY a.b"
    assert c("select distinct DATE_TRUNC('second', current_timestamp()) AS EventTime, 0 as Metric from sa1 union all select * from u;") == \
        "SELECT DISTINCT DATE_TRUNC('second', CURRENT_TIMESTAMP()) AS EventTime, 0 AS Metric FROM sa1 UNION ALL SELECT * FROM u"
    assert "NOT IN" in c("SELECT * FROM t WHERE x NOT IN (1,2) AND y BETWEEN 1 AND 3 AND z NOT LIKE '%a' AND w IS NOT NULL")
    assert c("SELECT CASE WHEN a > 1 THEN 'x' WHEN a > 0 THEN 'y' ELSE 'z' END AS k, CASE b WHEN 1 THEN 2 END FROM t") == \
        "SELECT IF((a > 1), 'x', IF((a > 0), 'y', 'z')) AS k, IF((b = 1), 2, NULL) FROM t"
    for bad in ("SELECT a FROM t ORDER BY a", "SELECT CASE WHEN a THEN 1 FROM t", "SELECT a FROM t RIGHT JOIN u ON a=b", "SELEC a"):
        with pytest.raises(dx.DxError):
            c(bad)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree only exists in the build container")
def test_parses_every_generated_statement_of_the_reference_goldens(dx):
    """Every statement in the reference's generated-SQL goldens goes through the splitter and the SQL
    parser (DataX.Flow.CodegenRules.Tests/CGen*.txt, cgen*.txt; Config tests' *-combined.txt)."""
    files = glob.glob(REF + "/Services/DataX.Flow/DataX.Flow.CodegenRules.Tests/[Cc][Gg]en*.txt")
    files += glob.glob(REF + "/Services/DataX.Config/**/*combined*.txt", recursive=True)
    assert len(files) >= 15
    n = 0
    for f in files:
        text = open(f, encoding="utf-8-sig").read()
        cmds, _ = dx.transform_split(text)
        for name, kind, sql in cmds:
            if kind != "Query":
                continue
            if "'DeviceId': DeviceId" in sql:
                # the fixtures' custom output template OT2 uses MAP('k': v), which is not Spark SQL either
                with pytest.raises(dx.DxError):
                    dx.sql_canonical(sql)
                continue
            dx.sql_canonical(sql)
            n += 1
    assert n > 200


def test_sql_parser_refuses_runaway_nesting_with_a_message(dx):
    # the parser and everything that walks its trees recurse: depth is bounded at parse time (a stack overflow would take the host down)
    assert dx.sql_canonical("SELECT " + "(" * 200 + "1" + ")" * 200 + " AS x FROM t") == "SELECT 1 AS x FROM t"
    dx.sql_canonical("SELECT a FROM t WHERE " + " AND ".join(["a > 1"] * 900))
    for sql in ("SELECT " + "(" * 100000 + "1" + ")" * 100000 + " AS x FROM t",
                "SELECT a FROM t WHERE " + "NOT " * 100000 + "a",
                "SELECT " + "-" * 100000 + "1 AS x FROM t",
                "SELECT a FROM t WHERE " + " AND ".join(["a > 1"] * 100000),
                "SELECT " + " + ".join(["a"] * 100000) + " AS x FROM t"):
        with pytest.raises(dx.DxError, match="deeper than"):
            dx.sql_canonical(sql)
