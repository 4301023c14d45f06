"""The oracle's restatement of the sink hand-off's to_json (oracle/dxo_json.py), pinned on what can be pinned without a JVM:
java.lang.Double.toString's documented outputs, Jackson's escape table, the timestamp pattern, and the struct / map / array rules
on small hand-built tables.  (Parity with Spark itself stays unpinned: NEEDS_SPARK_CONFIRMATION.md, rows J1-J8.)"""
import datetime
import json
import math
import random
import struct

import numpy as np

import dxo_flow
import dxo_json


def test_java_double_to_string_documented_values():
    # values from the Javadoc of Double.toString and from the JDK's own regression tests for it
    kat = {
        1.0: "1.0", 100.0: "100.0", 9999999.0: "9999999.0", 1.0e7: "1.0E7", 12345678.9: "1.23456789E7", 1.0e-3: "0.001",
        9.999e-4: "9.999E-4", 123456.789: "123456.789", 1.0e-5: "1.0E-5", 1.7976931348623157e308: "1.7976931348623157E308",
        4.9e-324: "4.9E-324", 0.1 + 0.2: "0.30000000000000004", 1e21: "1.0E21", 1e8: "1.0E8", 0.5: "0.5", 3.0e10: "3.0E10",
        23.45: "23.45", 2.2250738585072014e-308: "2.2250738585072014E-308", 1e23: "1.0E23", 100.125: "100.125", 0.001234: "0.001234",
    }
    for x, s in kat.items():
        assert dxo_json.java_double_string(x) == s, (x, s, dxo_json.java_double_string(x))
        assert dxo_json.java_double_string(-x) == "-" + s
    assert dxo_json.java_double_string(0.0) == "0.0" and dxo_json.java_double_string(-0.0) == "-0.0"
    assert dxo_json.java_double_string(float("nan")) == "NaN"
    assert dxo_json.java_double_string(float("inf")) == "Infinity" and dxo_json.java_double_string(float("-inf")) == "-Infinity"


def test_java_double_string_round_trips_and_is_shortest():
    rnd = random.Random(5)
    for _ in range(20000):
        x = struct.unpack("<d", struct.pack("<Q", rnd.getrandbits(64)))[0]
        if x != x or math.isinf(x):
            continue
        s = dxo_json.java_double_string(x)
        assert float(s) == x
        mant = s.split("E")[0].replace("-", "").replace(".", "").strip("0")
        assert len(mant) <= 17


def test_jackson_escapes():
    assert dxo_json.jackson_quote(b'a"b\\c') == b'"a\\"b\\\\c"'
    assert dxo_json.jackson_quote(bytes([8, 9, 10, 12, 13])) == b'"\\b\\t\\n\\f\\r"'
    assert dxo_json.jackson_quote(bytes([0, 1, 0x1F, 0x7F])) == b'"\\u0000\\u0001\\u001F\x7f"'
    assert dxo_json.jackson_quote("é/ü".encode()) == '"é/ü"'.encode()           # non-ASCII and '/' stay as they are
    for raw in (b"plain", b'q"\\\n\t\x01', "hé".encode()):
        assert json.loads(dxo_json.jackson_quote(raw)) == raw.decode()


def test_timestamp_patterns():
    assert dxo_json.json_timestamp(0) == "1970-01-01T00:00:00.000Z"
    us = 1_700_000_040_123_456
    d = datetime.datetime.fromtimestamp(us // 1_000_000, datetime.timezone.utc)
    assert dxo_json.json_timestamp(us) == d.strftime("%Y-%m-%dT%H:%M:%S") + ".123Z"
    assert dxo_json.json_timestamp(-1) == "1969-12-31T23:59:59.999Z"              # floor, not truncation
    assert dxo_json.cast_timestamp_string(us) == d.strftime("%Y-%m-%d %H:%M:%S") + ".123456"
    assert dxo_json.cast_timestamp_string(1_700_000_040_500_000).endswith(":00.5")
    assert dxo_json.cast_timestamp_string(1_700_000_040_000_000).endswith(":00")


def _tab(n):
    return dxo_flow.Tab(n)


def test_struct_skips_nulls_map_writes_them_and_types_unify():
    t = _tab(2)
    t.add("id", "long", np.array([7, 8]), np.array([True, False]))
    t.add("name", "string", np.array([b"x\"y", b"z"], dtype=object), np.array([True, True]))
    t.add("__DataX_FileInfo.inputPath", "string", np.array([b"p", b"p"], dtype=object), np.array([True, True]))
    t.add("deviceDetails.temperature", "double", np.array([20.5, 0.0]), np.array([True, False]))
    t.add("deviceDetails.ok", "bool", np.array([1, 0]), np.array([True, False]))
    t.add("m.a", "long", np.array([1, 2]), np.array([True, False]))
    t.add("m.b", "string", np.array([b"s", b"t"], dtype=object), np.array([True, True]))
    t.add("m.c", "double", np.array([1e7, 2.5]), np.array([True, True]))
    t.add("s.k.x", "long", np.array([5, 6]), np.array([False, True]))
    t.nest = {"m.a": (["a"], "m"), "m.b": (["b"], "m"), "m.c": (["c"], "m"), "s.k.x": (["k", "x"], "ss")}
    rows = dxo_json.to_json_rows(t)
    assert rows[0] == b'{"id":7,"name":"x\\"y","deviceDetails":{"temperature":20.5,"ok":true},"m":{"a":"1","b":"s","c":"1.0E7"},"s":{"k":{}}}'
    assert rows[1] == b'{"name":"z","m":{"a":null,"b":"t","c":"2.5"},"s":{"k":{"x":6}}}'
    for r in rows:
        json.loads(r)


def test_rules_array_and_single_map():
    t = _tab(3)
    t.add("Rules", "rules", np.array([[True, False, True], [False, False, False], [False, True, False]]), np.ones(3, bool))
    t.rule_table = [{"ruleId": "R0", "severity": "Critical"}, {"ruleId": "R1", "severity": "Low"}, {"ruleId": "R\"2", "severity": "x"}]
    rows = dxo_json.to_json_rows(t)
    assert rows[0] == b'{"Rules":[{"ruleId":"R0","severity":"Critical"},{"ruleId":"R\\"2","severity":"x"}]}'
    assert rows[1] == b'{"Rules":[]}'
    t1 = _tab(2)
    t1.add("r", "rules", np.array([[True], [False]]), np.ones(2, bool))
    t1.rule_table = [{"ruleId": "A"}]
    t1.rule_single = True
    assert dxo_json.to_json_rows(t1) == [b'{"r":{"ruleId":"A"}}', b"{}"]


def test_numeric_map_widens_long_to_double():
    t = _tab(1)
    t.add("m.a", "long", np.array([3]), np.array([True]))
    t.add("m.b", "double", np.array([0.25]), np.array([True]))
    t.nest = {"m.a": (["a"], "m"), "m.b": (["b"], "m")}
    assert dxo_json.to_json_rows(t) == [b'{"m":{"a":3.0,"b":0.25}}']
