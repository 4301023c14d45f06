"""Oracle (CPU restatement of Spark 2.4.5 from_json, CommonProcessorFactory.scala:93) against
ground truth it does not compute itself, hand-derived expectations and the reference's sample events."""
import json
import math
import os

import numpy as np

import dxo
from adversarial import to_ndjson

S2 = dxo.Schema(dxo.iot_schema_json(2))
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def params(shape=2, n_keys=1000, seed=20260102):
    return dxo.GenParams(seed, n_keys, 100, 1_700_000_000_000, 60_000, 60_000, shape, 10, 1, 0)


def test_generated_events_decode_to_their_fields():
    p = params()
    n = 20000
    data = dxo.gen_events(p, 5, n)
    truth = dxo.gen_fields(p, 5, n)
    got = dxo.decode(S2, data)
    assert got["__rowok"].all() and len(got["__rowok"]) == n
    for f in ("deviceId", "eventTime", "homeId", "status"):
        v, ok = got["deviceDetails." + f]
        assert ok.all() and (v == truth[f]).all(), f
    for f in ("temperature", "humidity"):
        v, ok = got["deviceDetails." + f]
        assert ok.all() and (v.view(np.int64) == truth[f].view(np.int64)).all(), f
    s, ok = got["deviceDetails.deviceType"]
    assert ok.all() and all(s[i] == dxo.TYPES[truth["deviceType"][i]] for i in range(n))
    # python's json agrees on every record (independent decoder, clean input)
    for i, line in enumerate(data.decode().splitlines()[:2000]):
        d = json.loads(line)["deviceDetails"]
        assert d["deviceId"] == truth["deviceId"][i] and d["temperature"] == truth["temperature"][i]


def one(rec, schema=S2):
    out = dxo.decode(schema, to_ndjson([rec]))
    assert len(out["__rowok"]) == 1
    return {k.split(".")[-1]: (v[0][0] if v[1][0] else None) for k, v in out.items() if k != "__rowok"}, bool(out["__rowok"][0])


def test_spark24_semantics_hand_cases():
    # DESIGN.md §3 — every expectation below is the written rule, not an output of the oracle
    row, ok = one('{"deviceDetails":{"deviceId":"","status":1}}')
    assert ok and row["deviceId"] is None and row["status"] == 1            # "" for long -> field NULL
    row, ok = one('{"deviceDetails":{"deviceId":"12","status":1}}')
    assert not ok and row["status"] is None                                   # non-empty string -> row NULL
    row, ok = one('{"deviceDetails":{"deviceId":1.0}}')
    assert not ok                                                             # float token for long
    row, ok = one('{"deviceDetails":{"deviceId":9223372036854775808}}')
    assert not ok                                                             # long overflow
    row, ok = one('{"deviceDetails":{"deviceId":-9223372036854775808}}')
    assert ok and row["deviceId"] == -2**63
    row, ok = one('{"deviceDetails":{"temperature":"NaN","humidity":"-Infinity"}}')
    assert ok and math.isnan(row["temperature"]) and row["humidity"] == -math.inf
    row, ok = one('{"deviceDetails":{"temperature":"nan"}}')
    assert not ok                                                             # 2.4.5 matches exact case only
    row, ok = one('{"deviceDetails":{"temperature":""}}')
    assert not ok                                                             # DoubleType's VALUE_STRING case throws
    row, ok = one('{"deviceDetails":{"deviceType":123,"deviceId":1}}')
    assert ok and row["deviceType"] == b"123"                                 # scalar token text into string field
    row, ok = one('{"deviceDetails":{"deviceId":1,"deviceId":2}}')
    assert ok and row["deviceId"] == 2                                        # duplicate key: last wins
    row, ok = one('{"deviceDetails":{"deviceId":1},"deviceDetails":null}')
    assert ok and row["deviceId"] is None
    row, ok = one('{"DeviceDetails":{"deviceId":1}}')
    assert ok and row["deviceId"] is None                                     # case-sensitive names
    row, ok = one('{"deviceDetails":{"deviceId":9}} trailing garbage')
    assert ok and row["deviceId"] == 9                                        # bytes after the first value ignored
    row, ok = one('{"deviceDetails":{"deviceId":9,}}')
    assert not ok
    row, ok = one('[{"deviceDetails":{"deviceId":11}}]')
    assert ok and row["deviceId"] == 11
    row, ok = one('[{"deviceDetails":{"deviceId":11}},{"deviceDetails":{"deviceId":12}}]')
    assert not ok
    row, ok = one('')
    assert not ok
    row, ok = one('{"deviceDetails":{"deviceType":"\\u00e9\\ud83d\\ude00"}}')
    assert ok and row["deviceType"] == "é😀".encode()
    row, ok = one("{'deviceDetails':{'deviceId':5}}")
    assert ok and row["deviceId"] == 5                                        # allowSingleQuotes default true
    row, ok = one('{"deviceDetails":{"deviceId":-0,"temperature":-0}}')
    assert ok and row["deviceId"] == 0 and math.copysign(1, row["temperature"]) == 1.0
    row, ok = one('{"deviceDetails":{"temperature":-0.0}}')
    assert ok and math.copysign(1, row["temperature"]) == -1.0
    row, ok = one('{"extra":[1,2,,3],"deviceDetails":{"deviceId":4}}')
    assert not ok                                                             # skipped values are still validated


def test_reference_sample_events():
    """Events shipped with the reference (DeploymentCloud/.../samples/iotDevice/iotsample.json:52,77,102 and
    iotsample-091610A6B869B8C9318E988F390F46AB.json), extracted by tests/golden/make_golden.py."""
    lines = open(os.path.join(GOLD, "reference_sample_events.ndjson"), "rb").read()
    expected = json.load(open(os.path.join(GOLD, "reference_sample_events.expected.json")))
    schema = dxo.Schema(dxo.iot_schema_json(1, event_time_type="string"))
    got = dxo.decode(schema, lines)
    n = len(expected)
    assert len(got["__rowok"]) == n and got["__rowok"].all()
    for i, e in enumerate(expected):
        assert got["deviceDetails.deviceId"][0][i] == e["deviceId"]
        assert got["deviceDetails.deviceType"][0][i] == e["deviceType"].encode()
        assert got["deviceDetails.eventTime"][0][i] == e["eventTime"].encode()
        assert got["deviceDetails.homeId"][0][i] == e["homeId"]
        assert got["deviceDetails.status"][0][i] == e["status"]


def test_double_rounding_against_python_float():
    import random
    from adversarial import random_number
    rng = random.Random(3)
    nums = [random_number(rng) for _ in range(20000)]
    recs = ['{"deviceDetails":{"temperature":%s}}' % s for s in nums]
    got = dxo.decode(S2, to_ndjson(recs))
    v, ok = got["deviceDetails.temperature"]
    for i, s in enumerate(nums):
        assert ok[i], s
        exp = float(s)
        if "." not in s and "e" not in s and exp == 0:
            exp = 0.0
        assert v[i] == exp or (math.isnan(v[i]) and math.isnan(exp)), (s, v[i], exp)
