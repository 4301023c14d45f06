"""Second opinion on the oracle: pyarrow's JSON reader and group-by are an independent implementation of
decode and count / sum / min / max / mean for CLEAN records.  They are not Spark (malformed input and NULL semantics
differ), so this is a tripwire for a bug shared by oracle/dxo_core.c and the CUDA decoder, both written from DESIGN.md §3:
the oracle's columns and aggregates must equal pyarrow's on generated S1 / S2 events."""
import io

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc
import pyarrow.json as pj
import pytest

import dxo
import dxo_flow

T0, I = 1_700_000_040_000, 60_000


def gen(shape, n, seed=20260110, keys=500):
    p = dxo.GenParams(seed, keys, max(1, keys // 10), T0, I, 60_000, shape, 10, 1, 0)
    return dxo.gen_events(p, 0, n)


@pytest.mark.parametrize("shape", [1, 2])
def test_oracle_decode_equals_pyarrow_on_clean_events(shape):
    data = gen(shape, 20_000)
    got = dxo.decode(dxo.Schema(dxo.iot_schema_json(shape)), data)
    tab = pj.read_json(io.BytesIO(data)).flatten()            # deviceDetails.* columns
    assert tab.num_rows == 20_000 and got["__rowok"].all()
    for name in tab.column_names:
        col = tab[name].combine_chunks()
        v, ok = got[name]
        assert ok.all(), name
        if pa.types.is_string(col.type):
            assert [x.decode() for x in v] == col.to_pylist(), name
        elif pa.types.is_floating(col.type):
            # both are correctly rounded binary64: bit-identical
            assert (np.asarray(col.to_numpy()).view(np.int64) == v.view(np.int64)).all(), name
        else:
            assert (np.asarray(col.to_numpy()) == v).all(), name


def test_oracle_group_by_equals_pyarrow_on_clean_events():
    data = gen(2, 30_000, keys=300)
    tab = pj.read_json(io.BytesIO(data)).flatten()
    ref = tab.group_by(["deviceDetails.deviceId", "deviceDetails.homeId"]).aggregate(
        [("deviceDetails.status", "count"), ("deviceDetails.status", "sum"), ("deviceDetails.status", "min"), ("deviceDetails.status", "max"),
         ("deviceDetails.temperature", "mean"), ("deviceDetails.temperature", "min"), ("deviceDetails.temperature", "max"),
         ("deviceDetails.humidity", "sum")])
    tr = """--DataXQuery--
G = SELECT deviceDetails.deviceId, deviceDetails.homeId, COUNT(*) AS N, SUM(deviceDetails.status) AS S, MIN(deviceDetails.status) AS Mn,
    MAX(deviceDetails.status) AS Mx, AVG(deviceDetails.temperature) AS Tavg, MIN(deviceDetails.temperature) AS Tmin,
    MAX(deviceDetails.temperature) AS Tmax, SUM(deviceDetails.humidity) AS Hsum FROM DataXProcessedInput
    GROUP BY deviceDetails.deviceId, deviceDetails.homeId
"""
    import dxgpu
    o = dxo_flow.Flow(dxgpu.flowgen.conf_text("pa"), dxo.iot_schema_json(2), ["Raw.*\n"], tr)
    res, _ = o.process(data, T0, I)
    g = res["G"]
    mine = {}
    cols = {c[0]: c for c in g.cols}
    for r in range(g.n):
        mine[(int(cols["deviceId"][2][r]), int(cols["homeId"][2][r]))] = tuple(cols[k][2][r] for k in ("N", "S", "Mn", "Mx", "Tavg", "Tmin", "Tmax", "Hsum"))
    theirs = {}
    d = ref.to_pydict()
    for r in range(ref.num_rows):
        theirs[(d["deviceDetails.deviceId"][r], d["deviceDetails.homeId"][r])] = (
            d["deviceDetails.status_count"][r], d["deviceDetails.status_sum"][r], d["deviceDetails.status_min"][r], d["deviceDetails.status_max"][r],
            d["deviceDetails.temperature_mean"][r], d["deviceDetails.temperature_min"][r], d["deviceDetails.temperature_max"][r],
            d["deviceDetails.humidity_sum"][r])
    assert mine.keys() == theirs.keys()
    assert g.approx == {"Tavg", "Hsum"}                    # what a comparison may be tolerant about, and nothing else
    for k, a in mine.items():
        b = theirs[k]
        assert tuple(int(x) for x in a[:4]) == tuple(b[:4]), k           # counts, integer sums, integer min / max: exact
        assert a[5] == b[5] and a[6] == b[6], k                           # double MIN / MAX: exact
        for x, y in ((a[4], b[4]), (a[7], b[7])):                         # order-dependent double sums: the stated tolerance
            assert abs(x - y) <= 1e-6 * max(abs(x), abs(y)), k
