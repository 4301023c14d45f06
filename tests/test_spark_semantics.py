"""The oracle against the written Spark 2.4.5 rules, one named case each (tests/spark_cases.py, NEEDS_SPARK_CONFIRMATION.md)."""
import pytest

import dxo
import dxo_flow
from flowcmp import oracle_rows
from spark_cases import DECODE, FLOW, same, sort_key

S2 = dxo.Schema(dxo.iot_schema_json(2))
T0, I = 1_700_000_040_000, 60_000
CONF = "datax.job.name=sem\ndatax.job.input.default.blobschemafile=s.json\ndatax.job.process.projection=p.txt\ndatax.job.process.transform=t.txt\n"


@pytest.mark.parametrize("name", sorted(DECODE))
def test_from_json(name):
    rec, row_ok, want = DECODE[name]
    out = dxo.decode(S2, rec.encode("utf-8") + b"\n")
    assert bool(out["__rowok"][0]) == row_ok
    for leaf, exp in want.items():
        v, ok = out["deviceDetails." + leaf]
        assert same(v[0] if ok[0] else None, exp), (leaf, v[0], ok[0], exp)
    if not row_ok:
        assert not any(out[k][1][0] for k in out if k != "__rowok")          # a NULL struct: every leaf NULL


@pytest.mark.parametrize("name", sorted(FLOW))
def test_sql(name):
    events, sql, want = FLOW[name]
    o = dxo_flow.Flow(CONF, dxo.iot_schema_json(2), ["Raw.*\n"], "--DataXQuery--\n" + sql + "\n")
    res, _ = o.process(("\n".join(events) + "\n").encode(), T0, I, now_ms=T0 + 1500)
    got = sorted(oracle_rows(res["T"]), key=sort_key)
    want = sorted(want, key=sort_key)
    assert len(got) == len(want), (got, want)
    for a, b in zip(got, want):
        assert len(a) == len(b) and all(same(x, y) for x, y in zip(a, b)), (got, want)
