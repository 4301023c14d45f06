"""CPU side of the code-generator goldens: the oracle executes every generated flow the reference's CodegenTests.cs
expects (tests/golden/codegen), and the alert shaping (MAP / STRUCT select items) comes out as nested leaves."""
import random

import pytest

import codegen_flows as cf
import dxo_flow


@pytest.mark.parametrize("name", cf.names())
def test_oracle_runs_generated_flow(name):
    if name in cf.NOT_SQL:
        with pytest.raises(ValueError):
            dxo_flow.Flow(cf.conf_text(name), cf.SCHEMA, cf.PROJECTION, cf.transform_text(name)).process(b"{}\n", cf.T0, cf.I)
        return
    o = dxo_flow.Flow(cf.conf_text(name), cf.SCHEMA, cf.PROJECTION, cf.transform_text(name))
    if name in cf.UNSUPPORTED:
        with pytest.raises(ValueError, match=cf.UNSUPPORTED[name]):
            o.process(cf.events(random.Random(3), 50, cf.T0), cf.T0, cf.I)
        return
    rng = random.Random(3)
    rows = {}
    for b in range(3):
        T = cf.T0 + b * cf.I
        ro, _ = o.process(cf.events(rng, 300, T), T, cf.I, now_ms=T + 2000)
        for k, v in ro.items():
            rows[k] = rows.get(k, 0) + v.n
    alerts = [k for k in rows if k.endswith("Alert") or k in ("T1", "T2", "DeviceCount", "House150GarageTotalTimeOpen")]
    assert alerts and all(rows[k] > 0 for k in alerts), rows


def test_alert_shaping_is_nested_leaves():
    o = dxo_flow.Flow(cf.conf_text("CGenAggRulDerTable.txt"), cf.SCHEMA, cf.PROJECTION, cf.transform_text("CGenAggRulDerTable.txt"))
    ro, _ = o.process(cf.events(random.Random(1), 200, cf.T0), cf.T0, cf.I)
    t = ro["ar3_2_1"]
    assert t.names()[:2] == ["pivots.DeviceId", "aggs.Humidity.AVG"] or "pivots.DeviceId" in t.names()
    assert t.nest["pivots.DeviceId"] == (["DeviceId"], "m")
    nested = [n for n in t.names() if n.startswith("aggs.")]
    assert nested and all(t.nest[n][1] == "mm" for n in nested)
    assert ro["T1"].nest == ro["ar4_2"].nest and ro["T1"].nest          # SELECT * keeps the shaping
