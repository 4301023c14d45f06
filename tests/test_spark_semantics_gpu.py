"""The CUDA path (through the C ABI) against the same written rules, one named case each (tests/spark_cases.py)."""
import pytest

import dxo
from flowcmp import gpu_rows
from spark_cases import DECODE, FLOW, same, sort_key

pytestmark = pytest.mark.gpu
SCHEMA2 = dxo.iot_schema_json(2)
T0, I = 1_700_000_040_000, 60_000


@pytest.mark.parametrize("name", sorted(DECODE))
def test_from_json(dx, ctx, name):
    rec, row_ok, want = DECODE[name]
    dec = dx.Decoder(ctx, SCHEMA2)
    t = dec.decode_host(rec.encode("utf-8") + b"\n")
    assert t.num_rows == 1 and dec.stats()["null_rows"] == (0 if row_ok else 1)
    for leaf, exp in want.items():
        v, ok = t.column("deviceDetails." + leaf)
        assert same(v[0] if ok[0] else None, exp), (leaf, v[0], ok[0], exp)
    if not row_ok:
        assert not any(t.column(c)[1][0] for c in t.columns)
    t.release()
    dec.close()


@pytest.mark.parametrize("name", sorted(FLOW))
def test_sql(dx, ctx, name):
    events, sql, want = FLOW[name]
    g = dx.CommonProcessorFactory.createProcessor(ctx, dx.flowgen.conf_text("sem"), SCHEMA2, ["Raw.*\n"], "--DataXQuery--\n" + sql + "\n")
    r = g.process_json(("\n".join(events) + "\n").encode(), T0, I, now_ms=T0 + 1500)
    got = sorted(gpu_rows(r.table("T"), dx), key=sort_key)
    want = sorted(want, key=sort_key)
    assert len(got) == len(want), (got, want)
    for a, b in zip(got, want):
        assert len(a) == len(b) and all(same(x, y) for x, y in zip(a, b)), (got, want)
    r.release()
    g.close()
