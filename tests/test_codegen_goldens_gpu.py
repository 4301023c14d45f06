"""Every flow the reference's code generator is tested to emit (DataX.Flow.CodegenRules.Tests/CGen*.txt, asserted by
CodegenTests.cs) is handed to the processor unchanged and compared with the CPU oracle table by table over
consecutive micro-batches: simple and aggregate rules, simple and aggregate alerts, metrics, back-ticked dotted
names, the MAP/STRUCT alert shaping."""
import random

import pytest

import codegen_flows as cf
import dxo_flow
from flowcmp import assert_same_rows, same_metrics

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", cf.names())
def test_generated_flow_matches_oracle(dx, ctx, name):
    if name in cf.NOT_SQL:
        # one statement of these two is not SQL (MAP('k': v) from the fixtures' custom template): an error, as in spark.sql
        with pytest.raises(dx.DxError):
            dx.CommonProcessorFactory.createProcessor(ctx, cf.conf_text(name), cf.SCHEMA, cf.PROJECTION, cf.transform_text(name))
        return
    g = dx.CommonProcessorFactory.createProcessor(ctx, cf.conf_text(name), cf.SCHEMA, cf.PROJECTION, cf.transform_text(name))
    o = dxo_flow.Flow(cf.conf_text(name), cf.SCHEMA, cf.PROJECTION, cf.transform_text(name))
    if name in cf.UNSUPPORTED:
        with pytest.raises(dx.DxError, match=cf.UNSUPPORTED[name]):
            g.process_json(cf.events(random.Random(3), 50, cf.T0), cf.T0, cf.I)
        g.close()
        return
    rng = random.Random(3)
    nested = 0
    for b in range(3):
        T = cf.T0 + b * cf.I
        data = cf.events(rng, 300, T)
        rg = g.process_json(data, T, cf.I, now_ms=T + 2000)
        ro, mo = o.process(data, T, cf.I, now_ms=T + 2000)
        assert same_metrics(rg.metrics, mo)
        assert set(rg.table_names) <= set(ro) and len(rg.table_names) >= len(ro) - 1
        for tname in rg.table_names:
            t = rg.table(tname)
            assert_same_rows(t, ro[tname], dx)
            for i, col in enumerate(t.columns):
                path, kinds = t.nesting(i)
                assert (path, kinds) == tuple(ro[tname].nest.get(col, ([], ""))), (tname, col)
                nested += bool(path)
        rg.release()
    if "aa3" in cf.transform_text(name) or "ar3" in cf.transform_text(name):
        assert nested > 0
    g.close()
