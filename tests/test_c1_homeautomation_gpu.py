"""BASELINE config #1: the reference's local sample flow (DeploymentLocal/sample/HomeAutomationLocal.json) —
projection + 5-minute window + reference-data join + user function + rules -> alerts — run through the
processor interface exactly as the reference loads it (conf naming the schema / projection / transform /
reference-data files), against the CPU oracle on the same 120-event micro-batches."""
import json
import os
import random

import pytest

import dxo_flow
from flowcmp import assert_same_rows, same_metrics

pytestmark = pytest.mark.gpu
C1 = os.path.join(os.path.dirname(__file__), "golden", "c1_homeautomation")
I = 60_000
T0 = 1_700_000_040_000


def who_opened(s):           # datax.sample.udf.UdfHelloWorld.call: "hello, " + p1
    return "hello, " + str(s if s is not None else "null")


def events(rng, n, now_ms):
    out = []
    for _ in range(n):
        out.append(json.dumps({"deviceDetails": {"deviceId": rng.choice([1, 2, 3, 4, 5, 6]),
                                                 "deviceType": rng.choice(["DoorLock", "WindowLock", "Heating"]),
                                                 "eventTime": now_ms - rng.randrange(1000), "homeId": rng.choice([32, 150, 25, 81]),
                                                 "status": rng.choice([0, 1])}}, separators=(",", ":")))
    return ("\n".join(out) + "\n").encode()


def test_home_automation_local_flow(dx, ctx):
    rd = lambda f: open(os.path.join(C1, f), encoding="utf-8-sig").read()
    g = dx.CommonProcessorFactory.createProcessorFromFiles(ctx, os.path.join(C1, "HomeAutomationLocal.conf"))
    g.register_udf("whoOpened", who_opened)
    o = dxo_flow.Flow(rd("HomeAutomationLocal.conf"), rd("inputschema.json"), [rd("projection.txt")], rd("HomeAutomationLocal-combined.txt"))
    o.add_refdata("myDevicesRefData", rd("devices.csv"))
    dxo_flow.UDFS["WHOOPENED"] = who_opened
    rng = random.Random(1)
    for b in range(100):            # 100 batches x 120 events (2 events/s, 60 s interval)
        T = T0 + b * I
        now = T + 1500 + b          # current_timestamp() of the batch: after the batch time, as in a live run
        data = events(rng, 120, now)
        rg = g.process_json(data, T, I, now_ms=now)
        ro, mo = o.process(data, T, I, now_ms=now)
        assert same_metrics(rg.metrics, mo)
        names = rg.table_names
        assert {"DeviceWindowedInput", "DeviceNotWindowedInputWithNameAndWho", "DeviceInfoTimeWindow", "DoorUnlockedSimpleAlert",
                "WindowOpenFiveMinWhileHeaterOnCombinedAlert", "Rules", "CLOSEAlert"} <= set(names)
        for name in names:
            assert_same_rows(rg.table(name), ro[name], dx)
        rg.release()
    # the sample flow does produce alerts on this data
    assert ro["DoorUnlockedSimpleAlert"].n > 0 and ro["CLOSEAlert"].n == 1
    g.close()


def test_home_automation_flow_from_its_gui_json(dx, ctx):
    """The same flow with its transform GENERATED from the flow's GUI JSON by the restated code generator (dx_codegen_flow:
    S450_PrepareTransformFile + RuleDefinitionGenerator + Engine.GenerateCode with the default templates) instead of the
    hand-derived text: the processor and the oracle both take the generated transform and must agree on every table."""
    from test_codegen_restatement import flow_transform
    code, _ = flow_transform()
    rd = lambda f: open(os.path.join(C1, f), encoding="utf-8-sig").read()
    g = dx.CommonProcessorFactory.createProcessor(ctx, rd("HomeAutomationLocal.conf"), rd("inputschema.json"), [rd("projection.txt")], code)
    g.add_refdata("myDevicesRefData", rd("devices.csv"))
    g.register_udf("whoOpened", who_opened)
    o = dxo_flow.Flow(rd("HomeAutomationLocal.conf"), rd("inputschema.json"), [rd("projection.txt")], code)
    o.add_refdata("myDevicesRefData", rd("devices.csv"))
    dxo_flow.UDFS["WHOOPENED"] = who_opened
    rng = random.Random(2)
    for b in range(30):
        T = T0 + b * I
        now = T + 1500 + b
        data = events(rng, 120, now)
        rg = g.process_json(data, T, I, now_ms=now)
        ro, mo = o.process(data, T, I, now_ms=now)
        assert same_metrics(rg.metrics, mo)
        assert {"DeviceWindowedInput", "DeviceInfoTimeWindow", "Rules", "CLOSEAlert"} <= set(rg.table_names)
        for name in rg.table_names:
            assert_same_rows(rg.table(name), ro[name], dx)
        rg.release()
    g.close()
