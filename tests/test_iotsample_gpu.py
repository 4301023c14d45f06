"""The reference's IoT sample flow, run from the reference's OWN generated transform text: the golden output of its
code generator (CodegenRules.Tests/CGenIoTSample.txt) is handed to the processor unchanged, together with the
config-generation goldens for projection and schema — windows, reference-data join, user function, state table,
COUNT(DISTINCT), joins between derived tables, alerts.  Checked against the CPU oracle over consecutive micro-batches."""
import datetime
import json
import os
import random

import pytest

import dxo_flow
from flowcmp import assert_same_rows, same_metrics

pytestmark = pytest.mark.gpu
D = os.path.join(os.path.dirname(__file__), "golden", "iotsample")
I = 60_000
T0 = 1_700_000_040_000


def who_opened(s):
    return "hello, " + str(s if s is not None else "null")


def events(rng, n, batch_ms):
    out = []
    for _ in range(n):
        t = datetime.datetime.fromtimestamp((batch_ms - 1 - rng.randrange(59_000)) / 1000, tz=datetime.timezone.utc)
        out.append(json.dumps({"deviceDetails": {"deviceId": rng.choice([1, 2, 3, 4, 6, 8]),
                                                 "deviceType": rng.choice(["DoorLock", "WindowLock", "Heating", "GarageDoorLock"]),
                                                 "eventTime": t.strftime("%m/%d/%Y %H:%M:%S"), "homeId": rng.choice([150, 150, 32, 25]),
                                                 "status": rng.choice([0, 1])}}, separators=(",", ":")))
    return ("\n".join(out) + "\n").encode()


def test_reference_generated_iot_sample_flow(dx, ctx):
    rd = lambda f: open(os.path.join(D, f), encoding="utf-8-sig").read()
    g = dx.CommonProcessorFactory.createProcessorFromFiles(ctx, os.path.join(D, "iotsample.conf"))
    g.register_udf("whoOpened", who_opened)
    o = dxo_flow.Flow(rd("iotsample.conf"), rd("inputschema.json"), [rd("projection.txt")], rd("iotsample-combined.txt"))
    o.add_refdata("myDevicesRefdata", rd("devices.csv"))
    dxo_flow.UDFS["WHOOPENED"] = who_opened
    rng = random.Random(5)
    produced = set()
    for b in range(10):
        T = T0 + b * I
        data = events(rng, 200, T)
        rg = g.process_json(data, T, I, now_ms=T + 2000)
        ro, mo = o.process(data, T, I, now_ms=T + 2000)
        assert same_metrics(rg.metrics, mo)
        assert len(rg.table_names) >= 28
        for name in rg.table_names:
            assert_same_rows(rg.table(name), ro[name], dx)
            if ro[name].n:
                produced.add(name)
        st, which = g.state_table("iotsample_GarageDoor_status_accumulated")
        assert which == o.state["iotsample_GarageDoor_status_accumulated"]["active"]
        rg.release()
    # the statements that carry the sample's alerts and metrics all produced rows at some point
    assert {"DoorUnlockedSimpleAlert", "GarageDoorAccumalator", "GarageDoorOpenInAnHour", "CLOSEAlert", "DeviceInfoTimeWindow",
            "WindowLockWindowedFiltered", "HeaterStateOneIsOn"} <= produced
    g.close()
