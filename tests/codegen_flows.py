"""Harness for running the reference's code-generator goldens (tests/golden/codegen/CGen*.txt) as flows.

The goldens are the transform text only; CodegenTests.cs never executes them.  To execute one, a flow needs an input
schema, a projection and — where the rule set targets a user-defined table (`ProcessAggregateAlerts(Foo)`,
UserCode*.txt) — the user's statement defining that table, which the goldens do not contain.  Those pieces are
supplied here, identically for the product and the oracle; the golden text itself is appended unchanged."""
import json
import os

D = os.path.join(os.path.dirname(__file__), "golden", "codegen")
I = 60_000
T0 = 1_700_000_040_000


def _f(name, ty):
    return {"name": name, "type": ty, "nullable": True, "metadata": {}}


def _s(*fields):
    return {"type": "struct", "fields": list(fields)}


# superset of every column the goldens read from DataXProcessedInput
SCHEMA = json.dumps(_s(
    _f("DeviceId", "long"), _f("Geo", "string"), _f("Temperature", "double"), _f("Humidity", "double"),
    _f("device", _s(_f("Temperature", "double"))),
    _f("deviceDetails", _s(_f("deviceId", "long"), _f("deviceType", "string"), _f("eventTime", "string"),
                           _f("homeId", "long"), _f("status", "long")))))
PROJECTION = ["Raw.*\ncurrent_timestamp() AS eventTimeStamp\n"]

# user statements the rule sets hang off (UserCodeAggAlDerTable.txt: ProcessAggregateAlerts(Foo) ...)
PREAMBLE = {
    "Foo": "--DataXQuery--\nFoo = SELECT DeviceId, Geo, Temperature, Humidity FROM DataXProcessedInput\n",
    "Bar": "--DataXQuery--\nBar = SELECT DeviceId, Geo, Temperature, Humidity FROM DataXProcessedInput\n",
}
# statements that are not Spark SQL (the fixtures' custom output template writes MAP('k': v)): spark.sql would throw
NOT_SQL = {"CGenNoCode.txt", "CGenSimRulNonTable.txt"}
# (CGenCreateMetric2 unions a string column with the state table's timestamp column: widened to string, as Spark does)
UNSUPPORTED = {}


def names():
    return sorted(f for f in os.listdir(D) if f.startswith("CGen") and f.endswith(".txt"))


def transform_text(name):
    text = open(os.path.join(D, name), encoding="utf-8-sig").read()
    pre = "".join(v for k, v in PREAMBLE.items() if (" " + k + " ") in text.replace("\n", " ").replace("(", " ").replace(")", " "))
    return pre + text


def conf_text(name):
    c = ["datax.job.name=" + name[:-4], "datax.job.process.timestampcolumn=eventTimeStamp", "datax.job.process.watermark=0 second"]
    if name == "CGenCreateMetric2.txt":
        c.append("datax.job.process.statetable.iotdevicesample_GarageDoor_accumulated.schema="
                 "deviceId long, deviceType string, homeId long, eventTime Timestamp, status long")
        c.append("datax.job.process.statetable.iotdevicesample_GarageDoor_accumulated.location=hdfs://x/")
    return "\n".join(c) + "\n"


def events(rng, n, batch_ms):
    out = []
    for _ in range(n):
        r = {"DeviceId": rng.choice([1, 2, 3, 7]), "Geo": rng.choice(["us", "eu", "ap"]),
             "Temperature": rng.choice([20.5, 39.0, 55.25, 85.0, 91.5, 99.0]), "Humidity": rng.choice([10.0, 35.5, 41.0, 77.0]),
             "device": {"Temperature": rng.choice([12.0, 39.5, 40.0, 41.5, 90.0])},
             "deviceDetails": {"deviceId": rng.choice([1, 2, 3, 4, 6, 8]), "deviceType": rng.choice(["DoorLock", "GarageDoorLock", "Heating"]),
                               "eventTime": "2023-11-14 22:%02d:%02d" % (rng.randrange(60), rng.randrange(60)),
                               "homeId": rng.choice([150, 150, 32, 25]), "status": rng.choice([0, 1, 1])}}
        if r["DeviceId"] == 7:          # one device runs hot and humid so the aggregate alerts fire
            r["Temperature"], r["Humidity"] = rng.choice([85.0, 91.5, 99.0]), rng.choice([41.0, 77.0])
        elif rng.random() < 0.05:
            r["Temperature"] = None
        if rng.random() < 0.05:
            del r["deviceDetails"]["status"]
        out.append(json.dumps(r, separators=(",", ":")))
    return ("\n".join(out) + "\n").encode()
