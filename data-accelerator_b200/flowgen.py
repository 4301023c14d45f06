"""Flow artefacts for the benchmark / parity configurations.

The reference produces these artefacts at design time (DataX.Config S450-S650 + CodegenRules);
that generator is a "next" row of the scope table, so the few templates needed here are written
out by hand in exactly the shape the reference emits:
  * ProcessRules  -> SELECT *, filterNull(Array(IF(cond, MAP(...), NULL), ...)) AS Rules FROM t
    (CodegenRules Engine.cs:232-266, 380-397; Rule.RulesObject Rule.cs:272-281;
     golden: DataX.Flow.CodegenRules.Tests/cgenDefault.txt:2-36)
  * <flow>.conf keys as flattened by flattenerConfig.json (golden: Local.Test/Resource/jobConfig.conf)
"""
import json
import math
import random

TYPES = ["DoorLock", "WindowLock", "Heating", "GarageDoorLock"]


def iot_schema_json(shape=2, event_time_type="long"):
    """Spark StructType JSON of the home-automation event (DeploymentLocal/sample/HomeAutomationLocal.json:288), plus the
    two double fields of shape S2 (DeploymentLocal/sample/BasicLocal.json:185)."""
    f = [{"name": "deviceId", "type": "long", "nullable": True, "metadata": {}},
         {"name": "deviceType", "type": "string", "nullable": True, "metadata": {}},
         {"name": "eventTime", "type": event_time_type, "nullable": True, "metadata": {}},
         {"name": "homeId", "type": "long", "nullable": True, "metadata": {}},
         {"name": "status", "type": "long", "nullable": True, "metadata": {}}]
    if shape == 2:
        f += [{"name": "temperature", "type": "double", "nullable": True, "metadata": {}},
              {"name": "humidity", "type": "double", "nullable": True, "metadata": {}}]
    return json.dumps({"type": "struct", "fields": [
        {"name": "deviceDetails", "type": {"type": "struct", "fields": f}, "nullable": True, "metadata": {}}]})


def sql_str(s):
    return "'" + s.replace("\\", "\\\\").replace("'", "\\'") + "'"


def make_rules(n_rules=200, seed=20260102, n_keys=1_000_000, n_homes=100_000):
    """C2 recipe: 60 % single thresholds, 30 % two-leaf AND/OR, 10 % three-leaf with a parenthesised
    group incl. deviceType = '...'; constants drawn so per-rule selectivity is log-uniform in [1e-4, 0.5]."""
    rng = random.Random(seed)

    def sel():
        return math.exp(rng.uniform(math.log(1e-4), math.log(0.5)))

    def leaf():
        k = rng.randrange(6)
        s = sel()
        if k == 0:   # temperature uniform in [5.10, 100.10)
            return "deviceDetails.temperature > %.2f" % (100.10 - s * 95.0)
        if k == 1:   # humidity uniform in [0, 100)
            return "deviceDetails.humidity <= %.2f" % (s * 100.0)
        if k == 2:
            return "deviceDetails.homeId = %d" % rng.randrange(1, n_homes + 1)
        if k == 3:
            return "deviceDetails.status = %d" % rng.randrange(2)
        if k == 4:
            return "deviceDetails.deviceId < %d" % max(2, int(s * n_keys))
        return "deviceDetails.deviceType = '%s'" % rng.choice(TYPES)

    rules = []
    for r in range(n_rules):
        u = r % 10
        if u < 6:
            cond = leaf()
        elif u < 9:
            cond = "%s %s %s" % (leaf(), rng.choice(["AND", "OR"]), leaf())
        else:
            cond = "%s AND (%s OR deviceDetails.deviceType = '%s')" % (leaf(), leaf(), rng.choice(TYPES))
        rules.append({"ruleId": "R%d" % (r + 1), "ruleDescription": "synthetic rule %d" % (r + 1),
                      "severity": rng.choice(["Critical", "Medium", "Low"]), "tagName": "Tag", "tag": "T%d" % (r % 7),
                      "condition": cond})
    return rules


def process_rules_sql(rules, target="Rules", table="DataXProcessedInput"):
    parts = []
    for r in rules:
        parts.append("IF(%s, MAP('ruleId', %s, 'ruleDescription', %s, 'severity', %s, %s, %s), NULL)" % (
            r["condition"], sql_str(r["ruleId"]), sql_str(r["ruleDescription"]), sql_str(r["severity"]),
            sql_str(r["tagName"]), sql_str(r["tag"])))
    return "--DataXQuery--\n%s = SELECT *, filterNull(Array(%s)) AS Rules FROM %s\n" % (target, ", ".join(parts), table)


def conf_text(name, timestamp_column=None, watermark=None, windows=(), state_tables=(), outputs=()):
    lines = ["datax.job.name=%s" % name,
             "datax.job.input.default.blobschemafile=inputschema.json",
             "datax.job.process.projection=projection.txt",
             "datax.job.process.transform=%s-combined.txt" % name]
    if timestamp_column:
        lines.append("datax.job.process.timestampcolumn=%s" % timestamp_column)
    if watermark:
        lines.append("datax.job.process.watermark=%s" % watermark)
    for wname, dur in windows:
        lines.append("datax.job.process.timewindow.%s.windowduration=%s" % (wname, dur))
    for tname, schema in state_tables:
        lines.append("datax.job.process.statetable.%s.schema=%s" % (tname, schema))
        lines.append("datax.job.process.statetable.%s.location=hdfs://local/%s/%s/" % (tname, name, tname))
    for o in outputs:
        lines.append("datax.job.output.%s.blob.compressiontype=none" % o)
    return "\n".join(lines) + "\n"
