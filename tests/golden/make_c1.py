#!/usr/bin/env python3
"""Derives the runtime artefacts of the reference's local sample flow (BASELINE config #1) from
DeploymentLocal/sample/HomeAutomationLocal.json, applying by hand the rewrites the reference's design-time
code generator performs (no .NET here): TIMEWINDOW('5 minutes') -> DataXProcessedInput_5minutes
(CodegenRules Engine.cs:595-626), statement terminators removed, OUTPUT ... TO removed (Engine.cs:495-556),
ProcessRules -> filterNull(Array(IF(cond, MAP(...), NULL))) (Engine.cs:232-266), ProcessAlerts -> sa1/sa2/<tag>Alert
(Engine.cs:200-230, golden CGenIoTSample.txt:384-413).  Run in the build container; the outputs are committed."""
import json
import os
import re

REF = "/root/reference/DeploymentLocal/sample/HomeAutomationLocal.json"
CSV = "/root/reference/DeploymentCloud/Deployment.DataX/Samples/usercontent/devices.csv"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "c1_homeautomation")

gui = json.load(open(REF, encoding="utf-8-sig"))["gui"]
props = gui["input"]["properties"]
open(os.path.join(OUT, "inputschema.json"), "w").write(props["inputSchemaFile"])
open(os.path.join(OUT, "projection.txt"), "w").write(props["normalizationSnippet"] + "\n")
sql = gui["process"]["queries"][0]
sql = re.sub(r"(\w+)\s+TIMEWINDOW\('5 minutes'\)", r"\1_5minutes", sql)
sql = re.sub(r"^OUTPUT .*$", "", sql, flags=re.M)
sql = sql.replace(";", "")
rules = []
for i, r in enumerate(gui["rules"]):
    p = r["properties"]
    rules.append({"ruleId": "R%d" % (i + 1), "ruleDescription": p["_S_ruleDescription"], "severity": p["_S_severity"],
                  "tagName": p["_S_tagName"], "tag": p["_S_tag"], "condition": p["_S_condition"], "isAlert": p["_S_isAlert"]})
q = lambda s: "'" + s.replace("'", "\\'") + "'"
rule_sql = "Rules = SELECT *, filterNull(Array(%s)) AS Rules FROM DataXProcessedInput" % ", ".join(
    "IF(%s, MAP('ruleId', %s, 'ruleDescription', %s, 'severity', %s, %s, %s), NULL)" % (
        r["condition"], q(r["ruleId"]), q(r["ruleDescription"]), q(r["severity"]), q(r["tagName"]), q(r["tag"])) for r in rules)
alert_sql = ""
k = 0
for r in rules:
    if not r["isAlert"]:
        continue
    k += 1
    alert_sql += ("--DataXQuery--\nsa1_1_%d = SELECT *, %s AS ruleId, %s AS ruleDescription, %s AS severity, %s AS %s FROM DataXProcessedInput WHERE %s\n"
                  % (k, q(r["ruleId"]), q(r["ruleDescription"]), q(r["severity"]), q(r["tag"]), r["tagName"], r["condition"]))
    alert_sql += "--DataXQuery--\nsa2_1_%d = SELECT * FROM sa1_1_%d\n" % (k, k)
    alert_sql += ("--DataXQuery--\n%sAlert = SELECT DISTINCT DATE_TRUNC('second', current_timestamp()) AS EventTime, '%sAlert' AS MetricName, 0 as Metric, "
                  "'%s' AS Product, %s AS Pivot1 FROM sa1_1_%d\n" % (r["tag"], r["tag"], gui["name"], q(r["ruleDescription"]), k))
sql = re.sub(r"Rules = ProcessRules\(DataXProcessedInput\)", rule_sql, sql)
sql = re.sub(r"--DataXQuery--\s*ProcessAlerts\(DataXProcessedInput\)\s*", alert_sql, sql)
open(os.path.join(OUT, "HomeAutomationLocal-combined.txt"), "w").write(sql.strip() + "\n")
conf = """datax.job.name=HomeAutomationLocal
datax.job.input.default.blobschemafile=inputschema.json
datax.job.input.default.streaming.intervalinseconds=%s
datax.job.input.default.referencedata.myDevicesRefData.path=devices.csv
datax.job.input.default.referencedata.myDevicesRefData.format=csv
datax.job.input.default.referencedata.myDevicesRefData.header=true
datax.job.input.default.referencedata.myDevicesRefData.delimiter=,
datax.job.process.projection=projection.txt
datax.job.process.transform=HomeAutomationLocal-combined.txt
datax.job.process.timestampcolumn=%s
datax.job.process.watermark=%s
datax.job.process.timewindow.DataXProcessedInput_5minutes.windowduration=5 minutes
datax.job.process.jar.udf.whoOpened.class=datax.sample.udf.UdfHelloWorld
datax.job.process.jar.udf.whoOpened.path=datax-udf-samples.jar
""" % (props["windowDuration"], gui["process"]["timestampColumn"], gui["process"]["watermark"])
open(os.path.join(OUT, "HomeAutomationLocal.conf"), "w").write(conf)
open(os.path.join(OUT, "devices.csv"), "w").write(open(CSV, encoding="utf-8-sig").read())
print(open(os.path.join(OUT, "HomeAutomationLocal-combined.txt")).read()[:600])
# the flow's GUI section as it is stored: the input of the restated code generator (dx_codegen_flow), which makes the transform above
# from it without the hand-applied rewrites of this script
json.dump({"gui": gui}, open(os.path.join(OUT, "flow_gui.json"), "w"), indent=1)
