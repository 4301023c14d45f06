#!/usr/bin/env python3
"""Extracts the reference's own sample events into committed fixtures.  Run in the build
container (needs /root/reference); the GPU box only ever reads the generated files."""
import json
import os
import re

REF = "/root/reference/DeploymentCloud/Deployment.DataX/Samples/samples"
HERE = os.path.dirname(os.path.abspath(__file__))

events = []
text = open(os.path.join(REF, "iotDevice/iotsample.json"), encoding="utf-8-sig").read()
for m in re.finditer(r'"dataStream":\s*"((?:[^"\\]|\\.)*)"', text):
    events.append(json.loads('"' + m.group(1) + '"'))
blob = open(os.path.join(REF, "iotsample-091610A6B869B8C9318E988F390F46AB.json"), encoding="utf-8-sig").read()
for line in blob.splitlines():
    line = line.strip()
    if line:
        events.append(json.loads(line)["Raw"])
with open(os.path.join(HERE, "reference_sample_events.ndjson"), "w") as f:
    for e in events:
        assert "\n" not in e
        f.write(e + "\n")
expected = [json.loads(e)["deviceDetails"] for e in events]
json.dump(expected, open(os.path.join(HERE, "reference_sample_events.expected.json"), "w"), indent=0)
print(len(events), "events")
