#!/usr/bin/env python3
"""Copies the reference's generated-SQL golden for the IoT sample flow and its companions into test fixtures:
  CGenIoTSample.txt   — expected output of CodegenTests.IoTSampleTest (DataX.Flow.CodegenRules.Tests/CodegenTests.cs),
                        i.e. exactly the transform text the reference's processor would be handed;
  the CREATE TABLE of the state table comes from UserCodeIoTSample.txt:2-3;
  projection / schema from DataX.Config.Local.Test/Resource (projection.txt, schema.json); devices.csv.
Run in the build container; the outputs are committed and are all the GPU box reads."""
import os
import re
import shutil

R = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "iotsample")
T = R + "/Services/DataX.Flow/DataX.Flow.CodegenRules.Tests"
L = R + "/Services/DataX.Config/DataX.Config.Local.Test/Resource"
shutil.copy(T + "/CGenIoTSample.txt", OUT + "/iotsample-combined.txt")
shutil.copy(L + "/projection.txt", OUT + "/projection.txt")
shutil.copy(L + "/schema.json", OUT + "/inputschema.json")
shutil.copy(R + "/DeploymentCloud/Deployment.DataX/Samples/usercontent/devices.csv", OUT + "/devices.csv")
user = open(T + "/UserCodeIoTSample.txt", encoding="utf-8-sig").read()
m = re.search(r"CREATE TABLE\s+(\w+)\s*\(([^)]*)\)", user, re.S)
conf = f"""datax.job.name=iotsample
datax.job.input.default.blobschemafile=inputschema.json
datax.job.input.default.referencedata.myDevicesRefdata.path=devices.csv
datax.job.input.default.referencedata.myDevicesRefdata.format=csv
datax.job.input.default.referencedata.myDevicesRefdata.header=true
datax.job.input.default.referencedata.myDevicesRefdata.delimiter=,
datax.job.process.projection=projection.txt
datax.job.process.transform=iotsample-combined.txt
datax.job.process.timestampcolumn=eventTimeStamp
datax.job.process.watermark=60 second
datax.job.process.timewindow.DataXProcessedInput_5minutes.windowduration=5 minutes
datax.job.process.statetable.{m.group(1)}.schema={' '.join(m.group(2).split())}
datax.job.process.statetable.{m.group(1)}.location=hdfs://mycluster/datax/iotsample/{m.group(1)}/
datax.job.process.jar.udf.whoOpened.class=datax.sample.udf.UdfHelloWorld
datax.job.process.jar.udf.whoOpened.path=datax-udf-samples.jar
"""
open(OUT + "/iotsample.conf", "w").write(conf)
print(conf)
