#!/usr/bin/env python3
"""Copies the expected outputs of the reference's code-generator tests (DataX.Flow.CodegenRules.Tests/CGen*.txt,
asserted by CodegenTests.cs) into test fixtures: each is exactly the transform text the reference's processor is
handed for that rule set.  CGenIoTSample.txt lives under ../iotsample (make_iotsample.py).
Run in the build container; the copies are committed and are all the GPU box reads."""
import glob
import os
import shutil

SRC = "/root/reference/Services/DataX.Flow/DataX.Flow.CodegenRules.Tests"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "codegen")
os.makedirs(OUT, exist_ok=True)
for f in sorted(glob.glob(SRC + "/CGen*.txt")):
    if os.path.basename(f) != "CGenIoTSample.txt":
        shutil.copy(f, OUT)
        print(os.path.basename(f))
