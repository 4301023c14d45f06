#!/usr/bin/env python3
"""Copies the INPUTS of the reference's code-generator tests (DataX.Flow.CodegenRules.Tests: user code files, Rules.json, the query and
output templates; the generator's embedded default templates) and the two expected outputs not yet under ../codegen, and writes the
case table of CodegenTests.cs (which user code, which product id, which expected file) as cases.json.  Run in the build container; the
copies are committed and are all the tests read."""
import glob
import json
import os
import re
import shutil

SRC = "/root/reference/Services/DataX.Flow/DataX.Flow.CodegenRules.Tests"
RES = "/root/reference/Services/DataX.Flow/DataX.Flow.CodegenRules/Resources"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "codegen_inputs")
os.makedirs(OUT, exist_ok=True)
for f in sorted(glob.glob(SRC + "/UserCode*.txt")) + [SRC + "/usercode.txt", SRC + "/Rules.json", SRC + "/QueryTemplates.xml", SRC + "/OutputTemplates.xml",
                                                      SRC + "/cgen.txt", SRC + "/cgenDefault.txt", SRC + "/CGenIoTSample.txt",
                                                      RES + "/defaultQueryTemplate.xml", RES + "/defaultOutputTemplate.xml"]:
    shutil.copy(f, OUT)
# the case table, from the test source: code file (or literal), templates given or default, product id, expected file
src = open(SRC + "/CodegenTests.cs", encoding="utf-8-sig").read()
cases = []
for body in re.split(r"\[TestMethod\]", src)[1:]:
    name = re.search(r"public void (\w+)", body).group(1)
    lit = re.search(r'string code = "(.*?)";', body)
    files = re.findall(r'new StreamReader\("([^"]+)"\)', body)
    user = None if lit else files[0]
    default_templates = re.search(r"GenerateCode\(code, rules, \"", body) is not None
    product = re.search(r'GenerateCode\([^;]*?"([^"]*)"\);', body).group(1)
    cases.append({"name": name, "user_code_file": user, "user_code": lit.group(1) if lit else None, "default_templates": default_templates,
                  "product": product, "expected": files[-1]})
json.dump(cases, open(os.path.join(OUT, "cases.json"), "w"), indent=1)
print(len(cases), "cases")
