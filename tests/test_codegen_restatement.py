"""The rules / flow code generator restated (dx_codegen_rules) against the reference's own test expectations:
every case of DataX.Flow.CodegenRules.Tests/CodegenTests.cs — user code + Rules.json + templates + product id -> the expected
generated transform — must come out byte for byte (fixtures copied by tests/golden/make_codegen_inputs.py).  No GPU."""
import ctypes as C
import json
import os

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
IN = os.path.join(HERE, "golden", "codegen_inputs")
OUT = os.path.join(HERE, "golden", "codegen")
CASES = json.load(open(os.path.join(IN, "cases.json")))


def read(name):
    for d in (IN, OUT, os.path.join(HERE, "golden", "iotsample")):
        p = os.path.join(d, name)
        if os.path.exists(p):
            return open(p, "rb").read().decode("utf-8-sig")          # what StreamReader.ReadToEnd gives: BOM dropped, line ends kept
    raise FileNotFoundError(name)


def generate(code, rules, qt, ot, product):
    from importlib import import_module
    L = import_module("data-accelerator_b200._native").lib()
    fn = L.dx_codegen_rules
    fn.restype = C.c_int
    fn.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_char_p, C.c_int64]
    out, info = C.create_string_buffer(1 << 20), C.create_string_buffer(1 << 16)
    rc = fn(code.encode(), rules.encode(), qt.encode(), ot.encode(), product.encode(), out, len(out), info, len(info))
    assert rc == 0, out.value.decode()
    return out.value.decode(), info.value.decode()


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_generated_code_equals_the_references_expectation(case):
    # the expectations were produced from working copies with CRLF line ends (the pretty-printer turns "\r" and "\n" into one space
    # each): the user code is handed over the same way
    code = case["user_code"] if case["user_code_file"] is None else read(case["user_code_file"]).replace("\r\n", "\n").replace("\n", "\r\n")
    qt = read("defaultQueryTemplate.xml" if case["default_templates"] else "QueryTemplates.xml")
    ot = read("defaultOutputTemplate.xml" if case["default_templates"] else "OutputTemplates.xml")
    got, info = generate(code, read("Rules.json"), qt, ot, case["product"])
    # the generator emits CRLF line ends; the reference's expectation files are stored with LF in its repository (git's text
    # normalisation), so the comparison is byte for byte after CRLF -> LF on both sides
    want = read(case["expected"]).replace("\r\n", "\n")
    got = got.replace("\r\n", "\n")
    assert got == want, "first difference at %d: %r vs %r" % (next((i for i in range(min(len(got), len(want))) if got[i] != want[i]), -1),
                                                              got[max(0, next((i for i in range(min(len(got), len(want))) if got[i] != want[i]), 0) - 60):][:160],
                                                              want[max(0, next((i for i in range(min(len(got), len(want))) if got[i] != want[i]), 0) - 60):][:160])


def test_side_results_outputs_tables_windows():
    got, info = generate(read("usercode.txt").replace("\n", "\r\n"), read("Rules.json"), read("QueryTemplates.xml"), read("OutputTemplates.xml"), "P1")
    lines = [l.split("\t") for l in info.splitlines()]
    assert ["table", "EventCount", "Count: INT"] in lines
    assert ["window", "DataXProcessedInput_50seconds", "50 seconds"] in lines and ["window", "DataXProcessedInput_40seconds", "40 seconds"] in lines
    assert any(l[0] == "output" and l[2] == "Metrics" for l in lines) and any(l[0] == "metric" for l in lines)
