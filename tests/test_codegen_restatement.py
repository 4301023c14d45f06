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


def flow_transform():
    from importlib import import_module
    L = import_module("data-accelerator_b200._native").lib()
    fn = L.dx_codegen_flow
    fn.restype, fn.argtypes = C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_char_p, C.c_int64]
    c1 = os.path.join(HERE, "golden", "c1_homeautomation")
    out, info = C.create_string_buffer(1 << 20), C.create_string_buffer(1 << 16)
    rc = fn(open(os.path.join(c1, "flow_gui.json"), "rb").read(), read("defaultQueryTemplate.xml").encode(), read("defaultOutputTemplate.xml").encode(),
            out, len(out), info, len(info))
    assert rc == 0, out.value.decode()
    return out.value.decode(), info.value.decode()


def test_home_automation_flow_json_to_transform():
    """The local sample flow's GUI JSON (DeploymentLocal/sample/HomeAutomationLocal.json) through the config generation's path
    (S450_PrepareTransformFile + RuleDefinitionGenerator + the generator with its default templates): the time window, the outputs
    and the statements come out, and the oracle computes from the generated transform exactly what it computes from the transform
    this repository used to derive by hand (tests/golden/make_c1.py) — rule ids aside, which the reference draws at random."""
    import json as _json
    import random
    import numpy as np
    import dxo_flow
    code, info = flow_transform()
    lines = [l.split("\t") for l in info.splitlines()]
    assert ["window", "DataXProcessedInput_5minutes", "5 minutes"] in lines
    assert ["output", "CLOSEAlert", "Metrics"] in lines and ["metric", "CLOSEAlert"] in lines
    assert "TIMEWINDOW" not in code and "ProcessRules" not in code and "OUTPUT" not in code and ";" not in code
    c1 = os.path.join(HERE, "golden", "c1_homeautomation")
    rd = lambda f: open(os.path.join(c1, f), encoding="utf-8-sig").read()
    flows = [dxo_flow.Flow(rd("HomeAutomationLocal.conf"), rd("inputschema.json"), [rd("projection.txt")], t) for t in (code, rd("HomeAutomationLocal-combined.txt"))]
    for f in flows:
        f.add_refdata("myDevicesRefData", rd("devices.csv"))
    dxo_flow.UDFS["WHOOPENED"] = lambda s: "hello, " + str(s if s is not None else "null")
    rng = random.Random(4)
    T0, I = 1_700_000_040_000, 60_000
    for b in range(8):
        T = T0 + b * I
        now = T + 1500
        ev = []
        for _ in range(120):
            ev.append(_json.dumps({"deviceDetails": {"deviceId": rng.choice([1, 2, 3, 4, 5, 6]), "deviceType": rng.choice(["DoorLock", "WindowLock", "Heating"]),
                                                     "eventTime": now - rng.randrange(1000), "homeId": rng.choice([32, 150, 25, 81]), "status": rng.choice([0, 1])}}, separators=(",", ":")))
        data = ("\n".join(ev) + "\n").encode()
        (ra, _), (rb, _) = [f.process(data, T, I, now_ms=now) for f in flows]
        assert set(ra) == set(rb)
        for name in ra:
            a, bt = ra[name], rb[name]
            assert a.n == bt.n and a.names() == bt.names(), name
            for ca, cb in zip(a.cols, bt.cols):
                if ca[0].lower() == "ruleid":
                    continue                      # "HomeAutomationLocal-rule-1" here, "R1" in the hand-derived text
                if ca[1] == "rules":
                    assert (np.asarray(ca[2]) == np.asarray(cb[2])).all(), name
                else:
                    assert sorted(map(repr, zip(ca[2].tolist() if ca[2] is not None else [None] * a.n, ca[3].tolist()))) == \
                           sorted(map(repr, zip(cb[2].tolist() if cb[2] is not None else [None] * bt.n, cb[3].tolist()))), (name, ca[0])


def test_unbalanced_parenthesis_in_user_code_is_an_error_not_a_crash():
    # the reference's formatter pops an empty stack here and throws; so does this one (found by mutating the goldens' user code)
    from importlib import import_module
    L = import_module("data-accelerator_b200._native").lib()
    out, info = C.create_string_buffer(1 << 16), C.create_string_buffer(1 << 12)
    L.dx_codegen_rules.restype = C.c_int
    L.dx_codegen_rules.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_char_p, C.c_int64]
    rc = L.dx_codegen_rules(b"T1 = ProcessRulesoo);(Foo);", read("Rules.json").encode(), read("QueryTemplates.xml").encode(),
                            read("OutputTemplates.xml").encode(), b"P1", out, len(out), info, len(info))
    assert rc != 0 and b"formatter" in out.value


def test_runaway_json_nesting_is_an_error_not_a_stack_overflow():
    from importlib import import_module
    L = import_module("data-accelerator_b200._native").lib()
    out, info = C.create_string_buffer(1 << 16), C.create_string_buffer(1 << 12)
    L.dx_codegen_rules.restype = C.c_int
    L.dx_codegen_rules.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_char_p, C.c_int64]
    rc = L.dx_codegen_rules(b"x", b"[" * 200000 + b"]" * 200000, b"<a/>", b"<a/>", b"P1", out, len(out), info, len(info))
    assert rc != 0 and b"nested deeper" in out.value
