"""The device's decimal->binary64 code (csrc/dx_numconv.cuh), compiled unchanged for the host,
against Python's correctly-rounded float() — which is what the reference gets from Jackson's
getDoubleValue inside from_json (CommonProcessorFactory.scala:93)."""
import ctypes
import math
import os
import random
import struct
import subprocess
from fractions import Fraction

import pytest

from adversarial import random_number

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def conv():
    src = os.path.join(ROOT, "tests", "native", "host_numconv.cpp")
    out = os.path.join(ROOT, "tests", "native", "libhostnumconv.so")
    subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", "-I" + os.path.join(ROOT, "data-accelerator_b200", "csrc"),
                           "-o", out, src])
    L = ctypes.CDLL(out)

    def f(s):
        o = ctypes.c_double()
        is_int = "." not in s and "e" not in s and "E" not in s
        assert L.numconv_parse(s.encode(), len(s), int(is_int), ctypes.byref(o)) == 1, s
        return o.value
    return f


def same(a, b):
    return struct.pack("<d", a) == struct.pack("<d", b)


def exact_decimal(fr: Fraction):
    """Exact decimal expansion of a dyadic rational."""
    n, d = fr.numerator, fr.denominator
    k = d.bit_length() - 1
    assert d == 1 << k
    digits = str(n * 5 ** k)
    if k == 0:
        return digits
    digits = digits.rjust(k + 1, "0")
    return digits[:-k] + "." + digits[-k:]


def test_random_numbers(conv):
    rng = random.Random(1)
    for _ in range(100_000):
        s = random_number(rng)
        exp = float(s)
        if "." not in s and "e" not in s and exp == 0:
            exp = 0.0
        assert same(conv(s), exp), s


def test_exact_midpoints_and_neighbours(conv):
    rng = random.Random(2)
    for _ in range(3000):
        e = rng.randrange(-1070, 1020)
        m = rng.getrandbits(52) | (1 << 52)
        x = math.ldexp(m, e - 52)
        if math.isinf(x) or x == 0:
            continue
        y = math.nextafter(x, math.inf)
        if math.isinf(y):
            continue
        mid = (Fraction(x) + Fraction(y)) / 2
        s = exact_decimal(mid)
        assert same(conv(s), float(s)), s[:60]                      # tie: round half to even
        above = s + "1" if "." in s else s + ".0000001"
        assert same(conv(above), y), s[:60]                         # just above the midpoint
        # just below: drop one unit in the last written place, then pad with nines
        if "." in s:
            ip, fp = s.split(".")
            v = str(int(ip + fp) - 1).rjust(len(ip + fp), "0")
            below = v[:len(ip)] + "." + v[len(ip):] + "99999"
        else:
            below = str(int(s) - 1) + ".99999"
        assert same(conv(below), x), s[:60]


def test_edges(conv):
    for s in ["0", "-0", "0.0", "-0.0", "1e-400", "1e400", "-1e400", "4.9e-324", "2.4703282292062327e-324",
              "2.4703282292062328e-324", "1.7976931348623157e308", "1.7976931348623158e308",
              "1.797693134862315807e308", "1.7976931348623159e308", "2.2250738585072011e-308",
              "2.2250738585072014e-308", "9007199254740993", "9007199254740992.5", "9007199254740993.0000000001",
              "0.1", "0.30000000000000004", "123456789012345678901234567890", "1" + "0" * 400,
              "0." + "0" * 400 + "1", "1e23", "8.41e21", "5e-324", "1e-323", "3e-324"]:
        exp = float(s)
        if s == "-0":
            exp = 0.0
        assert same(conv(s), exp), s
