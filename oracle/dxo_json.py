r"""dxo_json — CPU restatement of the sink hand-off's row serialisation.  TEST INFRASTRUCTURE: only tests/, smoke() and
bench.py's CPU legs import it; the product path never does.

What it restates:
  * OutputManager.sink (DataProcessing/datax-host/src/main/scala/datax/sink/OutputManager.scala:103-118): every column whose name does
    not start with "__DataX_" (OutputManager.scala:92) goes into to_json(struct(cols)); one boolean column per sink filter expression
    is appended (OutputManager.scala:53-61, 110-113) — the filters are evaluated by the flow oracle, not here.
  * to_json = Spark 2.4.5 JacksonGenerator (not in /root/reference: org.apache.spark:spark-sql_2.11:2.4.5, pom.xml:77-93):
      - a struct writes its fields in order and SKIPS the NULL ones; a map writes every key, a NULL value as `null`;
      - long -> decimal digits; boolean -> true / false; double -> Jackson writeNumber(double) = java.lang.Double.toString, with
        NaN / Infinity / -Infinity as quoted strings (JsonGenerator.Feature.QUOTE_NON_NUMERIC_NUMBERS is on by default);
      - string and field names -> Jackson's default escapes: \" \\ \b \t \n \f \r, other bytes below 0x20 as \u00XX with upper-case
        hex digits; everything else, non-ASCII included, is written as it is;
      - timestamp -> the default timestampFormat yyyy-MM-dd'T'HH:mm:ss.SSSXXX in the session time zone (UTC: ...Z), milliseconds
        floored from the microsecond value;
      - MAP(...) leaves of different types take Spark's common type first (TypeCoercion.FunctionArgumentConversion on CreateMap):
        any string among them -> all strings (Cast to string: Double.toString, true/false, 'yyyy-MM-dd HH:mm:ss[.f]'), long + double
        -> double.
    Parity at this boundary is unpinned by reference tests (no reference test serialises a row); the rules above are listed one by
    one in NEEDS_SPARK_CONFIRMATION.md.
  * java.lang.Double.toString: shortest decimal that round-trips (of two digits at least: the nearest 2-digit one when one digit would do), formatted d.ddd (1e-3 <= |x| < 1e7) or d.dddE±n.  JDK 8 (the
    reference's runtime) is known to print a longer or not-closest digit string for a small set of doubles (JDK-4511638, fixed in JDK 19);
    this restatement and the product print the shortest round-trip digits for those too.

A table row model note: a decoded struct (columns named a.b.c by the decoder) is written when at least one of its leaves is not NULL;
the table model does not keep the validity of intermediate structs, so {"a":{}} and a missing "a" both serialise as a missing "a".
"""
import calendar
import decimal
import math
import time as _time

import numpy as np

_ESC = {0x22: b'\\"', 0x5C: b"\\\\", 0x08: b"\\b", 0x09: b"\\t", 0x0A: b"\\n", 0x0C: b"\\f", 0x0D: b"\\r"}


def jackson_quote(b):
    """bytes (UTF-8) -> the quoted JSON string Jackson's writeString produces."""
    if isinstance(b, str):
        b = b.encode("utf-8")
    out = bytearray(b'"')
    for c in b:
        if c in _ESC:
            out += _ESC[c]
        elif c < 0x20:
            out += b"\\u00%02X" % c
        else:
            out.append(c)
    out += b'"'
    return bytes(out)


def java_double_string(x):
    """java.lang.Double.toString(x) with the shortest round-trip digits (see the module note on JDK-4511638)."""
    x = float(x)
    if x != x:
        return "NaN"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if x == 0.0:
        return "-0.0" if math.copysign(1.0, x) < 0 else "0.0"
    sign, digits, exp = decimal.Decimal(repr(abs(x))).as_tuple()
    ds = "".join(map(str, digits)).lstrip("0")
    stripped = ds.rstrip("0")
    exp += len(ds) - len(stripped)
    ds = stripped
    if len(ds) == 1:
        # Double.toString's specification: when the shortest decimal that rounds to x has ONE digit, the result is the 2-digit
        # decimal nearest to x among those that round to x (a tie goes to the even last digit): 4.9E-324, not 5.0E-324.
        from fractions import Fraction
        ax, fx = abs(x), Fraction(abs(x))
        best = None
        cands = [(c, exp - 1) for c in range(max(10, int(ds) * 10 - 9), min(99, int(ds) * 10 + 9) + 1)]
        if ds == "1":
            cands += [(c, exp - 2) for c in range(91, 100)]       # just below the power of ten: 9.9E-324 next to 1.0E-323
        for cand, ce in cands:
            try:
                if float("%de%d" % (cand, ce)) != ax:
                    continue
            except OverflowError:
                continue
            dist = abs(Fraction(cand) * Fraction(10) ** ce - fx)
            if best is None or dist < best[0] or (dist == best[0] and cand % 2 == 0):
                best = (dist, cand, ce)
        if best is not None and best[1] % 10 != 0:
            ds, exp = str(best[1]), best[2]
    e = len(ds) - 1 + exp                       # scientific exponent: x = d.ddd * 10^e
    if 0 <= e <= 6:
        s = ds[:e + 1].ljust(e + 1, "0") + "." + (ds[e + 1:] or "0")
    elif -3 <= e < 0:
        s = "0." + "0" * (-e - 1) + ds
    else:
        s = ds[0] + "." + (ds[1:] or "0") + "E" + str(e)
    return ("-" if x < 0 else "") + s


def json_timestamp(us):
    """yyyy-MM-dd'T'HH:mm:ss.SSSXXX in UTC for a microsecond timestamp."""
    ms = int(us) // 1000                         # floor, as java.sql.Timestamp.getTime() for a negative value with nanos
    s, frac = divmod(ms, 1000)
    g = _time.gmtime(s)
    return "%04d-%02d-%02dT%02d:%02d:%02d.%03dZ" % (g.tm_year, g.tm_mon, g.tm_mday, g.tm_hour, g.tm_min, g.tm_sec, frac)


def cast_timestamp_string(us):
    """Spark Cast(timestamp -> string): yyyy-MM-dd HH:mm:ss with the fractional second only when it is not zero, trailing zeros cut."""
    s, frac = divmod(int(us), 1_000_000)
    g = _time.gmtime(s)
    base = "%04d-%02d-%02d %02d:%02d:%02d" % (g.tm_year, g.tm_mon, g.tm_mday, g.tm_hour, g.tm_min, g.tm_sec)
    return base + ("." + ("%06d" % frac).rstrip("0") if frac else "")


def _scalar(ty, v, as_type=None):
    """one non-NULL value -> JSON bytes; as_type: the common type of a map's values when it differs from ty."""
    if as_type == "string" and ty != "string":
        if ty == "long":
            return jackson_quote(str(int(v)))
        if ty == "bool":
            return jackson_quote("true" if v else "false")
        if ty == "double":
            return jackson_quote(java_double_string(v))
        if ty == "ts":
            return jackson_quote(cast_timestamp_string(v))
    if as_type == "double" and ty == "long":
        ty, v = "double", float(int(v))
    if ty == "long":
        return str(int(v)).encode()
    if ty == "bool":
        return b"true" if v else b"false"
    if ty == "double":
        s = java_double_string(v)
        return jackson_quote(s) if s in ("NaN", "Infinity", "-Infinity") else s.encode()
    if ty == "ts":
        return jackson_quote(json_timestamp(v))
    if ty == "string":
        return jackson_quote(v)
    raise ValueError("to_json: type " + ty)


class _Node:
    def __init__(self, key, kind):
        self.key, self.kind, self.children = key, kind, []     # kind: 'd' decoded struct, 's' STRUCT(...), 'm' MAP(...), 'leaf'
        self.col = None


def _tree(tab):
    """columns -> top-level nodes; consecutive columns that share a path prefix of the same kind share the object."""
    from dxo_flow import is_metadata_column
    root = _Node(None, "s")
    nest = getattr(tab, "nest", {})
    for ci, (name, ty, v, ok) in enumerate(tab.cols):
        if name.startswith("__DataX_"):
            continue
        if name in nest:
            path, kinds = nest[name]
            alias = name[: len(name) - len(".".join(path)) - 1]
            keys, kk = [alias] + list(path), list(kinds)
        elif is_metadata_column(name):
            # DataXProperties / Properties / SystemProperties are map<string,string> columns: name = "<Column>.<key>"
            head, _, key = name.partition(".")
            keys, kk = [head, key], ["m"]
        else:
            keys = name.split(".")
            kk = ["d"] * (len(keys) - 1)
        node = root
        for depth, key in enumerate(keys[:-1]):
            last = node.children[-1] if node.children else None
            if last is None or last.kind == "leaf" or last.key != key or last.kind != kk[depth]:
                last = _Node(key, kk[depth])
                node.children.append(last)
            node = last
        leaf = _Node(keys[-1], "leaf")
        leaf.col = ci
        node.children.append(leaf)
    return root


def _common_type(tab, node):
    tys = {tab.cols[c.col][1] for c in node.children if c.kind == "leaf" and tab.cols[c.col][1] not in ("null", "rules")}
    if len(tys) <= 1:
        return None
    if "string" in tys:
        return "string"
    if tys == {"long", "double"}:
        return "double"
    return None


def _rules_bytes(tab, flags_row):
    table = tab.rule_table
    objs = []
    for r in np.nonzero(flags_row)[0]:
        objs.append(b"{" + b",".join(jackson_quote(k) + b":" + jackson_quote(v) for k, v in table[r].items()) + b"}")
    if getattr(tab, "rule_single", False):
        return objs[0] if objs else None          # IF(cond, MAP(...), NULL): the map, or a NULL field
    return b"[" + b",".join(objs) + b"]"          # filterNull(Array(...)): an array, empty when nothing matched


def _emit(tab, node, r, as_type=None):
    """-> bytes of the node's value for row r, or None when the value is NULL."""
    if node.kind == "leaf":
        name, ty, v, ok = tab.cols[node.col]
        if ty == "rules":
            return _rules_bytes(tab, v[r])
        if ty == "null" or v is None or not ok[r]:
            return None
        return _scalar(ty, v[r], as_type)
    common = _common_type(tab, node) if node.kind == "m" else None
    parts, any_value = [], False
    for c in node.children:
        b = _emit(tab, c, r, common)
        if b is None:
            if node.kind == "m":
                parts.append(jackson_quote(c.key) + b":null")
            continue
        any_value = True
        parts.append(jackson_quote(c.key) + b":" + b)
    if node.kind == "d" and not any_value:
        return None
    return b"{" + b",".join(parts) + b"}"


def to_json_rows(tab):
    """-> list of bytes, one JSON object per row (what to_json(struct(outputColumns)) yields)."""
    root = _tree(tab)
    return [_emit(tab, root, r) for r in range(tab.n)]
