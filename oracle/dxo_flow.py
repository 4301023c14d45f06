"""CPU ORACLE — flow level (test infrastructure, NOT product code; see dxo.py header).

Restates, in plain Python/numpy, the control flow of the reference's per-micro-batch processor and
the Spark SQL 2.4.5 semantics it relies on (PARITY UNPINNED at the Spark boundary):

  parse_transform / replace_table_names   datax/sql/TransformSqlParser.scala:27-104
  string_to_timestamp                     datax/utility/ConcurrentDateFormat.scala:14-62
  parse_duration                          TimeWindowHandler.scala:33-35 (scala Duration.create)
  Flow.process = processDataset           CommonProcessorFactory.scala:333-399
       project()                          CommonProcessorFactory.scala:90-103
       route()                            CommonProcessorFactory.scala:131-328 (window cache, eviction,
                                          the pastRdds.size>1 quirk, state-table overwrite + flip)
  SQL evaluation                          stands in for spark.sql(...) at CPF:257,271 — three-valued
                                          logic, NaN ordering, SUM/AVG/MIN/MAX/COUNT rules of SURVEY App. B
Its parser and evaluator share no code with the product's C++ front end.
"""
import calendar
import math
import re

import numpy as np

import dxo

# --------------------------------------------------------------------------------------------------
# TransformSQLParser (TransformSqlParser.scala:27-104)
# --------------------------------------------------------------------------------------------------


def parse_transform(text):
    commands, counts, buf = [], {}, []
    name = None

    def append(n):
        sql = " ".join(l for l in buf if l)
        commands.append((n, "Command" if n is None else "Query", sql))
        if n:
            if n in counts:
                raise ValueError(f"dataset name '{n}' has been created, please check the query to make sure it is not created again")
            counts[n] = 0
            for k in counts:
                if re.search(r"\b" + re.escape(k) + r"\b", sql):
                    counts[k] += 1

    for line in text.lstrip("﻿").split("\n"):
        line = line.rstrip("\r")
        if not line.strip():
            continue
        if re.fullmatch(r"^--DataXQuery--", line):
            if buf:
                append(name)
            name, buf = None, []
        elif re.match(r"^\s*--", line):
            continue
        elif not buf:
            m = re.match(r"^\s*([a-zA-Z0-9_]+)\s*=(.*)$", line)
            if m:
                name = m.group(1)
                buf.append(m.group(2).strip())
            else:
                buf.append(line.strip())
        else:
            buf.append(line.strip())
    if name is not None and buf:
        append(name)
    return commands, counts


# --------------------------------------------------------------------------------------------------
# stringToTimestamp (ConcurrentDateFormat.scala:38-62), UTC session zone
# --------------------------------------------------------------------------------------------------
def _epoch_lenient(y, mo, d, h, mi, s):
    m0 = mo - 1
    y += m0 // 12
    mo = m0 % 12 + 1
    return (calendar.timegm((y, mo, 1, 0, 0, 0)) // 86400 + (d - 1)) * 86400 + h * 3600 + mi * 60 + s


def _sdf(s, pattern):
    vals = {"y": 1970, "M": 1, "d": 1, "H": 0, "m": 0, "s": 0}
    i, p = 0, 0
    while p < len(pattern):
        c = pattern[p]
        if c == "'":
            q = pattern.index("'", p + 1)
            lit = pattern[p + 1:q]
            if s[i:i + len(lit)] != lit:
                return None
            i += len(lit)
            p = q + 1
        elif c in vals:
            while p < len(pattern) and pattern[p] == c:
                p += 1
            m = re.match(r"\d{1,18}", s[i:])
            if not m:
                return None
            vals[c] = int(m.group(0))
            i += len(m.group(0))
        else:
            if s[i:i + 1] != c:
                return None
            i += 1
            p += 1
    return _epoch_lenient(vals["y"], vals["M"], vals["d"], vals["H"], vals["m"], vals["s"]) * 1_000_000


def string_to_timestamp(s):
    if s is None or s == "":
        return None
    t = s.strip()
    m = re.fullmatch(r"(\d{4})-(\d{1,2})-(\d{1,2}) (\d+):(\d+):(\d+)(?:\.(\d{1,9}))?", t)   # java.sql.Timestamp.valueOf
    if m:
        y, mo, d, h, mi, se = (int(x) for x in m.groups()[:6])
        nanos = int((m.group(7) or "0").ljust(9, "0"))
        return _epoch_lenient(y, mo, d, h, mi, se) * 1_000_000 + nanos // 1000
    for pat in ("yyyy-MM-dd'T'HH:mm:ss'Z'", "MM/dd/yyyy HH:mm:ss"):
        v = _sdf(s, pat)
        if v is not None:
            return v
    return None


def parse_duration(s):
    m = re.fullmatch(r"\s*([\d.]+)\s*([a-zA-Z]+)\s*", s)
    if not m:
        raise ValueError(s)
    unit = m.group(2).lower()
    mul = {"d": 86400, "day": 86400, "days": 86400, "h": 3600, "hr": 3600, "hrs": 3600, "hour": 3600, "hours": 3600,
           "m": 60, "min": 60, "mins": 60, "minute": 60, "minutes": 60, "s": 1, "sec": 1, "secs": 1, "second": 1,
           "seconds": 1, "ms": 1e-3, "milli": 1e-3, "millis": 1e-3, "millisecond": 1e-3, "milliseconds": 1e-3}[unit]
    return int(math.floor(float(m.group(1)) * mul))


# --------------------------------------------------------------------------------------------------
# SQL subset: tokenizer + parser (independent of the product's)
# --------------------------------------------------------------------------------------------------
_TOK = re.compile(r"""\s*(?:(?P<num>\d+\.\d*(?:[eE][+-]?\d+)?|\.\d+(?:[eE][+-]?\d+)?|\d+(?:[eE][+-]?\d+)?)|(?P<id>[A-Za-z_][A-Za-z0-9_]*)|`(?P<qid>[^`]*)`|'(?P<str>(?:[^'\\]|\\.)*)'|"(?P<dstr>(?:[^"\\]|\\.)*)"|(?P<op><>|!=|<=|>=|==|\|\||[-+*/%(),.=<>;]))""")
_KW = {"SELECT", "FROM", "WHERE", "GROUP", "BY", "AS", "AND", "OR", "NOT", "IN", "BETWEEN", "LIKE", "IS", "JOIN", "INNER", "LEFT", "OUTER", "ON",
       "UNION", "ALL", "DISTINCT", "HAVING", "ORDER", "LIMIT", "CASE", "WHEN", "THEN", "ELSE", "END"}


def _unescape(s):
    out, i = [], 0
    while i < len(s):
        c = s[i]
        if c == "\\" and i + 1 < len(s):
            e = s[i + 1]
            out.append({"n": "\n", "t": "\t", "r": "\r", "%": "\\%", "_": "\\_"}.get(e, e))
            i += 2
        else:
            out.append(c)
            i += 1
    return "".join(out)


def tokenize(sql):
    toks, pos = [], 0
    while pos < len(sql):
        m = _TOK.match(sql, pos)
        if not m or m.end() == pos:
            if sql[pos:].strip() == "":
                break
            raise ValueError("cannot tokenize near " + sql[pos:pos + 20])
        pos = m.end()
        if m.group("num") is not None:
            toks.append(("num", m.group("num")))
        elif m.group("id") is not None:
            toks.append(("id", m.group("id")))
        elif m.group("qid") is not None:
            toks.append(("qid", m.group("qid")))
        elif m.group("str") is not None:
            toks.append(("str", _unescape(m.group("str"))))
        elif m.group("dstr") is not None:
            toks.append(("str", _unescape(m.group("dstr"))))
        else:
            toks.append(("op", m.group("op")))
    toks.append(("end", ""))
    return toks


class P:
    def __init__(self, sql):
        self.t = tokenize(sql)
        self.i = 0

    def peek(self, k=0):
        return self.t[min(self.i + k, len(self.t) - 1)]

    def take(self):
        x = self.peek()
        if self.i < len(self.t) - 1:
            self.i += 1
        return x

    def kw(self, w, k=0):
        a = self.peek(k)
        return a[0] == "id" and a[1].upper() == w

    def op(self, o, k=0):
        return self.peek(k) == ("op", o)

    def accept_kw(self, w):
        if self.kw(w):
            self.take()
            return True
        return False

    def accept_op(self, o):
        if self.op(o):
            self.take()
            return True
        return False

    def expect_op(self, o):
        if not self.accept_op(o):
            raise ValueError(f"expected {o} near {self.peek()}")

    def expect_kw(self, w):
        if not self.accept_kw(w):
            raise ValueError(f"expected {w} near {self.peek()}")

    # expression grammar -> tuples
    def primary(self):
        k, v = self.peek()
        if k == "num":
            self.take()
            return ("int", int(v)) if re.fullmatch(r"\d+", v) and int(v) < 2 ** 63 else ("dec", float(v))
        if k == "str":
            self.take()
            return ("str", v)
        if (k, v) == ("op", "("):
            self.take()
            e = self.expr()
            self.expect_op(")")
            return e
        if (k, v) == ("op", "*"):
            self.take()
            return ("star", [])
        if (k, v) == ("op", "-"):
            self.take()
            return ("neg", self.primary())
        if k in ("id", "qid"):
            u = v.upper()
            if k == "id":
                if u == "NULL":
                    self.take()
                    return ("null",)
                if u in ("TRUE", "FALSE"):
                    self.take()
                    return ("bool", u == "TRUE")
                if u == "CASE":
                    self.take()
                    subject = None if self.kw("WHEN") else self.expr()
                    arms = []
                    while self.accept_kw("WHEN"):
                        c = self.expr()
                        if subject is not None:
                            c = ("bin", "=", subject, c)
                        self.expect_kw("THEN")
                        arms.append((c, self.expr()))
                    if not arms:
                        raise ValueError("CASE without WHEN")
                    tail = self.expr() if self.accept_kw("ELSE") else ("null",)
                    self.expect_kw("END")
                    for c, v in reversed(arms):      # CaseWhen == nested IF (a NULL condition is not taken)
                        tail = ("func", "IF", [c, v, tail], False, False)
                    return tail
                if u == "CAST" and self.op("(", 1):
                    self.take(); self.take()
                    e = self.expr()
                    self.expect_kw("AS")
                    ty = self.take()[1].lower()
                    self.expect_op(")")
                    return ("cast", e, ty)
                if self.op("(", 1) and u not in _KW:
                    self.take(); self.take()
                    distinct = self.accept_kw("DISTINCT")
                    args, star = [], False
                    if not self.op(")"):
                        if self.op("*") and self.op(")", 1):
                            self.take()
                            star = True
                        else:
                            args.append(self.arg())
                            while self.accept_op(","):
                                args.append(self.arg())
                    self.expect_op(")")
                    return ("func", u, args, distinct, star)
            parts = [self.take()[1]]
            while self.op("."):
                if self.op("*", 1):
                    self.take(); self.take()
                    return ("star", parts)
                if self.peek(1)[0] not in ("id", "qid"):
                    break
                self.take()
                parts.append(self.take()[1])
            return ("col", parts)
        raise ValueError(f"unexpected token {self.peek()}")

    def arg(self):
        e = self.expr()
        if self.kw("AS"):
            self.take()
            return ("named", e, self.take()[1])      # STRUCT(x AS n)
        return e

    def mul(self):
        l = self.primary()
        while self.op("*") or self.op("/") or self.op("%"):
            o = self.take()[1]
            l = ("bin", o, l, self.primary())
        return l

    def add(self):
        l = self.mul()
        while self.op("+") or self.op("-"):
            o = self.take()[1]
            l = ("bin", o, l, self.mul())
        return l

    def cmp(self):
        l = self.add()
        while True:
            k, v = self.peek()
            if k == "op" and v in ("=", "==", "<>", "!=", "<", "<=", ">", ">="):
                self.take()
                v = {"==": "=", "!=": "<>"}.get(v, v)
                l = ("bin", v, l, self.add())
                continue
            neg = False
            if self.kw("NOT") and (self.kw("IN", 1) or self.kw("BETWEEN", 1) or self.kw("LIKE", 1)):
                self.take()
                neg = True
            if self.accept_kw("IN"):
                self.expect_op("(")
                items = [self.expr()]
                while self.accept_op(","):
                    items.append(self.expr())
                self.expect_op(")")
                l = ("in", l, items, neg)
                continue
            if self.accept_kw("BETWEEN"):
                lo = self.add()
                self.expect_kw("AND")
                hi = self.add()
                l = ("between", l, lo, hi, neg)
                continue
            if self.accept_kw("LIKE"):
                l = ("like", l, self.add(), neg)
                continue
            if self.kw("IS"):
                self.take()
                n = self.accept_kw("NOT")
                self.expect_kw("NULL")
                l = ("isnull", l, n)
                continue
            return l

    def not_(self):
        if self.kw("NOT"):
            self.take()
            return ("not", self.not_())
        return self.cmp()

    def and_(self):
        l = self.not_()
        while self.kw("AND"):
            self.take()
            l = ("bin", "AND", l, self.not_())
        return l

    def expr(self):
        l = self.and_()
        while self.kw("OR"):
            self.take()
            l = ("bin", "OR", l, self.and_())
        return l

    def alias(self):
        if self.accept_kw("AS"):
            return self.take()[1]
        k, v = self.peek()
        if (k == "id" and v.upper() not in _KW) or k == "qid":
            return self.take()[1]
        return None

    def select_item(self):
        e = self.expr()
        return e, self.alias()

    def select_core(self):
        self.expect_kw("SELECT")
        distinct = self.accept_kw("DISTINCT")
        items = [self.select_item()]
        while self.accept_op(","):
            items.append(self.select_item())
        self.expect_kw("FROM")
        frm = self.take()[1]
        frm_alias = self.alias()
        joins = []
        while True:
            outer = False
            if self.kw("LEFT") and (self.kw("JOIN", 1) or (self.kw("OUTER", 1) and self.kw("JOIN", 2))):
                self.take()
                self.accept_kw("OUTER")
                outer = True
            elif self.kw("INNER") and self.kw("JOIN", 1):
                self.take()
            if not self.accept_kw("JOIN"):
                break
            jt = self.take()[1]
            ja = self.alias()
            self.expect_kw("ON")
            joins.append((jt, ja, self.expr(), outer))
        where = self.expr() if self.accept_kw("WHERE") else None
        group = []
        if self.kw("GROUP"):
            self.take()
            self.expect_kw("BY")
            group.append(self.expr())
            while self.accept_op(","):
                group.append(self.expr())
        having = self.expr() if self.accept_kw("HAVING") else None
        return {"distinct": distinct, "items": items, "from": frm, "alias": frm_alias, "where": where, "group": group, "joins": joins,
                "having": having}

    def statement(self):
        sel = [self.select_core()]
        alls = []
        while self.kw("UNION"):
            self.take()
            alls.append(self.accept_kw("ALL"))
            sel.append(self.select_core())
        self.accept_op(";")
        if self.peek()[0] != "end":
            raise ValueError(f"trailing input near {self.peek()}")
        return sel, alls


# --------------------------------------------------------------------------------------------------
# tables and expression evaluation (Spark semantics)
# --------------------------------------------------------------------------------------------------
def is_metadata_column(name):
    return name.startswith(("__DataX_", "DataXProperties.", "Properties.", "SystemProperties."))


def java_timestamp_string(ms):
    """java.sql.Timestamp.toString in the UTC session: yyyy-mm-dd hh:mm:ss.f with at least one fractional digit."""
    s, frac = divmod(ms, 1000)
    t = calendar.timegm
    import time as _t
    g = _t.gmtime(s)
    base = "%04d-%02d-%02d %02d:%02d:%02d" % (g.tm_year, g.tm_mon, g.tm_mday, g.tm_hour, g.tm_min, g.tm_sec)
    return base + ("." + ("%03d" % frac).rstrip("0") if frac else ".0")


class Tab:
    """columns: list of (name, type, values, valid).  type in long|double|string|bool|ts|null|rules."""

    def __init__(self, n):
        self.n = n
        self.cols = []
        self.nest = {}       # MAP/STRUCT leaves: column name -> (key path, 'm'/'s' per level)
        self.approx = set()  # double columns whose value depends on summation order (SUM / AVG of doubles and what is computed
                             # from them): the only columns a comparison may give a tolerance; everything else is bit-exact

    def add(self, name, ty, vals, valid):
        self.cols.append([name, ty, vals, valid])

    def names(self):
        return [c[0] for c in self.cols]

    def find(self, parts, qualifiers=()):
        # Spark resolves names case-insensitively (spark.sql.caseSensitive=false)
        low = [p.lower() for p in parts]
        owners = getattr(self, "owners", None)
        if owners and len(low) >= 2 and low[0] in [o.lower() for o in owners]:
            n = ".".join(low[1:])
            hits = [c for c, o in zip(self.cols, owners) if o.lower() == low[0] and c[0].lower() == n]
            return hits[0] if len(hits) == 1 else None

        def exact(n):
            hits = [c for c in self.cols if c[0].lower() == n]
            return hits[0] if len(hits) == 1 else None
        quals = [q.lower() for q in qualifiers]
        for strip in range(len(low)):
            if strip > 0 and (strip > 1 or low[0] not in quals):
                break
            n = ".".join(low[strip:])
            h = exact(n)
            if h:
                return h
        return None      # a bare name never reaches into a struct: `deviceId` is not `deviceDetails.deviceId`

    def take(self, idx):
        t = Tab(len(idx))
        for name, ty, v, ok in self.cols:
            t.add(name, ty, v[idx] if v is not None else None, ok[idx])
        t.nest = dict(getattr(self, "nest", {}))
        t.approx = set(getattr(self, "approx", ()))
        if getattr(self, "owners", None):
            t.owners = list(self.owners)
        for a in ("rule_table", "rule_single"):
            if hasattr(self, a):
                setattr(t, a, getattr(self, a))
        return t


def concat_tabs(tabs):
    out = Tab(sum(t.n for t in tabs))
    for c in range(len(tabs[0].cols)):
        name = tabs[0].cols[c][0]
        ty = "null"
        for t in tabs:
            if t.cols[c][1] != "null":
                ty = t.cols[c][1]
                break
        kinds = {t.cols[c][1] for t in tabs}
        if ty == "ts" and "string" in kinds:
            ty = "string"
        vals, oks = [], []
        for t in tabs:
            _, cty, v, ok = t.cols[c]
            if cty == "ts" and ty == "string":     # Spark widens timestamp + string to string (TypeCoercion string promotion)
                v = np.array([ts_to_string(int(x)) if k else None for x, k in zip(v, ok)], dtype=object) if len(v) else np.zeros(0, dtype=object)
                cty = "string"
            if cty not in ("null", ty) and not {cty, ty} <= {"long", "bool"}:
                raise ValueError(f"UNION column '{name}' mixes types {ty} and {cty}")
            if cty == "null" or v is None:
                v = np.zeros(t.n, dtype=object if ty in ("string", "rules") else (np.float64 if ty == "double" else np.int64))
            vals.append(v)
            oks.append(ok)
        out.add(name, ty, np.concatenate(vals) if vals else np.zeros(0), np.concatenate(oks) if oks else np.zeros(0, bool))
    out.nest = dict(tabs[0].nest)
    for t in tabs:      # positional UNION: a column is order-dependent if it is in any branch
        for c, c0 in zip(t.cols, tabs[0].cols):
            if c[0] in getattr(t, "approx", ()):
                out.approx.add(c0[0])
    return out


def _col_refs(e, out):
    if isinstance(e, (list, tuple)):
        if e and e[0] == "col":
            out.append(e[1])
            return
        for x in e:
            _col_refs(x, out)


def _mark_approx(out, src, q, items):
    """an output column computed from an order-dependent input column is order-dependent too"""
    sa = getattr(src, "approx", ())
    if not sa:
        return
    for (e, alias), c in zip(items, out.cols):
        refs = []
        _col_refs(e, refs)
        for parts in refs:
            sc = src.find(parts, q)
            if sc is not None and sc[0] in sa:
                out.approx.add(c[0])


def ts_to_string(us):
    """CAST(timestamp AS STRING) as Spark 2.4 prints it (DateTimeUtils.timestampToString: java.sql.Timestamp.toString with a
    trailing ".0" removed), UTC session zone: yyyy-MM-dd HH:mm:ss plus the fraction without trailing zeros, if any."""
    import datetime
    sec, micro = divmod(us, 1_000_000)
    d = datetime.datetime(1970, 1, 1) + datetime.timedelta(seconds=sec)
    txt = d.strftime("%Y-%m-%d %H:%M:%S")
    if len(txt) < 19:                      # years below 1000 come out unpadded on some platforms
        txt = txt.rjust(19, "0")
    if micro:
        txt += (".%06d" % micro).rstrip("0")
    return txt.encode()


def _like_regex(pat):
    out, i = [], 0
    while i < len(pat):
        c = pat[i]
        if c == "\\" and i + 1 < len(pat):
            out.append(re.escape(pat[i + 1]))
            i += 2
            continue
        out.append(".*" if c == "%" else "." if c == "_" else re.escape(c))
        i += 1
    return re.compile("".join(out), re.S)


class Ev:
    def __init__(self, tab, qualifiers, now_us):
        self.t, self.q, self.now = tab, qualifiers, now_us

    def const(self, ty, v):
        n = self.t.n
        if ty == "string":
            a = np.empty(n, dtype=object)
            a[:] = v
            return ty, a, np.ones(n, bool)
        dt = np.float64 if ty == "double" else np.int64
        return ty, np.full(n, v, dtype=dt), np.ones(n, bool)

    def ev(self, e):
        k = e[0]
        n = self.t.n
        if k == "col":
            c = self.t.find(e[1], self.q)
            if c is None:
                raise KeyError(".".join(e[1]))
            if c[1] == "null":
                return "null", np.zeros(n, np.int64), np.zeros(n, bool)
            return c[1], c[2], c[3]
        if k == "named":
            return self.ev(e[1])
        if k == "int":
            return self.const("long", e[1])
        if k == "dec":
            return self.const("double", e[1])
        if k == "str":
            return self.const("string", e[1].encode("utf-8"))
        if k == "bool":
            return self.const("bool", int(e[1]))
        if k == "null":
            return "null", np.zeros(n, np.int64), np.zeros(n, bool)
        if k == "neg":
            ty, v, ok = self.ev(e[1])
            return ty, -v, ok
        if k == "not":
            ty, v, ok = self.ev(e[1])
            return "bool", (~v.astype(bool)).astype(np.int64), ok
        if k == "isnull":
            ty, v, ok = self.ev(e[1])
            r = ok if e[2] else ~ok
            return "bool", r.astype(np.int64), np.ones(n, bool)
        if k == "in":
            acc = None
            for item in e[2]:
                r = self.ev(("bin", "=", e[1], item))
                acc = r if acc is None else self.kleene("OR", acc, r)
            if e[3]:
                acc = ("bool", (~acc[1].astype(bool)).astype(np.int64), acc[2])
            return acc
        if k == "between":
            r = self.kleene("AND", self.ev(("bin", ">=", e[1], e[2])), self.ev(("bin", "<=", e[1], e[3])))
            if e[4]:
                r = ("bool", (~r[1].astype(bool)).astype(np.int64), r[2])
            return r
        if k == "like":
            ty, v, ok = self.ev(e[1])
            rx = _like_regex(e[2][1])
            r = np.array([bool(ok[i]) and rx.fullmatch(v[i].decode("utf-8", "surrogateescape")) is not None for i in range(n)])
            if e[3]:
                r = ~r
            return "bool", r.astype(np.int64), ok
        if k == "cast":
            ty, v, ok = self.ev(e[1])
            to = e[2]
            if to in ("double", "float"):
                return "double", v.astype(np.float64), ok
            if to in ("long", "bigint", "int", "integer"):
                if ty == "double":
                    with np.errstate(all="ignore"):
                        v = np.where(np.isnan(v), 0, np.clip(np.trunc(v), -9.2233720368547758e18, 9.2233720368547758e18)).astype(np.int64)
                return "long", v.astype(np.int64), ok
            if to in ("string", "varchar"):
                if ty == "string":
                    return ty, v, ok
                out = np.empty(n, dtype=object)
                for i in range(n):
                    out[i] = (ts_to_string(int(v[i])) if ty == "ts" else str(int(v[i])).encode()) if ok[i] else None
                return "string", out, ok
            if to == "timestamp":
                if ty == "ts":
                    return ty, v, ok
                if ty == "double":
                    return "ts", np.trunc(v * 1e6).astype(np.int64), ok
                return "ts", v.astype(np.int64) * 1_000_000, ok
            raise NotImplementedError("cast to " + to)
        if k == "func":
            return self.func(e)
        if k == "bin":
            o = e[1]
            if o in ("AND", "OR"):
                return self.kleene(o, self.ev(e[2]), self.ev(e[3]))
            a, b = self.ev(e[2]), self.ev(e[3])
            if o in ("=", "<>", "<", "<=", ">", ">="):
                return self.compare(o, a, b)
            return self.arith(o, a, b)
        raise NotImplementedError(str(e))

    def kleene(self, o, a, b):
        av, ak = a[1].astype(bool), a[2]
        bv, bk = b[1].astype(bool), b[2]
        if o == "AND":
            f = (ak & ~av) | (bk & ~bv)
            t = ak & av & bk & bv
            return "bool", t.astype(np.int64), f | t
        t = (ak & av) | (bk & bv)
        f = ak & ~av & bk & ~bv
        return "bool", t.astype(np.int64), f | t

    def compare(self, o, a, b):
        (at, av, ak), (bt, bv, bk) = a, b
        ok = ak & bk
        if at == "string" or bt == "string":
            if o not in ("=", "<>"):
                raise NotImplementedError("string ordering")
            eq = np.asarray(av == bv, dtype=bool) if len(av) else np.zeros(0, bool)
            return "bool", (eq if o == "=" else ~eq).astype(np.int64), ok
        if at == "double" or bt == "double":
            x, y = av.astype(np.float64), bv.astype(np.float64)
            xn, yn = np.isnan(x), np.isnan(y)
            with np.errstate(invalid="ignore"):
                lt = np.where(xn, False, np.where(yn, True, x < y))     # NaN is the greatest value
                eq = np.where(xn | yn, xn & yn, x == y)                 # NaN = NaN
        else:
            lt, eq = av < bv, av == bv
        r = {"=": eq, "<>": ~eq, "<": lt, "<=": lt | eq, ">": ~(lt | eq), ">=": ~lt}[o]
        return "bool", r.astype(np.int64), ok

    def arith(self, o, a, b):
        (at, av, ak), (bt, bv, bk) = a, b
        ok = ak & bk
        if o == "/" or at == "double" or bt == "double":
            x, y = av.astype(np.float64), bv.astype(np.float64)
            with np.errstate(all="ignore"):
                if o == "+":
                    r = x + y
                elif o == "-":
                    r = x - y
                elif o == "*":
                    r = x * y
                elif o == "/":
                    ok = ok & (y != 0)
                    r = np.where(y != 0, x / np.where(y == 0, 1, y), 0.0)
                else:
                    ok = ok & (y != 0)
                    r = np.where(y != 0, np.fmod(x, np.where(y == 0, 1, y)), 0.0)
            return "double", r, ok
        x, y = av.astype(np.int64), bv.astype(np.int64)
        with np.errstate(all="ignore"):
            if o == "+":
                r = x + y
            elif o == "-":
                r = x - y
            elif o == "*":
                r = x * y
            else:
                ok = ok & (y != 0)
                ys = np.where(y == 0, 1, y)
                r = np.where(ys == -1, 0, np.fmod(x, ys))
        rt = "ts" if "ts" in (at, bt) and o in ("+", "-") else "long"
        return rt, r, ok

    def func(self, e):
        _, fn, args, distinct, star = e
        n = self.t.n
        if fn == "IF":
            ct, cv, ck = self.ev(args[0])
            a, b = self.ev(args[1]), self.ev(args[2])
            c = ck & cv.astype(bool)
            ty = a[0] if a[0] != "null" else b[0]
            if a[0] != "null" and b[0] != "null" and a[0] != b[0] and "double" in (a[0], b[0]):
                ty = "double"
            av = a[1].astype(np.float64) if ty == "double" else a[1]
            bv = b[1].astype(np.float64) if ty == "double" else b[1]
            return ty, np.where(c, av, bv), np.where(c, a[2], b[2])
        if fn in ("CURRENT_TIMESTAMP", "NOW"):
            return self.const("ts", self.now)
        if fn == "UNIX_TIMESTAMP" and not args:
            return self.const("long", self.now // 1_000_000)
        if fn in ("TO_UNIX_TIMESTAMP", "UNIX_TIMESTAMP"):
            ty, v, ok = self.ev(args[0])
            assert ty == "ts"
            return "long", v // 1_000_000, ok
        if fn == "HOUR":
            ty, v, ok = self.ev(args[0])
            return "long", (v % 86_400_000_000) // 3_600_000_000, ok
        if fn == "DATE_TRUNC":
            assert args[0][0] == "str" and args[0][1].lower() == "second"
            ty, v, ok = self.ev(args[1])
            return "ts", v - (v % 1_000_000), ok
        if fn == "CONCAT" or fn in UDFS:
            vals = [self.ev(a) for a in args]
            out = np.empty(n, dtype=object)
            ok = np.ones(n, bool)
            for i in range(n):
                parts = []
                for ty, v, k in vals:
                    if not k[i]:
                        parts = None
                        break
                    parts.append(v[i] if ty == "string" else (ts_to_string(int(v[i])) if ty == "ts" else str(int(v[i])).encode()))
                if fn == "CONCAT":
                    if parts is None:
                        ok[i] = False
                        out[i] = None
                    else:
                        out[i] = b"".join(parts)
                elif len(vals) == 1:
                    r = UDFS[fn](None if parts is None else parts[0].decode("utf-8"))
                    ok[i] = r is not None
                    out[i] = None if r is None else r.encode("utf-8")
                else:
                    # 0 arguments: one value per batch; 2 or 3: a NULL argument gives NULL without a call (dxgpu.h, dx_flow_register_udfn)
                    if len(vals) == 0:
                        if i == 0:
                            const0 = UDFS[fn]()
                        r = const0
                    else:
                        r = None if parts is None else UDFS[fn](*[p.decode("utf-8") for p in parts])
                    ok[i] = r is not None
                    out[i] = None if r is None else r.encode("utf-8")
            return "string", out, ok
        if fn == "STRINGTOTIMESTAMP":
            ty, v, ok = self.ev(args[0])
            out = np.zeros(n, np.int64)
            ok2 = np.zeros(n, bool)
            for i in range(n):
                if ok[i]:
                    r = string_to_timestamp(v[i].decode("utf-8", "replace"))
                    if r is not None:
                        out[i], ok2[i] = r, True
            return "ts", out, ok2
        raise NotImplementedError(fn)


UDFS = {}      # upper-cased name -> python callable(*str) -> str | None (the flow's registered user functions, 0 to 3 arguments)


AGGS = {"COUNT", "SUM", "MIN", "MAX", "AVG"}


def _has_agg(e):
    if not isinstance(e, tuple):
        return False
    if e[0] == "func" and e[1] in AGGS:
        return True
    return any(_has_agg(x) for x in e[1:] if isinstance(x, tuple)) or any(
        _has_agg(y) for x in e[1:] if isinstance(x, list) for y in x if isinstance(y, tuple))


def _has_col(e):
    if not isinstance(e, tuple):
        return False
    if e[0] in ("col", "star"):
        return True
    return any(_has_col(x) for x in e[1:] if isinstance(x, tuple)) or any(
        _has_col(y) for x in e[1:] if isinstance(x, list) for y in x if isinstance(y, tuple))


def _item_name(e, alias):
    if alias:
        return alias
    if e[0] == "col":
        return e[1][-1]
    if e[0] == "str":
        return e[1]
    return repr(e)


_NESTED = ("MAP", "STRUCT", "NAMED_STRUCT")


def _flatten_nested(e, prefix, path, kinds, items, nest):
    """MAP('k', v, ..) / STRUCT(v AS n, ..) / NAMED_STRUCT('n', v, ..) -> one item per leaf named prefix.k1.k2...;
    nest[name] = (key path, 'm'/'s' per level).  (CodegenRules' alert shaping: pivots / aggs / result / RuleObject.)"""
    fn, args = e[1], e[2]
    step = 1 if fn == "STRUCT" else 2
    if step == 2 and len(args) % 2:
        raise ValueError(fn + " needs key/value pairs")
    for i in range(0, len(args), step):
        if fn == "STRUCT":
            val = args[i]
            if val[0] == "named":
                key, val = val[2], val[1]
            else:
                key = val[1][-1] if val[0] == "col" else "col%d" % (i + 1)
        else:
            if args[i][0] != "str":
                raise ValueError(fn + " keys must be string literals")
            key, val = args[i][1], args[i + 1]
        name = prefix + "." + key
        p2, k2 = path + [key], kinds + ("s" if fn == "STRUCT" else "m")
        if val[0] == "func" and val[1] in _NESTED:
            _flatten_nested(val, name, p2, k2, items, nest)
        else:
            items.append((val, name))
            nest[name] = (p2, k2)


def _rule_shape(e):
    """filterNull(Array(IF(c, MAP(...), NULL), ...)) | IF(c, MAP(...), NULL) -> [(cond, {k: v})]"""
    def one(x):
        if x[0] == "func" and x[1] == "IF" and len(x[2]) == 3 and x[2][2] == ("null",) and x[2][1][0] == "func" and x[2][1][1] == "MAP":
            m = x[2][1][2]
            d = {}
            for i in range(0, len(m), 2):
                if m[i][0] != "str" or m[i + 1][0] not in ("str", "int", "dec"):
                    return None
                d[m[i][1]] = str(m[i + 1][1])
            return x[2][0], d
        return None
    r = one(e)
    if r:
        return [r]
    if e[0] == "func" and e[1] == "FILTERNULL" and len(e[2]) == 1 and e[2][0][0] == "func" and e[2][0][1] == "ARRAY":
        out = [one(a) for a in e[2][0][2]]
        return out if all(out) else None
    return None


def _key_codes(cols):
    """(type, values, valid) list -> int group ids, with NULL keys forming one group."""
    n = len(cols[0][1]) if cols else 0
    if not cols:
        return np.zeros(n, np.int64), 1 if n else 0
    parts = []
    for ty, v, ok in cols:
        if ty == "string":
            uniq = {}
            code = np.array([uniq.setdefault(x, len(uniq)) if k else -1 for x, k in zip(v, ok)], dtype=np.int64)
        elif ty == "double":
            code = np.where(ok, v.view(np.int64), 0)      # Spark 2.4: -0.0 and 0.0 are different group keys
        else:
            code = np.where(ok, v.astype(np.int64), 0)
        parts.append(ok.astype(np.int64))
        parts.append(code)
    mat = np.stack(parts, axis=1) if n else np.zeros((0, len(parts)), np.int64)
    _, first, inv = np.unique(mat, axis=0, return_index=True, return_inverse=True)
    return inv.reshape(-1), len(first)


class CiDict(dict):
    """Table names are case-insensitive."""
    def __getitem__(self, k):
        return dict.__getitem__(self, k.lower())

    def __setitem__(self, k, v):
        dict.__setitem__(self, k.lower(), v)

    def __contains__(self, k):
        return dict.__contains__(self, k.lower())


def _flatten_and(e, out):
    if e[0] == "bin" and e[1] == "AND":
        _flatten_and(e[2], out)
        _flatten_and(e[3], out)
    else:
        out.append(e)


def _to_long(b):
    try:
        t = b.decode("utf-8").strip()
        if not re.fullmatch(r"[+-]?\d+", t):
            return None
        return int(t)
    except Exception:
        return None


def _join(core, views, now_us):
    """[INNER] JOIN ... ON a = b [AND c = d]: NULL keys never match; a string key against a numeric one is
    cast to the numeric type (unparsable -> NULL), as Spark's implicit cast does."""
    left = views[core["from"]]
    lname = core["alias"] or core["from"]
    owners = list(getattr(left, "owners", None) or [lname] * len(left.cols))
    acc = left
    for jt, ja, on, outer in core["joins"]:
        right = views[jt]
        rname = ja or jt
        acc.owners = owners
        conj = []
        _flatten_and(on, conj)
        lkeys, rkeys = [], []
        for c in conj:
            assert c[0] == "bin" and c[1] == "=" and c[2][0] == "col" and c[3][0] == "col", c
            a, b = acc.find(c[2][1]), right.find(c[3][1], [jt, rname])
            if a is None or b is None:
                a, b = acc.find(c[3][1]), right.find(c[2][1], [jt, rname])
            assert a is not None and b is not None, c
            lk = [(x if k else None) for x, k in zip(a[2], a[3])]
            rk = [(x if k else None) for x, k in zip(b[2], b[3])]
            if a[1] != "string" and b[1] == "string":
                rk = [None if x is None else _to_long(x) for x in rk]
            elif a[1] == "string" and b[1] != "string":
                lk = [None if x is None else _to_long(x) for x in lk]
            lkeys.append([None if x is None else (x if isinstance(x, bytes) else int(x)) for x in lk])
            rkeys.append([None if x is None else (x if isinstance(x, bytes) else int(x)) for x in rk])
        index = {}
        for r in range(right.n):
            k = tuple(col[r] for col in rkeys)
            if None not in k:
                index.setdefault(k, []).append(r)
        li, ri = [], []
        for l in range(acc.n):
            k = tuple(col[l] for col in lkeys)
            hits = () if None in k else index.get(k, ())
            for r in hits:
                li.append(l)
                ri.append(r)
            if outer and not hits:          # LEFT OUTER: the unmatched left row once, right side NULL
                li.append(l)
                ri.append(-1)
        li, ri = np.array(li, dtype=np.int64), np.array(ri, dtype=np.int64)
        out = Tab(len(li))
        for c in acc.cols:
            out.add(c[0], c[1], c[2][li] if c[2] is not None else None, c[3][li])
        miss = ri < 0
        rsafe = np.where(miss, 0, ri)
        for c in right.cols:
            if right.n == 0:
                vals = None if c[2] is None else np.zeros(len(ri), dtype=c[2].dtype)
                out.add(c[0], c[1], vals, np.zeros(len(ri), bool))
            else:
                out.add(c[0], c[1], c[2][rsafe] if c[2] is not None else None, c[3][rsafe] & ~miss)
        owners = owners + [rname] * len(right.cols)
        acc = out
    acc.owners = owners
    return acc


def _having_rewrite(e, inner, visible):
    """Aggregates and bare columns of a HAVING condition -> references to (possibly hidden) items of the aggregation."""
    if not isinstance(e, tuple):
        return e
    if (e[0] == "func" and e[1] in AGGS) or e[0] == "col":
        for x, alias in inner:
            if x == e:
                return ("col", [_item_name(x, alias)])
        if e[0] == "col" and len(e[1]) == 1 and any(_item_name(x, a) == e[1][0] for x, a in visible):
            return e
        name = "__having%d" % len(inner)
        inner.append((e, name))
        return ("col", [name])
    return tuple([_having_rewrite(y, inner, visible) for y in x] if isinstance(x, list) else _having_rewrite(x, inner, visible) for x in e)


def exec_core(core, views, now_us):
    if core.get("having") is not None:
        inner_items = list(core["items"])
        cond = _having_rewrite(core["having"], inner_items, core["items"])
        v2 = CiDict(views)
        v2["__having"] = exec_core(dict(core, items=inner_items, having=None), views, now_us)
        outer_items = [(("col", [_item_name(e, a)]), _item_name(e, a)) for e, a in core["items"]]
        return exec_core({"distinct": False, "items": outer_items, "from": "__having", "alias": None, "where": cond, "group": [],
                          "joins": [], "having": None}, v2, now_us)
    if any(e[0] == "func" and e[1] in _NESTED for e, _ in core["items"]):
        items, nest = [], {}
        for e, alias in core["items"]:
            if e[0] == "func" and e[1] in _NESTED:
                if not alias:
                    raise ValueError("a MAP/STRUCT select item needs an alias")
                _flatten_nested(e, alias, [], "", items, nest)
            else:
                items.append((e, alias))
        out = exec_core(dict(core, items=items), views, now_us)
        out.nest = dict(getattr(out, "nest", {}))
        for c in out.cols:
            if c[0] in nest:
                out.nest[c[0]] = nest[c[0]]
        return out
    if core.get("joins"):
        v2 = CiDict(views)
        v2["__join"] = _join(core, views, now_us)
        rest = dict(core, joins=[], alias=None)
        rest["from"] = "__join"
        return exec_core(rest, v2, now_us)
    src = views[core["from"]]
    q = [core["from"]] + ([core["alias"]] if core["alias"] else []) + ["Raw"]
    ev = Ev(src, q, now_us)
    if core["where"] is not None:
        ty, v, ok = ev.ev(core["where"])
        keep = np.nonzero(ok & v.astype(bool))[0]
        src = src.take(keep)
        ev = Ev(src, q, now_us)
    items = []
    for e, alias in core["items"]:
        if e[0] == "star" and (core["distinct"] or core["group"]):
            for c in src.cols:
                items.append((("col", [c[0]]), c[0]))
        else:
            items.append((e, alias))
    agg = core["distinct"] or core["group"] or any(_has_agg(e) for e, _ in items)
    if not agg:
        out = Tab(src.n)
        for e, alias in items:
            if e[0] == "star":
                parts = e[1]
                if core.get("projection") and parts == ["Raw"]:
                    cols = [c for c in src.cols if not is_metadata_column(c[0])]     # Raw.* : the decoded leaves, not the frame around Raw
                elif not parts or parts[0] in q:
                    cols = src.cols
                else:
                    prefix = ".".join(parts) + "."
                    if prefix.startswith("Raw."):
                        prefix = prefix[4:]
                    cols = [c for c in src.cols if c[0].startswith(prefix)]
                for c in cols:
                    out.add(*c)
                    if c[0] in getattr(src, "approx", ()):
                        out.approx.add(c[0])
                    if c[0] in getattr(src, "nest", {}):
                        out.nest = dict(getattr(out, "nest", {}))
                        out.nest[c[0]] = src.nest[c[0]]
                continue
            rs = _rule_shape(e)
            if rs:
                flags = np.zeros((src.n, len(rs)), dtype=bool)
                for r, (cond, _) in enumerate(rs):
                    ty, v, ok = ev.ev(cond)
                    flags[:, r] = ok & v.astype(bool)
                out.add(_item_name(e, alias), "rules", flags, np.ones(src.n, bool))
                out.rule_table = [d for _, d in rs]
                out.rule_single = not (e[0] == "func" and e[1] == "FILTERNULL")     # a lone IF(cond, MAP(...), NULL): a nullable map
                continue
            if e[0] == "col" and src.find(e[1], q) is None and (
                    e[1][0] in ("Properties", "SystemProperties", "__DataX_FileInfo", "DataXProperties") or
                    any(c[0].startswith(".".join(e[1]) + ".") and c[0] in src.nest for c in src.cols)):
                # a whole struct / map of transport metadata: its leaves renamed under the item's name; none -> a NULL column
                prefix = ".".join(e[1]) + "."
                hits = [c for c in src.cols if c[0].startswith(prefix)]
                for c in hits:
                    nm = _item_name(e, alias) + "." + c[0][len(prefix):]
                    out.add(nm, c[1], c[2], c[3])
                    if c[0] in src.nest:
                        out.nest[nm] = src.nest[c[0]]
                if not hits:
                    out.add(_item_name(e, alias), "null", None, np.zeros(src.n, bool))
                continue
            ty, v, ok = ev.ev(e)
            out.add(_item_name(e, alias), ty, v, ok)
            refs = []
            _col_refs(e, refs)
            if any((src.find(pp, q) or [None])[0] in getattr(src, "approx", ()) for pp in refs):
                out.approx.add(_item_name(e, alias))
        return out
    # aggregation
    gexprs = []
    if core["group"]:
        for g in core["group"]:
            ge = g
            if g[0] == "col" and src.find(g[1], q) is None and len(g[1]) == 1:
                for e, alias in items:
                    if _item_name(e, alias) == g[1][0] and not _has_agg(e):
                        ge = e
                        break
            gexprs.append(ge)
    elif core["distinct"]:
        gexprs = [e for e, _ in items]
    keycols = [ev.ev(g) for g in gexprs]
    gid, ng = _key_codes(keycols)
    if not gexprs:
        gid, ng = np.zeros(src.n, np.int64), 1          # global aggregate: one row even over zero rows
    first = np.full(ng, -1, np.int64)
    if src.n:
        order = np.arange(src.n)
        first_seen = np.full(ng, src.n, np.int64)
        np.minimum.at(first_seen, gid, order)
        first = first_seen
    out = Tab(ng)
    for e, alias in items:
        name = _item_name(e, alias)
        if e[0] == "func" and e[1] in AGGS:
            _, fn, args, distinct, star = e
            if distinct:
                assert fn == "COUNT"
                ty, v, ok = ev.ev(args[0])
                xk, _ = _key_codes([(ty, v, ok)])
                pairs = np.unique(np.stack([gid[ok], xk[ok]], axis=1), axis=0) if ok.any() else np.zeros((0, 2), np.int64)
                cnt = np.zeros(ng, np.int64)
                np.add.at(cnt, pairs[:, 0], 1)
                out.add(name, "long", cnt, np.ones(ng, bool))
                continue
            if star or (fn == "COUNT" and args and args[0][0] in ("int", "str", "dec", "bool")):
                cnt = np.zeros(ng, np.int64)
                np.add.at(cnt, gid, 1)
                out.add(name, "long", cnt, np.ones(ng, bool))
                continue
            ty, v, ok = ev.ev(args[0])
            cnt = np.zeros(ng, np.int64)
            np.add.at(cnt, gid[ok], 1)
            if fn == "COUNT":
                out.add(name, "long", cnt, np.ones(ng, bool))
            elif fn == "SUM":
                if ty == "double":
                    s = np.zeros(ng, np.float64)
                    np.add.at(s, gid[ok], v[ok])
                    out.add(name, "double", s, cnt > 0)
                    out.approx.add(name)            # the sum of doubles depends on the order rows arrive in
                else:
                    s = np.zeros(ng, np.int64)
                    with np.errstate(over="ignore"):
                        np.add.at(s, gid[ok], v[ok].astype(np.int64))
                    out.add(name, "long", s, cnt > 0)
            elif fn == "AVG":
                s = np.zeros(ng, np.float64)
                np.add.at(s, gid[ok], v[ok].astype(np.float64))
                with np.errstate(all="ignore"):
                    out.add(name, "double", np.where(cnt > 0, s / np.maximum(cnt, 1), 0.0), cnt > 0)
                if ty == "double":
                    out.approx.add(name)            # AVG(double) = order-dependent sum / count; AVG(long) sums exactly representable values up to 2^53
            else:
                if ty == "double":
                    x = v[ok]
                    key = np.where(np.isnan(x), np.inf, x)            # NaN greatest
                    isnan = np.isnan(x)
                    if fn == "MIN":
                        r = np.full(ng, np.inf)
                        np.minimum.at(r, gid[ok], key)
                        allnan = np.ones(ng, bool)
                        np.logical_and.at(allnan, gid[ok], isnan)
                        r = np.where(allnan & (cnt > 0), np.nan, r)
                    else:
                        r = np.full(ng, -np.inf)
                        np.maximum.at(r, gid[ok], key)
                        anynan = np.zeros(ng, bool)
                        np.logical_or.at(anynan, gid[ok], isnan)
                        r = np.where(anynan, np.nan, r)
                    out.add(name, "double", np.where(cnt > 0, r, 0.0), cnt > 0)
                else:
                    x = v[ok].astype(np.int64)
                    if fn == "MIN":
                        r = np.full(ng, np.iinfo(np.int64).max)
                        np.minimum.at(r, gid[ok], x)
                    else:
                        r = np.full(ng, np.iinfo(np.int64).min)
                        np.maximum.at(r, gid[ok], x)
                    out.add(name, ty, np.where(cnt > 0, r, 0), cnt > 0)
            continue
        hit = None
        for g, kc in zip(gexprs, keycols):
            if g == e or (g[0] == "col" and e[0] == "col" and src.find(g[1], q) is src.find(e[1], q) and src.find(e[1], q) is not None):
                hit = kc
                break
        if hit is None:
            if not _has_col(e):      # literals and foldable expressions (current_timestamp() ...) need no grouping
                ty, v, ok = Ev(Tab(ng), q, now_us).ev(e)
                out.add(name, ty, v, ok)
                continue
            raise ValueError(f"{e} is neither aggregated nor in GROUP BY")
        ty, v, ok = hit
        if ng and src.n:
            out.add(name, ty, v[first], ok[first])
        else:
            out.add(name, ty, v[:0], ok[:0])
    # an aggregate (MIN, MAX, group key) over an order-dependent input stays order-dependent
    for (e, alias), c in zip(items, out.cols):
        refs = []
        _col_refs(e, refs)
        if any((src.find(pp, q) or [None])[0] in getattr(src, "approx", ()) for pp in refs):
            out.approx.add(c[0])
    for (e, alias), c in zip(items, out.cols):      # DISTINCT / GROUP BY over MAP/STRUCT leaves keeps their shaping
        if e[0] == "col":
            sc = src.find(e[1], q)
            if sc is not None and sc[0] == c[0] and sc[0] in src.nest:
                out.nest[c[0]] = src.nest[sc[0]]
    return out


def exec_statement(sql, views, now_us):
    cores, alls = P(sql).statement()
    tabs = [exec_core(c, views, now_us) for c in cores]
    if len(tabs) == 1:
        return tabs[0]
    u = concat_tabs(tabs)
    if all(alls):
        return u
    v2 = CiDict(views)
    v2["__union"] = u
    return exec_core({"distinct": True, "items": [(("star", []), None)], "from": "__union", "alias": None, "where": None, "group": []}, v2, now_us)


# --------------------------------------------------------------------------------------------------
# the processor (CommonProcessorFactory.createProcessor / processDataset / project / route)
# --------------------------------------------------------------------------------------------------
_STATE_TYPES = {"long": "long", "bigint": "long", "int": "long", "integer": "long", "double": "double", "float": "double",
                "string": "string", "timestamp": "ts", "boolean": "bool"}


class Flow:
    def __init__(self, conf_text, schema_json, projection_steps, transform_text):
        self.conf = {}
        for line in conf_text.lstrip("﻿").splitlines():
            line = line.strip()
            if line and not line.startswith("#") and "=" in line:
                k, v = line.split("=", 1)
                self.conf[k.strip()] = v.strip()
        self.schema = dxo.Schema(schema_json)
        self.projections = [[P(l.strip()).select_item() for l in step.lstrip("﻿").splitlines() if l.strip()] for step in projection_steps]
        self.projections = [p for p in self.projections if p]
        self.commands, self.ref_counts = parse_transform(transform_text)
        root = "datax.job.process."
        self.windows = {}
        self.state = {}
        for k, v in self.conf.items():
            m = re.fullmatch(re.escape(root) + r"timewindow\.(.+)\.windowduration", k)
            if m:
                self.windows[m.group(1)] = parse_duration(v)
            m = re.fullmatch(re.escape(root) + r"statetable\.(.+)\.schema", k)
            if m:
                cols = []
                for part in v.split(","):
                    nm, ty = part.strip().replace(":", " ").split()[:2]
                    cols.append((nm, _STATE_TYPES[ty.lower()]))
                empty = Tab(0)
                for nm, ty in cols:
                    empty.add(nm, ty, np.zeros(0, dtype=object if ty == "string" else (np.float64 if ty == "double" else np.int64)), np.zeros(0, bool))
                self.state[m.group(1)] = {"schema": cols, "tables": {"A": empty, "B": empty}, "active": "A", "standby": "B"}
        wm, tc = self.conf.get(root + "watermark"), self.conf.get(root + "timestampcolumn")
        self.enabled = bool(self.windows) and wm is not None and tc is not None
        self.watermark = parse_duration(wm) if self.enabled else 0
        self.ts_col = tc
        self.max_window = max(self.windows.values()) if self.enabled else 0
        self.past = {}        # namespace -> {batch_ms: Tab}
        # transport metadata of the batch about to be processed (set_batch_info; None = FileInternal(), CPF:417) and the conf's
        # appendproperty.* entries (PropertiesHandler.scala:22-24)
        self.batch_info = None
        self.append_properties = [(k[len(root + "appendproperty."):], v) for k, v in self.conf.items() if k.startswith(root + "appendproperty.")]
        # outputs (OutputManager.getOperatators: one per name under datax.job.output.<name>.<sink>.*), in order of appearance
        self.outputs = []
        for k in self.conf:
            m = re.fullmatch(r"datax\.job\.output\.([^.]+)\..+", k)
            if m and m.group(1) not in self.outputs:
                self.outputs.append(m.group(1))

    def set_batch_info(self, info):
        """info: None or a dict with any of input_path, output_folders (dict), output_file_name, file_time_ms, rule_index_prefix,
        target, properties (dict), system_properties (dict), executor_id — the per-batch constants of dx_flow_set_batch_info."""
        self.batch_info = info

    def set_event_property(self, key, values, system=False):
        """Per-event Properties / SystemProperties values of the next batch (CPF:405-421, 628-654): bytes / str, or None."""
        if not hasattr(self, "event_props"):
            self.event_props = []
        name = ("SystemProperties." if system else "Properties.") + key
        vals = [None if v is None else (v.encode("utf-8") if isinstance(v, str) else bytes(v)) for v in values]
        self.event_props = [e for e in self.event_props if e[0] != name] + [(name, key, vals)]

    def _add_metadata(self, t, batch_ms, now_ms):
        b = self.batch_info or {}
        n = t.n

        def const(name, val, path, kinds, ty="string"):
            if val is None:
                t.add(name, "null", None, np.zeros(n, bool))
            elif ty == "ts":
                t.add(name, "ts", np.full(n, val, np.int64), np.ones(n, bool))
            else:
                v = np.empty(n, dtype=object)
                v[:] = val.encode("utf-8")
                t.add(name, "string", v, np.ones(n, bool))
            t.nest[name] = (path, kinds)
        fi = "__DataX_FileInfo"
        const(fi + ".inputPath", b.get("input_path"), ["inputPath"], "s")
        if b.get("output_folders"):
            for k, v in b["output_folders"].items():
                const(fi + ".outputFolders." + k, v, ["outputFolders", k], "sm")
        else:
            const(fi + ".outputFolders", None, ["outputFolders"], "s")
        const(fi + ".outputFileName", b.get("output_file_name"), ["outputFileName"], "s")
        const(fi + ".fileTime", None if b.get("file_time_ms") is None else b["file_time_ms"] * 1000, ["fileTime"], "s", "ts")
        const(fi + ".ruleIndexPrefix", b.get("rule_index_prefix"), ["ruleIndexPrefix"], "s")
        const(fi + ".target", b.get("target"), ["target"], "s")
        props = list(self.append_properties)

        def put(k, v):
            for i, (k2, _) in enumerate(props):
                if k2 == k:
                    props[i] = (k, v)
                    return
            props.append((k, v))
        put("BatchTime", java_timestamp_string(batch_ms))
        if b.get("file_time_ms") is not None:
            put("BlobTime", java_timestamp_string(b["file_time_ms"]))
        if b.get("output_file_name") is not None:
            put("BlobPathHint", b["output_file_name"])
        put("CPTime", java_timestamp_string(now_ms))
        put("CPExecutor", b.get("executor_id") or "driver")
        for k, v in props:
            const("DataXProperties." + k, v, [k], "m")
        for k, v in (b.get("properties") or {}).items():
            const("Properties." + k, v, [k], "m")
        for k, v in (b.get("system_properties") or {}).items():
            const("SystemProperties." + k, v, [k], "m")
        for name, key, vals in getattr(self, "event_props", []):
            if len(vals) != n:
                raise ValueError("event property '%s' was given for %d events, the batch has %d" % (name, len(vals), n))
            v = np.empty(n, dtype=object)
            v[:] = [x if x is not None else b"" for x in vals]
            col = [name, "string", v, np.array([x is not None for x in vals], dtype=bool)]
            hit = [i for i, c in enumerate(t.cols) if c[0] == name]
            if hit:
                t.cols[hit[0]] = col
            else:
                t.cols.append(col)
            t.nest[name] = ([key], "m")
        self.event_props = []

    def add_refdata(self, name, csv_text, delimiter=",", header=True):
        """ReferenceDataHandler / CSVUtil.scala:27-29: every column is a string; empty field -> NULL."""
        import csv
        import io
        rows = [r for r in csv.reader(io.StringIO(csv_text.lstrip("\ufeff")), delimiter=delimiter) if r]
        names = rows[0] if header else ["_c%d" % i for i in range(len(rows[0]))]
        body = rows[1:] if header else rows
        t = Tab(len(body))
        for c, nm in enumerate(names):
            v = np.empty(len(body), dtype=object)
            ok = np.zeros(len(body), bool)
            for r, row in enumerate(body):
                if c < len(row) and row[c] != "":
                    v[r], ok[r] = row[c].encode("utf-8"), True
            t.add(nm, "string", v, ok)
        if not hasattr(self, "refdata"):
            self.refdata = {}
        self.refdata[name] = t

    def decode(self, data, offsets=None):
        d = dxo.decode(self.schema, data, offsets)
        n = len(d["__rowok"])
        t = Tab(n)
        tmap = {dxo.LONG: "long", dxo.DOUBLE: "double", dxo.STRING: "string", dxo.BOOL: "bool"}
        for name, ty in zip(self.schema.leaf_names, self.schema.leaf_types):
            v, ok = d[name]
            t.add(name, tmap[ty], v, ok)
        return t

    def process(self, data, batch_ms, interval_ms, now_ms=None, namespace="", offsets=None):
        import time
        t_start = time.perf_counter()
        now_us = (batch_ms if now_ms is None else now_ms) * 1000
        metrics = {}
        projected = self.decode(data, offsets)
        # the frame around Raw (CPF:405-421, 458-461) and what project() adds before the steps (CPF:97)
        self._add_metadata(projected, batch_ms, batch_ms if now_ms is None else now_ms)
        if not self.projections:
            keep = Tab(projected.n)
            for c in projected.cols:
                if not is_metadata_column(c[0]) or c[0].startswith("__DataX_"):
                    keep.add(*c)
                    if c[0] in projected.nest:
                        keep.nest[c[0]] = projected.nest[c[0]]
            projected = keep
        for step in self.projections:
            core = {"distinct": False, "items": step, "from": "Raw", "alias": None, "where": None, "group": [], "projection": True}
            nxt = exec_core(core, {"Raw": projected}, now_us)
            for c in projected.cols:          # selectExpr(projection ++ preservedColumns) (CPF:95-100)
                if c[0].startswith("__DataX_") and c[0] not in nxt.names():
                    nxt.add(*c)
                    if c[0] in projected.nest:
                        nxt.nest[c[0]] = projected.nest[c[0]]
            projected = nxt
        metrics["Input_Normalized_Events_Count"] = float(projected.n)
        if dxo.last_limited:          # events made NULL rows by a limit of the implementation, not by Spark's rules
            metrics["Input_Unparsed_Events_Count"] = float(dxo.last_limited)
        views = CiDict()
        for rn, rt in getattr(self, 'refdata', {}).items():
            views[rn] = rt
        if self.enabled:
            T = batch_ms
            E = T - self.watermark * 1000
            S = E - self.max_window * 1000
            tsc = projected.find([self.ts_col])
            keep = np.nonzero(tsc[3] & (tsc[2] >= E * 1000))[0]
            rdd = projected.take(keep)
            metrics["Input_DataXProcessedInput_Events_Count"] = float(rdd.n)
            past = self.past.setdefault(namespace, {})
            cut = T - (self.watermark + self.max_window) * 1000
            for k in [k for k in past if k <= cut]:
                del past[k]
            union = concat_tabs([rdd] + list(past.values())) if len(past) > 1 else rdd     # CPF:198-201 quirk

            def ranged(lo_ms, hi_ms):
                c = union.find([self.ts_col])
                idx = np.nonzero(c[3] & (c[2] >= lo_ms * 1000) & (c[2] < hi_ms * 1000))[0]
                return union.take(idx)
            views["DataXProcessedInput_Window"] = ranged(S, E)
            for wname, dur in self.windows.items():
                views[wname] = ranged(E - dur * 1000, E)
            views["DataXProcessedInput"] = ranged(E - (interval_ms // 1000) * 1000, E)
            views["DataXProcessedInput_Batch"] = projected
            past[T] = rdd
        else:
            metrics["Input_DataXProcessedInput_Events_Count"] = float(projected.n)
            views["DataXProcessedInput"] = projected
        for name, st in self.state.items():
            views[name] = st["tables"][st["active"]]
        datasets = {}
        for name, kind, sql in self.commands:
            if kind != "Query":
                raise NotImplementedError("command statements")
            res = exec_statement(sql, views, now_us)
            if name in self.state:
                st = self.state[name]
                for c, (nm, _) in zip(res.cols, st["schema"]):
                    c[0] = nm
                st["tables"][st["standby"]] = res                      # INSERT OVERWRITE standby
                st["active"], st["standby"] = st["standby"], st["active"]   # flip
            views[name] = res
            datasets[name] = res
        for o in self.outputs:           # CPF:297-306 + OutputManager.scala:121-122: the sink stage's input count per output
            if o not in datasets:
                raise ValueError(f"could not find data set name '{o}' for output '{o}'")
            metrics[f"Output_{o}_Sink_InputEvents"] = float(datasets[o].n)
        datasets["DataXProcessedInput_Batch"] = projected
        metrics["Latency-Process"] = time.perf_counter() - t_start              # CPF:372-377 (wall-clock values: never compared)
        metrics["Latency-Batch"] = (time.time() * 1000.0 - batch_ms) / 1000.0
        return datasets, metrics


# --------------------------------------------------------------------------------------------------
# multi-GPU key ownership (stands in for Spark's hashpartitioning(keys) Exchange)
# --------------------------------------------------------------------------------------------------
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix64(z):
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def owner_of(key_cols, world):
    """key_cols: list of (int64 values, valid bool).  -> owner rank per row, the same function as the
    product's K5 kernel: 64-bit mix over the key values (NULL = fixed constant), multiply-shift to [0, world)."""
    n = len(key_cols[0][0])
    h = np.full(n, 0x6A09E667F3BCC909, dtype=np.uint64)
    for v, ok in key_cols:
        x = np.where(ok, v.astype(np.int64).view(np.uint64), np.uint64(0x5BD1E9955BD1E995))
        h = _mix64(h ^ x)
    return (((h >> np.uint64(32)) * np.uint64(world)) >> np.uint64(32)).astype(np.int64)
