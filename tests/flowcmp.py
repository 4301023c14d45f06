"""Compare a device Table (through the C ABI) with an oracle Tab as multisets of rows."""
import math

import numpy as np

REL_TOL = 1e-6      # ONLY for columns the oracle marks order-dependent (Tab.approx: SUM / AVG of doubles and what is computed
                    # from them; BASELINE.json north_star "1e-6 fp64, order-dependent sums").  Every other double — parsed
                    # values, MIN / MAX, computed columns such as temperature * 1.8 + 32 — is compared bit for bit.


def gpu_rows(table, dx):
    cols = []
    for i, name in enumerate(table.columns):
        t = table.column_type(i)
        if t == 0 or t == 7:
            cols.append([None] * table.num_rows)
            continue
        v, ok = table.column(i)
        if t == 6:
            nr = len(dx.rules_of(table, i))
            bits = np.unpackbits(v.view(np.uint8), axis=1, bitorder="little")[:, :nr].astype(bool) if table.num_rows else np.zeros((0, nr), bool)
            cols.append([tuple(np.nonzero(r)[0].tolist()) for r in bits])
        elif t == 2:
            cols.append([float(x) if k else None for x, k in zip(v, ok)])
        elif t == 3:
            cols.append([x if k else None for x, k in zip(v, ok)])
        else:
            cols.append([int(x) if k else None for x, k in zip(v, ok)])
    return [tuple(c[r] for c in cols) for r in range(table.num_rows)]


def oracle_rows(tab):
    cols = []
    for name, ty, v, ok in tab.cols:
        if ty == "null" or v is None:
            cols.append([None] * tab.n)
        elif ty == "rules":
            cols.append([tuple(np.nonzero(r)[0].tolist()) for r in v])
        elif ty == "double":
            cols.append([float(x) if k else None for x, k in zip(v, ok)])
        elif ty == "string":
            cols.append([x if k else None for x, k in zip(v, ok)])
        else:
            cols.append([int(x) if k else None for x, k in zip(v, ok)])
    return [tuple(c[r] for c in cols) for r in range(tab.n)]


def _key(row):
    out = []
    for x in row:
        if x is None:
            out.append((0, 0))
        elif isinstance(x, float):
            out.append((1, 0))                    # doubles do not take part in the ordering (tolerance)
        elif isinstance(x, bytes):
            out.append((2, x))
        elif isinstance(x, tuple):
            out.append((3, x))
        else:
            out.append((4, x))
    return out


def _fkey(row):
    return [x if isinstance(x, float) and not math.isnan(x) else 0.0 for x in row if isinstance(x, float)]


def _bits(x):
    import struct
    return struct.unpack("<q", struct.pack("<d", x))[0]


def assert_same_rows(table, tab, dx, ordered=False):
    assert table.columns == tab.names(), (table.columns, tab.names())
    g, o = gpu_rows(table, dx), oracle_rows(tab)
    assert len(g) == len(o), (len(g), len(o))
    approx = [name in getattr(tab, "approx", ()) for name in tab.names()]
    exact_f = [k for k, a in enumerate(approx) if not a]

    def exact_floats(row):       # exact double columns take part in the ordering (by bit pattern); tolerant ones only break ties
        return [_bits(row[k]) if isinstance(row[k], float) else 0 for k in exact_f]
    if not ordered:
        g.sort(key=lambda r: (_key(r), exact_floats(r), _fkey(r)))
        o.sort(key=lambda r: (_key(r), exact_floats(r), _fkey(r)))
    for a, b in zip(g, o):
        for k, (x, y) in enumerate(zip(a, b)):
            if isinstance(x, float) and isinstance(y, float):
                if math.isnan(x) or math.isnan(y):
                    assert math.isnan(x) and math.isnan(y), (tab.names()[k], a, b)
                elif approx[k]:
                    assert x == y or abs(x - y) <= REL_TOL * max(abs(x), abs(y)), (tab.names()[k], a, b)
                else:
                    assert _bits(x) == _bits(y), (tab.names()[k], "exact double column differs", a, b)
            else:
                assert x == y, (tab.names()[k], a, b)


def same_metrics(a, b):
    """The Map[String, Double] processDataset returns: same keys, same values — except Latency-Process / Latency-Batch
    (CommonProcessorFactory.scala:372-377), which are wall-clock readings: present, non-negative where meaningful, not compared."""
    if set(a) != set(b):
        return False
    for k in a:
        if k.startswith("Latency-"):
            if k == "Latency-Process" and not (a[k] >= 0.0 and b[k] >= 0.0):
                return False
            continue
        if a[k] != b[k]:
            return False
    return "Latency-Process" in a and "Latency-Batch" in a
