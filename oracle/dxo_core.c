/*
 * dxo_core.c — CPU ORACLE (test infrastructure, NOT product code).
 *
 * Plain-C restatement of the arithmetic on Data Accelerator's per-micro-batch hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this file's shared object.  The product (libdxgpu.so) never links or calls it.
 *
 * PARITY UNPINNED at the Spark boundary: the reference delegates JSON decode, predicate
 * evaluation and aggregation to Apache Spark SQL 2.4.5 (pom.xml:77-99, `provided`, not
 * vendored; no JVM in this image), and no reference test asserts rows produced by
 * project()/route().  What IS pinned (tests/test_oracle_kat.py): TransformSQLParserTests,
 * ConcurrentDateFormatTests, the sample events and the config goldens.
 *
 * Each function cites the reference call site it stands in for:
 *   dxo_decode      from_json(Raw, rawBlockSchema)      CommonProcessorFactory.scala:93
 *                   (Spark 2.4.5 JacksonParser semantics, written out in DESIGN.md §3)
 *   dxo_split       record framing of an NDJSON batch   CommonProcessorFactory.scala:458-461
 *   dxo_gen_*       synthetic events in the shape of    DataGenerator.scala:27-209 /
 *                   HomeAutomationLocal.json:288 (schema + allowedValues)
 *   dxo_group_agg   spark.sql("... GROUP BY ...")       CommonProcessorFactory.scala:271
 *                   (HashAggregateExec semantics: COUNT/SUM/MIN/MAX/AVG, NULL handling)
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>
#include <pthread.h>

/* ------------------------------------------------------------------------------------ */
/* schema model (built by oracle/dxo.py from the Spark StructType JSON)                  */
/* ------------------------------------------------------------------------------------ */
enum { DXO_LONG = 1, DXO_DOUBLE = 2, DXO_STRING = 3, DXO_BOOL = 4, DXO_STRUCT = 5 };

typedef struct {
    const char *name;
    int32_t name_len;
    int32_t type;
    int32_t target; /* leaf: flattened column index; struct: index into structs[] */
} dxo_field;

typedef struct {
    int32_t first_field;
    int32_t n_fields;
    int32_t first_leaf; /* leaves of this struct (recursively) are [first_leaf, first_leaf+n_leaves) */
    int32_t n_leaves;
} dxo_struct;

typedef struct {
    const dxo_struct *structs;
    const dxo_field *fields;
    int32_t n_leaves;
    /* per-row outputs for the row being parsed */
    int64_t *vals;   /* [n_leaves] (doubles bit-cast; strings: offset into heap) */
    int32_t *lens;   /* [n_leaves] string length */
    uint8_t *valid;  /* [n_leaves] */
    /* string heap */
    uint8_t *heap;
    int64_t heap_used;
    int64_t heap_cap;
    int limited;     /* the record being parsed hit a limit of the restatement, not a Spark rule */
} dxo_ctx;

typedef struct {
    const uint8_t *p, *end;
} cur_t;

static inline void skip_ws(cur_t *c) {
    while (c->p < c->end) {
        uint8_t b = *c->p;
        if (b == ' ' || b == '\t' || b == '\n' || b == '\r') c->p++;
        else break;
    }
}

/* Parses a JSON string token starting at the opening quote (either " or ', Spark's
 * allowSingleQuotes defaults to true).  Unescaped text is appended to out (if non-NULL).
 * Returns 0 on success. */
static int parse_string(cur_t *c, uint8_t *out, int64_t *out_len) {
    uint8_t q = *c->p++;
    int64_t n = 0;
    for (;;) {
        if (c->p >= c->end) return -1;
        uint8_t b = *c->p++;
        if (b == q) break;
        if (b < 0x20) return -1; /* unescaped control character: rejected */
        if (b != '\\') {
            if (out) out[n] = b;
            n++;
            continue;
        }
        if (c->p >= c->end) return -1;
        uint8_t e = *c->p++;
        uint32_t cp;
        switch (e) {
        case '"': cp = '"'; break;
        case '\\': cp = '\\'; break;
        case '/': cp = '/'; break;
        case 'b': cp = 8; break;
        case 'f': cp = 12; break;
        case 'n': cp = 10; break;
        case 'r': cp = 13; break;
        case 't': cp = 9; break;
        case '\'': if (q == '\'') { cp = '\''; break; } return -1;
        case 'u': {
            if (c->end - c->p < 4) return -1;
            cp = 0;
            for (int k = 0; k < 4; k++) {
                uint8_t h = *c->p++;
                uint32_t d;
                if (h >= '0' && h <= '9') d = h - '0';
                else if (h >= 'a' && h <= 'f') d = h - 'a' + 10;
                else if (h >= 'A' && h <= 'F') d = h - 'A' + 10;
                else return -1;
                cp = (cp << 4) | d;
            }
            if (cp >= 0xD800 && cp < 0xDC00 && c->end - c->p >= 6 && c->p[0] == '\\' && c->p[1] == 'u') {
                /* possible surrogate pair */
                uint32_t lo = 0; int ok = 1;
                for (int k = 0; k < 4; k++) {
                    uint8_t h = c->p[2 + k];
                    uint32_t d;
                    if (h >= '0' && h <= '9') d = h - '0';
                    else if (h >= 'a' && h <= 'f') d = h - 'a' + 10;
                    else if (h >= 'A' && h <= 'F') d = h - 'A' + 10;
                    else { ok = 0; d = 0; }
                    lo = (lo << 4) | d;
                }
                if (ok && lo >= 0xDC00 && lo < 0xE000) {
                    cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    c->p += 6;
                }
            }
            break;
        }
        default: return -1; /* backslash-any rejected */
        }
        /* UTF-8 encode; a lone surrogate is emitted as '?' (what Java's encoder does) */
        if (cp >= 0xD800 && cp < 0xE000) cp = '?';
        if (cp < 0x80) { if (out) out[n] = (uint8_t)cp; n += 1; }
        else if (cp < 0x800) {
            if (out) { out[n] = 0xC0 | (cp >> 6); out[n + 1] = 0x80 | (cp & 63); }
            n += 2;
        } else if (cp < 0x10000) {
            if (out) { out[n] = 0xE0 | (cp >> 12); out[n + 1] = 0x80 | ((cp >> 6) & 63); out[n + 2] = 0x80 | (cp & 63); }
            n += 3;
        } else {
            if (out) { out[n] = 0xF0 | (cp >> 18); out[n + 1] = 0x80 | ((cp >> 12) & 63);
                       out[n + 2] = 0x80 | ((cp >> 6) & 63); out[n + 3] = 0x80 | (cp & 63); }
            n += 4;
        }
    }
    *out_len = n;
    return 0;
}

/* Scans a JSON number token per the JSON grammar.  *is_float set when it has a fraction
 * or exponent.  Also accepts the non-numeric tokens Jackson admits under
 * ALLOW_NON_NUMERIC_NUMBERS (Spark default true): NaN, Infinity, -Infinity, INF, -INF. */
static int scan_number(cur_t *c, const uint8_t **beg, const uint8_t **endp, int *is_float, int *nonnum) {
    const uint8_t *s = c->p, *e = c->end;
    const uint8_t *p = s;
    *is_float = 0; *nonnum = 0;
    static const char *names[] = {"NaN", "Infinity", "-Infinity", "INF", "-INF"};
    for (int k = 0; k < 5; k++) {
        size_t L = strlen(names[k]);
        if ((size_t)(e - s) >= L && memcmp(s, names[k], L) == 0) {
            /* "-INF" must not shadow "-Infinity": names[] order handles it */
            *nonnum = k + 1; *is_float = 1; *beg = s; *endp = s + L; c->p = s + L; return 0;
        }
    }
    if (p < e && *p == '-') p++;
    if (p >= e) return -1;
    if (*p == '0') { p++; if (p < e && *p >= '0' && *p <= '9') return -1; /* leading zero */ }
    else if (*p >= '1' && *p <= '9') { while (p < e && *p >= '0' && *p <= '9') p++; }
    else return -1;
    if (p < e && *p == '.') {
        p++; *is_float = 1;
        if (p >= e || *p < '0' || *p > '9') return -1;
        while (p < e && *p >= '0' && *p <= '9') p++;
    }
    if (p < e && (*p == 'e' || *p == 'E')) {
        p++; *is_float = 1;
        if (p < e && (*p == '+' || *p == '-')) p++;
        if (p >= e || *p < '0' || *p > '9') return -1;
        while (p < e && *p >= '0' && *p <= '9') p++;
    }
    *beg = s; *endp = p; c->p = p;
    return 0;
}

static int match_lit(cur_t *c, const char *lit) {
    size_t L = strlen(lit);
    if ((size_t)(c->end - c->p) < L || memcmp(c->p, lit, L) != 0) return -1;
    c->p += L;
    return 0;
}

/* Skips (and validates) any JSON value: what Jackson's skipChildren() does for a field
 * that is not in the schema. */
static int skip_value(cur_t *c, int depth) {
    if (depth > 64) return -1;
    skip_ws(c);
    if (c->p >= c->end) return -1;
    uint8_t b = *c->p;
    if (b == '"' || b == '\'') { int64_t n; return parse_string(c, NULL, &n); }
    if (b == '{') {
        c->p++; skip_ws(c);
        if (c->p < c->end && *c->p == '}') { c->p++; return 0; }
        for (;;) {
            skip_ws(c);
            if (c->p >= c->end || (*c->p != '"' && *c->p != '\'')) return -1;
            int64_t n; if (parse_string(c, NULL, &n)) return -1;
            skip_ws(c);
            if (c->p >= c->end || *c->p != ':') return -1;
            c->p++;
            if (skip_value(c, depth + 1)) return -1;
            skip_ws(c);
            if (c->p >= c->end) return -1;
            if (*c->p == ',') { c->p++; continue; }
            if (*c->p == '}') { c->p++; return 0; }
            return -1;
        }
    }
    if (b == '[') {
        c->p++; skip_ws(c);
        if (c->p < c->end && *c->p == ']') { c->p++; return 0; }
        for (;;) {
            if (skip_value(c, depth + 1)) return -1;
            skip_ws(c);
            if (c->p >= c->end) return -1;
            if (*c->p == ',') { c->p++; continue; }
            if (*c->p == ']') { c->p++; return 0; }
            return -1;
        }
    }
    if (b == 't') return match_lit(c, "true");
    if (b == 'f') return match_lit(c, "false");
    if (b == 'n') return match_lit(c, "null");
    const uint8_t *s, *e; int isf, nn;
    return scan_number(c, &s, &e, &isf, &nn);
}

static void null_leaves(dxo_ctx *x, const dxo_struct *st) {
    for (int k = 0; k < st->n_leaves; k++) x->valid[st->first_leaf + k] = 0;
}

static int parse_object(dxo_ctx *x, cur_t *c, int sidx, int depth);

/* Converts the value at the cursor for a leaf field.  Returns 0 ok, -1 = BadRecord (the
 * whole row becomes NULL), following JacksonParser.makeConverter of Spark 2.4.5. */
static int parse_leaf(dxo_ctx *x, cur_t *c, const dxo_field *f) {
    int col = f->target;
    skip_ws(c);
    if (c->p >= c->end) return -1;
    uint8_t b = *c->p;
    if (b == 'n') { if (match_lit(c, "null")) return -1; x->valid[col] = 0; return 0; }
    if (b == '"' || b == '\'') {
        if (x->heap_used + (c->end - c->p) > x->heap_cap) return -2;
        int64_t n; uint8_t *dst = x->heap + x->heap_used;
        if (parse_string(c, dst, &n)) return -1;
        switch (f->type) {
        case DXO_STRING:
            x->vals[col] = x->heap_used; x->lens[col] = (int32_t)n; x->valid[col] = 1; x->heap_used += n; return 0;
        case DXO_LONG: case DXO_BOOL:
            if (n == 0) { x->valid[col] = 0; return 0; } /* failedConversion: "" => null field */
            return -1;
        case DXO_DOUBLE: {
            double d;
            if (n == 3 && memcmp(dst, "NaN", 3) == 0) d = NAN;
            else if (n == 8 && memcmp(dst, "Infinity", 8) == 0) d = INFINITY;
            else if (n == 9 && memcmp(dst, "-Infinity", 9) == 0) d = -INFINITY;
            else return -1; /* including "" : DoubleType's own VALUE_STRING case throws */
            memcpy(&x->vals[col], &d, 8); x->valid[col] = 1; return 0;
        }
        }
        return -1;
    }
    if (b == 't' || b == 'f') {
        int v = (b == 't');
        const uint8_t *s = c->p;
        if (match_lit(c, v ? "true" : "false")) return -1;
        if (f->type == DXO_BOOL) { x->vals[col] = v; x->valid[col] = 1; return 0; }
        if (f->type == DXO_STRING) {
            int64_t n = c->p - s;
            if (x->heap_used + n > x->heap_cap) return -2;
            memcpy(x->heap + x->heap_used, s, n);
            x->vals[col] = x->heap_used; x->lens[col] = (int32_t)n; x->valid[col] = 1; x->heap_used += n; return 0;
        }
        return -1;
    }
    if (b == '{' || b == '[') {
        /* string field: Spark re-serialises the structure; not restated */
        if (f->type == DXO_STRING) x->limited = 1;   /* Spark re-serialises the nested value; not restated: the row becomes NULL and is counted */
        return -1;
    }
    const uint8_t *s, *e; int isf, nn;
    if (scan_number(c, &s, &e, &isf, &nn)) return -1;
    switch (f->type) {
    case DXO_STRING: {
        int64_t n = e - s;
        if (x->heap_used + n > x->heap_cap) return -2;
        memcpy(x->heap + x->heap_used, s, n);
        x->vals[col] = x->heap_used; x->lens[col] = (int32_t)n; x->valid[col] = 1; x->heap_used += n; return 0;
    }
    case DXO_LONG: {
        if (isf) return -1;
        /* Jackson getLongValue: out of range => exception => row NULL */
        const uint8_t *p = s; int neg = 0;
        if (*p == '-') { neg = 1; p++; }
        uint64_t acc = 0;
        if (e - p > 19) return -1;
        for (; p < e; p++) {
            uint64_t d = *p - '0';
            if (acc > (UINT64_MAX - d) / 10) return -1;
            acc = acc * 10 + d;
        }
        if (!neg && acc > (uint64_t)INT64_MAX) return -1;
        if (neg && acc > (uint64_t)INT64_MAX + 1) return -1;
        x->vals[col] = neg ? (int64_t)(0 - acc) : (int64_t)acc; x->valid[col] = 1; return 0;
    }
    case DXO_DOUBLE: {
        double d;
        if (nn) d = (nn == 1) ? NAN : ((nn == 3 || nn == 5) ? -INFINITY : INFINITY);
        else {
            char buf[512]; size_t n = (size_t)(e - s);
            char *tmp = buf;
            if (n >= sizeof buf) tmp = (char *)malloc(n + 1);
            memcpy(tmp, s, n); tmp[n] = 0;
            d = strtod(tmp, NULL); /* glibc: correctly rounded, as Jackson's Double.parseDouble */
            if (tmp != buf) free(tmp);
            /* an integer token reaches getDoubleValue as an int/long first: "-0" => +0.0 */
            if (!isf && d == 0.0) d = 0.0;
        }
        memcpy(&x->vals[col], &d, 8); x->valid[col] = 1; return 0;
    }
    }
    return -1;
}

static int parse_object(dxo_ctx *x, cur_t *c, int sidx, int depth) {
    const dxo_struct *st = &x->structs[sidx];
    if (depth > 64) return -1;
    /* cursor is just past '{' */
    skip_ws(c);
    if (c->p < c->end && *c->p == '}') { c->p++; return 0; }
    uint8_t keybuf[256];
    for (;;) {
        skip_ws(c);
        if (c->p >= c->end || (*c->p != '"' && *c->p != '\'')) return -1;
        /* key: unescape into a bounded buffer; longer keys cannot match any field */
        cur_t k0 = *c; int64_t klen;
        if (parse_string(c, NULL, &klen)) return -1;
        const dxo_field *hit = NULL;
        if (klen <= (int64_t)sizeof keybuf) {
            parse_string(&k0, keybuf, &klen);
            for (int i = 0; i < st->n_fields; i++) {
                const dxo_field *f = &x->fields[st->first_field + i];
                if (f->name_len == klen && memcmp(f->name, keybuf, (size_t)klen) == 0) { hit = f; break; }
            }
        }
        skip_ws(c);
        if (c->p >= c->end || *c->p != ':') return -1;
        c->p++;
        if (!hit) { if (skip_value(c, depth + 1)) return -1; }
        else if (hit->type == DXO_STRUCT) {
            skip_ws(c);
            if (c->p >= c->end) return -1;
            const dxo_struct *child = &x->structs[hit->target];
            if (*c->p == 'n') { if (match_lit(c, "null")) return -1; null_leaves(x, child); }
            else if (*c->p == '{') {
                c->p++;
                null_leaves(x, child); /* duplicate key: last wins, start clean */
                int r = parse_object(x, c, hit->target, depth + 1);
                if (r) return r;
            } else return -1;
        } else {
            int r = parse_leaf(x, c, hit);
            if (r) return r;
        }
        skip_ws(c);
        if (c->p >= c->end) return -1;
        if (*c->p == ',') { c->p++; continue; }
        if (*c->p == '}') { c->p++; return 0; }
        return -1;
    }
}

/* One record.  Returns 1 = row ok, 0 = row NULL, <0 = oracle limitation hit. */
static int decode_record(dxo_ctx *x, const uint8_t *s, const uint8_t *e) {
    cur_t c = { s, e };
    int64_t heap0 = x->heap_used;
    for (int k = 0; k < x->n_leaves; k++) x->valid[k] = 0;
    skip_ws(&c);
    if (c.p >= c.end) return 0; /* empty / whitespace-only => NULL struct */
    int r = -1;
    if (*c.p == '{') { c.p++; r = parse_object(x, &c, 0, 0); }
    else if (*c.p == '[') {
        /* Spark 2.4 rootConverter: an array holding exactly one object => that row */
        c.p++; skip_ws(&c);
        if (c.p < c.end && *c.p == '{') {
            c.p++; r = parse_object(x, &c, 0, 1);
            if (r == 0) { skip_ws(&c); if (c.p < c.end && *c.p == ']') c.p++; else r = -1; }
        }
    }
    /* trailing bytes after the first value are never looked at */
    if (r == 0) return 1;
    x->heap_used = heap0;
    for (int k = 0; k < x->n_leaves; k++) x->valid[k] = 0;
    return (r == -1) ? 0 : r;
}

/*
 * dxo_decode: records i = json[offsets[i], offsets[i+1]) (offsets has n+1 entries), or
 * when ends != NULL, json[offsets[i], ends[i]).
 * Outputs are column-major: vals[leaf*n + row], lens[leaf*n + row], valid[leaf*n + row],
 * rowok[row]. Strings land in heap (caller allocates heap_cap >= total input bytes).
 * Returns number of rows decoded, or a negative limitation code.
 */
int64_t dxo_decode(const uint8_t *json, const int64_t *offsets, const int64_t *ends, int64_t n,
                   const dxo_struct *structs, const dxo_field *fields, int32_t n_leaves,
                   int64_t *vals, int32_t *lens, uint8_t *valid, uint8_t *rowok,
                   uint8_t *heap, int64_t heap_cap, int64_t *heap_used_out) {
    dxo_ctx x;
    x.structs = structs; x.fields = fields; x.n_leaves = n_leaves;
    x.vals = (int64_t *)calloc((size_t)n_leaves + 1, 8);
    x.lens = (int32_t *)calloc((size_t)n_leaves + 1, 4);
    x.valid = (uint8_t *)calloc((size_t)n_leaves + 1, 1);
    x.heap = heap; x.heap_used = 0; x.heap_cap = heap_cap;
    int64_t rc = n;
    for (int64_t i = 0; i < n; i++) {
        const uint8_t *s = json + offsets[i];
        const uint8_t *e = json + (ends ? ends[i] : offsets[i + 1]);
        x.limited = 0;
        int r = decode_record(&x, s, e);
        if (r < 0) { rc = r; break; }
        rowok[i] = (uint8_t)(r == 1 ? 1 : (x.limited ? 2 : 0));     /* 2: NULL row by a limit of the restatement */
        for (int k = 0; k < n_leaves; k++) {
            vals[(int64_t)k * n + i] = x.valid[k] ? x.vals[k] : 0;
            lens[(int64_t)k * n + i] = x.valid[k] ? x.lens[k] : 0;
            valid[(int64_t)k * n + i] = x.valid[k];
        }
    }
    if (heap_used_out) *heap_used_out = x.heap_used;
    free(x.vals); free(x.lens); free(x.valid);
    return rc;
}

/* Multi-threaded decode used only for the timed CPU baseline: each thread decodes a
 * contiguous slice of rows into the shared column arrays with a private string heap. */
typedef struct {
    const uint8_t *json; const int64_t *offsets; const int64_t *ends; int64_t lo, hi, n;
    const dxo_struct *structs; const dxo_field *fields; int32_t n_leaves;
    int64_t *vals; int32_t *lens; uint8_t *valid; uint8_t *rowok; int64_t rc;
    const uint8_t *leaf_is_string;   /* optional: string leaves report the FNV-1a hash of their bytes as the value */
} mt_arg;

static inline int64_t fnv1a64(const uint8_t *p, int64_t n) {
    uint64_t h = 0xcbf29ce484222325ull;
    for (int64_t i = 0; i < n; i++) { h ^= p[i]; h *= 0x100000001b3ull; }
    return (int64_t)h;
}

static void *mt_worker(void *p) {
    mt_arg *a = (mt_arg *)p;
    int64_t m = a->hi - a->lo;
    if (m <= 0) { a->rc = 0; return NULL; }
    int64_t bytes = (a->ends ? a->ends[a->hi - 1] : a->offsets[a->hi]) - a->offsets[a->lo];
    uint8_t *heap = (uint8_t *)malloc((size_t)bytes + 16);
    dxo_ctx x;
    x.structs = a->structs; x.fields = a->fields; x.n_leaves = a->n_leaves;
    x.vals = (int64_t *)calloc((size_t)a->n_leaves + 1, 8);
    x.lens = (int32_t *)calloc((size_t)a->n_leaves + 1, 4);
    x.valid = (uint8_t *)calloc((size_t)a->n_leaves + 1, 1);
    x.heap = heap; x.heap_cap = bytes + 16;
    a->rc = m;
    for (int64_t i = a->lo; i < a->hi; i++) {
        x.heap_used = 0; /* baseline keeps only the last string (dictionary coding is the product's job) */
        const uint8_t *s = a->json + a->offsets[i];
        const uint8_t *e = a->json + (a->ends ? a->ends[i] : a->offsets[i + 1]);
        x.limited = 0;
        int r = decode_record(&x, s, e);
        if (r < 0) { a->rc = r; break; }
        a->rowok[i] = (uint8_t)(r == 1 ? 1 : (x.limited ? 2 : 0));
        for (int k = 0; k < a->n_leaves; k++) {
            int64_t v = x.valid[k] ? x.vals[k] : 0;
            if (x.valid[k] && a->leaf_is_string && a->leaf_is_string[k]) v = fnv1a64(heap + x.vals[k], x.lens[k]);
            a->vals[(int64_t)k * a->n + i] = v;
            a->lens[(int64_t)k * a->n + i] = x.valid[k] ? x.lens[k] : 0;
            a->valid[(int64_t)k * a->n + i] = x.valid[k];
        }
    }
    free(x.vals); free(x.lens); free(x.valid); free(heap);
    return NULL;
}

/* String leaves come back as lens plus, when leaf_is_string marks them, the FNV-1a 64-bit hash of their bytes as the
 * value (a string identity the baseline's rule evaluation compares with the hashed literal; without the marks the
 * value is an offset within the record's private heap, meaningless across rows) — the baseline is for timing,
 * parity uses dxo_decode. */
int64_t dxo_decode_mt_h(const uint8_t *json, const int64_t *offsets, const int64_t *ends, int64_t n,
                        const dxo_struct *structs, const dxo_field *fields, int32_t n_leaves,
                        int64_t *vals, int32_t *lens, uint8_t *valid, uint8_t *rowok, int n_threads,
                        const uint8_t *leaf_is_string);
int64_t dxo_decode_mt(const uint8_t *json, const int64_t *offsets, const int64_t *ends, int64_t n,
                      const dxo_struct *structs, const dxo_field *fields, int32_t n_leaves,
                      int64_t *vals, int32_t *lens, uint8_t *valid, uint8_t *rowok, int n_threads) {
    return dxo_decode_mt_h(json, offsets, ends, n, structs, fields, n_leaves, vals, lens, valid, rowok, n_threads, NULL);
}
int64_t dxo_decode_mt_h(const uint8_t *json, const int64_t *offsets, const int64_t *ends, int64_t n,
                        const dxo_struct *structs, const dxo_field *fields, int32_t n_leaves,
                        int64_t *vals, int32_t *lens, uint8_t *valid, uint8_t *rowok, int n_threads,
                        const uint8_t *leaf_is_string) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256]; mt_arg args[256];
    for (int t = 0; t < n_threads; t++) {
        mt_arg *a = &args[t];
        a->json = json; a->offsets = offsets; a->ends = ends; a->n = n;
        a->lo = n * t / n_threads; a->hi = n * (t + 1) / n_threads;
        a->structs = structs; a->fields = fields; a->n_leaves = n_leaves;
        a->vals = vals; a->lens = lens; a->valid = valid; a->rowok = rowok; a->rc = 0;
        a->leaf_is_string = leaf_is_string;
        pthread_create(&th[t], NULL, mt_worker, a);
    }
    int64_t rc = n;
    for (int t = 0; t < n_threads; t++) { pthread_join(th[t], NULL); if (args[t].rc < 0) rc = args[t].rc; }
    return rc;
}

/* ------------------------------------------------------------------------------------ */
/* NDJSON framing: every '\n' ends a record; a final unterminated non-empty segment is   */
/* a record too.  starts/ends exclude the newline.  Returns the record count.            */
/* ------------------------------------------------------------------------------------ */
int64_t dxo_split(const uint8_t *bytes, int64_t n_bytes, int64_t *starts, int64_t *ends, int64_t cap) {
    int64_t n = 0, s = 0;
    for (int64_t i = 0; i < n_bytes; i++) {
        if (bytes[i] == '\n') {
            if (n < cap) { starts[n] = s; ends[n] = i; }
            n++; s = i + 1;
        }
    }
    if (s < n_bytes) { if (n < cap) { starts[n] = s; ends[n] = n_bytes; } n++; }
    return n;
}

/* ------------------------------------------------------------------------------------ */
/* Synthetic IoT events: event i is a pure function of (seed, i).         */
/* Same spec as the device generator in csrc/dx_gen.cu; written separately.              */
/* ------------------------------------------------------------------------------------ */
typedef struct {
    uint64_t seed;
    int64_t n_keys;        /* deviceId uniform in 1..n_keys */
    int64_t n_homes;       /* homeId = 1 + mix(deviceId) % n_homes */
    int64_t base_ms;       /* batch time (epoch ms) */
    int64_t interval_ms;
    int64_t watermark_ms;
    int32_t shape;         /* 1 = S1, 2 = S2 (adds temperature, humidity) */
    int32_t late_permille; /* events older than the watermark */
    int32_t future_permille;
    int32_t pad;
} dxo_gen_params;

static inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline uint64_t hk(const dxo_gen_params *g, uint64_t i, uint64_t k) { return mix64(g->seed ^ mix64(i * 16 + k)); }

static const char *const kTypes[4] = {"DoorLock", "WindowLock", "Heating", "GarageDoorLock"};

static int put_u64(char *dst, uint64_t v) {
    char tmp[24]; int n = 0;
    do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
    for (int k = 0; k < n; k++) dst[k] = tmp[n - 1 - k];
    return n;
}
static int put_i64(char *dst, int64_t v) {
    if (v < 0) { dst[0] = '-'; return 1 + put_u64(dst + 1, (uint64_t)0 - (uint64_t)v); }
    return put_u64(dst, (uint64_t)v);
}
static int put_str(char *dst, const char *s) { int n = (int)strlen(s); memcpy(dst, s, (size_t)n); return n; }

typedef struct { int64_t deviceId, homeId, eventTime, status; int32_t type; int32_t temp_c, hum_c; } dxo_event;

static void gen_fields(const dxo_gen_params *g, uint64_t i, dxo_event *ev) {
    ev->deviceId = 1 + (int64_t)(hk(g, i, 0) % (uint64_t)g->n_keys);
    ev->homeId = 1 + (int64_t)(mix64((uint64_t)ev->deviceId * 0x9E3779B97F4A7C15ull) % (uint64_t)g->n_homes);
    ev->type = (int32_t)(hk(g, i, 1) & 3);
    ev->status = (int64_t)(hk(g, i, 2) & 1);
    int64_t jitter = (int64_t)(hk(g, i, 3) % (uint64_t)g->interval_ms);
    int32_t cls = (int32_t)(hk(g, i, 4) % 1000);
    if (cls < g->late_permille) ev->eventTime = g->base_ms - g->watermark_ms - g->interval_ms - 1 - jitter;
    else if (cls >= 1000 - g->future_permille) ev->eventTime = g->base_ms + 1 + jitter;
    else ev->eventTime = g->base_ms - 1 - jitter;
    ev->temp_c = 510 + (int32_t)(hk(g, i, 5) % 9500);
    ev->hum_c = (int32_t)(hk(g, i, 6) % 10000);
}

static int gen_write(const dxo_gen_params *g, uint64_t i, char *d) {
    dxo_event ev; gen_fields(g, i, &ev);
    int n = 0;
    n += put_str(d + n, "{\"deviceDetails\":{\"deviceId\":");
    n += put_i64(d + n, ev.deviceId);
    n += put_str(d + n, ",\"deviceType\":\"");
    n += put_str(d + n, kTypes[ev.type]);
    n += put_str(d + n, "\",\"eventTime\":");
    n += put_i64(d + n, ev.eventTime);
    n += put_str(d + n, ",\"homeId\":");
    n += put_i64(d + n, ev.homeId);
    n += put_str(d + n, ",\"status\":");
    n += put_i64(d + n, ev.status);
    if (g->shape == 2) {
        n += put_str(d + n, ",\"temperature\":");
        n += put_u64(d + n, (uint64_t)(ev.temp_c / 100)); d[n++] = '.';
        d[n++] = (char)('0' + (ev.temp_c % 100) / 10); d[n++] = (char)('0' + ev.temp_c % 10);
        n += put_str(d + n, ",\"humidity\":");
        n += put_u64(d + n, (uint64_t)(ev.hum_c / 100)); d[n++] = '.';
        d[n++] = (char)('0' + (ev.hum_c % 100) / 10); d[n++] = (char)('0' + ev.hum_c % 10);
    }
    n += put_str(d + n, "}}\n");
    return n;
}

/* Writes events [first, first+n) as NDJSON into out (cap bytes); returns bytes written or -1. */
int64_t dxo_gen_events(const dxo_gen_params *g, int64_t first, int64_t n, uint8_t *out, int64_t cap) {
    int64_t used = 0; char tmp[320];
    for (int64_t i = 0; i < n; i++) {
        int L = gen_write(g, (uint64_t)(first + i), tmp);
        if (used + L > cap) return -1;
        memcpy(out + used, tmp, (size_t)L); used += L;
    }
    return used;
}

/* The parsed field values of the same events, for cross-checking decode against ground truth. */
void dxo_gen_fields(const dxo_gen_params *g, int64_t first, int64_t n, int64_t *deviceId, int32_t *type,
                    int64_t *eventTime, int64_t *homeId, int64_t *status, double *temp, double *hum) {
    for (int64_t i = 0; i < n; i++) {
        dxo_event ev; gen_fields(g, (uint64_t)(first + i), &ev);
        deviceId[i] = ev.deviceId; type[i] = ev.type; eventTime[i] = ev.eventTime;
        homeId[i] = ev.homeId; status[i] = ev.status;
        if (temp) temp[i] = (double)ev.temp_c / 100.0; /* == strtod("T.tt"): one correctly-rounded division */
        if (hum) hum[i] = (double)ev.hum_c / 100.0;
    }
}

/* ------------------------------------------------------------------------------------ */
/* Hash aggregate (Spark HashAggregateExec semantics) over up to 8 int64 key columns and */
/* one value column: COUNT(*), COUNT(v), SUM, MIN, MAX; AVG = sum/count on the caller.   */
/* Keys carry a validity byte (NULL keys form one group).  Value type: 1 long, 2 double. */
/* Output rows in first-seen order; returns the group count.                             */
/* ------------------------------------------------------------------------------------ */
typedef struct {
    int64_t count_star, count_v;
    int64_t sum_i; double sum_d;
    int64_t min_i, max_i; double min_d, max_d;
} dxo_agg;

static inline int dbl_lt(double a, double b) { /* Spark order: NaN greatest, -0.0 == 0.0 */
    if (isnan(a)) return 0;
    if (isnan(b)) return 1;
    return a < b;
}

int64_t dxo_group_agg(int64_t n, int32_t n_keys, const int64_t *const *keys, const uint8_t *const *key_valid,
                      int32_t vtype, const int64_t *vals, const uint8_t *vvalid, const uint8_t *row_sel,
                      int64_t *out_keys /* [n_keys][cap] */, uint8_t *out_key_valid, dxo_agg *out, int64_t cap) {
    int64_t tcap = 16; while (tcap < 2 * (cap < n ? cap : n) + 16) tcap <<= 1;
    int64_t *slot = (int64_t *)malloc((size_t)tcap * 8);
    for (int64_t i = 0; i < tcap; i++) slot[i] = -1;
    int64_t ng = 0;
    for (int64_t r = 0; r < n; r++) {
        if (row_sel && !row_sel[r]) continue;
        uint64_t h = 0x12345;
        for (int k = 0; k < n_keys; k++) {
            int v = key_valid ? key_valid[k][r] : 1;
            h = mix64(h ^ (v ? (uint64_t)keys[k][r] : 0xdeadbeefull) ^ ((uint64_t)v << 63));
        }
        int64_t p = (int64_t)(h & (uint64_t)(tcap - 1));
        int64_t g;
        for (;;) {
            g = slot[p];
            if (g < 0) break;
            int same = 1;
            for (int k = 0; k < n_keys && same; k++) {
                int v = key_valid ? key_valid[k][r] : 1;
                if (out_key_valid[(int64_t)k * cap + g] != v) same = 0;
                else if (v && out_keys[(int64_t)k * cap + g] != keys[k][r]) same = 0;
            }
            if (same) break;
            p = (p + 1) & (tcap - 1);
        }
        if (g < 0) {
            if (ng >= cap) { free(slot); return -1; }
            g = ng++; slot[p] = g;
            for (int k = 0; k < n_keys; k++) {
                int v = key_valid ? key_valid[k][r] : 1;
                out_key_valid[(int64_t)k * cap + g] = (uint8_t)v;
                out_keys[(int64_t)k * cap + g] = v ? keys[k][r] : 0;
            }
            memset(&out[g], 0, sizeof(dxo_agg));
        }
        dxo_agg *a = &out[g];
        a->count_star++;
        if (vals && (!vvalid || vvalid[r])) {
            if (vtype == 1) {
                int64_t v = vals[r];
                if (a->count_v == 0) { a->min_i = a->max_i = v; }
                else { if (v < a->min_i) a->min_i = v; if (v > a->max_i) a->max_i = v; }
                a->sum_i = (int64_t)((uint64_t)a->sum_i + (uint64_t)v); /* SUM(long) wraps */
                a->sum_d += (double)v;                                   /* AVG(long) sums as double */
            } else {
                double v; memcpy(&v, &vals[r], 8);
                if (a->count_v == 0) { a->min_d = a->max_d = v; }
                else { if (dbl_lt(v, a->min_d)) a->min_d = v; if (dbl_lt(a->max_d, v)) a->max_d = v; }
                a->sum_d += v;
            }
            a->count_v++;
        }
    }
    free(slot);
    return ng;
}
