// dx_decode.cuh — device-side JSON machinery for the K1 decode kernel: byte readers,
// Spark-2.4 `from_json` value conversion, Eisel-Lemire decimal->binary64, string dictionary.
//
// Stands in for from_json(Raw, rawBlockSchema) at CommonProcessorFactory.scala:93 (reference),
// whose arithmetic is Spark 2.4.5's JacksonParser; the exact rules restated here are listed in
// DESIGN.md §3 and mirrored (independently) by oracle/dxo_core.c.
#pragma once
#include "dx_common.h"

#define DX_MAX_LEAVES 31
#define DX_MAX_STRUCTS 16
#define DX_MAX_FIELDS 64
#define DX_NAMES_BYTES 1536
#define DX_MAX_SEGS 36
#define DX_LITS_BYTES 1536

// error/limitation bits accumulated in DecodeStats::flags (any non-zero bit fails the call)
#define DXF_DICT_OVERFLOW 1u
#define DXF_NUM_PRECISION 2u     // > 19 significant digits and the two bracketing roundings differ
#define DXF_STRUCT_AS_STRING 4u  // object/array token for a string field (Spark re-serialises it)
#define DXF_LONG_ESCAPED 8u      // escaped string longer than the unescape buffer
#define DXF_ROW_OVERFLOW 16u     // more records than the caller said
// Conditions that concern ONE record and that Spark would have parsed: the record becomes a NULL struct (as a malformed one
// does), is counted in DecodeStats::limited_rows and reported as a metric — one odd or hostile event must not fail the batch.
#define DXF_ROW_LIMITS (DXF_NUM_PRECISION | DXF_STRUCT_AS_STRING | DXF_LONG_ESCAPED)

struct DField { uint16_t name_off, name_len; int32_t type; int32_t target; };
struct DStruct { int32_t first_field, n_fields, first_leaf, n_leaves; };
// One template segment: a literal, then (type != 0) one leaf value.  All fields are ready-to-use 32-bit values so that the
// warp-uniform constant-bank reads in the fast path need no unpacking: the literal occupies words lit_word .. lit_word + n_words - 1
// of lits (n_words >= 1), the last of them compared under last_mask (all ones when the literal ends on a word boundary).
struct DSeg { uint32_t lit_word, n_words, lit_len, last_mask; int32_t type, leaf; };

struct DSchemaDev {
    DStruct structs[DX_MAX_STRUCTS];
    DField fields[DX_MAX_FIELDS];
    char names[DX_NAMES_BYTES];
    int32_t n_structs, n_fields, n_leaves;
    int32_t leaf_type[DX_MAX_LEAVES + 1];
    // fast-path template: the compact, schema-ordered serialisation of one event
    int32_t n_segs;
    DSeg segs[DX_MAX_SEGS];
    uint32_t lits[DX_LITS_BYTES / 4];   // literal bytes, each literal 4-byte aligned, zero padded
};

struct DecodeStats {
    unsigned long long rows, null_rows, fast_rows, flags;
    unsigned long long limited_rows;     // rows made NULL structs by a limit of this implementation (DXF_ROW_LIMITS), not by Spark's rules
};

struct DecodeArgs {
    const uint8_t* base_al;   // 16-byte aligned address at or below the first JSON byte
    uint32_t skew;            // first JSON byte = base_al + skew
    long long n_end;          // skew + n_bytes
    long long n_tiles;
    long long max_rows;
    const long long* offsets; // optional record offsets (n_rows+1), relative to the first JSON byte
    void* cols[DX_MAX_LEAVES + 1];
    void* mask;
    int32_t mask_width;       // 1 (<=7 leaves) or 4
    const DictDev* dicts;     // [n_leaves], used for string leaves
    const DSchemaDev* schema;
    unsigned long long* tile_state;
    unsigned int* tile_counter;
    DecodeStats* stats;
    // the record template (copy of DSchemaDev's): kernel parameters sit in the constant bank, so the segment
    // descriptors and literal words are read through the uniform data path instead of per-lane shared loads
    int32_t n_segs, pad_;
    DSeg segs[DX_MAX_SEGS];
    uint32_t lits[DX_LITS_BYTES / 4];
};
static_assert(sizeof(DecodeArgs) <= 4096, "kernel parameter block");

// ---------------------------------------------------------------------------------------
static __device__ __forceinline__ uint64_t dx_mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// ------------------------------- dictionary --------------------------------------------
static __device__ __forceinline__ bool dict_slot_match(const DictDev& d, uint64_t sw0, uint64_t sw1, uint32_t slen,
                                                uint32_t soff, uint64_t w0, uint64_t w1, uint32_t len,
                                                const uint8_t* bytes) {
    if (slen != len || sw0 != w0 || sw1 != w1) return false;
    if (len <= 16) return true;
    for (uint32_t k = 0; k < len; k++)
        if (__ldcg(d.heap + soff + k) != bytes[k]) return false;
    return true;
}

// bytes: generic pointer to the (unescaped) string; for len<=16 (w0,w1) are the packed bytes,
// otherwise w0 is a 64-bit hash of the bytes and w1 = 0.
static __device__ __noinline__ int32_t dict_get_or_insert(const DictDev& d, uint64_t w0, uint64_t w1, uint32_t len,
                                      const uint8_t* bytes, unsigned long long* err_flags) {
    uint64_t h = dx_mix64(w0 ^ dx_mix64(w1 + len));
    uint32_t i = (uint32_t)h & d.cap_mask;
    for (uint32_t probe = 0; probe <= d.cap_mask; probe++, i = (i + 1) & d.cap_mask) {
        DictSlot* s = d.slots + i;
        // cached read first: a READY slot never changes again, so an L1 hit is always valid
        const uint4 a = *reinterpret_cast<const uint4*>(s);
        const uint4 b = *(reinterpret_cast<const uint4*>(s) + 1);
        if (b.w == 2u) {
            uint64_t sw0 = ((uint64_t)a.y << 32) | a.x, sw1 = ((uint64_t)a.w << 32) | a.z;
            if (dict_slot_match(d, sw0, sw1, b.x, b.y, w0, w1, len, bytes)) return (int32_t)b.z;
            continue;
        }
        for (;;) {
            uint32_t st = *reinterpret_cast<volatile uint32_t*>(&s->state);
            if (st == 2u) {
                const uint4 a2 = __ldcg(reinterpret_cast<const uint4*>(s));
                const uint4 b2 = __ldcg(reinterpret_cast<const uint4*>(s) + 1);
                uint64_t sw0 = ((uint64_t)a2.y << 32) | a2.x, sw1 = ((uint64_t)a2.w << 32) | a2.z;
                if (dict_slot_match(d, sw0, sw1, b2.x, b2.y, w0, w1, len, bytes)) return (int32_t)b2.z;
                break; // occupied by another string: keep probing
            }
            if (st == 0u && atomicCAS(&s->state, 0u, 1u) == 0u) {
                uint32_t code = atomicAdd(&d.counters[0], 1u);
                uint32_t off = atomicAdd(&d.counters[1], len);
                bool ok = code < d.max_codes && (unsigned long long)off + len <= d.heap_cap;
                if (ok) {
                    for (uint32_t k = 0; k < len; k++) d.heap[off + k] = bytes[k];
                    d.code_off[code] = off; d.code_len[code] = len;
                } else {
                    atomicOr(err_flags, (unsigned long long)DXF_DICT_OVERFLOW);
                }
                s->w0 = w0; s->w1 = w1; s->len = len; s->heap_off = off; s->code = ok ? (int32_t)code : -1;
                __threadfence();
                *reinterpret_cast<volatile uint32_t*>(&s->state) = 2u;
                return ok ? (int32_t)code : -1;
            }
            // st == 1: another thread is filling this slot; wait for it
        }
    }
    atomicOr(err_flags, (unsigned long long)DXF_DICT_OVERFLOW);
    return -1;
}

static __device__ __forceinline__ void dict_pack(const uint8_t* bytes, uint32_t len, uint64_t& w0, uint64_t& w1) {
    w0 = 0; w1 = 0;
    if (len <= 16) {
        for (uint32_t k = 0; k < len; k++) {
            uint64_t b = bytes[k];
            if (k < 8) w0 |= b << (8 * k); else w1 |= b << (8 * (k - 8));
        }
    } else {
        uint64_t h = 0xcbf29ce484222325ull;
        for (uint32_t k = 0; k < len; k++) { h ^= bytes[k]; h *= 0x100000001b3ull; }
        w0 = h;
    }
}

#include "dx_numconv.cuh"

// ------------------------------- general (any-JSON) parser ------------------------------
struct GCur { const uint8_t* p; const uint8_t* end; };

static __device__ __forceinline__ void g_ws(GCur& c) {
    while (c.p < c.end) { uint8_t b = *c.p; if (b == ' ' || b == '\t' || b == '\n' || b == '\r') c.p++; else break; }
}

static __device__ __forceinline__ int g_hex(uint8_t h) {
    if (h >= '0' && h <= '9') return h - '0';
    if (h >= 'a' && h <= 'f') return h - 'a' + 10;
    if (h >= 'A' && h <= 'F') return h - 'A' + 10;
    return -1;
}

// String token at the opening quote.  Unescaped bytes go to out[0..cap) when out != nullptr
// (bytes beyond cap are counted but not stored).  *has_esc tells the caller whether the raw
// bytes between the quotes equal the unescaped text.  Returns false on a lexical error.
static __device__ bool g_string(GCur& c, uint8_t* out, uint32_t cap, uint32_t* out_len, bool* has_esc) {
    uint8_t q = *c.p++;
    uint32_t n = 0; bool esc = false;
    for (;;) {
        if (c.p >= c.end) return false;
        uint8_t b = *c.p++;
        if (b == q) break;
        if (b < 0x20) return false;
        if (b != '\\') { if (out && n < cap) out[n] = b; n++; continue; }
        esc = true;
        if (c.p >= c.end) return false;
        uint8_t e = *c.p++;
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
        case '\'': if (q == '\'') { cp = '\''; break; } return false;
        case 'u': {
            if (c.end - c.p < 4) return false;
            cp = 0;
            for (int k = 0; k < 4; k++) { int d = g_hex(*c.p++); if (d < 0) return false; cp = (cp << 4) | (uint32_t)d; }
            if (cp >= 0xD800 && cp < 0xDC00 && c.end - c.p >= 6 && c.p[0] == '\\' && c.p[1] == 'u') {
                uint32_t lo = 0; bool ok = true;
                for (int k = 0; k < 4; k++) { int d = g_hex(c.p[2 + k]); if (d < 0) { ok = false; d = 0; } lo = (lo << 4) | (uint32_t)d; }
                if (ok && lo >= 0xDC00 && lo < 0xE000) { cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00); c.p += 6; }
            }
            break;
        }
        default: return false;
        }
        if (cp >= 0xD800 && cp < 0xE000) cp = '?';
        uint8_t u[4]; uint32_t m;
        if (cp < 0x80) { u[0] = (uint8_t)cp; m = 1; }
        else if (cp < 0x800) { u[0] = 0xC0 | (cp >> 6); u[1] = 0x80 | (cp & 63); m = 2; }
        else if (cp < 0x10000) { u[0] = 0xE0 | (cp >> 12); u[1] = 0x80 | ((cp >> 6) & 63); u[2] = 0x80 | (cp & 63); m = 3; }
        else { u[0] = 0xF0 | (cp >> 18); u[1] = 0x80 | ((cp >> 12) & 63); u[2] = 0x80 | ((cp >> 6) & 63); u[3] = 0x80 | (cp & 63); m = 4; }
        for (uint32_t k = 0; k < m; k++) { if (out && n < cap) out[n] = u[k]; n++; }
    }
    *out_len = n; *has_esc = esc;
    return true;
}

static __device__ __forceinline__ bool g_lit(GCur& c, const char* lit, int L) {
    if (c.end - c.p < L) return false;
    for (int k = 0; k < L; k++) if (c.p[k] != (uint8_t)lit[k]) return false;
    c.p += L;
    return true;
}

// Number token per the JSON grammar, plus NaN / Infinity / -Infinity / INF / -INF
// (Jackson ALLOW_NON_NUMERIC_NUMBERS, on by default in Spark).  nonnum: 0 none, 1 NaN, 2 +inf, 3 -inf.
static __device__ bool g_number(GCur& c, const uint8_t** beg, const uint8_t** endp, bool* is_float, int* nonnum) {
    const uint8_t* s = c.p; const uint8_t* e = c.end; const uint8_t* p = s;
    *is_float = false; *nonnum = 0;
    if (g_lit(c, "NaN", 3)) { *nonnum = 1; *is_float = true; *beg = s; *endp = c.p; return true; }
    if (g_lit(c, "Infinity", 8)) { *nonnum = 2; *is_float = true; *beg = s; *endp = c.p; return true; }
    if (g_lit(c, "-Infinity", 9)) { *nonnum = 3; *is_float = true; *beg = s; *endp = c.p; return true; }
    if (g_lit(c, "INF", 3)) { *nonnum = 2; *is_float = true; *beg = s; *endp = c.p; return true; }
    if (g_lit(c, "-INF", 4)) { *nonnum = 3; *is_float = true; *beg = s; *endp = c.p; return true; }
    if (p < e && *p == '-') p++;
    if (p >= e) return false;
    if (*p == '0') { p++; if (p < e && *p >= '0' && *p <= '9') return false; }
    else if (*p >= '1' && *p <= '9') { while (p < e && *p >= '0' && *p <= '9') p++; }
    else return false;
    if (p < e && *p == '.') {
        p++; *is_float = true;
        if (p >= e || *p < '0' || *p > '9') return false;
        while (p < e && *p >= '0' && *p <= '9') p++;
    }
    if (p < e && (*p == 'e' || *p == 'E')) {
        p++; *is_float = true;
        if (p < e && (*p == '+' || *p == '-')) p++;
        if (p >= e || *p < '0' || *p > '9') return false;
        while (p < e && *p >= '0' && *p <= '9') p++;
    }
    *beg = s; *endp = p; c.p = p;
    return true;
}

// Validating skip of one value (Jackson skipChildren); iterative with a 64-deep bit stack.
static __device__ bool g_skip_value(GCur& c, int base_depth) {
    uint64_t kinds = 0; // bit d: 1 = object, 0 = array
    int depth = 0;
    bool need_value = true;
    for (;;) {
        if (need_value) {
            g_ws(c);
            if (c.p >= c.end) return false;
            uint8_t b = *c.p;
            if (b == '"' || b == '\'') { uint32_t n; bool esc; if (!g_string(c, nullptr, 0, &n, &esc)) return false; }
            else if (b == '{') {
                if (depth + base_depth >= 64) return false;
                c.p++; g_ws(c);
                if (c.p < c.end && *c.p == '}') { c.p++; }
                else {
                    kinds |= (1ull << depth); depth++;
                    // key
                    if (c.p >= c.end || (*c.p != '"' && *c.p != '\'')) return false;
                    uint32_t n; bool esc; if (!g_string(c, nullptr, 0, &n, &esc)) return false;
                    g_ws(c); if (c.p >= c.end || *c.p != ':') return false; c.p++;
                    continue;
                }
            } else if (b == '[') {
                if (depth + base_depth >= 64) return false;
                c.p++; g_ws(c);
                if (c.p < c.end && *c.p == ']') { c.p++; }
                else { kinds &= ~(1ull << depth); depth++; continue; }
            } else if (b == 't') { if (!g_lit(c, "true", 4)) return false; }
            else if (b == 'f') { if (!g_lit(c, "false", 5)) return false; }
            else if (b == 'n') { if (!g_lit(c, "null", 4)) return false; }
            else { const uint8_t *s, *e; bool f; int nn; if (!g_number(c, &s, &e, &f, &nn)) return false; }
        }
        // a value just ended
        if (depth == 0) return true;
        g_ws(c);
        if (c.p >= c.end) return false;
        bool is_obj = (kinds >> (depth - 1)) & 1;
        uint8_t b = *c.p++;
        if (b == ',') {
            if (is_obj) {
                g_ws(c);
                if (c.p >= c.end || (*c.p != '"' && *c.p != '\'')) return false;
                uint32_t n; bool esc; if (!g_string(c, nullptr, 0, &n, &esc)) return false;
                g_ws(c); if (c.p >= c.end || *c.p != ':') return false; c.p++;
            }
            need_value = true;
        } else if (b == (is_obj ? '}' : ']')) {
            depth--; need_value = false;
        } else return false;
    }
}

#define DX_UNESC_CAP 256

struct RowSink {
    const DecodeArgs* a;
    long long row;
    uint32_t valid;      // bit k: leaf k non-null
    uint32_t err;        // DXF_* bits
};

static __device__ __forceinline__ void sink_i64(RowSink& o, int leaf, long long v) {
    reinterpret_cast<long long*>(o.a->cols[leaf])[o.row] = v; o.valid |= 1u << leaf;
}
static __device__ __forceinline__ void sink_f64(RowSink& o, int leaf, double v) {
    reinterpret_cast<double*>(o.a->cols[leaf])[o.row] = v; o.valid |= 1u << leaf;
}
static __device__ __forceinline__ void sink_code(RowSink& o, int leaf, int32_t v) {
    reinterpret_cast<int32_t*>(o.a->cols[leaf])[o.row] = v; o.valid |= 1u << leaf;
}
static __device__ void sink_string(RowSink& o, int leaf, const uint8_t* bytes, uint32_t len) {
    uint64_t w0, w1; dict_pack(bytes, len, w0, w1);
    int32_t code = dict_get_or_insert(o.a->dicts[leaf], w0, w1, len, bytes, &o.a->stats->flags);
    sink_code(o, leaf, code);
}

// Leaf conversion (JacksonParser.makeConverter, Spark 2.4.5).  Returns 0 ok, 1 bad record.
static __device__ int g_leaf(GCur& c, const DField& f, RowSink& o) {
    const int leaf = f.target;
    g_ws(c);
    if (c.p >= c.end) return 1;
    uint8_t b = *c.p;
    if (b == 'n') { if (!g_lit(c, "null", 4)) return 1; o.valid &= ~(1u << leaf); return 0; }
    if (b == '"' || b == '\'') {
        const uint8_t* raw = c.p + 1;
        uint8_t buf[DX_UNESC_CAP];
        uint32_t n; bool esc;
        if (!g_string(c, buf, DX_UNESC_CAP, &n, &esc)) return 1;
        switch (f.type) {
        case DX_T_STR:
            if (!esc) sink_string(o, leaf, raw, n);
            else if (n <= DX_UNESC_CAP) sink_string(o, leaf, buf, n);
            else { o.err |= DXF_LONG_ESCAPED; return 1; }
            return 0;
        case DX_T_I64: case DX_T_BOOL:
            if (n == 0) { o.valid &= ~(1u << leaf); return 0; }
            return 1;
        case DX_T_F64: {
            const uint8_t* t = esc ? buf : raw;
            if (esc && n > DX_UNESC_CAP) return 1;
            double d;
            if (n == 3 && t[0] == 'N' && t[1] == 'a' && t[2] == 'N') d = __longlong_as_double(0x7FF8000000000000ll);
            else if (n == 8 && t[0] == 'I' && t[1] == 'n' && t[2] == 'f' && t[3] == 'i' && t[4] == 'n' && t[5] == 'i' && t[6] == 't' && t[7] == 'y')
                d = __longlong_as_double(0x7FF0000000000000ll);
            else if (n == 9 && t[0] == '-' && t[1] == 'I' && t[2] == 'n' && t[3] == 'f' && t[4] == 'i' && t[5] == 'n' && t[6] == 'i' && t[7] == 't' && t[8] == 'y')
                d = __longlong_as_double((long long)0xFFF0000000000000ull);
            else return 1;
            sink_f64(o, leaf, d); return 0;
        }
        }
        return 1;
    }
    if (b == 't' || b == 'f') {
        const uint8_t* s = c.p;
        bool v = (b == 't');
        if (!(v ? g_lit(c, "true", 4) : g_lit(c, "false", 5))) return 1;
        if (f.type == DX_T_BOOL) { sink_i64(o, leaf, v ? 1 : 0); return 0; }
        if (f.type == DX_T_STR) { sink_string(o, leaf, s, (uint32_t)(c.p - s)); return 0; }
        return 1;
    }
    if (b == '{' || b == '[') {
        if (f.type == DX_T_STR) o.err |= DXF_STRUCT_AS_STRING;
        return 1;
    }
    const uint8_t *s, *e; bool isf; int nn;
    if (!g_number(c, &s, &e, &isf, &nn)) return 1;
    switch (f.type) {
    case DX_T_STR: sink_string(o, leaf, s, (uint32_t)(e - s)); return 0;
    case DX_T_I64: {
        if (isf) return 1;
        const uint8_t* p = s; bool neg = false;
        if (*p == '-') { neg = true; p++; }
        if (e - p > 19) return 1;
        uint64_t acc = 0;
        for (; p < e; p++) {
            uint64_t d = *p - '0';
            if (acc > (0xFFFFFFFFFFFFFFFFull - d) / 10) return 1;
            acc = acc * 10 + d;
        }
        if (!neg && acc > 0x7FFFFFFFFFFFFFFFull) return 1;
        if (neg && acc > 0x8000000000000000ull) return 1;
        sink_i64(o, leaf, neg ? (long long)(0 - acc) : (long long)acc); return 0;
    }
    case DX_T_F64: {
        double d;
        if (nn == 1) d = __longlong_as_double(0x7FF8000000000000ll);
        else if (nn == 2) d = __longlong_as_double(0x7FF0000000000000ll);
        else if (nn == 3) d = __longlong_as_double((long long)0xFFF0000000000000ull);
        else if (!dx_number_to_double(s, e, !isf, &d)) { o.err |= DXF_NUM_PRECISION; return 1; }
        sink_f64(o, leaf, d); return 0;
    }
    }
    return 1;
}

static __device__ __forceinline__ uint32_t leaf_range_mask(const DStruct& st) {
    return (st.n_leaves >= 32 ? 0xFFFFFFFFu : ((1u << st.n_leaves) - 1u)) << st.first_leaf;
}

// Schema-directed object parse; cursor just past '{' of the root struct.  0 ok, 1 bad record.
static __device__ int g_object(GCur& c, const DSchemaDev* S, RowSink& o, int base_depth) {
    int sstack[8]; int sp = 0; int cur = 0;
    bool at_start = true;
    for (;;) {
        if (at_start) {
            g_ws(c);
            if (c.p < c.end && *c.p == '}') {
                c.p++;
                if (sp == 0) return 0;
                cur = sstack[--sp]; at_start = false;
                goto after_value;
            }
        }
        {
            g_ws(c);
            if (c.p >= c.end || (*c.p != '"' && *c.p != '\'')) return 1;
            const uint8_t* kraw = c.p + 1;
            uint8_t kbuf[64]; uint32_t klen; bool kesc;
            if (!g_string(c, kbuf, 64, &klen, &kesc)) return 1;
            const uint8_t* key = kesc ? kbuf : kraw;
            int hit = -1;
            if (!kesc || klen <= 64) {
                const DStruct& st = S->structs[cur];
                for (int i = 0; i < st.n_fields; i++) {
                    const DField& f = S->fields[st.first_field + i];
                    if (f.name_len != klen) continue;
                    bool eq = true;
                    for (uint32_t k = 0; k < klen; k++) if ((uint8_t)S->names[f.name_off + k] != key[k]) { eq = false; break; }
                    if (eq) { hit = st.first_field + i; break; }
                }
            }
            g_ws(c);
            if (c.p >= c.end || *c.p != ':') return 1;
            c.p++;
            if (hit < 0) { if (!g_skip_value(c, base_depth + sp + 1)) return 1; }
            else {
                const DField& f = S->fields[hit];
                if (f.type == 100) { // nested struct
                    g_ws(c);
                    if (c.p >= c.end) return 1;
                    const DStruct& child = S->structs[f.target];
                    if (*c.p == 'n') { if (!g_lit(c, "null", 4)) return 1; o.valid &= ~leaf_range_mask(child); }
                    else if (*c.p == '{') {
                        c.p++;
                        o.valid &= ~leaf_range_mask(child);
                        if (sp >= 8) return 1;
                        sstack[sp++] = cur; cur = f.target; at_start = true;
                        continue;
                    } else return 1;
                } else {
                    if (g_leaf(c, f, o)) return 1;
                }
            }
        }
    after_value:
        at_start = false;
        g_ws(c);
        if (c.p >= c.end) return 1;
        uint8_t b = *c.p++;
        if (b == ',') continue;
        if (b == '}') {
            if (sp == 0) return 0;
            cur = sstack[--sp];
            goto after_value;
        }
        return 1;
    }
}

// One record through the general path.  Returns true when the row is non-NULL.
static __device__ __noinline__ bool g_record(const uint8_t* s, const uint8_t* e, const DSchemaDev* S, RowSink& o) {
    GCur c{s, e};
    o.valid = 0;
    g_ws(c);
    if (c.p >= c.end) return false;
    int r = 1;
    if (*c.p == '{') { c.p++; r = g_object(c, S, o, 0); }
    else if (*c.p == '[') {
        c.p++; g_ws(c);
        if (c.p < c.end && *c.p == '{') {
            c.p++; r = g_object(c, S, o, 1);
            if (r == 0) { g_ws(c); if (c.p < c.end && *c.p == ']') c.p++; else r = 1; }
        }
    }
    if (r != 0) { o.valid = 0; return false; }
    return true;
}
