// dx_tojson.cu — K6: the rows of a result table serialised to NDJSON on the device (the sink hand-off).
//
// Reference: OutputManager.sink (datax-host/src/main/scala/datax/sink/OutputManager.scala:103-118) builds
// to_json(struct(<columns not starting with __DataX_>)) per row and hands the strings to the JSON sinks
// (SinkerUtil.outputAllEvents / outputFilteredEvents, datax-utility/.../SinkerUtil.scala:27-52: rows.map(_.getString(1))).
// to_json is Spark 2.4.5's JacksonGenerator; its rules as restated here:
//   * struct: fields in order, NULL fields skipped; map: every key, a NULL value as `null`;
//   * long -> digits, boolean -> true/false, double -> java.lang.Double.toString (NaN / Infinity / -Infinity quoted),
//     string and keys -> Jackson's default escapes (\" \\ \b \t \n \f \r, other bytes < 0x20 as \u00XX, upper-case hex),
//     timestamp -> yyyy-MM-dd'T'HH:mm:ss.SSSXXX in UTC;
//   * the values of one MAP(...) take Spark's common type first (any string -> all strings, long + double -> double);
//   * a rule-flag column is the array (or, for a lone IF(cond, MAP(...), NULL), the nullable map) of the fired rules' constant objects.
// Double.toString prints the shortest decimal that round-trips; the digits come from the Schubfach algorithm (R. Giulietti,
// "The Schubfach way to render doubles", the algorithm of JDK 19+; JDK 8 differs on a small known set, JDK-4511638: see
// NEEDS_SPARK_CONFIRMATION.md).
//
// Device form.  The host flattens the column tree into a short program (JOp list: open object, field, rule array, close object)
// with every key pre-escaped into a literal heap.  Pass 1 (k6_json_len) runs the program per row against a counting sink;
// an exclusive scan turns lengths into byte offsets; pass 2 (k6_json_write) runs it again against a writing sink.  One thread
// owns one row; the 128 rows of a CTA are adjacent in the output, so they are written into a shared-memory image of the CTA's
// output span laid out congruent to the global address mod 16, and the span leaves as ONE bulk (TMA) store plus < 16 ragged
// bytes on either side.  Spans that do not fit the stage (very long rows) are written by their threads directly.
// HBM traffic per row: the referenced column bytes + 4 B length + 8 B offset + the JSON bytes once.
#include "dx_common.h"
#include "dx_flow.h"
#include "dx_tojson.h"
#include "dx_pow10_table.h"
#include <cub/cub.cuh>

enum : uint8_t { JOP_OPEN = 1, JOP_CLOSE = 2, JOP_FIELD = 3, JOP_RULES = 4, JOP_LIT = 5 };
enum : uint8_t { JN_DECODED = 0, JN_STRUCT = 1, JN_MAP = 2 };
enum : uint8_t { JC_NONE = 0, JC_STRING = 1, JC_DOUBLE = 2 };

struct JOp {
    uint8_t op, type, cast, kind;
    uint8_t null_as_null, single, is_root, stride;   // stride 0: a constant column held as one element
    int32_t lit_off, lit_len;                        // key literal `"key":` (JOP_LIT: the whole text)
    const void* data; const void* vmask; int32_t vwidth, vbit;
    const uint8_t* heap; const uint32_t* code_off; const uint32_t* code_len;      // STR: the dictionary's text
    int32_t flag_words, n_rules; const int32_t* frag;                              // FLAGS: frag[2r], frag[2r+1] = literal-heap span of rule r's object
};

#define K6_NT 128
#define K6_STAGE (64 * 1024 - 128)
#define K6_MAX_DEPTH 8

// ---------------------------------------------------------------------------------------------------------------------
// sinks
// ---------------------------------------------------------------------------------------------------------------------
struct CountSink {
    uint32_t p = 0;
    __device__ __forceinline__ void put(uint8_t) { p++; }
    __device__ __forceinline__ void lit(const uint8_t*, int n) { p += (uint32_t)n; }
    __device__ __forceinline__ void set_last(uint8_t) {}
    static constexpr bool counting = true;
};
struct WriteSink {
    uint8_t* out; uint32_t p = 0;
    __device__ __forceinline__ void put(uint8_t c) { out[p++] = c; }
    __device__ __forceinline__ void lit(const uint8_t* s, int n) { for (int i = 0; i < n; i++) out[p + i] = s[i]; p += (uint32_t)n; }
    __device__ __forceinline__ void set_last(uint8_t c) { out[p - 1] = c; }
    static constexpr bool counting = false;
};

// ---------------------------------------------------------------------------------------------------------------------
// value formatting
// ---------------------------------------------------------------------------------------------------------------------
// decimal digits of v < 2^32 without a division loop
static __device__ __forceinline__ int j_ndigits32(uint32_t v) {
    return v < 10u ? 1 : v < 100u ? 2 : v < 1000u ? 3 : v < 10000u ? 4 : v < 100000u ? 5 : v < 1000000u ? 6 : v < 10000000u ? 7 : v < 100000000u ? 8 : v < 1000000000u ? 9 : 10;
}
template <class S> static __device__ __forceinline__ void emit_u32(S& s, uint32_t v) {
    const int n = j_ndigits32(v);
    if (S::counting) { s.p += (uint32_t)n; return; }
    char buf[10];
    for (int i = n - 1; i >= 0; i--) { buf[i] = (char)('0' + v % 10u); v /= 10u; }
    for (int i = 0; i < n; i++) s.put((uint8_t)buf[i]);
}
template <class S> static __device__ __forceinline__ void emit_u32_padded(S& s, uint32_t v, int width) {      // exactly `width` digits
    if (S::counting) { s.p += (uint32_t)width; return; }
    char buf[10];
    for (int i = width - 1; i >= 0; i--) { buf[i] = (char)('0' + v % 10u); v /= 10u; }
    for (int i = 0; i < width; i++) s.put((uint8_t)buf[i]);
}
// 64-bit values go out as up to three 32-bit pieces (32-bit arithmetic is a third of the instructions of 64-bit)
template <class S> static __device__ void emit_u64(S& s, unsigned long long v) {
    if (v < 1000000000ull) { emit_u32(s, (uint32_t)v); return; }
    const unsigned long long hi = v / 1000000000ull; const uint32_t lo = (uint32_t)(v - hi * 1000000000ull);
    if (hi < 1000000000ull) emit_u32(s, (uint32_t)hi);
    else { const unsigned long long top = hi / 1000000000ull; emit_u32(s, (uint32_t)top); emit_u32_padded(s, (uint32_t)(hi - top * 1000000000ull), 9); }
    emit_u32_padded(s, lo, 9);
}
template <class S> static __device__ void emit_i64(S& s, long long v) {
    if (v < 0) { s.put('-'); emit_u64(s, 0ull - (unsigned long long)v); } else emit_u64(s, (unsigned long long)v);
}
template <class S> static __device__ __forceinline__ void emit_2(S& s, int v) { s.put((uint8_t)('0' + v / 10)); s.put((uint8_t)('0' + v % 10)); }

// the decimal Double.toString prints for a finite, non-zero double: value = digits * 10^exp10 — the shortest that round-trips, of two
// digits at least, the nearest to the value among those (Schubfach, as in the JDK's DoubleToDecimal)
static __device__ __forceinline__ unsigned long long j_round_to_odd(unsigned long long ghi, unsigned long long glo, unsigned long long cp) {
    const unsigned long long xhi = __umul64hi(glo, cp);
    const unsigned long long ylo = ghi * cp, yhi = __umul64hi(ghi, cp);
    const unsigned long long y0 = ylo + xhi;
    const unsigned long long y1 = yhi + (y0 < ylo ? 1ull : 0ull);
    return y1 | (y0 > 1ull ? 1ull : 0ull);
}
static __device__ void j_shortest(unsigned long long bits, unsigned long long* digits, int* exp10) {
    const unsigned long long F = bits & ((1ull << 52) - 1ull);
    const int E = (int)((bits >> 52) & 0x7ffull);
    unsigned long long c; int q, dk = 0;
    if (E == 0) { c = F; q = 1 - 1075; if (c < 3ull) { c *= 10ull; dk = -1; } }     // the two smallest subnormals get their second digit this way (4.9E-324, 9.9E-324)
    else { c = F | (1ull << 52); q = E - 1075; }
    const bool is_even = (c & 1ull) == 0ull;
    const bool lower_closer = (F == 0ull && E > 1);
    const unsigned long long cbl = 4ull * c - 2ull + (lower_closer ? 1ull : 0ull), cb = 4ull * c, cbr = 4ull * c + 2ull;
    const int k = lower_closer ? (int)(((long long)q * 1262611ll - 524031ll) >> 22) : (int)(((long long)q * 1262611ll) >> 22);
    const int h = q + (int)(((long long)(-k) * 1741647ll) >> 19) + 1;
    const unsigned long long ghi = dx_pow10_128[-k - DX_POW10_KMIN][0], glo = dx_pow10_128[-k - DX_POW10_KMIN][1];
    const unsigned long long vbl = j_round_to_odd(ghi, glo, cbl << h), vb = j_round_to_odd(ghi, glo, cb << h), vbr = j_round_to_odd(ghi, glo, cbr << h);
    const unsigned long long lower = vbl + (is_even ? 0ull : 1ull), upper = vbr - (is_even ? 0ull : 1ull);
    const unsigned long long s = vb >> 2;
    if (s >= 100ull) {      // Double.toString wants two digits at least: a one-digit candidate is never tried
        const unsigned long long sp = s / 10ull;
        const bool up_in = lower <= 40ull * sp, wp_in = 40ull * sp + 40ull <= upper;
        if (up_in != wp_in) { *digits = sp + (wp_in ? 1ull : 0ull); *exp10 = k + 1 + dk; return; }
    }
    const bool u_in = lower <= 4ull * s, w_in = 4ull * s + 4ull <= upper;
    if (u_in != w_in) { *digits = s + (w_in ? 1ull : 0ull); *exp10 = k + dk; return; }
    const unsigned long long mid = 4ull * s + 2ull;
    const bool up = vb > mid || (vb == mid && (s & 1ull));
    *digits = s + (up ? 1ull : 0ull); *exp10 = k + dk;
}

// java.lang.Double.toString; quoted: inside a string-typed map (Cast(double AS string)); non-finite values are always quoted
// by Jackson when written as numbers
template <class S> static __device__ void emit_f64(S& s, double x, bool quoted) {
    const unsigned long long bits = (unsigned long long)__double_as_longlong(x);
    const bool neg = (bits >> 63) != 0ull;
    const unsigned long long mag = bits & 0x7fffffffffffffffull;
    if (mag >= 0x7ff0000000000000ull) {
        s.put('"');
        if (mag > 0x7ff0000000000000ull) { s.put('N'); s.put('a'); s.put('N'); }
        else { if (neg) s.put('-'); const char* t = "Infinity"; for (int i = 0; i < 8; i++) s.put((uint8_t)t[i]); }
        s.put('"');
        return;
    }
    if (quoted) s.put('"');
    if (neg) s.put('-');
    if (mag == 0ull) { s.put('0'); s.put('.'); s.put('0'); if (quoted) s.put('"'); return; }
    unsigned long long d; int e10;
    j_shortest(mag, &d, &e10);
    char buf[20]; int n = 0;
    {   // d < 10^17: two pieces of at most 9 and 8 digits, trailing zeros dropped and digits produced in 32-bit arithmetic
        uint32_t hi = (uint32_t)(d / 100000000ull), lo = (uint32_t)(d - (unsigned long long)hi * 100000000ull);
        if (lo == 0u) { e10 += 8; lo = hi; hi = 0u; }
        while (lo % 10u == 0u && (hi != 0u || lo != 0u)) {
            // a trailing zero of the low piece; when the high piece exists the low piece keeps its 8-digit place, so only its value shrinks
            if (hi) break;
            lo /= 10u; e10++;
        }
        if (hi) {
            int z = 0; uint32_t t = lo;
            while (t % 10u == 0u) { t /= 10u; z++; }                           // lo != 0 here
            e10 += z;
            for (int i = 0; i < 8 - z; i++) { buf[n++] = (char)('0' + t % 10u); t /= 10u; }
            do { buf[n++] = (char)('0' + hi % 10u); hi /= 10u; } while (hi);
        } else {
            do { buf[n++] = (char)('0' + lo % 10u); lo /= 10u; } while (lo);
        }
    }                                                                              // buf[n-1] is the first digit
    const int e = n - 1 + e10;                                                    // scientific exponent
    if (e >= 0 && e <= 6) {
        for (int i = 0; i <= e; i++) s.put(i < n ? (uint8_t)buf[n - 1 - i] : (uint8_t)'0');
        s.put('.');
        if (n > e + 1) { for (int i = e + 1; i < n; i++) s.put((uint8_t)buf[n - 1 - i]); } else s.put('0');
    } else if (e < 0 && e >= -3) {
        s.put('0'); s.put('.');
        for (int i = 0; i < -e - 1; i++) s.put('0');
        for (int i = 0; i < n; i++) s.put((uint8_t)buf[n - 1 - i]);
    } else {
        s.put((uint8_t)buf[n - 1]); s.put('.');
        if (n > 1) { for (int i = 1; i < n; i++) s.put((uint8_t)buf[n - 1 - i]); } else s.put('0');
        s.put('E');
        emit_i64(s, (long long)e);
    }
    if (quoted) s.put('"');
}

// days since 1970-01-01 -> civil date (proleptic Gregorian; H. Hinnant's days_from_civil inverse)
static __device__ __forceinline__ void j_civil(long long z, long long* y, int* m, int* d) {
    z += 719468;
    const long long era = (z >= 0 ? z : z - 146096) / 146097;
    const unsigned doe = (unsigned)(z - era * 146097);
    const unsigned yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
    const unsigned doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
    const unsigned mp = (5 * doy + 2) / 153;
    *d = (int)(doy - (153 * mp + 2) / 5 + 1);
    *m = (int)(mp < 10 ? mp + 3 : mp - 9);
    *y = (long long)yoe + era * 400 + (*m <= 2 ? 1 : 0);
}
static __device__ __forceinline__ long long j_floor_div(long long a, long long b) { long long q = a / b; return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q; }
template <class S> static __device__ void emit_date_time(S& s, long long secs, char sep) {
    const long long days = j_floor_div(secs, 86400);
    const int sod = (int)(secs - days * 86400);
    long long y; int m, d;
    j_civil(days, &y, &m, &d);
    if (y < 0) { s.put('-'); y = -y; }
    if (y < 1000) { s.put('0'); if (y < 100) s.put('0'); if (y < 10) s.put('0'); }
    emit_u64(s, (unsigned long long)y);
    s.put('-'); emit_2(s, m); s.put('-'); emit_2(s, d); s.put((uint8_t)sep);
    emit_2(s, sod / 3600); s.put(':'); emit_2(s, (sod / 60) % 60); s.put(':'); emit_2(s, sod % 60);
}
// to_json: yyyy-MM-dd'T'HH:mm:ss.SSSXXX (UTC -> Z), milliseconds floored
template <class S> static __device__ void emit_ts_json(S& s, long long us) {
    const long long ms = j_floor_div(us, 1000), secs = j_floor_div(ms, 1000);
    const int frac = (int)(ms - secs * 1000);
    s.put('"');
    emit_date_time(s, secs, 'T');
    s.put('.'); s.put((uint8_t)('0' + frac / 100)); emit_2(s, frac % 100); s.put('Z');
    s.put('"');
}
// Cast(timestamp AS string): yyyy-MM-dd HH:mm:ss[.ffffff without trailing zeros]
template <class S> static __device__ void emit_ts_cast(S& s, long long us) {
    const long long secs = j_floor_div(us, 1000000);
    int frac = (int)(us - secs * 1000000);
    s.put('"');
    emit_date_time(s, secs, ' ');
    if (frac) {
        s.put('.');
        int div = 100000;
        while (frac) { s.put((uint8_t)('0' + frac / div)); frac %= div; div /= 10; }
    }
    s.put('"');
}
template <class S> static __device__ void emit_string(S& s, const uint8_t* b, uint32_t n) {
    s.put('"');
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t c = b[i];
        if (c >= 0x20 && c != '"' && c != '\\') { s.put(c); continue; }
        s.put('\\');
        switch (c) {
        case '"': s.put('"'); break;
        case '\\': s.put('\\'); break;
        case 0x08: s.put('b'); break;
        case 0x09: s.put('t'); break;
        case 0x0A: s.put('n'); break;
        case 0x0C: s.put('f'); break;
        case 0x0D: s.put('r'); break;
        default: {
            const char* hx = "0123456789ABCDEF";
            s.put('u'); s.put('0'); s.put('0'); s.put((uint8_t)hx[c >> 4]); s.put((uint8_t)hx[c & 15]);
        }
        }
    }
    s.put('"');
}

// ---------------------------------------------------------------------------------------------------------------------
// the row program
// ---------------------------------------------------------------------------------------------------------------------
template <class S> static __device__ void json_row(const JOp* __restrict__ ops, int n_ops, const uint8_t* __restrict__ lits, long long row, long long n_rows, S& s) {
    // A decoded struct is written only if one of its leaves is: its opening is held back (entries [emitted, depth) of the stack)
    // until something inside it is about to be written.  Nothing is ever written past the row's final extent — the neighbouring
    // rows of the CTA are being written at the same time.
    int stack_op[K6_MAX_DEPTH]; uint32_t body_pos[K6_MAX_DEPTH];
    int depth = 0, emitted = 0;
    auto flush = [&]() {
        while (emitted < depth) {
            const JOp& o = ops[stack_op[emitted]];
            s.lit(lits + o.lit_off, o.lit_len); s.put('{');
            body_pos[emitted] = s.p; emitted++;
        }
    };
    for (int i = 0; i < n_ops; i++) {
        const JOp& op = ops[i];
        switch (op.op) {
        case JOP_OPEN:
            stack_op[depth++] = i;
            if (op.kind != JN_DECODED) flush();
            break;
        case JOP_CLOSE:
            depth--;
            if (emitted > depth) {
                emitted = depth;
                if (s.p == body_pos[depth]) s.put('}'); else s.set_last('}');
                s.put(op.is_root ? (uint8_t)'\n' : (uint8_t)',');
            }
            break;
        case JOP_LIT:
            flush();
            s.lit(lits + op.lit_off, op.lit_len);
            break;
        case JOP_FIELD: {
            const long long r = op.stride ? row : 0;
            bool valid = true;
            if (op.vmask) {
                const uint32_t m = op.vwidth == 1 ? (uint32_t)reinterpret_cast<const uint8_t*>(op.vmask)[r] : reinterpret_cast<const uint32_t*>(op.vmask)[r];
                valid = (m >> op.vbit) & 1u;
            }
            int32_t code = 0;
            if (valid && op.type == DX_T_STR) { code = reinterpret_cast<const int32_t*>(op.data)[r]; valid = code >= 0; }
            if (!valid) {
                if (op.null_as_null) { flush(); s.lit(lits + op.lit_off, op.lit_len); s.put('n'); s.put('u'); s.put('l'); s.put('l'); s.put(','); }
                break;
            }
            flush();
            s.lit(lits + op.lit_off, op.lit_len);
            const bool q = op.cast == JC_STRING;
            switch (op.type) {
            case DX_T_I64: {
                const long long v = reinterpret_cast<const long long*>(op.data)[r];
                if (op.cast == JC_DOUBLE) emit_f64(s, (double)v, false);
                else { if (q) s.put('"'); emit_i64(s, v); if (q) s.put('"'); }
                break;
            }
            case DX_T_BOOL: {
                const long long v = reinterpret_cast<const long long*>(op.data)[r];
                if (q) s.put('"');
                if (v) { s.put('t'); s.put('r'); s.put('u'); s.put('e'); } else { s.put('f'); s.put('a'); s.put('l'); s.put('s'); s.put('e'); }
                if (q) s.put('"');
                break;
            }
            case DX_T_F64: emit_f64(s, reinterpret_cast<const double*>(op.data)[r], q); break;
            case DX_T_TS: {
                const long long v = reinterpret_cast<const long long*>(op.data)[r];
                if (q) emit_ts_cast(s, v); else emit_ts_json(s, v);
                break;
            }
            case DX_T_STR: emit_string(s, op.heap + op.code_off[code], op.code_len[code]); break;
            default: s.put('n'); s.put('u'); s.put('l'); s.put('l');
            }
            s.put(',');
            break;
        }
        case JOP_RULES: {
            const uint32_t* fl = reinterpret_cast<const uint32_t*>(op.data);
            bool any = false;
            if (!op.single) { flush(); s.lit(lits + op.lit_off, op.lit_len); s.put('['); }
            for (int w = 0; w < op.flag_words; w++) {
                uint32_t bits = fl[(long long)w * n_rows + row];
                while (bits) {
                    const int b = __ffs((int)bits) - 1; bits &= bits - 1u;
                    const int rule = w * 32 + b;
                    if (rule >= op.n_rules) break;
                    if (op.single && !any) { flush(); s.lit(lits + op.lit_off, op.lit_len); }
                    s.lit(lits + op.frag[2 * rule], op.frag[2 * rule + 1]); s.put(',');
                    any = true;
                }
            }
            if (!op.single) { if (any) s.set_last(']'); else s.put(']'); s.put(','); }
            break;
        }
        }
    }
}

__global__ void __launch_bounds__(K6_NT) k6_json_len(const JOp* __restrict__ ops, int n_ops, const uint8_t* __restrict__ lits, long long n_rows, uint32_t* __restrict__ len) {
    const long long row = (long long)blockIdx.x * K6_NT + threadIdx.x;
    if (row > n_rows) return;
    if (row == n_rows) { len[row] = 0; return; }
    CountSink s;
    json_row(ops, n_ops, lits, row, n_rows, s);
    len[row] = s.p;
}

static __device__ __forceinline__ uint32_t k6_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(K6_NT) k6_json_write(const JOp* __restrict__ ops, int n_ops, const uint8_t* __restrict__ lits, long long n_rows,
                                                       const long long* __restrict__ off, uint8_t* __restrict__ out, uint32_t stage_cap) {
    extern __shared__ __align__(128) uint8_t stage[];
    const long long r0 = (long long)blockIdx.x * K6_NT;
    const long long r1 = r0 + K6_NT < n_rows ? r0 + K6_NT : n_rows;
    const long long row = r0 + threadIdx.x;
    const long long base = off[r0], span = off[r1] - base;
    const uint32_t skew = (uint32_t)(base & 15);
    const bool staged = span + skew <= (long long)stage_cap;          // CTA-uniform
    if (row < n_rows) {
        WriteSink s;
        s.out = staged ? stage + skew + (uint32_t)(off[row] - base) : out + off[row];
        json_row(ops, n_ops, lits, row, n_rows, s);
    }
    if (!staged) return;
    __syncthreads();
    // the span's image sits at stage + skew, congruent to the global address mod 16: ragged head, one bulk store, ragged tail
    uint32_t head = (16u - skew) & 15u; if (head > (uint32_t)span) head = (uint32_t)span;
    const uint32_t mid = ((uint32_t)span - head) & ~15u, tail = (uint32_t)span - head - mid;
    if (threadIdx.x < head) out[base + threadIdx.x] = stage[skew + threadIdx.x];
    if (threadIdx.x >= 32 && threadIdx.x - 32 < tail) out[base + head + mid + (threadIdx.x - 32)] = stage[skew + head + mid + (threadIdx.x - 32)];
    if (threadIdx.x == 64 && mid) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        uint32_t done = 0;
        while (done < mid) {
            uint32_t n = mid - done; if (n > 32768u) n = 32768u;
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                         ::"l"(out + base + head + done), "r"(k6_smem_u32(stage + skew + head + done)), "r"(n) : "memory");
            done += n;
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// host: column tree -> program
// ---------------------------------------------------------------------------------------------------------------------
struct ToI64 { __host__ __device__ long long operator()(uint32_t v) const { return (long long)v; } };
namespace {
struct JNode { std::string key; int kind = -1; int col = -1; std::vector<JNode> kids; };   // kind -1: leaf

std::string jackson_escape(const std::string& s) {
    std::string o;
    for (unsigned char c : s) {
        switch (c) {
        case '"': o += "\\\""; break;
        case '\\': o += "\\\\"; break;
        case 0x08: o += "\\b"; break;
        case 0x09: o += "\\t"; break;
        case 0x0A: o += "\\n"; break;
        case 0x0C: o += "\\f"; break;
        case 0x0D: o += "\\r"; break;
        default:
            if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04X", c); o += b; } else o.push_back((char)c);
        }
    }
    return o;
}

struct Planner {
    const DxTable* t;
    dx_json* js;
    std::vector<JOp> ops;
    std::string lits;
    std::vector<int32_t> frags;                 // all FLAGS columns' fragment spans, patched to device pointers after upload
    std::vector<std::pair<size_t, size_t>> frag_fix;   // (op index, first entry in frags)
    std::string err;

    int lit(const std::string& s, int32_t* off) { *off = (int32_t)lits.size(); lits += s; return (int32_t)s.size(); }
    void key_lit(JOp& op, const std::string& key, bool with_key) {
        if (with_key) op.lit_len = lit("\"" + jackson_escape(key) + "\":", &op.lit_off); else { op.lit_off = 0; op.lit_len = 0; }
    }
    bool leaf(const JNode& n, int cast, bool in_map) {
        const DxColumn& c = t->cols[n.col];
        JOp op; memset(&op, 0, sizeof op);
        const bool is_null = c.type == DX_T_NULL || c.all_null || c.type == 0;
        if (is_null) {
            if (in_map) { op.op = JOP_LIT; op.lit_len = lit("\"" + jackson_escape(n.key) + "\":null,", &op.lit_off); ops.push_back(op); }
            return true;
        }
        if (c.type == DX_T_FLAGS) {
            RuleTable* rt = static_cast<RuleTable*>(c.aux.get());
            if (!rt || !c.data) { err = "rule-flag column '" + c.name + "' has no rule table"; return false; }
            op.op = JOP_RULES; key_lit(op, n.key, true);
            op.data = c.data->p; op.flag_words = c.flag_words; op.n_rules = (int32_t)rt->rules.size(); op.single = rt->single ? 1 : 0;
            frag_fix.emplace_back(ops.size(), frags.size());
            for (auto& rule : rt->rules) {
                std::string o = "{";
                for (size_t i = 0; i < rule.size(); i++) { if (i) o += ","; o += "\"" + jackson_escape(rule[i].first) + "\":\"" + jackson_escape(rule[i].second) + "\""; }
                o += "}";
                int32_t off; const int len = lit(o, &off);
                frags.push_back(off); frags.push_back(len);
            }
            ops.push_back(op);
            return true;
        }
        if (c.type == DX_T_STR && (c.is_const || !c.data)) {      // a constant string: the whole field is a literal
            std::string v = "\"" + jackson_escape(c.const_s) + "\"";
            op.op = JOP_LIT; op.lit_len = lit("\"" + jackson_escape(n.key) + "\":" + v + ",", &op.lit_off); ops.push_back(op);
            return true;
        }
        op.op = JOP_FIELD; op.type = (uint8_t)c.type; op.cast = (uint8_t)cast; op.null_as_null = in_map ? 1 : 0; op.stride = 1;
        key_lit(op, n.key, true);
        if (c.is_const || !c.data) {
            // a constant number: one element on the device, formatted by the same code as every other value
            DevBufP b = make_buf(t->ctx, 8);
            long long bits = c.type == DX_T_F64 ? 0 : (long long)c.const_i;
            if (c.type == DX_T_F64) memcpy(&bits, &c.const_f, 8);
            js->keep.push_back(b); js->consts.push_back(bits);
            js->const_dst.push_back(b->p);
            op.data = b->p; op.stride = 0; op.vmask = nullptr;
        } else {
            op.data = c.data->p;
            if (c.vmask) { op.vmask = c.vmask->p; op.vwidth = c.vwidth; op.vbit = c.vbit; }
        }
        if (c.type == DX_T_STR) {
            if (!c.dict) { err = "string column '" + c.name + "' has no dictionary"; return false; }
            op.heap = c.dict->dev.heap; op.code_off = c.dict->dev.code_off; op.code_len = c.dict->dev.code_len;
        }
        ops.push_back(op);
        return true;
    }
    bool node(const JNode& n, bool is_root, bool in_map) {
        JOp op; memset(&op, 0, sizeof op);
        op.op = JOP_OPEN; op.kind = (uint8_t)n.kind; op.is_root = is_root; key_lit(op, n.key, !is_root);
        ops.push_back(op);
        // Spark's common type for the values of one MAP(...)
        int cast = JC_NONE;
        if (n.kind == JN_MAP) {
            std::set<int> tys;
            for (auto& k : n.kids) if (k.kind < 0) { const int ty = t->cols[k.col].type; if (ty != DX_T_NULL && ty != DX_T_FLAGS && ty != 0 && !t->cols[k.col].all_null) tys.insert(ty); }
            if (tys.size() > 1) { if (tys.count(DX_T_STR)) cast = JC_STRING; else if (tys.size() == 2 && tys.count(DX_T_I64) && tys.count(DX_T_F64)) cast = JC_DOUBLE; }
        }
        for (auto& k : n.kids) {
            if (k.kind < 0) {
                int c = JC_NONE;
                const int ty = t->cols[k.col].type;
                if (cast == JC_STRING && ty != DX_T_STR) c = JC_STRING;
                if (cast == JC_DOUBLE && ty == DX_T_I64) c = JC_DOUBLE;
                if (!leaf(k, c, n.kind == JN_MAP)) return false;
            } else if (!node(k, false, n.kind == JN_MAP)) return false;
        }
        memset(&op, 0, sizeof op);
        op.op = JOP_CLOSE; op.kind = (uint8_t)n.kind; op.is_root = is_root;
        ops.push_back(op);
        return true;
    }
};

std::vector<std::string> split_dots(const std::string& s) {
    std::vector<std::string> o; size_t a = 0;
    for (;;) { size_t b = s.find('.', a); if (b == std::string::npos) { o.push_back(s.substr(a)); break; } o.push_back(s.substr(a, b - a)); a = b + 1; }
    return o;
}
}  // namespace

int dx_json_build(const DxTable* t, dx_json** out) {
    dx_ctx* c = t->ctx;
    auto js = std::unique_ptr<dx_json>(new dx_json());
    js->ctx = c; js->n_rows = t->n_rows;
    // ---- the column tree (OutputManager.scala:92: internal columns are not output) ----
    JNode root; root.kind = JN_STRUCT;
    for (size_t ci = 0; ci < t->cols.size(); ci++) {
        const DxColumn& col = t->cols[ci];
        if (col.name.rfind("__DataX_", 0) == 0) continue;
        std::vector<std::string> keys; std::vector<int> kinds;
        if (!col.nest_path.empty()) {
            size_t plen = 0; for (auto& p : col.nest_path) plen += p.size() + 1;
            keys.push_back(col.name.substr(0, col.name.size() - plen));
            for (auto& p : col.nest_path) keys.push_back(p);
            for (char k : col.nest_kinds) kinds.push_back(k == 'm' ? JN_MAP : JN_STRUCT);
        } else if (dx_is_metadata_column(col.name)) {
            const size_t dot = col.name.find('.');
            keys = {col.name.substr(0, dot), col.name.substr(dot + 1)}; kinds = {JN_MAP};
        } else {
            keys = split_dots(col.name);
            kinds.assign(keys.size() - 1, JN_DECODED);
        }
        if (keys.size() > K6_MAX_DEPTH) return c->set_error(DX_ERR_UNSUPPORTED, "to_json: column '" + col.name + "' nests deeper than " + std::to_string(K6_MAX_DEPTH - 1) + " levels");
        JNode* n = &root;
        for (size_t d = 0; d + 1 < keys.size(); d++) {
            JNode* last = n->kids.empty() ? nullptr : &n->kids.back();
            if (!last || last->kind < 0 || last->key != keys[d] || last->kind != kinds[d]) { JNode nn; nn.key = keys[d]; nn.kind = kinds[d]; n->kids.push_back(nn); last = &n->kids.back(); }
            n = last;
        }
        JNode lf; lf.key = keys.back(); lf.col = (int)ci;
        n->kids.push_back(lf);
    }
    Planner P; P.t = t; P.js = js.get();
    if (!P.node(root, true, false)) return c->set_error(DX_ERR_UNSUPPORTED, "to_json: " + P.err);
    // ---- upload the program ----
    const int64_t n = t->n_rows;
    js->offsets = make_buf(c, (size_t)(n + 1) * 8);
    if (n == 0) { DX_CUDA(c, cudaMemsetAsync(js->offsets->p, 0, 8, c->stream)); js->n_bytes = 0; *out = js.release(); return DX_OK; }
    DevBufP d_lits = make_buf(c, P.lits.size() + 16), d_frags = make_buf(c, P.frags.size() * 4 + 16), d_ops = make_buf(c, P.ops.size() * sizeof(JOp));
    for (auto& fx : P.frag_fix) P.ops[fx.first].frag = d_frags->as<int32_t>() + fx.second;
    DX_CUDA(c, cudaMemcpyAsync(d_lits->p, P.lits.data(), P.lits.size(), cudaMemcpyHostToDevice, c->stream));
    if (!P.frags.empty()) DX_CUDA(c, cudaMemcpyAsync(d_frags->p, P.frags.data(), P.frags.size() * 4, cudaMemcpyHostToDevice, c->stream));
    DX_CUDA(c, cudaMemcpyAsync(d_ops->p, P.ops.data(), P.ops.size() * sizeof(JOp), cudaMemcpyHostToDevice, c->stream));
    for (size_t i = 0; i < js->consts.size(); i++) DX_CUDA(c, cudaMemcpyAsync(js->const_dst[i], &js->consts[i], 8, cudaMemcpyHostToDevice, c->stream));
    // ---- pass 1: lengths; scan; pass 2: bytes ----
    DevBufP d_len = make_buf(c, (size_t)(n + 1) * 4);
    const unsigned grid_len = (unsigned)((n + 1 + K6_NT - 1) / K6_NT), grid_w = (unsigned)((n + K6_NT - 1) / K6_NT);
    {
        StageTimer tm(c, "k6_json_len");
        k6_json_len<<<grid_len, K6_NT, 0, c->stream>>>(d_ops->as<JOp>(), (int)P.ops.size(), d_lits->as<uint8_t>(), n, d_len->as<uint32_t>());
        c->launches["k6_json_len"]++;
    }
    {
        StageTimer tm(c, "k6_json_scan");
        cub::TransformInputIterator<long long, ToI64, const uint32_t*> in(d_len->as<uint32_t>(), ToI64());
        size_t tmp_bytes = 0;
        DX_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, in, js->offsets->as<long long>(), (long long)(n + 1), c->stream));
        DevBufP tmp = make_buf(c, tmp_bytes + 16);
        DX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp->p, tmp_bytes, in, js->offsets->as<long long>(), (long long)(n + 1), c->stream));
        js->keep.push_back(tmp);
    }
    long long total = 0;
    DX_CUDA(c, cudaMemcpyAsync(&total, js->offsets->as<long long>() + n, 8, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    js->n_bytes = total;
    js->bytes = make_buf(c, (size_t)total + 16);
    if (c->k6_smem_configured < (size_t)K6_STAGE + 128) {
        DX_CUDA(c, cudaFuncSetAttribute(k6_json_write, cudaFuncAttributeMaxDynamicSharedMemorySize, K6_STAGE + 128));
        c->k6_smem_configured = (size_t)K6_STAGE + 128;
    }
    // the stage holds the output span of one CTA's 128 rows: sized from the mean row (known from the scan) with a third to spare,
    // so short rows leave room for more resident CTAs; a CTA whose span is larger than that writes straight to global memory
    size_t stage = (size_t)((double)total / (double)n * K6_NT * 1.35) + 512;
    stage = (stage + 1023) & ~size_t(1023);
    if (stage < 8192) stage = 8192;
    if (stage > K6_STAGE) stage = K6_STAGE;
    {
        StageTimer tm(c, "k6_json_write");
        k6_json_write<<<grid_w, K6_NT, stage + 128, c->stream>>>(d_ops->as<JOp>(), (int)P.ops.size(), d_lits->as<uint8_t>(), n, js->offsets->as<long long>(), js->bytes->as<uint8_t>(), (uint32_t)stage);
        c->launches["k6_json_write"]++;
    }
    DX_CUDA(c, cudaGetLastError());
    js->keep.push_back(d_lits); js->keep.push_back(d_frags); js->keep.push_back(d_ops); js->keep.push_back(d_len);
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    js->keep.clear();
    *out = js.release();
    return DX_OK;
}
