// dx_decode_fast.cuh — the template fast path of K1 (see dx_decode.cu): one lane parses one
// record out of the shared-memory tile, the 32 lanes of a warp walk the template segments in
// lock step (reconverging at every segment), literals are compared a word at a time through a
// carried funnel shift, digit runs and string bodies are classified eight bytes at a time (SWAR).
// The path accepts only the schema-ordered compact serialisation with plain integers, short
// decimals and escape-free strings; anything else reports false and the record is re-parsed from
// its first byte by the general Spark-conformant parser, so results never depend on which path ran.
#pragma once
#include "dx_decode.cuh"

static __device__ __forceinline__ uint64_t ld64u(const uint8_t* buf, uint32_t pos) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(buf + (pos & ~3u));
    const uint32_t sh = (pos & 3u) << 3;
    const uint32_t a = w[0], b = w[1], c = w[2];
    return ((uint64_t)__funnelshift_r(b, c, sh) << 32) | __funnelshift_r(a, b, sh);
}

// number of leading ASCII digits among the 4 bytes of w (byte 0 first)
static __device__ __forceinline__ int leading_digits4(uint32_t w) {
    const uint32_t nd = ((w + 0x46464646u) | (w - 0x30303030u)) & 0x80808080u;   // exact up to and including the first non-digit
    return __clz((int)__brev(nd)) >> 3;                    // trailing zeros / 8; no flag at all -> 32 / 8 = 4
}

// value of the first m (1..4) digits of the 32-bit window x
static __device__ __forceinline__ uint32_t digits4(uint32_t x, int m) {
    // right-align: the shift drops the bytes past the m-th, missing leading digits become zero (a borrow of the subtraction can
    // only start in a non-digit byte, i.e. above the digits, so the digit bytes are exact)
    const uint32_t t = (x - 0x30303030u) << ((4 - m) << 3);
    const uint32_t p = (t * 10u + (t >> 8)) & 0x00FF00FFu; // [d0 d1] in byte 0, [d2 d3] in byte 2
    return (p * 0x00640001u) >> 16;                        // (d0 d1) * 100 + (d2 d3)
}

static __constant__ uint32_t dx_pow10_u32[9] = {1u, 10u, 100u, 1000u, 10000u, 100000u, 1000000u, 10000000u, 100000000u};

// Parses a run of digits at pos, four bytes per step through a carried pair of aligned words.
// Returns the digit count (0 = none, -1 = more than 18), the value, the first digit and the byte that ends the run.
static __device__ __forceinline__ int parse_digit_run(const uint8_t* buf, uint32_t pos, unsigned long long* value, uint32_t* first_digit, uint32_t* term) {
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(buf + (pos & ~3u));
    const uint32_t sh = (pos & 3u) << 3;
    uint32_t a = wp[0], b = wp[1];
    uint32_t w = __funnelshift_r(a, b, sh);
    *first_digit = (w & 0xffu) - '0';
    int n = leading_digits4(w);
    if (n == 0) { *value = 0; *term = w & 0xffu; return 0; }
    uint32_t v32 = digits4(w, n); int nd = n;
    // up to 8 digits stay in 32-bit arithmetic
    if (n == 4) {
        a = b; b = wp[2]; w = __funnelshift_r(a, b, sh);
        n = leading_digits4(w);
        if (n) { v32 = v32 * dx_pow10_u32[n] + digits4(w, n); nd += n; }
    }
    unsigned long long v = v32;
    if (nd == 8) {
        int k = 3;
        for (;;) {
            a = b; b = wp[k++]; w = __funnelshift_r(a, b, sh);
            n = leading_digits4(w);
            if (n) { v = v * dx_pow10_u32[n] + digits4(w, n); nd += n; }
            if (n < 4) break;
            if (nd > 18) return -1;
        }
    }
    *value = v;
    *term = (w >> (n << 3)) & 0xffu;          // n < 4 here: the run ended inside the last window
    return nd > 18 ? -1 : nd;
}

// One record, executed by ALL lanes of the warp together (lanes without a record pass has=false).
static __device__ __forceinline__ bool fast_record_warp(const uint8_t* buf, uint32_t pos, uint32_t end, bool has, const DecodeArgs& a, RowSink& o) {
    bool ok = has;
    o.valid = 0;
    const int n_segs = a.n_segs;
    for (int si = 0; si < n_segs; si++) {
        __syncwarp();
        const DSeg sg = a.segs[si];                      // warp-uniform: constant-bank loads
        {
            // literal: a straight line of word compares entered at the literal's length (fall-through switch), each word one
            // shared load, a funnel shift and an XOR against a constant-bank operand; the last word goes under its mask.
            // Every lane runs the compare, lanes that already failed on the tile's first bytes: the dispatch is then
            // warp-uniform and the literal words are uniform operands instead of per-lane indexed loads.
            const uint32_t p0 = ok ? pos : 0u, sh = (p0 & 3u) << 3;
            const uint32_t* wp = reinterpret_cast<const uint32_t*>(buf + (p0 & ~3u));
            // the literal words are read as a.lits[index]: indexing the parameter block keeps them constant-bank loads (a pointer
            // into it that is advanced in a loop turns into generic global loads, one round of memory latency per segment)
            int li = (int)sg.lit_word;       // signed: the +i folds into the load's immediate offset
            uint32_t nf = sg.n_words - 1u, diff = 0;
#define DX_LIT_WORD(i) diff |= __funnelshift_r(wp[i], wp[(i) + 1], sh) ^ a.lits[li + (i)];
#pragma unroll 1
            while (nf > 8u) { DX_LIT_WORD(0) DX_LIT_WORD(1) DX_LIT_WORD(2) DX_LIT_WORD(3) DX_LIT_WORD(4) DX_LIT_WORD(5) DX_LIT_WORD(6) DX_LIT_WORD(7) wp += 8; li += 8; nf -= 8u; }
#undef DX_LIT_WORD
            // from the last word down, each step's low data word is the next step's high one: one shared load per literal word
            // (two registers swap the roles from case to case, so the fall-through needs no copy between the steps)
            uint32_t wa = wp[nf], wb = wa;
            diff |= (__funnelshift_r(wa, wp[nf + 1u], sh) ^ a.lits[li + (int)nf]) & sg.last_mask;
#define DX_LIT_EVEN(i) wb = wp[i]; diff |= __funnelshift_r(wb, wa, sh) ^ a.lits[li + (i)];     /* entered with nf even: high word in wa */
#define DX_LIT_ODD(i)  wa = wp[i]; diff |= __funnelshift_r(wa, wb, sh) ^ a.lits[li + (i)];     /* entered with nf odd: high word in wb */
            switch (nf) {
            case 8: DX_LIT_EVEN(7)
            case 7: DX_LIT_ODD(6)
            case 6: DX_LIT_EVEN(5)
            case 5: DX_LIT_ODD(4)
            case 4: DX_LIT_EVEN(3)
            case 3: DX_LIT_ODD(2)
            case 2: DX_LIT_EVEN(1)
            case 1: DX_LIT_ODD(0)
            default: break;
            }
#undef DX_LIT_EVEN
#undef DX_LIT_ODD
            ok = ok && diff == 0;
            pos += sg.lit_len;
        }
        if (sg.type == 0) break;
        if (ok) {
            switch (sg.type) {
            case DX_T_I64: {
                bool neg = buf[pos] == '-';
                if (neg) pos++;
                unsigned long long v; uint32_t d0, tm;
                const int nd = parse_digit_run(buf, pos, &v, &d0, &tm);
                if (nd <= 0 || (d0 == 0 && nd > 1)) { ok = false; break; }
                sink_i64(o, sg.leaf, neg ? -(long long)v : (long long)v);
                pos += nd;
                break;
            }
            case DX_T_F64: {
                bool neg = buf[pos] == '-';
                if (neg) pos++;
                unsigned long long ip, fp = 0; uint32_t d0, f0, tm;
                const int ni = parse_digit_run(buf, pos, &ip, &d0, &tm);
                if (ni <= 0 || (d0 == 0 && ni > 1)) { ok = false; break; }
                pos += ni;
                int nf = 0;
                if (tm == '.') {
                    nf = parse_digit_run(buf, pos + 1, &fp, &f0, &tm);
                    if (nf <= 0) { ok = false; break; }
                    pos += 1 + nf;
                }
                if (ni + nf > 15 || tm == 'e' || tm == 'E') { ok = false; break; }
                double x;
                if (nf == 0) x = (double)(neg ? -(long long)ip : (long long)ip);
                else {
                    unsigned long long m = ip;
                    if (nf > 8) { m = m * dx_pow10_u32[8]; m = m * dx_pow10_u32[nf - 8]; } else m = m * dx_pow10_u32[nf];
                    x = (double)(m + fp) / dx_pow10_exact[nf];
                    if (neg) x = -x;
                }
                sink_f64(o, sg.leaf, x);
                break;
            }
            case DX_T_BOOL: {
                const uint32_t w = ld32u(buf, pos);
                if (w == 0x65757274u) { sink_i64(o, sg.leaf, 1); pos += 4; }
                else if (w == 0x736c6166u && buf[pos + 4] == 'e') { sink_i64(o, sg.leaf, 0); pos += 5; }
                else ok = false;
                break;
            }
            case DX_T_STR: {
                if (buf[pos] != '"') { ok = false; break; }
                const uint32_t s0 = pos + 1;
                uint64_t w0 = 0, w1 = 0; uint32_t len = 0; bool bad = false;
                for (;;) {
                    const uint64_t w = ld64u(buf, s0 + len);
                    // first byte that is '"', '\\' or a control character
                    const uint64_t q = w ^ 0x2222222222222222ull, b = w ^ 0x5C5C5C5C5C5C5C5Cull;
                    const uint64_t zq = (q - 0x0101010101010101ull) & ~q, zb = (b - 0x0101010101010101ull) & ~b, zc = (w - 0x2020202020202020ull) & ~w;
                    const uint64_t stop = (zq | zb | zc) & 0x8080808080808080ull;
                    const int n = stop ? ((__ffsll((long long)stop) - 1) >> 3) : 8;
                    if (len == 0) w0 = w; else if (len == 8) w1 = w;
                    len += n;
                    if (n < 8) { bad = ((w >> (n << 3)) & 0xffu) != '"'; break; }
                    if (s0 + len >= end) { bad = true; break; }
                }
                if (bad) { ok = false; break; }
                if (len <= 16) {   // keep exactly the string's bytes in the packed identity
                    if (len < 8) { w0 &= (len ? (~0ull >> ((8 - len) << 3)) : 0ull); w1 = 0; }
                    else if (len < 16) w1 &= (len > 8 ? (~0ull >> ((16 - len) << 3)) : 0ull);
                } else dict_pack(buf + s0, len, w0, w1);
                const int32_t code = dict_get_or_insert(o.a->dicts[sg.leaf], w0, w1, len, buf + s0, &o.a->stats->flags);
                sink_code(o, sg.leaf, code);
                pos = s0 + len + 1;
                break;
            }
            default: ok = false;
            }
        }
    }
    return ok && pos == end;
}
