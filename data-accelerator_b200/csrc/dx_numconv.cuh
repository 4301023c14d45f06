// dx_numconv.cuh — decimal text -> binary64, correctly rounded (what Jackson's getDoubleValue /
// Double.parseDouble give Spark's from_json, reference call site CommonProcessorFactory.scala:93).
// Clinger fast path, Eisel-Lemire with a 128-bit power-of-five table, and an exact big-integer
// tie-break for inputs whose first 19 digits cannot decide the rounding.  Kept free of CUDA
// headers so tests/host_numconv.cpp can compile the very same code for the CPU test-suite.
#pragma once
#include <stdint.h>
#include "dx_pow5_table.h"

// ------------------------------- decimal -> double -------------------------------------
static __constant__ double dx_pow10_exact[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11,
                                              1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// Eisel-Lemire: w * 10^q -> nearest binary64 bits (w != 0).  Returns false if q out of the table.
static __device__ bool dx_eisel_lemire(uint64_t w, int q, uint64_t* bits_out) {
    if (q < DX_POW5_MIN_Q) { *bits_out = 0; return true; }
    if (q > DX_POW5_MAX_Q) { *bits_out = 0x7FF0000000000000ull; return true; }
    int lz = __clzll((long long)w);
    w <<= lz;
    const uint64_t* tp = dx_pow5_128 + 2 * (q - DX_POW5_MIN_Q);
    uint64_t hi = __umul64hi(w, tp[0]), lo = w * tp[0];
    const uint64_t precision_mask = 0xFFFFFFFFFFFFFFFFull >> 55;
    if ((hi & precision_mask) == precision_mask) {
        uint64_t hi2 = __umul64hi(w, tp[1]);
        lo += hi2;
        if (hi2 > lo) hi++;
    }
    int upperbit = (int)(hi >> 63);
    uint64_t mantissa = hi >> (upperbit + 64 - 52 - 3);
    int power2 = (((152170 + 65536) * q) >> 16) + 63 + upperbit - lz + 1023;
    if (power2 <= 0) {
        if (-power2 + 1 >= 64) { *bits_out = 0; return true; }
        mantissa >>= -power2 + 1;
        mantissa += (mantissa & 1);
        mantissa >>= 1;
        power2 = (mantissa < (1ull << 52)) ? 0 : 1;
        *bits_out = (mantissa & ~(1ull << 52)) | ((uint64_t)power2 << 52);
        return true;
    }
    if (lo <= 1 && q >= -4 && q <= 23 && (mantissa & 3) == 1) {
        if ((mantissa << (upperbit + 64 - 52 - 3)) == hi) mantissa &= ~1ull;
    }
    mantissa += (mantissa & 1);
    mantissa >>= 1;
    if (mantissa >= (2ull << 52)) { mantissa = 1ull << 52; power2++; }
    mantissa &= ~(1ull << 52);
    if (power2 >= 0x7FF) { *bits_out = 0x7FF0000000000000ull; return true; }
    *bits_out = mantissa | ((uint64_t)power2 << 52);
    return true;
}

// ---- exact tie-break for decimals the 19-digit window cannot decide -----------------------
// Fixed 4096-bit unsigned integer in local memory; only the rare undecided number gets here.
#define DX_BIG_LIMBS 128
#define DX_BIG_DIGITS 800   // an exact tie needs at most 767 significant digits
struct DxBig { uint32_t w[DX_BIG_LIMBS]; int n; };
static __device__ bool big_mul_add(DxBig& b, uint32_t m, uint32_t a) {
    uint64_t carry = a;
    for (int i = 0; i < b.n; i++) { uint64_t t = (uint64_t)b.w[i] * m + carry; b.w[i] = (uint32_t)t; carry = t >> 32; }
    if (carry) { if (b.n >= DX_BIG_LIMBS) return false; b.w[b.n++] = (uint32_t)carry; }
    return true;
}
static __device__ bool big_mul_pow5(DxBig& b, int k) {
    while (k >= 13) { if (!big_mul_add(b, 1220703125u, 0)) return false; k -= 13; }
    uint32_t m = 1; for (int i = 0; i < k; i++) m *= 5;
    return k ? big_mul_add(b, m, 0) : true;
}
static __device__ bool big_shl(DxBig& b, int bits) {
    int limbs = bits >> 5, r = bits & 31;
    if (b.n == 0) return true;
    if (b.n + limbs + 1 > DX_BIG_LIMBS) return false;
    if (r) {
        uint32_t carry = 0;
        for (int i = 0; i < b.n; i++) { uint32_t v = b.w[i]; b.w[i] = (v << r) | carry; carry = v >> (32 - r); }
        if (carry) b.w[b.n++] = carry;
    }
    if (limbs) {
        for (int i = b.n - 1; i >= 0; i--) b.w[i + limbs] = b.w[i];
        for (int i = 0; i < limbs; i++) b.w[i] = 0;
        b.n += limbs;
    }
    return true;
}
static __device__ int big_cmp(const DxBig& a, const DxBig& b) {
    if (a.n != b.n) return a.n < b.n ? -1 : 1;
    for (int i = a.n - 1; i >= 0; i--) if (a.w[i] != b.w[i]) return a.w[i] < b.w[i] ? -1 : 1;
    return 0;
}
// Compares |value of the number token s..e| with the midpoint between the positive double b1 and
// its successor.  -1 below, 0 exactly on it, +1 above, 2 = out of the bignum's range.
static __device__ __noinline__ int dx_cmp_to_midpoint(const uint8_t* s, const uint8_t* e, uint64_t b1) {
    DxBig L, R;
    L.n = 0;
    const uint8_t* p = s;
    if (*p == '-') p++;
    int int_sig = 0, lead_frac_zeros = 0, taken = 0; bool seen_point = false, seen_sig = false, sticky = false;
    uint32_t chunk = 0, chunk_mul = 1;
    for (; p < e; p++) {
        uint8_t c = *p;
        if (c == '.') { seen_point = true; continue; }
        if (c == 'e' || c == 'E') break;
        uint32_t d = c - '0';
        if (!seen_sig) { if (d == 0) { if (seen_point) lead_frac_zeros++; continue; } seen_sig = true; }
        if (!seen_point) int_sig++;
        if (taken < DX_BIG_DIGITS) {
            chunk = chunk * 10 + d; chunk_mul *= 10; taken++;
            if (chunk_mul == 1000000000u) {
                if (L.n == 0) { if (chunk) { L.w[0] = chunk; L.n = 1; } }
                else if (!big_mul_add(L, chunk_mul, chunk)) return 2;
                chunk = 0; chunk_mul = 1;
            }
        } else if (d) sticky = true;
    }
    if (chunk_mul > 1) {
        if (L.n == 0) { if (chunk) { L.w[0] = chunk; L.n = 1; } }
        else if (!big_mul_add(L, chunk_mul, chunk)) return 2;
    }
    int exp10 = 0;
    if (p < e) {
        p++;
        bool eneg = false;
        if (*p == '+') p++; else if (*p == '-') { eneg = true; p++; }
        int ev = 0;
        for (; p < e; p++) if (ev < 100000) ev = ev * 10 + (*p - '0');
        exp10 = eneg ? -ev : ev;
    }
    const int point_pos = int_sig ? int_sig : -lead_frac_zeros;
    const int p10 = point_pos - taken + exp10;
    const uint32_t be = (uint32_t)(b1 >> 52) & 0x7FFu;
    const uint64_t frac = b1 & ((1ull << 52) - 1);
    const uint64_t M = be ? (frac | (1ull << 52)) : frac;
    const int E = be ? (int)be - 1075 : -1074;
    const uint64_t mid = 2 * M + 1;
    R.w[0] = (uint32_t)mid; R.w[1] = (uint32_t)(mid >> 32); R.n = R.w[1] ? 2 : 1;
    int l2 = 0, r2 = 0;
    if (p10 > 0) { if (p10 > 1200 || !big_mul_pow5(L, p10)) return 2; l2 += p10; }
    else if (p10 < 0) { if (-p10 > 1200 || !big_mul_pow5(R, -p10)) return 2; r2 += -p10; }
    const int e2 = E - 1;
    if (e2 > 0) r2 += e2; else l2 += -e2;
    const int common = l2 < r2 ? l2 : r2;
    if (!big_shl(L, l2 - common) || !big_shl(R, r2 - common)) return 2;
    int c = big_cmp(L, R);
    if (c == 0 && sticky) c = 1;
    return c;
}

// s..e: a syntactically valid JSON number token.  is_int: no fraction/exponent.
// Returns false when the value cannot be decided with the 19-digit window (flagged by caller).
static __device__ bool dx_number_to_double(const uint8_t* s, const uint8_t* e, bool is_int, double* out) {
    const uint8_t* p = s;
    bool neg = false;
    if (*p == '-') { neg = true; p++; }
    uint64_t w = 0; int nd = 0; int dropped = 0; bool trunc_nz = false; int frac_digits = 0;
    bool in_frac = false;
    for (; p < e; p++) {
        uint8_t c = *p;
        if (c == '.') { in_frac = true; continue; }
        if (c == 'e' || c == 'E') break;
        uint32_t d = c - '0';
        if (nd == 0 && d == 0) { if (in_frac) frac_digits++; continue; } // leading zeros carry no digits
        if (nd < 19) { w = w * 10 + d; nd++; if (in_frac) frac_digits++; }
        else { if (!in_frac) dropped++; if (d) trunc_nz = true; }
    }
    int exp10 = 0;
    if (p < e) { // exponent
        p++;
        bool eneg = false;
        if (*p == '+') p++; else if (*p == '-') { eneg = true; p++; }
        int ev = 0;
        for (; p < e; p++) { if (ev < 100000) ev = ev * 10 + (*p - '0'); }
        exp10 = eneg ? -ev : ev;
    }
    exp10 += dropped - frac_digits;
    if (w == 0) { *out = neg ? -0.0 : 0.0; if (is_int) *out = 0.0; return true; }
    if (is_int && nd <= 18 && !trunc_nz && dropped == 0) {
        // Jackson parses an integer token as int/long first; (double) of that long
        long long v = neg ? -(long long)w : (long long)w;
        *out = (double)v; return true;
    }
    double r;
    if (!trunc_nz && w < (1ull << 53) && exp10 >= -22 && exp10 <= 22) {
        r = (double)w;
        if (exp10 < 0) r = r / dx_pow10_exact[-exp10]; else r = r * dx_pow10_exact[exp10];
    } else {
        uint64_t b1;
        dx_eisel_lemire(w, exp10, &b1);
        if (trunc_nz) {
            uint64_t b2;
            dx_eisel_lemire(w + 1, exp10, &b2);
            if (b1 != b2) {
                // the 19-digit window cannot decide: exact comparison against the midpoint
                int cmp = dx_cmp_to_midpoint(s, e, b1);
                if (cmp == 2) return false;
                if (cmp > 0 || (cmp == 0 && (b1 & 1))) b1 = b2;
            }
        }
        r = __longlong_as_double((long long)b1);
    }
    *out = neg ? -r : r;
    return true;
}

