// dx_expr_eval.cuh — device-side interpreter of expression programs (shared by the generic K2
// kernels in dx_expr.cu and the bucketed rule kernel in dx_rules.cu).  Semantics: see dx_expr.h.
#pragma once
#include "dx_expr.h"

// ------------------------------- device evaluation ---------------------------------------
struct EvalView {
    const uint32_t* code; const long long* consts; const ProgCol* cols; const ProgLut* luts;
};

static __device__ __forceinline__ int cmp_f_spark(double a, double b) {  // NaN greatest, NaN == NaN
    bool an = a != a, bn = b != b;
    if (an || bn) return an == bn ? 0 : (an ? 1 : -1);
    return a < b ? -1 : (a > b ? 1 : 0);
}

static __device__ void eval_program(const EvalView& v, uint32_t pc, long long row, long long& out, bool& out_null) {
    long long st[16]; uint32_t nulls = 0; int sp = 0;
    for (;; pc++) {
        const uint32_t ins = v.code[pc];
        const uint32_t op = DX_OPCODE(ins), arg = DX_OPARG(ins);
        switch (op) {
        case OP_END: out = st[0]; out_null = nulls & 1u; return;
        case OP_LOAD: {
            const ProgCol c = v.cols[arg];
            bool isnull = false;
            if (c.vmask) {
                uint32_t m = c.vwidth == 1 ? reinterpret_cast<const uint8_t*>(c.vmask)[row] : reinterpret_cast<const uint32_t*>(c.vmask)[row];
                isnull = !((m >> c.vbit) & 1u);
            }
            long long x = c.width == 4 ? (long long)reinterpret_cast<const int32_t*>(c.data)[row] : reinterpret_cast<const long long*>(c.data)[row];
            st[sp] = x; nulls = (nulls & ~(1u << sp)) | ((isnull ? 1u : 0u) << sp); sp++;
            break;
        }
        case OP_CONST: st[sp] = v.consts[arg]; nulls &= ~(1u << sp); sp++; break;
        case OP_NULL: st[sp] = 0; nulls |= 1u << sp; sp++; break;
        case OP_I2F: st[sp - 1] = __double_as_longlong((double)st[sp - 1]); break;
        case OP_F2I: {
            double d = __longlong_as_double(st[sp - 1]);
            long long r = (d != d) ? 0 : (d >= 9.2233720368547758e18 ? 0x7FFFFFFFFFFFFFFFll : (d <= -9.2233720368547758e18 ? (long long)0x8000000000000000ull : (long long)d));
            st[sp - 1] = r; break;
        }
        case OP_EQ_I: case OP_NE_I: case OP_LT_I: case OP_LE_I: case OP_GT_I: case OP_GE_I: {
            long long b = st[sp - 1], a = st[sp - 2]; sp--;
            bool r = op == OP_EQ_I ? a == b : op == OP_NE_I ? a != b : op == OP_LT_I ? a < b : op == OP_LE_I ? a <= b : op == OP_GT_I ? a > b : a >= b;
            uint32_t n = ((nulls >> sp) | (nulls >> (sp - 1))) & 1u;
            st[sp - 1] = r; nulls = (nulls & ~(3u << (sp - 1))) | (n << (sp - 1));
            break;
        }
        case OP_EQ_F: case OP_NE_F: case OP_LT_F: case OP_LE_F: case OP_GT_F: case OP_GE_F: {
            double b = __longlong_as_double(st[sp - 1]), a = __longlong_as_double(st[sp - 2]); sp--;
            int c = cmp_f_spark(a, b);
            bool r = op == OP_EQ_F ? c == 0 : op == OP_NE_F ? c != 0 : op == OP_LT_F ? c < 0 : op == OP_LE_F ? c <= 0 : op == OP_GT_F ? c > 0 : c >= 0;
            uint32_t n = ((nulls >> sp) | (nulls >> (sp - 1))) & 1u;
            st[sp - 1] = r; nulls = (nulls & ~(3u << (sp - 1))) | (n << (sp - 1));
            break;
        }
        case OP_AND: case OP_OR: {
            bool bn = (nulls >> (sp - 1)) & 1u, an = (nulls >> (sp - 2)) & 1u;
            bool b = st[sp - 1] != 0, a = st[sp - 2] != 0; sp--;
            bool r, n;
            if (op == OP_AND) { bool af = !an && !a, bf = !bn && !b; r = !(af || bf) && !(an || bn); n = !(af || bf) && (an || bn); }
            else { bool at = !an && a, bt = !bn && b; r = at || bt; n = !r && (an || bn); }
            st[sp - 1] = r; nulls = (nulls & ~(3u << (sp - 1))) | ((n ? 1u : 0u) << (sp - 1));
            break;
        }
        case OP_NOT: st[sp - 1] = !st[sp - 1]; break;
        case OP_ISNULL: st[sp - 1] = (nulls >> (sp - 1)) & 1u; nulls &= ~(1u << (sp - 1)); break;
        case OP_ISNOTNULL: st[sp - 1] = !((nulls >> (sp - 1)) & 1u); nulls &= ~(1u << (sp - 1)); break;
        case OP_ADD_I: case OP_SUB_I: case OP_MUL_I: case OP_MOD_I: {
            unsigned long long b = (unsigned long long)st[sp - 1], a = (unsigned long long)st[sp - 2]; sp--;
            long long r; uint32_t n = ((nulls >> sp) | (nulls >> (sp - 1))) & 1u;
            if (op == OP_ADD_I) r = (long long)(a + b); else if (op == OP_SUB_I) r = (long long)(a - b);
            else if (op == OP_MUL_I) r = (long long)(a * b);
            else { long long sb = (long long)b; if (sb == 0) { r = 0; n = 1; } else if (sb == -1) r = 0; else r = (long long)a % sb; }
            st[sp - 1] = r; nulls = (nulls & ~(3u << (sp - 1))) | (n << (sp - 1));
            break;
        }
        case OP_ADD_F: case OP_SUB_F: case OP_MUL_F: case OP_DIV_F: case OP_MOD_F: {
            double b = __longlong_as_double(st[sp - 1]), a = __longlong_as_double(st[sp - 2]); sp--;
            uint32_t n = ((nulls >> sp) | (nulls >> (sp - 1))) & 1u;
            double r;
            if (op == OP_ADD_F) r = a + b; else if (op == OP_SUB_F) r = a - b; else if (op == OP_MUL_F) r = a * b;
            else if (op == OP_DIV_F) { if (b == 0.0) { r = 0; n = 1; } else r = a / b; }   // Spark: x / 0 is NULL
            else { if (b == 0.0) { r = 0; n = 1; } else r = fmod(a, b); }
            st[sp - 1] = __double_as_longlong(r); nulls = (nulls & ~(3u << (sp - 1))) | (n << (sp - 1));
            break;
        }
        case OP_NEG_I: st[sp - 1] = (long long)(0ull - (unsigned long long)st[sp - 1]); break;
        case OP_NEG_F: st[sp - 1] = __double_as_longlong(-__longlong_as_double(st[sp - 1])); break;
        case OP_LUT: {
            long long code = st[sp - 1];
            const ProgLut l = v.luts[arg];
            st[sp - 1] = (code >= 0 && code < (long long)l.size) ? l.bits[code] : 0;
            break;
        }
        case OP_SELECT: {
            bool cn = (nulls >> (sp - 3)) & 1u; bool c = !cn && st[sp - 3] != 0;
            long long r = c ? st[sp - 2] : st[sp - 1];
            uint32_t n = c ? (nulls >> (sp - 2)) & 1u : (nulls >> (sp - 1)) & 1u;
            sp -= 2; st[sp - 1] = r; nulls = (nulls & ~(7u << (sp - 1))) | (n << (sp - 1));
            break;
        }
        case OP_HOUR: {
            long long us = st[sp - 1];
            long long day_us = 86400000000ll;
            long long r = us % day_us; if (r < 0) r += day_us;
            st[sp - 1] = r / 3600000000ll; break;
        }
        case OP_TRUNC_SEC: {
            long long us = st[sp - 1];
            long long r = us % 1000000ll; if (r < 0) r += 1000000ll;
            st[sp - 1] = us - r; break;
        }
        case OP_TS_TO_SEC: {
            long long us = st[sp - 1];
            long long r = us % 1000000ll; if (r < 0) r += 1000000ll;
            st[sp - 1] = (us - r) / 1000000ll; break;
        }
        default: out = 0; out_null = true; return;
        }
    }
}

