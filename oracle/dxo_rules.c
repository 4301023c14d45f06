/* dxo_rules.c — CPU restatement of rule evaluation, threaded (TEST INFRASTRUCTURE / CPU BASELINE ONLY).
 *
 * What it restates: the per-row evaluation of the conditions inside
 *   SELECT *, filterNull(Array(IF(cond_r, MAP(...), NULL), ...)) AS Rules FROM DataXProcessedInput
 * (CodegenRules Engine.cs:232-266, 380-397; filterNull UdfInitializer.scala:16-18) that Spark performs row by row:
 * bit r of a row is set iff cond_r is TRUE under SQL three-valued logic (NULL -> not set).  Conditions follow the
 * designer's grammar (flowHelpers.js:93-217): column/literal comparisons joined by AND / OR / NOT with parentheses.
 * The numpy evaluator in dxo_flow.py is the oracle of record; this one exists so the CPU baseline uses every host core
 * (SURVEY.md 8(d) "one thread per host core") and is checked against the numpy evaluator in tests/test_oracle_flow.py.
 *
 * Programs are postfix; the Python side (dxo.compile_rules) resolves columns, types and string literals (to
 * dictionary codes) so the inner loop is typed.  Product code never links or loads this file. */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { R_LOAD_I = 1, R_LOAD_F, R_CONST_I, R_CONST_F, R_CONST_NULL, R_I2F_TOP, R_I2F_SECOND, R_CMP_I, R_CMP_F, R_AND, R_OR, R_NOT, R_ISNULL, R_ISNOTNULL };
enum { C_EQ = 0, C_NE, C_LT, C_LE, C_GT, C_GE };

typedef struct { int32_t op, arg; int64_t i; double f; } dxo_rop;

typedef struct {
    const dxo_rop* code; const int32_t* rule_start; int n_rules;
    const void* const* cols; const uint8_t* const* valid;
    int64_t row0, row1; uint32_t* flags; int words;
} rules_job;

/* Spark orders doubles with NaN greater than everything and NaN = NaN */
static inline int cmp_f(double a, double b, int op) {
    int an = isnan(a), bn = isnan(b), lt, eq;
    if (an || bn) { eq = an && bn; lt = !an && bn; }
    else { eq = a == b; lt = a < b; }
    switch (op) { case C_EQ: return eq; case C_NE: return !eq; case C_LT: return lt; case C_LE: return lt || eq; case C_GT: return !lt && !eq; default: return !lt; }
}
static inline int cmp_i(int64_t a, int64_t b, int op) {
    switch (op) { case C_EQ: return a == b; case C_NE: return a != b; case C_LT: return a < b; case C_LE: return a <= b; case C_GT: return a > b; default: return a >= b; }
}

static void* rules_worker(void* p) {
    rules_job* j = (rules_job*)p;
    union { int64_t i; double f; } st[64]; uint8_t nul[64];
    for (int64_t row = j->row0; row < j->row1; row++) {
        uint32_t* out = j->flags + row * j->words;
        for (int w = 0; w < j->words; w++) out[w] = 0;
        for (int r = 0; r < j->n_rules; r++) {
            int sp = 0;
            for (int32_t pc = j->rule_start[r]; pc < j->rule_start[r + 1]; pc++) {
                const dxo_rop* o = &j->code[pc];
                switch (o->op) {
                case R_LOAD_I: st[sp].i = ((const int64_t*)j->cols[o->arg])[row]; nul[sp] = j->valid[o->arg] ? !j->valid[o->arg][row] : 0; sp++; break;
                case R_LOAD_F: st[sp].f = ((const double*)j->cols[o->arg])[row]; nul[sp] = j->valid[o->arg] ? !j->valid[o->arg][row] : 0; sp++; break;
                case R_CONST_I: st[sp].i = o->i; nul[sp] = 0; sp++; break;
                case R_CONST_F: st[sp].f = o->f; nul[sp] = 0; sp++; break;
                case R_CONST_NULL: st[sp].i = 0; nul[sp] = 1; sp++; break;
                case R_I2F_TOP: st[sp - 1].f = (double)st[sp - 1].i; break;
                case R_I2F_SECOND: st[sp - 2].f = (double)st[sp - 2].i; break;
                case R_CMP_I: sp--; st[sp - 1].i = cmp_i(st[sp - 1].i, st[sp].i, o->arg); nul[sp - 1] |= nul[sp]; break;
                case R_CMP_F: sp--; st[sp - 1].i = cmp_f(st[sp - 1].f, st[sp].f, o->arg); nul[sp - 1] |= nul[sp]; break;
                case R_AND: {   /* Kleene: FALSE dominates NULL */
                    sp--; int a = !nul[sp - 1] && st[sp - 1].i, b = !nul[sp] && st[sp].i;
                    int af = !nul[sp - 1] && !st[sp - 1].i, bf = !nul[sp] && !st[sp].i;
                    if (af || bf) { st[sp - 1].i = 0; nul[sp - 1] = 0; } else if (a && b) { st[sp - 1].i = 1; nul[sp - 1] = 0; } else { st[sp - 1].i = 0; nul[sp - 1] = 1; }
                    break; }
                case R_OR: {    /* Kleene: TRUE dominates NULL */
                    sp--; int a = !nul[sp - 1] && st[sp - 1].i, b = !nul[sp] && st[sp].i;
                    int af = !nul[sp - 1] && !st[sp - 1].i, bf = !nul[sp] && !st[sp].i;
                    if (a || b) { st[sp - 1].i = 1; nul[sp - 1] = 0; } else if (af && bf) { st[sp - 1].i = 0; nul[sp - 1] = 0; } else { st[sp - 1].i = 0; nul[sp - 1] = 1; }
                    break; }
                case R_NOT: st[sp - 1].i = !st[sp - 1].i; break;
                case R_ISNULL: st[sp - 1].i = nul[sp - 1]; nul[sp - 1] = 0; break;
                case R_ISNOTNULL: st[sp - 1].i = !nul[sp - 1]; nul[sp - 1] = 0; break;
                default: break;
                }
            }
            if (sp == 1 && !nul[0] && st[0].i) out[r >> 5] |= 1u << (r & 31);
        }
    }
    return NULL;
}

/* flags: row-major, `words` uint32 per row.  Returns 0, or -1 on a malformed program. */
int dxo_eval_rules(const dxo_rop* code, const int32_t* rule_start, int n_rules, const void* const* cols, const uint8_t* const* valid,
                   int64_t n_rows, uint32_t* flags, int words, int n_threads) {
    if (n_rules > words * 32) return -1;
    for (int r = 0; r < n_rules; r++) {      /* stack discipline checked once, not per row */
        int sp = 0;
        for (int32_t pc = rule_start[r]; pc < rule_start[r + 1]; pc++) {
            switch (code[pc].op) {
            case R_LOAD_I: case R_LOAD_F: case R_CONST_I: case R_CONST_F: case R_CONST_NULL: sp++; break;
            case R_CMP_I: case R_CMP_F: case R_AND: case R_OR: if (sp < 2) return -1; sp--; break;
            case R_I2F_SECOND: if (sp < 2) return -1; break;
            case R_I2F_TOP: case R_NOT: case R_ISNULL: case R_ISNOTNULL: if (sp < 1) return -1; break;
            default: return -1;
            }
            if (sp > 60) return -1;
        }
        if (sp != 1) return -1;
    }
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256]; rules_job jobs[256];
    int64_t per = (n_rows + n_threads - 1) / n_threads;
    int started = 0;
    for (int t = 0; t < n_threads; t++) {
        int64_t a = per * t, b = a + per; if (b > n_rows) b = n_rows; if (a >= b) break;
        jobs[t] = (rules_job){code, rule_start, n_rules, cols, valid, a, b, flags, words};
        if (pthread_create(&th[t], NULL, rules_worker, &jobs[t]) != 0) { rules_worker(&jobs[t]); th[t] = 0; }
        started = t + 1;
    }
    for (int t = 0; t < started; t++) if (th[t]) pthread_join(th[t], NULL);
    return 0;
}
