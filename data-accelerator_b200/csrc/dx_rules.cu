// dx_rules.cu — K2 for rule sets: bucketed, bit-sliced predicate evaluation.
//
// What it replaces: Spark evaluating R `IF(cond_r, MAP(...), NULL)` expressions per row for the
// rules CodegenRules compiles into SQL (reference: DataX.Flow.CodegenRules/Engine.cs:232-266,
// executed by spark.sql at CommonProcessorFactory.scala:271; rule grammar from the GUI:
// Website/Packages/datax-pipeline/src/modules/flowDefinition/flowHelpers.js:93-217).
//
// A naive evaluator costs O(R x leaves) operations per event and is ALU-bound by ~100x.  Here:
//   * every leaf predicate that depends on ONE column (`col op const`, IN, BETWEEN, LIKE on a
//     dictionary column, IS [NOT] NULL) is grouped by (column, comparison domain); the column's
//     constants are sorted once on the host; per event ONE binary search per column yields a
//     bucket, and a precomputed table row gives the truth of all that column's leaves at once;
//   * table rows are laid out in RULE ORDER per leaf slot (bit r of slot j = truth of rule r's
//     j-th distinct leaf), so OR-ing the rows of all columns assembles R-bit vectors directly in
//     output order — no bit permutation;
//   * each rule's AND/OR/NOT tree is put in disjunctive normal form over its (<= 4) slots; rules
//     with the same DNF shape form a class and are combined with 32-rules-per-instruction word
//     operations under the class's membership mask.  Three-valued logic is exact: tables hold a
//     TRUE plane and (only if some rule negates) a FALSE plane; NULL inputs select a bucket whose
//     rows are neither.
//   * anything that is not a one-column predicate (col-vs-col, arithmetic, IF ...) becomes a
//     "generic leaf" evaluated by the bytecode interpreter inside the same kernel; a rule that
//     does not fit (more than 4 distinct leaves, DNF blow-up) becomes one generic leaf as a whole.
#include "dx_rules.h"
#include "dx_expr_eval.cuh"
#include <cmath>
#include <array>

using namespace dxsql;

#define RB_MAX_SPACES 32
#define RB_MAX_CLASSES 16
#define RB_MAX_TERMS 8
#define RB_JMAX 4
#define RB_WG 8                      // flag words handled per pass (256 rules)
#define RB_SMEM_TABLE_BYTES (160 * 1024)
#ifndef K2_MIN_CTAS
#define K2_MIN_CTAS 4
#endif

enum { DOM_I64 = 0, DOM_F64 = 1, DOM_F64_FROM_I64 = 2, DOM_CODE = 3, DOM_GENERIC = 4 };

struct RSpaceDev {
    ProgCol col;
    int32_t domain;
    int32_t n_consts;        // numeric domains
    int32_t const_off;       // index into consts[]
    int32_t lut_off, lut_size, other_bucket;   // DOM_CODE: code -> bucket
    int32_t null_bucket;
    int32_t row_base;        // first table row of this space
    uint32_t slot_mask;      // slots (bits 0..J-1) some rule places a leaf of this space in
    uint32_t prog_start;     // DOM_GENERIC
    int32_t pad;             // numeric domains: largest power of two <= n_consts (binary-search first step)
};
struct RClassDev {
    int32_t n_terms;
    uint8_t use[RB_MAX_TERMS], neg[RB_MAX_TERMS];
    int32_t mask_off;        // index into class_masks[] (Wp words)
    int32_t shape;           // 0 = generic DNF loop; 1 a, 2 a&b, 3 a|b, 4 a&(b|c), 5 a&b&c (no negation)
    int32_t pad;             // keeps sizeof a multiple of 8: the shared-memory arrays behind it hold 64-bit words
};
static_assert(sizeof(RSpaceDev) % 8 == 0 && sizeof(RClassDev) % 8 == 0, "shared layout needs 8-byte multiples");
struct RulesDev {
    const RSpaceDev* spaces; int32_t n_spaces;
    const RClassDev* classes; int32_t n_classes;
    const long long* consts; int32_t n_consts;
    const int32_t* luts; int32_t n_luts;
    const uint32_t* class_masks;
    const uint32_t* table_t; const uint32_t* table_f;   // [rows][J][Wp]
    int32_t total_rows, J, Wp, words;
    int32_t table_in_smem;
    const void* shared_vmask; int32_t shared_vwidth;   // validity mask array shared by most spaces (loaded once per row)
    int32_t staged, stage_off, stage_buf_bytes;         // column tiles staged in shared memory by bulk copies (see K2_TILE)
    // generic leaves
    const uint32_t* code; const long long* gconsts; const ProgCol* gcols; const ProgLut* gluts; int32_t n_gcols;
};

// branch-free lower bound: number of constants < v, in ceil(log2(m+1)) fixed steps
static __device__ __forceinline__ int bucket_i64(const long long* c, int m, int top, long long v) {
    int lo = 0;
    for (int step = top; step; step >>= 1) { const int idx = lo + step; if (idx <= m && c[idx - 1] < v) lo = idx; }
    return 2 * lo + ((lo < m && c[lo] == v) ? 1 : 0);
}
static __device__ __forceinline__ int bucket_f64(const long long* c, int m, int top, double v) {
    if (v != v) return 2 * m;                                   // NaN is greater than every constant
    int lo = 0;
    for (int step = top; step; step >>= 1) { const int idx = lo + step; if (idx <= m && __longlong_as_double(c[idx - 1]) < v) lo = idx; }
    return 2 * lo + ((lo < m && __longlong_as_double(c[lo]) == v) ? 1 : 0);
}

// a leaf the bucket planner does not take (arbitrary expression): the bytecode interpreter, kept out of line
static __device__ __noinline__ int generic_bucket(const EvalView& gv, uint32_t prog_start, long long row) {
    long long val; bool isnull;
    eval_program(gv, prog_start, row, val, isnull);
    return isnull ? 2 : (val != 0 ? 1 : 0);
}

// ---- bulk-copy staging of the column tiles (STAGED kernels) ----
#define K2_TILE 256                  // rows per tile = threads per CTA
#define K2_SLOT 2048                 // bytes per staged column slot (256 rows x 8 bytes)
#ifndef K2_NBUF
#define K2_NBUF 1                    // staged tiles per CTA: with one, the resident CTAs of an SM cover each other's copies
#endif
static __device__ __forceinline__ uint32_t k2_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
static __device__ __forceinline__ void k2_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(k2_smem_u32(bar)), "r"(count) : "memory");
}
static __device__ __forceinline__ void k2_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(k2_smem_u32(bar)), "r"(bytes) : "memory");
}
static __device__ __forceinline__ void k2_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(k2_smem_u32(dst)), "l"(src), "r"(bytes), "r"(k2_smem_u32(bar)) : "memory");
}
static __device__ __forceinline__ bool k2_mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(k2_smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}

template <int J, bool NEEDF, int NSMAX, bool STAGED>
__global__ void __launch_bounds__(256, K2_MIN_CTAS) k2_rules_bucket(const __grid_constant__ RulesDev D, long long n_rows, const long long* sel, uint32_t* flags) {
    extern __shared__ __align__(16) uint8_t rsm[];
    // shared layout: spaces | classes | class masks | consts | luts | [table_t | table_f]
    RSpaceDev* s_spaces = reinterpret_cast<RSpaceDev*>(rsm);
    RClassDev* s_classes = reinterpret_cast<RClassDev*>(s_spaces + D.n_spaces);
    uint32_t* s_cmask = reinterpret_cast<uint32_t*>(s_classes + D.n_classes);
    long long* s_consts = reinterpret_cast<long long*>(s_cmask + ((D.n_classes * D.Wp + 1) & ~1));
    int32_t* s_luts = reinterpret_cast<int32_t*>(s_consts + D.n_consts);
    uint32_t* s_tab = reinterpret_cast<uint32_t*>((reinterpret_cast<uintptr_t>(s_luts + ((D.n_luts + 3) & ~3)) + 15) & ~uintptr_t(15));
    for (int i = threadIdx.x; i < D.n_spaces; i += blockDim.x) s_spaces[i] = D.spaces[i];
    for (int i = threadIdx.x; i < D.n_classes; i += blockDim.x) s_classes[i] = D.classes[i];
    for (int i = threadIdx.x; i < D.n_classes * D.Wp; i += blockDim.x) s_cmask[i] = D.class_masks[i];
    for (int i = threadIdx.x; i < D.n_consts; i += blockDim.x) s_consts[i] = D.consts[i];
    for (int i = threadIdx.x; i < D.n_luts; i += blockDim.x) s_luts[i] = D.luts[i];
    const int row_words = J * D.Wp;
    const long long tab_words = (long long)D.total_rows * row_words;
    const uint32_t* tab_t = D.table_t; const uint32_t* tab_f = D.table_f;
    if (D.table_in_smem) {
        const uint4* src = reinterpret_cast<const uint4*>(D.table_t); uint4* dst = reinterpret_cast<uint4*>(s_tab);
        for (long long i = threadIdx.x; i < tab_words / 4; i += blockDim.x) dst[i] = src[i];
        tab_t = s_tab;
        if (NEEDF) {
            const uint4* srcf = reinterpret_cast<const uint4*>(D.table_f); uint4* dstf = reinterpret_cast<uint4*>(s_tab + tab_words);
            for (long long i = threadIdx.x; i < tab_words / 4; i += blockDim.x) dstf[i] = srcf[i];
            tab_f = s_tab + tab_words;
        }
    }
    __syncthreads();
    EvalView gv{D.code, D.gconsts, D.gcols, D.gluts};
    const int n_wg = D.Wp / RB_WG;
    const int ns = D.n_spaces;
    // STAGED: the CTA walks tiles of K2_TILE rows; one thread moves every column slice of the next tile (and the shared
    // validity bytes) into shared memory with bulk copies while the current tile is evaluated, so the column reads of
    // phase A are shared-memory loads instead of global loads each warp waits ~1 us for.  Full tiles only: the last,
    // ragged tile (bulk sizes must be multiples of 16 bytes) reads global memory directly.
    __shared__ uint64_t s_bar[2];
    uint8_t* s_stage = rsm + D.stage_off;
    const long long n_tiles = (n_rows + K2_TILE - 1) / K2_TILE;
    auto issue_tile = [&](long long t, int buf) {          // thread 0
        const long long row0 = t * K2_TILE;
        if (row0 + K2_TILE > n_rows) return;               // ragged tile: not staged
        uint8_t* dst = s_stage + (size_t)buf * D.stage_buf_bytes;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        uint32_t total = D.shared_vmask ? (uint32_t)(K2_TILE * D.shared_vwidth) : 0u;
        for (int s = 0; s < D.n_spaces; s++) total += (uint32_t)(K2_TILE * s_spaces[s].col.width);
        k2_mbar_expect_tx(&s_bar[buf], total);
        for (int s = 0; s < D.n_spaces; s++) {
            const int w = s_spaces[s].col.width;
            k2_bulk_g2s(dst + s * K2_SLOT, reinterpret_cast<const uint8_t*>(s_spaces[s].col.data) + row0 * w, (uint32_t)(K2_TILE * w), &s_bar[buf]);
        }
        if (D.shared_vmask)
            k2_bulk_g2s(dst + D.n_spaces * K2_SLOT, reinterpret_cast<const uint8_t*>(D.shared_vmask) + row0 * D.shared_vwidth,
                        (uint32_t)(K2_TILE * D.shared_vwidth), &s_bar[buf]);
    };
    if (STAGED) {
        if (threadIdx.x == 0) { k2_mbar_init(&s_bar[0], 1); k2_mbar_init(&s_bar[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        __syncthreads();
        if (threadIdx.x == 0 && blockIdx.x < n_tiles) issue_tile(blockIdx.x, 0);
    }
    uint32_t it = 0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it++) {
        const long long i = tile * K2_TILE + threadIdx.x;
        const bool full = STAGED && tile * K2_TILE + K2_TILE <= n_rows;
        const uint32_t sbuf = it % K2_NBUF;
        const uint8_t* stg = s_stage + (size_t)sbuf * D.stage_buf_bytes;
        if (STAGED) {
            if (K2_NBUF == 2 && threadIdx.x == 0 && tile + gridDim.x < n_tiles) issue_tile(tile + gridDim.x, (int)((it + 1) & 1u));
            if (full) { while (!k2_mbar_try_wait(&s_bar[sbuf], (it / K2_NBUF) & 1u)) {} }
        }
        if (i < n_rows) {
        const long long row = sel ? sel[i] : i;
        uint8_t* stw = s_stage + (size_t)sbuf * D.stage_buf_bytes;
        if (STAGED && !full) {
            // ragged last tile: every thread puts its own row's values where the bulk copies would have (it reads only those back)
            for (int s = 0; s < ns; s++) {
                const RSpaceDev& sp = s_spaces[s];
                if (sp.domain == DOM_CODE) reinterpret_cast<int32_t*>(stw + s * K2_SLOT)[threadIdx.x] = reinterpret_cast<const int32_t*>(sp.col.data)[row];
                else reinterpret_cast<long long*>(stw + s * K2_SLOT)[threadIdx.x] = reinterpret_cast<const long long*>(sp.col.data)[row];
            }
            if (D.shared_vmask) {
                if (D.shared_vwidth == 1) stw[ns * K2_SLOT + threadIdx.x] = reinterpret_cast<const uint8_t*>(D.shared_vmask)[row];
                else reinterpret_cast<uint32_t*>(stw + ns * K2_SLOT)[threadIdx.x] = reinterpret_cast<const uint32_t*>(D.shared_vmask)[row];
            }
        }
        uint32_t shared_mask = 0xFFFFFFFFu;
        if (D.shared_vmask) {
            if (STAGED) shared_mask = D.shared_vwidth == 1 ? stg[ns * K2_SLOT + threadIdx.x] : reinterpret_cast<const uint32_t*>(stg + ns * K2_SLOT)[threadIdx.x];
            else shared_mask = D.shared_vwidth == 1 ? reinterpret_cast<const uint8_t*>(D.shared_vmask)[row] : reinterpret_cast<const uint32_t*>(D.shared_vmask)[row];
        }
        // phase A (direct kernels): every column value this rule set looks at, issued back to back (memory-level parallelism);
        // the staged kernels read the values from the shared-memory tile where they are needed
        long long vv[STAGED ? 1 : NSMAX];
        if (!STAGED) {
#pragma unroll
            for (int s = 0; s < NSMAX; s++) {
                if (s >= ns) break;
                const RSpaceDev& sp = s_spaces[s];
                long long v = 0;
                if (sp.domain == DOM_CODE) v = reinterpret_cast<const int32_t*>(sp.col.data)[row];
                else if (sp.domain != DOM_GENERIC) v = reinterpret_cast<const long long*>(sp.col.data)[row];
                vv[s] = v;
            }
        }
        // phase B: one bucket per space -> first table word of that bucket's row
        auto bucket_row = [&](int s) -> int {
            const RSpaceDev& sp = s_spaces[s];
            int b;
            if (sp.domain == DOM_GENERIC) b = generic_bucket(gv, sp.prog_start, row);
            else {
                bool valid = true;
                if (sp.col.vmask) {
                    uint32_t m = shared_mask;
                    if (!STAGED && (!D.shared_vmask || sp.col.vmask != D.shared_vmask))
                        m = sp.col.vwidth == 1 ? reinterpret_cast<const uint8_t*>(sp.col.vmask)[row] : reinterpret_cast<const uint32_t*>(sp.col.vmask)[row];
                    valid = (m >> sp.col.vbit) & 1u;
                }
                long long v;
                if (!STAGED) v = vv[STAGED ? 0 : s];
                else if (sp.domain == DOM_CODE) v = reinterpret_cast<const int32_t*>(stg + s * K2_SLOT)[threadIdx.x];
                else v = reinterpret_cast<const long long*>(stg + s * K2_SLOT)[threadIdx.x];
                if (!valid) b = sp.null_bucket;
                else if (sp.domain == DOM_CODE) b = (v >= 0 && v < sp.lut_size) ? s_luts[sp.lut_off + (int)v] : sp.other_bucket;
                else if (sp.domain == DOM_I64) b = bucket_i64(s_consts + sp.const_off, sp.n_consts, sp.pad, v);
                else if (sp.domain == DOM_F64) b = bucket_f64(s_consts + sp.const_off, sp.n_consts, sp.pad, __longlong_as_double(v));
                else b = bucket_f64(s_consts + sp.const_off, sp.n_consts, sp.pad, (double)v);
            }
            return (sp.row_base + b) * row_words;
        };
        uint32_t T[J][RB_WG], F[NEEDF ? J : 1][RB_WG];
        // phase C: OR the bucket's truth-table row into the slots the space feeds
        auto or_rows = [&](int s, int base) {
            const uint32_t sm = s_spaces[s].slot_mask;
#pragma unroll
            for (int j = 0; j < J; j++) {
                if ((sm >> j) & 1u) {
                    const uint4 a = *reinterpret_cast<const uint4*>(tab_t + base + j * D.Wp);
                    const uint4 b4 = *reinterpret_cast<const uint4*>(tab_t + base + j * D.Wp + 4);
                    T[j][0] |= a.x; T[j][1] |= a.y; T[j][2] |= a.z; T[j][3] |= a.w;
                    T[j][4] |= b4.x; T[j][5] |= b4.y; T[j][6] |= b4.z; T[j][7] |= b4.w;
                    if (NEEDF) {
                        const uint4 c = *reinterpret_cast<const uint4*>(tab_f + base + j * D.Wp);
                        const uint4 d = *reinterpret_cast<const uint4*>(tab_f + base + j * D.Wp + 4);
                        F[j][0] |= c.x; F[j][1] |= c.y; F[j][2] |= c.z; F[j][3] |= c.w;
                        F[j][4] |= d.x; F[j][5] |= d.y; F[j][6] |= d.z; F[j][7] |= d.w;
                    }
                }
            }
        };
        auto clear_slots = [&]() {
#pragma unroll
            for (int j = 0; j < J; j++)
#pragma unroll
                for (int w = 0; w < RB_WG; w++) { T[j][w] = 0; if (NEEDF) F[j][w] = 0; }
        };
        // phase D: class combination and the flag words of word group wg
        auto combine_store = [&](int wg) {
            uint32_t out[RB_WG];
#pragma unroll
            for (int w = 0; w < RB_WG; w++) out[w] = 0;
            for (int k = 0; k < D.n_classes; k++) {
                const RClassDev& cl = s_classes[k];
                const uint32_t* cm = s_cmask + cl.mask_off + wg * RB_WG;
                // the common rule shapes as straight-line word operations; everything else through the DNF loop
                switch (cl.shape) {
                case 1:   // a
#pragma unroll
                    for (int w = 0; w < RB_WG; w++) out[w] |= cm[w] & T[0][w];
                    break;
                case 2:   // a AND b
                    if (J >= 2) {
#pragma unroll
                        for (int w = 0; w < RB_WG; w++) out[w] |= cm[w] & T[0][w] & T[J >= 2 ? 1 : 0][w];
                    }
                    break;
                case 3:   // a OR b
                    if (J >= 2) {
#pragma unroll
                        for (int w = 0; w < RB_WG; w++) out[w] |= cm[w] & (T[0][w] | T[J >= 2 ? 1 : 0][w]);
                    }
                    break;
                case 4:   // a AND (b OR c)
                    if (J >= 3) {
#pragma unroll
                        for (int w = 0; w < RB_WG; w++) out[w] |= cm[w] & T[0][w] & (T[J >= 3 ? 1 : 0][w] | T[J >= 3 ? 2 : 0][w]);
                    }
                    break;
                case 5:   // a AND b AND c
                    if (J >= 3) {
#pragma unroll
                        for (int w = 0; w < RB_WG; w++) out[w] |= cm[w] & T[0][w] & T[J >= 3 ? 1 : 0][w] & T[J >= 3 ? 2 : 0][w];
                    }
                    break;
                default:
                    for (int t = 0; t < cl.n_terms; t++) {
                        const uint32_t use = cl.use[t], neg = cl.neg[t];
                        uint32_t acc[RB_WG];
#pragma unroll
                        for (int w = 0; w < RB_WG; w++) acc[w] = cm[w];
#pragma unroll
                        for (int j = 0; j < J; j++) {
                            if ((use >> j) & 1u) {
                                if (NEEDF && ((neg >> j) & 1u)) {
#pragma unroll
                                    for (int w = 0; w < RB_WG; w++) acc[w] &= F[j][w];
                                } else {
#pragma unroll
                                    for (int w = 0; w < RB_WG; w++) acc[w] &= T[j][w];
                                }
                            }
                        }
#pragma unroll
                        for (int w = 0; w < RB_WG; w++) out[w] |= acc[w];
                    }
                }
            }
#pragma unroll
            for (int w = 0; w < RB_WG; w++) {
                const int gw = wg * RB_WG + w;
                if (gw < D.words) flags[(long long)gw * n_rows + i] = out[w];
            }
        };
        if (n_wg == 1) {
            // up to 256 rules: bucket search and row OR per space back to back, nothing kept per space
            clear_slots();
#pragma unroll
            for (int s = 0; s < NSMAX; s++) {
                if (s >= ns) break;
                or_rows(s, bucket_row(s));
            }
            combine_store(0);
        } else {
            int brow[NSMAX];
#pragma unroll
            for (int s = 0; s < NSMAX; s++) {
                if (s >= ns) break;
                brow[s] = bucket_row(s);
            }
            for (int wg = 0; wg < n_wg; wg++) {
                clear_slots();
#pragma unroll
                for (int s = 0; s < NSMAX; s++) {
                    if (s >= ns) break;
                    or_rows(s, brow[s] + wg * RB_WG);
                }
                combine_store(wg);
            }
        }
        }
        if (STAGED) {
            __syncthreads();        // the tile's buffer is about to be refilled: every reader must be done
            if (K2_NBUF == 1 && threadIdx.x == 0 && tile + gridDim.x < n_tiles) issue_tile(tile + gridDim.x, 0);
        }
    }
}

// =======================================================================================
// k2_rules_fast — the same bucketed, bit-sliced evaluation with everything a rule set has in common hoisted out of the row
// loop.  The class kernel above costs ~1500 thread-instructions per row for 200 rules (ncu, round 1: 47 warp-instructions per
// event, a third of them descriptor loads, predicate tests and the per-class switch).  Here
//   * the per-space descriptors ride in the kernel parameter block, read at compile-time offsets of the constant bank;
//   * constants are padded to 2^k - 1 entries, so a bucket search is k steps of {LDS, compare, predicated add} with no bound
//     test (integer columns whose constants fit 32 bits search 4-byte keys after clamping the value);
//   * every rule is put in ONE normal form, out = (a1 AND a2 AND ...) AND (o1 OR o2 OR ...), either part optional (a; a AND b;
//     a OR b; a AND (b OR c); x IN (...) AND y; ...).  Each space keeps at most two R-bit vectors per bucket: A, the conjunction
//     of the rule's AND-leaves that live in this space (all ones where the rule has none), and O, the disjunction of its
//     OR-leaves here (zero where it has none) — leaves of one rule in the same column (BETWEEN, IN lists) are folded into the
//     table.  A row is then out = AND_s A_s[bucket_s] & (OR_s O_s[bucket_s] | no-OR-group bias): at most two table reads per
//     referenced column and one LOP3 per flag word at the end, whatever the number of rule shapes;
//   * table entries are padded to an odd number of 16-byte slots, so random buckets spread over all shared-memory banks
//     (entries of 128 bytes put every lane's first read on the same four banks: 8-way conflicts);
//   * flag words are a template parameter: 200 rules work on 7 words, not on a padded 8.
// Rule sets with NOT over a leaf, more than four distinct leaves, DNF shapes outside the normal form, LIKE over a growing dictionary,
// computed leaves or more than 256 rules keep the class kernel.  Monotone formulas need only the TRUE plane for exact
// three-valued logic: TRUE(a AND b) = TRUE(a) & TRUE(b), TRUE(a OR b) = TRUE(a) | TRUE(b).
// =======================================================================================
#define RF_MAX_SPACES 8
#ifndef RF_NT
#define RF_NT 256                    // rows per tile = threads per CTA
#endif
#ifndef RF_NBUF
#define RF_NBUF 2                    // staged tiles per CTA
#endif
#ifndef RF_MIN_CTAS
#define RF_MIN_CTAS 4
#endif
enum { RF_F64 = 0, RF_F64_FROM_I64 = 1, RF_I32 = 2, RF_I64 = 3, RF_CODE = 4, RF_IDX = 5 };
struct RFSpace {
    const void* data;
    int32_t domain, width;
    int32_t m;               // real constants
    int32_t top;             // RF_I64: first search step in BYTES of the padded constant array; 0: no constants
    int32_t const_off;       // byte offset in the image of: the 2^KS 32-bit search keys (RF_F64*, RF_I32), the padded int64
                             // constants (RF_I64), the value -> bucket table of uint16 (RF_IDX)
    int32_t exact_off;       // RF_F64*: the m constants as doubles, read only when a value's key equals a constant's
    int32_t idx_span;        // RF_IDX: last index of the table (values are clamped to [idx_base, idx_base + idx_span])
    long long idx_base;
    int32_t lut_off, lut_size, other_bucket;      // RF_CODE: code -> bucket
    int32_t null_bucket;
    int32_t tab_off;         // byte offset of this space's truth table in the image
    int32_t row_bytes;       // bytes per bucket: the A and / or O vectors (WS words each), padded to an odd number of 16-byte slots
    int32_t role_mask;       // bit 0: has an A vector, bit 1: has an O vector
    uint32_t vbitmask;       // this column's bit in the shared validity mask; 0: never NULL
    int32_t stage_off;       // byte offset of the column's slot in a stage buffer
};
struct RFDev {
    RFSpace sp[RF_MAX_SPACES];
    int32_t ns, words, image_bytes, stage_bytes;
    int32_t search_steps, pad0;              // steps of the key search the instantiation that runs makes (6 or 8)
    const uint8_t* image;
    const void* vmask; int32_t vwidth, vmask_stage_off;
    uint32_t bias_o[8];      // rules without an OR group
};

template <int W, int KS, bool STAGED>
__global__ void __launch_bounds__(RF_NT, RF_MIN_CTAS) k2_rules_fast(const __grid_constant__ RFDev D, long long n_rows, const long long* sel, uint32_t* flags) {
    constexpr int WS = W <= 1 ? 1 : (W <= 2 ? 2 : (W <= 4 ? 4 : 8));      // words a (row, role) entry occupies
    extern __shared__ __align__(128) uint8_t fsm[];
    __shared__ uint64_t f_bar[RF_NBUF];
    const int tid = threadIdx.x;
    {
        const uint4* src = reinterpret_cast<const uint4*>(D.image); uint4* dst = reinterpret_cast<uint4*>(fsm);
        for (int i = tid; i < (D.image_bytes >> 4); i += RF_NT) dst[i] = src[i];
    }
    uint8_t* const stage0 = fsm + ((D.image_bytes + 127) & ~127);
    const long long n_tiles = (n_rows + RF_NT - 1) / RF_NT;
    auto issue_tile = [&](long long t, int buf) {          // thread 0
        const long long row0 = t * RF_NT;
        if (row0 + RF_NT > n_rows) return;                 // ragged tile: read directly
        uint8_t* dst = stage0 + (size_t)buf * D.stage_bytes;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        uint32_t total = D.vmask ? (uint32_t)(RF_NT * D.vwidth) : 0u;
#pragma unroll
        for (int s = 0; s < RF_MAX_SPACES; s++) if (s < D.ns) total += (uint32_t)(RF_NT * D.sp[s].width);
        k2_mbar_expect_tx(&f_bar[buf], total);
#pragma unroll
        for (int s = 0; s < RF_MAX_SPACES; s++) if (s < D.ns)
            k2_bulk_g2s(dst + D.sp[s].stage_off, reinterpret_cast<const uint8_t*>(D.sp[s].data) + row0 * D.sp[s].width, (uint32_t)(RF_NT * D.sp[s].width), &f_bar[buf]);
        if (D.vmask) k2_bulk_g2s(dst + D.vmask_stage_off, reinterpret_cast<const uint8_t*>(D.vmask) + row0 * D.vwidth, (uint32_t)(RF_NT * D.vwidth), &f_bar[buf]);
    };
    if (STAGED) {
        if (tid == 0) {
            for (int k = 0; k < RF_NBUF; k++) k2_mbar_init(&f_bar[k], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    __syncthreads();
    if (STAGED && tid == 0 && blockIdx.x < n_tiles) issue_tile(blockIdx.x, 0);
    uint32_t it = 0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, it++) {
        const long long i = tile * RF_NT + tid;
        const bool full = STAGED && tile * RF_NT + RF_NT <= n_rows;
        const uint32_t sbuf = it % RF_NBUF;
        const uint8_t* stg = stage0 + (size_t)sbuf * D.stage_bytes;
        if (STAGED) {
            if (RF_NBUF > 1 && tid == 0 && tile + gridDim.x < n_tiles) issue_tile(tile + gridDim.x, (int)((it + 1) % RF_NBUF));
            if (full) { while (!k2_mbar_try_wait(&f_bar[sbuf], (it / RF_NBUF) & 1u)) {} }
        }
        if (i < n_rows) {
            const long long row = sel ? sel[i] : i;
            uint32_t vm = 0xFFFFFFFFu;
            if (D.vmask) {
                if (full) vm = D.vwidth == 1 ? stg[D.vmask_stage_off + tid] : reinterpret_cast<const uint32_t*>(stg + D.vmask_stage_off)[tid];
                else vm = D.vwidth == 1 ? reinterpret_cast<const uint8_t*>(D.vmask)[row] : reinterpret_cast<const uint32_t*>(D.vmask)[row];
            }
            uint32_t TA[W], TO[W];
#pragma unroll
            for (int w = 0; w < W; w++) { TA[w] = 0xFFFFFFFFu; TO[w] = D.bias_o[w]; }
#pragma unroll
            for (int s = 0; s < RF_MAX_SPACES; s++) {
                if (s >= D.ns) break;
                const RFSpace& sp = D.sp[s];
                long long v;
                if (sp.width == 4) v = full ? (long long)reinterpret_cast<const int32_t*>(stg + sp.stage_off)[tid] : (long long)reinterpret_cast<const int32_t*>(sp.data)[row];
                else v = full ? reinterpret_cast<const long long*>(stg + sp.stage_off)[tid] : reinterpret_cast<const long long*>(sp.data)[row];
                int b;
                const uint8_t* c = fsm + sp.const_off;
                if (sp.domain > RF_I32) {
                if (sp.domain == RF_CODE) {
                    b = (v >= 0 && v < sp.lut_size) ? reinterpret_cast<const int32_t*>(fsm + sp.lut_off)[(int)v] : sp.other_bucket;
                } else if (sp.domain == RF_IDX) {
                    // a small integer range (status codes, hours, enumerations): the bucket of every value in it, tabulated
                    long long d = v - sp.idx_base;
                    d = d < 0 ? 0 : (d > sp.idx_span ? (long long)sp.idx_span : d);
                    b = reinterpret_cast<const uint16_t*>(c)[(int)d];
                } else {      // RF_I64
                    int lo = 0;
                    for (int step = sp.top; step >= 8; step >>= 1) if (*reinterpret_cast<const long long*>(c + lo + step - 8) < v) lo += step;
                    const bool eq = lo < sp.m * 8 && *reinterpret_cast<const long long*>(c + lo) == v;
                    b = (lo >> 2) + (eq ? 1 : 0);                      // 2 * (lo / 8) + eq
                }
                } else {
                    // One search for every other numeric column: KS fully unrolled steps over 2^KS - 1 order-preserving 32-bit keys
                    // (padded with 0xFFFFFFFF, which no value's key exceeds).  An int64 value is clamped into the int32 range the
                    // constants lie strictly inside; a double is represented by the top half of its order-preserving bit pattern,
                    // which decides every comparison except against a constant with the same top half — only then are the exact
                    // doubles compared.
                    uint32_t key; double x = 0.0; bool isnan = false;
                    const bool is_int = sp.domain == RF_I32;
                    if (is_int) {
                        const int xi = v < -2147483647ll - 1 ? (-2147483647 - 1) : (v > 2147483647ll ? 2147483647 : (int)v);
                        key = (uint32_t)xi ^ 0x80000000u;
                    } else {
                        x = (sp.domain == RF_F64 ? __longlong_as_double(v) : (double)v) + 0.0;      // -0.0 -> +0.0
                        isnan = x != x;
                        const int hi = __double2hiint(x);
                        key = (uint32_t)hi ^ ((uint32_t)(hi >> 31) | 0x80000000u);
                    }
                    const uint32_t* ck = reinterpret_cast<const uint32_t*>(c);
                    uint32_t lo = 0;
#pragma unroll
                    for (int k = KS - 1; k >= 0; k--) if (ck[lo + (1u << k) - 1u] < key) lo += 1u << k;
                    bool eq = false;
                    if ((int)lo < sp.m && ck[lo] == key) {
                        if (is_int) eq = true;
                        else {
                            const double* cd = reinterpret_cast<const double*>(fsm + sp.exact_off);
                            while ((int)lo < sp.m && ck[lo] == key && cd[lo] < x) lo++;
                            eq = (int)lo < sp.m && cd[lo] == x;
                        }
                    }
                    b = (int)(2u * lo) + (eq ? 1 : 0);
                    if (isnan) b = 2 * sp.m;                           // NaN is greater than every constant
                }
                if (~vm & sp.vbitmask) b = sp.null_bucket;
                const uint8_t* tr = fsm + sp.tab_off + b * sp.row_bytes;
                uint32_t x8[8];
                auto fetch = [&]() {
                    if (WS == 8) {
                        const uint4 a = *reinterpret_cast<const uint4*>(tr), b4 = *reinterpret_cast<const uint4*>(tr + 16);
                        x8[0] = a.x; x8[1] = a.y; x8[2] = a.z; x8[3] = a.w; x8[4] = b4.x; x8[5] = b4.y; x8[6] = b4.z; x8[7] = b4.w;
                    } else if (WS == 4) {
                        const uint4 a = *reinterpret_cast<const uint4*>(tr);
                        x8[0] = a.x; x8[1] = a.y; x8[2] = a.z; x8[3] = a.w;
                    } else if (WS == 2) {
                        const uint2 a = *reinterpret_cast<const uint2*>(tr);
                        x8[0] = a.x; x8[1] = a.y;
                    } else x8[0] = *reinterpret_cast<const uint32_t*>(tr);
                    tr += WS * 4;
                };
                if (sp.role_mask & 1) {
                    fetch();
#pragma unroll
                    for (int w = 0; w < W; w++) TA[w] &= x8[w];
                }
                if (sp.role_mask & 2) {
                    fetch();
#pragma unroll
                    for (int w = 0; w < W; w++) TO[w] |= x8[w];
                }
            }
            uint32_t* op = flags + i;
#pragma unroll
            for (int w = 0; w < W; w++) { *op = TA[w] & TO[w]; op += n_rows; }
        }
        if (STAGED) {
            __syncthreads();        // the tile's buffer is refilled next: every reader must be done
            if (RF_NBUF == 1 && tid == 0 && tile + gridDim.x < n_tiles) issue_tile(tile + gridDim.x, 0);
        }
    }
}

// =======================================================================================
// host-side planner
// =======================================================================================
namespace {

struct BNode { int kind; std::vector<BNode> ch; int leaf = -1; };   // kind: 0 leaf, 1 AND, 2 OR, 3 NOT
struct Lit { int leaf; bool neg; bool operator<(const Lit& o) const { return leaf != o.leaf ? leaf < o.leaf : neg < o.neg; } bool operator==(const Lit& o) const { return leaf == o.leaf && neg == o.neg; } };
typedef std::vector<Lit> Term;

struct LeafDef {
    int space = -1;
    int kind = 0;             // 0 numeric compare, 1 code compare, 2 like, 3 is-null, 4 generic
    int op = 0;               // 0 = 1 <> 2 < 3 <= 4 > 5 >=
    long long cbits = 0;      // numeric constant (int64 or double bits) before sorting
    int cidx = -1;            // index among the space's sorted constants / literal index for codes
    int code = -1;
    bool negated = false;     // IS NOT NULL / NOT LIKE
    std::string pattern;
};
struct SpaceDef {
    int table_col = -1, domain = 0;
    std::vector<long long> consts;          // sorted distinct (numeric domains)
    std::vector<int> lit_codes;             // DOM_CODE: literal codes -> bucket index
    bool per_code = false;                  // DOM_CODE with LIKE: one bucket per dictionary code
    DxDictP dict;
    int n_buckets = 0, null_bucket = 0, other_bucket = 0, row_base = 0;
    uint32_t slot_mask = 0;
    int prog = -1;                          // DOM_GENERIC
};

struct Planner {
    dx_ctx* ctx; const BindScope& sc;
    std::vector<LeafDef> leaves; std::map<std::string, int> leaf_by_key;
    std::vector<SpaceDef> spaces; std::map<std::pair<int, int>, int> space_by_key;
    ExprPrograms G;
    std::string err;

    Planner(dx_ctx* c, const BindScope& s) : ctx(c), sc(s) {}

    int space_for(int col, int domain) {
        auto key = std::make_pair(col, domain);
        auto it = space_by_key.find(key);
        if (it != space_by_key.end()) return it->second;
        SpaceDef sp; sp.table_col = col; sp.domain = domain; sp.dict = sc.table->cols[col].dict;
        spaces.push_back(sp); space_by_key[key] = (int)spaces.size() - 1;
        return (int)spaces.size() - 1;
    }
    int any_space_for_column(int col) {
        for (size_t s = 0; s < spaces.size(); s++) if (spaces[s].table_col == col && spaces[s].domain != DOM_GENERIC) return (int)s;
        int t = sc.table->cols[col].type;
        return space_for(col, t == DX_T_STR ? DOM_CODE : (t == DX_T_F64 ? DOM_F64 : DOM_I64));
    }
    int generic_leaf(const ExprP& e) {
        std::string key = "G:" + to_string(e);
        auto it = leaf_by_key.find(key);
        if (it != leaf_by_key.end()) return it->second;
        int prog = (int)G.prog_start.size();
        if (dx_compile_expr(ctx, sc, e, &G)) { err = ctx->last_error; return -1; }
        SpaceDef sp; sp.domain = DOM_GENERIC; sp.prog = prog;
        spaces.push_back(sp);
        LeafDef l; l.kind = 4; l.space = (int)spaces.size() - 1;
        leaves.push_back(l); leaf_by_key[key] = (int)leaves.size() - 1;
        return (int)leaves.size() - 1;
    }
    // column with device data, or -1
    int data_column(const ExprP& e) {
        if (e->kind != EK::Col) return -1;
        std::string er; int ci = dx_resolve_column(sc, e->parts, &er);
        if (ci < 0) return -1;
        const DxColumn& c = sc.table->cols[ci];
        if (!c.data || c.all_null || c.type == DX_T_FLAGS || c.type == DX_T_NULL) return -1;
        return ci;
    }
    static bool numeric_literal(const ExprP& e, bool* is_dec, long long* iv, double* dv) {
        const Expr* x = e.get(); bool neg = false;
        if (x->kind == EK::Neg && (x->args[0]->kind == EK::Int || x->args[0]->kind == EK::Dec)) { neg = true; x = x->args[0].get(); }
        if (x->kind == EK::Int) { *is_dec = false; *iv = neg ? -x->ival : x->ival; *dv = (double)*iv; return true; }
        if (x->kind == EK::Dec) { *is_dec = true; *dv = neg ? -x->dval : x->dval; *iv = 0; return true; }
        return false;
    }
    int leaf_for(const ExprP& e) {
        std::string key = to_string(e);
        auto it = leaf_by_key.find(key);
        if (it != leaf_by_key.end()) return it->second;
        LeafDef l; bool ok = false;
        if (e->kind == EK::Bin) {
            static const char* ops[] = {"=", "<>", "<", "<=", ">", ">="};
            static const int flip[] = {0, 1, 4, 5, 2, 3};
            int op = -1; for (int k = 0; k < 6; k++) if (e->sval == ops[k]) op = k;
            if (op >= 0) {
                int lc = data_column(e->args[0]), rc = data_column(e->args[1]);
                const ExprP* lit = nullptr; int col = -1;
                if (lc >= 0 && rc < 0) { col = lc; lit = &e->args[1]; }
                else if (rc >= 0 && lc < 0) { col = rc; lit = &e->args[0]; op = flip[op]; }
                if (col >= 0) {
                    const DxColumn& c = sc.table->cols[col];
                    bool is_dec; long long iv; double dv;
                    if (c.type == DX_T_STR) {
                        if ((*lit)->kind == EK::Str && op <= 1 && c.dict) {
                            int32_t code;
                            if (c.dict->insert_host((*lit)->sval, &code)) { err = ctx->last_error; return -1; }
                            l.kind = 1; l.op = op; l.code = code; l.space = space_for(col, DOM_CODE); ok = true;
                        }
                    } else if (numeric_literal(*lit, &is_dec, &iv, &dv) && c.type != DX_T_TS) {
                        l.kind = 0; l.op = op;
                        if (c.type == DX_T_F64) { l.space = space_for(col, DOM_F64); memcpy(&l.cbits, &dv, 8); }
                        else if (is_dec) { l.space = space_for(col, DOM_F64_FROM_I64); memcpy(&l.cbits, &dv, 8); }
                        else { l.space = space_for(col, DOM_I64); l.cbits = iv; }
                        ok = true;
                    }
                }
            }
        } else if (e->kind == EK::IsNull) {
            int col = data_column(e->args[0]);
            if (col >= 0) { l.kind = 3; l.negated = e->negated; l.space = any_space_for_column(col); ok = true; }
        } else if (e->kind == EK::Like) {
            int col = data_column(e->args[0]);
            if (col >= 0 && sc.table->cols[col].type == DX_T_STR && sc.table->cols[col].dict && e->args[1]->kind == EK::Str) {
                if (sc.table->cols[col].dict->sync_to_host()) { err = ctx->last_error; return -1; }
                if (sc.table->cols[col].dict->host_strings.size() <= 4096) {
                    l.kind = 2; l.negated = e->negated; l.pattern = e->args[1]->sval; l.space = space_for(col, DOM_CODE);
                    spaces[l.space].per_code = true; ok = true;
                }
            }
        }
        if (!ok) return generic_leaf(e);
        leaves.push_back(l); leaf_by_key[key] = (int)leaves.size() - 1;
        return (int)leaves.size() - 1;
    }
    bool build(const ExprP& e, BNode* out) {
        if (e->kind == EK::Bin && (e->sval == "AND" || e->sval == "OR")) {
            out->kind = e->sval == "AND" ? 1 : 2; out->ch.resize(2);
            return build(e->args[0], &out->ch[0]) && build(e->args[1], &out->ch[1]);
        }
        if (e->kind == EK::Not) { out->kind = 3; out->ch.resize(1); return build(e->args[0], &out->ch[0]); }
        if (e->kind == EK::In && e->args.size() >= 2) {
            BNode orn; orn.kind = 2;
            for (size_t i = 1; i < e->args.size(); i++) {
                auto eq = std::make_shared<Expr>(); eq->kind = EK::Bin; eq->sval = "="; eq->args = {e->args[0], e->args[i]};
                BNode c; c.kind = 0; c.leaf = leaf_for(eq); if (c.leaf < 0) return false; orn.ch.push_back(c);
            }
            if (orn.ch.size() == 1) orn = orn.ch[0];
            if (e->negated) { out->kind = 3; out->ch.push_back(orn); } else *out = orn;
            return true;
        }
        if (e->kind == EK::Between) {
            auto ge = std::make_shared<Expr>(); ge->kind = EK::Bin; ge->sval = ">="; ge->args = {e->args[0], e->args[1]};
            auto le = std::make_shared<Expr>(); le->kind = EK::Bin; le->sval = "<="; le->args = {e->args[0], e->args[2]};
            BNode an; an.kind = 1; an.ch.resize(2);
            an.ch[0].kind = 0; an.ch[0].leaf = leaf_for(ge); an.ch[1].kind = 0; an.ch[1].leaf = leaf_for(le);
            if (an.ch[0].leaf < 0 || an.ch[1].leaf < 0) return false;
            if (e->negated) { out->kind = 3; out->ch.push_back(an); } else *out = an;
            return true;
        }
        out->kind = 0; out->leaf = leaf_for(e);
        return out->leaf >= 0;
    }
    // DNF of a tree with negation pushed to the leaves (valid in Kleene logic); false on blow-up
    static bool dnf(const BNode& n, bool neg, std::vector<Term>* out) {
        int kind = n.kind;
        if (kind == 3) return dnf(n.ch[0], !neg, out);
        if (kind == 0) { out->push_back(Term{Lit{n.leaf, neg}}); return true; }
        bool is_and = (kind == 1) != neg;     // De Morgan
        std::vector<std::vector<Term>> parts(n.ch.size());
        for (size_t i = 0; i < n.ch.size(); i++) if (!dnf(n.ch[i], neg, &parts[i])) return false;
        if (!is_and) { for (auto& p : parts) for (auto& t : p) out->push_back(t); }
        else {
            std::vector<Term> acc{Term{}};
            for (auto& p : parts) {
                std::vector<Term> next;
                for (auto& a : acc) for (auto& t : p) {
                    Term m = a; for (auto& l : t) if (std::find(m.begin(), m.end(), l) == m.end()) m.push_back(l);
                    next.push_back(m);
                    if (next.size() > 64) return false;
                }
                acc.swap(next);
            }
            for (auto& t : acc) out->push_back(t);
        }
        return out->size() <= 64;
    }
};

// truth of a leaf in a bucket: 1 TRUE, 0 FALSE, 2 NULL
static int leaf_truth(const LeafDef& l, const SpaceDef& sp, int b) {
    if (sp.domain == DOM_GENERIC) return b == 1 ? 1 : (b == 0 ? 0 : 2);
    if (l.kind == 3) { bool isnull = b == sp.null_bucket; return (isnull != l.negated) ? 1 : 0; }
    if (b == sp.null_bucket) return 2;
    if (l.kind == 0) {
        int pos = b >> 1, eq = b & 1, ci = l.cidx, cmp;      // cmp = sign(value - const)
        if (eq) cmp = pos == ci ? 0 : (ci > pos ? -1 : 1);
        else cmp = ci >= pos ? -1 : 1;
        switch (l.op) { case 0: return cmp == 0; case 1: return cmp != 0; case 2: return cmp < 0; case 3: return cmp <= 0; case 4: return cmp > 0; default: return cmp >= 0; }
    }
    if (l.kind == 1) {
        bool same;
        if (sp.per_code) same = (b == l.code);
        else same = (b < (int)sp.lit_codes.size() && sp.lit_codes[b] == l.code);
        return (l.op == 0) == same ? 1 : 0;
    }
    return 0;
}

// SQL LIKE on one dictionary entry (same matcher as the generic path)
static bool like_match2(const std::string& s, const std::string& p) {
    size_t n = s.size(), m = p.size();
    std::vector<size_t> cps; for (size_t i = 0; i < n; i++) if ((s[i] & 0xC0) != 0x80) cps.push_back(i);
    cps.push_back(n); size_t ns = cps.size() - 1;
    struct PT { int kind; std::string lit; }; std::vector<PT> pt;
    for (size_t i = 0; i < m;) {
        char c = p[i];
        if (c == '\\' && i + 1 < m) { i++; size_t j = i + 1; while (j < m && (p[j] & 0xC0) == 0x80) j++; pt.push_back({0, p.substr(i, j - i)}); i = j; }
        else if (c == '%') { pt.push_back({2, ""}); i++; }
        else if (c == '_') { pt.push_back({1, ""}); i++; }
        else { size_t j = i + 1; while (j < m && (p[j] & 0xC0) == 0x80) j++; pt.push_back({0, p.substr(i, j - i)}); i = j; }
    }
    size_t np = pt.size(); std::vector<char> dp((ns + 1) * (np + 1), 0);
    auto at = [&](size_t i, size_t j) -> char& { return dp[i * (np + 1) + j]; };
    at(0, 0) = 1;
    for (size_t j = 1; j <= np; j++) at(0, j) = at(0, j - 1) && pt[j - 1].kind == 2;
    for (size_t i = 1; i <= ns; i++) for (size_t j = 1; j <= np; j++) {
        const PT& t = pt[j - 1];
        if (t.kind == 2) at(i, j) = at(i, j - 1) || at(i - 1, j);
        else if (t.kind == 1) at(i, j) = at(i - 1, j - 1);
        else at(i, j) = at(i - 1, j - 1) && s.compare(cps[i - 1], cps[i] - cps[i - 1], t.lit) == 0;
    }
    return at(ns, np) != 0;
}

} // namespace

template <int J, bool NEEDF, int NSMAX, bool STAGED>
static void launch_rules(dx_ctx* ctx, const RulesDev& D, size_t smem, int64_t n_rows, const int64_t* sel, uint32_t* flags) {
    cudaFuncSetAttribute(k2_rules_bucket<J, NEEDF, NSMAX, STAGED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k2_rules_bucket<J, NEEDF, NSMAX, STAGED>, 256, smem);
    if (occ < 1) occ = 1;
    long long blocks = (n_rows + 255) / 256, cap = (long long)ctx->sm_count * occ;
    if (blocks > cap) blocks = cap;
    k2_rules_bucket<J, NEEDF, NSMAX, STAGED><<<(unsigned)blocks, 256, smem, ctx->stream>>>(D, n_rows, reinterpret_cast<const long long*>(sel), flags);
}

template <int W, int KS, bool STAGED>
static void launch_fast_w(dx_ctx* ctx, const RFDev& D, size_t smem, int64_t n_rows, const int64_t* sel, uint32_t* flags) {
    cudaFuncSetAttribute(k2_rules_fast<W, KS, STAGED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k2_rules_fast<W, KS, STAGED>, RF_NT, smem);
    if (occ < 1) occ = 1;
    long long blocks = (n_rows + RF_NT - 1) / RF_NT, cap = (long long)ctx->sm_count * occ;
    if (blocks > cap) blocks = cap;
    k2_rules_fast<W, KS, STAGED><<<(unsigned)blocks, RF_NT, smem, ctx->stream>>>(D, n_rows, reinterpret_cast<const long long*>(sel), flags);
}
// instantiated for every flag-word count from 1 to 8 and for searches of 6 or 8 steps
template <int KS, bool STAGED>
static void launch_fast_k(dx_ctx* ctx, const RFDev& D, size_t smem, int64_t n_rows, const int64_t* sel, uint32_t* flags) {
    switch (D.words) {
    case 1: launch_fast_w<1, KS, STAGED>(ctx, D, smem, n_rows, sel, flags); break;
    case 2: launch_fast_w<2, KS, STAGED>(ctx, D, smem, n_rows, sel, flags); break;
    case 3: launch_fast_w<3, KS, STAGED>(ctx, D, smem, n_rows, sel, flags); break;
    case 4: launch_fast_w<4, KS, STAGED>(ctx, D, smem, n_rows, sel, flags); break;
    case 5: launch_fast_w<5, KS, STAGED>(ctx, D, smem, n_rows, sel, flags); break;
    case 6: launch_fast_w<6, KS, STAGED>(ctx, D, smem, n_rows, sel, flags); break;
    case 7: launch_fast_w<7, KS, STAGED>(ctx, D, smem, n_rows, sel, flags); break;
    default: launch_fast_w<8, KS, STAGED>(ctx, D, smem, n_rows, sel, flags); break;
    }
}
template <bool STAGED>
static void launch_fast(dx_ctx* ctx, const RFDev& D, size_t smem, int64_t n_rows, const int64_t* sel, uint32_t* flags) {
    if (D.search_steps <= 6) launch_fast_k<6, STAGED>(ctx, D, smem, n_rows, sel, flags);
    else launch_fast_k<8, STAGED>(ctx, D, smem, n_rows, sel, flags);
}
// fills the per-batch parts of a fast plan (column pointers, validity mask, staging layout) and launches it
static int fast_launch(dx_ctx* ctx, RFDev D, const std::vector<ProgCol>& cols, int64_t n_rows, const int64_t* sel, uint32_t* flags, bool no_staging, bool* ok) {
    *ok = false;
    const void* vm = nullptr; int vw = 0;
    for (size_t k = 0; k < cols.size(); k++) {
        D.sp[k].data = cols[k].data; D.sp[k].width = cols[k].width;
        if (cols[k].width != (D.sp[k].domain == RF_CODE ? 4 : 8)) { ctx->counters["k2_fast_reject_width"]++; return DX_OK; }
        if (cols[k].vmask) {
            if (vm && cols[k].vmask != vm) { ctx->counters["k2_fast_reject_vmask"]++; return DX_OK; }          // one validity array for all referenced columns (the decoder's), or the class kernel
            vm = cols[k].vmask; vw = cols[k].vwidth; D.sp[k].vbitmask = 1u << cols[k].vbit;
        } else D.sp[k].vbitmask = 0;
    }
    if (vm && vw != 1 && vw != 4) return DX_OK;
    D.vmask = vm; D.vwidth = vw;
    bool staged = !no_staging && !sel && n_rows >= RF_NT;
    int off = 0;
    for (size_t k = 0; k < cols.size(); k++) {
        if (reinterpret_cast<uintptr_t>(cols[k].data) & 15) staged = false;
        D.sp[k].stage_off = off; off += RF_NT * cols[k].width;
    }
    if (vm && (reinterpret_cast<uintptr_t>(vm) & 15)) staged = false;
    D.vmask_stage_off = off; off += vm ? RF_NT * vw : 0;
    D.stage_bytes = (off + 127) & ~127;
    const size_t smem = (size_t)((D.image_bytes + 127) & ~127) + (staged ? (size_t)RF_NBUF * D.stage_bytes : 0);
    if (smem > 110 * 1024) { ctx->counters["k2_fast_reject_smem"]++; return DX_OK; }
    StageTimer tm(ctx, "k2_rules_bucket");
    if (staged) launch_fast<true>(ctx, D, smem, n_rows, sel, flags); else launch_fast<false>(ctx, D, smem, n_rows, sel, flags);
    ctx->launches["k2_rules_fast"]++;
    ctx->counters["k2_rules_fast_launches"]++;
    *ok = true;
    return DX_OK;
}

// A planned rule set kept across micro-batches: everything but the column pointers is a function of the statement text,
// the column types and the dictionaries' literal codes, so a steady-state batch only refreshes the pointers and launches.
struct PlanCache {
    std::vector<RSpaceDev> sdev; std::vector<int> space_col; std::vector<std::string> space_name; std::vector<int> space_type;
    std::vector<const void*> space_dict;
    DevBuf b_spaces, b_classes, b_cmask, b_consts, b_luts, b_tt, b_tf;
    RulesDev D; size_t smem = 0; int J = 1; bool needf = false; int R = 0, words = 0;
    bool fast = false; RFDev FD; DevBuf b_image;        // the normal-form plan (k2_rules_fast) when the rule set fits it
};

static void rules_shared_vmask(RulesDev& D, const std::vector<RSpaceDev>& sdev) {
    // the validity array most spaces share (the decoder writes one mask byte per row for all leaves)
    std::map<const void*, int> votes; const void* best = nullptr; int bw = 0, bc = 0;
    for (auto& d : sdev) if (d.domain != DOM_GENERIC && d.col.vmask) { int c = ++votes[d.col.vmask]; if (c > bc) { bc = c; best = d.col.vmask; bw = d.col.vwidth; } }
    D.shared_vmask = bc >= 2 ? best : nullptr; D.shared_vwidth = bw;
}

// Whether the column tiles can be staged by bulk copies: rows taken in order, at most 8 plain column spaces, every column
// (and the one validity array they share) 16-byte aligned, no space with a validity array of its own.
static bool rules_can_stage(const RulesDev& D, const std::vector<RSpaceDev>& sdev, int64_t n_rows, const int64_t* sel, bool no_staging) {
    if (sel || n_rows < K2_TILE || sdev.size() > 8 || no_staging) return false;
    for (auto& d : sdev) {
        if (d.domain == DOM_GENERIC || !d.col.data || (reinterpret_cast<uintptr_t>(d.col.data) & 15)) return false;
        if (d.col.width != (d.domain == DOM_CODE ? 4 : 8)) return false;
        if (d.col.vmask && d.col.vmask != D.shared_vmask) return false;
    }
    if (D.shared_vmask && ((reinterpret_cast<uintptr_t>(D.shared_vmask) & 15) || (D.shared_vwidth != 1 && D.shared_vwidth != 4))) return false;
    return true;
}

static void rules_launch(dx_ctx* ctx, const RulesDev& D0, const std::vector<RSpaceDev>& sdev, int J, bool needf, size_t smem, int64_t n_rows, const int64_t* sel, uint32_t* flags, bool no_staging) {
    StageTimer tm(ctx, "k2_rules_bucket");
    RulesDev D = D0;
    const bool staged = rules_can_stage(D, sdev, n_rows, sel, no_staging);
    D.staged = staged ? 1 : 0; D.stage_off = 0; D.stage_buf_bytes = 0;
    if (staged) {
        D.stage_off = (int32_t)((smem + 127) & ~size_t(127));
        D.stage_buf_bytes = (int32_t)(sdev.size() * K2_SLOT + 1024);
        smem = (size_t)D.stage_off + K2_NBUF * (size_t)D.stage_buf_bytes;
    }
#define RB_LAUNCH_NS(JJ, FF) do { if (staged) launch_rules<JJ, FF, 8, true>(ctx, D, smem, n_rows, sel, flags); \
        else if (D.n_spaces <= 8) launch_rules<JJ, FF, 8, false>(ctx, D, smem, n_rows, sel, flags); \
        else launch_rules<JJ, FF, RB_MAX_SPACES, false>(ctx, D, smem, n_rows, sel, flags); } while (0)
#define RB_LAUNCH(JJ) do { if (needf) RB_LAUNCH_NS(JJ, true); else RB_LAUNCH_NS(JJ, false); } while (0)
    if (J == 1) RB_LAUNCH(1); else if (J == 2) RB_LAUNCH(2); else if (J == 3) RB_LAUNCH(3); else RB_LAUNCH(4);
#undef RB_LAUNCH
#undef RB_LAUNCH_NS
    ctx->launches["k2_rules_bucket"]++;
}

int dx_rules_eval(dx_ctx* ctx, const BindScope& sc, const std::vector<ExprP>& conds, int64_t n_rows, const int64_t* sel,
                  uint32_t* flags, int words, bool* used_fast, std::shared_ptr<void>* cache, int mode) {
    *used_fast = false;
    if (n_rows <= 0 || conds.empty()) return DX_OK;
    const int R = (int)conds.size();
    if (cache && *cache) {
        PlanCache* pc = static_cast<PlanCache*>(cache->get());
        bool ok = pc->R == R && pc->words == words;
        for (size_t s2 = 0; ok && s2 < pc->sdev.size(); s2++) {
            const int ci = pc->space_col[s2];
            if (ci < 0 || ci >= (int)sc.table->cols.size()) { ok = false; break; }
            const DxColumn& c = sc.table->cols[ci];
            ok = c.name == pc->space_name[s2] && c.type == pc->space_type[s2] && c.data && !c.all_null && (const void*)c.dict.get() == pc->space_dict[s2];
            if (ok) { ProgCol& pcol = pc->sdev[s2].col; pcol.data = c.data->p; pcol.vmask = c.vmask ? c.vmask->p : nullptr; pcol.vwidth = c.vwidth; pcol.vbit = c.vbit; pcol.width = dx_type_width(c.type); }
        }
        if (ok && pc->fast && !(mode & DX_RULES_NO_FAST)) {
            std::vector<ProgCol> cols; for (auto& d : pc->sdev) cols.push_back(d.col);
            bool launched = false;
            int rc = fast_launch(ctx, pc->FD, cols, n_rows, sel, flags, (mode & DX_RULES_NO_STAGING) != 0, &launched); if (rc) return rc;
            if (launched) { DX_CUDA(ctx, cudaGetLastError()); *used_fast = true; return DX_OK; }
        }
        if (ok && pc->b_spaces.p) {
            DX_CUDA(ctx, cudaMemcpyAsync(pc->b_spaces.p, pc->sdev.data(), pc->sdev.size() * sizeof(RSpaceDev), cudaMemcpyHostToDevice, ctx->stream));
            rules_shared_vmask(pc->D, pc->sdev);
            rules_launch(ctx, pc->D, pc->sdev, pc->J, pc->needf, pc->smem, n_rows, sel, flags, (mode & DX_RULES_NO_STAGING) != 0);
            DX_CUDA(ctx, cudaGetLastError());
            *used_fast = true;
            return DX_OK;
        }
        cache->reset();        // schema or dictionaries changed (or the cached plan has no class tables): plan again
    }
    Planner P(ctx, sc);
    // per rule: DNF over leaf ids
    std::vector<std::vector<Term>> rule_terms(R);
    for (int r = 0; r < R; r++) {
        BNode root; std::vector<Term> terms; bool fits = false;
        size_t leaves_before = P.leaves.size(), spaces_before = P.spaces.size();
        (void)leaves_before; (void)spaces_before;
        if (P.build(conds[r], &root) && Planner::dnf(root, false, &terms)) {
            std::vector<int> distinct;
            for (auto& t : terms) for (auto& l : t) if (std::find(distinct.begin(), distinct.end(), l.leaf) == distinct.end()) distinct.push_back(l.leaf);
            fits = distinct.size() <= RB_JMAX && terms.size() <= RB_MAX_TERMS;
        }
        if (!P.err.empty()) return ctx->set_error(DX_ERR_UNSUPPORTED, P.err);
        if (!fits) {
            int g = P.generic_leaf(conds[r]);
            if (g < 0) return ctx->set_error(DX_ERR_UNSUPPORTED, P.err);
            terms.assign(1, Term{Lit{g, false}});
        }
        rule_terms[r] = terms;
    }
    if ((int)P.spaces.size() > RB_MAX_SPACES) return DX_OK;     // caller falls back to the generic kernel
    // finalise spaces: constants, buckets
    for (auto& l : P.leaves) {
        SpaceDef& sp = P.spaces[l.space];
        if (l.kind == 0) sp.consts.push_back(l.cbits);
        if (l.kind == 1 && std::find(sp.lit_codes.begin(), sp.lit_codes.end(), l.code) == sp.lit_codes.end()) sp.lit_codes.push_back(l.code);
    }
    std::vector<long long> all_consts; std::vector<int32_t> all_luts;
    std::vector<RSpaceDev> sdev(P.spaces.size());
    int total_rows = 0;
    for (size_t s = 0; s < P.spaces.size(); s++) {
        SpaceDef& sp = P.spaces[s]; RSpaceDev& d = sdev[s]; memset(&d, 0, sizeof d);
        d.domain = sp.domain;
        if (sp.domain == DOM_GENERIC) { sp.n_buckets = 3; sp.null_bucket = 2; d.prog_start = P.G.prog_start[sp.prog]; }
        else {
            const DxColumn& c = sc.table->cols[sp.table_col];
            d.col.data = c.data->p; d.col.vmask = c.vmask ? c.vmask->p : nullptr; d.col.vwidth = c.vwidth; d.col.vbit = c.vbit; d.col.width = dx_type_width(c.type);
            if (sp.domain == DOM_CODE) {
                if (sp.per_code) {
                    int n = (int)sp.dict->host_strings.size();
                    sp.other_bucket = n; sp.null_bucket = n + 1; sp.n_buckets = n + 2;
                    d.lut_off = (int)all_luts.size(); d.lut_size = n;
                    for (int k = 0; k < n; k++) all_luts.push_back(k);
                } else {
                    int n = (int)sp.lit_codes.size(), mx = -1;
                    for (int cde : sp.lit_codes) mx = std::max(mx, cde);
                    sp.other_bucket = n; sp.null_bucket = n + 1; sp.n_buckets = n + 2;
                    d.lut_off = (int)all_luts.size(); d.lut_size = mx + 1;
                    for (int k = 0; k <= mx; k++) { int b = n; for (int q = 0; q < n; q++) if (sp.lit_codes[q] == k) b = q; all_luts.push_back(b); }
                }
            } else {
                if (sp.domain == DOM_I64) std::sort(sp.consts.begin(), sp.consts.end());
                else std::sort(sp.consts.begin(), sp.consts.end(), [](long long a, long long b) { double x, y; memcpy(&x, &a, 8); memcpy(&y, &b, 8); return x < y; });
                if (sp.domain == DOM_I64) sp.consts.erase(std::unique(sp.consts.begin(), sp.consts.end()), sp.consts.end());
                else sp.consts.erase(std::unique(sp.consts.begin(), sp.consts.end(), [](long long a, long long b) { double x, y; memcpy(&x, &a, 8); memcpy(&y, &b, 8); return x == y; }), sp.consts.end());
                int m = (int)sp.consts.size();
                sp.null_bucket = 2 * m + 1; sp.n_buckets = 2 * m + 2;
                d.const_off = (int)all_consts.size(); d.n_consts = m;
                { int top = 0; if (m > 0) { top = 1; while (top * 2 <= m) top *= 2; } d.pad = top; }
                for (long long v : sp.consts) all_consts.push_back(v);
            }
        }
        d.other_bucket = sp.other_bucket; d.null_bucket = sp.null_bucket;
        sp.row_base = total_rows; d.row_base = total_rows; total_rows += sp.n_buckets;
    }
    for (auto& l : P.leaves) {
        SpaceDef& sp = P.spaces[l.space];
        if (l.kind == 0) {
            for (size_t k = 0; k < sp.consts.size(); k++) {
                bool same;
                if (sp.domain == DOM_I64) same = sp.consts[k] == l.cbits;
                else { double x, y; memcpy(&x, &sp.consts[k], 8); memcpy(&y, &l.cbits, 8); same = x == y; }
                if (same) l.cidx = (int)k;
            }
        }
    }
    // ---- the normal-form plan (k2_rules_fast), when every rule fits out = p AND s AND (q OR r) ----
    std::shared_ptr<PlanCache> fpc;
    if (!(mode & DX_RULES_NO_FAST) && words <= 8 && P.spaces.size() <= RF_MAX_SPACES) {
        bool fits = true;
        for (auto& sp : P.spaces) if (sp.domain == DOM_GENERIC || sp.per_code) fits = false;
        std::vector<std::vector<int>> and_leaves(R), or_leaves(R);
        for (int r = 0; r < R && fits; r++) {
            for (auto& t : rule_terms[r]) for (auto& l : t) if (l.neg) fits = false;
            if (!fits) break;
            // terms as sorted leaf sets, without duplicates and without absorbed terms (a OR (a AND b) = a also holds in Kleene logic)
            std::vector<std::vector<int>> terms;
            for (auto& t : rule_terms[r]) { std::vector<int> ls; for (auto& l : t) ls.push_back(l.leaf); std::sort(ls.begin(), ls.end()); ls.erase(std::unique(ls.begin(), ls.end()), ls.end()); terms.push_back(ls); }
            for (size_t a = 0; a < terms.size();) {
                bool absorbed = false;
                for (size_t b = 0; b < terms.size() && !absorbed; b++)
                    if (b != a && std::includes(terms[a].begin(), terms[a].end(), terms[b].begin(), terms[b].end()) && (terms[a] != terms[b] || b < a)) absorbed = true;
                if (absorbed) terms.erase(terms.begin() + a); else a++;
            }
            if (terms.empty()) { fits = false; break; }
            if (terms.size() == 1) { and_leaves[r] = terms[0]; continue; }
            // (common leaves) AND (one further leaf per term)
            std::vector<int> common = terms[0];
            for (size_t t = 1; t < terms.size(); t++) {
                std::vector<int> keep;
                for (int l : common) if (std::find(terms[t].begin(), terms[t].end(), l) != terms[t].end()) keep.push_back(l);
                common.swap(keep);
            }
            for (auto& t : terms) {
                if (t.size() != common.size() + 1) { fits = false; break; }
                for (int l : t) if (std::find(common.begin(), common.end(), l) == common.end()) or_leaves[r].push_back(l);
            }
            and_leaves[r] = common;
        }
        if (!fits) ctx->counters["k2_fast_reject_shape"]++;
        if (fits) {
            const int W = words, WS = W <= 1 ? 1 : (W <= 2 ? 2 : (W <= 4 ? 4 : 8));
            RFDev FD; memset(&FD, 0, sizeof FD);
            FD.ns = (int)P.spaces.size(); FD.words = W;
            for (int r = 0; r < R; r++) {
                if (or_leaves[r].empty()) FD.bias_o[r / 32] |= 1u << (r % 32);
                for (int l : and_leaves[r]) FD.sp[P.leaves[l].space].role_mask |= 1;
                for (int l : or_leaves[r]) FD.sp[P.leaves[l].space].role_mask |= 2;
            }
            std::vector<uint8_t> image;
            auto align_to = [&](size_t a) { while (image.size() % a) image.push_back(0); };
            int max_steps = 0; std::vector<int> key_spaces;
            for (size_t s2 = 0; s2 < P.spaces.size() && fits; s2++) {
                SpaceDef& sp = P.spaces[s2]; RFSpace& d = FD.sp[s2];
                d.null_bucket = sp.null_bucket; d.other_bucket = sp.other_bucket; d.vbitmask = 0;
                if (sp.domain == DOM_CODE) {
                    d.domain = RF_CODE;
                    align_to(4);
                    d.lut_off = (int)image.size(); d.lut_size = sdev[s2].lut_size;
                    for (int k = 0; k < sdev[s2].lut_size; k++) { int32_t v = all_luts[sdev[s2].lut_off + k]; const uint8_t* b = reinterpret_cast<const uint8_t*>(&v); image.insert(image.end(), b, b + 4); }
                } else {
                    const int m = (int)sp.consts.size();
                    d.m = m;
                    bool narrow = sp.domain == DOM_I64;
                    if (narrow) for (long long cv : sp.consts) if (cv <= -2147483647ll || cv >= 2147483647ll) narrow = false;
                    auto put = [&](const void* ptr, size_t n) { const uint8_t* b = reinterpret_cast<const uint8_t*>(ptr); image.insert(image.end(), b, b + n); };
                    if (sp.domain == DOM_I64 && m > 0 && sp.consts.back() - sp.consts.front() <= 1020 && 2 * m + 2 < 65536) {
                        // small range: bucket of every integer from one below the least constant to one above the greatest
                        d.domain = RF_IDX; d.idx_base = sp.consts.front() - 1; d.idx_span = (int)(sp.consts.back() + 1 - d.idx_base);
                        align_to(4);
                        d.const_off = (int)image.size();
                        for (long long u = d.idx_base; u <= d.idx_base + d.idx_span; u++) {
                            const int lo = (int)(std::lower_bound(sp.consts.begin(), sp.consts.end(), u) - sp.consts.begin());
                            const uint16_t bk = (uint16_t)(2 * lo + ((lo < m && sp.consts[lo] == u) ? 1 : 0));
                            put(&bk, 2);
                        }
                    } else if (sp.domain == DOM_I64 && !narrow) {
                        d.domain = RF_I64;
                        int Pm = 0; while (Pm < m) Pm = Pm * 2 + 1;                       // 2^k - 1 >= m
                        d.top = ((Pm + 1) / 2) * 8; if (m == 0) d.top = 0;
                        align_to(8);
                        d.const_off = (int)image.size();
                        for (int k = 0; k <= Pm; k++) { long long v = k < m ? sp.consts[k] : 0x7FFFFFFFFFFFFFFFll; put(&v, 8); }   // one extra entry: the equality probe
                    } else {
                        d.domain = sp.domain == DOM_F64 ? RF_F64 : (sp.domain == DOM_F64_FROM_I64 ? RF_F64_FROM_I64 : RF_I32);
                        int ks = 0; while ((1 << ks) - 1 < m) ks++;
                        max_steps = std::max(max_steps, ks);
                        key_spaces.push_back((int)s2);
                        if (d.domain != RF_I32) {
                            align_to(8);
                            d.exact_off = (int)image.size();
                            for (int k = 0; k < m; k++) { double cv; memcpy(&cv, &sp.consts[k], 8); cv += 0.0; put(&cv, 8); }
                        }
                    }
                }
            }
            if (max_steps > 8) { fits = false; ctx->counters["k2_fast_reject_consts"]++; }
            FD.search_steps = max_steps <= 6 ? 6 : 8;
            for (int s2 : key_spaces) {           // 2^KS search keys per space (the last one is only ever the equality probe's)
                SpaceDef& sp = P.spaces[s2]; RFSpace& d = FD.sp[s2];
                align_to(4);
                d.const_off = (int)image.size();
                for (int k = 0; k < (1 << FD.search_steps); k++) {
                    uint32_t key = 0xFFFFFFFFu;
                    if (k < d.m) {
                        if (d.domain == RF_I32) key = (uint32_t)(int32_t)sp.consts[k] ^ 0x80000000u;
                        else { double cv; memcpy(&cv, &sp.consts[k], 8); cv += 0.0; long long bits; memcpy(&bits, &cv, 8); const int hi = (int)(bits >> 32); key = (uint32_t)hi ^ ((uint32_t)(hi >> 31) | 0x80000000u); }
                    }
                    const uint8_t* b = reinterpret_cast<const uint8_t*>(&key); image.insert(image.end(), b, b + 4);
                }
            }
            if (fits) {
            // truth tables: per space and bucket the A vector (if any rule has an AND-leaf here) then the O vector, WS words each
            for (size_t s2 = 0; s2 < P.spaces.size(); s2++) {
                SpaceDef& sp = P.spaces[s2]; RFSpace& d = FD.sp[s2];
                const int n_vec = (d.role_mask & 1) + ((d.role_mask >> 1) & 1);
                int slots = (n_vec * WS * 4 + 15) / 16; if (slots > 0 && slots % 2 == 0) slots++;      // odd number of 16-byte slots
                d.row_bytes = slots * 16;
                align_to(16);
                d.tab_off = (int)image.size();
                image.resize(image.size() + (size_t)sp.n_buckets * d.row_bytes, 0);
                if (d.role_mask & 1)        // A starts as all ones: rules without an AND-leaf in this space do not constrain it
                    for (int b = 0; b < sp.n_buckets; b++) memset(image.data() + d.tab_off + (size_t)b * d.row_bytes, 0xFF, (size_t)WS * 4);
            }
            for (int r = 0; r < R; r++) {
                for (size_t s2 = 0; s2 < P.spaces.size(); s2++) {
                    const SpaceDef& sp = P.spaces[s2]; const RFSpace& d = FD.sp[s2];
                    std::vector<int> al, ol;
                    for (int l : and_leaves[r]) if (P.leaves[l].space == (int)s2) al.push_back(l);
                    for (int l : or_leaves[r]) if (P.leaves[l].space == (int)s2) ol.push_back(l);
                    for (int b = 0; b < sp.n_buckets; b++) {
                        uint8_t* row = image.data() + d.tab_off + (size_t)b * d.row_bytes;
                        if (!al.empty()) {
                            bool all_true = true; for (int l : al) if (leaf_truth(P.leaves[l], sp, b) != 1) all_true = false;
                            if (!all_true) reinterpret_cast<uint32_t*>(row)[r / 32] &= ~(1u << (r % 32));
                        }
                        if (!ol.empty()) {
                            bool any_true = false; for (int l : ol) if (leaf_truth(P.leaves[l], sp, b) == 1) any_true = true;
                            if (any_true) reinterpret_cast<uint32_t*>(row + ((d.role_mask & 1) ? WS * 4 : 0))[r / 32] |= 1u << (r % 32);
                        }
                    }
                }
            }
            // (bits past the last rule: A vectors hold ones there, O vectors and the bias zeros, so the output bits are zero)
            align_to(16);
            FD.image_bytes = (int)image.size();
            if (image.size() > 96 * 1024) ctx->counters["k2_fast_reject_image"]++;
            if (image.size() <= 96 * 1024) {
                DevBuf b_image(ctx, image.size() + 64);
                if (!b_image.p) return ctx->set_error(DX_ERR_CUDA, "rule table allocation failed");
                DX_CUDA(ctx, cudaMemcpyAsync(b_image.p, image.data(), image.size(), cudaMemcpyHostToDevice, ctx->stream));
                FD.image = b_image.as<uint8_t>();
                std::vector<ProgCol> cols; for (auto& d : sdev) cols.push_back(d.col);
                bool launched = false;
                int rc2 = fast_launch(ctx, FD, cols, n_rows, sel, flags, (mode & DX_RULES_NO_STAGING) != 0, &launched); if (rc2) return rc2;
                if (launched) {
                    DX_CUDA(ctx, cudaGetLastError());
                    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));     // `image` (host) was the source of an asynchronous copy
                    if (cache) {
                        fpc = std::make_shared<PlanCache>();
                        fpc->sdev = sdev; fpc->R = R; fpc->words = words; fpc->fast = true; fpc->FD = FD; fpc->b_image = std::move(b_image);
                        for (auto& sp : P.spaces) {
                            const DxColumn& c = sc.table->cols[sp.table_col];
                            fpc->space_col.push_back(sp.table_col); fpc->space_name.push_back(c.name); fpc->space_type.push_back(c.type); fpc->space_dict.push_back(c.dict.get());
                        }
                        *cache = fpc;
                    }
                    *used_fast = true;
                    return DX_OK;
                }
            }
            }
        }
    }
    // slots and classes
    int J = 1; bool needf = false;
    struct RulePlan { std::vector<int> slot_leaf; std::vector<std::pair<uint8_t, uint8_t>> terms; int cls = -1; };
    std::vector<RulePlan> plans(R);
    std::map<std::string, int> class_by_sig; std::vector<RClassDev> classes;
    for (int r = 0; r < R; r++) {
        RulePlan& rp = plans[r];
        for (auto& t : rule_terms[r]) {
            uint8_t use = 0, neg = 0;
            for (auto& l : t) {
                int j = (int)(std::find(rp.slot_leaf.begin(), rp.slot_leaf.end(), l.leaf) - rp.slot_leaf.begin());
                if (j == (int)rp.slot_leaf.size()) rp.slot_leaf.push_back(l.leaf);
                use |= 1u << j; if (l.neg) { neg |= 1u << j; needf = true; }
            }
            rp.terms.emplace_back(use, neg);
        }
        std::sort(rp.terms.begin(), rp.terms.end());
        rp.terms.erase(std::unique(rp.terms.begin(), rp.terms.end()), rp.terms.end());
        J = std::max(J, (int)rp.slot_leaf.size());
        std::string sig;
        for (auto& t : rp.terms) { sig += std::to_string(t.first) + "/" + std::to_string(t.second) + ";"; }
        auto it = class_by_sig.find(sig);
        if (it == class_by_sig.end()) {
            if ((int)classes.size() >= RB_MAX_CLASSES) return DX_OK;     // fall back
            RClassDev c; memset(&c, 0, sizeof c); c.n_terms = (int)rp.terms.size();
            for (size_t t = 0; t < rp.terms.size(); t++) { c.use[t] = rp.terms[t].first; c.neg[t] = rp.terms[t].second; }
            c.shape = 0;
            if (sig == "1/0;") c.shape = 1;
            else if (sig == "3/0;") c.shape = 2;
            else if (sig == "1/0;2/0;") c.shape = 3;
            else if (sig == "3/0;5/0;") c.shape = 4;
            else if (sig == "7/0;") c.shape = 5;
            classes.push_back(c); class_by_sig[sig] = (int)classes.size() - 1; rp.cls = (int)classes.size() - 1;
        } else rp.cls = it->second;
    }
    const int Wp = ((words + RB_WG - 1) / RB_WG) * RB_WG;
    std::vector<uint32_t> cmask((size_t)classes.size() * Wp, 0);
    for (size_t k = 0; k < classes.size(); k++) classes[k].mask_off = (int)k * Wp;
    for (int r = 0; r < R; r++) cmask[(size_t)plans[r].cls * Wp + r / 32] |= 1u << (r % 32);
    // truth tables in rule order per slot
    const size_t row_words = (size_t)J * Wp;
    std::vector<uint32_t> tt((size_t)total_rows * row_words, 0), tf(needf ? (size_t)total_rows * row_words : 0, 0);
    std::map<std::pair<int, int>, std::vector<uint8_t>> like_cache;   // (space, leaf) -> per-code result
    for (int r = 0; r < R; r++) {
        for (size_t j = 0; j < plans[r].slot_leaf.size(); j++) {
            const int li = plans[r].slot_leaf[j];
            const LeafDef& l = P.leaves[li]; SpaceDef& sp = P.spaces[l.space];
            sp.slot_mask |= 1u << j;
            for (int b = 0; b < sp.n_buckets; b++) {
                int tv;
                if (l.kind == 2) {
                    if (b == sp.null_bucket) tv = 2;
                    else if (b == sp.other_bucket) tv = l.negated ? 1 : 0;    // codes that appear later: rebuilt next batch before use
                    else { bool m = like_match2(sp.dict->host_strings[b], l.pattern); tv = (m != l.negated) ? 1 : 0; }
                } else tv = leaf_truth(l, sp, b);
                size_t off = ((size_t)(sp.row_base + b) * J + j) * Wp + r / 32;
                if (tv == 1) tt[off] |= 1u << (r % 32);
                else if (tv == 0 && needf) tf[off] |= 1u << (r % 32);
            }
        }
    }
    for (size_t s = 0; s < P.spaces.size(); s++) sdev[s].slot_mask = P.spaces[s].slot_mask;
    // a per-code LIKE space must see every code of this batch: the dictionary was synced when the leaf was planned,
    // i.e. after this batch was decoded, so the table covers all codes present in the columns.
    // ---- upload
    DevPrograms GD;
    if (!P.G.prog_start.empty()) { int rc = dx_upload_programs(ctx, &P.G, &GD); if (rc) return rc; }
    if (P.G.cols.size() > 32) return DX_OK;
    auto up = [&](const void* src, size_t bytes, DevBuf* dst) -> int {
        *dst = DevBuf(ctx, bytes + 64);
        if (!dst->p) return ctx->set_error(DX_ERR_CUDA, "rule table allocation failed");
        if (bytes) DX_CUDA(ctx, cudaMemcpyAsync(dst->p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        return DX_OK;
    };
    DevBuf b_spaces, b_classes, b_cmask, b_consts, b_luts, b_tt, b_tf;
    int rc;
    if ((rc = up(sdev.data(), sdev.size() * sizeof(RSpaceDev), &b_spaces))) return rc;
    if ((rc = up(classes.data(), classes.size() * sizeof(RClassDev), &b_classes))) return rc;
    if ((rc = up(cmask.data(), cmask.size() * 4, &b_cmask))) return rc;
    if ((rc = up(all_consts.data(), all_consts.size() * 8, &b_consts))) return rc;
    if ((rc = up(all_luts.data(), all_luts.size() * 4, &b_luts))) return rc;
    if ((rc = up(tt.data(), tt.size() * 4, &b_tt))) return rc;
    if ((rc = up(tf.data(), tf.size() * 4, &b_tf))) return rc;
    RulesDev D; memset(&D, 0, sizeof D);
    D.spaces = b_spaces.as<RSpaceDev>(); D.n_spaces = (int)sdev.size();
    D.classes = b_classes.as<RClassDev>(); D.n_classes = (int)classes.size();
    D.consts = b_consts.as<long long>(); D.n_consts = (int)all_consts.size();
    D.luts = b_luts.as<int32_t>(); D.n_luts = (int)all_luts.size();
    D.class_masks = b_cmask.as<uint32_t>();
    D.table_t = b_tt.as<uint32_t>(); D.table_f = b_tf.as<uint32_t>();
    D.total_rows = total_rows; D.J = J; D.Wp = Wp; D.words = words;
    rules_shared_vmask(D, sdev);
    if (!P.G.prog_start.empty()) {
        D.code = GD.code.as<uint32_t>(); D.gconsts = GD.consts.as<long long>(); D.gcols = GD.cols.as<ProgCol>(); D.gluts = GD.luts.as<ProgLut>(); D.n_gcols = GD.n_cols;
    }
    size_t fixed = sdev.size() * sizeof(RSpaceDev) + classes.size() * sizeof(RClassDev) + ((cmask.size() + 1) & ~size_t(1)) * 4 +
                   all_consts.size() * 8 + ((all_luts.size() + 3) & ~size_t(3)) * 4 + 64;
    size_t tab_bytes = tt.size() * 4 * (needf ? 2 : 1);
    D.table_in_smem = (fixed + tab_bytes <= RB_SMEM_TABLE_BYTES) ? 1 : 0;
    size_t smem = fixed + (D.table_in_smem ? tab_bytes : 0);
    if (fixed > 64 * 1024) return DX_OK;       // pathological constant count: generic kernel
    rules_launch(ctx, D, sdev, J, needf, smem, n_rows, sel, flags, (mode & DX_RULES_NO_STAGING) != 0);
    DX_CUDA(ctx, cudaGetLastError());
    bool cacheable = cache != nullptr && P.G.prog_start.empty();
    for (auto& sp : P.spaces) if (sp.per_code) cacheable = false;       // per-code LIKE tables follow the dictionary
    if (cacheable) {
        auto pc = std::make_shared<PlanCache>();
        pc->sdev = sdev; pc->D = D; pc->smem = smem; pc->J = J; pc->needf = needf; pc->R = R; pc->words = words;
        for (auto& sp : P.spaces) {
            const DxColumn& c = sc.table->cols[sp.table_col];
            pc->space_col.push_back(sp.table_col); pc->space_name.push_back(c.name); pc->space_type.push_back(c.type); pc->space_dict.push_back(c.dict.get());
        }
        pc->b_spaces = std::move(b_spaces); pc->b_classes = std::move(b_classes); pc->b_cmask = std::move(b_cmask); pc->b_consts = std::move(b_consts);
        pc->b_luts = std::move(b_luts); pc->b_tt = std::move(b_tt); pc->b_tf = std::move(b_tf);
        *cache = pc;           // the device tables live on with the plan; nothing to wait for
    } else {
        DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));     // the uploaded tables die with this scope
    }
    *used_fast = true;
    return DX_OK;
}

// =======================================================================================
// per-rule hit counts of a flag bitmap ("200-bit flags/event + per-rule counts")
// =======================================================================================
// Bit-sliced counting: a lane never looks at single bits of the words it reads.  It adds them, 32 rules at a time, into a
// vertical counter (one register per binary digit: ones, twos, fours, then five more) with carry-save adders — 7 three-input
// adders (14 LOP3) take in eight flag words — and only when the counter is about to overflow, or at the end, are the digits
// of each of the 32 rule positions put back together and summed over the warp (REDUX).  About three instructions per flag
// word, against four per BIT for the ballot form this replaces (2.5 ms per 25 M events x 7 words); it runs at memory speed.
struct RcAcc { uint32_t ones, twos, fours, c[5]; };
static __device__ __forceinline__ void rc_csa(uint32_t& hi, uint32_t& lo, uint32_t a, uint32_t b, uint32_t c) {
    lo = a ^ b ^ c; hi = (a & b) | (a & c) | (b & c);
}
static __device__ __forceinline__ void rc_ripple(RcAcc& A, uint32_t carry) {       // adds `carry` at weight 8
#pragma unroll
    for (int l = 0; l < 5; l++) { const uint32_t t = A.c[l] & carry; A.c[l] ^= carry; carry = t; }
}
static __device__ __forceinline__ void rc_add8(RcAcc& A, const uint32_t* x) {
    uint32_t ta, tb, fa, fb, ea;
    rc_csa(ta, A.ones, A.ones, x[0], x[1]);
    rc_csa(tb, A.ones, A.ones, x[2], x[3]);
    rc_csa(fa, A.twos, A.twos, ta, tb);
    rc_csa(ta, A.ones, A.ones, x[4], x[5]);
    rc_csa(tb, A.ones, A.ones, x[6], x[7]);
    rc_csa(fb, A.twos, A.twos, ta, tb);
    rc_csa(ea, A.fours, A.fours, fa, fb);
    rc_ripple(A, ea);
}
static __device__ __forceinline__ void rc_add1(RcAcc& A, uint32_t x) {
    uint32_t t = A.ones & x; A.ones ^= x; x = t;
    t = A.twos & x; A.twos ^= x; x = t;
    t = A.fours & x; A.fours ^= x; x = t;
    rc_ripple(A, x);
}
// digits -> per-rule counts, summed over the warp; lane b keeps rule position b.  The counter holds at most 255 per position.
static __device__ __forceinline__ void rc_flush(RcAcc& A, unsigned int& mine, int lane) {
#pragma unroll
    for (int b = 0; b < 32; b++) {
        const unsigned int v = ((A.ones >> b) & 1u) + (((A.twos >> b) & 1u) << 1) + (((A.fours >> b) & 1u) << 2) + (((A.c[0] >> b) & 1u) << 3) +
                               (((A.c[1] >> b) & 1u) << 4) + (((A.c[2] >> b) & 1u) << 5) + (((A.c[3] >> b) & 1u) << 6) + (((A.c[4] >> b) & 1u) << 7);
        const unsigned int tot = __reduce_add_sync(0xffffffffu, v);
        if (lane == b) mine += tot;
    }
    A.ones = A.twos = A.fours = 0;
#pragma unroll
    for (int l = 0; l < 5; l++) A.c[l] = 0;
}
__global__ void __launch_bounds__(256) k2_rule_counts(const uint32_t* flags, long long n_rows, int words, unsigned long long* counts) {
    __shared__ unsigned int s_cnt[32];
    const int lane = threadIdx.x & 31, w = blockIdx.y;
    if (threadIdx.x < 32) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    unsigned int mine = 0;                      // lane b keeps the count of bit b
    const uint32_t* col = flags + (size_t)w * n_rows;
    RcAcc A; A.ones = A.twos = A.fours = 0;
#pragma unroll
    for (int l = 0; l < 5; l++) A.c[l] = 0;
    int held = 0;                               // words added since the last flush (per lane; the same for every lane of a warp)
    // rows before the first 16-byte boundary of this word's column, then 16-byte vectors (two per step), then the tail
    long long head = (long long)(((16 - (reinterpret_cast<uintptr_t>(col) & 15)) & 15) >> 2);
    if (head > n_rows) head = n_rows;
    const long long n_vec = (n_rows - head) >> 2, tail0 = head + (n_vec << 2);
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x, gsz = (long long)gridDim.x * blockDim.x;
    if (gid < head) rc_add1(A, col[gid]);
    if (gid < n_rows - tail0) rc_add1(A, col[tail0 + gid]);
    held = 2;
    const uint4* vec = reinterpret_cast<const uint4*>(col + head);
    // every lane of a warp makes the same number of trips (the flush is a warp-wide reduction): lanes past the end add zeros
    const long long warp0 = gid - lane;
    for (long long v0 = warp0; v0 < n_vec; v0 += 2 * gsz) {
        const long long va = v0 + lane, vb = v0 + gsz + lane;
        uint4 a = make_uint4(0, 0, 0, 0), b4 = make_uint4(0, 0, 0, 0);
        if (va < n_vec) a = __ldcs(vec + va);
        if (vb < n_vec) b4 = __ldcs(vec + vb);
        const uint32_t x[8] = {a.x, a.y, a.z, a.w, b4.x, b4.y, b4.z, b4.w};
        rc_add8(A, x);
        held += 8;
        if (held > 240) { rc_flush(A, mine, lane); held = 0; }
    }
    rc_flush(A, mine, lane);
    if (mine) atomicAdd(&s_cnt[lane], mine);
    __syncthreads();
    if (threadIdx.x < 32 && s_cnt[threadIdx.x]) atomicAdd(&counts[w * 32 + threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
}

int dx_rule_counts(dx_ctx* ctx, const uint32_t* flags, int64_t n_rows, int words, int n_rules, int64_t* counts_host) {
    for (int r = 0; r < n_rules; r++) counts_host[r] = 0;
    if (n_rows <= 0 || n_rules <= 0) return DX_OK;
    DevBuf d(ctx, (size_t)words * 32 * 8 + 16);
    if (!d.p) return ctx->set_error(DX_ERR_CUDA, "rule count allocation failed");
    DX_CUDA(ctx, cudaMemsetAsync(d.p, 0, (size_t)words * 32 * 8, ctx->stream));
    long long bx = (n_rows + 8191) / 8192; if (bx > ctx->sm_count * 4) bx = ctx->sm_count * 4; if (bx < 1) bx = 1;
    {
        StageTimer tm(ctx, "k2_rule_counts");
        k2_rule_counts<<<dim3((unsigned)bx, (unsigned)words), 256, 0, ctx->stream>>>(flags, n_rows, words, d.as<unsigned long long>());
    }
    ctx->launches["k2_rule_counts"]++;
    DX_CUDA(ctx, cudaGetLastError());
    std::vector<unsigned long long> h((size_t)words * 32);
    DX_CUDA(ctx, cudaMemcpyAsync(h.data(), d.p, h.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int r = 0; r < n_rules; r++) counts_host[r] = (int64_t)h[r];
    return DX_OK;
}
