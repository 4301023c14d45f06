// dx_expr.h — expression programs: SQL expression trees compiled to a small stack bytecode that
// the K2 kernels evaluate column-wise with Spark's three-valued logic.
//
// Semantics restated from Spark SQL 2.4.5 (the engine behind spark.sql(...) at
// CommonProcessorFactory.scala:257,271 and selectExpr at :100): comparisons with NULL are NULL,
// AND/OR/NOT are Kleene, WHERE keeps only TRUE, NaN = NaN is TRUE and NaN is the greatest double,
// long-vs-double comparisons are done in double, `/` is always a double division.
#pragma once
#include "dx_common.h"
#include "dx_sql.h"

enum DxOp : uint32_t {
    OP_LOAD = 1, OP_CONST, OP_NULL, OP_I2F,
    OP_EQ_I, OP_NE_I, OP_LT_I, OP_LE_I, OP_GT_I, OP_GE_I,
    OP_EQ_F, OP_NE_F, OP_LT_F, OP_LE_F, OP_GT_F, OP_GE_F,
    OP_AND, OP_OR, OP_NOT, OP_ISNULL, OP_ISNOTNULL,
    OP_ADD_I, OP_SUB_I, OP_MUL_I, OP_MOD_I, OP_NEG_I,
    OP_ADD_F, OP_SUB_F, OP_MUL_F, OP_DIV_F, OP_MOD_F, OP_NEG_F,
    OP_LUT,          // pop dictionary code -> push bool from LUT[arg]
    OP_SELECT,       // pop else, then, cond -> cond IS TRUE ? then : else
    OP_HOUR,         // timestamp (us) -> hour of day, UTC session zone
    OP_TRUNC_SEC,    // timestamp (us) -> truncated to the second
    OP_TS_TO_SEC,    // timestamp (us) -> unix seconds (floor)
    OP_F2I,          // double -> long (CAST semantics: truncation toward zero)
    OP_END
};
#define DX_OPCODE(x) ((x) & 0xffu)
#define DX_OPARG(x) ((x) >> 8)

struct ProgCol {             // device-visible column descriptor
    const void* data;
    const void* vmask;       // nullptr: never NULL
    int32_t vwidth, vbit;
    int32_t width;           // 4 (dictionary code) or 8
    int32_t pad;
};
struct ProgLut { const uint8_t* bits; uint32_t size; uint32_t pad; };   // one byte per dictionary code

// Host-side compiled program set (one or many expressions over the same input columns)
struct ExprPrograms {
    std::vector<uint32_t> code;          // all programs back to back, each ends with OP_END
    std::vector<uint32_t> prog_start;    // offset of each program in code
    std::vector<int32_t> result_type;    // DxType of each program's result
    std::vector<long long> consts;
    std::vector<ProgCol> cols;
    std::vector<int> col_index;          // table column index behind cols[i]
    struct LutSpec { int table_col; DxDictP dict; std::string pattern; bool negated; int kind; DevBufP bits; uint32_t built_for = 0; uint32_t cap = 0; };
    std::vector<LutSpec> luts;           // kind 0 = LIKE pattern
    std::vector<DxDictP> result_dict;    // per program: dictionary of a STR result (else null)
    int max_stack = 0;
};

struct BindScope {
    const DxTable* table;
    std::vector<std::string> qualifiers;   // names this table answers to ("DataXProcessedInput", alias, "Raw")
    long long now_us = 0;                  // current_timestamp() of the statement being compiled
    int table2_offset = -1;                // joins: columns >= offset belong to the second table
};

// Resolves a dotted name against a table: full dotted match first, then with a leading
// qualifier stripped, then by unique last-part match (Spark resolves struct fields by path;
// after a SELECT the output column is named by the last part).  Returns -1 if absent/ambiguous.
int dx_resolve_column(const BindScope& sc, const std::vector<std::string>& parts, std::string* err);

// Appends one compiled expression to P.  Returns DX_OK or sets ctx error.
int dx_compile_expr(dx_ctx* ctx, const BindScope& sc, const dxsql::ExprP& e, ExprPrograms* P);

// Device-side packing of a program set.
struct DevPrograms {
    DevBuf code, consts, cols, luts, starts;
    int n_progs = 0, n_code = 0, n_consts = 0, n_cols = 0, n_luts = 0;
    std::vector<uint32_t> host_starts;   // offset of each program in code (host copy of `starts`)
};
int dx_upload_programs(dx_ctx* ctx, ExprPrograms* P, DevPrograms* D);   // also (re)builds LUTs for grown dictionaries

// kernels (dx_expr.cu)
// flags[row * words + p/32] bit p%32 = program p is TRUE on that row.  sel (optional): row indirection.
int dx_eval_flags(dx_ctx* ctx, const DevPrograms& D, int64_t n_rows, const int64_t* sel, uint32_t* flags, int words);
// value column + validity bit for program `prog`
int dx_eval_column(dx_ctx* ctx, const DevPrograms& D, int prog, int64_t n_rows, const int64_t* sel,
                   void* out, int out_width, uint32_t* vmask32, int vbit);
// gather: out[i] = in[sel[i]] for a column (+ its validity bit into vmask32)
int dx_gather_column(dx_ctx* ctx, const DxColumn& in, int64_t n_out, const int64_t* sel, void* out, uint32_t* vmask32, int vbit);
// the same for up to DX_GATHER_MAX columns of one source table that share the selection and the destination mask word: the
// selection is read once per row and the validity bits leave as one reduction
#define DX_GATHER_MAX 8
struct GatherJobs {
    int n = 0;
    const void* in[DX_GATHER_MAX]; const void* vmask[DX_GATHER_MAX]; void* out[DX_GATHER_MAX];
    uint8_t width[DX_GATHER_MAX], vwidth[DX_GATHER_MAX], vbit_in[DX_GATHER_MAX], vbit[DX_GATHER_MAX];
};
int dx_gather_columns(dx_ctx* ctx, const GatherJobs& jobs, int64_t n_out, const int64_t* sel, uint32_t* vmask32);
// row selection: indices of rows whose flag bit 0 is set -> sel (device), count to host
int dx_select_rows(dx_ctx* ctx, const uint32_t* flags, int words, int64_t n_rows, DevBufP* sel_out, int64_t* n_sel);
