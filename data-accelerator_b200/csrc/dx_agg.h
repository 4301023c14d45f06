// dx_agg.h — K3/K4 hash aggregation: GROUP BY keys with COUNT/SUM/MIN/MAX/AVG accumulators.
//
// Stands in for Spark's HashAggregateExec (partial + final) behind every "... GROUP BY ..." that
// the flow SQL sends through spark.sql at CommonProcessorFactory.scala:271 — window aggregates
// over DataXProcessedInput_<window> views, aggregate rules (CodegenRules Engine.cs:268-354),
// DISTINCT — and, for state tables, the aggregate-upsert merge (StateTableHandler.scala:99-104).
#pragma once
#include "dx_common.h"
#include "dx_expr.h"

enum DxAcc : int32_t { ACC_COUNT_STAR = 1, ACC_COUNT = 2, ACC_SUM_I = 3, ACC_SUM_F = 4, ACC_MIN_I = 5, ACC_MAX_I = 6, ACC_MIN_F = 7, ACC_MAX_F = 8 };

#define DX_AGG_MAX_KEYS 16
#define DX_AGG_MAX_ACCS 16

struct AggAccDesc { int32_t kind; int32_t col; };   // col: index into the value column list (-1 for COUNT(*))

struct AggInput {                 // one chunk of input rows
    ProgCol keys[DX_AGG_MAX_KEYS];
    ProgCol vals[DX_AGG_MAX_ACCS];
    const long long* sel;         // optional row selection
    long long n_rows;             // rows to visit (length of sel if given)
    ProgCol range_col;            // optional fused time-range filter: keep rows with lo <= range_col < hi (NULL fails)
    long long range_lo, range_hi;
    int32_t has_range, pad;
    // pane partials: when pane_us > 0 the LAST key is not loaded but computed as floor(range_col / pane_us)
    // (rows whose range_col is NULL are skipped: they belong to no window)
    long long pane_us;
};

struct AggTableDev {
    unsigned long long* slots;    // capacity * slot_words
    unsigned long long cap_mask;
    int32_t slot_words, n_keys, n_accs, n_vals;
    AggAccDesc accs[DX_AGG_MAX_ACCS];
    unsigned long long* counters; // [0] groups, [1] overflow flag
    unsigned long long max_groups;
};

struct AggTable {
    dx_ctx* ctx = nullptr;
    AggTableDev dev;
    DevBuf slots, counters;
    unsigned long long capacity = 0;
    int init(dx_ctx* c, int n_keys, int n_vals, const std::vector<AggAccDesc>& accs, unsigned long long capacity_pow2);
    int update(const AggInput& in);                    // launches the aggregation kernel for one chunk
    int finish(unsigned long long* n_groups, bool* overflow);
    // writes group rows: keys[k][g] (int64), key_null[g] (bit k set = NULL), acc[a][g] (raw 64-bit), acc_cnt...
    int extract(unsigned long long n_groups, long long* const* key_out, unsigned long long* keynull_out, unsigned long long* const* acc_out);
};

int dx_agg_keynull_to_valid(dx_ctx* c, const unsigned long long* keynull, long long n, uint32_t* valid);
