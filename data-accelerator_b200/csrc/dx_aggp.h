// dx_aggp.h — K3P: partitioned shared-memory aggregation for the window pane partials.
//
// The global-memory hash table of dx_agg.cu pays a random 64-bit atomic in L2 per accumulator per row.  For the pane
// partials (dx_flow.h: PanePartial) the same work is done the way BASELINE.json's north_star asks for stage 3 —
// "shared-memory hash ... window aggregation": rows are first scattered into P partitions by the hash of their group
// key (sequential HBM traffic), then one CTA per partition aggregates its rows in a shared-memory hash table and writes
// the groups out contiguously.  Partials stay partition-ordered, so combining the partials of the cached batches of a
// window (CommonProcessorFactory.scala:203-227 rescans those batches instead) needs no second scatter: CTA p reads
// partition p of every partial.
#pragma once
#include "dx_agg.h"

struct AggPPartial {
    DevBufP recs;                 // n groups x gw 8-byte words: [hash][keys nk][pane][key-null bits][accumulators na], partition-ordered
    DevBufP part_base, part_cnt;  // P entries each (uint32): first group and group count of partition p
    int64_t n = 0;
    int P = 0, gw = 0;
};

// whether the shared-memory table takes nk group keys (+ pane) and na accumulators
bool dx_aggp_fits(int nk, int na);

// the group keys (in.keys[0..nk)), the pane key computed from in.range_col / in.pane_us, and the value columns of one
// cached batch -> partition-ordered partial.  *overflow: a partition held more groups than the shared table takes
// (the caller retries with more partitions).
int dx_aggp_build(dx_ctx* ctx, const AggInput& in, int nk, int nv, const std::vector<AggAccDesc>& accs, int P, AggPPartial* out, bool* overflow,
                  long long* stride_hint = nullptr, long long* hint_rows = nullptr, bool force_two_pass = false);
// stride_hint / hint_rows (in/out, per statement signature): records of the fullest partition and the row count of the
// previous build; with them the scatter runs as one pass into fixed partition slices.

// partials of one signature (same nk / accumulators / P), panes in [pane_lo, pane_hi) -> groups as column arrays
// (key_out[k][g], keynull_out[g] with bit k set = NULL, acc_out[a][g] raw 64-bit, as AggTable::extract writes them).
// src_acc[a]: index of accumulator a among the partials' accumulators (partials shared between statements may carry more).
// The arrays must hold max_groups entries; *n_groups is the number written.
int dx_aggp_combine(dx_ctx* ctx, const std::vector<const AggPPartial*>& parts, int nk, const std::vector<AggAccDesc>& accs, const std::vector<int>& src_acc,
                    long long pane_lo, long long pane_hi, long long max_groups, long long* const* key_out,
                    unsigned long long* keynull_out, unsigned long long* const* acc_out, long long* n_groups, bool* overflow);
