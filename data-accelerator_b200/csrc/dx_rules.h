// dx_rules.h — bucketed, bit-sliced rule evaluation (K2 fast path); see dx_rules.cu.
#pragma once
#include "dx_expr.h"
#include <memory>

// Evaluates conds[r] for every row (through sel if given) into word-major flag words
// (flags[w * n_rows + i], bit r%32 of word r/32).  *used_fast = false means the rule set did not
// fit the planner's limits and nothing was written: the caller uses the generic kernel instead.
int dx_rules_eval(dx_ctx* ctx, const BindScope& sc, const std::vector<dxsql::ExprP>& conds, int64_t n_rows, const int64_t* sel,
                  uint32_t* flags, int words, bool* used_fast, std::shared_ptr<void>* plan_cache = nullptr, int mode = 0);
// mode bits (conf keys of the flow, for A/B runs and the kernel-agreement tests)
#define DX_RULES_NO_STAGING 1   /* datax.job.process.gpu.rulesstaging=false: column tiles read from global memory, not staged by bulk copies */
#define DX_RULES_NO_FAST 2      /* datax.job.process.gpu.rulesfast=false: the per-class kernel even where the normal-form kernel applies */
// plan_cache: where the caller keeps this statement's plan between micro-batches (device tables included); it is
// rebuilt when a referenced column changes type or dictionary.

// hits per rule over a word-major flag bitmap -> counts_host[0..n_rules)
int dx_rule_counts(dx_ctx* ctx, const uint32_t* flags, int64_t n_rows, int words, int n_rules, int64_t* counts_host);
