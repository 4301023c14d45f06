// dx_join.h — inner equi-join of two device tables (see dx_join.cu).
#pragma once
#include "dx_common.h"
#include "dx_expr.h"

#define DX_JOIN_MAX_KEYS 4
struct JoinSide { ProgCol keys[DX_JOIN_MAX_KEYS]; int32_t n_keys, pad; };

// (left row, right row) index pairs of all matches; NULL keys never match.
int dx_join_pairs(dx_ctx* ctx, const JoinSide& L, int64_t n_left, const int64_t* sel_left, const JoinSide& R, int64_t n_right,
                  DevBufP* left_idx, DevBufP* right_idx, int64_t* n_pairs, bool left_outer = false);
