// dx_shuffle.h — K5: key-hash partition/pack and unpack for the multi-GPU exchange (see dx_shuffle.cu).
#pragma once
#include "dx_common.h"
#include "dx_expr.h"

#define DX_SHUFFLE_MAX_WORLD 64
#define DX_SHUFFLE_MAX_COLS 32

struct ShuffleArgs {
    ProgCol cols[DX_SHUFFLE_MAX_COLS];
    int32_t key_idx[8];      // positions (within cols) of the group-key columns
    int32_t n_cols, n_keys, world, pad;
    long long n_rows;
};
struct UnpackCols { long long* p[DX_SHUFFLE_MAX_COLS]; };

// Packs the rows of t (columns `cols`, by table index) grouped by owner rank into packed_dev
// (n_rows * 8 * (n_cols + 1) bytes, caller-owned); counts_host[r] = rows destined to rank r.
int dx_shuffle_pack(dx_ctx* ctx, const DxTable* t, const std::vector<int>& cols, const std::vector<int>& key_pos, int world,
                    void* packed_dev, int64_t* counts_host);
// Rebuilds a table from n_rows received records; with n_segs > 0 the records lie in n_segs runs (seg_first / seg_count, in
// records) of the buffer, one block per sender, and n_rows is ignored.
int dx_shuffle_unpack(dx_ctx* ctx, const std::vector<std::string>& names, const std::vector<int>& types, const void* packed_dev, int64_t n_rows,
                      DxTableP* out, int n_segs = 0, const int64_t* seg_first = nullptr, const int64_t* seg_count = nullptr);

// rows per owner rank only
int dx_shuffle_count(dx_ctx* ctx, const DxTable* t, const std::vector<int>& cols, const std::vector<int>& key_pos, int world, int64_t* counts_host);
// pack fused with the exchange: records go straight into the owners' receive buffers (peer pointers over NVLink)
// block_records (optional): this rank's block in owner r's buffer holds block_records[r] records — the layout that needs no
// count exchange before the put; sent (optional): records written per owner.  A block that is too small is an error.
int dx_shuffle_put(dx_ctx* ctx, const DxTable* t, const std::vector<int>& cols, const std::vector<int>& key_pos, int world,
                   void* const* peer_bufs, const int64_t* first_record, const int64_t* block_records = nullptr, int64_t* sent = nullptr);
