// dx_tojson.h — K6: result-table rows -> NDJSON on the device (see dx_tojson.cu).
#pragma once
#include "dx_common.h"

struct dx_json {
    dx_ctx* ctx = nullptr;
    int64_t n_rows = 0, n_bytes = 0;
    DevBufP bytes;                  // n_bytes: row r is bytes[offsets[r] .. offsets[r+1] - 1), followed by '\n'
    DevBufP offsets;                // n_rows + 1 int64
    std::vector<DevBufP> keep;      // program buffers, alive until the kernels have run
    std::vector<long long> consts; std::vector<void*> const_dst;     // constant columns: one element each, uploaded with the program
};
int dx_json_build(const DxTable* t, dx_json** out);
