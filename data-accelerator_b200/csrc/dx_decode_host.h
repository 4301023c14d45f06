// dx_decode_host.h — host-side handle of the K1 decoder.
#pragma once
#include "dx_decode.cuh"

struct dx_decoder {
    dx_ctx* ctx = nullptr;
    DSchemaDev hs;
    std::vector<std::string> leaf_names;   // dotted paths, e.g. deviceDetails.deviceId
    std::vector<int> leaf_types;
    std::vector<DxDictP> dicts;            // per leaf; null for non-string leaves
    DevBuf dev_schema, dev_dicts, dev_stats, tile_state, staging, staging_off;
    DecodeStats last_stats{0, 0, 0, 0, 0};
    uint32_t dict_capacity = 1u << 16;     // slots per string column (power of two)
    uint32_t dict_heap_bytes = 4u << 20;
    int dict_retries = 0;                  // decode passes repeated after a dictionary grew (decode_device)
    int ctas_per_sm = 1;
    int occ_by_tile[18] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1};   // resident CTAs per SM of each tile-size instantiation of K1: [0..9) NDJSON, [9..18) caller-framed
    int init(dx_ctx* c, const std::string& schema_json);
    int decode_device(const uint8_t* json_dev, int64_t n_bytes, const int64_t* offsets_dev, int64_t n_events, DxTableP* out);
    int decode_host(const uint8_t* json, int64_t n_bytes, const int64_t* offsets, int64_t n_events, DxTableP* out);
    // allocation-free variant: caller-owned outputs, kernel scratch in a fixed in-library device array
    int decode_noalloc(const uint8_t* json_dev, int64_t n_bytes, int64_t n_events, void* const* column_ptrs, int n_columns,
                       void* mask_ptr, int64_t* rows_out, int64_t* null_rows_out);
};
