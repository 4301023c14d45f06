// dx_parquet.h — Parquet subset for state tables on disk (see dx_parquet.cu).
#pragma once
#include "dx_common.h"

struct dx_parquet {
    struct Col {
        std::string name; int type = 0, pq_type = 0, converted = -1;
        std::vector<int64_t> values;        // long / double bits / boolean 0|1 / timestamp in microseconds
        std::vector<uint8_t> valid;
        std::vector<int64_t> offsets;       // strings: n_rows + 1
        std::string bytes;
    };
    int64_t n_rows = 0;
    std::vector<Col> cols;
};
int dx_parquet_read_impl(const std::string& path_or_dir, dx_parquet** out, std::string* err);
int dx_parquet_write_impl(const std::string& path, const std::vector<std::string>& names, const std::vector<int>& types, int64_t n_rows,
                          const void* const* data, const uint8_t* const* valid, const uint8_t* const* string_bytes, std::string* err);
int dx_state_metadata_read(const std::string& location, std::string* active, std::string* standby, std::string* err);
int dx_state_metadata_write(const std::string& location, const std::string& active, const std::string& standby, std::string* err);
void dx_clear_directory_files(const std::string& dir);
