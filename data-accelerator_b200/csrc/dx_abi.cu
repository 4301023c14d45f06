// dx_abi.cu — extern "C" entry points of libdxgpu.so (see include/dxgpu.h).
#include "dx_common.h"
#include "dx_abi_types.h"
#include <memory>
#include "dx_decode_host.h"
#include "dx_flow.h"
#include "dx_tojson.h"
#include "dx_inflate.h"
#include "dx_parquet.h"
#include "dx_codegen.h"
#include <sys/syscall.h>
#include <unistd.h>
#include "dx_rules.h"
#include "dx_shuffle.h"

int dx_gen_device(dx_ctx* c, const dx_gen_params* g, int64_t first, int64_t n, uint8_t** bytes_out,
                  int64_t* n_bytes_out, int64_t** offsets_out);

// Every table handed across the ABI is a heap-allocated shared_ptr holder.
struct TableHandle { DxTableP t; };
static inline const DxTable* T(const dx_table* h) { return reinterpret_cast<const TableHandle*>(h)->t.get(); }
dx_table* dx_wrap_table(const DxTableP& t) { return reinterpret_cast<dx_table*>(new TableHandle{t}); }
DxTableP dx_unwrap_table(const dx_table* h) { return reinterpret_cast<const TableHandle*>(h)->t; }

extern "C" {

int dx_ctx_create(int device_ordinal, dx_ctx** out) {
    if (!out) return DX_ERR_ARG;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device_ordinal < 0 || device_ordinal >= n) return DX_ERR_CUDA;
    if (cudaSetDevice(device_ordinal) != cudaSuccess) return DX_ERR_CUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device_ordinal) != cudaSuccess) return DX_ERR_CUDA;
    if (prop.major != 10) return DX_ERR_UNSUPPORTED;   // sm_100a binary only; no other code path exists
    dx_ctx* c = new dx_ctx();
    c->device = device_ordinal; c->pool.device = device_ordinal; c->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return DX_ERR_CUDA; }
    *out = c;
    return DX_OK;
}
void dx_ctx_destroy(dx_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaStreamDestroy(c->stream);
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    delete c;
}
const char* dx_ctx_last_error(const dx_ctx* c) { return c ? c->last_error.c_str() : "null context"; }
int dx_ctx_set_timing(dx_ctx* c, int enabled) { c->timing = enabled != 0; return DX_OK; }
int dx_ctx_take_stage_ms(dx_ctx* c, const char* stage, double* total_ms, int64_t* samples) {
    auto it = c->stage_ms.find(stage);
    double t = 0; int64_t n = 0;
    if (it != c->stage_ms.end()) { for (float v : it->second) t += v; n = (int64_t)it->second.size(); it->second.clear(); }
    if (total_ms) *total_ms = t;
    if (samples) *samples = n;
    return DX_OK;
}
int64_t dx_ctx_take_launches(dx_ctx* c) {
    int64_t n = 0;
    for (auto& kv : c->launches) { n += kv.second; kv.second = 0; }
    return n;
}
int64_t dx_ctx_take_counter(dx_ctx* c, const char* name) {
    auto it = c->counters.find(name);
    if (it == c->counters.end()) return 0;
    int64_t v = it->second; it->second = 0; return v;
}
void* dx_ctx_stream(dx_ctx* c) { return reinterpret_cast<void*>(c->stream); }
int dx_ctx_synchronize(dx_ctx* c) { DX_CUDA(c, cudaStreamSynchronize(c->stream)); return DX_OK; }
int dx_dev_alloc(dx_ctx* c, int64_t bytes, void** out) {
    if (bytes < 0 || !out) return c->set_error(DX_ERR_ARG, "dx_dev_alloc: bad arguments");
    cudaSetDevice(c->device);
    void* p = c->pool.alloc((size_t)bytes + 64, true);      // the caller's blocks are allocation bases (CUDA IPC exports them)
    if (!p) return c->set_error(DX_ERR_CUDA, "device allocation failed");
    *out = p; return DX_OK;
}
void dx_dev_free(dx_ctx* c, void* p) { if (c && p) c->pool.release(p); }
// NUMA node the GPU hangs off (sysfs; -1 when the platform does not say, e.g. a single-node VM)
int dx_ctx_numa_node(dx_ctx* c) {
    char bus[32] = {0};
    if (!c || cudaDeviceGetPCIBusId(bus, sizeof bus, c->device) != cudaSuccess) return -1;
    for (char* p = bus; *p; p++) *p = (char)tolower((unsigned char)*p);
    const std::string path = std::string("/sys/bus/pci/devices/") + bus + "/numa_node";
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return -1;
    int node = -1;
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
    return node;
}
int dx_host_alloc(dx_ctx* c, int64_t bytes, void** out) {
    if (bytes < 0 || !out) return c->set_error(DX_ERR_ARG, "dx_host_alloc: bad arguments");
    cudaSetDevice(c->device);
    // The pages are taken from the memory of the GPU's own NUMA node (preferred, not bound: a full node still allocates): a
    // copy from the other socket's memory crosses the inter-socket link and, with one rank per GPU, eight of them share it.
    const int node = dx_ctx_numa_node(c);
    bool policy_set = false;
    if (node >= 0 && node < 1024) {
        unsigned long mask[16] = {0};
        mask[node / 64] |= 1ul << (node % 64);
        policy_set = syscall(SYS_set_mempolicy, 1 /* MPOL_PREFERRED */, mask, 1025ul) == 0;
    }
    const cudaError_t e = cudaHostAlloc(out, (size_t)bytes + 64, cudaHostAllocDefault);
    if (policy_set) syscall(SYS_set_mempolicy, 0 /* MPOL_DEFAULT */, nullptr, 0ul);
    DX_CUDA(c, e);
    return DX_OK;
}
void dx_host_free(dx_ctx* c, void* p) { if (c && p) { cudaSetDevice(c->device); cudaFreeHost(p); } }
int dx_copy_d2h_async(dx_ctx* c, void* dst, const void* src, int64_t bytes) {
    DX_CUDA(c, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost, c->stream));
    return DX_OK;
}
int dx_copy_h2d(dx_ctx* c, void* dst, const void* src, int64_t bytes) {
    DX_CUDA(c, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyHostToDevice, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream)); return DX_OK;
}
int dx_copy_d2h(dx_ctx* c, void* dst, const void* src, int64_t bytes) {
    DX_CUDA(c, cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream)); return DX_OK;
}

// ---- decoder -------------------------------------------------------------------------------
int dx_decoder_create(dx_ctx* c, const char* schema_json, dx_decoder** out) {
    if (!c || !schema_json || !out) return DX_ERR_ARG;
    cudaSetDevice(c->device);
    dx_decoder* d = new dx_decoder();
    int rc = d->init(c, schema_json);
    if (rc) { delete d; return rc; }
    *out = d; return DX_OK;
}
void dx_decoder_destroy(dx_decoder* d) { delete d; }
int dx_decoder_decode(dx_decoder* d, const uint8_t* json, int64_t n_bytes, const int64_t* offsets,
                      int64_t n_events, int memory_kind, dx_table** out) {
    if (!d || !out || (n_bytes > 0 && !json)) return DX_ERR_ARG;
    cudaSetDevice(d->ctx->device);
    DxTableP t;
    int rc = (memory_kind == DX_MEM_DEVICE) ? d->decode_device(json, n_bytes, offsets, n_events, &t)
                                            : d->decode_host(json, n_bytes, offsets, n_events, &t);
    if (rc) return rc;
    *out = dx_wrap_table(t);
    return DX_OK;
}
// Decode for callers that must not allocate device memory in their per-batch path (see the block
// comment above dx_decoder::decode_noalloc in dx_decode.cu).  No pool, no cudaMalloc/cudaFree here.
int dx_decoder_decode_noalloc(dx_decoder* d, const uint8_t* json_dev, int64_t n_bytes, int64_t n_events,
                              void* const* column_ptrs, int n_columns, void* mask_ptr, int64_t* rows, int64_t* null_rows) {
    if (!d) return DX_ERR_ARG;
    if (cudaSetDevice(d->ctx->device) != cudaSuccess) return d->ctx->set_error(DX_ERR_CUDA, "cudaSetDevice failed");
    return d->decode_noalloc(json_dev, n_bytes, n_events, column_ptrs, n_columns, mask_ptr, rows, null_rows);
}
int dx_decoder_dict_size(dx_decoder* d, int column, int64_t* n_codes) {
    if (!d || !n_codes) return DX_ERR_ARG;
    if (column < 0 || column >= (int)d->dicts.size() || !d->dicts[column]) return d->ctx->set_error(DX_ERR_ARG, "not a string column of this decoder");
    int prev = -1; cudaGetDevice(&prev);
    cudaSetDevice(d->ctx->device);
    int rc = d->dicts[column]->sync_to_host();
    if (prev >= 0) cudaSetDevice(prev);
    if (rc) return rc;
    *n_codes = (int64_t)d->dicts[column]->host_strings.size();
    return DX_OK;
}
int dx_decoder_dict_string(dx_decoder* d, int column, int64_t code, const char** bytes, int64_t* len) {
    if (!d || !bytes || !len) return DX_ERR_ARG;
    if (column < 0 || column >= (int)d->dicts.size() || !d->dicts[column]) return d->ctx->set_error(DX_ERR_ARG, "not a string column of this decoder");
    const auto& hs = d->dicts[column]->host_strings;
    if (code < 0 || code >= (int64_t)hs.size()) return d->ctx->set_error(DX_ERR_ARG, "dictionary code out of range (call dx_decoder_dict_size first)");
    *bytes = hs[(size_t)code].data(); *len = (int64_t)hs[(size_t)code].size();
    return DX_OK;
}
int64_t dx_decoder_limited_rows(const dx_decoder* d) { return d ? (int64_t)d->last_stats.limited_rows : 0; }
int dx_decoder_stats(const dx_decoder* d, int64_t* rows, int64_t* null_rows, int64_t* fast_rows) {
    if (rows) *rows = (int64_t)d->last_stats.rows;
    if (null_rows) *null_rows = (int64_t)d->last_stats.null_rows;
    if (fast_rows) *fast_rows = (int64_t)d->last_stats.fast_rows;
    return DX_OK;
}

// ---- tables --------------------------------------------------------------------------------
int64_t dx_table_num_rows(const dx_table* h) { return T(h)->n_rows; }
int dx_table_num_columns(const dx_table* h) { return (int)T(h)->cols.size(); }
const char* dx_table_column_name(const dx_table* h, int col) {
    const DxTable* t = T(h);
    return (col >= 0 && col < (int)t->cols.size()) ? t->cols[col].name.c_str() : nullptr;
}
int dx_table_column_nesting(const dx_table* h, int col, const char** kinds) {
    const DxTable* t = T(h);
    if (col < 0 || col >= (int)t->cols.size()) return -1;
    if (kinds) *kinds = t->cols[col].nest_kinds.c_str();
    return (int)t->cols[col].nest_path.size();
}
const char* dx_table_column_path(const dx_table* h, int col, int level) {
    const DxTable* t = T(h);
    if (col < 0 || col >= (int)t->cols.size()) return nullptr;
    const auto& p = t->cols[col].nest_path;
    return (level >= 0 && level < (int)p.size()) ? p[level].c_str() : nullptr;
}
int dx_table_column_type(const dx_table* h, int col) {
    const DxTable* t = T(h);
    return (col >= 0 && col < (int)t->cols.size()) ? t->cols[col].type : 0;
}
int dx_table_find_column(const dx_table* h, const char* name) { return T(h)->find(name); }
int dx_table_flag_words(const dx_table* h, int col) {
    const DxTable* t = T(h);
    return (col >= 0 && col < (int)t->cols.size()) ? t->cols[col].flag_words : 0;
}
int dx_table_column_device(const dx_table* h, int col, const void** data, const void** vmask, int* vw, int* vb) {
    const DxTable* t = T(h);
    if (col < 0 || col >= (int)t->cols.size()) return t->ctx->set_error(DX_ERR_ARG, "column index out of range");
    const DxColumn& c = t->cols[col];
    if (data) *data = c.data ? c.data->p : nullptr;
    if (vmask) *vmask = c.vmask ? c.vmask->p : nullptr;
    if (vw) *vw = c.vwidth;
    if (vb) *vb = c.vbit;
    return DX_OK;
}

int dx_table_column_to_host(const dx_table* h, int col, void* data, uint8_t* valid) {
    const DxTable* t = T(h);
    dx_ctx* c = t->ctx;
    if (col < 0 || col >= (int)t->cols.size()) return c->set_error(DX_ERR_ARG, "column index out of range");
    cudaSetDevice(c->device);
    const DxColumn& k = t->cols[col];
    const int64_t n = t->n_rows;
    size_t w = (k.type == DX_T_FLAGS) ? (size_t)k.flag_words * 4 : (size_t)dx_type_width(k.type);
    if (data && n > 0) {
        if (k.data) DX_CUDA(c, cudaMemcpyAsync(data, k.data->p, (size_t)n * w, cudaMemcpyDeviceToHost, c->stream));
        else {
            // constant column: replicate on the host side of the ABI (sink hand-off convenience)
            for (int64_t i = 0; i < n; i++) {
                if (k.type == DX_T_F64) reinterpret_cast<double*>(data)[i] = k.const_f;
                else if (k.type == DX_T_STR) reinterpret_cast<int32_t*>(data)[i] = 0;   // constant string: one-entry dictionary
                else reinterpret_cast<int64_t*>(data)[i] = k.const_i;
            }
        }
    }
    if (valid && n > 0) {
        if (k.all_null) memset(valid, 0, (size_t)n);
        else if (!k.vmask) memset(valid, 1, (size_t)n);
        else {
            std::vector<uint8_t> m((size_t)n * k.vwidth);
            DX_CUDA(c, cudaMemcpyAsync(m.data(), k.vmask->p, m.size(), cudaMemcpyDeviceToHost, c->stream));
            DX_CUDA(c, cudaStreamSynchronize(c->stream));
            for (int64_t i = 0; i < n; i++) {
                uint32_t v = (k.vwidth == 1) ? m[i] : reinterpret_cast<const uint32_t*>(m.data())[i];
                valid[i] = (v >> k.vbit) & 1u;
            }
        }
    }
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    return DX_OK;
}

int dx_table_dict_size(const dx_table* h, int col, int64_t* n_codes) {
    const DxTable* t = T(h);
    if (col >= 0 && col < (int)t->cols.size() && t->cols[col].type == DX_T_STR && !t->cols[col].data) { *n_codes = 1; return DX_OK; }
    if (col >= 0 && col < (int)t->cols.size() && t->cols[col].type == DX_T_STR && !t->cols[col].dict) { *n_codes = 0; return DX_OK; }   // no string seen yet
    if (col < 0 || col >= (int)t->cols.size() || !t->cols[col].dict) return t->ctx->set_error(DX_ERR_ARG, "not a dictionary column");
    cudaSetDevice(t->ctx->device);
    int rc = t->cols[col].dict->sync_to_host();
    if (rc) return rc;
    *n_codes = (int64_t)t->cols[col].dict->host_strings.size();
    return DX_OK;
}
int dx_table_dict_string(const dx_table* h, int col, int64_t code, const char** bytes, int64_t* len) {
    const DxTable* t = T(h);
    if (col >= 0 && col < (int)t->cols.size() && t->cols[col].type == DX_T_STR && !t->cols[col].data && code == 0) {
        *bytes = t->cols[col].const_s.data(); *len = (int64_t)t->cols[col].const_s.size(); return DX_OK;
    }
    if (col < 0 || col >= (int)t->cols.size() || !t->cols[col].dict) return t->ctx->set_error(DX_ERR_ARG, "not a dictionary column");
    const auto& hs = t->cols[col].dict->host_strings;
    if (code < 0 || code >= (int64_t)hs.size()) return t->ctx->set_error(DX_ERR_ARG, "dictionary code out of range (call dx_table_dict_size first)");
    *bytes = hs[(size_t)code].data(); *len = (int64_t)hs[(size_t)code].size();
    return DX_OK;
}
void dx_table_release(dx_table* h) { delete reinterpret_cast<TableHandle*>(h); }

int dx_table_rule_count(const dx_table* h, int col) {
    const DxTable* t = T(h);
    if (col < 0 || col >= (int)t->cols.size() || !t->cols[col].aux) return 0;
    return (int)static_cast<RuleTable*>(t->cols[col].aux.get())->rules.size();
}
int dx_table_rule_hits(const dx_table* h, int col, int64_t* counts, int n_counts) {
    const DxTable* t = T(h);
    if (!counts || col < 0 || col >= (int)t->cols.size()) return DX_ERR_ARG;
    const DxColumn& c = t->cols[col];
    if (c.type != DX_T_FLAGS || !c.aux) return t->ctx->set_error(DX_ERR_ARG, "column '" + c.name + "' is not a rule-flag column");
    const int R = (int)static_cast<RuleTable*>(c.aux.get())->rules.size();
    if (n_counts < R) return t->ctx->set_error(DX_ERR_ARG, "rule hit counts need " + std::to_string(R) + " entries");
    cudaSetDevice(t->ctx->device);
    return dx_rule_counts(t->ctx, c.data->as<uint32_t>(), t->n_rows, c.flag_words, R, counts);
}
static std::string json_escape(const std::string& s) {
    std::string o;
    for (unsigned char c : s) {
        if (c == '"' || c == '\\') { o.push_back('\\'); o.push_back((char)c); }
        else if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
        else o.push_back((char)c);
    }
    return o;
}
int dx_table_rule_json(const dx_table* h, int col, int rule, const char** json, int64_t* len) {
    const DxTable* t = T(h);
    if (col < 0 || col >= (int)t->cols.size() || !t->cols[col].aux) return t->ctx->set_error(DX_ERR_ARG, "not a rule-flag column");
    RuleTable* rt = static_cast<RuleTable*>(t->cols[col].aux.get());
    if (rule < 0 || rule >= (int)rt->rules.size()) return t->ctx->set_error(DX_ERR_ARG, "rule index out of range");
    static thread_local std::string buf;
    buf = "{";
    for (size_t i = 0; i < rt->rules[rule].size(); i++) {
        if (i) buf += ",";
        buf += "\"" + json_escape(rt->rules[rule][i].first) + "\":\"" + json_escape(rt->rules[rule][i].second) + "\"";
    }
    buf += "}";
    *json = buf.c_str(); *len = (int64_t)buf.size();
    return DX_OK;
}

// ---- design time: rules + user code -> transform text (dx_codegen.cu) ---------------------------
static int codegen_out(int rc, const dx_codegen_result& res, const std::string& err, char* code_out, int64_t code_cap, char* info_out, int64_t info_cap) {
    if (rc) { snprintf(code_out, (size_t)code_cap, "%s", err.c_str()); return rc; }
    if ((int64_t)res.code.size() + 1 > code_cap) { snprintf(code_out, (size_t)code_cap, "generated code needs %zu bytes", res.code.size() + 1); return DX_ERR_CAPACITY; }
    memcpy(code_out, res.code.c_str(), res.code.size() + 1);
    if (info_out && info_cap > 0) {
        std::string info;
        for (auto& o : res.outputs) info += "output\t" + o.first + "\t" + o.second + "\n";
        for (auto& t : res.tables) info += "table\t" + t.first + "\t" + t.second + "\n";
        for (auto& w : res.windows) info += "window\t" + w.first + "\t" + w.second + "\n";
        for (auto& m : res.metrics) info += "metric\t" + m + "\n";
        snprintf(info_out, (size_t)info_cap, "%s", info.c_str());
    }
    return DX_OK;
}
int dx_codegen_rules(const char* user_code, const char* rules_json, const char* query_templates_xml, const char* output_templates_xml,
                     const char* product_id, char* code_out, int64_t code_cap, char* info_out, int64_t info_cap) {
    if (!user_code || !rules_json || !query_templates_xml || !output_templates_xml || !code_out || code_cap < 1) return DX_ERR_ARG;
    dx_codegen_result res; std::string err;
    int rc = dx_codegen_impl(user_code, rules_json, query_templates_xml, output_templates_xml, product_id ? product_id : "", &res, &err);
    return codegen_out(rc, res, err, code_out, code_cap, info_out, info_cap);
}
int dx_codegen_flow(const char* flow_json, const char* query_templates_xml, const char* output_templates_xml, char* code_out, int64_t code_cap,
                    char* info_out, int64_t info_cap) {
    if (!flow_json || !query_templates_xml || !output_templates_xml || !code_out || code_cap < 1) return DX_ERR_ARG;
    dx_codegen_result res; std::string err;
    int rc = dx_codegen_flow_impl(flow_json, query_templates_xml, output_templates_xml, &res, &err);
    return codegen_out(rc, res, err, code_out, code_cap, info_out, info_cap);
}

// ---- state tables on disk: Parquet A/B + metadata.info (dx_parquet.cu) -------------------------
static void pq_err(char* err, int64_t cap, const std::string& m) { if (err && cap > 0) { snprintf(err, (size_t)cap, "%s", m.c_str()); } }
int dx_parquet_write(const char* path, int n_columns, const char* const* names, const int* types, int64_t n_rows, const void* const* column_data,
                     const uint8_t* const* column_valid, const uint8_t* const* string_bytes, char* err, int64_t err_cap) {
    if (!path || n_columns < 1 || !names || !types || n_rows < 0 || !column_data) { pq_err(err, err_cap, "dx_parquet_write: bad arguments"); return DX_ERR_ARG; }
    std::vector<std::string> nm; std::vector<int> ty;
    for (int i = 0; i < n_columns; i++) { nm.push_back(names[i]); ty.push_back(types[i]); }
    std::string e; int rc = dx_parquet_write_impl(path, nm, ty, n_rows, column_data, column_valid, string_bytes, &e);
    if (rc) pq_err(err, err_cap, e);
    return rc;
}
int dx_parquet_read(const char* path_or_dir, dx_parquet** out, char* err, int64_t err_cap) {
    if (!path_or_dir || !out) return DX_ERR_ARG;
    std::string e; int rc = dx_parquet_read_impl(path_or_dir, out, &e);
    if (rc) pq_err(err, err_cap, e);
    return rc;
}
int64_t dx_parquet_num_rows(const dx_parquet* p) { return p ? p->n_rows : 0; }
int dx_parquet_num_columns(const dx_parquet* p) { return p ? (int)p->cols.size() : 0; }
const char* dx_parquet_column_name(const dx_parquet* p, int c) { return p && c >= 0 && c < (int)p->cols.size() ? p->cols[c].name.c_str() : nullptr; }
int dx_parquet_column_type(const dx_parquet* p, int c) { return p && c >= 0 && c < (int)p->cols.size() ? p->cols[c].type : 0; }
int dx_parquet_column(const dx_parquet* p, int c, const void** data, const uint8_t** valid, const uint8_t** string_bytes) {
    if (!p || c < 0 || c >= (int)p->cols.size()) return DX_ERR_ARG;
    const dx_parquet::Col& col = p->cols[c];
    if (data) *data = col.type == DX_T_STR ? (const void*)col.offsets.data() : (const void*)col.values.data();
    if (valid) *valid = col.valid.data();
    if (string_bytes) *string_bytes = col.type == DX_T_STR ? reinterpret_cast<const uint8_t*>(col.bytes.data()) : nullptr;
    return DX_OK;
}
void dx_parquet_release(dx_parquet* p) { delete p; }

int dx_flow_state_save(dx_flow* f, const char* table, const char* location) {
    if (!f || !table || !location) return DX_ERR_ARG;
    dx_ctx* c = f->ctx;
    cudaSetDevice(c->device);
    auto it = f->state_tables.find(table);
    if (it == f->state_tables.end()) return c->set_error(DX_ERR_ARG, std::string("no state table '") + table + "'");
    StateTable& st = it->second;
    const DxTableP t = st.phys[st.active];
    const std::string loc = std::string(location).substr(0, std::string(location).find_last_not_of('/') + 1);
    const std::string active = st.active ? "B" : "A", standby = st.active ? "A" : "B";
    std::vector<std::string> names; std::vector<int> types;
    for (auto& kv : st.schema) { names.push_back(kv.first); types.push_back(kv.second); }
    const int64_t n = t ? t->n_rows : 0;
    // the columns on the host, strings as offsets + bytes
    std::vector<std::vector<int64_t>> vals(names.size()); std::vector<std::vector<uint8_t>> valid(names.size()); std::vector<std::string> bytes(names.size());
    for (size_t k = 0; k < names.size(); k++) {
        valid[k].assign((size_t)(n > 0 ? n : 1), 0);
        if (types[k] == DX_T_STR) {
            vals[k].assign((size_t)n + 1, 0);
            if (n > 0) {
                if ((size_t)k >= t->cols.size()) return c->set_error(DX_ERR_STATE, "state table has fewer columns than its schema");
                std::vector<int32_t> codes((size_t)n);
                dx_table* h = dx_wrap_table(t);
                int rc = dx_table_column_to_host(h, (int)k, codes.data(), valid[k].data());
                dx_table_release(h);
                if (rc) return rc;
                const DxColumn& col = t->cols[k];
                if (col.dict) { rc = col.dict->sync_to_host(); if (rc) return rc; }
                for (int64_t r = 0; r < n; r++) {
                    if (valid[k][(size_t)r] && col.dict && codes[(size_t)r] >= 0 && codes[(size_t)r] < (int32_t)col.dict->host_strings.size()) bytes[k] += col.dict->host_strings[(size_t)codes[(size_t)r]];
                    else if (valid[k][(size_t)r] && !col.dict) bytes[k] += col.const_s;
                    vals[k][(size_t)r + 1] = (int64_t)bytes[k].size();
                }
            }
        } else {
            vals[k].assign((size_t)(n > 0 ? n : 1), 0);
            if (n > 0) { dx_table* h = dx_wrap_table(t); int rc = dx_table_column_to_host(h, (int)k, vals[k].data(), valid[k].data()); dx_table_release(h); if (rc) return rc; }
        }
    }
    std::vector<const void*> dp; std::vector<const uint8_t*> vp, sp;
    for (size_t k = 0; k < names.size(); k++) { dp.push_back(vals[k].data()); vp.push_back(valid[k].data()); sp.push_back(reinterpret_cast<const uint8_t*>(bytes[k].data())); }
    std::string e;
    dx_clear_directory_files(loc + "/" + active);
    int rc = dx_parquet_write_impl(loc + "/" + active + "/part-00000.parquet", names, types, n, dp.data(), vp.data(), sp.data(), &e);
    if (rc) return c->set_error(rc, "state table '" + st.name + "': " + e);
    rc = dx_state_metadata_write(loc, active, standby, &e);
    if (rc) return c->set_error(rc, "state table '" + st.name + "': " + e);
    return DX_OK;
}
int dx_flow_state_load(dx_flow* f, const char* table, const char* location) {
    if (!f || !table || !location) return DX_ERR_ARG;
    dx_ctx* c = f->ctx;
    auto it = f->state_tables.find(table);
    if (it == f->state_tables.end()) return c->set_error(DX_ERR_ARG, std::string("no state table '") + table + "'");
    StateTable& st = it->second;
    const std::string loc = std::string(location).substr(0, std::string(location).find_last_not_of('/') + 1);
    std::string active, standby, e;
    int rc = dx_state_metadata_read(loc, &active, &standby, &e);
    if (rc) return c->set_error(rc, e);
    if (active != "A" && active != "B") return c->set_error(DX_ERR_PARSE, "metadata.info of '" + st.name + "': active=" + active);
    dx_parquet* pq = nullptr;
    rc = dx_parquet_read_impl(loc + "/" + active, &pq, &e);
    if (rc && e.find("cannot open") == std::string::npos && !e.empty()) return c->set_error(rc, "state table '" + st.name + "': " + e);
    std::unique_ptr<dx_parquet> hold(pq);
    if (!pq || pq->cols.empty()) { st.active = active == "B" ? 1 : 0; return DX_OK; }        // CREATE TABLE IF NOT EXISTS over an empty location: an empty table
    if (pq->cols.size() != st.schema.size()) return c->set_error(DX_ERR_ARG, "state table '" + st.name + "': the files have " + std::to_string(pq->cols.size()) + " columns, the schema " + std::to_string(st.schema.size()));
    std::vector<const void*> dp; std::vector<const uint8_t*> vp, sp;
    for (size_t k = 0; k < pq->cols.size(); k++) {
        const dx_parquet::Col& col = pq->cols[k];
        const int want = st.schema[k].second;
        const bool ok = col.type == want || (want == DX_T_I64 && col.type == DX_T_BOOL) || (want == DX_T_BOOL && col.type == DX_T_I64);
        if (!ok) return c->set_error(DX_ERR_ARG, "state table '" + st.name + "': column '" + st.schema[k].first + "' has another type in the files");
        dp.push_back(col.type == DX_T_STR ? (const void*)col.offsets.data() : (const void*)col.values.data());
        vp.push_back(col.valid.data()); sp.push_back(reinterpret_cast<const uint8_t*>(col.bytes.data()));
    }
    return dx_flow_state_import(f, table, pq->n_rows, (int)pq->cols.size(), dp.data(), vp.data(), sp.data(), active == "B" ? 1 : 0);
}

// ---- blob input: gzip members inflated on the device (dx_inflate.cu) --------------------------
int dx_gunzip_device(dx_ctx* c, const uint8_t* gz, const int64_t* member_offsets, int n_members, int memory_kind, void** out_dev, int64_t* n_bytes) {
    if (!c || !out_dev || !n_bytes) return DX_ERR_ARG;
    cudaSetDevice(c->device);
    DevBufP buf;
    int rc = dx_gunzip(c, gz, member_offsets, n_members, memory_kind, &buf, n_bytes);
    if (rc) return rc;
    *out_dev = buf->p; buf->p = nullptr;          // ownership passes to the caller (dx_dev_free)
    return DX_OK;
}

// ---- sink hand-off: rows -> JSON on the device (dx_tojson.cu) --------------------------------
int dx_table_to_json(const dx_table* h, dx_json** out) {
    const DxTable* t = T(h);
    if (!t || !out) return DX_ERR_ARG;
    cudaSetDevice(t->ctx->device);
    try { return dx_json_build(t, out); } catch (std::exception& e) { return t->ctx->set_error(DX_ERR_UNSUPPORTED, e.what()); }
}
int64_t dx_json_num_rows(const dx_json* j) { return j ? j->n_rows : 0; }
int64_t dx_json_num_bytes(const dx_json* j) { return j ? j->n_bytes : 0; }
int dx_json_device(const dx_json* j, const uint8_t** bytes, const int64_t** offsets) {
    if (!j) return DX_ERR_ARG;
    if (bytes) *bytes = j->bytes ? j->bytes->as<uint8_t>() : nullptr;
    if (offsets) *offsets = j->offsets->as<int64_t>();
    return DX_OK;
}
int dx_json_to_host(const dx_json* j, uint8_t* bytes, int64_t* offsets) {
    if (!j) return DX_ERR_ARG;
    dx_ctx* c = j->ctx;
    cudaSetDevice(c->device);
    if (bytes && j->n_bytes) DX_CUDA(c, cudaMemcpyAsync(bytes, j->bytes->p, (size_t)j->n_bytes, cudaMemcpyDeviceToHost, c->stream));
    if (offsets) DX_CUDA(c, cudaMemcpyAsync(offsets, j->offsets->p, (size_t)(j->n_rows + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    return DX_OK;
}
void dx_json_release(dx_json* j) { delete j; }
int dx_table_filter_flags(const dx_table* h, const char* expr, uint8_t* flags) {
    const DxTable* t = T(h);
    if (!t || !expr || (!flags && t->n_rows)) return DX_ERR_ARG;
    dx_ctx* c = t->ctx;
    cudaSetDevice(c->device);
    try {
        dxsql::Parser ps(expr);
        dxsql::ExprP e = ps.condition();
        BindScope sc; sc.table = t;
        ExprPrograms P; DevPrograms D;
        int rc = dx_compile_expr(c, sc, e, &P); if (rc) return rc;
        if (P.result_type[0] != DX_T_BOOL) return c->set_error(DX_ERR_UNSUPPORTED, std::string("sink filter '") + expr + "' is not a boolean expression");
        rc = dx_upload_programs(c, &P, &D); if (rc) return rc;
        if (t->n_rows == 0) return DX_OK;
        DevBuf fl(c, (size_t)t->n_rows * 4);
        rc = dx_eval_flags(c, D, t->n_rows, nullptr, fl.as<uint32_t>(), 1); if (rc) return rc;
        std::vector<uint32_t> hf((size_t)t->n_rows);
        DX_CUDA(c, cudaMemcpyAsync(hf.data(), fl.p, hf.size() * 4, cudaMemcpyDeviceToHost, c->stream));
        DX_CUDA(c, cudaStreamSynchronize(c->stream));
        for (int64_t i = 0; i < t->n_rows; i++) flags[i] = (uint8_t)(hf[(size_t)i] & 1u);
        return DX_OK;
    } catch (std::exception& ex) { return c->set_error(DX_ERR_PARSE, std::string("sink filter: ") + ex.what()); }
}

// ---- flow ------------------------------------------------------------------------------------
int dx_flow_create(dx_ctx* c, const char* conf_text, const char* schema_json, const char* const* projection_steps, int n_steps,
                   const char* transform_text, dx_flow** out) {
    if (!c || !schema_json || !out || n_steps < 0) return DX_ERR_ARG;
    cudaSetDevice(c->device);
    try { return dx_flow_build(c, conf_text, schema_json, projection_steps, n_steps, transform_text, out); }
    catch (std::exception& e) { return c->set_error(DX_ERR_PARSE, e.what()); }
}
void dx_flow_destroy(dx_flow* f) { delete f; }
int dx_flow_process_batch(dx_flow* f, const char* name_space, const uint8_t* json, int64_t n_bytes, const int64_t* offsets, int64_t n_events,
                          int memory_kind, int64_t batch_time_ms, int64_t interval_ms, int64_t output_partition_time_ms, int64_t now_ms, dx_result** out) {
    if (!f || !out || (n_bytes > 0 && !json)) return DX_ERR_ARG;
    (void)output_partition_time_ms;   // consumed by the sinks on the host side of the boundary
    cudaSetDevice(f->ctx->device);
    try { return f->process(name_space ? name_space : "", json, n_bytes, offsets, n_events, memory_kind, batch_time_ms, interval_ms, now_ms, out); }
    catch (std::exception& e) { return f->ctx->set_error(DX_ERR_STATE, e.what()); }
}
int dx_flow_set_event_property(dx_flow* f, int which, const char* key, const uint8_t* bytes, const int64_t* offsets, const uint8_t* valid, int64_t n_events) {
    if (!f) return DX_ERR_ARG;
    cudaSetDevice(f->ctx->device);
    try { return f->set_event_property(which, key, bytes, offsets, valid, n_events); } catch (std::exception& e) { return f->ctx->set_error(DX_ERR_ARG, e.what()); }
}
int dx_flow_set_batch_info(dx_flow* f, const dx_batch_info* info) {
    if (!f) return DX_ERR_ARG;
    dx_flow::BatchInfo b;
    if (info) {
        auto set = [](const char* src, bool* has, std::string* dst) { *has = src != nullptr; if (src) *dst = src; };
        set(info->input_path, &b.has_input_path, &b.input_path);
        set(info->output_file_name, &b.has_output_file_name, &b.output_file_name);
        set(info->rule_index_prefix, &b.has_rule_index_prefix, &b.rule_index_prefix);
        set(info->target, &b.has_target, &b.target);
        b.has_file_time = info->has_file_time != 0; b.file_time_ms = info->file_time_ms;
        b.has_output_folders = info->n_output_folders >= 0;
        for (int i = 0; i < info->n_output_folders; i++) b.output_folders.emplace_back(info->output_folder_keys[i], info->output_folder_values[i]);
        for (int i = 0; i < info->n_properties; i++) b.properties.emplace_back(info->property_keys[i], info->property_values[i]);
        for (int i = 0; i < info->n_system_properties; i++) b.system_properties.emplace_back(info->system_property_keys[i], info->system_property_values[i]);
        if (info->executor_id) b.executor_id = info->executor_id;
    }
    f->batch_info = b;
    return DX_OK;
}

// ---- ingest running ahead of the processor -------------------------------------------------------
struct dx_staged { dx_ctx* ctx = nullptr; DevBuf buf; int64_t n_bytes = 0; cudaEvent_t done = nullptr; };
int dx_flow_stage_batch(dx_flow* f, const uint8_t* json_host, int64_t n_bytes, dx_staged** out) {
    if (!f || !out || n_bytes < 0 || (n_bytes > 0 && !json_host)) return DX_ERR_ARG;
    dx_ctx* c = f->ctx;
    cudaSetDevice(c->device);
    if (!c->copy_stream) DX_CUDA(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    std::unique_ptr<dx_staged> st(new dx_staged());
    st->ctx = c; st->n_bytes = n_bytes;
    st->buf = DevBuf(c, (size_t)n_bytes + 64);
    if (!st->buf.p) return c->set_error(DX_ERR_CUDA, "dx_flow_stage_batch: device allocation failed");
    DX_CUDA(c, cudaEventCreateWithFlags(&st->done, cudaEventDisableTiming));
    if (n_bytes) DX_CUDA(c, cudaMemcpyAsync(st->buf.p, json_host, (size_t)n_bytes, cudaMemcpyHostToDevice, c->copy_stream));
    DX_CUDA(c, cudaEventRecord(st->done, c->copy_stream));
    *out = st.release();
    return DX_OK;
}
void dx_staged_release(dx_staged* st) {
    if (!st) return;
    cudaSetDevice(st->ctx->device);
    if (st->done) { cudaEventSynchronize(st->done); cudaEventDestroy(st->done); }      // the copy must not outlive its target
    delete st;
}
int dx_flow_process_staged(dx_flow* f, const char* name_space, dx_staged* st, int64_t n_events, int64_t batch_time_ms, int64_t interval_ms,
                           int64_t output_partition_time_ms, int64_t now_ms, dx_result** out) {
    if (!f || !out || !st || st->ctx != f->ctx) return DX_ERR_ARG;
    (void)output_partition_time_ms;
    dx_ctx* c = f->ctx;
    cudaSetDevice(c->device);
    DX_CUDA(c, cudaStreamWaitEvent(c->stream, st->done, 0));      // the processor's stream waits on the device; the host does not
    int rc;
    try { rc = f->process(name_space ? name_space : "", st->buf.as<uint8_t>(), st->n_bytes, nullptr, n_events, DX_MEM_DEVICE, batch_time_ms, interval_ms, now_ms, out); }
    catch (std::exception& e) { rc = c->set_error(DX_ERR_STATE, e.what()); }
    cudaStreamSynchronize(c->stream);       // every reader of the staged bytes is done before the buffer returns to the pool
    dx_staged_release(st);
    return rc;
}
int dx_result_metrics(dx_result* r, const char* const** keys, const double** vals, int* n) {
    *keys = r->metric_key_ptrs.data(); *vals = r->metric_vals.data(); *n = (int)r->metric_vals.size();
    return DX_OK;
}
int dx_result_num_tables(const dx_result* r) { return (int)r->dataset_order.size(); }
const char* dx_result_table_name(const dx_result* r, int i) { return (i >= 0 && i < (int)r->dataset_order.size()) ? r->dataset_order[i].c_str() : nullptr; }
int dx_result_table(dx_result* r, const char* dataset, dx_table** out) {
    auto it = r->datasets.find(dataset);
    if (it == r->datasets.end()) return r->flow->ctx->set_error(DX_ERR_ARG, std::string("could not find data set name '") + dataset + "'");
    *out = dx_wrap_table(it->second);
    return DX_OK;
}
void dx_result_release(dx_result* r) { delete r; }
int dx_flow_state_export(dx_flow* f, const char* table, dx_table** out, int* active_is_b) {
    auto it = f->state_tables.find(table);
    if (it == f->state_tables.end()) return f->ctx->set_error(DX_ERR_ARG, std::string("no state table '") + table + "'");
    *out = dx_wrap_table(it->second.phys[it->second.active]);
    if (active_is_b) *active_is_b = it->second.active;
    return DX_OK;
}

int dx_flow_state_import(dx_flow* f, const char* table, int64_t n_rows, int n_columns, const void* const* column_data,
                         const uint8_t* const* column_valid, const uint8_t* const* string_bytes, int active_is_b) {
    if (!f || !table || n_rows < 0 || (n_rows > 0 && !column_data)) return DX_ERR_ARG;
    dx_ctx* c = f->ctx;
    cudaSetDevice(c->device);
    auto it = f->state_tables.find(table);
    if (it == f->state_tables.end()) return c->set_error(DX_ERR_ARG, std::string("no state table '") + table + "'");
    StateTable& st = it->second;
    if (n_columns != (int)st.schema.size())
        return c->set_error(DX_ERR_ARG, "state table '" + st.name + "' has " + std::to_string(st.schema.size()) + " columns, " + std::to_string(n_columns) + " given");
    auto t = std::make_shared<DxTable>(); t->ctx = c; t->n_rows = n_rows;
    const size_t nr = (size_t)(n_rows > 0 ? n_rows : 1);
    // one 32-bit validity word per row, bit k for column k (more than 32 columns: a second array)
    std::vector<DevBufP> masks((st.schema.size() + 31) / 32);
    std::vector<std::vector<uint32_t>> hmask(masks.size(), std::vector<uint32_t>(nr, 0));
    for (size_t k = 0; k < st.schema.size(); k++) {
        DxColumn col; col.name = st.schema[k].first; col.type = st.schema[k].second;
        const int w = dx_type_width(col.type);
        col.data = make_buf(c, nr * w);
        if (!col.data->p) return c->set_error(DX_ERR_CUDA, "state import allocation failed");
        if (n_rows > 0 && !column_data[k]) return c->set_error(DX_ERR_ARG, "state import: column '" + col.name + "' has no data");
        const uint8_t* valid = column_valid ? column_valid[k] : nullptr;
        for (int64_t r = 0; r < n_rows; r++) if (!valid || valid[r]) hmask[k / 32][r] |= 1u << (k % 32);
        if (col.type == DX_T_STR) {
            if (n_rows > 0 && (!string_bytes || !string_bytes[k])) return c->set_error(DX_ERR_ARG, "state import: string column '" + col.name + "' needs its bytes");
            col.dict = std::make_shared<DxDict>();
            int rc = col.dict->init(c, 1u << 16, 1u << 22); if (rc) return rc;
            const int64_t* off = reinterpret_cast<const int64_t*>(column_data[k]);
            std::vector<int32_t> codes(nr, -1);
            std::map<std::string, int32_t> seen;
            for (int64_t r = 0; r < n_rows; r++) {
                if (valid && !valid[r]) continue;
                if (off[r + 1] < off[r]) return c->set_error(DX_ERR_ARG, "state import: offsets of '" + col.name + "' are not ascending");
                std::string sv(reinterpret_cast<const char*>(string_bytes[k]) + off[r], (size_t)(off[r + 1] - off[r]));
                auto hit = seen.find(sv);
                if (hit == seen.end()) { int32_t code; rc = col.dict->insert_host(sv, &code); if (rc) return rc; hit = seen.emplace(sv, code).first; }
                codes[r] = hit->second;
            }
            DX_CUDA(c, cudaMemcpyAsync(col.data->p, codes.data(), nr * 4, cudaMemcpyHostToDevice, c->stream));
            DX_CUDA(c, cudaStreamSynchronize(c->stream));
        } else if (n_rows > 0) {
            DX_CUDA(c, cudaMemcpyAsync(col.data->p, column_data[k], nr * w, cudaMemcpyHostToDevice, c->stream));
            DX_CUDA(c, cudaStreamSynchronize(c->stream));
        }
        col.vwidth = 4; col.vbit = (int)(k % 32);
        t->cols.push_back(col);
    }
    for (size_t m = 0; m < masks.size(); m++) {
        masks[m] = make_buf(c, nr * 4);
        if (!masks[m]->p) return c->set_error(DX_ERR_CUDA, "state import allocation failed");
        DX_CUDA(c, cudaMemcpyAsync(masks[m]->p, hmask[m].data(), nr * 4, cudaMemcpyHostToDevice, c->stream));
    }
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    for (size_t k = 0; k < t->cols.size(); k++) t->cols[k].vmask = masks[k / 32];
    st.active = active_is_b ? 1 : 0;
    st.phys[st.active] = t;
    return DX_OK;
}

int dx_flow_add_refdata(dx_flow* f, const char* name, const char* csv_text, char delimiter, int has_header) {
    if (!f || !name || !csv_text) return DX_ERR_ARG;
    cudaSetDevice(f->ctx->device);
    return dx_flow_add_refdata_impl(f, name, csv_text, delimiter ? delimiter : ',', has_header != 0);
}
int dx_flow_register_udf(dx_flow* f, const char* name, dx_udf1_fn fn, void* user) {
    if (!f || !name || !fn) return DX_ERR_ARG;
    dx_flow::Udf u; u.fn = fn; u.user = user; u.n_args = 1;
    f->udfs[dxsql::upper(name)] = u;
    return DX_OK;
}
int dx_flow_register_udfn(dx_flow* f, const char* name, int n_args, dx_udfn_fn fn, void* user, int refresh_each_batch) {
    if (!f || !name || !fn) return DX_ERR_ARG;
    if (n_args < 0 || n_args > 3) return f->ctx->set_error(DX_ERR_UNSUPPORTED, "user functions take 0 to 3 arguments (DynamicUDF.Generator0..3)");
    dx_flow::Udf u; u.fnn = fn; u.user = user; u.n_args = n_args; u.refresh = refresh_each_batch != 0;
    f->udfs[dxsql::upper(name)] = u;
    return DX_OK;
}
int dx_flow_set_interval_hook(dx_flow* f, dx_interval_fn fn, void* user) {
    if (!f) return DX_ERR_ARG;
    f->on_interval = fn; f->on_interval_user = user;
    return DX_OK;
}
int dx_flow_set_normalizer(dx_flow* f, dx_normalizer_fn fn, void* user) {
    if (!f) return DX_ERR_ARG;
    f->normalizer = fn; f->normalizer_user = user;
    return DX_OK;
}

// ---- multi-GPU: projection only, route only, key-hash shuffle ------------------------------------
int dx_flow_project(dx_flow* f, const uint8_t* json, int64_t n_bytes, const int64_t* offsets, int64_t n_events, int memory_kind,
                    int64_t now_ms, dx_table** out) {
    if (!f || !out || (n_bytes > 0 && !json)) return DX_ERR_ARG;
    cudaSetDevice(f->ctx->device);
    DxTableP t;
    try { int rc = f->project(json, n_bytes, offsets, n_events, memory_kind, now_ms, &t); if (rc) return rc; }
    catch (std::exception& e) { return f->ctx->set_error(DX_ERR_STATE, e.what()); }
    *out = dx_wrap_table(t);
    return DX_OK;
}
int dx_flow_process_table(dx_flow* f, const char* name_space, const dx_table* projected, int64_t batch_time_ms, int64_t interval_ms,
                          int64_t now_ms, dx_result** out) {
    if (!f || !projected || !out) return DX_ERR_ARG;
    cudaSetDevice(f->ctx->device);
    try { return f->route(name_space ? name_space : "", dx_unwrap_table(projected), batch_time_ms, interval_ms, now_ms, out); }
    catch (std::exception& e) { return f->ctx->set_error(DX_ERR_STATE, e.what()); }
}
static int shuffle_columns(const DxTable* t, const char* const* cols, int n_cols, const char* const* key_cols, int n_keys,
                           std::vector<int>* idx_out, std::vector<int>* kpos_out);

int dx_table_shuffle_count(const dx_table* h, const char* const* cols, int n_cols, const char* const* key_cols, int n_keys, int world, int64_t* counts) {
    const DxTable* t = T(h);
    cudaSetDevice(t->ctx->device);
    if (!counts) return t->ctx->set_error(DX_ERR_ARG, "shuffle_count: bad arguments");
    std::vector<int> idx, kpos;
    int rc = shuffle_columns(t, cols, n_cols, key_cols, n_keys, &idx, &kpos); if (rc) return rc;
    return dx_shuffle_count(t->ctx, t, idx, kpos, world, counts);
}
int dx_table_shuffle_put(const dx_table* h, const char* const* cols, int n_cols, const char* const* key_cols, int n_keys, int world,
                         void* const* peer_bufs, const int64_t* first_record) {
    const DxTable* t = T(h);
    cudaSetDevice(t->ctx->device);
    if (!peer_bufs || !first_record) return t->ctx->set_error(DX_ERR_ARG, "shuffle_put: bad arguments");
    std::vector<int> idx, kpos;
    int rc = shuffle_columns(t, cols, n_cols, key_cols, n_keys, &idx, &kpos); if (rc) return rc;
    return dx_shuffle_put(t->ctx, t, idx, kpos, world, peer_bufs, first_record);
}

int dx_table_shuffle_put_blocks(const dx_table* h, const char* const* cols, int n_cols, const char* const* key_cols, int n_keys, int world,
                                void* const* peer_bufs, const int64_t* first_record, const int64_t* block_records, int64_t* sent) {
    const DxTable* t = T(h);
    cudaSetDevice(t->ctx->device);
    if (!peer_bufs || !first_record || !block_records || !sent) return t->ctx->set_error(DX_ERR_ARG, "shuffle_put_blocks: bad arguments");
    std::vector<int> idx, kpos;
    int rc = shuffle_columns(t, cols, n_cols, key_cols, n_keys, &idx, &kpos); if (rc) return rc;
    return dx_shuffle_put(t->ctx, t, idx, kpos, world, peer_bufs, first_record, block_records, sent);
}
int dx_ipc_export(dx_ctx* c, const void* dev_ptr, uint8_t* handle64) {
    if (!c || !dev_ptr || !handle64) return DX_ERR_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaSetDevice(c->device);
    cudaIpcMemHandle_t hnd;
    DX_CUDA(c, cudaIpcGetMemHandle(&hnd, const_cast<void*>(dev_ptr)));
    memcpy(handle64, &hnd, 64);
    return DX_OK;
}
int dx_ipc_open(dx_ctx* c, const uint8_t* handle64, void** dev_ptr) {
    if (!c || !handle64 || !dev_ptr) return DX_ERR_ARG;
    cudaSetDevice(c->device);
    cudaIpcMemHandle_t hnd; memcpy(&hnd, handle64, 64);
    DX_CUDA(c, cudaIpcOpenMemHandle(dev_ptr, hnd, cudaIpcMemLazyEnablePeerAccess));
    return DX_OK;
}
int dx_ipc_close(dx_ctx* c, void* dev_ptr) {
    if (!c || !dev_ptr) return DX_ERR_ARG;
    cudaSetDevice(c->device);
    DX_CUDA(c, cudaIpcCloseMemHandle(dev_ptr));
    return DX_OK;
}

static int shuffle_columns(const DxTable* t, const char* const* cols, int n_cols, const char* const* key_cols, int n_keys,
                           std::vector<int>* idx_out, std::vector<int>* kpos_out) {
    dx_ctx* c = t->ctx;
    if (n_cols < 1 || n_keys < 0 || n_keys > 8) return c->set_error(DX_ERR_ARG, "shuffle: bad arguments");
    BindScope sc; sc.table = t;
    std::vector<int> idx, kpos;
    for (int i = 0; i < n_cols; i++) {
        std::string err; std::vector<std::string> parts;
        std::string nm = cols[i]; size_t p0 = 0;
        while (true) { size_t d = nm.find('.', p0); if (d == std::string::npos) { parts.push_back(nm.substr(p0)); break; } parts.push_back(nm.substr(p0, d - p0)); p0 = d + 1; }
        int ci = dx_resolve_column(sc, parts, &err);
        if (ci < 0) return c->set_error(DX_ERR_ARG, "shuffle: " + err);
        idx.push_back(ci);
    }
    for (int k = 0; k < n_keys; k++) {
        int pos = -1;
        for (int i = 0; i < n_cols; i++) if (std::string(cols[i]) == key_cols[k]) pos = i;
        if (pos < 0) return c->set_error(DX_ERR_ARG, std::string("shuffle: key column '") + key_cols[k] + "' is not among the sent columns");
        kpos.push_back(pos);
    }
    *idx_out = idx; *kpos_out = kpos;
    return DX_OK;
}

int dx_table_shuffle_pack(const dx_table* h, const char* const* cols, int n_cols, const char* const* key_cols, int n_keys, int world,
                          void* packed_dev, int64_t* counts, int64_t* rec_bytes) {
    const DxTable* t = T(h);
    dx_ctx* c = t->ctx;
    cudaSetDevice(c->device);
    if (n_cols < 1 || n_keys < 0 || n_keys > 8 || !counts) return c->set_error(DX_ERR_ARG, "shuffle_pack: bad arguments");
    BindScope sc; sc.table = t;
    std::vector<int> idx, kpos;
    for (int i = 0; i < n_cols; i++) {
        std::string err; std::vector<std::string> parts;
        std::string nm = cols[i]; size_t p0 = 0;
        while (true) { size_t d = nm.find('.', p0); if (d == std::string::npos) { parts.push_back(nm.substr(p0)); break; } parts.push_back(nm.substr(p0, d - p0)); p0 = d + 1; }
        int ci = dx_resolve_column(sc, parts, &err);
        if (ci < 0) return c->set_error(DX_ERR_ARG, "shuffle_pack: " + err);
        idx.push_back(ci);
    }
    for (int k = 0; k < n_keys; k++) {
        int pos = -1;
        for (int i = 0; i < n_cols; i++) if (std::string(cols[i]) == key_cols[k]) pos = i;
        if (pos < 0) return c->set_error(DX_ERR_ARG, std::string("shuffle_pack: key column '") + key_cols[k] + "' is not among the sent columns");
        kpos.push_back(pos);
    }
    if (rec_bytes) *rec_bytes = 8 * (int64_t)(n_cols + 1);
    if (t->n_rows > 0 && !packed_dev) return c->set_error(DX_ERR_ARG, "shuffle_pack: null output buffer");
    return dx_shuffle_pack(c, t, idx, kpos, world, packed_dev, counts);
}
int dx_table_from_packed(dx_ctx* c, const char* const* col_names, const int* col_types, int n_cols, const void* packed_dev, int64_t n_rows,
                         dx_table** out) {
    if (!c || !out || n_cols < 1 || n_rows < 0) return DX_ERR_ARG;
    cudaSetDevice(c->device);
    std::vector<std::string> names; std::vector<int> types;
    for (int i = 0; i < n_cols; i++) { names.push_back(col_names[i]); types.push_back(col_types[i]); }
    DxTableP t;
    int rc = dx_shuffle_unpack(c, names, types, packed_dev, n_rows, &t);
    if (rc) return rc;
    *out = dx_wrap_table(t);
    return DX_OK;
}

int dx_table_from_packed_blocks(dx_ctx* c, const char* const* col_names, const int* col_types, int n_cols, const void* packed_dev,
                                int n_blocks, const int64_t* block_first, const int64_t* block_count, dx_table** out) {
    if (!c || !out || n_cols < 1 || n_blocks < 1 || !block_first || !block_count) return DX_ERR_ARG;
    cudaSetDevice(c->device);
    std::vector<std::string> names; std::vector<int> types;
    for (int i = 0; i < n_cols; i++) { names.push_back(col_names[i]); types.push_back(col_types[i]); }
    DxTableP t;
    int rc = dx_shuffle_unpack(c, names, types, packed_dev, 0, &t, n_blocks, block_first, block_count);
    if (rc) return rc;
    *out = dx_wrap_table(t);
    return DX_OK;
}

// ---- host-side parsers -------------------------------------------------------------------------
static int put_text(const std::string& s, char* buf, int64_t cap) {
    if (cap <= 0) return DX_ERR_ARG;
    size_t n = std::min<size_t>(s.size(), (size_t)cap - 1);
    memcpy(buf, s.data(), n); buf[n] = 0;
    return DX_OK;
}
int dx_sql_canonical(const char* sql, char* buf, int64_t cap) {
    try { dxsql::Parser p(sql); auto st = p.statement(); return put_text(dxsql::to_string(st), buf, cap); }
    catch (std::exception& e) { put_text(e.what(), buf, cap); return DX_ERR_PARSE; }
}
int dx_transform_split(const char* text, char* buf, int64_t cap) {
    ParsedTransform pt; std::string err;
    int rc = dx_parse_transform(text, &pt, &err);
    if (rc) { put_text(err, buf, cap); return rc; }
    std::string out;
    for (auto& c : pt.commands) out += c.name + "\t" + (c.is_query ? "Query" : "Command") + "\t" + c.text + "\n";
    for (auto& kv : pt.view_reference_count) out += "#ref\t" + kv.first + "\t" + std::to_string(kv.second) + "\n";
    return put_text(out, buf, cap);
}
int dx_string_to_timestamp(const char* s, int64_t* us) {
    long long v = 0;
    if (!s || !dx_string_to_timestamp_us(s, &v)) return 0;
    *us = v; return 1;
}
int64_t dx_duration_seconds(const char* s) { bool ok = false; long long v = dx_parse_duration_seconds(s ? s : "", &ok); return ok ? v : -1; }

// ---- generator -----------------------------------------------------------------------------
int dx_gen_events(dx_ctx* c, const dx_gen_params* p, int64_t first, int64_t n, uint8_t** bytes, int64_t* n_bytes, int64_t** offsets) {
    if (!c || !p || !bytes || !n_bytes) return DX_ERR_ARG;
    cudaSetDevice(c->device);
    return dx_gen_device(c, p, first, n, bytes, n_bytes, offsets);
}

} // extern "C"
