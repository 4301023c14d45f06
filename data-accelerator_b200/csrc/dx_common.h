// dx_common.h — shared host/device definitions for libdxgpu (sm_100a only).
#pragma once
#include <nvtx3/nvToolsExt.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include <map>
#include <memory>
#include <mutex>

#define DX_OK 0
#define DX_ERR_CUDA 1
#define DX_ERR_ARG 2
#define DX_ERR_UNSUPPORTED 3   // construct outside the supported subset: fail loudly, never a CPU path
#define DX_ERR_PARSE 4
#define DX_ERR_CAPACITY 5
#define DX_ERR_STATE 6

// column physical types
enum DxType : int32_t {
    DX_T_I64 = 1,   // long/int/short/byte
    DX_T_F64 = 2,   // double/float
    DX_T_STR = 3,   // int32 dictionary code
    DX_T_BOOL = 4,  // stored as int64 0/1
    DX_T_TS = 5,    // timestamp, int64 microseconds since epoch (UTC)
    DX_T_FLAGS = 6, // rule flag bitmap: words_per_row uint32 per row
    DX_T_NULL = 7,  // untyped NULL literal column
};

static inline int dx_type_width(int t) { return t == DX_T_STR ? 4 : 8; }

struct dx_ctx;

// Device memory pool.  Memory comes from the driver in large slabs (each at least half of what is already held, so a
// growing working set — a window cache filling up over its first hour of batches — costs O(log) cudaMalloc calls, not several
// per micro-batch) and is handed out best-fit from an address-ordered free list with splitting and coalescing; a steady-state
// micro-batch loop performs no allocation calls at all.  Blocks for the caller (dx_dev_alloc) are dedicated allocations, so
// their addresses are allocation bases (what CUDA IPC exports).
struct DxPool {
    std::map<char*, size_t> slabs;                  // base -> bytes, as obtained from cudaMalloc
    std::map<char*, size_t> free_by_addr;           // free blocks; neighbours inside one slab are merged
    std::multimap<size_t, char*> free_by_size;
    std::map<void*, size_t> live;
    size_t bytes_reserved = 0;
    std::mutex mu;
    int device = 0;
    void* alloc(size_t bytes, bool dedicated = false);
    void release(void* p);
    void trim();                                    // returns every slab that is entirely free to the driver
    ~DxPool();
private:
    void add_free(char* p, size_t n);
    void take_free(std::map<char*, size_t>::iterator it);
    char* new_slab(size_t bytes);
};

struct dx_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;            // created on first use: host->device copies that run ahead of the processor (dx_flow_stage_batch)
    DxPool pool;
    std::string last_error;
    int sm_count = 148;
    // stage timing (CUDA events on this ctx's stream)
    bool timing = false;
    std::map<std::string, std::vector<float>> stage_ms;
    std::map<std::string, long long> launches;
    std::map<std::string, long long> counters;     // work counters (rows visited by K3, rows moved by K4 ...)
    size_t k3p_scatter_smem_configured = 0;
    size_t k5_put_smem_configured = 0;
    size_t k7_smem_configured = 0;                  // ... k7_inflate ...
    size_t k6_smem_configured = 0;                  // dynamic shared memory k6_json_write is configured for on this device
    size_t k3p_smem_configured = 0;                 // dynamic shared memory k3p_aggregate is configured for on this device
    int set_error(int code, const std::string& msg) { last_error = msg; return code; }
};

#define DX_CUDA(ctx, call)                                                                      \
    do {                                                                                        \
        cudaError_t _e = (call);                                                                \
        if (_e != cudaSuccess) {                                                                \
            char _b[512];                                                                       \
            snprintf(_b, sizeof _b, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e),    \
                     __FILE__, __LINE__);                                                       \
            return (ctx)->set_error(DX_ERR_CUDA, _b);                                           \
        }                                                                                       \
    } while (0)

// RAII handle for pooled device memory.
struct DevBuf {
    dx_ctx* ctx = nullptr;
    void* p = nullptr;
    size_t bytes = 0;
    DevBuf() {}
    DevBuf(dx_ctx* c, size_t n) : ctx(c), bytes(n) { p = n ? c->pool.alloc(n) : nullptr; }
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept { ctx = o.ctx; p = o.p; bytes = o.bytes; o.p = nullptr; o.bytes = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        if (this != &o) { reset(); ctx = o.ctx; p = o.p; bytes = o.bytes; o.p = nullptr; o.bytes = 0; }
        return *this;
    }
    void reset() { if (p && ctx) ctx->pool.release(p); p = nullptr; bytes = 0; }
    ~DevBuf() { reset(); }
    template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};
typedef std::shared_ptr<DevBuf> DevBufP;
static inline DevBufP make_buf(dx_ctx* c, size_t n) { return std::make_shared<DevBuf>(c, n); }

// ---------------------------------------------------------------------------------------
// String dictionary (one per string column lineage).  Device-resident open-addressing
// table; codes are dense int32 assigned on first sight.  Strings of <= 16 bytes are
// identified exactly by their packed bytes (w0,w1,len); longer ones by a 64-bit hash plus
// a byte compare against the heap.
// ---------------------------------------------------------------------------------------
struct DictSlot {          // 32 bytes = one L2 sector
    uint64_t w0, w1;
    uint32_t len;
    uint32_t heap_off;
    int32_t code;
    uint32_t state;        // 0 empty, 1 being filled, 2 ready
};
struct DictDev {
    DictSlot* slots;
    uint32_t cap_mask;     // capacity - 1 (power of two)
    uint32_t max_codes;
    uint8_t* heap;
    uint32_t heap_cap;
    uint32_t* counters;    // [0] = n_codes, [1] = heap_used, [2] = overflow flag
    uint32_t* code_off;    // [max_codes]
    uint32_t* code_len;    // [max_codes]
};
struct DxDict {
    dx_ctx* ctx;
    DictDev dev;
    DevBuf slots, heap, counters, code_off, code_len;
    // host mirror, refreshed on demand
    std::vector<std::string> host_strings;
    std::map<std::string, int32_t> seeded;   // literals already entered from the host
    int init(dx_ctx* c, uint32_t capacity_pow2, uint32_t heap_bytes);
    int insert_host(const std::string& s, int32_t* code_out);   // pre-seed literals
    int insert_host_many(const std::vector<std::string>& strs, std::vector<int32_t>* codes_out);   // one launch for all of them; grows the table when it is full
    int sync_to_host();                                          // pulls new entries
    // A column of unbounded cardinality (event ids, CAST(timestamp AS STRING)) outgrows any fixed table.  grow() rebuilds the
    // device table four times larger from the host mirror, every string keeping its code, so a streaming flow never dies of a
    // full dictionary; callers re-run the kernel that reported the overflow.  overflowed(): the device counters say so.
    int grow();
    int overflowed(bool* yes);
    int lookup_host(const std::string& s) const;                 // -1 if absent (after sync)
};
typedef std::shared_ptr<DxDict> DxDictP;

// ---------------------------------------------------------------------------------------
// Device-resident columnar table.  A column's validity is bit `vbit` of the per-row mask
// array `vmask` (element width vwidth bytes); several columns may share one mask array
// (the decoder writes one byte per row holding all leaf bits) and forwarded columns keep
// pointing at their producer's buffers (zero copy).
// ---------------------------------------------------------------------------------------
struct DxColumn {
    std::string name;        // dotted path for decoded leaves, alias for derived columns
    int32_t type = 0;
    DevBufP data;            // null => constant column
    DevBufP vmask;           // null => never NULL (or always NULL if all_null)
    int32_t vwidth = 0;      // 1 or 4
    int32_t vbit = 0;
    bool is_const = false;
    bool all_null = false;
    int64_t const_i = 0;     // I64/BOOL/TS/STR(code) constant
    double const_f = 0;
    std::string const_s;     // STR constant text (also entered into dict when one is attached)
    DxDictP dict;            // STR columns
    int32_t flag_words = 0;  // FLAGS columns: uint32 words per row
    std::shared_ptr<void> aux; // FLAGS: the static rule table
    // leaf of a MAP(...) / STRUCT(...) select item: the key path below the item's alias and one
    // 'm' (map) or 's' (struct) per level; the column's name is alias + "." + path joined by "."
    std::vector<std::string> nest_path;
    std::string nest_kinds;
};

struct DxTable {
    dx_ctx* ctx = nullptr;
    std::string name;
    int64_t n_rows = 0;
    std::vector<DxColumn> cols;
    std::vector<std::string> col_owner;   // join results: the table (name or alias) each column came from, for qualified references
    bool has_ts_bounds = false;     // window cache: bounds of the flow's timestamp column over this table's rows
    long long ts_min = 0, ts_max = 0;
    // window cache: rows whose timestamp is NULL or below ts_cut are not part of this cached batch (the late-event drop of
    // CPF:161-167 kept as a predicate instead of a filtered copy); every reader of a cached batch applies it with its time range
    bool has_cut = false; long long ts_cut = 0;
    int find(const std::string& n) const {
        for (size_t i = 0; i < cols.size(); i++) if (cols[i].name == n) return (int)i;
        return -1;
    }
};
typedef std::shared_ptr<DxTable> DxTableP;

// stage timing helper
// NVTX range for the life of the object: every kernel stage (through StageTimer) and every phase of a micro-batch (project,
// route, each statement) shows up by name on a profiler's timeline; without a tool attached the calls are no-ops.
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    explicit NvtxRange(const std::string& name) { nvtxRangePushA(name.c_str()); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete; NvtxRange& operator=(const NvtxRange&) = delete;
};

struct StageTimer {
    dx_ctx* ctx; const char* name; cudaEvent_t a = nullptr, b = nullptr; NvtxRange range;
    StageTimer(dx_ctx* c, const char* n) : ctx(c), name(n), range(n) {
        if (ctx->timing) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, ctx->stream); }
    }
    void stop() {
        if (ctx->timing && a) {
            cudaEventRecord(b, ctx->stream); cudaEventSynchronize(b);
            float ms = 0; cudaEventElapsedTime(&ms, a, b);
            ctx->stage_ms[name].push_back(ms);
            cudaEventDestroy(a); cudaEventDestroy(b); a = b = nullptr;
        }
    }
    ~StageTimer() { stop(); }
};
