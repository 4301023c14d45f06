// dx_decode.cu — K1: NDJSON / offset-framed JSON events -> columnar buffers (sm_100a).
//
// Replaces from_json(inputNormalizerUdf(Raw), rawBlockSchema) in project()
// (reference: datax/processor/CommonProcessorFactory.scala:90-103, schema from
// datax/input/SchemaFile.scala:22-34) for the supported leaf types.
//
// Shape of the kernel (one persistent CTA loop, several CTAs per SM so that one CTA's bulk
// load overlaps another's parsing):
//   1. tile t (dynamic index from an atomic counter) = 32 KiB of the byte stream plus an
//      overlap; moved global->shared with cp.async.bulk (TMA bulk copy) completing on an mbarrier;
//   2. warp-parallel structural scan: every lane classifies 16 bytes at a time, newline masks
//      are counted and prefix-summed with warp shuffles -> list of record starts in the tile;
//   3. the tile's record count is published and the tile's first row index obtained by a
//      decoupled look-back over earlier tiles (single pass over the input, no pre-count pass);
//   4. one thread per record: schema-ordered template fast path (word-wise literal compare,
//      digit loops) with fall-through to a general Spark-2.4-conformant parser in the same
//      kernel; lanes hold consecutive rows, so every column store is a coalesced 128/256 B line.
// The kernel template lives in dx_decode_kernel.cuh; this file instantiates the NDJSON kernels and holds the host side,
// dx_decode_framed.cu instantiates the kernels for caller-provided record offsets (step 2 replaced by a search in the offsets).
#include "dx_decode.cuh"
#include "dx_json.h"
#include "dx_decode_host.h"

#include "dx_decode_kernel.cuh"

// Counts records of an NDJSON buffer (used only when the caller does not state n_events).
__global__ void k1_count_records(const uint8_t* bytes, long long n, unsigned long long* out) {
    unsigned long long c = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        c += (bytes[i] == '\n') ? 1 : 0;
    for (int s = 16; s > 0; s >>= 1) c += __shfl_xor_sync(0xffffffffu, c, s);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// =======================================================================================
// host side
// =======================================================================================
static const int k1_tiles[] = {K1_TS(16384), K1_TS(20480), K1_TS(24576), K1_TS(28672), K1_TS(32768), K1_TS(36864), K1_TS(40960), K1_TS(45056), K1_TS(49152)};
// tile bytes for a batch: the largest instantiated size that still keeps the expected records per tile just under the CTA's thread count
static int k1_pick_tile(long long n_bytes, long long n_events) {
    if (n_events <= 0 || n_bytes <= 0) return K1_T_DEFAULT;
    const double avg = (double)n_bytes / (double)n_events;
    int best = k1_tiles[0];
    for (int t : k1_tiles) if ((double)t / avg <= (double)(K1_NT - (K1_NT >= 128 ? 10 : 3))) best = t;
    return best;
}
template <int T> static int k1_setup_one(dx_ctx* c, int* occ_out) {      // occ_out[0]: NDJSON kernel, occ_out[9]: framed kernel
    DX_CUDA(c, cudaFuncSetAttribute(k1_decode<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k1_smem_bytes(T)));
    int occ = 0;
    DX_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k1_decode<T, false>, K1_NT, k1_smem_bytes(T)));
    occ_out[0] = occ > 0 ? occ : 1;
    DX_CUDA(c, (cudaError_t)k1_framed_setup(T, &occ));
    occ_out[9] = occ > 0 ? occ : 1;
    return DX_OK;
}
static int k1_launch(dx_ctx* c, int tile, const int* occ_by_tile, const DecodeArgs& a) {
    int idx = 0; for (int i = 0; i < 9; i++) if (k1_tiles[i] == tile) idx = i;
    const bool framed = a.offsets != nullptr;
    long long grid = (long long)c->sm_count * occ_by_tile[idx + (framed ? 9 : 0)];
    if (grid > a.n_tiles) grid = a.n_tiles;
    const size_t smem = k1_smem_bytes(tile);
    if (framed) { k1_framed_launch(tile, (unsigned)grid, c->stream, a); c->launches["k1_decode"]++; return DX_OK; }
    switch (tile) {
    case K1_TS(16384): k1_decode<K1_TS(16384), false><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case K1_TS(20480): k1_decode<K1_TS(20480), false><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case K1_TS(24576): k1_decode<K1_TS(24576), false><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case K1_TS(28672): k1_decode<K1_TS(28672), false><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case K1_TS(36864): k1_decode<K1_TS(36864), false><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case K1_TS(40960): k1_decode<K1_TS(40960), false><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case K1_TS(45056): k1_decode<K1_TS(45056), false><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case K1_TS(49152): k1_decode<K1_TS(49152), false><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    default: k1_decode<K1_TS(32768), false><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    }
    c->launches["k1_decode"]++;
    return DX_OK;
}

static int map_spark_type(const std::string& t) {
    if (t == "long" || t == "integer" || t == "int" || t == "short" || t == "byte") return DX_T_I64;
    if (t == "double" || t == "float") return DX_T_F64;
    if (t == "string") return DX_T_STR;
    if (t == "boolean") return DX_T_BOOL;
    return 0;
}

struct SchemaBuilder {
    DSchemaDev* S;
    std::vector<std::string>* leaf_names;
    std::string names;
    std::string err;
    // Fast-path template pieces
    std::vector<std::string> seg_lits; std::vector<int> seg_types; std::vector<int> seg_leaves;
    std::string pending;

    // Reserves the struct entry first so children structs get higher indices; fields of one
    // struct are contiguous.
    int add_struct(const JVal& st, const std::string& prefix) {
        const JVal* fields = st.get("fields");
        if (!fields || fields->kind != JVal::ARR) { err = "struct without fields"; return -1; }
        if (S->n_structs >= DX_MAX_STRUCTS) { err = "too many nested structs"; return -1; }
        int sidx = S->n_structs++;
        int nf = (int)fields->arr.size();
        if (S->n_fields + nf > DX_MAX_FIELDS) { err = "too many fields"; return -1; }
        int first = S->n_fields; S->n_fields += nf;
        S->structs[sidx].first_field = first; S->structs[sidx].n_fields = nf;
        S->structs[sidx].first_leaf = S->n_leaves;
        pending += "{";
        for (int i = 0; i < nf; i++) {
            const JVal& f = fields->arr[i];
            const JVal* nm = f.get("name"); const JVal* ty = f.get("type");
            if (!nm || !ty || nm->kind != JVal::STR) { err = "field without name/type"; return -1; }
            DField& df = S->fields[first + i];
            if (names.size() + nm->str.size() > DX_NAMES_BYTES) { err = "field names too long"; return -1; }
            df.name_off = (uint16_t)names.size(); df.name_len = (uint16_t)nm->str.size();
            names += nm->str;
            pending += (i ? ",\"" : "\"") + nm->str + "\":";
            if (ty->kind == JVal::STR) {
                int t = map_spark_type(ty->str);
                if (!t) { err = "unsupported field type '" + ty->str + "' for " + prefix + nm->str; return -1; }
                if (S->n_leaves >= DX_MAX_LEAVES) { err = "more than 31 leaf fields"; return -1; }
                df.type = t; df.target = S->n_leaves;
                S->leaf_type[S->n_leaves] = t;
                leaf_names->push_back(prefix + nm->str);
                seg_lits.push_back(pending); seg_types.push_back(t); seg_leaves.push_back(S->n_leaves);
                pending.clear();
                S->n_leaves++;
            } else if (ty->kind == JVal::OBJ && ty->get("type") && ty->get("type")->str == "struct") {
                df.type = 100;
                int child = add_struct(*ty, prefix + nm->str + ".");
                if (child < 0) return -1;
                df.target = child;
            } else { err = "unsupported field type (array/map) for " + prefix + nm->str; return -1; }
        }
        pending += "}";
        S->structs[sidx].n_leaves = S->n_leaves - S->structs[sidx].first_leaf;
        return sidx;
    }
};

int dx_decoder::init(dx_ctx* c, const std::string& schema_json) {
    ctx = c;
    memset(&hs, 0, sizeof hs);
    JVal root;
    try { root = dx_parse_json(schema_json); } catch (std::exception& e) { return c->set_error(DX_ERR_PARSE, e.what()); }
    SchemaBuilder b; b.S = &hs; b.leaf_names = &leaf_names;
    if (b.add_struct(root, "") < 0) return c->set_error(DX_ERR_UNSUPPORTED, "input schema: " + b.err);
    memcpy(hs.names, b.names.data(), b.names.size());
    // template: literals before each leaf value, then the closing braces
    b.seg_lits.push_back(b.pending); b.seg_types.push_back(0); b.seg_leaves.push_back(0);
    size_t lit_used = 0; hs.n_segs = 0;
    bool tmpl_ok = b.seg_lits.size() <= DX_MAX_SEGS;
    for (size_t i = 0; tmpl_ok && i < b.seg_lits.size(); i++) {
        size_t L = b.seg_lits[i].size(), Lp = (L + 3) & ~size_t(3);
        if (lit_used + Lp > DX_LITS_BYTES) { tmpl_ok = false; break; }
        memcpy(reinterpret_cast<char*>(hs.lits) + lit_used, b.seg_lits[i].data(), L);
        if (L == 0) { tmpl_ok = false; break; }
        hs.segs[i].lit_word = (uint32_t)(lit_used >> 2); hs.segs[i].n_words = (uint32_t)(Lp >> 2); hs.segs[i].lit_len = (uint32_t)L;
        hs.segs[i].type = b.seg_types[i]; hs.segs[i].leaf = b.seg_leaves[i];
        hs.segs[i].last_mask = (L & 3) ? ((1u << ((L & 3) << 3)) - 1u) : 0xffffffffu;
        lit_used += Lp; hs.n_segs++;
    }
    if (!tmpl_ok) { hs.n_segs = 1; hs.segs[0] = DSeg{0u, 1u, 4u, 0xffffffffu, 0, 0}; memcpy(hs.lits, "\x01\x01\x01\x01", 4); }
    for (int k = 0; k < hs.n_leaves; k++) leaf_types.push_back(hs.leaf_type[k]);
    dev_schema = DevBuf(c, sizeof(DSchemaDev));
    DX_CUDA(c, cudaMemcpyAsync(dev_schema.p, &hs, sizeof hs, cudaMemcpyHostToDevice, c->stream));
    dicts.resize(hs.n_leaves);
    std::vector<DictDev> dd(hs.n_leaves);
    memset(dd.data(), 0, dd.size() * sizeof(DictDev));
    for (int k = 0; k < hs.n_leaves; k++) if (hs.leaf_type[k] == DX_T_STR) {
        dicts[k] = std::make_shared<DxDict>();
        int rc = dicts[k]->init(c, dict_capacity, dict_heap_bytes);
        if (rc) return rc;
        dd[k] = dicts[k]->dev;
    }
    dev_dicts = DevBuf(c, sizeof(DictDev) * (hs.n_leaves + 1));
    DX_CUDA(c, cudaMemcpyAsync(dev_dicts.p, dd.data(), sizeof(DictDev) * hs.n_leaves, cudaMemcpyHostToDevice, c->stream));
    dev_stats = DevBuf(c, sizeof(DecodeStats) + 16);
    { int rc2;
      if ((rc2 = k1_setup_one<K1_TS(16384)>(c, &occ_by_tile[0])) || (rc2 = k1_setup_one<K1_TS(20480)>(c, &occ_by_tile[1])) || (rc2 = k1_setup_one<K1_TS(24576)>(c, &occ_by_tile[2])) ||
          (rc2 = k1_setup_one<K1_TS(28672)>(c, &occ_by_tile[3])) || (rc2 = k1_setup_one<K1_TS(32768)>(c, &occ_by_tile[4])) || (rc2 = k1_setup_one<K1_TS(36864)>(c, &occ_by_tile[5])) ||
          (rc2 = k1_setup_one<K1_TS(40960)>(c, &occ_by_tile[6])) || (rc2 = k1_setup_one<K1_TS(45056)>(c, &occ_by_tile[7])) || (rc2 = k1_setup_one<K1_TS(49152)>(c, &occ_by_tile[8]))) return rc2; }
    ctas_per_sm = occ_by_tile[4];
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    return DX_OK;
}

// json_dev: device pointer to the first JSON byte.  offsets_dev: optional device array of
// n_events+1 record offsets.  n_events < 0 with NDJSON framing => counted on the device first.
int dx_decoder::decode_device(const uint8_t* json_dev, int64_t n_bytes, const int64_t* offsets_dev,
                              int64_t n_events, DxTableP* out) {
    dx_ctx* c = ctx;
    if (n_bytes < 0) return c->set_error(DX_ERR_ARG, "negative byte count");
    if (n_events < 0) {
        if (offsets_dev) return c->set_error(DX_ERR_ARG, "n_events required with offsets");
        DX_CUDA(c, cudaMemsetAsync(dev_stats.p, 0, sizeof(DecodeStats) + 16, c->stream));
        unsigned long long* cnt = reinterpret_cast<unsigned long long*>(dev_stats.as<uint8_t>() + sizeof(DecodeStats));
        if (n_bytes > 0) {
            k1_count_records<<<c->sm_count * 8, 256, 0, c->stream>>>(json_dev, n_bytes, cnt);
            c->launches["k1_count_records"]++;
        }
        unsigned long long h = 0; uint8_t last = '\n';
        DX_CUDA(c, cudaMemcpyAsync(&h, cnt, 8, cudaMemcpyDeviceToHost, c->stream));
        if (n_bytes > 0) DX_CUDA(c, cudaMemcpyAsync(&last, json_dev + n_bytes - 1, 1, cudaMemcpyDeviceToHost, c->stream));
        DX_CUDA(c, cudaStreamSynchronize(c->stream));
        n_events = (int64_t)h + ((n_bytes > 0 && last != '\n') ? 1 : 0);
    }
    auto t = std::make_shared<DxTable>();
    t->ctx = c; t->n_rows = n_events; t->name = "Raw";
    const int nl = hs.n_leaves;
    const int mw = (nl <= 7) ? 1 : 4;
    DevBufP mask = make_buf(c, (size_t)(n_events > 0 ? n_events : 1) * mw);
    DecodeArgs a; memset(&a, 0, sizeof a);
    a.n_segs = hs.n_segs; memcpy(a.segs, hs.segs, sizeof a.segs); memcpy(a.lits, hs.lits, sizeof a.lits);
    for (int k = 0; k < nl; k++) {
        DxColumn col;
        col.name = leaf_names[k]; col.type = hs.leaf_type[k];
        col.data = make_buf(c, (size_t)(n_events > 0 ? n_events : 1) * dx_type_width(col.type));
        col.vmask = mask; col.vwidth = mw; col.vbit = k;
        col.dict = dicts[k];
        a.cols[k] = col.data->p;
        t->cols.push_back(std::move(col));
    }
    if (n_events == 0 || n_bytes == 0) {
        if (n_events > 0) DX_CUDA(c, cudaMemsetAsync(mask->p, 0, (size_t)n_events * mw, c->stream));
        last_stats = DecodeStats{(unsigned long long)n_events, (unsigned long long)n_events, 0, 0, 0};
        *out = t; return DX_OK;
    }
    uintptr_t addr = reinterpret_cast<uintptr_t>(json_dev);
    a.skew = (uint32_t)(addr & 15);
    a.base_al = reinterpret_cast<const uint8_t*>(addr - a.skew);
    a.n_end = (long long)a.skew + n_bytes;
    const int tile = k1_pick_tile(n_bytes, n_events);
    a.n_tiles = (a.n_end + tile - 1) / tile;
    a.max_rows = n_events;
    a.offsets = reinterpret_cast<const long long*>(offsets_dev);
    a.mask = mask->p; a.mask_width = mw;
    a.dicts = dev_dicts.as<DictDev>();
    a.schema = dev_schema.as<DSchemaDev>();
    size_t need = (size_t)(a.n_tiles + 2) * 8 + 16;
    if (tile_state.bytes < need) tile_state = DevBuf(c, need * 2);
    DX_CUDA(c, cudaMemsetAsync(tile_state.p, 0, need, c->stream));
    a.tile_counter = tile_state.as<unsigned int>();
    a.tile_state = tile_state.as<unsigned long long>() + 2;
    DX_CUDA(c, cudaMemsetAsync(dev_stats.p, 0, sizeof(DecodeStats), c->stream));
    a.stats = dev_stats.as<DecodeStats>();
    // offsets mode writes every row; NDJSON mode too unless the caller overstated n_events — clear the mask
    // so never-written rows read as NULL rows
    {
        StageTimer tm(c, "k1_decode");
        k1_launch(c, tile, occ_by_tile, a);
    }
    DX_CUDA(c, cudaGetLastError());
    DX_CUDA(c, cudaMemcpyAsync(&last_stats, dev_stats.p, sizeof(DecodeStats), cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    if ((last_stats.flags & DXF_DICT_OVERFLOW) && dict_retries < 6) {
        // a string column has more distinct values than its dictionary holds: rebuild the full dictionaries four times larger
        // (every string keeps its code) and decode this batch again — a stream of new values must not stop the flow
        bool grew = false;
        std::vector<DictDev> dd(hs.n_leaves); memset(dd.data(), 0, sizeof(DictDev) * hs.n_leaves);
        for (int k = 0; k < hs.n_leaves; k++) {
            if (!dicts[k]) continue;
            bool full = false; int rc2 = dicts[k]->overflowed(&full); if (rc2) return rc2;
            if (full) { rc2 = dicts[k]->grow(); if (rc2) return rc2; grew = true; }
            dd[k] = dicts[k]->dev;
        }
        if (grew) {
            DX_CUDA(c, cudaMemcpyAsync(dev_dicts.p, dd.data(), sizeof(DictDev) * hs.n_leaves, cudaMemcpyHostToDevice, c->stream));
            DX_CUDA(c, cudaStreamSynchronize(c->stream));
            dict_retries++;
            int rc3 = decode_device(json_dev, n_bytes, offsets_dev, n_events, out);
            dict_retries = 0;
            return rc3;
        }
    }
    if (last_stats.flags & ~(unsigned long long)DXF_ROW_LIMITS) {      // per-record limits NULL the record (limited_rows); only these fail the call
        char b[256];
        snprintf(b, sizeof b, "decode failed: flags=0x%llx (1 dictionary overflow, 16 more records than n_events; "
                 "2 / 4 / 8 only mark records this implementation could not take: %d-byte escaped strings, > 19-digit ties, nested values under a string field)",
                 last_stats.flags, DX_UNESC_CAP);
        return c->set_error((last_stats.flags & DXF_DICT_OVERFLOW) ? DX_ERR_CAPACITY : DX_ERR_UNSUPPORTED, b);
    }
    if ((int64_t)last_stats.rows != n_events) {
        char b[160];
        snprintf(b, sizeof b, "decode: caller stated %lld events but the batch frames %llu records",
                 (long long)n_events, last_stats.rows);
        return c->set_error(DX_ERR_ARG, b);
    }
    *out = t;
    return DX_OK;
}

int dx_decoder::decode_host(const uint8_t* json, int64_t n_bytes, const int64_t* offsets, int64_t n_events, DxTableP* out) {
    dx_ctx* c = ctx;
    size_t need = (size_t)n_bytes + 64;
    if (staging.bytes < need) staging = DevBuf(c, need + need / 4);
    if (n_bytes > 0) DX_CUDA(c, cudaMemcpyAsync(staging.p, json, (size_t)n_bytes, cudaMemcpyHostToDevice, c->stream));
    const int64_t* off_dev = nullptr;
    if (offsets) {
        size_t ob = (size_t)(n_events + 1) * 8;
        if (staging_off.bytes < ob) staging_off = DevBuf(c, ob + ob / 4);
        DX_CUDA(c, cudaMemcpyAsync(staging_off.p, offsets, ob, cudaMemcpyHostToDevice, c->stream));
        off_dev = staging_off.as<int64_t>();
    }
    return decode_device(staging.as<uint8_t>(), n_bytes, off_dev, n_events, out);
}

// =======================================================================================
// Allocation-free decode.
//
// Some hosts of this library size all of their device memory when the job starts and must not
// allocate afterwards: a Spark executor that shares the GPU with other flows cannot afford a
// cudaMalloc in its per-batch path (cudaMalloc/cudaFree synchronise the whole device, stalling
// every other stream, and can fail late when memory is tight), and a latency-bound caller wants a
// per-batch call whose cost never includes the driver's allocator.  For them this variant of K1
// keeps the kernel's scratch — the tile counter, the decode statistics and the per-tile look-back
// states — in one fixed array that lives inside the loaded library image on every GPU that loads
// it (DX_NOALLOC_SCRATCH_BYTES, about 512 KiB, held until the process exits).  The caller passes
// the input bytes and every output buffer; nothing is taken from the pool or the driver.
//
// The array is shared by all calls on a device and keeps the previous call's contents, so each
// call first clears the part it is going to use; two calls whose GPU work overlaps (different
// streams or threads) would race on it — callers serialise.
// =======================================================================================
#define DX_NOALLOC_MAX_TILES 65536                                   /* x 32 KiB tiles = 2 GiB of JSON per call */
#define DX_NOALLOC_HEADER_WORDS 8                                    /* [0] tile counter, [2..6] DecodeStats */
#define DX_NOALLOC_SCRATCH_WORDS (DX_NOALLOC_HEADER_WORDS + DX_NOALLOC_MAX_TILES + 2)
#define DX_NOALLOC_SCRATCH_BYTES (DX_NOALLOC_SCRATCH_WORDS * 8)
static __device__ __align__(256) unsigned long long dx_noalloc_scratch[DX_NOALLOC_SCRATCH_WORDS];

#define DX_NOALLOC_MAX_DEVICES 64
static std::mutex g_noalloc_mu;
static unsigned long long* g_noalloc_dev_ptr[DX_NOALLOC_MAX_DEVICES];   // device address of the array, per GPU

// In host code `dx_noalloc_scratch` is only a handle for the runtime, not a device address; the
// address the kernel needs comes from cudaGetSymbolAddress, asked once per device with that device
// current.  The first call on a device therefore does this one-time set-up (no allocation, but a
// runtime query): make that first call before starting any stream capture.
static int noalloc_scratch_address(dx_ctx* c, unsigned long long** out) {
    if (c->device < 0 || c->device >= DX_NOALLOC_MAX_DEVICES) return c->set_error(DX_ERR_ARG, "device ordinal out of range");
    std::lock_guard<std::mutex> g(g_noalloc_mu);
    if (!g_noalloc_dev_ptr[c->device]) {
        void* p = nullptr;                        // the caller has made c->device current
        cudaError_t e = cudaGetSymbolAddress(&p, dx_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            return c->set_error(DX_ERR_CUDA, std::string("cudaGetSymbolAddress(dx_noalloc_scratch) failed: ") + cudaGetErrorString(e));
        }
        g_noalloc_dev_ptr[c->device] = reinterpret_cast<unsigned long long*>(p);
    }
    *out = g_noalloc_dev_ptr[c->device];
    return DX_OK;
}

int dx_decoder::decode_noalloc(const uint8_t* json_dev, int64_t n_bytes, int64_t n_events, void* const* column_ptrs,
                               int n_columns, void* mask_ptr, int64_t* rows_out, int64_t* null_rows_out) {
    dx_ctx* c = ctx;
    const int nl = hs.n_leaves;
    if (n_bytes < 0 || n_events < 0 || (n_bytes > 0 && !json_dev) || !column_ptrs || !mask_ptr)
        return c->set_error(DX_ERR_ARG, "decode_noalloc: bad arguments (n_events must be stated; NDJSON framing only)");
    if (n_columns != nl) {
        char b[160]; snprintf(b, sizeof b, "decode_noalloc: %d column buffers passed, the schema has %d leaf columns", n_columns, nl);
        return c->set_error(DX_ERR_ARG, b);
    }
    for (int k = 0; k < nl; k++) if (!column_ptrs[k]) return c->set_error(DX_ERR_ARG, "decode_noalloc: null column buffer");
    if (rows_out) *rows_out = 0;
    if (null_rows_out) *null_rows_out = 0;
    const int mw = (nl <= 7) ? 1 : 4;
    if (n_bytes == 0) {
        if (n_events != 0) return c->set_error(DX_ERR_ARG, "decode_noalloc: empty input but n_events > 0");
        return DX_OK;
    }
    DecodeArgs a; memset(&a, 0, sizeof a);
    a.n_segs = hs.n_segs; memcpy(a.segs, hs.segs, sizeof a.segs); memcpy(a.lits, hs.lits, sizeof a.lits);
    uintptr_t addr = reinterpret_cast<uintptr_t>(json_dev);
    a.skew = (uint32_t)(addr & 15);
    a.base_al = reinterpret_cast<const uint8_t*>(addr - a.skew);
    a.n_end = (long long)a.skew + n_bytes;
    const int tile = K1_T_DEFAULT;      // fixed: the scratch array's capacity is stated in tiles of this size
    a.n_tiles = (a.n_end + tile - 1) / tile;
    if (a.n_tiles > DX_NOALLOC_MAX_TILES) {
        char b[256];
        snprintf(b, sizeof b, "decode_noalloc: a batch of %lld bytes needs %lld tiles; this entry point's fixed scratch handles at most %d tiles "
                 "(%lld bytes) per call — split the batch or use dx_decoder_decode", (long long)n_bytes, a.n_tiles, DX_NOALLOC_MAX_TILES,
                 (long long)DX_NOALLOC_MAX_TILES * K1_T_DEFAULT);
        return c->set_error(DX_ERR_CAPACITY, b);
    }
    // run on the context's device and hand the calling thread its own current device back on every way out
    struct DeviceGuard { int prev = -1; ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); } } guard;
    cudaGetDevice(&guard.prev);
    DX_CUDA(c, cudaSetDevice(c->device));
    unsigned long long* scratch = nullptr;
    int rc = noalloc_scratch_address(c, &scratch);
    if (rc) return rc;
    // the array still holds the last call's counter, statistics and tile states: clear what this call reads
    DX_CUDA(c, cudaMemsetAsync(scratch, 0, (size_t)(DX_NOALLOC_HEADER_WORDS + a.n_tiles + 2) * 8, c->stream));
    a.max_rows = n_events;
    a.offsets = nullptr;
    for (int k = 0; k < nl; k++) a.cols[k] = column_ptrs[k];
    a.mask = mask_ptr; a.mask_width = mw;
    a.dicts = dev_dicts.as<DictDev>();
    a.schema = dev_schema.as<DSchemaDev>();
    a.tile_counter = reinterpret_cast<unsigned int*>(scratch);
    a.stats = reinterpret_cast<DecodeStats*>(scratch + 2);
    a.tile_state = scratch + DX_NOALLOC_HEADER_WORDS;
    k1_launch(c, tile, occ_by_tile, a);
    DX_CUDA(c, cudaGetLastError());
    DecodeStats st;
    DX_CUDA(c, cudaMemcpyAsync(&st, a.stats, sizeof st, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    last_stats = st;
    if (st.flags & ~(unsigned long long)DXF_ROW_LIMITS) {
        char b[200];
        snprintf(b, sizeof b, "decode_noalloc failed: flags=0x%llx (1 dictionary overflow, 16 more records than n_events)", st.flags);
        return c->set_error((st.flags & DXF_DICT_OVERFLOW) ? DX_ERR_CAPACITY : DX_ERR_UNSUPPORTED, b);
    }
    if ((int64_t)st.rows != n_events) {
        char b[160];
        snprintf(b, sizeof b, "decode_noalloc: caller stated %lld events but the batch frames %llu records", (long long)n_events, st.rows);
        return c->set_error(DX_ERR_ARG, b);
    }
    if (rows_out) *rows_out = (int64_t)st.rows;
    if (null_rows_out) *null_rows_out = (int64_t)st.null_rows;
    return DX_OK;
}
