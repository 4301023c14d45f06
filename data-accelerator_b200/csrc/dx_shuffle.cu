// dx_shuffle.cu — K5: key-hash partition + pack, and unpack, around the multi-GPU all-to-all.
//
// What it replaces: the Exchange(hashpartitioning(keys)) Spark plans between the partial and final
// HashAggregateExec of every GROUP BY the flow SQL sends through spark.sql
// (CommonProcessorFactory.scala:271).  On one 8xB200 box events are decoded and
// rule-checked where they arrive; rows that feed keyed window aggregates are sent to the GPU that
// owns hash64(group key) % G, so every key's window cache and state live on exactly one GPU.
// The exchange itself is one NCCL all-to-all over NVLink, issued by the host harness
// (torch.distributed.all_to_all_single) on the buffers these kernels fill and read.
//
// Record layout on the wire: n_cols x 8 B values (dictionary-free columns only) + 8 B validity bits.
#include "dx_shuffle.h"

static __device__ __forceinline__ unsigned long long smix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static __device__ __forceinline__ bool sh_load(const ProgCol& c, long long row, long long& v) {
    if (c.vmask) {
        uint32_t m = c.vwidth == 1 ? reinterpret_cast<const uint8_t*>(c.vmask)[row] : reinterpret_cast<const uint32_t*>(c.vmask)[row];
        if (!((m >> c.vbit) & 1u)) { v = 0; return false; }
    }
    v = c.width == 4 ? (long long)reinterpret_cast<const int32_t*>(c.data)[row] : reinterpret_cast<const long long*>(c.data)[row];
    return true;
}
// owner(key) — the same function on every rank: a 64-bit mix over the key values (NULL hashes as a
// fixed constant), reduced to [0, G) by multiply-shift
static __device__ __forceinline__ int owner_of(const ShuffleArgs& a, long long row) {
    unsigned long long h = 0x6A09E667F3BCC909ull;
    for (int k = 0; k < a.n_keys; k++) {
        long long v; bool ok = sh_load(a.cols[a.key_idx[k]], row, v);
        h = smix64(h ^ (ok ? (unsigned long long)v : 0x5BD1E9955BD1E995ull));
    }
    return (int)(((h >> 32) * (unsigned long long)a.world) >> 32);
}

__global__ void __launch_bounds__(256) k5_count(const __grid_constant__ ShuffleArgs a, unsigned long long* counts) {
    __shared__ unsigned int s_cnt[DX_SHUFFLE_MAX_WORLD];
    for (int i = threadIdx.x; i < a.world; i += blockDim.x) s_cnt[i] = 0;
    __syncthreads();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n_rows; i += (long long)gridDim.x * blockDim.x)
        atomicAdd(&s_cnt[owner_of(a, i)], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < a.world; i += blockDim.x) if (s_cnt[i]) atomicAdd(&counts[i], (unsigned long long)s_cnt[i]);
}

// Rows are written grouped by destination; within a destination the order is arbitrary (the
// consumer is a group-by).  One global atomic per (warp, destination) instead of per row.
__global__ void __launch_bounds__(256) k5_scatter(const __grid_constant__ ShuffleArgs a, const unsigned long long* base, unsigned long long* cursor,
                                                  unsigned long long* out) {
    const int rec_words = a.n_cols + 1;
    const int lane = threadIdx.x & 31;
    for (long long i0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) - lane; i0 < a.n_rows; i0 += (long long)gridDim.x * blockDim.x) {
        const long long i = i0 + lane;
        const bool has = i < a.n_rows;
        const int dst = has ? owner_of(a, i) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, dst);
        const int leader = __ffs(peers) - 1;
        const int rank_in_group = __popc(peers & ((1u << lane) - 1u));
        unsigned long long start = 0;
        if (has && lane == leader) start = atomicAdd(&cursor[dst], (unsigned long long)__popc(peers));
        start = __shfl_sync(0xffffffffu, start, leader);
        if (!has) continue;
        unsigned long long* rec = out + (base[dst] + start + rank_in_group) * rec_words;
        unsigned long long valid = 0;
        for (int c = 0; c < a.n_cols; c++) {
            long long v; if (sh_load(a.cols[c], i, v)) valid |= 1ull << c;
            rec[c] = (unsigned long long)v;
        }
        rec[a.n_cols] = valid;
    }
}

// K5 fused with the exchange: the record of every row is written straight into the receive buffer of the GPU that owns
// its key — a peer pointer mapped over NVLink (CUDA IPC) — at the offset this rank was assigned there, so the hash, the
// column gather and the transfer overlap row by row and neither a send buffer nor a separate all-to-all pass exists.
struct PutTargets {
    unsigned long long* base[DX_SHUFFLE_MAX_WORLD]; unsigned long long first[DX_SHUFFLE_MAX_WORLD];
    unsigned long long limit[DX_SHUFFLE_MAX_WORLD];      // records this rank's block holds in that owner's buffer (>= 1): nothing is written past it
};
__global__ void __launch_bounds__(256) k5_scatter_put(const __grid_constant__ ShuffleArgs a, const __grid_constant__ PutTargets tg, unsigned long long* cursor) {
    const int rec_words = a.n_cols + 1;
    const int lane = threadIdx.x & 31;
    for (long long i0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) - lane; i0 < a.n_rows; i0 += (long long)gridDim.x * blockDim.x) {
        const long long i = i0 + lane;
        const bool has = i < a.n_rows;
        const int dst = has ? owner_of(a, i) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, dst);
        const int leader = __ffs(peers) - 1;
        const int rank_in_group = __popc(peers & ((1u << lane) - 1u));
        unsigned long long start = 0;
        if (has && lane == leader) start = atomicAdd(&cursor[dst], (unsigned long long)__popc(peers));
        start = __shfl_sync(0xffffffffu, start, leader);
        if (!has) continue;
        unsigned long long* rec = tg.base[dst] + (tg.first[dst] + start + rank_in_group) * rec_words;
        unsigned long long valid = 0;
        for (int c = 0; c < a.n_cols; c++) {
            long long v; if (sh_load(a.cols[c], i, v)) valid |= 1ull << c;
            rec[c] = (unsigned long long)v;
        }
        rec[a.n_cols] = valid;
    }
}

// The same with coalesced remote stores: the lanes of a warp that share an owner first lay their records side by side in
// shared memory, then the whole warp copies that run to the owner's buffer with consecutive 8-byte stores (a 40-byte
// record written field by field from one lane would cross NVLink as five separate partial-line writes).
__global__ void __launch_bounds__(256) k5_scatter_put_staged(const __grid_constant__ ShuffleArgs a, const __grid_constant__ PutTargets tg, unsigned long long* cursor) {
    extern __shared__ unsigned long long s_stage[];          // per warp: 32 records
    const int rec_words = a.n_cols + 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long* stage = s_stage + (size_t)warp * 32 * rec_words;
    for (long long i0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) - lane; i0 < a.n_rows; i0 += (long long)gridDim.x * blockDim.x) {
        const long long i = i0 + lane;
        const bool has = i < a.n_rows;
        const int dst = has ? owner_of(a, i) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, dst);
        const int leader = __ffs(peers) - 1;
        const int rank_in_group = __popc(peers & ((1u << lane) - 1u));
        unsigned long long start = 0;
        if (has && lane == leader) start = atomicAdd(&cursor[dst], (unsigned long long)__popc(peers));
        start = __shfl_sync(0xffffffffu, start, leader);
        // one group (owner) at a time, in order of its first lane
        unsigned todo = __ballot_sync(0xffffffffu, has && lane == leader);
        while (todo) {
            const int gl = __ffs(todo) - 1; todo &= todo - 1;
            const unsigned gmask = __shfl_sync(0xffffffffu, peers, gl);
            const int gdst = __shfl_sync(0xffffffffu, dst, gl);
            const unsigned long long gstart = __shfl_sync(0xffffffffu, start, gl);
            if (has && dst == gdst) {
                unsigned long long* r = stage + (size_t)rank_in_group * rec_words;
                unsigned long long valid = 0;
                for (int c = 0; c < a.n_cols; c++) {
                    long long v; if (sh_load(a.cols[c], i, v)) valid |= 1ull << c;
                    r[c] = (unsigned long long)v;
                }
                r[a.n_cols] = valid;
            }
            __syncwarp();
            const int n_words = __popc(gmask) * rec_words;
            unsigned long long* out = tg.base[gdst] + (tg.first[gdst] + gstart) * rec_words;
            for (int w = lane; w < n_words; w += 32) out[w] = stage[w];
            __syncwarp();
        }
    }
}

// The form that normally runs, for the fused put and for the local pack alike: a CTA takes a tile of PT_TILE rows, counts them
// per owner in shared memory (one shared atomic per warp and owner), reserves each owner's run in that owner's buffer with ONE
// global atomic per owner per tile, lays the records down in shared memory sorted by owner, and copies the staging area out
// with consecutive lanes writing consecutive 8-byte words: every owner receives one contiguous run per tile (128 records =
// 5 KB at 8 owners), so the stores that cross NVLink are full lines instead of the per-warp fragments of a few records the
// warp-staged kernel above produces, and the source columns are read four rows per thread with the loads in flight together.
#ifndef PT_NT
#define PT_NT 256
#endif
#ifndef PT_RPT
#define PT_RPT 2
#endif
#define PT_TILE (PT_NT * PT_RPT)
__global__ void __launch_bounds__(PT_NT) k5_put_tiled(const __grid_constant__ ShuffleArgs a, const __grid_constant__ PutTargets tg, unsigned long long* cursor) {
    extern __shared__ __align__(16) unsigned long long pt_stage[];       // PT_TILE x rec_words words, then PT_TILE destination pointers
    __shared__ unsigned int s_cnt[DX_SHUFFLE_MAX_WORLD], s_base[DX_SHUFFLE_MAX_WORLD + 1];
    __shared__ unsigned long long s_gstart[DX_SHUFFLE_MAX_WORLD];
    const int rw = a.n_cols + 1, tid = threadIdx.x, lane = tid & 31;
    unsigned long long** s_dst = reinterpret_cast<unsigned long long**>(pt_stage + (size_t)PT_TILE * rw);
    const long long n_tiles = (a.n_rows + PT_TILE - 1) / PT_TILE;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        if (tid < a.world) s_cnt[tid] = 0;
        __syncthreads();
        int dst[PT_RPT]; unsigned int rank[PT_RPT];
#pragma unroll
        for (int j = 0; j < PT_RPT; j++) {
            const long long i = tile * PT_TILE + (long long)j * PT_NT + tid;
            dst[j] = i < a.n_rows ? owner_of(a, i) : -1;
        }
#pragma unroll
        for (int j = 0; j < PT_RPT; j++) {
            const unsigned peers = __match_any_sync(0xffffffffu, dst[j]);
            const int leader = __ffs(peers) - 1;
            unsigned int start = 0;
            if (dst[j] >= 0 && lane == leader) start = atomicAdd(&s_cnt[dst[j]], (unsigned int)__popc(peers));
            start = __shfl_sync(0xffffffffu, start, leader);
            rank[j] = start + __popc(peers & ((1u << lane) - 1u));
        }
        __syncthreads();
        if (tid == 0) { unsigned int acc = 0; for (int d = 0; d < a.world; d++) { s_base[d] = acc; acc += s_cnt[d]; } s_base[a.world] = acc; }
        if (tid < a.world && s_cnt[tid]) s_gstart[tid] = atomicAdd(&cursor[tid], (unsigned long long)s_cnt[tid]);
        __syncthreads();
#pragma unroll
        for (int j = 0; j < PT_RPT; j++) {
            if (dst[j] < 0) continue;
            const long long i = tile * PT_TILE + (long long)j * PT_NT + tid;
            const unsigned int slot = s_base[dst[j]] + rank[j];
            unsigned long long* r = pt_stage + (size_t)slot * rw;
            unsigned long long valid = 0;
            for (int c = 0; c < a.n_cols; c++) {
                long long v; if (sh_load(a.cols[c], i, v)) valid |= 1ull << c;
                r[c] = (unsigned long long)v;
            }
            r[a.n_cols] = valid;
            // a record past the block's end is written over the block's last record: the call fails anyway (the cursor says so),
            // and no store ever leaves the block
            const unsigned long long at = s_gstart[dst[j]] + rank[j], lim = tg.limit[dst[j]];
            s_dst[slot] = tg.base[dst[j]] + (tg.first[dst[j]] + (at < lim ? at : lim - 1ull)) * rw;
        }
        __syncthreads();
        {   // no division: word index = thread id modulo the next power of two >= the record length (the lanes past it idle)
            int sh = 0; while ((1 << sh) < rw) sh++;
            const unsigned int w = tid & ((1u << sh) - 1u), r0 = tid >> sh, rstep = PT_NT >> sh, n_slots = s_base[a.world];
            if (w < (unsigned int)rw)
                for (unsigned int slot = r0; slot < n_slots; slot += rstep) s_dst[slot][w] = pt_stage[slot * rw + w];
        }
        __syncthreads();
    }
}
static size_t pt_smem(int rec_words) { return (size_t)PT_TILE * rec_words * 8 + (size_t)PT_TILE * 8; }
static int pt_launch(dx_ctx* ctx, const ShuffleArgs& a, const PutTargets& tg, unsigned long long* cursor) {
    const size_t smem = pt_smem(a.n_cols + 1);
    if (smem > ctx->k5_put_smem_configured) {
        DX_CUDA(ctx, cudaFuncSetAttribute(k5_put_tiled, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ctx->k5_put_smem_configured = smem;
    }
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k5_put_tiled, PT_NT, smem);
    if (occ < 1) occ = 1;
    long long blocks = (a.n_rows + PT_TILE - 1) / PT_TILE, cap = (long long)ctx->sm_count * occ;
    if (blocks > cap) blocks = cap;
    k5_put_tiled<<<(unsigned)blocks, PT_NT, smem, ctx->stream>>>(a, tg, cursor);
    return DX_OK;
}

// segs: the received records lie in up to DX_SHUFFLE_MAX_WORLD runs (one block per sender); row i of the table is record
// first[s] + (i - pre[s]) of the run s with pre[s] <= i < pre[s + 1]
struct UnpackSegs { int n; long long first[DX_SHUFFLE_MAX_WORLD]; long long pre[DX_SHUFFLE_MAX_WORLD + 1]; };
__global__ void __launch_bounds__(256) k5_unpack(const unsigned long long* packed, long long n_rows, int n_cols, UnpackCols cols, uint32_t* vmask32, const __grid_constant__ UnpackSegs segs) {
    const int rec_words = n_cols + 1;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_rows; i += (long long)gridDim.x * blockDim.x) {
        long long src = i;
        if (segs.n == 1) src = segs.first[0] + i;
        else if (segs.n > 1) {
            int sg = 0;
#pragma unroll
            for (int k = 1; k < DX_SHUFFLE_MAX_WORLD; k++) if (k < segs.n && i >= segs.pre[k]) sg = k;
            src = segs.first[sg] + (i - segs.pre[sg]);
        }
        const unsigned long long* rec = packed + src * rec_words;
        for (int c = 0; c < n_cols; c++) cols.p[c][i] = (long long)rec[c];
        vmask32[i] = (uint32_t)rec[n_cols];
    }
}

static unsigned grid_of(dx_ctx* c, long long n) {
    long long b = (n + 255) / 256, cap = (long long)c->sm_count * 8;
    return (unsigned)(b > cap ? cap : (b < 1 ? 1 : b));
}

static int shuffle_args(dx_ctx* ctx, const DxTable* t, const std::vector<int>& cols, const std::vector<int>& key_pos, int world, ShuffleArgs* out);

// rows per owner rank (k5_count only)
int dx_shuffle_count(dx_ctx* ctx, const DxTable* t, const std::vector<int>& cols, const std::vector<int>& key_pos, int world, int64_t* counts_host) {
    ShuffleArgs a; int rc = shuffle_args(ctx, t, cols, key_pos, world, &a); if (rc) return rc;
    for (int r = 0; r < world; r++) counts_host[r] = 0;
    if (t->n_rows == 0) return DX_OK;
    DevBuf st(ctx, DX_SHUFFLE_MAX_WORLD * 8);
    DX_CUDA(ctx, cudaMemsetAsync(st.p, 0, DX_SHUFFLE_MAX_WORLD * 8, ctx->stream));
    {
        StageTimer tm(ctx, "k5_count");
        k5_count<<<grid_of(ctx, t->n_rows), 256, 0, ctx->stream>>>(a, st.as<unsigned long long>());
        ctx->launches["k5_count"]++;
    }
    unsigned long long h[DX_SHUFFLE_MAX_WORLD];
    DX_CUDA(ctx, cudaMemcpyAsync(h, st.p, world * 8, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int r = 0; r < world; r++) counts_host[r] = (int64_t)h[r];
    return DX_OK;
}

// records written directly to each owner's buffer: peer_bufs[r] = rank r's receive buffer as mapped in this process,
// first_record[r] = where this rank's block starts there
int dx_shuffle_put(dx_ctx* ctx, const DxTable* t, const std::vector<int>& cols, const std::vector<int>& key_pos, int world,
                   void* const* peer_bufs, const int64_t* first_record, const int64_t* block_records, int64_t* sent) {
    ShuffleArgs a; int rc = shuffle_args(ctx, t, cols, key_pos, world, &a); if (rc) return rc;
    if (sent) for (int r = 0; r < world; r++) sent[r] = 0;
    if (t->n_rows == 0) return DX_OK;
    PutTargets tg; memset(&tg, 0, sizeof tg);
    for (int r = 0; r < world; r++) {
        tg.base[r] = reinterpret_cast<unsigned long long*>(peer_bufs[r]); tg.first[r] = (unsigned long long)first_record[r];
        tg.limit[r] = block_records ? (unsigned long long)(block_records[r] > 0 ? block_records[r] : 1) : ~0ull;
    }
    const int rec_words = a.n_cols + 1;
    if (block_records && rec_words > 12) return ctx->set_error(DX_ERR_UNSUPPORTED, "shuffle put into fixed blocks: at most 11 columns");
    DevBuf cur(ctx, DX_SHUFFLE_MAX_WORLD * 8);
    DX_CUDA(ctx, cudaMemsetAsync(cur.p, 0, DX_SHUFFLE_MAX_WORLD * 8, ctx->stream));
    {
        StageTimer tm(ctx, "k5_scatter_put");
        if (rec_words <= 12) { rc = pt_launch(ctx, a, tg, cur.as<unsigned long long>()); if (rc) return rc; }
        else if (rec_words <= 16) k5_scatter_put_staged<<<grid_of(ctx, t->n_rows), 256, (size_t)8 * 32 * rec_words * 8, ctx->stream>>>(a, tg, cur.as<unsigned long long>());
        else k5_scatter_put<<<grid_of(ctx, t->n_rows), 256, 0, ctx->stream>>>(a, tg, cur.as<unsigned long long>());
        ctx->launches["k5_scatter_put"]++;
    }
    DX_CUDA(ctx, cudaGetLastError());
    unsigned long long h[DX_SHUFFLE_MAX_WORLD];
    DX_CUDA(ctx, cudaMemcpyAsync(h, cur.p, DX_SHUFFLE_MAX_WORLD * 8, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));      // the peers may read once every rank is past this point (the caller's barrier)
    for (int r = 0; r < world; r++) {
        if (sent) sent[r] = (int64_t)h[r];
        if (block_records && h[r] > (unsigned long long)block_records[r])
            return ctx->set_error(DX_ERR_CAPACITY, "shuffle put: " + std::to_string(h[r]) + " records for rank " + std::to_string(r) + " exceed this rank's block of " + std::to_string(block_records[r]));
    }
    return DX_OK;
}

static int shuffle_args(dx_ctx* ctx, const DxTable* t, const std::vector<int>& cols, const std::vector<int>& key_pos, int world, ShuffleArgs* out) {
    if (world < 1 || world > DX_SHUFFLE_MAX_WORLD) return ctx->set_error(DX_ERR_ARG, "shuffle: world size out of range");
    if (cols.empty() || (int)cols.size() > DX_SHUFFLE_MAX_COLS) return ctx->set_error(DX_ERR_ARG, "shuffle: 1..32 columns");
    ShuffleArgs a; memset(&a, 0, sizeof a);
    a.n_cols = (int)cols.size(); a.n_keys = (int)key_pos.size(); a.world = world; a.n_rows = t->n_rows;
    for (size_t c = 0; c < cols.size(); c++) {
        const DxColumn& col = t->cols[cols[c]];
        if (col.type == DX_T_STR || col.type == DX_T_FLAGS) return ctx->set_error(DX_ERR_UNSUPPORTED, "shuffle: dictionary-coded and flag columns are rank-local and cannot be sent ('" + col.name + "')");
        if (!col.data || col.all_null) return ctx->set_error(DX_ERR_UNSUPPORTED, "shuffle: constant column '" + col.name + "'");
        a.cols[c].data = col.data->p; a.cols[c].vmask = col.vmask ? col.vmask->p : nullptr; a.cols[c].vwidth = col.vwidth; a.cols[c].vbit = col.vbit; a.cols[c].width = dx_type_width(col.type);
    }
    for (size_t k = 0; k < key_pos.size(); k++) a.key_idx[k] = key_pos[k];
    *out = a;
    return DX_OK;
}

int dx_shuffle_pack(dx_ctx* ctx, const DxTable* t, const std::vector<int>& cols, const std::vector<int>& key_pos, int world,
                    void* packed_dev, int64_t* counts_host) {
    if (world < 1 || world > DX_SHUFFLE_MAX_WORLD) return ctx->set_error(DX_ERR_ARG, "shuffle: world size out of range");
    if (cols.empty() || (int)cols.size() > DX_SHUFFLE_MAX_COLS) return ctx->set_error(DX_ERR_ARG, "shuffle: 1..32 columns");
    ShuffleArgs a; memset(&a, 0, sizeof a);
    a.n_cols = (int)cols.size(); a.n_keys = (int)key_pos.size(); a.world = world; a.n_rows = t->n_rows;
    for (size_t c = 0; c < cols.size(); c++) {
        const DxColumn& col = t->cols[cols[c]];
        if (col.type == DX_T_STR || col.type == DX_T_FLAGS) return ctx->set_error(DX_ERR_UNSUPPORTED, "shuffle: dictionary-coded and flag columns are rank-local and cannot be sent ('" + col.name + "')");
        if (!col.data || col.all_null) return ctx->set_error(DX_ERR_UNSUPPORTED, "shuffle: constant column '" + col.name + "'");
        a.cols[c].data = col.data->p; a.cols[c].vmask = col.vmask ? col.vmask->p : nullptr; a.cols[c].vwidth = col.vwidth; a.cols[c].vbit = col.vbit; a.cols[c].width = dx_type_width(col.type);
    }
    for (size_t k = 0; k < key_pos.size(); k++) a.key_idx[k] = key_pos[k];
    for (int r = 0; r < world; r++) counts_host[r] = 0;
    if (t->n_rows == 0) return DX_OK;
    DevBuf st(ctx, 3 * DX_SHUFFLE_MAX_WORLD * 8);
    unsigned long long* counts = st.as<unsigned long long>();
    unsigned long long* base = counts + DX_SHUFFLE_MAX_WORLD;
    unsigned long long* cursor = base + DX_SHUFFLE_MAX_WORLD;
    DX_CUDA(ctx, cudaMemsetAsync(st.p, 0, 3 * DX_SHUFFLE_MAX_WORLD * 8, ctx->stream));
    {
        StageTimer tm(ctx, "k5_count");
        k5_count<<<grid_of(ctx, t->n_rows), 256, 0, ctx->stream>>>(a, counts);
        ctx->launches["k5_count"]++;
    }
    unsigned long long h[DX_SHUFFLE_MAX_WORLD];
    DX_CUDA(ctx, cudaMemcpyAsync(h, counts, world * 8, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    unsigned long long hb[DX_SHUFFLE_MAX_WORLD], acc = 0;
    for (int r = 0; r < world; r++) { hb[r] = acc; acc += h[r]; counts_host[r] = (int64_t)h[r]; }
    DX_CUDA(ctx, cudaMemcpyAsync(base, hb, world * 8, cudaMemcpyHostToDevice, ctx->stream));
    {
        StageTimer tm(ctx, "k5_scatter");
        if (a.n_cols + 1 <= 12) {      // the tiled kernel with every owner's run in the one local buffer
            PutTargets tg; memset(&tg, 0, sizeof tg);
            for (int r = 0; r < world; r++) { tg.base[r] = reinterpret_cast<unsigned long long*>(packed_dev); tg.first[r] = hb[r]; tg.limit[r] = ~0ull; }
            int rc2 = pt_launch(ctx, a, tg, cursor); if (rc2) return rc2;
        } else k5_scatter<<<grid_of(ctx, t->n_rows), 256, 0, ctx->stream>>>(a, base, cursor, reinterpret_cast<unsigned long long*>(packed_dev));
        ctx->launches["k5_scatter"]++;
    }
    DX_CUDA(ctx, cudaGetLastError());
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return DX_OK;
}

int dx_shuffle_unpack(dx_ctx* ctx, const std::vector<std::string>& names, const std::vector<int>& types, const void* packed_dev, int64_t n_rows,
                      DxTableP* out, int n_segs, const int64_t* seg_first, const int64_t* seg_count) {
    UnpackSegs sg; memset(&sg, 0, sizeof sg);
    if (n_segs > 0) {
        if (n_segs > DX_SHUFFLE_MAX_WORLD || !seg_first || !seg_count) return ctx->set_error(DX_ERR_ARG, "unpack: bad segment list");
        sg.n = n_segs; n_rows = 0;
        for (int k = 0; k < n_segs; k++) { sg.first[k] = seg_first[k]; sg.pre[k] = n_rows; n_rows += seg_count[k]; }
        sg.pre[n_segs] = n_rows;
    }
    if (names.empty() || names.size() != types.size() || (int)names.size() > DX_SHUFFLE_MAX_COLS) return ctx->set_error(DX_ERR_ARG, "unpack: 1..32 columns");
    auto t = std::make_shared<DxTable>(); t->ctx = ctx; t->n_rows = n_rows;
    DevBufP mask = make_buf(ctx, (size_t)(n_rows > 0 ? n_rows : 1) * 4);
    UnpackCols uc; memset(&uc, 0, sizeof uc);
    for (size_t c = 0; c < names.size(); c++) {
        if (types[c] == DX_T_STR || types[c] == DX_T_FLAGS) return ctx->set_error(DX_ERR_UNSUPPORTED, "unpack: dictionary-coded and flag columns cannot be received");
        DxColumn col; col.name = names[c]; col.type = types[c];
        col.data = make_buf(ctx, (size_t)(n_rows > 0 ? n_rows : 1) * 8);
        col.vmask = mask; col.vwidth = 4; col.vbit = (int)c;
        uc.p[c] = col.data->as<long long>();
        t->cols.push_back(col);
    }
    if (n_rows > 0) {
        StageTimer tm(ctx, "k5_unpack");
        k5_unpack<<<grid_of(ctx, n_rows), 256, 0, ctx->stream>>>(reinterpret_cast<const unsigned long long*>(packed_dev), n_rows, (int)names.size(), uc, mask->as<uint32_t>(), sg);
        ctx->launches["k5_unpack"]++;
        DX_CUDA(ctx, cudaGetLastError());
    }
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out = t;
    return DX_OK;
}
