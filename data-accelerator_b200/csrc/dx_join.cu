// dx_join.cu — inner equi-join of two device tables.
//
// Stands in for the joins the flow SQL sends through spark.sql (CommonProcessorFactory.scala:271):
// events x reference data (ReferenceDataHandler.scala:42-60: CSV temp views, all-string columns) and
// joins between derived tables (HomeAutomationLocal.json:326: DeviceInfoTimeWindow INNER JOIN ... ON
// a.EventTime = b.MaxEventTime).  Keys that are NULL never match; matches are emitted as (left row,
// right row) pairs and both sides' columns are gathered through them.
#include "dx_join.h"
#include <cub/device/device_scan.cuh>

static __device__ __forceinline__ unsigned long long jmix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static __device__ __forceinline__ bool j_load(const ProgCol& c, long long row, long long& v) {
    if (c.vmask) {
        uint32_t m = c.vwidth == 1 ? reinterpret_cast<const uint8_t*>(c.vmask)[row] : reinterpret_cast<const uint32_t*>(c.vmask)[row];
        if (!((m >> c.vbit) & 1u)) { v = 0; return false; }
    }
    v = c.width == 4 ? (long long)reinterpret_cast<const int32_t*>(c.data)[row] : reinterpret_cast<const long long*>(c.data)[row];
    return true;
}
static __device__ __forceinline__ bool j_key(const JoinSide& s, long long row, long long* kv, unsigned long long* h) {
    unsigned long long x = 0x51ED270B27B4F3CFull;
    for (int k = 0; k < s.n_keys; k++) { if (!j_load(s.keys[k], row, kv[k])) return false; x = jmix64(x ^ (unsigned long long)kv[k]); }
    *h = x; return true;
}

// build: head[slot] = last inserted right row with that hash bucket, next[row] = previous one
__global__ void k_join_build(const __grid_constant__ JoinSide R, long long n, long long* head, unsigned long long mask, long long* next) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        long long kv[DX_JOIN_MAX_KEYS]; unsigned long long h;
        next[i] = -1;
        if (!j_key(R, i, kv, &h)) continue;
        next[i] = (long long)atomicExch(reinterpret_cast<unsigned long long*>(&head[h & mask]), (unsigned long long)i);
    }
}

// probe: pass 0 counts the matches of every left row, pass 1 writes the pairs at the scanned offsets
__global__ void k_join_probe(const __grid_constant__ JoinSide L, const __grid_constant__ JoinSide R, long long n_left, const long long* sel,
                             const long long* head, unsigned long long mask, const long long* next, long long* counts_or_offsets,
                             long long* out_left, long long* out_right, int pass, int left_outer) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_left; i += (long long)gridDim.x * blockDim.x) {
        const long long row = sel ? sel[i] : i;
        long long kv[DX_JOIN_MAX_KEYS]; unsigned long long h;
        long long cnt = 0;
        const long long base = pass ? counts_or_offsets[i] : 0;
        if (j_key(L, row, kv, &h)) {
            for (long long r = head[h & mask]; r >= 0; r = next[r]) {
                bool same = true;
                for (int k = 0; k < R.n_keys && same; k++) { long long rv; same = j_load(R.keys[k], r, rv) && rv == kv[k]; }
                if (same) { if (pass) { out_left[base + cnt] = row; out_right[base + cnt] = r; } cnt++; }
            }
        }
        if (left_outer && cnt == 0) {       // LEFT OUTER: an unmatched left row is kept once, its right side NULL (index -1)
            if (pass) { out_left[base] = row; out_right[base] = -1; }
            cnt = 1;
        }
        if (!pass) counts_or_offsets[i] = cnt;
    }
}

static unsigned grid_of(dx_ctx* c, long long n) {
    long long b = (n + 255) / 256, cap = (long long)c->sm_count * 8;
    return (unsigned)(b > cap ? cap : (b < 1 ? 1 : b));
}

int dx_join_pairs(dx_ctx* ctx, const JoinSide& L, int64_t n_left, const int64_t* sel_left, const JoinSide& R, int64_t n_right,
                  DevBufP* left_idx, DevBufP* right_idx, int64_t* n_pairs, bool left_outer) {
    *n_pairs = 0;
    *left_idx = make_buf(ctx, 8); *right_idx = make_buf(ctx, 8);
    if (n_left <= 0 || (n_right <= 0 && !left_outer)) return DX_OK;
    unsigned long long cap = 64; while (cap < (unsigned long long)(n_right > 0 ? n_right : 1) * 2) cap <<= 1;
    DevBuf head(ctx, cap * 8), next(ctx, (size_t)(n_right > 0 ? n_right : 1) * 8), counts(ctx, (size_t)(n_left + 1) * 8), offs(ctx, (size_t)(n_left + 1) * 8);
    if (!head.p || !next.p || !counts.p || !offs.p) return ctx->set_error(DX_ERR_CUDA, "join allocation failed");
    DX_CUDA(ctx, cudaMemsetAsync(head.p, 0xFF, cap * 8, ctx->stream));
    DX_CUDA(ctx, cudaMemsetAsync(counts.p, 0, (size_t)(n_left + 1) * 8, ctx->stream));
    if (n_right > 0) k_join_build<<<grid_of(ctx, n_right), 256, 0, ctx->stream>>>(R, n_right, head.as<long long>(), cap - 1, next.as<long long>());
    ctx->launches["k_join_build"]++;
    {
        StageTimer tm(ctx, "k_join_probe");
        k_join_probe<<<grid_of(ctx, n_left), 256, 0, ctx->stream>>>(L, R, n_left, reinterpret_cast<const long long*>(sel_left), head.as<long long>(), cap - 1,
                                                                    next.as<long long>(), counts.as<long long>(), nullptr, nullptr, 0, left_outer ? 1 : 0);
        ctx->launches["k_join_probe"]++;
    }
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts.as<long long>(), offs.as<long long>(), n_left + 1, ctx->stream);
    DevBuf tmp(ctx, tmp_bytes + 16);
    DX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, counts.as<long long>(), offs.as<long long>(), n_left + 1, ctx->stream));
    long long total = 0;
    DX_CUDA(ctx, cudaMemcpyAsync(&total, offs.as<long long>() + n_left, 8, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *n_pairs = total;
    if (total == 0) return DX_OK;
    *left_idx = make_buf(ctx, (size_t)total * 8); *right_idx = make_buf(ctx, (size_t)total * 8);
    if (!(*left_idx)->p || !(*right_idx)->p) return ctx->set_error(DX_ERR_CUDA, "join output allocation failed");
    k_join_probe<<<grid_of(ctx, n_left), 256, 0, ctx->stream>>>(L, R, n_left, reinterpret_cast<const long long*>(sel_left), head.as<long long>(), cap - 1,
                                                                next.as<long long>(), offs.as<long long>(), (*left_idx)->as<long long>(), (*right_idx)->as<long long>(), 1, left_outer ? 1 : 0);
    ctx->launches["k_join_probe"]++;
    DX_CUDA(ctx, cudaGetLastError());
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return DX_OK;
}
