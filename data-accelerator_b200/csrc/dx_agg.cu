// dx_agg.cu — K3 (window / group-by aggregation) and the aggregate-upsert half of K4.
// One open-addressing hash table in HBM/L2 per statement; every event probes by the 64-bit
// hash of its group key and updates the group's accumulators with native 64-bit atomics.
// See dx_agg.h for the reference call sites this replaces.
#include "dx_agg.h"

static __device__ __forceinline__ unsigned long long amix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static __device__ __forceinline__ long long f64_order(long long bits) {   // involution: double order <-> signed order
    double d = __longlong_as_double(bits);
    if (d != d) bits = 0x7FF8000000000000ll;          // every NaN is the one greatest value (Spark ordering)
    return bits < 0 ? (bits ^ 0x7FFFFFFFFFFFFFFFll) : bits;
}
static __device__ __forceinline__ bool load_val(const ProgCol& c, long long row, long long& v) {
    if (c.vmask) {
        uint32_t m = c.vwidth == 1 ? reinterpret_cast<const uint8_t*>(c.vmask)[row] : reinterpret_cast<const uint32_t*>(c.vmask)[row];
        if (!((m >> c.vbit) & 1u)) { v = 0; return false; }
    }
    v = c.width == 4 ? (long long)reinterpret_cast<const int32_t*>(c.data)[row] : reinterpret_cast<const long long*>(c.data)[row];
    return true;
}

// KMAX bounds the unrolled key loops (4, 8 or 16) so that narrow group-bys do not carry the registers of wide ones
template <int KMAX>
__global__ void __launch_bounds__(256) k3_hash_agg(const __grid_constant__ AggTableDev T, const __grid_constant__ AggInput in) {
    const int nk = T.n_keys, na = T.n_accs, sw = T.slot_words;
    // Every thread of a warp makes the same number of trips, so the warp can be re-joined (__syncwarp) between the divergent
    // slot search and the accumulator updates; without it the updates run with a handful of active lanes.
    for (long long i0 = (long long)blockIdx.x * blockDim.x; i0 < in.n_rows; i0 += (long long)gridDim.x * blockDim.x) {
        const long long i = i0 + threadIdx.x;
        if (*reinterpret_cast<volatile unsigned long long*>(&T.counters[1])) return;   // table declared too small: the host retries with a larger one
        bool has = i < in.n_rows;
        const long long row = has ? (in.sel ? in.sel[i] : i) : 0;
        if (has && in.has_range) {     // the window view's [start, end) filter, fused (CommonProcessorFactory.scala:203-227)
            long long t;
            if (!load_val(in.range_col, row, t) || t < in.range_lo || t >= in.range_hi) has = false;
        }
        long long pane = 0;
        if (has && in.pane_us) {
            long long t;
            if (!load_val(in.range_col, row, t)) has = false;
            else pane = t >= 0 ? t / in.pane_us : -((-t + in.pane_us - 1) / in.pane_us);
        }
        unsigned long long* slot = nullptr;
        if (has) {
        long long kv[KMAX];
        unsigned long long nullbits = 0, h = 0x2545F4914F6CDD1Dull;
#pragma unroll
        for (int k = 0; k < KMAX; k++) {
            if (k < nk) {
                long long v; bool ok;
                if (in.pane_us && k == nk - 1) { v = pane; ok = true; } else ok = load_val(in.keys[k], row, v);
                kv[k] = v; if (!ok) nullbits |= 1ull << k;
                h = amix64(h ^ (unsigned long long)v);
            }
        }
        h = amix64(h ^ nullbits);
        const unsigned long long tag = h | 0x8000000000000000ull;     // never 0 (empty) or 1 (being filled)
        unsigned long long idx = (h >> 3) & T.cap_mask;
        for (unsigned long long probes = 0; probes <= T.cap_mask && probes < 256; ) {   // a probe run this long means the table is crowded: give up, grow
            unsigned long long* s = T.slots + idx * sw;
            unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(s);
            if (cur == tag) {
                // The filler publishes the tag after a __threadfence, so the key words are in L2 once the tag is seen — but this
                // SM's L1 may still hold the sector from before they were written (a slot can straddle the 32-byte sector of a
                // neighbour that was compared earlier).  Read them from L2 (ld.global.cg), never through L1.
                bool same = (__ldcg(s + nk + 1) == nullbits);
#pragma unroll
                for (int k = 0; k < KMAX; k++) if (k < nk) same = same && ((long long)__ldcg(s + 1 + k) == kv[k]);
                if (same) { slot = s; break; }
            } else if (cur == 0ull) {
                if (atomicCAS(s, 0ull, 1ull) == 0ull) {
#pragma unroll
                    for (int k = 0; k < KMAX; k++) if (k < nk) s[1 + k] = (unsigned long long)kv[k];
                    s[nk + 1] = nullbits;
                    for (int a = 0; a < na; a++) {
                        const int kind = T.accs[a].kind;
                        unsigned long long init = 0;
                        if (kind == ACC_MIN_I || kind == ACC_MIN_F) init = 0x7FFFFFFFFFFFFFFFull;
                        else if (kind == ACC_MAX_I || kind == ACC_MAX_F) init = 0x8000000000000000ull;
                        s[nk + 2 + a] = init;
                    }
                    __threadfence();
                    *reinterpret_cast<volatile unsigned long long*>(s) = tag;
                    unsigned long long g = atomicAdd(&T.counters[0], 1ull);
                    if (g + 1 > T.max_groups) T.counters[1] = 1ull;
                    slot = s; break;
                }
                continue;   // lost the race: look at the same slot again
            } else if (cur == 1ull) {
                continue;   // being filled by another thread
            }
            idx = (idx + 1) & T.cap_mask; probes++;
        }
        if (!slot) T.counters[1] = 1ull;
        }
        __syncwarp();
        if (!slot) continue;
        unsigned long long* acc = slot + nk + 2;
        for (int a = 0; a < na; a++) {
            const int kind = T.accs[a].kind;
            if (kind == ACC_COUNT_STAR) { atomicAdd(&acc[a], 1ull); continue; }
            long long v;
            if (!load_val(in.vals[T.accs[a].col], row, v)) continue;
            switch (kind) {
            case ACC_COUNT: atomicAdd(&acc[a], 1ull); break;
            case ACC_SUM_I: atomicAdd(&acc[a], (unsigned long long)v); break;
            case ACC_SUM_F: atomicAdd(reinterpret_cast<double*>(&acc[a]), __longlong_as_double(v)); break;
            case ACC_MIN_I: atomicMin(reinterpret_cast<long long*>(&acc[a]), v); break;
            case ACC_MAX_I: atomicMax(reinterpret_cast<long long*>(&acc[a]), v); break;
            case ACC_MIN_F: atomicMin(reinterpret_cast<long long*>(&acc[a]), f64_order(v)); break;
            case ACC_MAX_F: atomicMax(reinterpret_cast<long long*>(&acc[a]), f64_order(v)); break;
            }
        }
    }
}

__global__ void k3_extract(const AggTableDev T, unsigned long long capacity, unsigned long long* out_count,
                           long long* const* key_out, unsigned long long* keynull_out, unsigned long long* const* acc_out) {
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < capacity; i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long* s = T.slots + i * T.slot_words;
        if (s[0] < 2ull) continue;
        unsigned long long g = atomicAdd(out_count, 1ull);
        for (int k = 0; k < T.n_keys; k++) key_out[k][g] = (long long)s[1 + k];
        keynull_out[g] = s[T.n_keys + 1];
        for (int a = 0; a < T.n_accs; a++) {
            unsigned long long v = s[T.n_keys + 2 + a];
            const int kind = T.accs[a].kind;
            if (kind == ACC_MIN_F || kind == ACC_MAX_F) v = (unsigned long long)f64_order((long long)v);
            acc_out[a][g] = v;
        }
    }
}

// group null bits (bit k set = key k is NULL) -> a 32-bit validity word per group (bit k set = key k present)
__global__ void k3_keynull_to_valid(const unsigned long long* keynull, long long n, uint32_t* valid) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) valid[i] = ~(uint32_t)keynull[i];
}
int dx_agg_keynull_to_valid(dx_ctx* c, const unsigned long long* keynull, long long n, uint32_t* valid) {
    if (n <= 0) return DX_OK;
    k3_keynull_to_valid<<<(unsigned)std::min<long long>((n + 255) / 256, 4096), 256, 0, c->stream>>>(keynull, n, valid);
    c->launches["k3_keynull_to_valid"]++;
    DX_CUDA(c, cudaGetLastError());
    return DX_OK;
}

int AggTable::init(dx_ctx* c, int n_keys, int n_vals, const std::vector<AggAccDesc>& accs, unsigned long long capacity_pow2) {
    ctx = c;
    if (n_keys > DX_AGG_MAX_KEYS) return c->set_error(DX_ERR_UNSUPPORTED, "more than 16 group-by keys");
    if ((int)accs.size() > DX_AGG_MAX_ACCS || n_vals > DX_AGG_MAX_ACCS) return c->set_error(DX_ERR_UNSUPPORTED, "more than 16 aggregate accumulators");
    unsigned long long cap = 64; while (cap < capacity_pow2) cap <<= 1;
    capacity = cap;
    memset(&dev, 0, sizeof dev);
    dev.n_keys = n_keys; dev.n_accs = (int)accs.size(); dev.n_vals = n_vals;
    dev.slot_words = 2 + n_keys + (int)accs.size();
    for (size_t a = 0; a < accs.size(); a++) dev.accs[a] = accs[a];
    slots = DevBuf(c, (size_t)cap * dev.slot_words * 8);
    counters = DevBuf(c, 64);
    if (!slots.p || !counters.p) return c->set_error(DX_ERR_CUDA, "aggregation table allocation failed");
    DX_CUDA(c, cudaMemsetAsync(slots.p, 0, (size_t)cap * dev.slot_words * 8, c->stream));
    DX_CUDA(c, cudaMemsetAsync(counters.p, 0, 64, c->stream));
    dev.slots = slots.as<unsigned long long>(); dev.cap_mask = cap - 1;
    dev.counters = counters.as<unsigned long long>();
    dev.max_groups = cap - cap / 4;     // load factor <= 0.75
    c->launches["memset"] += 0;
    return DX_OK;
}

int AggTable::update(const AggInput& in) {
    if (in.n_rows <= 0) return DX_OK;
    long long blocks = (in.n_rows + 255) / 256;
    long long capb = (long long)ctx->sm_count * 8; if (blocks > capb) blocks = capb;
    ctx->counters["k3_rows"] += in.n_rows;
    StageTimer tm(ctx, "k3_hash_agg");
    if (dev.n_keys <= 4) k3_hash_agg<4><<<(unsigned)blocks, 256, 0, ctx->stream>>>(dev, in);
    else if (dev.n_keys <= 8) k3_hash_agg<8><<<(unsigned)blocks, 256, 0, ctx->stream>>>(dev, in);
    else k3_hash_agg<16><<<(unsigned)blocks, 256, 0, ctx->stream>>>(dev, in);
    ctx->launches["k3_hash_agg"]++;
    DX_CUDA(ctx, cudaGetLastError());
    return DX_OK;
}

int AggTable::finish(unsigned long long* n_groups, bool* overflow) {
    unsigned long long h[2] = {0, 0};
    DX_CUDA(ctx, cudaMemcpyAsync(h, counters.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *n_groups = h[0]; *overflow = h[1] != 0;
    return DX_OK;
}

int AggTable::extract(unsigned long long n_groups, long long* const* key_out, unsigned long long* keynull_out, unsigned long long* const* acc_out) {
    if (n_groups == 0) return DX_OK;
    DevBuf ptrs(ctx, (DX_AGG_MAX_KEYS + DX_AGG_MAX_ACCS) * 8 + 16);
    void* host_ptrs[DX_AGG_MAX_KEYS + DX_AGG_MAX_ACCS];
    for (int k = 0; k < dev.n_keys; k++) host_ptrs[k] = key_out[k];
    for (int a = 0; a < dev.n_accs; a++) host_ptrs[DX_AGG_MAX_KEYS + a] = acc_out[a];
    DX_CUDA(ctx, cudaMemcpyAsync(ptrs.p, host_ptrs, sizeof host_ptrs, cudaMemcpyHostToDevice, ctx->stream));
    DX_CUDA(ctx, cudaMemsetAsync(counters.as<unsigned long long>() + 2, 0, 8, ctx->stream));
    long long blocks = (long long)((capacity + 255) / 256);
    long long capb = (long long)ctx->sm_count * 8; if (blocks > capb) blocks = capb;
    k3_extract<<<(unsigned)blocks, 256, 0, ctx->stream>>>(dev, capacity, counters.as<unsigned long long>() + 2,
        reinterpret_cast<long long* const*>(ptrs.p), keynull_out,
        reinterpret_cast<unsigned long long* const*>(ptrs.as<void*>() + DX_AGG_MAX_KEYS));
    ctx->launches["k3_extract"]++;
    DX_CUDA(ctx, cudaGetLastError());
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return DX_OK;
}
