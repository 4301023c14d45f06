// dx_aggp.cu — K3P: partitioned shared-memory aggregation (see dx_aggp.h).
#include "dx_aggp.h"
#include <cub/device/device_scan.cuh>
#include <chrono>
#include <cmath>
static inline long long now_us() { return std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

#ifndef K3P_SLOTS
#define K3P_SLOTS 512           // shared hash-table slots per CTA (512: four CTAs per SM; measured best of 256 / 512 / 1024)
#endif
#define K3P_NT 256
#define K3P_MAX_GROUPS (K3P_SLOTS * 7 / 10)      // per partition; above this the caller re-partitions finer

static __device__ __forceinline__ unsigned long long pmix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static __device__ __forceinline__ long long p_f64_order(long long bits) {
    double d = __longlong_as_double(bits);
    if (d != d) bits = 0x7FF8000000000000ll;
    return bits < 0 ? (bits ^ 0x7FFFFFFFFFFFFFFFll) : bits;
}
static __device__ __forceinline__ bool p_load(const ProgCol& c, long long row, long long& v) {
    if (c.vmask) {
        uint32_t m = c.vwidth == 1 ? reinterpret_cast<const uint8_t*>(c.vmask)[row] : reinterpret_cast<const uint32_t*>(c.vmask)[row];
        if (!((m >> c.vbit) & 1u)) { v = 0; return false; }
    }
    v = c.width == 4 ? (long long)reinterpret_cast<const int32_t*>(c.data)[row] : reinterpret_cast<const long long*>(c.data)[row];
    return true;
}

// floor(t / d) for d > 0 without the 64-bit integer division routine (~100 instructions): a double-precision estimate, exact
// after one remainder step for |t| < 2^52 (timestamps in microseconds are < 2^51); wider values take the integer division.
static __device__ __forceinline__ long long p_floor_div(long long t, long long d, double inv) {
    if (t > -(1ll << 52) && t < (1ll << 52)) {
        long long q = (long long)floor((double)t * inv);
        const long long r = t - q * d;
        if (r < 0) q--; else if (r >= d) q++;
        return q;
    }
    return t >= 0 ? t / d : -((-t + d - 1) / d);
}

// ---- phase 1: rows -> records [hash][keys nk][pane][bits][vals nv], grouped by partition = top bits of the key hash ----
// bits: low 16 = key-null bits, bit 16+v = value v present.  pass 0 counts, pass 1 writes.
// pass 0 counts, pass 1 writes at bases[p] + cursor (the exact two-pass form); pass 2 is the single pass: every partition
// owns a fixed slice of `stride` records (sized from the row count with headroom, see dx_aggp_build), the cursor alone places
// the record, and a partition that outgrows its slice raises *overflow (the host then falls back to the two-pass form).
__global__ void __launch_bounds__(256) k3p_scatter(const __grid_constant__ AggInput in, int nk, int nv, int pbits, int pass,
                                                   unsigned int* counts, const unsigned int* bases, unsigned int* cursors, unsigned long long* recs,
                                                   unsigned int stride, unsigned long long* overflow) {
    const int rw = nk + nv + 3;
    const double inv_pane = 1.0 / (double)in.pane_us;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < in.n_rows; i += (long long)gridDim.x * blockDim.x) {
        const long long row = in.sel ? in.sel[i] : i;
        long long t;
        if (!p_load(in.range_col, row, t)) continue;                  // no timestamp: in no window
        if (in.has_range && (t < in.range_lo || t >= in.range_hi)) continue;      // a cached batch's late-event cut
        const long long pane = p_floor_div(t, in.pane_us, inv_pane);
        long long kv[DX_AGG_MAX_KEYS];
        unsigned long long bits = 0, h = 0x2545F4914F6CDD1Dull;
        for (int k = 0; k < nk; k++) {
            long long v; const bool ok = p_load(in.keys[k], row, v);
            kv[k] = v; if (!ok) bits |= 1ull << k;
            h = pmix64(h ^ (unsigned long long)v);
        }
        h = pmix64(h ^ bits);
        const unsigned int p = pbits ? (unsigned int)(h >> (64 - pbits)) : 0u;
        if (pass == 0) { atomicAdd(&counts[p], 1u); continue; }
        unsigned int pos;
        if (pass == 2) {
            const unsigned int at = atomicAdd(&cursors[p], 1u);
            if (at >= stride) { *overflow = 1ull; continue; }
            pos = p * stride + at;
        } else pos = bases[p] + atomicAdd(&cursors[p], 1u);
        unsigned long long* r = recs + (size_t)pos * rw;
        r[0] = h;
        for (int k = 0; k < nk; k++) r[1 + k] = (unsigned long long)kv[k];
        r[1 + nk] = (unsigned long long)pane;
        for (int v = 0; v < nv; v++) {
            long long x; if (p_load(in.vals[v], row, x)) bits |= 1ull << (16 + v);
            r[3 + nk + v] = (unsigned long long)x;
        }
        r[2 + nk] = bits;
    }
}

// ---- phase 1, staged form (the one that normally runs): up to 4 group keys and 4 value columns -------------------------
// A record written by the thread that built it is six (rw) 8-byte stores to a line no neighbouring lane touches: the
// load/store unit sees 32 separate sectors per store instruction and the kernel crawls at ~8 G rows/s, every warp waiting
// on it (ncu: long scoreboard 99.8, lg throttle 25).  Here a CTA takes tiles of SC_TILE rows; every thread builds its rows'
// records into a shared-memory staging area (row-major) and claims their slots with independent atomics, and after a barrier
// the whole CTA copies the staged words out, consecutive lanes writing consecutive words of the same record: a store
// instruction then covers ~5 records (5-6 sectors) instead of 32, with 16-byte stores when the record length allows.
#ifndef SC_NT
#define SC_NT 256
#endif
#ifndef SC_RPT
#define SC_RPT 2
#endif
#define SC_TILE (SC_NT * SC_RPT)
#define SC_MAXK 4
#define SC_MAXV 4
static __device__ __forceinline__ bool p_load_u(const ProgCol& c, long long row, long long& v) {
    // branch-free on the hot path: mask and value loads are independent of each other
    uint32_t m = 0xFFFFFFFFu;
    if (c.vmask) m = c.vwidth == 1 ? (uint32_t)__ldg(reinterpret_cast<const uint8_t*>(c.vmask) + row) : __ldg(reinterpret_cast<const uint32_t*>(c.vmask) + row);
    long long x = c.width == 4 ? (long long)__ldg(reinterpret_cast<const int32_t*>(c.data) + row) : __ldg(reinterpret_cast<const long long*>(c.data) + row);
    const bool ok = (m >> c.vbit) & 1u;
    v = ok ? x : 0;
    return ok;
}
__global__ void __launch_bounds__(SC_NT) k3p_scatter_staged(const __grid_constant__ AggInput in, int nk, int nv, int pbits, int pass,
                                                            unsigned int* counts, const unsigned int* bases, unsigned int* cursors,
                                                            unsigned long long* recs, unsigned int stride, unsigned long long* overflow) {
    extern __shared__ __align__(16) unsigned long long s_stage[];      // SC_TILE x rw words, then SC_TILE positions (uint32)
    const int rw = nk + nv + 3, tid = threadIdx.x;
    unsigned int* s_pos = reinterpret_cast<unsigned int*>(s_stage + (size_t)SC_TILE * rw);
    const double inv_pane = 1.0 / (double)in.pane_us;
    const long long n_tiles = (in.n_rows + SC_TILE - 1) / SC_TILE;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        unsigned int part[SC_RPT]; bool live[SC_RPT];
#pragma unroll
        for (int j = 0; j < SC_RPT; j++) {
            const long long i = tile * SC_TILE + (long long)j * SC_NT + tid;
            live[j] = false; part[j] = 0;
            if (i >= in.n_rows) continue;
            const long long row = in.sel ? in.sel[i] : i;
            long long t;
            const bool has_t = p_load_u(in.range_col, row, t);
            long long kv[SC_MAXK], xv[SC_MAXV];
            unsigned long long bits = 0, h = 0x2545F4914F6CDD1Dull;
#pragma unroll
            for (int k = 0; k < SC_MAXK; k++) if (k < nk) { if (!p_load_u(in.keys[k], row, kv[k])) bits |= 1ull << k; }
#pragma unroll
            for (int v = 0; v < SC_MAXV; v++) if (v < nv) { if (p_load_u(in.vals[v], row, xv[v])) bits |= 1ull << (16 + v); }
            if (!has_t || (in.has_range && (t < in.range_lo || t >= in.range_hi))) continue;      // no timestamp: in no window; below a cached batch's late-event cut: not in the batch
#pragma unroll
            for (int k = 0; k < SC_MAXK; k++) if (k < nk) h = pmix64(h ^ (unsigned long long)kv[k]);
            h = pmix64(h ^ (bits & 0xFFFFull));
            const long long pane = p_floor_div(t, in.pane_us, inv_pane);
            part[j] = pbits ? (unsigned int)(h >> (64 - pbits)) : 0u;
            live[j] = true;
            if (pass == 0) continue;
            unsigned long long* r = s_stage + (size_t)(j * SC_NT + tid) * rw;
            r[0] = h;
#pragma unroll
            for (int k = 0; k < SC_MAXK; k++) if (k < nk) r[1 + k] = (unsigned long long)kv[k];
            r[1 + nk] = (unsigned long long)pane;
            r[2 + nk] = bits;
#pragma unroll
            for (int v = 0; v < SC_MAXV; v++) if (v < nv) r[3 + nk + v] = (unsigned long long)xv[v];
        }
        if (pass == 0) {
#pragma unroll
            for (int j = 0; j < SC_RPT; j++) if (live[j]) atomicAdd(&counts[part[j]], 1u);
            continue;
        }
        // slots: SC_RPT independent atomics in flight per thread
        unsigned int at[SC_RPT];
#pragma unroll
        for (int j = 0; j < SC_RPT; j++) at[j] = live[j] ? atomicAdd(&cursors[part[j]], 1u) : 0u;
#pragma unroll
        for (int j = 0; j < SC_RPT; j++) {
            unsigned int pos = 0xFFFFFFFFu;
            if (live[j]) {
                if (pass == 2) { if (at[j] >= stride) *overflow = 1ull; else pos = part[j] * stride + at[j]; }
                else pos = bases[part[j]] + at[j];
            }
            s_pos[j * SC_NT + tid] = pos;
        }
        __syncthreads();
        // copy-out without a division: a thread's word index within a record is its id modulo the next power of two >= the
        // record length (the lanes past the record's end idle: 1/4 of them for 6 words, 1/8 for 7)
        if ((rw & 1) == 0) {
            const int rq = rw >> 1;                                    // 16-byte chunks per record
            int sh = 0; while ((1 << sh) < rq) sh++;
            const unsigned int w = tid & ((1u << sh) - 1u), r0 = tid >> sh, rstep = SC_NT >> sh;
            const uint4* src = reinterpret_cast<const uint4*>(s_stage);
            uint4* dst = reinterpret_cast<uint4*>(recs);
            if (w < (unsigned int)rq)
                for (unsigned int rec = r0; rec < SC_TILE; rec += rstep) {
                    const unsigned int pos = s_pos[rec];
                    if (pos != 0xFFFFFFFFu) dst[(size_t)pos * rq + w] = src[rec * rq + w];
                }
        } else {
            int sh = 0; while ((1 << sh) < rw) sh++;
            const unsigned int w = tid & ((1u << sh) - 1u), r0 = tid >> sh, rstep = SC_NT >> sh;
            if (w < (unsigned int)rw)
                for (unsigned int rec = r0; rec < SC_TILE; rec += rstep) {
                    const unsigned int pos = s_pos[rec];
                    if (pos != 0xFFFFFFFFu) recs[(size_t)pos * rw + w] = s_stage[rec * rw + w];
                }
        }
        __syncthreads();
    }
}
static size_t sc_smem(int rw) { return (size_t)SC_TILE * rw * 8 + (size_t)SC_TILE * 4; }

// ---- phase 2: one CTA per partition, shared-memory hash table --------------------------------------------------
// part_base == nullptr: partition p is the fixed slice [p * stride, p * stride + min(part_cnt[p], stride))
struct P2Seg { const unsigned long long* recs; const unsigned int* part_base; const unsigned int* part_cnt; unsigned long long stride; };
struct P2Args {
    const P2Seg* segs; int n_seg;
    int rw;                 // words per input record
    int nkk;                // identity key words at [1, 1+nkk) (build: group keys + pane; combine: group keys)
    int pane_in_key;        // build: the last identity word is the pane, mixed into the table tag
    int bits_word, val0, nv;
    int range_word; long long range_lo, range_hi;    // combine: keep records with lo <= rec[range_word] < hi (-1: none)
    int na; AggAccDesc accs[DX_AGG_MAX_ACCS];         // col indexes the value words; ACC_COUNT_STAR has none
    int check_valid;        // build: honour the value-present bits; combine: every value word counts
    unsigned long long* out; int gw;                  // output groups: [hash][keys nkk][key-null bits][accs na]
    unsigned int* out_base; unsigned int* out_cnt; unsigned long long* counters;   // [0] groups written, [1] overflow, [2] output too small
    unsigned long long out_cap;                       // groups the output holds
};

// 64-bit shared-memory atomics have no native form on sm_100 (add, min, max and double add all compile to LDS + CAS loops,
// SASS ATOMS.CAST.SPIN.64).  Integer sums therefore go through two native 32-bit adds with an exact carry (the wraps of the
// low word are counted one by one, whatever order the adds land in), counts through one, and MIN / MAX look before they
// swap: after a group's first few rows almost no row changes its extremes.
static __device__ __forceinline__ void s_add64(unsigned long long* p, unsigned long long v) {
    unsigned int* w = reinterpret_cast<unsigned int*>(p);
    const unsigned int lo = (unsigned int)v, hi = (unsigned int)(v >> 32);
    unsigned int carry = 0;
    if (lo) { const unsigned int old = atomicAdd(&w[0], lo); carry = (old + lo < old) ? 1u : 0u; }
    if (hi + carry) atomicAdd(&w[1], hi + carry);
}
static __device__ __forceinline__ void s_min64(unsigned long long* p, long long v) {
    if (v < *reinterpret_cast<volatile long long*>(p)) atomicMin(reinterpret_cast<long long*>(p), v);
}
static __device__ __forceinline__ void s_max64(unsigned long long* p, long long v) {
    if (v > *reinterpret_cast<volatile long long*>(p)) atomicMax(reinterpret_cast<long long*>(p), v);
}

__global__ void __launch_bounds__(K3P_NT) k3p_aggregate(const __grid_constant__ P2Args A) {
    extern __shared__ __align__(16) unsigned long long tab[];      // K3P_SLOTS x sw: [tag][hash][keys nkk][bits][accs na]
    __shared__ unsigned int s_groups, s_over, s_base;
    __shared__ unsigned int s_wcnt[K3P_NT / 32];
    const int sw = A.gw + 1, tid = threadIdx.x;
    const unsigned int p = blockIdx.x;
    for (int i = tid; i < K3P_SLOTS; i += K3P_NT) tab[(size_t)i * sw] = 0ull;
    if (tid == 0) { s_groups = 0; s_over = 0; }
    __syncthreads();
    // The partition's records of up to K3P_NT segments at a time form one flat index space: the segment descriptors are
    // fetched by one thread each (one round of global latency instead of one per segment) and every thread walks the
    // concatenation, so short segments (a window over many cached batches) still keep the whole CTA busy.
    __shared__ unsigned int s_pref[K3P_NT + 1];
    __shared__ const unsigned long long* s_ptr[K3P_NT];
    for (int s0 = 0; s0 < A.n_seg; s0 += K3P_NT) {
        const int ns = min(K3P_NT, A.n_seg - s0);
        __syncthreads();
        unsigned int mycnt = 0;
        if (tid < ns) {
            const P2Seg sg = A.segs[s0 + tid];
            mycnt = sg.part_cnt[p];
            if (sg.part_base) s_ptr[tid] = sg.recs + (size_t)sg.part_base[p] * A.rw;
            else { if (mycnt > sg.stride) mycnt = (unsigned int)sg.stride; s_ptr[tid] = sg.recs + (size_t)p * sg.stride * A.rw; }
        }
        {   // exclusive prefix of the counts over the CTA
            unsigned int incl = mycnt;
            const int ln = tid & 31, wp = tid >> 5;
            for (int d = 1; d < 32; d <<= 1) { unsigned int t2 = __shfl_up_sync(0xffffffffu, incl, d); if (ln >= d) incl += t2; }
            if (ln == 31) s_wcnt[wp] = incl;
            __syncthreads();
            unsigned int wb = 0; for (int w = 0; w < wp; w++) wb += s_wcnt[w];
            s_pref[tid + 1] = wb + incl;
            if (tid == 0) s_pref[0] = 0;
            __syncthreads();
        }
        const unsigned int total_recs = s_pref[ns];
        if (tid == 0 && A.pane_in_key) atomicMax(&A.counters[4], (unsigned long long)total_recs);     // build: the fullest partition sizes the next batch's slices
        int seg = 0;
        // every thread makes the same number of trips, so the warp can be re-joined (__syncwarp) between the divergent slot
        // search and the accumulator updates: without it the updates run with a handful of active lanes
        for (unsigned int i0 = 0; i0 < total_recs; i0 += K3P_NT) {
            const unsigned int i = i0 + tid;
            bool has = i < total_recs;
            const unsigned long long* r = nullptr;
            if (has) {
                while (s_pref[seg + 1] <= i) seg++;
                r = s_ptr[seg] + (size_t)(i - s_pref[seg]) * A.rw;
                if (A.range_word >= 0) { const long long pn = (long long)r[A.range_word]; if (pn < A.range_lo || pn >= A.range_hi) has = false; }
                if (*reinterpret_cast<volatile unsigned int*>(&s_over)) has = false;
            }
            unsigned long long* slot = nullptr; unsigned long long bits = 0;
            if (has) {
            const unsigned long long h0 = r[0], knull = (bits = r[A.bits_word]) & 0xFFFFull;
            unsigned long long tag = h0;
            if (A.pane_in_key) tag = pmix64(h0 ^ pmix64(r[A.nkk]));
            tag |= 0x8000000000000000ull;
            unsigned int idx = (unsigned int)(tag >> 11) & (K3P_SLOTS - 1);
            for (int probes = 0; probes < K3P_SLOTS; ) {
                unsigned long long* sl = tab + (size_t)idx * sw;
                const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(sl);
                if (cur == tag) {
                    const volatile unsigned long long* vs = sl;     // ordered after the volatile read of the tag
                    bool same = vs[2 + A.nkk] == knull;
                    for (int k = 0; k < A.nkk && same; k++) same = vs[2 + k] == r[1 + k];
                    if (same) { slot = sl; break; }
                } else if (cur == 0ull) {
                    if (atomicCAS(sl, 0ull, 1ull) == 0ull) {
                        sl[1] = h0;
                        for (int k = 0; k < A.nkk; k++) sl[2 + k] = r[1 + k];
                        sl[2 + A.nkk] = knull;
                        for (int a = 0; a < A.na; a++) {
                            const int kind = A.accs[a].kind;
                            unsigned long long init = 0;
                            if (kind == ACC_MIN_I || kind == ACC_MIN_F) init = 0x7FFFFFFFFFFFFFFFull;
                            else if (kind == ACC_MAX_I || kind == ACC_MAX_F) init = 0x8000000000000000ull;
                            sl[3 + A.nkk + a] = init;
                        }
                        __threadfence_block();
                        *reinterpret_cast<volatile unsigned long long*>(sl) = tag;
                        if (atomicAdd(&s_groups, 1u) + 1 > K3P_MAX_GROUPS) s_over = 1;
                        slot = sl; break;
                    }
                    continue;
                } else if (cur == 1ull) continue;
                idx = (idx + 1) & (K3P_SLOTS - 1); probes++;
            }
            if (!slot) s_over = 1;
            }
            __syncwarp();
            if (!slot) continue;
            unsigned long long* acc = slot + 3 + A.nkk;
            for (int a = 0; a < A.na; a++) {
                const int kind = A.accs[a].kind;
                if (kind == ACC_COUNT_STAR) { s_add64(&acc[a], 1ull); continue; }
                const int c = A.accs[a].col;
                if (A.check_valid && !((bits >> (16 + c)) & 1ull)) continue;
                const long long v = (long long)r[A.val0 + c];
                switch (kind) {
                case ACC_COUNT: s_add64(&acc[a], 1ull); break;
                case ACC_SUM_I: s_add64(&acc[a], (unsigned long long)v); break;
                case ACC_SUM_F: atomicAdd(reinterpret_cast<double*>(&acc[a]), __longlong_as_double(v)); break;
                case ACC_MIN_I: s_min64(&acc[a], v); break;
                case ACC_MAX_I: s_max64(&acc[a], v); break;
                case ACC_MIN_F: s_min64(&acc[a], p_f64_order(v)); break;
                case ACC_MAX_F: s_max64(&acc[a], p_f64_order(v)); break;
                }
            }
        }
    }
    __syncthreads();
    if (s_over) { if (tid == 0) { A.counters[1] = 1ull; A.out_base[p] = 0; A.out_cnt[p] = 0; } return; }
    // compact the occupied slots into one contiguous run of the output (slot order within the partition)
    constexpr int PER = K3P_SLOTS / K3P_NT;
    unsigned int mine = 0;
    for (int j = 0; j < PER; j++) if (tab[(size_t)(tid * PER + j) * sw] >= 2ull) mine++;
    unsigned int inc = mine;
    const int lane = tid & 31, warp = tid >> 5;
    for (int d = 1; d < 32; d <<= 1) { unsigned int t2 = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t2; }
    if (lane == 31) s_wcnt[warp] = inc;
    __syncthreads();
    unsigned int wbase = 0, total = 0;
    for (int w = 0; w < K3P_NT / 32; w++) { if (w < warp) wbase += s_wcnt[w]; total += s_wcnt[w]; }
    if (tid == 0) { s_base = (unsigned int)atomicAdd(&A.counters[0], (unsigned long long)total); A.out_base[p] = s_base; A.out_cnt[p] = total; }
    __syncthreads();
    if ((unsigned long long)s_base + total > A.out_cap) { if (tid == 0) A.counters[2] = 1ull; return; }   // counted, not written: the caller retries with room
    unsigned int at = s_base + wbase + inc - mine;
    for (int j = 0; j < PER; j++) {
        const unsigned long long* sl = tab + (size_t)(tid * PER + j) * sw;
        if (sl[0] < 2ull) continue;
        unsigned long long* o = A.out + (size_t)at * A.gw;
        for (int w = 0; w < A.gw; w++) {
            unsigned long long v = sl[1 + w];
            const int a = w - (2 + A.nkk);
            if (a >= 0 && (A.accs[a].kind == ACC_MIN_F || A.accs[a].kind == ACC_MAX_F)) v = (unsigned long long)p_f64_order((long long)v);
            o[w] = v;
        }
        at++;
    }
}

// groups [hash][keys nk][pane?][bits][accs] -> the column arrays exec_agg finalises from
__global__ void k3p_to_columns(const unsigned long long* recs, long long n, int gw, int nk, int bits_word, int acc0, int na,
                               long long* const* key_out, unsigned long long* keynull_out, unsigned long long* const* acc_out) {
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n; g += (long long)gridDim.x * blockDim.x) {
        const unsigned long long* r = recs + (size_t)g * gw;
        for (int k = 0; k < nk; k++) key_out[k][g] = (long long)r[1 + k];
        keynull_out[g] = r[bits_word] & 0xFFFFull;
        for (int a = 0; a < na; a++) acc_out[a][g] = r[acc0 + a];
    }
}

static size_t p2_smem(int gw) { return (size_t)K3P_SLOTS * (gw + 1) * 8; }

static int p2_launch(dx_ctx* ctx, const P2Args& A, int P) {
    const size_t smem = p2_smem(A.gw);
    // the attribute belongs to the (function, device) pair: remembered per context, i.e. per device, not process-wide
    if (smem > ctx->k3p_smem_configured) {
        DX_CUDA(ctx, cudaFuncSetAttribute(k3p_aggregate, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ctx->k3p_smem_configured = smem;
    }
    StageTimer tm(ctx, "k3p_aggregate");
    k3p_aggregate<<<(unsigned)P, K3P_NT, smem, ctx->stream>>>(A);
    ctx->launches["k3p_aggregate"]++;
    DX_CUDA(ctx, cudaGetLastError());
    return DX_OK;
}

bool dx_aggp_fits(int nk, int na) { return p2_smem(1 + (nk + 1) + 1 + na) <= 200 * 1024 && nk + 1 <= 15 && na <= DX_AGG_MAX_ACCS; }

int dx_aggp_build(dx_ctx* ctx, const AggInput& in_, int nk, int nv, const std::vector<AggAccDesc>& accs, int P, AggPPartial* out, bool* overflow,
                  long long* stride_hint, long long* hint_rows, bool force_two_pass) {
    *overflow = false;
    const long long t_b0 = now_us();
    out->n = 0; out->P = P; out->gw = 1 + (nk + 1) + 1 + (int)accs.size();
    int pbits = 0; while ((1 << pbits) < P) pbits++;
    if ((1 << pbits) != P) return ctx->set_error(DX_ERR_ARG, "partition count must be a power of two");
    AggInput in = in_;
    const int rw = nk + nv + 3;
    DevBuf counts(ctx, (size_t)P * 4 + 16), bases(ctx, (size_t)P * 4 + 16), cursors(ctx, (size_t)P * 4 + 16), ctr(ctx, 64);
    out->part_base = make_buf(ctx, (size_t)P * 4 + 16); out->part_cnt = make_buf(ctx, (size_t)P * 4 + 16);
    if (!counts.p || !bases.p || !cursors.p || !ctr.p || !out->part_base->p || !out->part_cnt->p) return ctx->set_error(DX_ERR_CUDA, "pane partial allocation failed");
    DX_CUDA(ctx, cudaMemsetAsync(counts.p, 0, (size_t)P * 4, ctx->stream));
    DX_CUDA(ctx, cudaMemsetAsync(cursors.p, 0, (size_t)P * 4, ctx->stream));
    DX_CUDA(ctx, cudaMemsetAsync(ctr.p, 0, 64, ctx->stream));
    if (in.n_rows <= 0) {
        DX_CUDA(ctx, cudaMemsetAsync(out->part_base->p, 0, (size_t)P * 4, ctx->stream));
        DX_CUDA(ctx, cudaMemsetAsync(out->part_cnt->p, 0, (size_t)P * 4, ctx->stream));
        out->recs = make_buf(ctx, 64);
        return DX_OK;
    }
    long long blocks = (in.n_rows + 255) / 256, capb = (long long)ctx->sm_count * 8; if (blocks > capb) blocks = capb;
    ctx->counters["k3_rows"] += in.n_rows;
    const bool staged = nk <= SC_MAXK && nv <= SC_MAXV;
    const size_t sc_bytes = sc_smem(rw);
    long long sblocks = (in.n_rows + SC_TILE - 1) / SC_TILE;
    if (staged) {
        if (sc_bytes > ctx->k3p_scatter_smem_configured) {
            DX_CUDA(ctx, cudaFuncSetAttribute(k3p_scatter_staged, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sc_bytes));
            ctx->k3p_scatter_smem_configured = sc_bytes;
        }
        int occ = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k3p_scatter_staged, SC_NT, sc_bytes);
        if (occ < 1) occ = 1;
        const long long scap = (long long)ctx->sm_count * occ; if (sblocks > scap) sblocks = scap;
    }
    auto scatter = [&](int pass, unsigned int* cnt, const unsigned int* bas, unsigned int* cur, unsigned long long* rec, unsigned int strd, unsigned long long* ovf) {
        StageTimer tm(ctx, "k3p_scatter");
        if (staged) k3p_scatter_staged<<<(unsigned)sblocks, SC_NT, sc_bytes, ctx->stream>>>(in, nk, nv, pbits, pass, cnt, bas, cur, rec, strd, ovf);
        else k3p_scatter<<<(unsigned)blocks, 256, 0, ctx->stream>>>(in, nk, nv, pbits, pass, cnt, bas, cur, rec, strd, ovf);
        ctx->launches["k3p_scatter"]++;
    };
    // Single pass when possible: every partition gets a fixed slice of records and the cursor alone places a record — no
    // counting pass, no scan.  A partition that outgrows its slice (keys skewed since the last batch) raises a flag that comes
    // back with the aggregation counters below, and the exact two-pass form runs instead.
    // (rows of one key land in one partition, so the spread is wider than a binomial's: the slice is sized from the fullest
    // partition the previous batch of this statement produced, with 25 % headroom; the first batch runs the two-pass form)
    const unsigned long long stride = stride_hint ? (unsigned long long)((double)*stride_hint * ((double)in.n_rows / (double)std::max<long long>(1, hint_rows ? *hint_rows : in.n_rows)) * 1.25 + 64.0) : 0ull;
    bool single_pass = !force_two_pass && stride_hint && *stride_hint > 0 && stride * (unsigned long long)P < (1ull << 32) &&
                       stride * (unsigned long long)P <= (unsigned long long)in.n_rows * 8 + 128ull * P;
    DevBuf recs;
    ctx->counters[single_pass ? "k3p_single_pass_builds" : "k3p_two_pass_builds"]++;
    if (single_pass) {
        recs = DevBuf(ctx, (size_t)stride * P * rw * 8 + 64);
        if (!recs.p) return ctx->set_error(DX_ERR_CUDA, "pane partial allocation failed");
        scatter(2, nullptr, nullptr, cursors.as<unsigned int>(), recs.as<unsigned long long>(), (unsigned int)stride, ctr.as<unsigned long long>() + 3);
    } else {
        scatter(0, counts.as<unsigned int>(), nullptr, nullptr, nullptr, 0u, nullptr);
        size_t tmp_bytes = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts.as<unsigned int>(), bases.as<unsigned int>(), P, ctx->stream);
        DevBuf tmp(ctx, tmp_bytes + 16);
        DX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, counts.as<unsigned int>(), bases.as<unsigned int>(), P, ctx->stream));
        // every row with a timestamp becomes one record: at most n_rows of them
        recs = DevBuf(ctx, (size_t)in.n_rows * rw * 8 + 64);
        if (!recs.p) return ctx->set_error(DX_ERR_CUDA, "pane partial allocation failed");
        scatter(1, nullptr, bases.as<unsigned int>(), cursors.as<unsigned int>(), recs.as<unsigned long long>(), 0u, nullptr);
    }
    DX_CUDA(ctx, cudaGetLastError());
    // aggregate each partition by (keys, pane)
    P2Seg hseg{recs.as<unsigned long long>(), single_pass ? nullptr : bases.as<unsigned int>(), single_pass ? cursors.as<unsigned int>() : counts.as<unsigned int>(), stride};
    DevBuf dseg(ctx, sizeof(P2Seg));
    DX_CUDA(ctx, cudaMemcpyAsync(dseg.p, &hseg, sizeof hseg, cudaMemcpyHostToDevice, ctx->stream));
    // groups <= records; the output is sized by P * K3P_MAX_GROUPS or the row count, whichever is smaller
    const long long max_groups = std::min<long long>(in.n_rows, (long long)P * K3P_MAX_GROUPS);
    DevBufP groups = make_buf(ctx, (size_t)max_groups * out->gw * 8 + 64);
    if (!groups->p) return ctx->set_error(DX_ERR_CUDA, "pane partial allocation failed");
    P2Args A; memset(&A, 0, sizeof A);
    A.segs = dseg.as<P2Seg>(); A.n_seg = 1; A.rw = rw; A.nkk = nk + 1; A.pane_in_key = 1;
    A.bits_word = 2 + nk; A.val0 = 3 + nk; A.nv = nv; A.range_word = -1;
    A.na = (int)accs.size(); for (size_t a = 0; a < accs.size(); a++) A.accs[a] = accs[a];
    A.check_valid = 1;
    A.out = groups->as<unsigned long long>(); A.gw = out->gw; A.out_cap = (unsigned long long)max_groups;
    A.out_base = out->part_base->as<unsigned int>(); A.out_cnt = out->part_cnt->as<unsigned int>(); A.counters = ctr.as<unsigned long long>();
    int rc = p2_launch(ctx, A, P); if (rc) return rc;
    unsigned long long h[5] = {0, 0, 0, 0, 0};
    DX_CUDA(ctx, cudaMemcpyAsync(h, ctr.p, 40, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->counters["host_us:k3p_build_total_to_agg"] += now_us() - t_b0;
    if (h[3]) {      // a partition outgrew its fixed slice (skewed keys): the exact two-pass scatter
        ctx->counters["k3p_two_pass_fallbacks"]++;
        return dx_aggp_build(ctx, in_, nk, nv, accs, P, out, overflow, stride_hint, hint_rows, true);
    }
    if (stride_hint) { *stride_hint = (long long)h[4]; if (hint_rows) *hint_rows = in.n_rows; }      // fullest partition of this batch
    if (h[1] || h[2]) { *overflow = true; return DX_OK; }
    out->n = (int64_t)h[0];
    // keep only what the groups need (the worst-case buffer can be far larger than the partial)
    if ((size_t)out->n * 2 < (size_t)max_groups) {
        DevBufP tight = make_buf(ctx, (size_t)(out->n > 0 ? out->n : 1) * out->gw * 8 + 64);
        if (!tight->p) return ctx->set_error(DX_ERR_CUDA, "pane partial allocation failed");
        DX_CUDA(ctx, cudaMemcpyAsync(tight->p, groups->p, (size_t)out->n * out->gw * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        groups = tight;
    }
    out->recs = groups;
    return DX_OK;
}

int dx_aggp_combine(dx_ctx* ctx, const std::vector<const AggPPartial*>& parts, int nk, const std::vector<AggAccDesc>& accs, const std::vector<int>& src_acc,
                    long long pane_lo, long long pane_hi, long long max_groups, long long* const* key_out,
                    unsigned long long* keynull_out, unsigned long long* const* acc_out, long long* n_groups, bool* overflow) {
    *overflow = false; *n_groups = 0;
    if (parts.empty()) return DX_OK;
    const int P = parts[0]->P, gw_in = parts[0]->gw, na = (int)accs.size();
    std::vector<P2Seg> hsegs;
    for (auto* pp : parts) {
        if (pp->P != P || pp->gw != gw_in) return ctx->set_error(DX_ERR_STATE, "pane partials of different shapes cannot be combined");
        if (pp->n > 0) hsegs.push_back(P2Seg{pp->recs->as<unsigned long long>(), pp->part_base->as<unsigned int>(), pp->part_cnt->as<unsigned int>(), 0ull});
    }
    if (hsegs.empty()) return DX_OK;
    const int gw_out = 1 + nk + 1 + na;
    DevBuf dseg(ctx, hsegs.size() * sizeof(P2Seg)), ctr(ctx, 64), obase(ctx, (size_t)P * 4 + 16), ocnt(ctx, (size_t)P * 4 + 16);
    const long long cap = std::min<long long>(max_groups, (long long)P * K3P_MAX_GROUPS);
    DevBuf groups(ctx, (size_t)cap * gw_out * 8 + 64);
    if (!dseg.p || !ctr.p || !obase.p || !ocnt.p || !groups.p) return ctx->set_error(DX_ERR_CUDA, "pane combine allocation failed");
    DX_CUDA(ctx, cudaMemcpyAsync(dseg.p, hsegs.data(), hsegs.size() * sizeof(P2Seg), cudaMemcpyHostToDevice, ctx->stream));
    DX_CUDA(ctx, cudaMemsetAsync(ctr.p, 0, 64, ctx->stream));
    P2Args A; memset(&A, 0, sizeof A);
    A.segs = dseg.as<P2Seg>(); A.n_seg = (int)hsegs.size(); A.rw = gw_in; A.nkk = nk; A.pane_in_key = 0;
    A.range_word = 1 + nk; A.range_lo = pane_lo; A.range_hi = pane_hi;
    A.bits_word = 2 + nk; A.val0 = 3 + nk; A.nv = na;
    A.na = na;
    for (int a = 0; a < na; a++) {      // COUNTs become sums of the partial counts; everything else keeps its operator
        A.accs[a].kind = (accs[a].kind == ACC_COUNT_STAR || accs[a].kind == ACC_COUNT) ? ACC_SUM_I : accs[a].kind;
        A.accs[a].col = src_acc[a];          // where this accumulator sits among the partial's (the partial may carry more)
    }
    A.check_valid = 0;
    A.out = groups.as<unsigned long long>(); A.gw = gw_out; A.out_cap = (unsigned long long)cap;
    A.out_base = obase.as<unsigned int>(); A.out_cnt = ocnt.as<unsigned int>(); A.counters = ctr.as<unsigned long long>();
    long long rows = 0; for (auto* pp : parts) rows += pp->n;
    ctx->counters["k3_rows"] += rows;
    int rc = p2_launch(ctx, A, P); if (rc) return rc;
    unsigned long long h[3] = {0, 0, 0};
    DX_CUDA(ctx, cudaMemcpyAsync(h, ctr.p, 24, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (h[1]) { *overflow = true; return DX_OK; }
    if (h[2] || (long long)h[0] > max_groups) { *overflow = true; *n_groups = -1; return DX_OK; }   // -1: the output arrays were too small, not the shared tables
    *n_groups = (long long)h[0];
    if (*n_groups == 0) return DX_OK;
    // column arrays for the caller
    DevBuf ptrs(ctx, (DX_AGG_MAX_KEYS + DX_AGG_MAX_ACCS) * 8 + 16);
    void* host_ptrs[DX_AGG_MAX_KEYS + DX_AGG_MAX_ACCS]; memset(host_ptrs, 0, sizeof host_ptrs);
    for (int k = 0; k < nk; k++) host_ptrs[k] = key_out[k];
    for (int a = 0; a < na; a++) host_ptrs[DX_AGG_MAX_KEYS + a] = acc_out[a];
    DX_CUDA(ctx, cudaMemcpyAsync(ptrs.p, host_ptrs, sizeof host_ptrs, cudaMemcpyHostToDevice, ctx->stream));
    k3p_to_columns<<<(unsigned)std::min<long long>((*n_groups + 255) / 256, 4096), 256, 0, ctx->stream>>>(
        groups.as<unsigned long long>(), *n_groups, gw_out, nk, 1 + nk, 2 + nk, na,
        reinterpret_cast<long long* const*>(ptrs.p), keynull_out, reinterpret_cast<unsigned long long* const*>(ptrs.as<void*>() + DX_AGG_MAX_KEYS));
    ctx->launches["k3p_to_columns"]++;
    DX_CUDA(ctx, cudaGetLastError());
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return DX_OK;
}
