// dx_decode_kernel.cuh — the K1 kernel template (see dx_decode.cu for the description).  Included by dx_decode.cu, which
// instantiates the NDJSON kernels (FRAMED = false: structural scan inside the kernel), and by dx_decode_framed.cu, which
// instantiates the kernels for caller-provided record offsets (FRAMED = true); two translation units so they build in parallel.
#pragma once
#include "dx_decode.cuh"

#ifndef K1_NT
#define K1_NT 256           // threads per CTA
#endif
#define K1_TS(x) ((x) * K1_NT / 256)     // tile sizes are stated for 256 threads and scale with the CTA
#ifndef K1_T_DEFAULT
#define K1_T_DEFAULT K1_TS(32768)  // tile bytes when the record length is unknown (see k1_pick_tile)
#endif
#ifndef K1_OVL
#define K1_OVL 2048         // overlap bytes past the tile (records starting in the tile may end here)
#endif
#define K1_PAD 32
#define K1_SC 1024          // (caller-framed kernels only) kept for the shared-memory layout
#define K1_WCAP 126         // record starts a warp holds per round (a slice normally has ~30 records)
#ifndef K1_STAGES
#define K1_STAGES 1         // shared-memory tile stages; 2 = prefetch the next tile while parsing (measured slower: the resident CTAs of an SM already overlap copy and parse)
#endif
#ifndef K1_MIN_CTAS_FRAMED
#define K1_MIN_CTAS_FRAMED 5   // caller-provided framing: no scan state; the record parse alone runs best at 48 registers (2.74 -> 2.51 ms per 25 M events)
#endif
#ifndef K1_MIN_CTAS
#define K1_MIN_CTAS 4       // resident CTAs per SM the register allocation is held to (64 registers; measured 3 / 4 / 5: 3.27 / 3.03 / 3.03 ms per 25 M events)
#endif

// ------------------------------- PTX wrappers ------------------------------------------
static __device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
static __device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
static __device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
static __device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
static __device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
static __device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
static __device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// ------------------------------- shared-memory readers ---------------------------------
static __device__ __forceinline__ uint32_t ld32u(const uint8_t* buf, uint32_t pos) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(buf + (pos & ~3u));
    return __funnelshift_r(w[0], w[1], (pos & 3u) << 3);
}
#include "dx_decode_fast.cuh"

static __device__ __forceinline__ unsigned long long ld_state(const unsigned long long* p) {
    return *reinterpret_cast<const volatile unsigned long long*>(p);
}

// Decoupled look-back over per-tile record counts; executed by one full warp.
static __device__ long long tile_lookback(unsigned long long* state, long long tile, unsigned long long my_count) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned long long VAL = (1ull << 62) - 1;
    if (tile == 0) {
        if (lane == 0) atomicExch(&state[0], (2ull << 62) | my_count);
        return 0;
    }
    if (lane == 0) atomicExch(&state[tile], (1ull << 62) | my_count);
    long long excl = 0;
    long long j = tile - 1;
    for (;;) {
        long long idx = j - lane;
        unsigned long long v;
        do {
            v = (idx >= 0) ? ld_state(&state[idx]) : (2ull << 62);
        } while (__any_sync(full, (v >> 62) == 0));
        unsigned pmask = __ballot_sync(full, (v >> 62) == 2);
        int first = pmask ? (__ffs(pmask) - 1) : 31;
        long long contrib = (lane <= first) ? (long long)(v & VAL) : 0;
        for (int s = 16; s > 0; s >>= 1) contrib += __shfl_xor_sync(full, contrib, s);
        excl += contrib;
        if (pmask) break;
        j -= 32;
    }
    if (lane == 0) atomicExch(&state[tile], (2ull << 62) | (unsigned long long)(excl + (long long)my_count));
    return excl;
}

// Only the first and the last tile of a batch can hold newline candidates outside the data; kept out of line so the nine
// unrolled scan rounds stay compact in the instruction cache.
static __device__ __noinline__ uint32_t k1_edge_mask(uint32_t m, long long q0, long long n_end, long long skew) {
    if (q0 + 16 >= n_end) {
        for (int i = 0; i < 16; i++) if (q0 + i + 1 >= n_end) m &= ~(1u << i);
    }
    if (q0 < skew) {
        for (int i = 0; i < 16; i++) if (q0 + i < skew) m &= ~(1u << i);
    }
    return m;
}

// The tile size is a template parameter chosen per launch from the batch's mean record length, so that a tile holds just
// under K1_NT records and the one-lane-per-record parse keeps (nearly) every lane of every warp busy.
template <int K1_T, bool FRAMED>
__global__ void __launch_bounds__(K1_NT, FRAMED ? K1_MIN_CTAS_FRAMED : K1_MIN_CTAS) k1_decode(const __grid_constant__ DecodeArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int TILE_BYTES = K1_T + K1_OVL + K1_PAD;                     // one staged tile (multiple of 16)
    uint16_t* s_starts = reinterpret_cast<uint16_t*>(smem + K1_STAGES * TILE_BYTES);   // per warp: chunk list (K1_T / 128 entries) + K1_WCAP + 2 record starts
    __shared__ uint64_t s_bar[K1_STAGES];
    __shared__ long long s_tile;
    __shared__ long long s_rowbase;
    __shared__ uint32_t s_wtot[K1_NT / 32], s_wtail[K1_NT / 32], s_next;
    __shared__ long long s_r0;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const DSchemaDev* S = a.schema;
    if (tid == 0) { for (int k = 0; k < K1_STAGES; k++) mbar_init(&s_bar[k], 1); fence_mbar_init(); }
    __syncthreads();

    const uint32_t n_leaves = (uint32_t)S->n_leaves;
    const uint32_t all_valid = (n_leaves >= 32) ? 0xFFFFFFFFu : ((1u << n_leaves) - 1u);
    uint32_t c_rows = 0, c_null = 0, c_fast = 0, c_err = 0, c_lim = 0;    // per thread (a thread sees far fewer than 2^32 records); widened when added to the statistics
    uint32_t phase_bits = 0;     // bit k: parity the next completion of stage k will have

    // Stages one tile: bulk (TMA) copy of the 16-byte-aligned part onto the stage's mbarrier, the ragged
    // tail (< 16 bytes) and the end-of-data sentinel by ordinary stores.  All threads call it.
    auto stage_tile = [&](int slot, long long tile) {
        uint8_t* dst = smem + slot * TILE_BYTES;
        const long long lo_ = tile * (long long)K1_T;
        long long av = a.n_end - lo_; if (av > K1_T + K1_OVL) av = K1_T + K1_OVL;
        const uint32_t bulk_ = (uint32_t)av & ~15u;
        if (tid == 0) {
            fence_proxy_async();
            if (bulk_) {
                mbar_expect_tx(&s_bar[slot], bulk_);
                uint32_t off = 0;
                while (off < bulk_) {
                    uint32_t n = bulk_ - off; if (n > 16384u) n = 16384u;
                    bulk_g2s(dst + off, a.base_al + lo_ + off, n, &s_bar[slot]);
                    off += n;
                }
            }
        }
        if (tid < (int)((uint32_t)av - bulk_)) dst[bulk_ + tid] = a.base_al[lo_ + bulk_ + tid];
        if (tid == (K1_NT > 32 ? 32 : 0) && lo_ + av == a.n_end) dst[av] = '\n';
    };

    if (tid == 0) s_tile = (long long)atomicAdd(a.tile_counter, 1u);
    __syncthreads();
    long long tile = s_tile;
    int slot = 0;
    if (tile < a.n_tiles) stage_tile(0, tile);
    for (;;) {
        if (tile >= a.n_tiles) break;
        long long next_tile = a.n_tiles;
        if (K1_STAGES > 1) {
            // claim the following tile now and start its copy into the other stage, so it lands while this one is parsed
            __syncthreads();
            if (tid == 0) s_tile = (long long)atomicAdd(a.tile_counter, 1u);
            __syncthreads();
            next_tile = s_tile;
            if (next_tile < a.n_tiles) stage_tile(slot ^ 1, next_tile);
        }
        uint8_t* buf = smem + slot * TILE_BYTES;
        const long long lo = tile * (long long)K1_T;                 // absolute (relative to base_al)
        long long avail = a.n_end - lo;                              // bytes of data from lo on
        if (avail > K1_T + K1_OVL) avail = K1_T + K1_OVL;
        const uint32_t bulk = (uint32_t)avail & ~15u;
        if (bulk) { while (!mbar_try_wait(&s_bar[slot], (phase_bits >> slot) & 1u)) {} phase_bits ^= 1u << slot; }
        __syncthreads();

        long long n_rec = 0; long long rec0 = 0;   // offsets mode: records [rec0, rec0+n_rec)
        const bool first_virtual = (tile == 0 && (long long)a.skew < a.n_end);
        if (!FRAMED) {
            // --- structural scan and parse, warp by warp: warp w owns the records whose preceding newline q lies in its slice
            // [w * SLICE, (w + 1) * SLICE) of the tile (newline q owns the record starting at q + 1).  A warp scans its own slice,
            // lists its own record starts and parses its own records: the only thing the warps of a CTA exchange is their record
            // counts (for the row numbers), so a tile costs two CTA barriers between load and parse instead of a CTA-wide scan,
            // prefix sum and record redistribution.
            // Scan: newlines are sparse (one per ~150 bytes), so a lane first only decides whether its 16 bytes hold ANY 0x0A
            // (one subtraction per word; never a miss, rarely a false alarm); the chunks that do are listed in shared memory by ballot rank,
            // and the exact per-byte masks, the data-bounds masking of a batch's first and last tile and the prefix sums run once
            // per slice over that list with every lane busy, not once per round with three lanes busy.
            constexpr int NWARP = K1_NT / 32;
            constexpr int SLICE = K1_T / NWARP;               // bytes per warp: a multiple of 512
            constexpr int ROUNDS = SLICE / 512;
            constexpr int LCAP = SLICE / 16;                  // chunks per slice: the list cannot overflow
            uint16_t* w_list = s_starts + warp * (LCAP + K1_WCAP + 2);
            uint16_t* w_starts = w_list + LCAP;               // K1_WCAP + 2 record starts of the round being parsed
            const bool edge_tile = lo < (long long)a.skew || lo + (long long)K1_T + 16 >= a.n_end;
            const uint32_t lt_mask = (1u << lane) - 1u;
            uint32_t n_list = 0;
#pragma unroll
            for (int r = 0; r < ROUNDS; r++) {
                const int chunk = warp * LCAP + r * 32 + lane;
                const uint4 v = *reinterpret_cast<const uint4*>(buf + chunk * 16);
                // "may hold a newline": a byte below 0x10 borrows and leaves its top bit set whatever the borrow coming in, so a newline is
                // never missed; bytes from 0x90 up (UTF-8 tails) and tabs also flag their chunk, which only costs them the exact mask below
                const uint32_t z = ((v.x - 0x10101010u) | (v.y - 0x10101010u) | (v.z - 0x10101010u) | (v.w - 0x10101010u)) & 0x80808080u;
                const uint32_t bal = __ballot_sync(0xffffffffu, z != 0);
                if (z) w_list[n_list + __popc(bal & lt_mask)] = (uint16_t)chunk;
                n_list += __popc(bal);
            }
            __syncwarp();
            // exact newline mask of one listed chunk (bit i: byte i is a newline that starts a record of this batch)
            auto chunk_mask = [&](uint32_t chunk) -> uint32_t {
                const uint4 v = *reinterpret_cast<const uint4*>(buf + chunk * 16);
                const uint32_t ws[4] = {v.x, v.y, v.z, v.w};
                uint32_t m = 0;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t x = ws[k] ^ 0x0A0A0A0Au;                 // zero byte where '\n'
                    uint32_t t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;           // exact per-byte zero test
                    t = ~(t | x | 0x7F7F7F7Fu);                             // 0x80 in each byte that was zero
                    m |= ((t * 0x00204081u) >> 28) << (k * 4);              // gather the four flags
                }
                if (edge_tile) m = k1_edge_mask(m, lo + (long long)chunk * 16, a.n_end, (long long)a.skew);
                return m;
            };
            const bool first_virtual_w = first_virtual && warp == 0;       // the batch's first record has no newline before it
            // one pass over the list: counts the warp's records and stores the starts of ranks [round0, round0 + K1_WCAP]
            auto list_pass = [&](uint32_t round0) -> uint32_t {
                uint32_t total = first_virtual_w ? 1u : 0u;
                if (first_virtual_w && round0 == 0 && lane == 0) w_starts[0] = (uint16_t)a.skew;
                for (uint32_t e0 = 0; e0 < n_list; e0 += 32) {
                    const uint32_t e = e0 + lane;
                    uint32_t chunk = 0, m = 0;
                    if (e < n_list) { chunk = w_list[e]; m = chunk_mask(chunk); }
                    const uint32_t c = __popc(m);
                    uint32_t incl = c;
#pragma unroll
                    for (int sft = 1; sft < 32; sft <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, sft); if (lane >= sft) incl += t; }
                    uint32_t rel = total + incl - c - round0;              // wraps to a huge value for ranks of earlier rounds
                    const uint32_t pos0 = chunk * 16 + 1;
                    while (m) {
                        const int i = __ffs(m) - 1; m &= m - 1;
                        if (rel <= K1_WCAP) w_starts[rel] = (uint16_t)(pos0 + i);
                        rel++;
                    }
                    total += __shfl_sync(0xffffffffu, incl, 31);
                }
                return total;
            };
            const uint32_t total_w = list_pass(0);
            // the warp's last record ends at the first newline at or past the slice end: found by the whole warp, 16 bytes per lane
            uint32_t w_tail = 0xFFFFFFFFu;
            {
                const uint32_t lim = (uint32_t)avail + ((lo + avail == a.n_end) ? 1u : 0u);
                for (uint32_t c0 = (uint32_t)(warp + 1) * LCAP; c0 * 16 < lim && w_tail == 0xFFFFFFFFu; c0 += 32) {
                    const uint32_t chunk = c0 + lane;
                    uint32_t m = 0;
                    if (chunk * 16 < lim) {
                        const uint4 v = *reinterpret_cast<const uint4*>(buf + chunk * 16);
                        const uint32_t ws[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            const uint32_t x = ws[k] ^ 0x0A0A0A0Au;
                            uint32_t t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
                            t = ~(t | x | 0x7F7F7F7Fu);
                            m |= ((t * 0x00204081u) >> 28) << (k * 4);
                        }
                        if (chunk * 16 + 16 > lim) m &= (1u << (lim - chunk * 16)) - 1u;
                    }
                    const unsigned who = __ballot_sync(0xffffffffu, m != 0);
                    if (who) {
                        const int src = __ffs(who) - 1;
                        const uint32_t mm = __shfl_sync(0xffffffffu, m, src);
                        w_tail = (c0 + src) * 16 + (__ffs(mm) - 1);
                    }
                }
            }
            if (lane == 0) { s_wtot[warp] = total_w; s_wtail[warp] = w_tail; }
            if (tid == 0) s_next = 0;
            __syncthreads();
            uint32_t wbase = 0, total = 0, wmax = 0;
#pragma unroll
            for (int w = 0; w < NWARP; w++) { const uint32_t t = s_wtot[w]; if (w < warp) wbase += t; total += t; wmax = t > wmax ? t : wmax; }
            n_rec = total;
            if (warp == 0) {
#ifdef K1_NO_LOOKBACK    /* tuning experiment only: what the row-index look-back costs (row numbers are wrong in this build) */
                long long rb = tile * 200;
#else
                long long rb = tile_lookback(a.tile_state, tile, (unsigned long long)n_rec);
#endif
                if (lane == 0) s_rowbase = rb;
            }
            __syncthreads();
            // Parse.  Normally no slice holds more than K1_WCAP records: the tile's records then form one flat list (warp w's
            // starts follow warp w-1's) and the warps draw rounds of 32 consecutive records from a shared counter, so every
            // round but the tile's last runs with 32 busy lanes wherever the slice boundaries fall.  A tile of very short
            // records keeps the records with the warp that listed them, K1_WCAP at a time.
            const bool flat = wmax <= K1_WCAP;
            const long long row_base = s_rowbase;
            uint32_t round0 = 0; int base = 0;
            for (;;) {
                uint32_t idx, rank, tot_w, tail; const uint16_t* ws; bool has;
                if (flat) {
                    uint32_t b = 0;
                    if (lane == 0) b = atomicAdd(&s_next, 32u);
                    b = __shfl_sync(0xffffffffu, b, 0);
                    if (b >= total) break;
                    rank = b + lane; has = rank < total;
                    uint32_t w = 0, pre = 0;
#pragma unroll
                    for (int k = 0; k < NWARP - 1; k++) { const uint32_t t = s_wtot[k]; if (w == (uint32_t)k && rank >= pre + t) { pre += t; w = k + 1; } }
                    ws = s_starts + w * (LCAP + K1_WCAP + 2) + LCAP; idx = rank - pre; tot_w = s_wtot[w]; tail = s_wtail[w];
                } else {
                    if (round0 >= total_w) break;
                    if (base == 0 && round0 > 0) { __syncwarp(); list_pass(round0); }          // the next K1_WCAP starts of this slice
                    __syncwarp();
                    const int cnt = (int)((total_w - round0 > K1_WCAP) ? K1_WCAP : (total_w - round0));
                    idx = base + lane; has = (int)idx < cnt; rank = wbase + round0 + idx; ws = w_starts; tot_w = total_w - round0; tail = w_tail;
                    base += 32; if (base >= cnt) { base = 0; round0 += K1_WCAP; }
                }
                {
                    uint32_t st = 0, en = 0; bool in_smem = true; long long en_abs = 0;
                    if (has) {
                        st = ws[idx];
                        if (idx + 1 < tot_w) en = (uint32_t)ws[idx + 1] - 1u;
                        else {
                            // last record owned by its warp: its newline lies at or past the slice end
                            const uint32_t lim = (uint32_t)avail + ((lo + avail == a.n_end) ? 1u : 0u);
                            if (tail != 0xFFFFFFFFu) en = tail;
                            else if (lo + avail == a.n_end) {
                                // the data ends inside the loaded bytes: the record runs to the final newline or the end of data
                                const uint32_t av = (uint32_t)avail;
                                en = (av > st && buf[av - 1] == '\n') ? av - 1 : av;
                            } else {
                                in_smem = false;
                                long long g = lo + lim;
                                while (g < a.n_end && a.base_al[g] != '\n') g++;
                                en_abs = g;
                            }
                        }
                    }
                    RowSink o; o.a = &a; o.row = row_base + rank; o.valid = 0; o.err = 0;
                    const bool overflow = has && o.row >= a.max_rows;
                    if (overflow) c_err |= DXF_ROW_OVERFLOW;
                    const bool live = has && !overflow;
#ifdef K1_SKIP_PARSE     /* tuning experiment only (tools/build_variant.sh): cost of scan + row indexing without the record parse */
                    bool ok = live; o.valid = all_valid; if (live) reinterpret_cast<long long*>(a.cols[0])[o.row] = st + en;
#else
                    bool ok = fast_record_warp(buf, st, en, live && in_smem, a, o);
#endif
                    if (live) {
                        if (ok) c_fast++;
                        else if (in_smem) ok = g_record(buf + st, buf + en, S, o);
                        else ok = g_record(a.base_al + lo + st, a.base_al + en_abs, S, o);
                        c_rows++; if (!ok) c_null++;
                        c_err |= o.err; if (o.err & DXF_ROW_LIMITS) c_lim++;
                        if (o.valid != all_valid) {
                            for (uint32_t k = 0; k < n_leaves; k++) if (!((o.valid >> k) & 1u)) {
                                if (S->leaf_type[k] == DX_T_STR) reinterpret_cast<int32_t*>(a.cols[k])[o.row] = -1;
                                else reinterpret_cast<long long*>(a.cols[k])[o.row] = 0;
                            }
                        }
                        if (a.mask_width == 1) reinterpret_cast<uint8_t*>(a.mask)[o.row] = (uint8_t)(o.valid | (ok ? 0x80u : 0u));
                        else reinterpret_cast<uint32_t*>(a.mask)[o.row] = o.valid | (ok ? 0x80000000u : 0u);
                    }
                }
            }
        } else {
            // --- caller-provided framing: records whose first byte lies in [lo, hi) ---
            if (warp == 0) {
                // first record r with skew + offsets[r] >= target (offsets ascending, n_rows+1 entries): a 33-way search, the 32
                // lanes probing evenly spaced entries per round, so a tile costs ~5 dependent global loads instead of ~25
                auto lower_bound = [&](long long l, long long h, long long target) -> long long {
                    while (l < h) {
                        const long long step = (h - l + 32) / 33;
                        const long long idx = l + (long long)(lane + 1) * step - 1;
                        const bool ge = idx >= h || (long long)a.skew + a.offsets[idx] >= target;
                        const unsigned b = __ballot_sync(0xffffffffu, ge);
                        const int f = b ? __ffs(b) - 1 : 32;
                        const long long pf = l + (long long)(f + 1) * step - 1;          // first probe that is >= target (if f < 32)
                        const long long nl = f > 0 ? l + (long long)f * step : l;          // one past the last probe that is < target
                        long long nh = h;
                        if (f < 32 && pf < h) nh = pf;
                        l = nl; h = nh;
                    }
                    return l;
                };
                const long long r_lo = lower_bound(0, a.max_rows, lo);
                const long long r_hi = lower_bound(r_lo, a.max_rows, lo + K1_T);
                if (lane == 0) { s_r0 = r_lo; s_rowbase = r_hi; }
            }
            __syncthreads();
            rec0 = s_r0; n_rec = s_rowbase - rec0;
            for (long long base = 0; base < n_rec; base += K1_NT) {
                const long long idx = base + tid;
                const bool has = idx < n_rec;
                const long long r = rec0 + (has ? idx : 0);
                long long st_abs = 0, en_abs = 0; uint32_t st = 0, en = 0, en_f = 0; bool in_smem = false;
                if (has) {
                    st_abs = (long long)a.skew + a.offsets[r];
                    en_abs = (long long)a.skew + a.offsets[r + 1];
                    in_smem = en_abs - lo <= avail;
                    if (in_smem) {
                        st = (uint32_t)(st_abs - lo); en = (uint32_t)(en_abs - lo);
                        // a trailing newline inside the frame is whitespace; the fast path wants the bare record
                        en_f = en; if (en_f > st && buf[en_f - 1] == '\n') en_f--;
                    }
                }
                RowSink o; o.a = &a; o.row = r; o.valid = 0; o.err = 0;
                bool ok = fast_record_warp(buf, st, en_f, has && in_smem, a, o);
                if (has) {
                    if (ok) c_fast++;
                    else if (in_smem) ok = g_record(buf + st, buf + en, S, o);
                    else ok = g_record(a.base_al + st_abs, a.base_al + en_abs, S, o);
                    c_rows++; if (!ok) c_null++;
                    c_err |= o.err; if (o.err & DXF_ROW_LIMITS) c_lim++;
                    if (o.valid != all_valid) {
                        for (uint32_t k = 0; k < n_leaves; k++) if (!((o.valid >> k) & 1u)) {
                            if (S->leaf_type[k] == DX_T_STR) reinterpret_cast<int32_t*>(a.cols[k])[o.row] = -1;
                            else reinterpret_cast<long long*>(a.cols[k])[o.row] = 0;
                        }
                    }
                    if (a.mask_width == 1) reinterpret_cast<uint8_t*>(a.mask)[o.row] = (uint8_t)(o.valid | (ok ? 0x80u : 0u));
                    else reinterpret_cast<uint32_t*>(a.mask)[o.row] = o.valid | (ok ? 0x80000000u : 0u);
                }
            }
        }
        // the next tile is claimed by thread 0 on its way into the barrier that ends this tile (every thread copied s_tile into a
        // register two barriers ago): the atomic's round trip overlaps the other warps' last records and the tile costs one barrier less
        if (K1_STAGES == 1 && tid == 0) s_tile = (long long)atomicAdd(a.tile_counter, 1u);
        __syncthreads();
        if (K1_STAGES > 1) { tile = next_tile; slot ^= 1; }
        else {
            tile = s_tile;
            if (tile < a.n_tiles) stage_tile(0, tile);
        }
    }
    // per-warp reduction of the counters, one atomic each
    for (int s = 16; s > 0; s >>= 1) {
        c_rows += __shfl_xor_sync(0xffffffffu, c_rows, s);
        c_null += __shfl_xor_sync(0xffffffffu, c_null, s);
        c_fast += __shfl_xor_sync(0xffffffffu, c_fast, s);
        c_err |= __shfl_xor_sync(0xffffffffu, c_err, s);
        c_lim += __shfl_xor_sync(0xffffffffu, c_lim, s);
    }
    if (lane == 0) {
        if (c_rows) atomicAdd(&a.stats->rows, (unsigned long long)c_rows);
        if (c_null) atomicAdd(&a.stats->null_rows, (unsigned long long)c_null);
        if (c_fast) atomicAdd(&a.stats->fast_rows, (unsigned long long)c_fast);
        if (c_err) atomicOr(&a.stats->flags, (unsigned long long)c_err);
        if (c_lim) atomicAdd(&a.stats->limited_rows, (unsigned long long)c_lim);
    }
}

static inline size_t k1_smem_bytes(int tile) {
    return K1_STAGES * (tile + K1_OVL + K1_PAD) + (size_t)(K1_NT / 32) * (tile / (K1_NT / 32) / 16 + K1_WCAP + 2) * 2 + 64;
}

// dx_decode_framed.cu
int k1_framed_setup(int tile, int* occ);                  // 0 or a cudaError_t
void k1_framed_launch(int tile, unsigned grid, cudaStream_t stream, const DecodeArgs& a);
