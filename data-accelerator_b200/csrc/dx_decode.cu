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
#include "dx_decode.cuh"
#include "dx_json.h"
#include "dx_decode_host.h"

#ifndef K1_T_DEFAULT
#define K1_T_DEFAULT 32768  // tile bytes when the record length is unknown (a multiple of 4096; see k1_pick_tile)
#endif
#define K1_OVL 2048         // overlap bytes past the tile (records starting in the tile may end here)
#define K1_PAD 32
#ifndef K1_NT
#define K1_NT 256           // threads per CTA
#endif
#define K1_SC 1024          // record starts held in shared memory per round
#ifndef K1_STAGES
#define K1_STAGES 1         // shared-memory tile stages; 2 = prefetch the next tile while parsing (measured slower: 3 CTAs/SM already overlap copy and parse)
#endif
#ifndef K1_MIN_CTAS
#define K1_MIN_CTAS 4       // resident CTAs per SM the register allocation is held to (64 registers; measured 3 / 4 / 5: 3.27 / 3.03 / 3.03 ms per 25 M events)
#endif

// ------------------------------- PTX wrappers ------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// ------------------------------- shared-memory readers ---------------------------------
__device__ __forceinline__ uint32_t ld32u(const uint8_t* buf, uint32_t pos) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(buf + (pos & ~3u));
    return __funnelshift_r(w[0], w[1], (pos & 3u) << 3);
}
#include "dx_decode_fast.cuh"

__device__ __forceinline__ unsigned long long ld_state(const unsigned long long* p) {
    return *reinterpret_cast<const volatile unsigned long long*>(p);
}

// Decoupled look-back over per-tile record counts; executed by one full warp.
__device__ long long tile_lookback(unsigned long long* state, long long tile, unsigned long long my_count) {
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
__device__ __noinline__ uint32_t k1_edge_mask(uint32_t m, long long q0, long long n_end, long long skew) {
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
template <int K1_T>
__global__ void __launch_bounds__(K1_NT, K1_MIN_CTAS) k1_decode(const __grid_constant__ DecodeArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int TILE_BYTES = K1_T + K1_OVL + K1_PAD;                     // one staged tile (multiple of 16)
    uint16_t* s_starts = reinterpret_cast<uint16_t*>(smem + K1_STAGES * TILE_BYTES);   // K1_SC + 2
    __shared__ uint64_t s_bar[K1_STAGES];
    __shared__ long long s_tile;
    __shared__ long long s_rowbase;
    __shared__ uint32_t s_wtot[K1_NT / 32];
    __shared__ long long s_r0;
    __shared__ uint32_t s_tail_nl;   // first newline at or past the tile end (shared-memory position), ~0u if none loaded

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const DSchemaDev* S = a.schema;
    if (tid == 0) { for (int k = 0; k < K1_STAGES; k++) mbar_init(&s_bar[k], 1); fence_mbar_init(); }
    __syncthreads();

    const uint32_t n_leaves = (uint32_t)S->n_leaves;
    const uint32_t all_valid = (n_leaves >= 32) ? 0xFFFFFFFFu : ((1u << n_leaves) - 1u);
    uint32_t c_rows = 0, c_null = 0, c_fast = 0, c_err = 0;    // per thread (a thread sees far fewer than 2^32 records); widened when added to the statistics
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
        if (tid == 32 && lo_ + av == a.n_end) dst[av] = '\n';
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
        if (a.offsets == nullptr) {
            // --- structural scan: newline q in [lo, hi) owns the record starting at q+1 ---
            // warp w scans chunks [w*CPW, (w+1)*CPW), lane-strided so shared loads are conflict-free
            constexpr int CPW = (K1_T / 16) / (K1_NT / 32);   // 256
            constexpr int ROUNDS = CPW / 32;                  // 8
            const bool edge_tile = lo < (long long)a.skew || lo + (long long)K1_T + 16 >= a.n_end;
            // A tile with more than K1_SC records (records shorter than ~40 bytes) is parsed in rounds of K1_SC; every round
            // repeats the scan instead of keeping its masks and ranks, so those registers are free while records are parsed.
            for (uint32_t round0 = 0; ; round0 += K1_SC) {
            if (round0 > 0) __syncthreads();      // the previous round's readers of s_starts / s_wtot are done
            uint32_t cnt_total = 0;
            uint32_t excl_in_warp[ROUNDS];
            uint32_t my_masks[(ROUNDS + 1) / 2];
#pragma unroll
            for (int r = 0; r < ROUNDS; r++) {
                const int chunk = warp * CPW + r * 32 + lane;
                const uint4 v = *reinterpret_cast<const uint4*>(buf + chunk * 16);
                uint32_t m = 0;
                const uint32_t ws[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t x = ws[k] ^ 0x0A0A0A0Au;                 // zero byte where '\n'
                    uint32_t t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;           // exact per-byte zero test
                    t = ~(t | x | 0x7F7F7F7Fu);                             // 0x80 in each byte that was zero
                    m |= ((t * 0x00204081u) >> 28) << (k * 4);              // gather the four flags
                }
                // validity of each candidate q = lo + chunk*16 + i : start q+1 must be data (< n_end), q >= skew
                if (edge_tile) m = k1_edge_mask(m, lo + (long long)chunk * 16, a.n_end, (long long)a.skew);
                if (r & 1) my_masks[r >> 1] |= m << 16; else my_masks[r >> 1] = m;
                excl_in_warp[r] = __popc(m);          // this lane's count for now; turned into a rank below
            }
            {
                // one prefix sum for three rounds at a time: a lane's count is <= 16 and a round's total <= 512, so three 10-bit
                // fields share a word and the rounds' shuffle chains run side by side instead of one after the other
                constexpr int NW = (ROUNDS + 2) / 3;
                uint32_t pk[NW];
#pragma unroll
                for (int w = 0; w < NW; w++) {
                    pk[w] = excl_in_warp[3 * w];
                    if (3 * w + 1 < ROUNDS) pk[w] |= excl_in_warp[3 * w + 1] << 10;
                    if (3 * w + 2 < ROUNDS) pk[w] |= excl_in_warp[3 * w + 2] << 20;
                }
#pragma unroll
                for (int s = 1; s < 32; s <<= 1) {
#pragma unroll
                    for (int w = 0; w < NW; w++) { const uint32_t t = __shfl_up_sync(0xffffffffu, pk[w], s); if (lane >= s) pk[w] += t; }
                }
#pragma unroll
                for (int r = 0; r < ROUNDS; r++) {
                    const uint32_t incl = (pk[r / 3] >> (10 * (r % 3))) & 1023u;
                    const uint32_t tot = (__shfl_sync(0xffffffffu, pk[r / 3], 31) >> (10 * (r % 3))) & 1023u;
                    excl_in_warp[r] = cnt_total + incl - excl_in_warp[r];
                    cnt_total += tot;
                }
            }
            if (lane == 0) s_wtot[warp] = cnt_total;
            if (warp == K1_NT / 32 - 1 && round0 == 0) {
                // the tile's last record ends at the first newline at or past the tile end: find it with the
                // whole warp (16 bytes per lane per round) instead of one lane walking bytes
                const uint32_t lim = (uint32_t)avail + ((lo + avail == a.n_end) ? 1u : 0u);
                uint32_t found = 0xFFFFFFFFu;
                for (uint32_t c0 = K1_T / 16; c0 * 16 < lim && found == 0xFFFFFFFFu; c0 += 32) {
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
                        found = (c0 + src) * 16 + (__ffs(mm) - 1);
                    }
                }
                if (lane == 0) s_tail_nl = found;
            }
            __syncthreads();
            uint32_t wbase = first_virtual ? 1u : 0u, total = first_virtual ? 1u : 0u;
#pragma unroll
            for (int w = 0; w < K1_NT / 32; w++) { uint32_t t = s_wtot[w]; if (w < warp) wbase += t; total += t; }
            n_rec = total;
            if (warp == 0 && round0 == 0) {
#ifdef K1_NO_LOOKBACK    /* tuning experiment only: what the row-index look-back costs (row numbers are wrong in this build) */
                long long rb = tile * 200;
#else
                long long rb = tile_lookback(a.tile_state, tile, (unsigned long long)n_rec);
#endif
                if (lane == 0) s_rowbase = rb;
            }
            {
                if (first_virtual && tid == 0 && round0 == 0) s_starts[0] = (uint16_t)a.skew;
#pragma unroll
                for (int r = 0; r < ROUNDS; r++) {
                    uint32_t m = (r & 1) ? (my_masks[r >> 1] >> 16) : (my_masks[r >> 1] & 0xffffu);
                    uint32_t rel = wbase + excl_in_warp[r] - round0;     // wraps to a huge value for earlier rounds' records
                    const uint32_t pos0 = (uint32_t)(warp * CPW + r * 32 + lane) * 16 + 1;
                    while (m) {
                        const int i = __ffs(m) - 1; m &= m - 1;
                        if (rel <= K1_SC) s_starts[rel] = (uint16_t)(pos0 + i);
                        rel++;
                    }
                }
                __syncthreads();
                const long long row_base = s_rowbase;
                const int cnt = (int)((n_rec - round0 > K1_SC) ? K1_SC : (n_rec - round0));
                for (int base = 0; base < cnt; base += K1_NT) {
                    const int idx = base + tid;
                    const bool has = idx < cnt;
                    uint32_t st = 0, en = 0; bool in_smem = true; long long en_abs = 0;
                    const long long rank = round0 + idx;
                    if (has) {
                        st = s_starts[idx];
                        if (rank + 1 < n_rec) en = (uint32_t)s_starts[idx + 1] - 1u;
                        else {
                            // last record owned by this tile: its newline lies at or past the tile end
                            const uint32_t lim = (uint32_t)avail + ((lo + avail == a.n_end) ? 1u : 0u);
                            const uint32_t p = s_tail_nl;
                            if (p != 0xFFFFFFFFu) en = p;
                            else if (lo + avail == a.n_end) {
                                // the data ends inside the tile proper: the record runs to the final newline or the end of data
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
                        c_err |= o.err;
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
                if ((long long)round0 + K1_SC >= n_rec) break;
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
                    c_err |= o.err;
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
        __syncthreads();
        if (K1_STAGES > 1) { tile = next_tile; slot ^= 1; }
        else {
            if (tid == 0) s_tile = (long long)atomicAdd(a.tile_counter, 1u);
            __syncthreads();
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
    }
    if (lane == 0) {
        if (c_rows) atomicAdd(&a.stats->rows, (unsigned long long)c_rows);
        if (c_null) atomicAdd(&a.stats->null_rows, (unsigned long long)c_null);
        if (c_fast) atomicAdd(&a.stats->fast_rows, (unsigned long long)c_fast);
        if (c_err) atomicOr(&a.stats->flags, (unsigned long long)c_err);
    }
}

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
static size_t k1_smem_bytes(int tile) {
    return K1_STAGES * (tile + K1_OVL + K1_PAD) + (K1_SC + 2) * 2 + 64;
}
static const int k1_tiles[] = {16384, 20480, 24576, 28672, 32768, 36864, 40960, 45056, 49152};
// tile bytes for a batch: the largest instantiated size that still keeps the expected records per tile under ~246
static int k1_pick_tile(long long n_bytes, long long n_events) {
    if (n_events <= 0 || n_bytes <= 0) return K1_T_DEFAULT;
    const double avg = (double)n_bytes / (double)n_events;
    int best = k1_tiles[0];
    for (int t : k1_tiles) if ((double)t / avg <= 246.0) best = t;
    return best;
}
template <int T> static int k1_setup_one(dx_ctx* c, int* occ_out) {
    DX_CUDA(c, cudaFuncSetAttribute(k1_decode<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k1_smem_bytes(T)));
    int occ = 0;
    DX_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k1_decode<T>, K1_NT, k1_smem_bytes(T)));
    *occ_out = occ > 0 ? occ : 1;
    return DX_OK;
}
static int k1_launch(dx_ctx* c, int tile, const int* occ_by_tile, const DecodeArgs& a) {
    int idx = 0; for (int i = 0; i < 9; i++) if (k1_tiles[i] == tile) idx = i;
    long long grid = (long long)c->sm_count * occ_by_tile[idx];
    if (grid > a.n_tiles) grid = a.n_tiles;
    const size_t smem = k1_smem_bytes(tile);
    switch (tile) {
    case 16384: k1_decode<16384><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case 20480: k1_decode<20480><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case 24576: k1_decode<24576><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case 28672: k1_decode<28672><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case 36864: k1_decode<36864><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case 40960: k1_decode<40960><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case 45056: k1_decode<45056><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    case 49152: k1_decode<49152><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
    default: k1_decode<32768><<<(unsigned)grid, K1_NT, smem, c->stream>>>(a); break;
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
        hs.segs[i].lit_off = (uint16_t)lit_used; hs.segs[i].lit_len = (uint16_t)L;
        hs.segs[i].type = (int16_t)b.seg_types[i]; hs.segs[i].leaf = (int16_t)b.seg_leaves[i];
        hs.segs[i].tail_mask = (L & 3) ? ((1u << ((L & 3) << 3)) - 1u) : 0u;
        lit_used += Lp; hs.n_segs++;
    }
    if (!tmpl_ok) { hs.n_segs = 1; hs.segs[0].lit_off = 0; hs.segs[0].lit_len = 4; hs.segs[0].type = 0; hs.segs[0].tail_mask = 0; memcpy(hs.lits, "\x01\x01\x01\x01", 4); }
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
      if ((rc2 = k1_setup_one<16384>(c, &occ_by_tile[0])) || (rc2 = k1_setup_one<20480>(c, &occ_by_tile[1])) || (rc2 = k1_setup_one<24576>(c, &occ_by_tile[2])) ||
          (rc2 = k1_setup_one<28672>(c, &occ_by_tile[3])) || (rc2 = k1_setup_one<32768>(c, &occ_by_tile[4])) || (rc2 = k1_setup_one<36864>(c, &occ_by_tile[5])) ||
          (rc2 = k1_setup_one<40960>(c, &occ_by_tile[6])) || (rc2 = k1_setup_one<45056>(c, &occ_by_tile[7])) || (rc2 = k1_setup_one<49152>(c, &occ_by_tile[8]))) return rc2; }
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
        last_stats = DecodeStats{(unsigned long long)n_events, (unsigned long long)n_events, 0, 0};
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
    if (last_stats.flags) {
        char b[256];
        snprintf(b, sizeof b, "decode failed: flags=0x%llx (1 dictionary overflow, 2 number beyond 19-digit window, "
                 "4 object/array into string field, 8 escaped string > %d bytes, 16 more records than n_events)",
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
#define DX_NOALLOC_HEADER_WORDS 8                                    /* [0] tile counter, [2..5] DecodeStats */
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
        DX_CUDA(c, cudaSetDevice(c->device));
        void* p = nullptr;
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
    unsigned long long* scratch = nullptr;
    int rc = noalloc_scratch_address(c, &scratch);
    if (rc) return rc;
    DX_CUDA(c, cudaSetDevice(c->device));
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
    if (st.flags) {
        char b[200];
        snprintf(b, sizeof b, "decode_noalloc failed: flags=0x%llx (1 dictionary overflow, 2 number beyond 19-digit window, "
                 "4 object/array into string field, 8 long escaped string, 16 more records than n_events)", st.flags);
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
