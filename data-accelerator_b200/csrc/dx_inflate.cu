// dx_inflate.cu — K7: gzip members (the blob input's files) inflated on the device (the step before the path).
//
// Reference: the blob input reads gzip-compressed JSON-lines files through Hadoop's codec before processJson sees a line
// (datax/input/BlobPointerInput.scala, datax/processor/BlobPointerProcessor.scala; CPF:463-515 hands the lines to project()).
// Here the compressed files of a micro-batch go to the device as they are and come out as one NDJSON buffer for K1.
//
// One WARP per gzip member.  Every lane runs the same bit reader and Huffman decode (uniform control flow: the loads are
// broadcasts, nothing diverges), so a literal costs what it costs one thread, and a match — most of the bytes of repetitive
// JSON — is copied by all 32 lanes (`out[pos + i] = out[pos - dist + i mod dist]`, which is also right when the match overlaps
// its own output).  Code tables live in shared memory per warp: a 10-bit direct table for literal/length codes and an 8-bit one
// for distance codes, longer codes through the canonical count/symbol walk.  Output offsets come from the members' ISIZE
// trailers (RFC 1952), so members inflate independently; the decoded length is checked against ISIZE.  The CRC-32 of the
// trailer is not verified on the device (documented in DESIGN.md).
#include "dx_common.h"
#include "dx_inflate.h"
#include <cub/cub.cuh>

#define INF_WARPS 4
#define INF_FAST_L 10
#define INF_FAST_D 8
#define INF_ERR_HEADER 1u
#define INF_ERR_DATA 2u
#define INF_ERR_SIZE 4u

struct InfTables {
    uint16_t fast_l[1 << INF_FAST_L];      // (symbol << 4) | length; 0 = longer than the table's bits (or unused)
    uint16_t fast_d[1 << INF_FAST_D];
    uint16_t sym_l[288], sym_d[32];        // symbols ordered by (code length, symbol)
    uint16_t cnt_l[16], cnt_d[16];
    uint8_t lens[320];
};

static __constant__ uint16_t inf_len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
static __constant__ uint8_t inf_len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
static __constant__ uint16_t inf_dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
static __constant__ uint8_t inf_dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
static __constant__ uint8_t inf_cl_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

// LSB-first bit reader; every lane holds the same state (the loads are broadcasts).  Only real bytes enter the buffer: past the
// end the missing bits read as zero and cnt goes negative, which the caller checks.
struct InfBits {
    const uint8_t* p; const uint8_t* end;
    unsigned long long buf; int cnt;
    __device__ __forceinline__ void refill() { while (cnt <= 56 && p < end) { buf |= (unsigned long long)(*p++) << cnt; cnt += 8; } }
    __device__ __forceinline__ uint32_t peek(int n) const { return (uint32_t)(buf & ((1ull << n) - 1ull)); }
    __device__ __forceinline__ void drop(int n) { buf >>= n; cnt -= n; }
    __device__ __forceinline__ uint32_t take(int n) { if (cnt < n) refill(); const uint32_t v = peek(n); drop(n); return v; }
};

// Builds the canonical code of `n` symbols with lengths `lens` (0 = unused): count / symbol arrays and the direct table.
// Executed by every lane (same inputs, same stores: the warp writes each entry once per lane with the same value); lane-strided
// where it pays.  Returns false for an over-subscribed set of lengths.
static __device__ bool inf_build(const uint8_t* lens, int n, uint16_t* cnt, uint16_t* sym, uint16_t* fast, int fast_bits, int lane) {
    if (lane == 0) {
        for (int l = 0; l < 16; l++) cnt[l] = 0;
        for (int s = 0; s < n; s++) cnt[lens[s]]++;
    }
    __syncwarp();
    int left = 1;
    for (int l = 1; l < 16; l++) { left <<= 1; left -= cnt[l]; if (left < 0) return false; }
    uint16_t offs[16]; offs[1] = 0;
    for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + cnt[l];
    if (lane == 0) for (int s = 0; s < n; s++) if (lens[s]) sym[offs[lens[s]]++] = (uint16_t)s;
    for (int i = lane; i < (1 << fast_bits); i += 32) fast[i] = 0;
    __syncwarp();
    // direct table: canonical codes are assigned in (length, symbol) order; the bit reader is LSB-first, so a code's bits are reversed
    if (lane == 0) {
        unsigned code = 0; int idx = 0;
        for (int l = 1; l <= fast_bits; l++) {
            for (int k = 0; k < cnt[l]; k++, idx++, code++) {
                const unsigned rev = __brev(code) >> (32 - l);
                for (unsigned e = rev; e < (1u << fast_bits); e += 1u << l) fast[e] = (uint16_t)((sym[idx] << 4) | l);
            }
            code <<= 1;
        }
    }
    __syncwarp();
    return true;
}

// one symbol: direct table, else the canonical walk (RFC 1951 3.2.2) over the longer lengths
static __device__ __forceinline__ int inf_decode(InfBits& b, const uint16_t* cnt, const uint16_t* sym, const uint16_t* fast, int fast_bits) {
    if (b.cnt < 15) b.refill();
    const uint16_t e = fast[b.peek(fast_bits)];
    if (e) { b.drop(e & 15); return e >> 4; }
    int code = 0, first = 0, index = 0;
    unsigned long long bits = b.buf;
    for (int l = 1; l <= 15; l++) {
        code |= (int)(bits & 1ull); bits >>= 1;
        const int c = cnt[l];
        if (code - c < first) { b.drop(l); return sym[index + (code - first)]; }
        index += c; first += c; first <<= 1; code <<= 1;
    }
    return -1;
}

__global__ void k7_gzip_sizes(const uint8_t* __restrict__ gz, const long long* __restrict__ off, int n, unsigned int* __restrict__ sizes, unsigned int* __restrict__ err) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m > n) return;
    if (m == n) { sizes[m] = 0; return; }
    const long long a = off[m], e = off[m + 1];
    if (e - a < 18 || gz[a] != 0x1f || gz[a + 1] != 0x8b || gz[a + 2] != 8) { sizes[m] = 0; atomicOr(err, INF_ERR_HEADER); return; }
    sizes[m] = (unsigned)gz[e - 4] | ((unsigned)gz[e - 3] << 8) | ((unsigned)gz[e - 2] << 16) | ((unsigned)gz[e - 1] << 24);
}

__global__ void __launch_bounds__(INF_WARPS * 32) k7_inflate(const uint8_t* __restrict__ gz, const long long* __restrict__ off, int n,
                                                             const long long* __restrict__ out_off, uint8_t* __restrict__ out, unsigned int* __restrict__ err) {
    __shared__ InfTables s_tab[INF_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int m = blockIdx.x * INF_WARPS + warp;
    if (m >= n) return;
    InfTables& T = s_tab[warp];
    const uint8_t* g = gz + off[m];
    const long long glen = off[m + 1] - off[m] - 8;                                  // the trailer is not deflate data
    // ---- gzip header (RFC 1952): fixed 10 bytes, then FEXTRA / FNAME / FCOMMENT / FHCRC as flagged
    long long hp = 10;
    const unsigned flg = g[3];
    if (flg & 4) { if (hp + 2 > glen) { if (!lane) atomicOr(err, INF_ERR_HEADER); return; } hp += 2 + (g[hp] | (g[hp + 1] << 8)); }
    if (flg & 8) { while (hp < glen && g[hp]) hp++; hp++; }
    if (flg & 16) { while (hp < glen && g[hp]) hp++; hp++; }
    if (flg & 2) hp += 2;
    if (hp > glen) { if (!lane) atomicOr(err, INF_ERR_HEADER); return; }
    InfBits b; b.p = g + hp; b.end = g + glen; b.buf = 0; b.cnt = 0;
    uint8_t* o = out + out_off[m];
    const long long want = out_off[m + 1] - out_off[m];
    long long pos = 0;
    bool bad = false;
    for (bool last = false; !last && !bad;) {
        last = b.take(1) != 0;
        const uint32_t type = b.take(2);
        if (type == 0) {                                    // stored block: byte-aligned LEN, NLEN, bytes
            b.drop(b.cnt & 7);
            if (b.cnt < 32) b.refill();
            const uint32_t len = b.take(16), nlen = b.take(16);
            if (b.cnt < 0 || (len ^ 0xffffu) != nlen) { bad = true; break; }
            const uint8_t* src = b.p - (b.cnt >> 3);        // the bytes still in the bit buffer come first
            if (src + len > b.end || pos + (long long)len > want) { bad = true; break; }
            for (uint32_t i = lane; i < len; i += 32) o[pos + i] = src[i];
            pos += len;
            b.p = src + len; b.buf = 0; b.cnt = 0;
            __syncwarp();
            continue;
        }
        if (type == 3) { bad = true; break; }
        if (type == 1) {                                    // fixed codes
            for (int s = lane; s < 288; s += 32) T.lens[s] = s < 144 ? 8 : (s < 256 ? 9 : (s < 280 ? 7 : 8));
            __syncwarp();
            inf_build(T.lens, 288, T.cnt_l, T.sym_l, T.fast_l, INF_FAST_L, lane);
            for (int s = lane; s < 30; s += 32) T.lens[s] = 5;
            __syncwarp();
            inf_build(T.lens, 30, T.cnt_d, T.sym_d, T.fast_d, INF_FAST_D, lane);
        } else {                                            // dynamic codes
            const int hlit = (int)b.take(5) + 257, hdist = (int)b.take(5) + 1, hclen = (int)b.take(4) + 4;
            if (hlit > 286 || hdist > 30) { bad = true; break; }
            uint8_t cl[19];
            for (int i = 0; i < 19; i++) cl[i] = 0;
            for (int i = 0; i < hclen; i++) cl[inf_cl_order[i]] = (uint8_t)b.take(3);
            __syncwarp();
            if (lane == 0) for (int i = 0; i < 19; i++) T.lens[i] = cl[i];
            __syncwarp();
            // the code-length code shares the distance tables' storage (rebuilt below)
            if (!inf_build(T.lens, 19, T.cnt_d, T.sym_d, T.fast_d, 7, lane)) { bad = true; break; }
            int idx = 0;
            while (idx < hlit + hdist) {
                const int s = inf_decode(b, T.cnt_d, T.sym_d, T.fast_d, 7);
                if (s < 0) { bad = true; break; }
                if (s < 16) { if (lane == 0) T.lens[idx] = (uint8_t)s; idx++; continue; }
                int rep, val = 0;
                if (s == 16) { if (idx == 0) { bad = true; break; } __syncwarp(); val = T.lens[idx - 1]; rep = 3 + (int)b.take(2); }
                else if (s == 17) rep = 3 + (int)b.take(3);
                else rep = 11 + (int)b.take(7);
                if (idx + rep > hlit + hdist) { bad = true; break; }
                __syncwarp();
                if (lane == 0) for (int k = 0; k < rep; k++) T.lens[idx + k] = (uint8_t)val;
                idx += rep;
                __syncwarp();
            }
            if (bad) break;
            __syncwarp();
            if (T.lens[256] == 0) { bad = true; break; }
            // distance lengths follow the literal/length ones: move them out before the first table is built over lens
            uint8_t dl[30];
            for (int i = 0; i < 30; i++) dl[i] = i < hdist ? T.lens[hlit + i] : 0;
            __syncwarp();
            if (lane == 0) for (int i = hlit; i < 288; i++) T.lens[i] = 0;
            __syncwarp();
            if (!inf_build(T.lens, 288, T.cnt_l, T.sym_l, T.fast_l, INF_FAST_L, lane)) { bad = true; break; }
            if (lane == 0) for (int i = 0; i < 30; i++) T.lens[i] = dl[i];
            __syncwarp();
            if (!inf_build(T.lens, 30, T.cnt_d, T.sym_d, T.fast_d, INF_FAST_D, lane)) { bad = true; break; }
        }
        // ---- the block's symbols (the window is the output itself; a shared-memory window was measured: the same 27 MB/s per warp —
        // a warp is bound by the dependent decode chain, ~480 cycles per symbol — and a third of the resident warps)
        for (;;) {
            if (b.cnt < 0) { bad = true; break; }                 // ran past the end of the member's data
            const int s = inf_decode(b, T.cnt_l, T.sym_l, T.fast_l, INF_FAST_L);
            if (s < 0) { bad = true; break; }
            if (s < 256) {
                if (pos >= want) { bad = true; break; }
                if (lane == 0) o[pos] = (uint8_t)s;
                pos++;
                continue;
            }
            if (s == 256) break;
            const int li = s - 257;
            if (li >= 29) { bad = true; break; }
            const int len = inf_len_base[li] + (int)b.take(inf_len_extra[li]);
            const int ds = inf_decode(b, T.cnt_d, T.sym_d, T.fast_d, INF_FAST_D);
            if (ds < 0 || ds >= 30) { bad = true; break; }
            const int dist = inf_dist_base[ds] + (int)b.take(inf_dist_extra[ds]);
            if ((long long)dist > pos || pos + len > want) { bad = true; break; }
            __syncwarp();                                  // the bytes the match reads were written by other lanes
            const uint8_t* src = o + pos - dist;
            if (dist >= len) { for (int i = lane; i < len; i += 32) o[pos + i] = src[i]; }
            else { for (int i = lane; i < len; i += 32) o[pos + i] = src[i % dist]; }
            pos += len;
            __syncwarp();
        }
    }
    if (bad || b.cnt < 0) { if (!lane) atomicOr(err, INF_ERR_DATA); return; }
    if (pos != want && !lane) atomicOr(err, INF_ERR_SIZE);
}

struct InfToI64 { __host__ __device__ long long operator()(unsigned int v) const { return (long long)v; } };

int dx_gunzip(dx_ctx* ctx, const uint8_t* gz, const int64_t* member_offsets, int n_members, int memory_kind, DevBufP* out, int64_t* n_bytes) {
    if (n_members < 0 || (n_members > 0 && (!gz || !member_offsets)) || !out || !n_bytes) return ctx->set_error(DX_ERR_ARG, "dx_gunzip: bad arguments");
    *n_bytes = 0;
    if (n_members == 0) { *out = make_buf(ctx, 16); return DX_OK; }
    const int64_t total_in = member_offsets[n_members];
    for (int m = 0; m < n_members; m++) if (member_offsets[m + 1] < member_offsets[m]) return ctx->set_error(DX_ERR_ARG, "dx_gunzip: member offsets are not ascending");
    DevBufP d_gz;
    const uint8_t* gz_dev = gz;
    if (memory_kind == 0) {
        d_gz = make_buf(ctx, (size_t)total_in + 16);
        DX_CUDA(ctx, cudaMemcpyAsync(d_gz->p, gz, (size_t)total_in, cudaMemcpyHostToDevice, ctx->stream));
        gz_dev = d_gz->as<uint8_t>();
    }
    DevBuf d_off(ctx, (size_t)(n_members + 1) * 8), d_sizes(ctx, (size_t)(n_members + 1) * 4), d_ooff(ctx, (size_t)(n_members + 1) * 8), d_err(ctx, 16);
    DX_CUDA(ctx, cudaMemcpyAsync(d_off.p, member_offsets, (size_t)(n_members + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    DX_CUDA(ctx, cudaMemsetAsync(d_err.p, 0, 16, ctx->stream));
    k7_gzip_sizes<<<(n_members + 1 + 127) / 128, 128, 0, ctx->stream>>>(gz_dev, d_off.as<long long>(), n_members, d_sizes.as<unsigned int>(), d_err.as<unsigned int>());
    {
        cub::TransformInputIterator<long long, InfToI64, const unsigned int*> in(d_sizes.as<unsigned int>(), InfToI64());
        size_t tmp_bytes = 0;
        DX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, in, d_ooff.as<long long>(), n_members + 1, ctx->stream));
        DevBuf tmp(ctx, tmp_bytes + 16);
        DX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, in, d_ooff.as<long long>(), n_members + 1, ctx->stream));
        DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    long long total = 0; unsigned int e = 0;
    DX_CUDA(ctx, cudaMemcpyAsync(&total, d_ooff.as<long long>() + n_members, 8, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaMemcpyAsync(&e, d_err.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (e) return ctx->set_error(DX_ERR_PARSE, "dx_gunzip: a member does not start with a gzip header (1f 8b 08) or is shorter than 18 bytes");
    DevBufP o = make_buf(ctx, (size_t)total + 64);
    if (!o->p) return ctx->set_error(DX_ERR_CUDA, "dx_gunzip: output allocation failed");
    {
        StageTimer tm(ctx, "k7_inflate");
        k7_inflate<<<(n_members + INF_WARPS - 1) / INF_WARPS, INF_WARPS * 32, 0, ctx->stream>>>(gz_dev, d_off.as<long long>(), n_members, d_ooff.as<long long>(), o->as<uint8_t>(), d_err.as<unsigned int>());
        ctx->launches["k7_inflate"]++;
    }
    DX_CUDA(ctx, cudaGetLastError());
    DX_CUDA(ctx, cudaMemcpyAsync(&e, d_err.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (e & INF_ERR_HEADER) return ctx->set_error(DX_ERR_PARSE, "dx_gunzip: damaged gzip header");
    if (e & INF_ERR_DATA) return ctx->set_error(DX_ERR_PARSE, "dx_gunzip: damaged deflate data");
    if (e & INF_ERR_SIZE) return ctx->set_error(DX_ERR_PARSE, "dx_gunzip: a member's decoded length differs from its ISIZE trailer");
    *out = o; *n_bytes = total;
    return DX_OK;
}
