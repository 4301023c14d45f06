// dx_gen.cu — synthetic IoT event generator on the device (measurement tooling, not part of the
// flow path).  Event i is a pure function of (seed, i); the byte-identical CPU twin lives in
// oracle/dxo_core.c (dxo_gen_events).  Shapes follow the reference's simulated home-automation
// events: DataGenerator.scala:27-209 driven by the schema in HomeAutomationLocal.json:288, and
// SURVEY.md §8(d) (S1 = 5 fields, S2 = S1 + temperature, humidity).
#include "dx_common.h"
#include "dx_abi_types.h"
#include <cub/device/device_scan.cuh>

static __device__ __forceinline__ uint64_t gmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static __device__ __forceinline__ uint64_t ghk(const dx_gen_params& g, uint64_t i, uint64_t k) { return gmix64(g.seed ^ gmix64(i * 16 + k)); }

struct GenEvent { long long deviceId, homeId, eventTime, status; int type, temp_c, hum_c; };

static __device__ __forceinline__ void gen_fields(const dx_gen_params& g, uint64_t i, GenEvent& ev) {
    ev.deviceId = 1 + (long long)(ghk(g, i, 0) % (uint64_t)g.n_keys);
    ev.homeId = 1 + (long long)(gmix64((uint64_t)ev.deviceId * 0x9E3779B97F4A7C15ull) % (uint64_t)g.n_homes);
    ev.type = (int)(ghk(g, i, 1) & 3);
    ev.status = (long long)(ghk(g, i, 2) & 1);
    long long jitter = (long long)(ghk(g, i, 3) % (uint64_t)g.interval_ms);
    int cls = (int)(ghk(g, i, 4) % 1000);
    if (cls < g.late_permille) ev.eventTime = g.base_ms - g.watermark_ms - g.interval_ms - 1 - jitter;
    else if (cls >= 1000 - g.future_permille) ev.eventTime = g.base_ms + 1 + jitter;
    else ev.eventTime = g.base_ms - 1 - jitter;
    ev.temp_c = 510 + (int)(ghk(g, i, 5) % 9500);
    ev.hum_c = (int)(ghk(g, i, 6) % 10000);
}

static __device__ __forceinline__ int ndigits(unsigned long long v) { int n = 1; while (v >= 10) { v /= 10; n++; } return n; }
static __device__ __forceinline__ int len_i64(long long v) { return v < 0 ? 1 + ndigits(0ull - (unsigned long long)v) : ndigits((unsigned long long)v); }

static __device__ const char gk_types[4][16] = {"DoorLock", "WindowLock", "Heating", "GarageDoorLock"};
static __device__ const int gk_type_len[4] = {8, 10, 7, 14};

static __device__ __forceinline__ int put_u64(uint8_t* d, unsigned long long v) {
    int n = ndigits(v);
    for (int k = n - 1; k >= 0; k--) { d[k] = (uint8_t)('0' + v % 10); v /= 10; }
    return n;
}
static __device__ __forceinline__ int put_i64(uint8_t* d, long long v) {
    if (v < 0) { d[0] = '-'; return 1 + put_u64(d + 1, 0ull - (unsigned long long)v); }
    return put_u64(d, (unsigned long long)v);
}
static __device__ __forceinline__ int put_lit(uint8_t* d, const char* s, int n) { for (int k = 0; k < n; k++) d[k] = (uint8_t)s[k]; return n; }
#define PUT(s) n += put_lit(d + n, s, (int)sizeof(s) - 1)

__global__ void gen_lengths(dx_gen_params g, long long first, long long n, int* lens) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    GenEvent ev; gen_fields(g, (uint64_t)(first + i), ev);
    constexpr int LIT1 = sizeof("{\"deviceDetails\":{\"deviceId\":") - 1 + sizeof(",\"deviceType\":\"") - 1 +
                         sizeof("\",\"eventTime\":") - 1 + sizeof(",\"homeId\":") - 1 + sizeof(",\"status\":") - 1 +
                         sizeof("}}\n") - 1;
    constexpr int LIT2 = sizeof(",\"temperature\":") - 1 + sizeof(",\"humidity\":") - 1;
    int L = LIT1 + len_i64(ev.deviceId) + gk_type_len[ev.type] + len_i64(ev.eventTime) + len_i64(ev.homeId) + 1;
    if (g.shape == 2) L += LIT2 + ndigits((unsigned)(ev.temp_c / 100)) + 3 + ndigits((unsigned)(ev.hum_c / 100)) + 3;
    lens[i] = L;
}

__global__ void gen_write(dx_gen_params g, long long first, long long count, const long long* offs, uint8_t* out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    GenEvent ev; gen_fields(g, (uint64_t)(first + i), ev);
    uint8_t* d = out + offs[i];
    int n = 0;
    PUT("{\"deviceDetails\":{\"deviceId\":");
    n += put_i64(d + n, ev.deviceId);
    PUT(",\"deviceType\":\"");
    n += put_lit(d + n, gk_types[ev.type], gk_type_len[ev.type]);
    PUT("\",\"eventTime\":");
    n += put_i64(d + n, ev.eventTime);
    PUT(",\"homeId\":");
    n += put_i64(d + n, ev.homeId);
    PUT(",\"status\":");
    n += put_i64(d + n, ev.status);
    if (g.shape == 2) {
        PUT(",\"temperature\":");
        n += put_u64(d + n, (unsigned)(ev.temp_c / 100)); d[n++] = '.';
        d[n++] = (uint8_t)('0' + (ev.temp_c % 100) / 10); d[n++] = (uint8_t)('0' + ev.temp_c % 10);
        PUT(",\"humidity\":");
        n += put_u64(d + n, (unsigned)(ev.hum_c / 100)); d[n++] = '.';
        d[n++] = (uint8_t)('0' + (ev.hum_c % 100) / 10); d[n++] = (uint8_t)('0' + ev.hum_c % 10);
    }
    PUT("}}\n");
}

__global__ void gen_widen(const int* lens, long long n, long long* out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = lens[i];
}

// Generates events [first, first+n) as NDJSON into a pooled device buffer.  offsets_out (optional)
// receives a device array of n+1 record offsets.  Caller releases both with dx_dev_free.
int dx_gen_device(dx_ctx* c, const dx_gen_params* g, int64_t first, int64_t n, uint8_t** bytes_out,
                  int64_t* n_bytes_out, int64_t** offsets_out) {
    if (n <= 0 || g->n_keys <= 0 || g->n_homes <= 0 || g->interval_ms <= 0) return c->set_error(DX_ERR_ARG, "bad generator parameters");
    DevBuf lens(c, (size_t)n * 4), wide(c, (size_t)(n + 1) * 8);
    int64_t* offs = reinterpret_cast<int64_t*>(c->pool.alloc((size_t)(n + 1) * 8));
    if (!lens.p || !wide.p || !offs) return c->set_error(DX_ERR_CUDA, "generator allocation failed");
    unsigned blocks = (unsigned)((n + 255) / 256);
    gen_lengths<<<blocks, 256, 0, c->stream>>>(*g, first, n, lens.as<int>());
    DX_CUDA(c, cudaMemsetAsync(wide.p, 0, (size_t)(n + 1) * 8, c->stream));
    gen_widen<<<blocks, 256, 0, c->stream>>>(lens.as<int>(), n, wide.as<long long>());
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, wide.as<long long>(), reinterpret_cast<long long*>(offs), n + 1, c->stream);
    DevBuf tmp(c, tmp_bytes + 16);
    DX_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, wide.as<long long>(), reinterpret_cast<long long*>(offs), n + 1, c->stream));
    long long total = 0;
    DX_CUDA(c, cudaMemcpyAsync(&total, offs + n, 8, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    uint8_t* out = reinterpret_cast<uint8_t*>(c->pool.alloc((size_t)total + 64));
    if (!out) { c->pool.release(offs); return c->set_error(DX_ERR_CUDA, "generator allocation failed"); }
    gen_write<<<blocks, 256, 0, c->stream>>>(*g, first, n, reinterpret_cast<const long long*>(offs), out);
    DX_CUDA(c, cudaGetLastError());
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    *bytes_out = out; *n_bytes_out = total;
    if (offsets_out) *offsets_out = offs; else c->pool.release(offs);
    return DX_OK;
}
