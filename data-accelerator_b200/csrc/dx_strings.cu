// dx_strings.cu — string-valued expressions over dictionary columns.
//
// Covers the string work the reference's flows ask of Spark SQL (executed by spark.sql at
// CommonProcessorFactory.scala:271 / selectExpr at :100):
//   * CONCAT(...) and CAST(<long> AS STRING)            -> k_strbuild: bytes assembled per row, entered
//     (e.g. CGenIoTSample.txt:223-229, HomeAutomationLocal.json:326)   into a fresh device dictionary
//   * stringToTimestamp(col) (ConcurrentDateFormat.scala:38-62), LIKE, user UDFs over a string column
//     (ExtendedUDFHandler.scala:39-103, JarUDFHandler.scala:42-62) -> evaluated ONCE PER DISTINCT
//     DICTIONARY ENTRY on the host side of the boundary (user code cannot run on the GPU), kept as a
//     device lookup table indexed by code, applied to the rows by k_lut_apply.
#include "dx_strings.h"
#include "dx_decode.cuh"

__global__ void __launch_bounds__(256) k_strbuild(const __grid_constant__ StrBuildArgs a, int32_t* out_codes, uint32_t* vmask32, int vbit,
                                                  unsigned long long* err_flags) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n_rows; i += (long long)gridDim.x * blockDim.x) {
        const long long row = a.sel ? a.sel[i] : i;
        uint8_t buf[DX_STRBUILD_MAX];
        uint32_t n = 0; bool isnull = false, too_long = false;
        for (int p = 0; p < a.n_parts && !isnull; p++) {
            const StrPart& sp = a.parts[p];
            if (sp.kind == SP_LIT) {
                for (uint32_t k = 0; k < sp.len; k++) { if (n < DX_STRBUILD_MAX) buf[n] = a.lit_bytes[sp.off + k]; else too_long = true; n++; }
                continue;
            }
            bool valid = true;
            if (sp.col.vmask) {
                uint32_t m = sp.col.vwidth == 1 ? reinterpret_cast<const uint8_t*>(sp.col.vmask)[row] : reinterpret_cast<const uint32_t*>(sp.col.vmask)[row];
                valid = (m >> sp.col.vbit) & 1u;
            }
            if (!valid) { isnull = true; break; }          // CONCAT / CAST of NULL is NULL
            if (sp.kind == SP_STRCOL) {
                const int32_t code = reinterpret_cast<const int32_t*>(sp.col.data)[row];
                if (code < 0) { isnull = true; break; }
                const uint32_t off = sp.dict.code_off[code], len = sp.dict.code_len[code];
                for (uint32_t k = 0; k < len; k++) { if (n < DX_STRBUILD_MAX) buf[n] = __ldcg(sp.dict.heap + off + k); else too_long = true; n++; }
            } else if (sp.kind == SP_TS) {
                // CAST(timestamp AS STRING) (Spark 2.4 DateTimeUtils.timestampToString): yyyy-MM-dd HH:mm:ss, then the fraction
                // without trailing zeros, only if there is one
                const long long us = reinterpret_cast<const long long*>(sp.col.data)[row];
                long long days = us / 86400000000ll, rem = us % 86400000000ll;
                if (rem < 0) { rem += 86400000000ll; days -= 1; }
                // civil date from days since 1970-01-01 (proleptic Gregorian)
                const long long z = days + 719468;
                const long long era = (z >= 0 ? z : z - 146096) / 146097;
                const unsigned doe = (unsigned)(z - era * 146097);
                const unsigned yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
                long long y = (long long)yoe + era * 400;
                const unsigned doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
                const unsigned mp = (5 * doy + 2) / 153;
                const unsigned d = doy - (153 * mp + 2) / 5 + 1;
                const unsigned mth = mp < 10 ? mp + 3 : mp - 9;
                if (mth <= 2) y += 1;
                const unsigned secs = (unsigned)(rem / 1000000ll), micro = (unsigned)(rem % 1000000ll);
                uint8_t tmp[40]; int m = 0;
                unsigned long long yy = y < 0 ? (unsigned long long)(-y) : (unsigned long long)y;
                if (y < 0) tmp[m++] = '-';
                { uint8_t dg[20]; int nd = 0; do { dg[nd++] = (uint8_t)('0' + yy % 10); yy /= 10; } while (yy); while (nd < 4) dg[nd++] = '0'; while (nd) tmp[m++] = dg[--nd]; }
                tmp[m++] = '-'; tmp[m++] = (uint8_t)('0' + mth / 10); tmp[m++] = (uint8_t)('0' + mth % 10);
                tmp[m++] = '-'; tmp[m++] = (uint8_t)('0' + d / 10); tmp[m++] = (uint8_t)('0' + d % 10);
                const unsigned hh = secs / 3600, mi = (secs / 60) % 60, ss = secs % 60;
                tmp[m++] = ' '; tmp[m++] = (uint8_t)('0' + hh / 10); tmp[m++] = (uint8_t)('0' + hh % 10);
                tmp[m++] = ':'; tmp[m++] = (uint8_t)('0' + mi / 10); tmp[m++] = (uint8_t)('0' + mi % 10);
                tmp[m++] = ':'; tmp[m++] = (uint8_t)('0' + ss / 10); tmp[m++] = (uint8_t)('0' + ss % 10);
                if (micro) {
                    tmp[m++] = '.';
                    unsigned f = micro, div = 100000;
                    while (f) { tmp[m++] = (uint8_t)('0' + f / div); f %= div; div /= 10; }
                }
                for (int k = 0; k < m; k++) { if (n < DX_STRBUILD_MAX) buf[n] = tmp[k]; else too_long = true; n++; }
            } else {    // SP_I64: decimal text, as Java's Long.toString
                long long v = reinterpret_cast<const long long*>(sp.col.data)[row];
                unsigned long long u = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
                uint8_t tmp[20]; int nd = 0;
                do { tmp[nd++] = (uint8_t)('0' + u % 10); u /= 10; } while (u);
                if (v < 0) { if (n < DX_STRBUILD_MAX) buf[n] = '-'; else too_long = true; n++; }
                for (int k = nd - 1; k >= 0; k--) { if (n < DX_STRBUILD_MAX) buf[n] = tmp[k]; else too_long = true; n++; }
            }
        }
        if (too_long) { atomicOr(err_flags, 32ull); isnull = true; }
        int32_t code = -1;
        if (!isnull) {
            uint64_t w0, w1; dict_pack(buf, n, w0, w1);
            code = dict_get_or_insert(a.out_dict, w0, w1, n, buf, err_flags);
        }
        out_codes[i] = code;
        if (!isnull && code >= 0) atomicOr(&vmask32[i], 1u << vbit);
    }
}

__global__ void __launch_bounds__(256) k_lut_apply(ProgCol in, const long long* lut_vals, const uint8_t* lut_valid, uint32_t lut_size, long long n_rows,
                                                   const long long* sel, void* out, int out_width, uint32_t* vmask32, int vbit) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_rows; i += (long long)gridDim.x * blockDim.x) {
        const long long row = sel ? sel[i] : i;
        bool valid = true;
        if (in.vmask) {
            uint32_t m = in.vwidth == 1 ? reinterpret_cast<const uint8_t*>(in.vmask)[row] : reinterpret_cast<const uint32_t*>(in.vmask)[row];
            valid = (m >> in.vbit) & 1u;
        }
        long long v = 0;
        if (valid) {
            const int32_t code = reinterpret_cast<const int32_t*>(in.data)[row];
            if (code >= 0 && (uint32_t)code < lut_size && lut_valid[code]) v = lut_vals[code]; else valid = false;
        }
        if (out_width == 4) reinterpret_cast<int32_t*>(out)[i] = valid ? (int32_t)v : -1; else reinterpret_cast<long long*>(out)[i] = v;
        if (valid) atomicOr(&vmask32[i], 1u << vbit);
    }
}

static unsigned grid_of(dx_ctx* c, long long n) {
    long long b = (n + 255) / 256, cap = (long long)c->sm_count * 8;
    return (unsigned)(b > cap ? cap : (b < 1 ? 1 : b));
}

int dx_strbuild(dx_ctx* ctx, StrBuildArgs& a, const std::string& lit_bytes, const DxDictP& out_dict, int32_t* out_codes, uint32_t* vmask32, int vbit) {
    if (a.n_rows <= 0) return DX_OK;
    DevBuf lits(ctx, lit_bytes.size() + 16), err(ctx, 16);
    if (!lit_bytes.empty()) DX_CUDA(ctx, cudaMemcpyAsync(lits.p, lit_bytes.data(), lit_bytes.size(), cudaMemcpyHostToDevice, ctx->stream));
    DX_CUDA(ctx, cudaMemsetAsync(err.p, 0, 16, ctx->stream));
    a.lit_bytes = lits.as<uint8_t>();
    for (int attempt = 0;; attempt++) {
        a.out_dict = out_dict->dev;
        {
            StageTimer tm(ctx, "k_strbuild");
            k_strbuild<<<grid_of(ctx, a.n_rows), 256, 0, ctx->stream>>>(a, out_codes, vmask32, vbit, err.as<unsigned long long>());
            ctx->launches["k_strbuild"]++;
        }
        DX_CUDA(ctx, cudaGetLastError());
        unsigned long long h = 0;
        DX_CUDA(ctx, cudaMemcpyAsync(&h, err.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
        DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (h & 32ull) return ctx->set_error(DX_ERR_UNSUPPORTED, "string expression produced a value longer than 256 bytes");
        if (!h) return DX_OK;
        // the result dictionary is full: rebuild it larger (codes are kept) and build the strings again
        if (attempt >= 6) return ctx->set_error(DX_ERR_CAPACITY, "string dictionary overflow while building strings");
        int rc = out_dict->grow(); if (rc) return rc;
        DX_CUDA(ctx, cudaMemsetAsync(err.p, 0, 16, ctx->stream));
        // (vmask32 bits set by the failed pass are set again by the next: every row writes the same validity)
    }
}

int dx_lut_apply(dx_ctx* ctx, const DxColumn& in, const LutCache& lut, int64_t n_rows, const int64_t* sel, void* out, int out_width,
                 uint32_t* vmask32, int vbit) {
    if (n_rows <= 0) return DX_OK;
    ProgCol pc; pc.data = in.data->p; pc.vmask = in.vmask ? in.vmask->p : nullptr; pc.vwidth = in.vwidth; pc.vbit = in.vbit; pc.width = 4; pc.pad = 0;
    StageTimer tm(ctx, "k_lut_apply");
    k_lut_apply<<<grid_of(ctx, n_rows), 256, 0, ctx->stream>>>(pc, lut.vals.as<long long>(), lut.valid.as<uint8_t>(), lut.built, n_rows,
                                                              reinterpret_cast<const long long*>(sel), out, out_width, vmask32, vbit);
    ctx->launches["k_lut_apply"]++;
    DX_CUDA(ctx, cudaGetLastError());
    return DX_OK;
}

// Brings a per-dictionary lookup table up to date: fn is called once for every dictionary entry
// that appeared since the last call.
int dx_lut_refresh(dx_ctx* ctx, LutCache* lut, const DxDictP& dict, const std::function<bool(const std::string&, long long*)>& fn) {
    int rc = dict->sync_to_host(); if (rc) return rc;
    const uint32_t n = (uint32_t)dict->host_strings.size();
    if (!lut->vals.p || lut->cap < dict->dev.max_codes) {      // first use, or the dictionary has grown since
        DevBuf nv(ctx, (size_t)dict->dev.max_codes * 8 + 16), nok(ctx, (size_t)dict->dev.max_codes + 16);
        if (!nv.p || !nok.p) return ctx->set_error(DX_ERR_CUDA, "lookup table allocation failed");
        if (lut->vals.p && lut->built) {
            DX_CUDA(ctx, cudaMemcpyAsync(nv.p, lut->vals.p, (size_t)lut->built * 8, cudaMemcpyDeviceToDevice, ctx->stream));
            DX_CUDA(ctx, cudaMemcpyAsync(nok.p, lut->valid.p, (size_t)lut->built, cudaMemcpyDeviceToDevice, ctx->stream));
            DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        }
        lut->vals = std::move(nv); lut->valid = std::move(nok); lut->cap = dict->dev.max_codes;
    }
    if (n > lut->built) {
        std::vector<long long> v(n - lut->built); std::vector<uint8_t> ok(n - lut->built);
        for (uint32_t i = lut->built; i < n; i++) { long long x = 0; ok[i - lut->built] = fn(dict->host_strings[i], &x) ? 1 : 0; v[i - lut->built] = x; }
        DX_CUDA(ctx, cudaMemcpyAsync(lut->vals.as<long long>() + lut->built, v.data(), v.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
        DX_CUDA(ctx, cudaMemcpyAsync(lut->valid.as<uint8_t>() + lut->built, ok.data(), ok.size(), cudaMemcpyHostToDevice, ctx->stream));
        DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        lut->built = n;
    }
    return DX_OK;
}
