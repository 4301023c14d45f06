// dx_strings.h — string-valued expressions over dictionary columns (see dx_strings.cu).
#pragma once
#include "dx_common.h"
#include "dx_expr.h"
#include <functional>

#define DX_STRBUILD_MAX 256
#define DX_STRBUILD_MAX_PARTS 16
enum { SP_LIT = 0, SP_STRCOL = 1, SP_I64 = 2, SP_TS = 3 };   // SP_TS: timestamp (us) as Spark casts it to string, UTC session zone

struct StrPart {
    int32_t kind; uint32_t off, len; int32_t pad;     // SP_LIT: bytes [off, off+len) of lit_bytes
    ProgCol col;                                      // SP_STRCOL / SP_I64
    DictDev dict;                                     // SP_STRCOL: the column's dictionary
};
struct StrBuildArgs {
    StrPart parts[DX_STRBUILD_MAX_PARTS];
    int32_t n_parts, pad;
    long long n_rows;
    const long long* sel;
    const uint8_t* lit_bytes;
    DictDev out_dict;
};

// device table indexed by dictionary code: a 64-bit value and a validity byte per code
struct LutCache {
    DevBuf vals, valid;
    uint32_t built = 0;          // codes [0, built) are filled
    uint32_t cap = 0;            // codes the two arrays hold (the source dictionary's max_codes when they were allocated)
    DxDictP out_dict;            // string-valued tables (UDF results): dictionary of the results
    DxDictP src_dict;            // kept alive where the cache key is the source dictionary's address
};

int dx_strbuild(dx_ctx* ctx, StrBuildArgs& a, const std::string& lit_bytes, const DxDictP& out_dict, int32_t* out_codes, uint32_t* vmask32, int vbit);
int dx_lut_apply(dx_ctx* ctx, const DxColumn& in, const LutCache& lut, int64_t n_rows, const int64_t* sel, void* out, int out_width,
                 uint32_t* vmask32, int vbit);
int dx_lut_refresh(dx_ctx* ctx, LutCache* lut, const DxDictP& dict, const std::function<bool(const std::string&, long long*)>& fn);
