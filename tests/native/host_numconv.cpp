// Host build of csrc/dx_numconv.cuh for the CPU test-suite: the device code is compiled unchanged
// with the CUDA intrinsics it uses mapped to their host equivalents.
#include <stdint.h>
#include <string.h>
#define __device__
#define __constant__ const
#define __forceinline__ inline
#define __noinline__
static inline int __clzll(long long x) { return x == 0 ? 64 : __builtin_clzll((unsigned long long)x); }
static inline uint64_t __umul64hi(uint64_t a, uint64_t b) { return (uint64_t)(((unsigned __int128)a * b) >> 64); }
static inline double __longlong_as_double(long long v) { double d; memcpy(&d, &v, 8); return d; }
#include "dx_numconv.cuh"

extern "C" int numconv_parse(const char* s, int len, int is_int, double* out) {
    return dx_number_to_double((const uint8_t*)s, (const uint8_t*)s + len, is_int != 0, out) ? 1 : 0;
}
