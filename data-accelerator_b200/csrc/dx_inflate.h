// dx_inflate.h — K7: gzip members inflated on the device (see dx_inflate.cu).
#pragma once
#include "dx_common.h"
// gz: n_members gzip members back to back (member m = bytes [member_offsets[m], member_offsets[m+1])), in host (memory_kind 0) or
// device (1) memory.  *out: one device buffer with the members' contents in order (pool memory; release with the pool).
int dx_gunzip(dx_ctx* ctx, const uint8_t* gz, const int64_t* member_offsets, int n_members, int memory_kind, DevBufP* out, int64_t* n_bytes);
