// dx_decode_framed.cu — K1 for caller-provided record offsets (FRAMED = true instantiations of dx_decode_kernel.cuh):
// no structural scan, so the kernel holds no scan state and is built for 5 resident CTAs per SM.
#include "dx_decode_kernel.cuh"

template <int T> static int setup_one(int* occ) {
    cudaError_t e = cudaFuncSetAttribute(k1_decode<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k1_smem_bytes(T));
    if (e != cudaSuccess) return (int)e;
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, k1_decode<T, true>, K1_NT, k1_smem_bytes(T));
}

int k1_framed_setup(int tile, int* occ) {
    switch (tile) {
    case K1_TS(16384): return setup_one<K1_TS(16384)>(occ);
    case K1_TS(20480): return setup_one<K1_TS(20480)>(occ);
    case K1_TS(24576): return setup_one<K1_TS(24576)>(occ);
    case K1_TS(28672): return setup_one<K1_TS(28672)>(occ);
    case K1_TS(36864): return setup_one<K1_TS(36864)>(occ);
    case K1_TS(40960): return setup_one<K1_TS(40960)>(occ);
    case K1_TS(45056): return setup_one<K1_TS(45056)>(occ);
    case K1_TS(49152): return setup_one<K1_TS(49152)>(occ);
    default: return setup_one<K1_TS(32768)>(occ);
    }
}

void k1_framed_launch(int tile, unsigned grid, cudaStream_t stream, const DecodeArgs& a) {
    const size_t smem = k1_smem_bytes(tile);
    switch (tile) {
    case K1_TS(16384): k1_decode<K1_TS(16384), true><<<grid, K1_NT, smem, stream>>>(a); break;
    case K1_TS(20480): k1_decode<K1_TS(20480), true><<<grid, K1_NT, smem, stream>>>(a); break;
    case K1_TS(24576): k1_decode<K1_TS(24576), true><<<grid, K1_NT, smem, stream>>>(a); break;
    case K1_TS(28672): k1_decode<K1_TS(28672), true><<<grid, K1_NT, smem, stream>>>(a); break;
    case K1_TS(36864): k1_decode<K1_TS(36864), true><<<grid, K1_NT, smem, stream>>>(a); break;
    case K1_TS(40960): k1_decode<K1_TS(40960), true><<<grid, K1_NT, smem, stream>>>(a); break;
    case K1_TS(45056): k1_decode<K1_TS(45056), true><<<grid, K1_NT, smem, stream>>>(a); break;
    case K1_TS(49152): k1_decode<K1_TS(49152), true><<<grid, K1_NT, smem, stream>>>(a); break;
    default: k1_decode<K1_TS(32768), true><<<grid, K1_NT, smem, stream>>>(a); break;
    }
}
