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
    case 16384: return setup_one<16384>(occ);
    case 20480: return setup_one<20480>(occ);
    case 24576: return setup_one<24576>(occ);
    case 28672: return setup_one<28672>(occ);
    case 36864: return setup_one<36864>(occ);
    case 40960: return setup_one<40960>(occ);
    case 45056: return setup_one<45056>(occ);
    case 49152: return setup_one<49152>(occ);
    default: return setup_one<32768>(occ);
    }
}

void k1_framed_launch(int tile, unsigned grid, cudaStream_t stream, const DecodeArgs& a) {
    const size_t smem = k1_smem_bytes(tile);
    switch (tile) {
    case 16384: k1_decode<16384, true><<<grid, K1_NT, smem, stream>>>(a); break;
    case 20480: k1_decode<20480, true><<<grid, K1_NT, smem, stream>>>(a); break;
    case 24576: k1_decode<24576, true><<<grid, K1_NT, smem, stream>>>(a); break;
    case 28672: k1_decode<28672, true><<<grid, K1_NT, smem, stream>>>(a); break;
    case 36864: k1_decode<36864, true><<<grid, K1_NT, smem, stream>>>(a); break;
    case 40960: k1_decode<40960, true><<<grid, K1_NT, smem, stream>>>(a); break;
    case 45056: k1_decode<45056, true><<<grid, K1_NT, smem, stream>>>(a); break;
    case 49152: k1_decode<49152, true><<<grid, K1_NT, smem, stream>>>(a); break;
    default: k1_decode<32768, true><<<grid, K1_NT, smem, stream>>>(a); break;
    }
}
