// dequant.cu -- block dequantisation to fp32 (bulk kernel + launcher); the arithmetic lives in
// dequant.cuh and is bit-exact with ggml's dequantize_row_* (gguf-py quants.py:395-572).
//
// HBM-bound byte work: one thread produces four consecutive outputs from one or two packed bytes and
// writes them as a float4 (coalesced 16-byte stores; the 4.5-8.5 bit/weight reads are a small
// fraction of the 32 bit/weight written).  Blocks are read in place at their 2-byte alignment.
#include "dequant.cuh"

#include "common.h"
#include "ops.h"

namespace b2 {

template <int TYPE>
__global__ void dequant_kernel(const uint8_t* __restrict__ src, float4* __restrict__ dst, int64_t n4) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < n4; g += (int64_t)gridDim.x * blockDim.x)
        dst[g] = dq4(TYPE, src, g);
}

int launch_dequantize(int type, const void* src, float* dst, int64_t n, cudaStream_t st) {
    const TypeTraits* t = type_traits(type);
    if (!t) {
        set_error("b2_dequantize: unsupported ggml type %d", type);
        return B2_EUNSUP;
    }
    if (n < 0 || n % t->block_elems || n % 4) {
        set_error("b2_dequantize: n=%lld is not a whole number of %s blocks (and of 4)", (long long)n, t->name);
        return B2_EINVAL;
    }
    if (n == 0) return 0;
    if (!src || !dst) {
        set_error("b2_dequantize: NULL pointer");
        return B2_EINVAL;
    }
    if (((uintptr_t)dst & 15) || ((uintptr_t)src & (type == B2_TYPE_F32 ? 3 : 1))) {
        set_error("b2_dequantize: dst must be 16-byte aligned and src aligned to its element size");
        return B2_EINVAL;
    }
    const int threads = 256;
    const int64_t n4 = n / 4;
    int64_t blocks = (n4 + threads - 1) / threads;
    if (blocks > 148 * 16) blocks = 148 * 16;  // grid-stride above 16 CTAs per SM
    const uint8_t* s = (const uint8_t*)src;
    float4* d = (float4*)dst;
    switch (type) {
        case B2_TYPE_F32: dequant_kernel<B2_TYPE_F32><<<(int)blocks, threads, 0, st>>>(s, d, n4); break;
        case B2_TYPE_F16: dequant_kernel<B2_TYPE_F16><<<(int)blocks, threads, 0, st>>>(s, d, n4); break;
        case B2_TYPE_BF16: dequant_kernel<B2_TYPE_BF16><<<(int)blocks, threads, 0, st>>>(s, d, n4); break;
        case B2_TYPE_Q8_0: dequant_kernel<B2_TYPE_Q8_0><<<(int)blocks, threads, 0, st>>>(s, d, n4); break;
        case B2_TYPE_Q4_K: dequant_kernel<B2_TYPE_Q4_K><<<(int)blocks, threads, 0, st>>>(s, d, n4); break;
        case B2_TYPE_Q5_K: dequant_kernel<B2_TYPE_Q5_K><<<(int)blocks, threads, 0, st>>>(s, d, n4); break;
        case B2_TYPE_Q6_K: dequant_kernel<B2_TYPE_Q6_K><<<(int)blocks, threads, 0, st>>>(s, d, n4); break;
    }
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b2
