// dequant.cuh -- per-group (4 consecutive elements) block dequantisation, shared by the bulk
// dequantise kernels (dequant.cu) and the embedding-row lookup (misc.cu).  Bit-exact with ggml's
// dequantize_row_* as restated by gguf-py (quants.py:395-401 Q8_0, :475-522 Q4_K, :525-549 Q5_K,
// :552-572 Q6_K): the association is (d*sc)*q - (dmin*m); every product is exact in fp32 (SURVEY
// Appendix B) so mul+sub and fma give identical bits.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b2.h"

namespace b2 {

__device__ __forceinline__ void scale_min_k4(int j, const uint8_t* q, int* sc, int* m) {
    if (j < 4) {
        *sc = q[j] & 63;
        *m = q[j + 4] & 63;
    } else {
        *sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        *m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
    }
}

// elements [4g, 4g+4) of a run of Q8_0 blocks starting at src
__device__ __forceinline__ float4 dq4_q8_0(const uint8_t* __restrict__ src, int64_t g) {
    const int64_t e = g * 4;
    const uint8_t* blk = src + (e >> 5) * 34;
    const float d = __half2float(*(const __half*)blk);
    const int8_t* q = (const int8_t*)(blk + 2 + (e & 31));
    return make_float4(q[0] * d, q[1] * d, q[2] * d, q[3] * d);
}

template <bool Q5>
__device__ __forceinline__ float4 dq4_q45_K(const uint8_t* __restrict__ src, int64_t g) {
    constexpr int BB = Q5 ? 176 : 144;
    const int64_t e = g * 4;
    const uint8_t* blk = src + (e >> 8) * BB;
    const int ie = (int)(e & 255);   // element inside the super-block
    const int s = ie >> 5;           // sub-block 0..7
    const int l = ie & 31;           // position inside the sub-block (multiple of 4)
    const float d = __half2float(*(const __half*)blk);
    const float dmin = __half2float(*(const __half*)(blk + 2));
    int sc, m;
    scale_min_k4(s, blk + 4, &sc, &m);
    const float ds = d * (float)sc, dm = dmin * (float)m;
    const uint8_t* qs = blk + (Q5 ? 48 : 16) + (s >> 1) * 32 + l;
    const int shift = (s & 1) * 4;
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int q = (qs[i] >> shift) & 0xF;
        if (Q5) q |= ((blk[16 + l + i] >> s) & 1) << 4;
        o[i] = ds * (float)q - dm;
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

__device__ __forceinline__ float4 dq4_q6_K(const uint8_t* __restrict__ src, int64_t g) {
    const int64_t e = g * 4;
    const uint8_t* blk = src + (e >> 8) * 210;
    const int ie = (int)(e & 255);
    const int h = ie >> 7;         // half 0/1
    const int j = (ie >> 5) & 3;   // 32-row inside the half
    const int l = ie & 31;
    const float d = __half2float(*(const __half*)(blk + 208));
    const int sc = ((const int8_t*)(blk + 192))[ie >> 4];
    const float ds = d * (float)sc;
    const uint8_t* ql = blk + 64 * h + 32 * (j & 1) + l;
    const uint8_t* qh = blk + 128 + 32 * h + l;
    const int lsh = (j >> 1) * 4, hsh = 2 * j;
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int q = (((ql[i] >> lsh) & 0xF) | (((qh[i] >> hsh) & 3) << 4)) - 32;
        o[i] = ds * (float)q;
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

__device__ __forceinline__ float4 dq4_f16(const uint8_t* __restrict__ src, int64_t g) {
    const __half* h = (const __half*)src + g * 4;
    return make_float4(__half2float(h[0]), __half2float(h[1]), __half2float(h[2]), __half2float(h[3]));
}
__device__ __forceinline__ float4 dq4_bf16(const uint8_t* __restrict__ src, int64_t g) {
    const uint16_t* h = (const uint16_t*)src + g * 4;
    return make_float4(__uint_as_float((uint32_t)h[0] << 16), __uint_as_float((uint32_t)h[1] << 16),
                       __uint_as_float((uint32_t)h[2] << 16), __uint_as_float((uint32_t)h[3] << 16));
}
__device__ __forceinline__ float4 dq4_f32(const uint8_t* __restrict__ src, int64_t g) {
    const float* f = (const float*)src + g * 4;
    return make_float4(f[0], f[1], f[2], f[3]);
}

// type-dispatched group dequantise (type is uniform across the grid)
__device__ __forceinline__ float4 dq4(int type, const uint8_t* __restrict__ src, int64_t g) {
    switch (type) {
        case B2_TYPE_Q8_0: return dq4_q8_0(src, g);
        case B2_TYPE_Q4_K: return dq4_q45_K<false>(src, g);
        case B2_TYPE_Q5_K: return dq4_q45_K<true>(src, g);
        case B2_TYPE_Q6_K: return dq4_q6_K(src, g);
        case B2_TYPE_F16: return dq4_f16(src, g);
        case B2_TYPE_BF16: return dq4_bf16(src, g);
        default: return dq4_f32(src, g);
    }
}

}  // namespace b2
