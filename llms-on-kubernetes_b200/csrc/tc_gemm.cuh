// tc_gemm.cuh -- prompt-phase mat-mul on the 5th-generation tensor cores (tc_gemm.cu)
//
// Y[T][N] (+)= X[T][K] . W[N][K]^T for T up to thousands of prompt tokens (SURVEY 8a row a12 with n_tokens = T,
// the reference's ggml_compute_forward_mul_mat prompt path).  The quantised matrix is dequantised to bf16 tiles'
// source (launch_dequant_bf16, bit-exact values rounded once to bf16), the activations are normalised and rounded
// to bf16, and one persistent kernel streams both through TMA (128-byte swizzle) into shared memory and
// multiplies them with tcgen05.mma (128 x BN x 16, fp32 accumulators in tensor memory).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2 {

constexpr int TC_BM = 128;   // weight rows per tile = UMMA M = tensor-memory lanes
constexpr int TC_BK = 64;    // K elements per pipeline stage = one 128-byte swizzled row of bf16

enum { TC_EPI_STORE = 0, TC_EPI_ADD = 1 };

struct TcGemmArgs {
    const __nv_bfloat16* w;   // [N][K] dequantised weights
    const __nv_bfloat16* x;   // [T][K] activations
    float* out;               // [T][ldo]; element (t, n) at out[t * ldo + n]
    const float* bias;        // [N] or nullptr (F32)
    int N, K, T;
    int64_t ldo;
    int epi;                  // TC_EPI_STORE: out = acc (+ bias); TC_EPI_ADD: out += acc (+ bias)
};

// K must be a multiple of TC_BK; w and x 16-byte aligned
int launch_tc_gemm(const TcGemmArgs& a, cudaStream_t st);

// n elements of `type` blocks -> bf16 (round-to-nearest-even of the exact ggml dequantised value)
int launch_dequant_bf16(int type, const void* src, __nv_bfloat16* dst, int64_t n, cudaStream_t st);

// y[t][:] = bf16(x[t][:] * rsqrt(mean(x[t]^2) + eps) * w[:]) when w != nullptr, bf16(x[t][:]) otherwise
int launch_rows_to_bf16(const float* x, int64_t ldx, const float* norm_w, float eps, int n, int rows, __nv_bfloat16* y, cudaStream_t st);

}  // namespace b2
