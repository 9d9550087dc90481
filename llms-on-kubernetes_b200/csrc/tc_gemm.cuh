// tc_gemm.cuh -- prompt-phase mat-mul on the 5th-generation tensor cores (tc_gemm.cu)
//
// Y[T][N] (+)= X[T][K] . W[N][K]^T for T up to thousands of prompt tokens (SURVEY 8a row a12 with n_tokens = T,
// the reference's ggml_compute_forward_mul_mat prompt path).  The quantised matrix is dequantised to the fp16 source
// of the tiles (launch_dequant_f16: ggml's exact values rounded once to fp16), the activations are normalised,
// quantised as the reference quantises them and rounded to fp16 (launch_rows_to_f16), and one persistent kernel
// streams both through TMA (128-byte swizzle) into shared memory and multiplies them with tcgen05.mma
// (128 x BN x 16, fp32 accumulators in tensor memory).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2 {

constexpr int TC_BM = 128;   // weight rows per tile = UMMA M = tensor-memory lanes
constexpr int TC_BK = 64;    // K elements per pipeline stage = one 128-byte swizzled row of fp16

enum { TC_EPI_STORE = 0, TC_EPI_ADD = 1 };

struct TcGemmArgs {
    const __half* w;          // [N][K] dequantised weights
    const __half* x;          // [T][K] activations
    float* out;               // [T][ldo]; element (t, n) at out[t * ldo + n]
    const float* bias;        // [N] or nullptr (F32)
    int N, K, T;
    int64_t ldo;
    int epi;                  // TC_EPI_STORE: out = acc (+ bias); TC_EPI_ADD: out += acc (+ bias)
    // split-K workspace (optional): with few output tiles (short prompts, decode batches, narrow matrices) the launcher
    // cuts K across `splits` CTAs per tile, partial sums go to ws[splits][T][N] and a second kernel reduces them in a
    // fixed order
    float* ws;
    size_t ws_bytes;
    int splits;               // set by launch_tc_gemm
    int defer_reduce;         // set by launch_tc_gemm
};
size_t tc_gemm_ws_bytes(int N, int K, int T);   // workspace that lets launch_tc_gemm split K for this shape (0: no split)

// a mat-mul output whose split-K reduction is left to the kernel that consumes it: the sum over s < splits of
// ws[s * slice + t * N + n]; splits <= 1: nothing pending, the output buffer holds the result
struct TcParts {
    const float* ws;
    int splits;
    size_t slice;
};

// K must be a multiple of TC_BK; w and x 16-byte aligned.  deferred != nullptr: when the launcher splits K it skips the
// reduction kernel (and the epilogue: `out` is NOT touched) and describes the slices in *deferred for the consumer
// (launch_rows_to_f16's `pend`, launch_swiglu_to_f16, the QKV post kernel); without a split *deferred says so
int launch_tc_gemm(const TcGemmArgs& a, cudaStream_t st, TcParts* deferred = nullptr);

// n elements of `type` blocks -> fp16 (round-to-nearest-even of the exact ggml dequantised value, saturating)
int launch_dequant_f16(int type, const void* src, __half* dst, int64_t n, cudaStream_t st);

// y[t][:] = fp16(Q(x[t][:] * rsqrt(mean(x[t]^2) + eps) * w[:]))  (no norm when w == nullptr), where Q quantises and
// dequantises the row as the reference's mat-mul would: act_fmt ACT_Q8K (K-quant weights: 256-element Q8_K blocks),
// ACT_Q80 (Q8_0 weights: 32-element Q8_0 blocks), or < 0 for no quantisation (plain rounding).  pend: a deferred
// residual mat-mul (out += ...) into x: its slices are added to x[t][:] first (x is written back)
int launch_rows_to_f16(float* x, int64_t ldx, const float* norm_w, float eps, int n, int rows, int act_fmt, __half* y, cudaStream_t st,
                       TcParts pend = TcParts{nullptr, 1, 0}, bool perm = false);   // perm: see TcqGemmArgs::x_perm

// y[t][:] = fp16(Q(silu(gate) * up)) of the stacked gate | up output gu[T][2 d_ff] (or its deferred slices)
int launch_swiglu_to_f16(const float* gu, TcParts parts, int d_ff, int T, int act_fmt, __half* y, float* h_last, cudaStream_t st, bool perm = false);

}  // namespace b2
