// tcq_gemm.cuh -- mat-mul of QUANTISED weights on the 5th-generation tensor cores with the dequantisation fused
// into the producer (tcq_gemm.cu): raw GGUF blocks -> TMA -> shared memory -> dequantising warps -> 128-byte-swizzled
// fp16 tiles in shared memory -> tcgen05.mma -> tensor memory.
//
// Y[T][N] = X[T][K] . W[N][K]^T for T of a decode batch (2 .. 128 rows; SURVEY 8a row a12 with n_tokens = T), W a stack
// of up to three GGUF tensors of types Q4_K / Q6_K (the Q | K | V or gate | up matrices of a layer), read from HBM once
// per pass in their quantised form.  The values the tensor cores see are the ones tc_gemm.cu's resident fp16 copy
// holds (ggml's dequantised value rounded once to fp16), so both paths give the same sums.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "tc_gemm.cuh"

namespace b2 {

constexpr int TCQ_MAX_SEGS = 3;
constexpr int TCQ_MAX_T = 128;

struct TcqSeg {
    const void* w;   // [rows][K] blocks of `type`
    int type;        // B2_TYPE_Q4_K or B2_TYPE_Q6_K
    int rows;
};

struct TcqGemmArgs {
    TcqSeg seg[TCQ_MAX_SEGS];   // stacked along N in this order
    int n_seg;
    const __half* x;            // [T][K] activations (launch_rows_to_f16 / launch_swiglu_to_f16)
    int K, T;
    // 1: x was written with `perm` (elements 1 and 2 of every group of four swapped) and the kernel converts the codes in
    // fp16 arithmetic (scale and minimum rounded to fp16 before the multiply-add: <= 2 ulp of the fp16 weight instead of the
    // exact path's 1/2 ulp; 1.6 instead of 3.4 instructions per weight).  0: natural order, exact conversion.
    int x_perm;
    // The result is always left as slices: out(t, n) = sum over s < splits of ws[s * T * N + t * N + n] (N = the stacked
    // rows), to be added up in slice order by the consumer (TcParts) or by tcq_reduce.  Slices a tile does not need hold zeros.
    float* ws;
    size_t ws_bytes;
};

bool tcq_fast_conversion();                                         // the default; B2_TCQ_EXACT=1: exact conversion, natural K order
bool tcq_supported(int type, int K);                                // Q4_K: K a multiple of 256; Q6_K: of 2048 (16-byte row stride)
// host arithmetic of the stream-K decomposition (no GPU needed): see tcq_gemm.cu
int tcq_plan(const int* seg_rows, int n_seg, int K, int n_sms, int* tiles_out, int* grid_out, int* slices_out);
void tcq_share(int tiles, int nblk, int grid, int c, long long* lo, long long* hi);
int tcq_first_cta_of_tile(int tiles, int nblk, int grid, int tile);
size_t tcq_ws_bytes(int N, int K, int T);                           // upper bound of the slices' size for this shape
// runs the mat-mul; *parts describes the slices (splits >= 1)
int launch_tcq_gemm(const TcqGemmArgs& a, cudaStream_t st, TcParts* parts);
// out[t * ldo + n] (+)= sum of the slices (+ bias[n]): for outputs no consumer kernel adds up itself
int launch_tcq_reduce(const TcParts& parts, int T, int N, float* out, int64_t ldo, const float* bias, int epi, cudaStream_t st);

}  // namespace b2
