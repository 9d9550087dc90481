// gemm.cuh -- batched quantised mat-mul Y[B x N] = X[B x K] . W[N x K]^T for 2 <= B <= 32 columns
// (batched decode and chunked prefill; SURVEY 8a row a12 with B in {8, 32} and prefill).
//
// Same arithmetic as the single-column mat-vec (gemv.cu): activations quantised to Q8_K, integer dot
// products per 32-element sub-block, fp32 scales -- but the dot products run on the tensor cores
// (mma.sync m16n8k32 u8 x s8 -> s32: 16 weight rows x 8 columns per instruction), so one pass over the
// weights serves up to 32 columns.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "gemv.cuh"

namespace b2 {

constexpr int GEMM_CHUNK_BLOCKS = 8;                          // K-chunk of one pipeline stage: 8 super-blocks
constexpr int GEMM_CHUNK_K = GEMM_CHUNK_BLOCKS * 256;         // = 2048 elements
constexpr int GEMM_TILE_ROWS = 16;                            // weight rows per tile = M of the MMA
constexpr int GEMM_TILE_PAIRS = GEMM_TILE_ROWS / 2;
constexpr int GEMM_MAX_COLS = 32;
constexpr int GEMM_Q_PITCH = GEMM_CHUNK_K + 32;               // bytes between columns of a chunk (8 words mod 32 banks)

// Quantised activations of one mat-mul input, as the kernel's producer warp copies them (one bulk copy
// per K-chunk): chunk c = { int8 q[NB][GEMM_Q_PITCH]; float d[8][NB]; uint8 bl[8][NB][8]; int8 bh[8][NB][8] }
// with NB = columns rounded up to 8; bl/bh = low 7 bits / the rest of the sum of each 32 quantised values.
__host__ __device__ inline int gemm_nb(int cols) { return cols <= 8 ? 8 : (cols <= 16 ? 16 : 32); }   // column groups: 1, 2 or 4
__host__ __device__ inline size_t gemm_act_chunk_bytes(int nb) {
    return (size_t)nb * GEMM_Q_PITCH + (size_t)GEMM_CHUNK_BLOCKS * nb * (4 + 8 + 8);
}
// Q8_0 weights take Q8_0 activations (32-element blocks, fp16-rounded scale each) in K-chunks of 32 blocks:
// chunk c = { int8 q[NB][GEMM80_Q_PITCH]; float d[32][NB] }
constexpr int GEMM80_CHUNK_BLOCKS = 32;
constexpr int GEMM80_CHUNK_K = GEMM80_CHUNK_BLOCKS * 32;      // = 1024 elements
constexpr int GEMM80_Q_PITCH = GEMM80_CHUNK_K + 32;
__host__ __device__ inline size_t gemm80_act_chunk_bytes(int nb) { return (size_t)nb * GEMM80_Q_PITCH + (size_t)GEMM80_CHUNK_BLOCKS * nb * 4; }
// act_fmt: ACT_Q8K (K-quant weights) or ACT_Q80 (Q8_0 weights)
inline size_t gemm_act_bytes(int K, int cols, int act_fmt = ACT_Q8K) {
    if (act_fmt == ACT_Q80) return (size_t)((K + GEMM80_CHUNK_K - 1) / GEMM80_CHUNK_K) * gemm80_act_chunk_bytes(gemm_nb(cols));
    return (size_t)((K + GEMM_CHUNK_K - 1) / GEMM_CHUNK_K) * gemm_act_chunk_bytes(gemm_nb(cols));
}

// per-column bookkeeping of a batch: which sequence slot and position a column belongs to
struct GemmCols {
    int n;                 // columns in use (<= GEMM_MAX_COLS)
    const int* pos;        // [n] device: position of each column's token (RoPE, KV write)
    const int* slot;       // [n] device: sequence slot of each column (KV write), NULL = slot 0
};

// One launch: up to three row segments (as GemvSeg) sharing the input.  Outputs are [column][row] with the
// given column strides; the fused epilogues are those of the mat-vec.
struct GemmArgs {
    GemvSeg seg[3];
    int n_seg, n_tiles;          // tiles of GEMM_TILE_ROWS rows
    int K;
    int act_fmt;                 // ACT_Q8K / ACT_Q80, set by gemm_finalize from the segment types (all segments agree)
    const uint8_t* act;          // quantised activations (gemm_act_bytes)
    GemmCols cols;
    int64_t out_stride;          // floats between columns of seg[].out (and of the residual input for EPI_RESID)
    int64_t slot_stride;         // halfs between sequence slots in seg[].cache (kv_table == nullptr)
    // paged KV: row `pos` of sequence slot s lives at token kv_table[s * kv_pps + pos / 256] * 256 + pos % 256 of the
    // cache (seg[].n_ctx = tokens in the whole pool); nullptr: slot s is the contiguous window at s * slot_stride
    const int32_t* kv_table;
    int kv_pps;
    const float2* rope_tab;
    int rope_half;
};

// X[cols][K] fp32 (column stride x_stride) -> quantised chunks; optional RMS norm (norm_w != NULL)
int launch_act_quant(const float* x, int64_t x_stride, const float* norm_w, float eps, int K, int cols, uint8_t* act, cudaStream_t st,
                     int act_fmt = ACT_Q8K);
int gemm_finalize(GemmArgs& a);
bool gemm_supports_type(int type);   // Q4_K, Q5_K, Q6_K, Q8_0
int gemm_act_for_type(int type);     // ACT_Q8K / ACT_Q80 / -1
int launch_gemm(const GemmArgs& a, cudaStream_t st);

}  // namespace b2
