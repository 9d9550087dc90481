// gemv.cuh -- launch descriptor of the dequant-fused mat-vec kernel (gemv.cu)
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2 {

// how the two rows a consumer warp works on are chosen
enum { PAIR_ADJ = 0,    // rows 2p, 2p+1 of matrix A           (plain, llama-style RoPE pairs, V)
       PAIR_SPLIT = 1,  // row p of matrix A and row p of B    (ffn_gate / ffn_up -> SwiGLU)
       PAIR_NEOX = 2 }; // rows h*hd+i and h*hd+i+hd/2 of A    (qwen2-style RoPE pairs)

enum { EPI_STORE = 0,   // out[row] = v (+bias)
       EPI_RESID = 1,   // out[row] += v                        (attn_output, ffn_down: residual add)
       EPI_SWIGLU = 2,  // out[p] = silu(v_a) * v_b
       EPI_ROPE_Q = 3,  // rotate the pair, store fp32 q
       EPI_ROPE_K = 4,  // rotate the pair, store fp16 into the K cache at the current position
       EPI_V = 5,       // store fp16 into the V cache at the current position
       EPI_RESID_TP = 6 };  // out[row] = out[row] * res_scale + v: a tensor-parallel shard's partial sum plus its 1/tp share
                            // of the residual, so that the all-reduce over the ranks yields residual + sum in place

enum { PRO_PLAIN = 0,   // activation = x
       PRO_NORM = 1 };  // activation = rmsnorm(x) * norm_w

enum { ACT_Q8K = 0, ACT_Q80 = 1 };

struct GemvSeg {
    const uint8_t* wa;
    const uint8_t* wb;
    const float* bias;   // indexed by row of A (and of B for PAIR_SPLIT: bias_b)
    const float* bias_b;
    float* out;
    __half* cache;       // [n_kv][n_ctx][head_dim] for EPI_ROPE_K / EPI_V
    int type, pairing, epi;
    int n_pairs, n_rows, tile0;
    int head_dim, n_ctx;
    float res_scale;     // EPI_RESID_TP
    // EPI_RESID_TP inside the persistent kernel of a tensor-parallel shard: instead of updating `out` the epilogue stores
    // res * res_scale + partial at the same row of every rank's inbox (peer memory, plain stores); out is only read
    float* peer_out[8];
    int n_peer;          // 0: update `out` in place (the all-reduce follows as a separate call)
    int peer_local;      // 1: peer_out[0] is this GPU's own memory (the peers fetch the row: a device-scope fence is enough)
};

struct GemvArgs {
    GemvSeg seg[3];
    int n_seg, n_tiles;
    const float* x;        // [K] fp32 activation
    const float* norm_w;   // [K] for PRO_NORM
    float eps;
    int K, prologue;
    const int* pos_ptr;    // device-resident position (RoPE / cache epilogues), may be NULL otherwise
    const float2* rope_tab;  // [n_ctx][rope_half] (cos, sin)
    int rope_half;
    // Tensor-parallel shard inside the persistent kernel (step2.cu): the activation is the sum, in rank order, of x_nparts
    // vectors of K floats at x_parts + r * K -- the partial residual streams the ranks' row-split epilogues stored into this
    // rank's inbox -- once xg_flags[r] (written by rank r) have reached the stage's sequence number.  The stage also leaves
    // the sum in x (the residual the next row-split epilogue reads).  x_nparts == 0: plain x.
    const float* x_parts;
    int x_nparts, xg_k;
    const unsigned int* xg_flags;
    unsigned int* xg_peer_flags[8];
    // pull form of the exchange (OP_XCHG): rank r's partial row is read from x_part_ptr[r] -- this rank's own buffer or the
    // peer-mapped address of rank r's -- instead of from x_parts + r * K (the inbox the ranks stored into); NULL: inbox form
    const float* x_part_ptr[8];
    const unsigned int* xg_epoch;
};

constexpr int GEMV_CONSUMER_WARPS = 8;
constexpr int GEMV_THREADS = (GEMV_CONSUMER_WARPS + 3) * 32;  // + producer, header decoder, epilogue warps
constexpr int GEMV_TILE_PAIRS = GEMV_CONSUMER_WARPS / 2;  // two consumer warps per row pair (K split)
constexpr int GEMV_STAGES = 3;

// fills tile0 / n_tiles from n_pairs; returns 0 or a negative B2_E* code after validating shapes
int gemv_finalize(GemvArgs& a, int act);
int gemv_act_for_type(int type);  // ACT_* or -1
size_t gemv_smem_bytes(int act, int K);
int launch_gemv(const GemvArgs& a, int act, cudaStream_t st, bool pdl);

}  // namespace b2
