// ops.h -- host-side launchers of the CUDA kernels (internal; the public surface is include/b2.h)
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "common.h"

namespace b2 {

// dequant.cu
int launch_dequantize(int type, const void* src, float* dst, int64_t n, cudaStream_t st);

// attn.cu
constexpr int ATTN_SPLIT_TOKENS = 1024;  // tokens one CTA handles per pass (shared-memory score buffer)
constexpr int ATTN_MAX_SPLITS = 64;
int attn_max_splits(int n_ctx);
size_t attn_scratch_bytes(int n_head, int head_dim, int max_splits);
// pos_ptr != NULL: attend positions [0, *pos_ptr] (device-resident position, graph-replayable);
// otherwise attend [0, n_tokens_imm).
int launch_attn_decode(const float* q, const __half* kc, const __half* vc, float* out, int n_head, int n_kv, int head_dim,
                       int n_ctx, const int* pos_ptr, int n_tokens_imm, int max_splits, void* scratch, cudaStream_t st, bool pdl,
                       int n_cols = 1, const int* col_slot = nullptr, int64_t slot_stride = 0, int split_tokens = ATTN_SPLIT_TOKENS,
                       int cols_cap = 0 /* columns the scratch was sized for (>= n_cols; 0 = n_cols) */);
// split_tokens: tokens per CTA (<= ATTN_SPLIT_TOKENS); the batched path uses shorter splits for more CTAs per column
constexpr int ATTN_BATCH_SPLIT_TOKENS = 256;
// n_cols > 1: a batch of queries q[n_cols][n_head][hd] -> out, column j at position pos_ptr[j] of sequence slot
// col_slot[j] (caches slot_stride halfs apart); scratch of attn_scratch_bytes_batch()
size_t attn_scratch_bytes_batch(int n_head, int head_dim, int max_splits, int n_cols);

// attn_batch.cu: the same batched decode attention on the tensor cores (the step kernel's work item per (column, kv
// head, 256-token chunk)); scratch and cols_cap as for launch_attn_decode's batched form, max_chunks * 256 >= n_ctx
bool attn_batch_supported(int head_dim, int n_ctx, int max_chunks);
// n_ctx: rows per kv head in kc / vc (the stride).  kv_table != nullptr (paged cache): position p of slot s is row
// kv_table[s * kv_pps + p / 256] * 256 + p % 256; col_slot / slot_stride then only name the slot
int launch_attn_batch(const float* q, const __half* kc, const __half* vc, float* out, int n_head, int n_kv, int head_dim, int n_ctx,
                      const int* col_pos, const int* col_slot, int64_t slot_stride, int n_cols, int max_chunks, void* scratch, int cols_cap,
                      cudaStream_t st, const int* kv_table = nullptr, int kv_pps = 0);

// misc.cu
// x[d] = dequantised row *token_ptr of the embedding table (row a10)
int launch_embed(int type, const void* table, int64_t row_bytes, int d_model, int vocab, const int32_t* token_ptr, float* x,
                 cudaStream_t st, bool pdl);
constexpr int ARGMAX_BLOCKS = 64;
size_t argmax_scratch_bytes();
// argmax over logits -> *token_out; if step_out != NULL also appends the token to step_out[*step_ptr]
// and advances *pos_ptr / *step_ptr (the device-resident decode loop's bookkeeping)
int launch_argmax(const float* logits, int64_t n, int32_t* token_out, void* scratch, int32_t* pos_ptr, int32_t* step_ptr,
                  int32_t* step_out, int step_cap, cudaStream_t st, bool pdl);
int launch_rmsnorm(const float* x, const float* w, float* y, int64_t n, int64_t rows, float eps, cudaStream_t st);
int launch_rope(float* x, int n_heads, int head_dim, int n_dims, int mode, int pos, float freq_base, const float* freq_factors,
                cudaStream_t st);

// prefill_attn.cu: causal attention of T consecutive prompt positions [pos0, pos0 + T) of one sequence over its fp16
// cache (kc, vc: the slot's and layer's [n_kv][n_ctx][head_dim]); q, out: [T][n_head * head_dim] fp32
// pages != nullptr (paged cache): the sequence's page list -- position p is row pages[p / 256] * 256 + p % 256 of the
// n_ctx rows per kv head
int launch_prefill_attn(const float* q, const __half* kc, const __half* vc, float* out, int n_head, int n_kv, int head_dim, int n_ctx, int T, int pos0,
                        cudaStream_t st, const int* pages = nullptr);

// prefill_ops.cu: the elementwise kernels of the prompt-phase (tensor-core) path, over all T rows at once
struct PrefillQkvArgs {
    const float* qkv;            // [T][nq + 2 nk]: Q | K | V mat-mul outputs
    const float *bq, *bk, *bv;   // optional biases (qwen2)
    float* q;                    // [T][nq] rotated queries
    __half *k_cache, *v_cache;   // the slot's and layer's cache [n_kv][n_ctx][head_dim]
    const float2* rope_tab;      // [n_ctx][rope_half] (cos, sin)
    int nq, nk, head_dim, n_ctx, pos0, rope_half, neox;
    // batch form: row t is its own sequence -- position col_pos[t], cache of slot col_slot[t] (k_cache / v_cache then
    // point at slot 0); nullptr: row t is position pos0 + t of the one sequence k_cache / v_cache belong to
    const int32_t *col_pos, *col_slot;
    int64_t slot_stride;
    // paged cache: position p of slot s (the column's, or `slot` in the prompt form) is row
    // kv_table[s * kv_pps + p / 256] * 256 + p % 256 of the n_ctx rows per kv head; nullptr: contiguous slots
    const int32_t* kv_table;
    int kv_pps, slot;
    // the mat-mul's split-K reduction deferred to this kernel (tc_gemm.cuh TcParts): splits > 1 -> element (t, col) is
    // the sum over s of parts_ws[s * parts_slice + t * (nq + 2 nk) + col] and qkv is not read
    const float* parts_ws;
    int parts_splits;
    size_t parts_slice;
};
int launch_embed_rows(int type, const void* table, int64_t row_bytes, int d_model, int vocab, const int32_t* tokens, int T, float* x, cudaStream_t st);
int launch_prefill_qkv_post(const PrefillQkvArgs& a, int T, cudaStream_t st);
int launch_argmax_rows(const float* logits, int64_t n, int rows, int32_t* out, cudaStream_t st);

// sample.cu: device-side sampling, one CTA per logits row.  only_sampled: rows whose parameters are greedy are left
// untouched (the caller's argmax already wrote them).  draw_i32 (optional, instead of draw_dev): one device int32 used as
// the draw index of every row (the single-sequence loop's position); echo: also store the token at echo_out[*echo_step - 1]
// (the device-resident loop's output list).
constexpr int SAMPLE_MAX_K = B2_SAMPLE_MAX_K;
int launch_sample_rows(const float* logits, int64_t vocab, int64_t row_stride, int rows, const b2_sampling* params_dev, int64_t param_stride,
                       const uint64_t* draw_dev, int32_t* out_dev, bool only_sampled, cudaStream_t st, const int32_t* draw_i32 = nullptr,
                       int32_t* echo_out = nullptr, const int32_t* echo_step = nullptr, int echo_cap = 0);

// tp.cu: the tensor-parallel exchange (NCCL through dlopen)
int tp_unique_id(void* out, size_t cap);                                   // -> bytes written (128) or an error
int tp_comm_init(const void* id_bytes, size_t len, int rank, int size, void** comm_out);
void tp_comm_free(void* comm);
int tp_allreduce_inplace(void* comm, float* x, int n, cudaStream_t st);            // x = allreduce_sum(x)
// peer memory for the exchange inside the persistent kernel: every rank's zeroed buffer of `bytes`, mapped into every rank
int tp_peer_buffers(void* comm, int rank, int size, size_t bytes, void** local_out, void** peers, cudaStream_t st);
void tp_peer_release(int rank, int size, void* local, void** peers);
int tp_barrier(void* comm, float* scratch_dev, cudaStream_t st);

}  // namespace b2
