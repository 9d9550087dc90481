// model.h -- weights in HBM + one decode context (internal; public surface is include/b2.h)
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <memory>
#include <vector>

#include <mutex>

#include "gguf.h"

struct b2_model {
    struct W {
        const uint8_t* p = nullptr;  // device pointer to the raw blocks
        int type = 0;
        int64_t rows = 0, cols = 0;
    };
    struct Layer {
        W attn_norm, wq, wk, wv, wo, ffn_norm, gate, up, down;
        const float *bq = nullptr, *bk = nullptr, *bv = nullptr;
    };
    std::unique_ptr<b2::Gguf> g;
    b2_gguf gguf_handle;   // non-owning view handed out by b2_model_gguf()
    int device = 0;
    uint8_t* d_alloc = nullptr;  // allocation (guard bytes both sides)
    uint8_t* d_data = nullptr;   // image of the GGUF data section, byte for byte
    b2_model_info info{};
    W tok_embd, output, output_norm;
    const float* rope_freqs = nullptr;
    std::vector<float> rope_freqs_host;
    std::vector<Layer> layers;
    // tensor parallelism: this process holds shard tp_rank of tp_size.  info.n_head / n_kv / d_ff are then
    // the shard's; attn_output and ffn_down point at compact copies of their K slices (tp_allocs), the row-split
    // matrices at their row range inside the uploaded file image
    int tp_rank = 0, tp_size = 1;
    std::vector<void*> tp_allocs;
    // fp16 copies of the layer matrices for the tensor-core prompt / wide-batch path (tc_gemm.cu), built on first
    // use when they fit the budget (B2_F16_CACHE_GB, default 64 GB and at most half of the free HBM): per layer the
    // stacked Q | K | V, O, stacked gate | up, down.  Without it the path dequantises each matrix per pass.
    void* f16_cache = nullptr;
    bool f16_tried = false;
    struct F16Layer { size_t qkv, o, gate_up, down; };   // element offsets into f16_cache
    size_t f16_output = 0;                                // the lm_head matrix
    std::vector<F16Layer> f16_off;
    std::mutex f16_mu;
};

constexpr int BATCH_MAX_COLS = 256;
constexpr int KV_PAGE = 256;   // positions per page of the KV pool = the attention kernels' chunk

struct b2_ctx {
    b2_model* m = nullptr;
    int n_ctx = 0;
    cudaStream_t stream = nullptr;
    __half *k_cache = nullptr, *v_cache = nullptr;  // [layer][n_kv][n_ctx][head_dim]
    float *x = nullptr, *q = nullptr, *attn_out = nullptr, *h = nullptr, *logits = nullptr;
    int32_t *d_tok = nullptr, *d_pos = nullptr, *d_step = nullptr, *d_out = nullptr, *d_prompt = nullptr;
    int out_cap = 0, prompt_cap = 0;
    float2* rope_tab = nullptr;
    void *attn_scratch = nullptr, *argmax_scratch = nullptr;
    int max_splits = 1;
    int32_t* h_pin = nullptr;  // pinned host staging (token ids in / out)
    int h_pin_cap = 0;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    bool use_graph = true, use_pdl = true;
    int launches_per_step = 0, matvec_launches = 0;
    // persistent per-token kernel (step.cu): program + scratch; use_step false = one launch per stage
    bool use_step = true;
    void* d_ops = nullptr;   // StepOp[n_ops]
    int n_ops = 0, step_act = -1, step_max_k = 0;
    int step_version = 2, step_l2_ahead = 0, step_fine_op = -1;   // step2.cu unless B2_STEP_V=1 or a K it does not take
    float* step_attn_partials = nullptr;
    unsigned int* step_counters = nullptr;   // [1] argmax arrivals, [8 .. 8 + n_head) attention arrivals
    unsigned int* step_sync = nullptr;       // grid barrier: arrival counter + flags (STEP_SYNC_BYTES)
    unsigned long long* step_trace = nullptr;   // B2_STEP_TRACE=1: per-stage time stamps of CTA 0
    std::vector<int> step_desc;                 // per stage: kind, K, tiles (for the trace dump)
    float* step_argmax_val = nullptr;
    int* step_argmax_idx = nullptr;
    int pos = 0;
    // device-side sampling (b2_ctx_set_sampler): per-slot parameters (temperature <= 0: greedy), their device copy, and the
    // per-column parameter / draw-index arrays of the current batched pass (pinned host staging + device)
    std::vector<b2_sampling> samplers;
    int n_samplers = 0;                     // slots with a non-greedy sampler
    b2_sampling* d_samplers = nullptr;      // [n_slots]
    b2_sampling *h_col_params = nullptr, *d_col_params = nullptr;   // [BATCH_MAX_COLS]
    uint64_t *h_col_draw = nullptr, *d_col_draw = nullptr;          // [BATCH_MAX_COLS]
    // tensor parallelism: NCCL communicator (b2_allreduce_init)
    void* tp_comm = nullptr;
    // ... and the exchange inside the persistent kernel: this rank's inbox + flags, every rank's mapping of them, launch counter
    void* tp_local = nullptr;
    void* tp_peers[8] = {};
    size_t tp_flags_off = 0;
    unsigned int* tp_epoch = nullptr;
    // batched path: up to BATCH_MAX_COLS tokens per pass over the weights, each with its own (sequence slot,
    // position) -- up to GEMM_MAX_COLS columns on the int8 kernel (gemm.cu), wider batches (and any batch of at least
    // tc_batch_min() columns when the model's fp16 copy is resident) on the tensor-core pass (tc_gemm.cu).  KV caches
    // of slot s start s * slot_stride halfs after slot 0's.
    int n_slots = 1;
    int64_t slot_stride = 0;
    // Paged KV: the cache is one pool of kv_tokens rows per (layer, kv head), [layer][kv head][kv_tokens][head_dim], cut
    // into pages of KV_PAGE positions.  kv_table[slot][i] (host mirror + device copy) is the page holding positions
    // [i * KV_PAGE, (i + 1) * KV_PAGE) of the slot, or -1.  Slot 0 owns pages 0 .. pages_per_seq - 1 for good (identity),
    // so the single-sequence kernels address it as a contiguous cache with row stride kv_tokens.  With a pool large
    // enough for every slot at full length (the default when it fits) every slot is mapped like that up front
    // (kv_dynamic = false, slot_stride = pages_per_seq * KV_PAGE rows); a smaller pool (B2_KV_POOL_TOKENS, or the
    // HBM cap) maps pages on demand and releases them with b2_ctx_slot_truncate.
    int kv_tokens = 0, pages_per_seq = 0, n_pages = 0;
    bool kv_dynamic = false;
    std::vector<int32_t> kv_table;     // [n_slots][pages_per_seq]
    std::vector<int32_t> kv_free;      // unmapped pages (stack)
    int32_t* d_kv_table = nullptr;
    bool kv_table_dirty = false;
    struct Batch {
        float *x = nullptr, *q = nullptr, *attn_out = nullptr, *h = nullptr, *logits = nullptr;
        uint8_t* act = nullptr;                                   // quantised activations of the current mat-mul input
        int32_t *tok = nullptr, *pos = nullptr, *slot = nullptr, *out = nullptr;   // [cap] device
        void *attn_scratch = nullptr, *argmax_scratch = nullptr;
        int cap = 0;
        // one CUDA graph per (columns, logit columns): ~310 launches per pass replayed as one; an entry is
        // captured the second time its shape is used (the first use runs eagerly)
        struct G { cudaGraphExec_t exec = nullptr; bool seen = false; };
        std::vector<G> graphs;   // index = (path * (cap + 1) + n) * (cap + 1) + n_logits  (path 1: the fused-dequantisation mat-mul)
    } bt;
    // prompt-phase path (tc_gemm.cu, prefill_attn.cu): whole chunks of a prompt per pass over fp16 copies of the
    // weights, allocated on the first long prompt
    struct Prefill {
        int cap = 0;                       // rows (prompt tokens per chunk)
        float *x = nullptr, *qkv = nullptr, *q = nullptr, *attn_out = nullptr, *gu = nullptr;
        void *xb = nullptr, *wb = nullptr; // fp16: activations [cap][max K], one dequantised matrix
        int32_t* tok = nullptr;            // [cap] device
        int32_t *col_pos = nullptr, *col_slot = nullptr;   // [cap] device: per-row position / slot of a several-prompt pass
        float* ws = nullptr;               // split-K workspace of the mat-muls (passes of <= 128 rows)
        size_t ws_bytes = 0;
    } pf;
};
