/*
 * b2.h -- C ABI of the B200-native GGUF inference path (libb2.so).
 *
 * The reference (graz-dev/llms-on-kubernetes) has no in-process plugin / FFI for this path: its
 * boundary is the process `llama-server --host --port --model --alias`
 * (ramalama-models/helm-chart/templates/model-deployments.yaml:26-35) speaking OpenAI-style HTTP to
 * the API gateway (ramalama-models/helm-chart/templates/api-gateway.yaml:63-104).  This header is the
 * "thin C-ABI layer" BASELINE.json's north_star asks for underneath that process: the llama-server
 * binary in this repo (llms-on-kubernetes_b200/csrc/server) and the pytest/ctypes harness both call
 * exactly these symbols.  Each entry point cites the stage of the reference's hot path it replaces
 * (the reference-side implementation of rows a8-a16 is the un-vendored
 * llama-server launched at model-deployments.yaml:26-35).
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every function returns 0 on success and
 * a negative B2_E* code on failure, with b2_last_error() giving a thread-local message; the caller
 * owns host buffers, the library owns device buffers it allocates; every kernel-launching entry point
 * takes an explicit cudaStream_t (passed as void*; NULL = the legacy default stream); no exceptions
 * cross the boundary; there is NO CPU fallback -- without a CUDA device the compute entry points fail
 * with B2_ECUDA.
 */
#ifndef B2_H
#define B2_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_ABI_VERSION 1

enum {
    B2_OK = 0,
    B2_EINVAL = -1,   /* bad argument (shape, type, NULL)            */
    B2_EIO = -2,      /* file cannot be opened / read                */
    B2_EFORMAT = -3,  /* not a GGUF file, or malformed               */
    B2_EUNSUP = -4,   /* valid GGUF, but arch / tensor type unsupported */
    B2_ECUDA = -5,    /* CUDA runtime error (no device, OOM, launch) */
    B2_ERANGE = -6,   /* position / token / size outside limits      */
    B2_ENOTFOUND = -7,
    B2_ENOMEM = -8    /* the KV pool has no free page (b2_ctx_slot_truncate)  */
};

/* ggml tensor type ids as stored in GGUF (gguf-py constants.py:4059-4093) */
enum {
    B2_TYPE_F32 = 0, B2_TYPE_F16 = 1, B2_TYPE_Q8_0 = 8, B2_TYPE_Q4_K = 12,
    B2_TYPE_Q5_K = 13, B2_TYPE_Q6_K = 14, B2_TYPE_BF16 = 30
};

/* GGUF metadata value types (gguf-py constants.py:4157-4170) */
enum {
    B2_KV_U8 = 0, B2_KV_I8, B2_KV_U16, B2_KV_I16, B2_KV_U32, B2_KV_I32, B2_KV_F32, B2_KV_BOOL,
    B2_KV_STRING, B2_KV_ARRAY, B2_KV_U64, B2_KV_I64, B2_KV_F64
};

typedef struct b2_gguf b2_gguf;    /* a parsed GGUF file (host side, mmap'd)          */
typedef struct b2_model b2_model;  /* weights resident in HBM, raw blocks as on disk  */
typedef struct b2_ctx b2_ctx;      /* one decode context: KV cache, scratch, CUDA graph */

const char* b2_last_error(void);
int b2_abi_version(void);
/* number of CUDA devices visible, or a negative error (used by the server to fail loudly) */
int b2_device_count(void);

/* ---------------------------------------------------------------------------------------------
 * GGUF container  ("GGUF load"; layout = gguf-py gguf_reader.py:132-185,259-287)
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    const char* name;      /* NUL-terminated copy owned by the b2_gguf */
    uint32_t n_dims;
    uint64_t dims[4];      /* fastest-varying first, as on disk */
    uint32_t type;         /* ggml type id */
    uint64_t offset;       /* relative to the data section */
    uint64_t n_bytes;
} b2_tensor_info;

int b2_gguf_open(const char* path, b2_gguf** out);
void b2_gguf_close(b2_gguf* g);
uint32_t b2_gguf_version(const b2_gguf* g);
uint32_t b2_gguf_alignment(const b2_gguf* g);
uint64_t b2_gguf_data_offset(const b2_gguf* g);      /* absolute file offset of the tensor data */
int64_t b2_gguf_n_kv(const b2_gguf* g);
int64_t b2_gguf_n_tensors(const b2_gguf* g);
const char* b2_gguf_kv_key(const b2_gguf* g, int64_t i);
int b2_gguf_kv_type(const b2_gguf* g, int64_t i);        /* B2_KV_* */
int b2_gguf_kv_array_type(const b2_gguf* g, int64_t i);  /* element type when B2_KV_ARRAY */
int64_t b2_gguf_kv_array_len(const b2_gguf* g, int64_t i);
int64_t b2_gguf_find_kv(const b2_gguf* g, const char* key);  /* index or B2_ENOTFOUND */
/* scalar accessors convert between integer widths / float widths; strings return a pointer into
 * the mapping plus a length (not NUL-terminated) */
int b2_gguf_kv_u64(const b2_gguf* g, int64_t i, uint64_t* out);
int b2_gguf_kv_i64(const b2_gguf* g, int64_t i, int64_t* out);
int b2_gguf_kv_f64(const b2_gguf* g, int64_t i, double* out);
int b2_gguf_kv_str(const b2_gguf* g, int64_t i, const char** s, uint64_t* len);
/* array element accessors (element j of array-valued KV i) */
int b2_gguf_kv_arr_str(const b2_gguf* g, int64_t i, int64_t j, const char** s, uint64_t* len);
int b2_gguf_kv_arr_f64(const b2_gguf* g, int64_t i, int64_t j, double* out);
int b2_gguf_kv_arr_i64(const b2_gguf* g, int64_t i, int64_t j, int64_t* out);
int b2_gguf_tensor_info(const b2_gguf* g, int64_t i, b2_tensor_info* out);
int64_t b2_gguf_find_tensor(const b2_gguf* g, const char* name);
/* host pointer to the raw bytes of tensor i inside the mapping */
const void* b2_gguf_tensor_data(const b2_gguf* g, int64_t i);

/* ---------------------------------------------------------------------------------------------
 * Device memory helpers for FFI callers without a CUDA binding of their own
 * ------------------------------------------------------------------------------------------- */
int b2_dev_malloc(void** out, size_t bytes);
int b2_dev_free(void* p);
int b2_dev_memset(void* p, int value, size_t bytes, void* stream);
int b2_memcpy_h2d(void* dst_dev, const void* src_host, size_t bytes, void* stream);
int b2_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes, void* stream);
int b2_stream_sync(void* stream);

/* ---------------------------------------------------------------------------------------------
 * Block-level operators on DEVICE pointers
 * ------------------------------------------------------------------------------------------- */
/* bytes of one row of k elements of `type`, or negative if k is not a whole number of blocks */
int64_t b2_row_bytes(int type, int64_t k);

/* Dequantise n elements to fp32, bit-exact with ggml's dequantize_row_* (SURVEY 8a rows a8/a10;
 * block math gguf-py quants.py:395-401,475-522,525-549,552-572).  src_dev needs 2-byte alignment only. */
int b2_dequantize(int type, const void* src_dev, float* dst_dev, int64_t n, void* stream);
/* same, host buffers in and out (upload, kernel, download) */
int b2_dequantize_host(int type, const void* src_host, float* dst_host, int64_t n);

/* y[N] = W[N x K] . x[K]   ("quantised matmul"; ggml_compute_forward_mul_mat with
 * n_tokens = 1).  W: N rows of K elements in `type` blocks, exactly as in the GGUF.  x, y fp32.  The
 * activation row is quantised on the fly to Q8_K (K-quants) or Q8_0 (Q8_0 weights) and the dot product
 * is integer, as on ggml's CPU path.  B > 1 runs B independent columns: x is [B][K], y is [B][N]. */
int b2_gemv(int type, const void* W_dev, const float* x_dev, float* y_dev, int64_t N, int64_t K, int B, void* stream);

/* Y[B][N] = X[B][K] . W[N x K]^T for 1 <= B <= 32 columns in ONE pass over the weights (same row a12 with
 * n_tokens = B: batched decode and chunked prefill).  Same arithmetic as b2_gemv -- Q8_K activations, integer
 * sub-block dot products, fp32 scales -- with the dot products on the tensor cores (mma.sync int8).
 * Q4_K, Q5_K, Q6_K and Q8_0 weights (Q8_0 with Q8_0 activations, as b2_gemv).  scratch_dev: b2_gemm_scratch_bytes(K, B)
 * bytes of device memory, 16-byte aligned, for the quantised activations. */
size_t b2_gemm_scratch_bytes(int64_t K, int B);
int b2_gemm(int type, const void* W_dev, const float* x_dev, float* y_dev, int64_t N, int64_t K, int B, void* scratch_dev, void* stream);

/* Y[T][N] (+)= X[T][K] . W[N x K]^T for a whole prompt chunk of T tokens on the 5th-generation tensor cores (row a12
 * with n_tokens = T, ggml_compute_forward_mul_mat's prompt path; BASELINE north_star "dequant tile staging through
 * TMA/shared memory into tcgen05").  The blocks are dequantised exactly as ggml does and rounded once to fp16; the
 * activations are quantised as the reference quantises them for this weight type (Q8_K / Q8_0 blocks) and d * q is
 * rounded to fp16; products accumulate in fp32 in tensor memory.  accumulate != 0 adds into y (the residual form).
 * K a multiple of 64; scratch_dev: b2_tc_gemm_scratch_bytes(N, K, T) bytes, 256-byte aligned (the fp16 copies of W
 * and X). */
size_t b2_tc_gemm_scratch_bytes(int64_t N, int64_t K, int T);
int b2_tc_gemm(int type, const void* W_dev, const float* x_dev, float* y_dev, int64_t N, int64_t K, int T, int accumulate, void* scratch_dev,
               void* stream);

/* The same product for a decode batch (T <= 128 rows) with the dequantisation FUSED into the kernel's producer: the
 * quantised blocks go HBM -> TMA -> shared memory as they are, dequantising warps write the fp16 tiles tcgen05.mma
 * reads, so W crosses HBM once per call in its 4.5 / 6.56-bit form and no fp16 copy of it exists (row a12 with
 * 2 <= n_tokens <= 128; north_star "dequant fused into the TMA -> smem -> tcgen05 producer").  Same operand values as
 * b2_tc_gemm, so the results differ only by fp32 summation order.  Q4_K with K a multiple of 256, Q6_K with K a multiple
 * of 2048 (rows of 210-byte blocks then keep the 16-byte stride a tensor map needs), N a multiple of 4; B2_EUNSUP otherwise.  scratch_dev: b2_tcq_gemm_scratch_bytes(N, K, T) bytes, 256-byte aligned. */
size_t b2_tcq_gemm_scratch_bytes(int64_t N, int64_t K, int T);
/* (row a12, host side) How b2_tcq_gemm cuts a stack of n_seg tensors (seg_rows[i] rows each, K columns) over n_sms CTAs: host arithmetic, no GPU
 * needed.  The (128-row tile, 256-column block) pairs are numbered tile-major and every CTA takes an equal contiguous
 * share; a tile that several CTAs touch is summed from that many slices.  *tiles_out, *grid_out (CTAs), *slices_out (the
 * most CTAs any tile meets), shares_out[2 c], [2 c + 1] = CTA c's first step and end (room for 2 n_sms values),
 * first_cta_out[t] = the first CTA that touches tile t (room for *tiles_out values); any output may be NULL. */
int b2_tcq_plan(const int32_t* seg_rows, int n_seg, int64_t K, int n_sms, int32_t* tiles_out, int32_t* grid_out, int32_t* slices_out,
                int64_t* shares_out, int32_t* first_cta_out);
int b2_tcq_gemm(int type, const void* W_dev, const float* x_dev, float* y_dev, int64_t N, int64_t K, int T, int accumulate, void* scratch_dev,
                void* stream);

/* y = x * rsqrt(mean(x^2) + eps) * w  over rows of length n   (row a11, ggml rms_norm + mul) */
int b2_rmsnorm(const float* x_dev, const float* w_dev, float* y_dev, int64_t n, int64_t rows, float eps, void* stream);

/* Rotary embedding in place on x[n_heads][head_dim] for one position (row a13, ggml rope).
 * mode 0 = adjacent pairs (`llama`), 2 = NEOX pairs (`qwen2`).  freq_factors_dev may be NULL. */
int b2_rope(float* x_dev, int n_heads, int head_dim, int n_dims, int mode, int pos, float freq_base,
            const float* freq_factors_dev, void* stream);

/* Single-query attention over an fp16 KV cache (row a14).  q: [n_head][head_dim] fp32;
 * k_cache, v_cache: [n_kv][n_ctx][head_dim] fp16; attends positions [0, n_tokens).  out: [n_head][head_dim].
 * scratch_dev: b2_attn_scratch_bytes(n_head, head_dim, n_tokens) bytes of device memory, zeroed once by the caller. */
size_t b2_attn_scratch_bytes(int n_head, int head_dim, int n_tokens);
int b2_attn_decode(const float* q_dev, const void* k_cache_dev, const void* v_cache_dev, float* out_dev,
                   int n_head, int n_kv, int head_dim, int n_ctx, int n_tokens, void* scratch_dev, void* stream);

/* The same attention for callers that cannot allocate device memory (code inside a CUDA-graph capture or a
 * latency-critical loop where cudaMalloc would synchronise the device, plugins running under a fixed memory
 * budget, hosts whose own allocator owns all of HBM).  No counterpart in the reference: its scratch is a
 * fixed array inside libb2.so (about 133 KiB on every GPU that loads the library), so there is no
 * scratch argument.  Limits: n_head <= 64, head_dim <= 128, n_tokens <= 4096 -- larger requests fail with
 * B2_ERANGE; use b2_attn_decode for those (n_ctx, the row stride of the caches, has no limit of its own beyond
 * n_tokens <= n_ctx; sizes that are zero, negative or otherwise invalid fail with B2_EINVAL before the array is
 * touched).  The first call on each device looks up the array's device
 * address; make that first call outside any stream capture.  All calls on a device share the array: calls
 * whose kernels overlap on the GPU race.  Calls no cudaMalloc / cudaFree. */
int b2_attn_decode_static(const float* q_dev, const void* k_cache_dev, const void* v_cache_dev, float* out_dev,
                          int n_head, int n_kv, int head_dim, int n_ctx, int n_tokens, void* stream);

/* argmax over n logits, lowest index wins ties (row a16 greedy).  scratch_dev: b2_argmax_scratch_bytes(n),
 * zeroed once by the caller. */
size_t b2_argmax_scratch_bytes(int64_t n);
int b2_sample_greedy(const float* logits_dev, int64_t n, int32_t* token_out_dev, void* scratch_dev, void* stream);

/* Sampling parameters as llama-server applies them (row a16; the server's defaults when a request omits them are
 * temperature 0.8, top_k 40, top_p 0.95, min_p 0.05).  temperature <= 0 or top_k == 1 is greedy. */
typedef struct b2_sampling {
    float temperature, top_p, min_p;
    int32_t top_k;     /* <= 0: no top-k limit of its own (the sampler considers the B2_SAMPLE_MAX_K most probable tokens) */
    uint64_t seed;     /* one Philox4x32-10 draw per (seed, draw index): the same pair always picks the same token */
} b2_sampling;
#define B2_SAMPLE_MAX_K 1024
/* Device-side sampling of n_rows logits rows (row r at logits_dev + r * row_stride floats; row_stride 0 = the same row
 * for every draw): top-k -> softmax with temperature -> min-p -> top-p -> one draw, in llama-server's order.  Row r
 * uses params_dev[r * param_stride] (param_stride 0 = one parameter set for all rows) and draw index draw_dev[r]
 * (NULL: r).  Replaces the host sampler of the reference's llama-server: the logits never leave the GPU. */
int b2_sample(const float* logits_dev, int n_rows, int64_t vocab, int64_t row_stride, const b2_sampling* params_dev,
              int64_t param_stride, const uint64_t* draw_dev, int32_t* tokens_out_dev, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Model + decode context  (rows a8-a16 end to end; replaces llama_model_load / llama_decode inside
 * the llama-server the chart launches)
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    int32_t arch;       /* 0 = llama, 1 = qwen2 */
    int32_t n_layer, d_model, d_ff, n_head, n_kv, head_dim, vocab, n_ctx_train, rope_dim;
    float eps, rope_base;
    int32_t bos_id, eos_id, eot_id;
    int32_t tied;           /* output.weight absent -> token_embd used as lm_head */
    uint64_t weight_bytes;  /* bytes resident in HBM */
    uint64_t stream_bytes;  /* bytes one decode step reads (SURVEY 8d W_stream) */
} b2_model_info;

/* Upload the GGUF's tensor data to `device` unmodified.  tp_rank/tp_size: 0/1 for a full replica.
 * tp_size > 1 (the 70B configuration: one process per GPU): this process keeps shard tp_rank -- attn_q/k/v
 * by heads and ffn_gate/up by ff rows (row ranges), attn_output and ffn_down along K on super-block boundaries; norms,
 * embeddings and the lm_head stay whole.  b2_model_get_info then reports the shard's n_head / n_kv / d_ff and
 * stream_bytes.  A context of a sharded model decodes one sequence, one kernel per stage with an NCCL all-reduce after
 * each row-split mat-vec, and must be given the communicator before the first decode: rank 0 calls
 * b2_allreduce_unique_id (128 bytes), the launcher hands the same bytes to every rank, every rank calls
 * b2_allreduce_init(ctx, id, 128) (collective).  libnccl.so.2 is loaded at that point, not linked. */
int b2_model_load(const char* path, int device, int tp_rank, int tp_size, b2_model** out);
int b2_allreduce_unique_id(void* id_out, size_t cap);   /* -> bytes written (128), or a negative status */
int b2_allreduce_init(b2_ctx* ctx, const void* id, size_t id_bytes);
void b2_model_free(b2_model* m);
int b2_model_get_info(const b2_model* m, b2_model_info* out);
const b2_gguf* b2_model_gguf(const b2_model* m);
/* device pointer + info of a named tensor (for tests that check the upload is byte-identical) */
int b2_model_tensor(const b2_model* m, const char* name, const void** dev_ptr, b2_tensor_info* info);

/* n_ctx: KV-cache capacity in tokens (0 = min(model training context, 8192)).
 * One DECODING context per GPU at a time: the per-token kernel is persistent (a grid barrier between its stages), all
 * of its CTAs must be resident together (two per SM).  Another context decoding concurrently on the same GPU -- a second
 * process under MPS, or an overlapping stream of this process -- can keep CTAs off the SMs; every wait in the kernel is
 * bounded, so the call then fails with B2_ECUDA instead of hanging, but the step is lost.  Contexts whose calls do not
 * overlap may share a GPU. */
int b2_ctx_create(b2_model* m, int n_ctx, b2_ctx** out);
void b2_ctx_free(b2_ctx* c);
int b2_ctx_n_ctx(const b2_ctx* c);
/* forget the sequence (positions restart at 0) */
int b2_ctx_reset(b2_ctx* c);

/* Feed n tokens at positions pos0 .. pos0+n-1 (host array).  If logits_out_host != NULL it receives the
 * vocab-sized fp32 logits after the LAST token.  If next_token_out != NULL it receives the greedy
 * continuation.  Host buffers; the H2D / D2H copies are part of the call. */
int b2_ctx_decode(b2_ctx* c, const int32_t* tokens, int n, int pos0, float* logits_out_host, int32_t* next_token_out);

/* Sampling for sequence slot `slot` (p == NULL or a greedy parameter set: back to greedy).  While set, every call that
 * returns a "next token" for that slot -- b2_ctx_decode, b2_ctx_generate_greedy (slot 0), b2_ctx_decode_batch,
 * b2_ctx_prefill, b2_ctx_prefill_multi -- returns a token drawn on the device with these parameters; the draw index is
 * the position the new token will take, so a (seed, prompt) pair reproduces its continuation. */
int b2_ctx_set_sampler(b2_ctx* c, int slot, const b2_sampling* p);

/* A context with n_slots independent sequences (KV caches of n_ctx tokens each); b2_ctx_create = 1 slot.
 * The single-sequence calls above use slot 0. */
int b2_ctx_create_slots(b2_model* m, int n_ctx, int n_slots, b2_ctx** out);
int b2_ctx_n_slots(const b2_ctx* c);

/* Widest batch b2_ctx_decode_batch accepts for this context: 32 columns on the int8 tensor-core kernel; 256 when the
 * model's fp16 copy is resident (built on the first call that needs it, budget B2_F16_CACHE_GB, default 64 GB and at
 * most half of the free HBM), where batches of 16 and more columns run one HBM sweep of fp16 tiles through
 * tcgen05.mma.  0 when the model has matrices the batched kernels do not compute. */
int b2_ctx_max_batch(b2_ctx* ctx);

/* The KV cache of a context with several slots is one pool of pages of 256 positions (paged KV).
 * When the pool holds every slot at full length -- the default while n_slots x n_ctx fits 60 % of the free HBM -- all
 * pages are mapped up front and these calls are no-ops.  A smaller pool (that cap, or B2_KV_POOL_TOKENS) maps a page
 * when a batch or prompt first touches it (B2_ENOMEM when none is free) and
 * b2_ctx_slot_truncate(ctx, slot, n_keep) gives back the pages of `slot` beyond its first n_keep positions (0: all) --
 * call it when a sequence ends or is cut back to a shared prefix.  Slot 0 (the single-sequence calls' cache) is never
 * paged out.  b2_ctx_kv_pages reports the pool (returns 1 when pages are mapped on demand, 0 when statically). */
int b2_ctx_slot_truncate(b2_ctx* ctx, int slot, int n_keep);
int b2_ctx_kv_pages(const b2_ctx* ctx, int* total_out, int* free_out, int* page_tokens_out);

/* One pass over the weights for a batch of 1 <= n <= 32 tokens (llama.cpp's llama_batch: row a12 with
 * n_tokens = n): column j is token tokens[j] at position pos[j] of sequence slot slots[j] (slots == NULL: all
 * slot 0).  Batched decode = one token of each of n sequences; chunked prefill = n consecutive positions of one
 * sequence (a column attends to the columns before it).  K/V of every column are written to its slot's cache.
 * last_only = 0: next_tokens_out[n] / logits_out_host[n][vocab] receive every column's greedy token / logits;
 * last_only = 1: only the last column's (one entry).  Either output may be NULL. */
int b2_ctx_decode_batch(b2_ctx* c, const int32_t* tokens, const int32_t* slots, const int32_t* pos, int n, int last_only,
                        int32_t* next_tokens_out, float* logits_out_host);
/* fp32 logits [vocab] of output column col of the last b2_ctx_decode_batch (for callers that sample only some
 * columns on the host instead of copying all of them) */
int b2_ctx_batch_logits(b2_ctx* c, int col, float* logits_out_host);

/* A prompt (or the rest of one after a reused prefix) of ONE sequence: n tokens at positions pos0 .. pos0 + n - 1 of
 * sequence slot `slot`, processed in whole chunks on the tensor-core prompt path (b2_tc_gemm's kernel per matrix,
 * causal attention over the chunk; llama_decode with a single-sequence batch of n tokens).  Fills the slot's KV
 * cache; next_token_out / logits_out_host (either may be NULL) receive the greedy token and the logits of the LAST
 * position.  B2_EUNSUP when the model's shapes or types are outside the path (use b2_ctx_decode_batch). */
int b2_ctx_prefill(b2_ctx* ctx, int slot, const int32_t* tokens, int n, int pos0, int32_t* next_token_out, float* logits_out_host);

/* The prompts (or prompt tails) of several sequences in ONE pass over the weights (llama_decode with a batch holding
 * several sequences' prompt tokens): segment i is lens[i] tokens at positions pos0[i] .. of slot slots[i] (distinct
 * slots), the token ids of all segments concatenated in `tokens`; at most 2048 rows (B2_PREFILL_CHUNK) and 256
 * segments per call.  next_tokens_out[n_seq] / logits_out_host[n_seq][vocab] (either may be NULL): greedy token and
 * logits of every segment's LAST position.  Same kernels as b2_ctx_prefill, causal attention per segment. */
int b2_ctx_prefill_multi(b2_ctx* ctx, int n_seq, const int32_t* slots, const int32_t* pos0, const int32_t* lens, const int32_t* tokens,
                         int32_t* next_tokens_out, float* logits_out_host);

/* Greedy generation entirely on the device: starting from the state left by b2_ctx_decode (whose greedy
 * token t0 the caller already has), run n_steps decode steps, each consuming the previous argmax without a
 * host round trip, then copy the n_steps newly produced token ids t1..tn to tokens_out_host. */
int b2_ctx_generate_greedy(b2_ctx* c, int n_steps, int32_t* tokens_out_host);

/* Same loop without the final copy or synchronisation, for device-side timing: enqueue n_steps steps on
 * the context's stream.  b2_ctx_stream() returns that cudaStream_t so callers can record events on it. */
int b2_ctx_enqueue_steps(b2_ctx* c, int n_steps);
void* b2_ctx_stream(b2_ctx* c);
int b2_ctx_sync(b2_ctx* c);
/* Read back state for parity tests.  Activation buffers as the last decode step left them: the residual
 * stream after the last layer, and the last layer's rotated query, attention output and SwiGLU output.
 * Returns the number of floats written (cap = capacity of out_host in floats). */
enum { B2_ACT_HIDDEN = 0, B2_ACT_Q = 1, B2_ACT_ATTN_OUT = 2, B2_ACT_FFN_H = 3, B2_ACT_LOGITS = 4 };
int b2_ctx_get_activation(b2_ctx* c, int which, float* out_host, int64_t cap);
/* fp16 K and V rows [n_kv * head_dim] of one layer at one position */
int b2_ctx_get_kv(b2_ctx* c, int layer, int pos, uint16_t* k_out_host, uint16_t* v_out_host);
int b2_ctx_pos(const b2_ctx* c);
/* kernels launched per decode step (for bench.py's gpu_launches) */
int b2_ctx_launches_per_step(const b2_ctx* c);

/* ---------------------------------------------------------------------------------------------
 * Measurement helpers (no counterpart on the reference's request path; the reference's own tool for
 * this is llama-bench).  Both time with CUDA events recorded on the context's stream.
 * ------------------------------------------------------------------------------------------- */
/* n_steps greedy decode steps (as b2_ctx_enqueue_steps) -> elapsed milliseconds */
int b2_ctx_time_steps(b2_ctx* c, int n_steps, float* ms_out);
/* only the mat-vec launches of a decode step, `reps` times back to back: elapsed milliseconds, the weight
 * bytes those launches read per repetition and their number.  Leaves the context reset to position 0. */
int b2_ctx_time_matvec_only(b2_ctx* c, int reps, float* ms_out, uint64_t* bytes_per_rep_out, int* launches_per_rep_out);

#ifdef __cplusplus
}
#endif
#endif /* B2_H */
