// capi.cu -- the extern "C" surface of include/b2.h that is not in model.cu: error slot, GGUF
// accessors, device-memory helpers and the stand-alone operators.
#include <cuda_runtime.h>

#include <cstring>

#include "gemm.cuh"
#include "gemv.cuh"
#include "gguf.h"
#include "ops.h"
#include "tc_gemm.cuh"
#include "tcq_gemm.cuh"

using namespace b2;

extern "C" {

const char* b2_last_error(void) { return get_error(); }
int b2_abi_version(void) { return B2_ABI_VERSION; }

int b2_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("no usable CUDA device: %s", cudaGetErrorString(e));
        return B2_ECUDA;
    }
    return n;
}

// ---- GGUF ---------------------------------------------------------------------------------------
int b2_gguf_open(const char* path, b2_gguf** out) {
    if (!path || !out) {
        set_error("b2_gguf_open: NULL argument");
        return B2_EINVAL;
    }
    *out = nullptr;
    Gguf* g = nullptr;
    int rc = Gguf::open(path, &g);
    if (rc) return rc;
    b2_gguf* h = new b2_gguf();
    h->g = g;
    h->owned = true;
    *out = h;
    return 0;
}
void b2_gguf_close(b2_gguf* h) {
    if (!h) return;
    if (h->owned) {
        delete h->g;
        delete h;
    }
}
uint32_t b2_gguf_version(const b2_gguf* h) { return h->g->version; }
uint32_t b2_gguf_alignment(const b2_gguf* h) { return h->g->alignment; }
uint64_t b2_gguf_data_offset(const b2_gguf* h) { return h->g->data_offset; }
int64_t b2_gguf_n_kv(const b2_gguf* h) { return (int64_t)h->g->kvs.size(); }
int64_t b2_gguf_n_tensors(const b2_gguf* h) { return (int64_t)h->g->tensors.size(); }
static bool kv_ok(const b2_gguf* h, int64_t i) { return h && i >= 0 && i < (int64_t)h->g->kvs.size(); }
const char* b2_gguf_kv_key(const b2_gguf* h, int64_t i) { return kv_ok(h, i) ? h->g->kvs[i].key.c_str() : nullptr; }
int b2_gguf_kv_type(const b2_gguf* h, int64_t i) { return kv_ok(h, i) ? (int)h->g->kvs[i].type : B2_EINVAL; }
int b2_gguf_kv_array_type(const b2_gguf* h, int64_t i) {
    return kv_ok(h, i) && h->g->kvs[i].type == B2_KV_ARRAY ? (int)h->g->kvs[i].arr_type : B2_EINVAL;
}
int64_t b2_gguf_kv_array_len(const b2_gguf* h, int64_t i) {
    return kv_ok(h, i) && h->g->kvs[i].type == B2_KV_ARRAY ? (int64_t)h->g->kvs[i].arr_len : B2_EINVAL;
}
int64_t b2_gguf_find_kv(const b2_gguf* h, const char* key) {
    const int64_t i = h && key ? h->g->find_kv(key) : -1;
    return i < 0 ? B2_ENOTFOUND : i;
}
int b2_gguf_kv_u64(const b2_gguf* h, int64_t i, uint64_t* out) { return kv_ok(h, i) && out ? h->g->kv_u64(i, out) : B2_EINVAL; }
int b2_gguf_kv_i64(const b2_gguf* h, int64_t i, int64_t* out) { return kv_ok(h, i) && out ? h->g->kv_i64(i, out) : B2_EINVAL; }
int b2_gguf_kv_f64(const b2_gguf* h, int64_t i, double* out) { return kv_ok(h, i) && out ? h->g->kv_f64(i, out) : B2_EINVAL; }
int b2_gguf_kv_str(const b2_gguf* h, int64_t i, const char** s, uint64_t* len) {
    return kv_ok(h, i) && s && len ? h->g->kv_str(i, s, len) : B2_EINVAL;
}
int b2_gguf_kv_arr_str(const b2_gguf* h, int64_t i, int64_t j, const char** s, uint64_t* len) {
    return kv_ok(h, i) && s && len ? h->g->arr_str(i, j, s, len) : B2_EINVAL;
}
int b2_gguf_kv_arr_f64(const b2_gguf* h, int64_t i, int64_t j, double* out) { return kv_ok(h, i) && out ? h->g->arr_f64(i, j, out) : B2_EINVAL; }
int b2_gguf_kv_arr_i64(const b2_gguf* h, int64_t i, int64_t j, int64_t* out) { return kv_ok(h, i) && out ? h->g->arr_i64(i, j, out) : B2_EINVAL; }
int b2_gguf_tensor_info(const b2_gguf* h, int64_t i, b2_tensor_info* out) {
    if (!h || !out || i < 0 || i >= (int64_t)h->g->tensors.size()) return B2_EINVAL;
    const auto& t = h->g->tensors[i];
    out->name = t.name.c_str();
    out->n_dims = t.n_dims;
    for (int d = 0; d < 4; d++) out->dims[d] = t.dims[d];
    out->type = t.type;
    out->offset = t.offset;
    out->n_bytes = t.n_bytes;
    return 0;
}
int64_t b2_gguf_find_tensor(const b2_gguf* h, const char* name) {
    const int64_t i = h && name ? h->g->find_tensor(name) : -1;
    return i < 0 ? B2_ENOTFOUND : i;
}
const void* b2_gguf_tensor_data(const b2_gguf* h, int64_t i) {
    if (!h || i < 0 || i >= (int64_t)h->g->tensors.size()) return nullptr;
    return h->g->tensor_data(i);
}

// ---- device memory ------------------------------------------------------------------------------
int b2_dev_malloc(void** out, size_t bytes) {
    if (!out) return B2_EINVAL;
    B2_CUDA_CHECK(cudaMalloc(out, bytes ? bytes : 1));
    return 0;
}
int b2_dev_free(void* p) {
    B2_CUDA_CHECK(cudaFree(p));
    return 0;
}
int b2_dev_memset(void* p, int value, size_t bytes, void* stream) {
    B2_CUDA_CHECK(cudaMemsetAsync(p, value, bytes, (cudaStream_t)stream));
    return 0;
}
int b2_memcpy_h2d(void* dst, const void* src, size_t bytes, void* stream) {
    B2_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return 0;
}
int b2_memcpy_d2h(void* dst, const void* src, size_t bytes, void* stream) {
    B2_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return 0;
}
int b2_stream_sync(void* stream) {
    B2_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}

// ---- operators ----------------------------------------------------------------------------------
int64_t b2_row_bytes(int type, int64_t k) { return row_bytes(type, k); }

int b2_dequantize(int type, const void* src_dev, float* dst_dev, int64_t n, void* stream) {
    return launch_dequantize(type, src_dev, dst_dev, n, (cudaStream_t)stream);
}

int b2_dequantize_host(int type, const void* src_host, float* dst_host, int64_t n) {
    const TypeTraits* t = type_traits(type);
    if (!t || n < 0 || n % t->block_elems) {
        set_error("b2_dequantize_host: bad type %d or n=%lld", type, (long long)n);
        return B2_EINVAL;
    }
    if (n == 0) return 0;
    const size_t in_bytes = (size_t)(n / t->block_elems) * t->block_bytes;
    void* d_in = nullptr;
    float* d_out = nullptr;
    B2_CUDA_CHECK(cudaMalloc(&d_in, in_bytes));
    cudaError_t e = cudaMalloc((void**)&d_out, (size_t)n * 4);
    if (e != cudaSuccess) {
        cudaFree(d_in);
        B2_CUDA_CHECK(e);
    }
    int rc = 0;
    e = cudaMemcpy(d_in, src_host, in_bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) rc = launch_dequantize(type, d_in, d_out, n, nullptr);
    if (e == cudaSuccess && rc == 0) e = cudaMemcpy(dst_host, d_out, (size_t)n * 4, cudaMemcpyDeviceToHost);
    cudaFree(d_in);
    cudaFree(d_out);
    if (rc) return rc;
    B2_CUDA_CHECK(e);
    return 0;
}

int b2_gemv(int type, const void* W_dev, const float* x_dev, float* y_dev, int64_t N, int64_t K, int B, void* stream) {
    const int act = gemv_act_for_type(type);
    if (act < 0) {
        set_error("b2_gemv: ggml type %d is not a supported weight type (Q4_K, Q5_K, Q6_K, Q8_0)", type);
        return B2_EUNSUP;
    }
    if (!W_dev || !x_dev || !y_dev || N <= 0 || K <= 0 || B <= 0 || N > 0x3fffffff || K > 0x3fffffff) {
        set_error("b2_gemv: bad arguments (N=%lld K=%lld B=%d)", (long long)N, (long long)K, B);
        return B2_EINVAL;
    }
    for (int b = 0; b < B; b++) {
        GemvArgs a{};
        a.x = x_dev + (size_t)b * K;
        a.prologue = PRO_PLAIN;
        a.K = (int)K;
        GemvSeg& s = a.seg[0];
        s.wa = (const uint8_t*)W_dev;
        s.type = type;
        s.pairing = PAIR_ADJ;
        s.epi = EPI_STORE;
        s.n_rows = (int)N;
        s.n_pairs = (int)((N + 1) / 2);
        s.out = y_dev + (size_t)b * N;
        a.n_seg = 1;
        int rc = gemv_finalize(a, act);
        if (rc) return rc;
        if ((rc = launch_gemv(a, act, (cudaStream_t)stream, false))) return rc;
    }
    return 0;
}

size_t b2_gemm_scratch_bytes(int64_t K, int B) {
    if (K <= 0 || B < 1 || B > GEMM_MAX_COLS) return 0;
    const size_t a = gemm_act_bytes((int)K, B, ACT_Q8K), b = gemm_act_bytes((int)K, B, ACT_Q80);
    return a > b ? a : b;
}

int b2_gemm(int type, const void* W_dev, const float* x_dev, float* y_dev, int64_t N, int64_t K, int B, void* scratch_dev, void* stream) {
    if (!W_dev || !x_dev || !y_dev || !scratch_dev || N <= 0 || K <= 0 || B < 1 || B > GEMM_MAX_COLS || N > 0x3fffffff || K > 0x3fffffff) {
        set_error("b2_gemm: bad arguments (N=%lld K=%lld B=%d)", (long long)N, (long long)K, B);
        return B2_EINVAL;
    }
    const int fmt = gemm_act_for_type(type);
    if (fmt < 0) {
        set_error("b2_gemm: ggml type %d is not computed by the batched kernel (Q4_K, Q5_K, Q6_K, Q8_0); use b2_gemv", type);
        return B2_EUNSUP;
    }
    int rc = launch_act_quant(x_dev, K, nullptr, 0.0f, (int)K, B, (uint8_t*)scratch_dev, (cudaStream_t)stream, fmt);
    if (rc) return rc;
    GemmArgs a{};
    a.K = (int)K;
    a.act = (const uint8_t*)scratch_dev;
    a.cols.n = B;
    a.out_stride = N;
    GemvSeg& s = a.seg[0];
    s.wa = (const uint8_t*)W_dev;
    s.type = type;
    s.pairing = PAIR_ADJ;
    s.epi = EPI_STORE;
    s.n_rows = (int)N;
    s.n_pairs = (int)((N + 1) / 2);
    s.out = y_dev;
    a.n_seg = 1;
    if ((rc = gemm_finalize(a))) return rc;
    return launch_gemm(a, (cudaStream_t)stream);
}

size_t b2_tc_gemm_scratch_bytes(int64_t N, int64_t K, int T) {
    if (N <= 0 || K <= 0 || T <= 0) return 0;
    return ((size_t)N * K * 2 + 255) / 256 * 256 + (size_t)T * K * 2;
}

int b2_tc_gemm(int type, const void* W_dev, const float* x_dev, float* y_dev, int64_t N, int64_t K, int T, int accumulate, void* scratch_dev,
               void* stream) {
    if (!W_dev || !x_dev || !y_dev || !scratch_dev || N <= 0 || K <= 0 || T <= 0 || N > 0x3fffffff || K > 0x3fffffff || K % TC_BK ||
        ((uintptr_t)scratch_dev & 255)) {
        set_error("b2_tc_gemm: bad arguments (N=%lld K=%lld T=%d; K must be a multiple of %d, scratch 256-byte aligned)", (long long)N, (long long)K, T,
                  TC_BK);
        return B2_EINVAL;
    }
    __half* wb = (__half*)scratch_dev;
    __half* xb = (__half*)((uint8_t*)scratch_dev + ((size_t)N * K * 2 + 255) / 256 * 256);
    int rc;
    if ((rc = launch_dequant_f16(type, W_dev, wb, N * K, (cudaStream_t)stream))) return rc;
    // activations quantised as the reference's mat-mul with this weight type quantises them (Q8_K / Q8_0), then fp16
    int fmt = gemm_act_for_type(type);
    if ((fmt == ACT_Q8K && K % 256) || (fmt == ACT_Q80 && K % 32)) fmt = -1;
    if ((rc = launch_rows_to_f16(const_cast<float*>(x_dev), K, nullptr, 0.0f, (int)K, T, fmt, xb, (cudaStream_t)stream))) return rc;
    TcGemmArgs a{};
    a.w = wb;
    a.x = xb;
    a.out = y_dev;
    a.N = (int)N;
    a.K = (int)K;
    a.T = T;
    a.ldo = N;
    a.epi = accumulate ? TC_EPI_ADD : TC_EPI_STORE;
    return launch_tc_gemm(a, (cudaStream_t)stream);
}

int b2_tcq_plan(const int32_t* seg_rows, int n_seg, int64_t K, int n_sms, int32_t* tiles_out, int32_t* grid_out, int32_t* slices_out,
                int64_t* shares_out, int32_t* first_cta_out) {
    if (K <= 0 || K > 0x3fffffff) return B2_EINVAL;
    int tiles = 0, grid = 0, slices = 0;
    const int rc = tcq_plan(seg_rows, n_seg, (int)K, n_sms, &tiles, &grid, &slices);
    if (rc) {
        set_error("b2_tcq_plan: bad arguments");
        return rc;
    }
    if (tiles_out) *tiles_out = tiles;
    if (grid_out) *grid_out = grid;
    if (slices_out) *slices_out = slices;
    const int nblk = (int)(K / 256);
    if (shares_out)
        for (int c = 0; c < grid; c++) {
            long long lo, hi;
            tcq_share(tiles, nblk, grid, c, &lo, &hi);
            shares_out[2 * c] = lo;
            shares_out[2 * c + 1] = hi;
        }
    if (first_cta_out)
        for (int t = 0; t < tiles; t++) first_cta_out[t] = tcq_first_cta_of_tile(tiles, nblk, grid, t);
    return 0;
}

size_t b2_tcq_gemm_scratch_bytes(int64_t N, int64_t K, int T) {
    if (N <= 0 || K <= 0 || T <= 0) return 0;
    return ((size_t)T * K * 2 + 255) / 256 * 256 + tcq_ws_bytes((int)N, (int)K, T);
}

int b2_tcq_gemm(int type, const void* W_dev, const float* x_dev, float* y_dev, int64_t N, int64_t K, int T, int accumulate, void* scratch_dev,
                void* stream) {
    if (!W_dev || !x_dev || !y_dev || !scratch_dev || N <= 0 || K <= 0 || T <= 0 || N > 0x3fffffff || K > 0x3fffffff || N % 4 ||
        ((uintptr_t)scratch_dev & 255)) {
        set_error("b2_tcq_gemm: bad arguments (N=%lld K=%lld T=%d; N a multiple of 4, scratch 256-byte aligned)", (long long)N, (long long)K, T);
        return B2_EINVAL;
    }
    if (!tcq_supported(type, (int)K) || T > TCQ_MAX_T) {
        set_error("b2_tcq_gemm: type %d, K=%lld, T=%d is outside the fused path (Q4_K / Q6_K, K a multiple of 256, T <= %d)", type, (long long)K, T,
                  TCQ_MAX_T);
        return B2_EUNSUP;
    }
    __half* xb = (__half*)scratch_dev;
    float* ws = (float*)((uint8_t*)scratch_dev + ((size_t)T * K * 2 + 255) / 256 * 256);
    int rc;
    const bool perm = tcq_fast_conversion();
    if ((rc = launch_rows_to_f16(const_cast<float*>(x_dev), K, nullptr, 0.0f, (int)K, T, ACT_Q8K, xb, (cudaStream_t)stream, TcParts{nullptr, 1, 0}, perm)))
        return rc;
    TcqGemmArgs a{};
    a.x_perm = perm ? 1 : 0;
    a.seg[0] = TcqSeg{W_dev, type, (int)N};
    a.n_seg = 1;
    a.x = xb;
    a.K = (int)K;
    a.T = T;
    a.ws = ws;
    a.ws_bytes = tcq_ws_bytes((int)N, (int)K, T);
    TcParts parts;
    if ((rc = launch_tcq_gemm(a, (cudaStream_t)stream, &parts))) return rc;
    return launch_tcq_reduce(parts, T, (int)N, y_dev, N, nullptr, accumulate ? TC_EPI_ADD : TC_EPI_STORE, (cudaStream_t)stream);
}

int b2_rmsnorm(const float* x_dev, const float* w_dev, float* y_dev, int64_t n, int64_t rows, float eps, void* stream) {
    if (!x_dev || !w_dev || !y_dev) {
        set_error("b2_rmsnorm: NULL pointer");
        return B2_EINVAL;
    }
    return launch_rmsnorm(x_dev, w_dev, y_dev, n, rows, eps, (cudaStream_t)stream);
}

int b2_rope(float* x_dev, int n_heads, int head_dim, int n_dims, int mode, int pos, float freq_base, const float* freq_factors_dev,
            void* stream) {
    if (!x_dev) {
        set_error("b2_rope: NULL pointer");
        return B2_EINVAL;
    }
    return launch_rope(x_dev, n_heads, head_dim, n_dims, mode, pos, freq_base, freq_factors_dev, (cudaStream_t)stream);
}

size_t b2_attn_scratch_bytes(int n_head, int head_dim, int n_tokens) {
    return attn_scratch_bytes(n_head, head_dim, attn_max_splits(n_tokens));
}

int b2_attn_decode(const float* q_dev, const void* k_cache_dev, const void* v_cache_dev, float* out_dev, int n_head, int n_kv,
                   int head_dim, int n_ctx, int n_tokens, void* scratch_dev, void* stream) {
    if (!q_dev || !k_cache_dev || !v_cache_dev || !out_dev || !scratch_dev) {
        set_error("b2_attn_decode: NULL pointer");
        return B2_EINVAL;
    }
    return launch_attn_decode(q_dev, (const __half*)k_cache_dev, (const __half*)v_cache_dev, out_dev, n_head, n_kv, head_dim, n_ctx,
                              nullptr, n_tokens, attn_max_splits(n_tokens), scratch_dev, (cudaStream_t)stream, false);
}

size_t b2_argmax_scratch_bytes(int64_t n) {
    (void)n;
    return argmax_scratch_bytes();
}

int b2_sample_greedy(const float* logits_dev, int64_t n, int32_t* token_out_dev, void* scratch_dev, void* stream) {
    if (!logits_dev || !token_out_dev || !scratch_dev) {
        set_error("b2_sample_greedy: NULL pointer");
        return B2_EINVAL;
    }
    return launch_argmax(logits_dev, n, token_out_dev, scratch_dev, nullptr, nullptr, nullptr, 0, (cudaStream_t)stream, false);
}

int b2_sample(const float* logits_dev, int n_rows, int64_t vocab, int64_t row_stride, const b2_sampling* params_dev, int64_t param_stride,
              const uint64_t* draw_dev, int32_t* tokens_out_dev, void* stream) {
    if (row_stride < 0 || param_stride < 0 || (row_stride != 0 && row_stride < vocab)) {
        set_error("b2_sample: row_stride=%lld param_stride=%lld (row_stride is 0 or at least vocab)", (long long)row_stride, (long long)param_stride);
        return B2_EINVAL;
    }
    return launch_sample_rows(logits_dev, vocab, row_stride, n_rows, params_dev, param_stride, draw_dev, tokens_out_dev, false, (cudaStream_t)stream);
}

}  // extern "C"
