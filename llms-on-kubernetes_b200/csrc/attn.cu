// attn.cu -- single-query (decode) attention over the fp16 KV cache, split along the sequence
// (8a row a14).
//
// Numerics follow the classic (non-flash) ggml CPU path restated in oracle/ggml_ref.c: K and V are
// stored fp16, the query and the soft-max probabilities are rounded to fp16 before their dot
// products (mul_mat converts src1 to the F16 vec_dot type), accumulation is fp32, the soft-max is
// exp(s - max) / sum.  With one split (n_tokens <= ATTN_SPLIT_TOKENS) that is reproduced exactly up
// to summation order; with several splits each split normalises by its own maximum and the partial
// results are merged with the usual log-sum-exp correction (probabilities are then rounded to fp16
// before that correction -- a difference of at most one fp16 ulp on a probability, covered by the
// stated logits tolerance).
//
// Layout: cache [n_kv][n_ctx][head_dim] so one KV head's rows are contiguous: a CTA streams
// head_dim*2 bytes per token with 16-byte loads, fully coalesced.  One CTA per (kv head, split)
// serves all n_head/n_kv query heads of the group, so the cache is read once per group.
// Scratch: per (head, split) partial {max, sum, out[head_dim]} plus one arrival counter per kv head;
// the last CTA of a kv head to arrive merges the partials and resets the counter, so the scratch is
// reusable without a memset between launches.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <mutex>

#include "common.h"
#include "device_utils.cuh"
#include "ops.h"

namespace b2 {

namespace {

constexpr int ATTN_THREADS = 256;
constexpr int ATTN_WARPS = ATTN_THREADS / 32;
constexpr int MAX_GROUP = 8;   // query heads per kv head handled by one CTA
constexpr int MAX_HD = 256;

// partial record: [0] = max, [1] = sum, [2..2+hd) = unnormalised output
__device__ __forceinline__ size_t part_stride(int hd) { return (size_t)hd + 2; }

// MG = query heads per kv head the registers are sized for (4 keeps two CTAs per SM; 8 for wider GQA groups)
template <int HD, int MG>
__global__ void __launch_bounds__(ATTN_THREADS, MG <= 4 ? 2 : 1) attn_decode_kernel(const float* __restrict__ q, const __half* __restrict__ kc,
                                                                   const __half* __restrict__ vc, float* __restrict__ out,
                                                                   int n_head, int n_kv, int n_ctx, const int* __restrict__ n_tokens_ptr,
                                                                   int n_tokens_imm, int max_splits, float* __restrict__ partials,
                                                                   unsigned int* __restrict__ counters, const int* __restrict__ col_slot,
                                                                   long long slot_stride, int split_tokens) {
    pdl_launch_dependents();
    pdl_wait();
    constexpr int LPT = HD / 8;              // lanes per token row (8 halves = 16 bytes per lane)
    constexpr int TPW = 32 / LPT;            // tokens per warp load
    const int kvh = blockIdx.x, split = blockIdx.y;
    const int g = n_head / n_kv;
    // blockIdx.z = column of a batch: its own query / output row, position, sequence slot and scratch
    const int col = blockIdx.z;
    q += (size_t)col * n_head * HD;
    out += (size_t)col * n_head * HD;
    if (col_slot) {
        kc += (size_t)col_slot[col] * slot_stride;
        vc += (size_t)col_slot[col] * slot_stride;
    }
    partials += (size_t)col * n_head * max_splits * part_stride(HD);
    counters += (size_t)col * n_kv;
    const int T = n_tokens_ptr ? n_tokens_ptr[col] + 1 : n_tokens_imm;   // device position + the token just written
    const int n_splits = min(max_splits, (T + split_tokens - 1) / split_tokens);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    __shared__ float s_scores[MG][ATTN_SPLIT_TOKENS];
    __shared__ float s_red[MG][ATTN_WARPS];
    __shared__ float s_out[MG][HD];
    __shared__ float s_m[MG], s_l[MG];
    __shared__ int s_last;

    if (split < n_splits) {
        const int t0 = split * split_tokens;
        const int t1 = min(T, (split == n_splits - 1) ? T : t0 + split_tokens);
        // tokens beyond max_splits*ATTN_SPLIT_TOKENS fall into the last split: cover them in passes
        const __half* kbase = kc + (size_t)kvh * n_ctx * HD;
        const __half* vbase = vc + (size_t)kvh * n_ctx * HD;
        const int sub = lane / LPT, ln = lane % LPT;   // which token of the warp load, which 8-dim slice
        const float scale = 1.0f / sqrtf((float)HD);

        // query slice for this lane, rounded to fp16 like ggml's F16 vec_dot operand
        float qf[MG][8];
#pragma unroll
        for (int h = 0; h < MG; h++)
            if (h < g) {
                const float* qp = q + ((size_t)(kvh * g + h)) * HD + ln * 8;
#pragma unroll
                for (int i = 0; i < 8; i++) qf[h][i] = __half2float(__float2half_rn(qp[i]));
            }

        float run_max[MG], run_sum[MG], acc[MG][8];
#pragma unroll
        for (int h = 0; h < MG; h++) {
            run_max[h] = -INFINITY;
            run_sum[h] = 0.0f;
#pragma unroll
            for (int i = 0; i < 8; i++) acc[h][i] = 0.0f;
        }

        for (int p0 = t0; p0 < t1; p0 += split_tokens) {
            const int p1 = min(t1, p0 + split_tokens);
            const int n = p1 - p0;
            // ---- scores s[h][t] = scale * <k_t, q_h> ----
            for (int tb = warp * TPW; tb < n; tb += ATTN_WARPS * TPW) {
                const int t = tb + sub;
                float d[MG];
#pragma unroll
                for (int h = 0; h < MG; h++) d[h] = 0.0f;
                if (t < n) {
                    const uint4 kv = *(const uint4*)(kbase + (size_t)(p0 + t) * HD + ln * 8);
                    const __half2* k2 = (const __half2*)&kv;
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        const float2 kf = __half22float2(k2[i]);
#pragma unroll
                        for (int h = 0; h < MG; h++)
                            if (h < g) d[h] += kf.x * qf[h][2 * i] + kf.y * qf[h][2 * i + 1];
                    }
                }
#pragma unroll
                for (int h = 0; h < MG; h++)
                    if (h < g) {
#pragma unroll
                        for (int o = LPT / 2; o > 0; o >>= 1) d[h] += __shfl_xor_sync(0xffffffffu, d[h], o);
                        if (ln == 0 && t < n) s_scores[h][t] = d[h] * scale;
                    }
            }
            __syncthreads();
            // ---- pass maximum, probabilities (fp32 exp, fp16-rounded for the V product), sum ----
            float pmax[MG], psum[MG];
#pragma unroll
            for (int h = 0; h < MG; h++)
                if (h < g) {
                    float m = -INFINITY;
                    for (int t = tid; t < n; t += ATTN_THREADS) m = fmaxf(m, s_scores[h][t]);
                    m = warp_max(m);
                    if (lane == 0) s_red[h][warp] = m;
                }
            __syncthreads();
#pragma unroll
            for (int h = 0; h < MG; h++)
                if (h < g) {
                    float m = s_red[h][0];
#pragma unroll
                    for (int w = 1; w < ATTN_WARPS; w++) m = fmaxf(m, s_red[h][w]);
                    pmax[h] = m;
                }
            __syncthreads();
#pragma unroll
            for (int h = 0; h < MG; h++)
                if (h < g) {
                    float s = 0.0f;
                    for (int t = tid; t < n; t += ATTN_THREADS) {
                        const float e = expf(s_scores[h][t] - pmax[h]);
                        s_scores[h][t] = e;
                        s += e;
                    }
                    s = warp_sum(s);
                    if (lane == 0) s_red[h][warp] = s;
                }
            __syncthreads();
#pragma unroll
            for (int h = 0; h < MG; h++)
                if (h < g) {
                    float s = 0.0f;
#pragma unroll
                    for (int w = 0; w < ATTN_WARPS; w++) s += s_red[h][w];
                    psum[h] = s;
                }
            // single pass over the whole sequence: normalise before rounding, exactly like the oracle
            const bool exact = (n_splits == 1 && p0 == t0 && p1 == t1);
            // ---- o[h][d] += p[h][t] * v[t][d] ----
            float corr[MG], pn[MG];
#pragma unroll
            for (int h = 0; h < MG; h++)
                if (h < g) {
                    const float nm = fmaxf(run_max[h], pmax[h]);
                    corr[h] = run_max[h] == -INFINITY ? 0.0f : expf(run_max[h] - nm);
                    pn[h] = exact ? 1.0f / psum[h] : expf(pmax[h] - nm);
                    run_sum[h] = run_sum[h] * corr[h] + (exact ? 1.0f : psum[h] * expf(pmax[h] - nm));
                    run_max[h] = nm;
#pragma unroll
                    for (int i = 0; i < 8; i++) acc[h][i] *= corr[h];
                }
            for (int tb = warp * TPW; tb < n; tb += ATTN_WARPS * TPW) {
                const int t = tb + sub;
                if (t < n) {
                    const uint4 vv = *(const uint4*)(vbase + (size_t)(p0 + t) * HD + ln * 8);
                    const __half2* v2 = (const __half2*)&vv;
                    float vf[8];
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        const float2 f = __half22float2(v2[i]);
                        vf[2 * i] = f.x;
                        vf[2 * i + 1] = f.y;
                    }
#pragma unroll
                    for (int h = 0; h < MG; h++)
                        if (h < g) {
                            const float e = s_scores[h][t];
                            const float pr = exact ? __half2float(__float2half_rn(e * pn[h]))
                                                   : __half2float(__float2half_rn(e)) * pn[h];
#pragma unroll
                            for (int i = 0; i < 8; i++) acc[h][i] += pr * vf[i];
                        }
                }
            }
            __syncthreads();
        }
        // ---- reduce acc over the token sub-lanes of a warp, then over warps ----
#pragma unroll
        for (int h = 0; h < MG; h++)
            if (h < g) {
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    float v = acc[h][i];
#pragma unroll
                    for (int o = LPT; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                    acc[h][i] = v;
                }
            }
        for (int i = tid; i < MG * HD; i += ATTN_THREADS) (&s_out[0][0])[i] = 0.0f;
        __syncthreads();
        for (int w = 0; w < ATTN_WARPS; w++) {
            if (warp == w && sub == 0) {
#pragma unroll
                for (int h = 0; h < MG; h++)
                    if (h < g) {
#pragma unroll
                        for (int i = 0; i < 8; i++) s_out[h][ln * 8 + i] += acc[h][i];
                    }
            }
            __syncthreads();
        }
        if (tid == 0) {
#pragma unroll
            for (int h = 0; h < MG; h++)
                if (h < g) {
                    s_m[h] = run_max[h];
                    s_l[h] = run_sum[h];
                }
        }
        __syncthreads();
        if (n_splits == 1) {
            // s_l == 1 when the whole sequence fitted one pass (already normalised, the exact path)
            for (int i = tid; i < g * HD; i += ATTN_THREADS) {
                const int h = i / HD, d = i - h * HD;
                out[((size_t)(kvh * g + h)) * HD + d] = s_l[h] == 1.0f ? s_out[h][d] : s_out[h][d] / s_l[h];
            }
            return;
        }
        // write the partial record {max, sum, out[HD]}
        for (int i = tid; i < g * HD; i += ATTN_THREADS) {
            const int h = i / HD, d = i - h * HD;
            float* rec = partials + ((size_t)(kvh * g + h) * max_splits + split) * part_stride(HD);
            rec[2 + d] = s_out[h][d];
        }
        if (tid < g) {
            float* rec = partials + ((size_t)(kvh * g + tid) * max_splits + split) * part_stride(HD);
            rec[0] = s_m[tid];
            rec[1] = s_l[tid];
        }
    }
    if (n_splits == 1) return;   // CTAs beyond the single split have nothing to do
    // ---- last CTA of this kv head merges ----
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned int arrived = atomicAdd(&counters[kvh], 1u) + 1u;
        s_last = (arrived == (unsigned int)gridDim.y);
        if (s_last) counters[kvh] = 0u;   // reusable by the next launch
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int i = tid; i < g * HD; i += ATTN_THREADS) {
        const int h = i / HD, d = i - h * HD;
        const float* base = partials + ((size_t)(kvh * g + h) * max_splits) * part_stride(HD);
        float m = -INFINITY;
        for (int s = 0; s < n_splits; s++) m = fmaxf(m, base[s * part_stride(HD)]);
        float den = 0.0f, num = 0.0f;
        for (int s = 0; s < n_splits; s++) {
            const float* rec = base + s * part_stride(HD);
            const float w = expf(rec[0] - m);
            den += rec[1] * w;
            num += rec[2 + d] * w;
        }
        out[((size_t)(kvh * g + h)) * HD + d] = num / den;
    }
}

}  // namespace

int attn_max_splits(int n_ctx) {
    int s = (n_ctx + ATTN_SPLIT_TOKENS - 1) / ATTN_SPLIT_TOKENS;
    if (s > ATTN_MAX_SPLITS) s = ATTN_MAX_SPLITS;
    return s < 1 ? 1 : s;
}

size_t attn_scratch_bytes(int n_head, int head_dim, int max_splits) {
    // partial records, then one counter per kv head (n_head is an upper bound for n_kv), 256-aligned
    size_t part = (size_t)n_head * max_splits * ((size_t)head_dim + 2) * sizeof(float);
    part = (part + 255) & ~(size_t)255;
    return part + (((size_t)n_head * sizeof(unsigned int) + 255) & ~(size_t)255);
}

size_t attn_scratch_bytes_batch(int n_head, int head_dim, int max_splits, int n_cols) {
    size_t part = (size_t)n_cols * n_head * max_splits * ((size_t)head_dim + 2) * sizeof(float);
    part = (part + 255) & ~(size_t)255;
    return part + (((size_t)n_cols * n_head * sizeof(unsigned int) + 255) & ~(size_t)255);
}

int launch_attn_decode(const float* q, const __half* kc, const __half* vc, float* out, int n_head, int n_kv, int head_dim,
                       int n_ctx, const int* pos_ptr, int n_tokens_imm, int max_splits, void* scratch, cudaStream_t st, bool pdl,
                       int n_cols, const int* col_slot, int64_t slot_stride, int split_tokens, int cols_cap) {
    if (cols_cap < n_cols) cols_cap = n_cols;
    if (split_tokens < 16 || split_tokens > ATTN_SPLIT_TOKENS) {
        set_error("attn_decode: split_tokens=%d outside [16, %d]", split_tokens, ATTN_SPLIT_TOKENS);
        return B2_EINVAL;
    }
    if (n_cols < 1 || (n_cols > 1 && !pos_ptr)) {
        set_error("attn_decode: a batch of %d columns needs per-column positions", n_cols);
        return B2_EINVAL;
    }
    if (n_head <= 0 || n_kv <= 0 || n_head % n_kv || n_head / n_kv > MAX_GROUP) {
        set_error("attn_decode: n_head=%d n_kv=%d unsupported (group size must divide and be <= %d)", n_head, n_kv, MAX_GROUP);
        return B2_EUNSUP;
    }
    if (head_dim != 64 && head_dim != 128) {
        set_error("attn_decode: head_dim=%d unsupported (64 or 128)", head_dim);
        return B2_EUNSUP;
    }
    if (!pos_ptr && (n_tokens_imm < 1 || n_tokens_imm > n_ctx)) {
        set_error("attn_decode: n_tokens=%d outside [1, n_ctx=%d]", n_tokens_imm, n_ctx);
        return B2_ERANGE;
    }
    if (max_splits < 1 || max_splits > ATTN_MAX_SPLITS) {
        set_error("attn_decode: max_splits=%d out of range", max_splits);
        return B2_EINVAL;
    }
    // the arrival counters follow the partial records of cols_cap columns -- the capacity the scratch was sized and
    // zeroed for (attn_scratch_bytes_batch), NOT this launch's column count: a smaller batch must find its counters
    // where every other batch size leaves them at zero, not inside a region earlier launches wrote records to
    size_t part = (size_t)cols_cap * n_head * max_splits * ((size_t)head_dim + 2) * sizeof(float);
    part = (part + 255) & ~(size_t)255;
    float* partials = (float*)scratch;
    unsigned int* counters = (unsigned int*)((uint8_t*)scratch + part);
    // with an immediate length the grid can be sized to the splits actually needed
    int gy = max_splits;
    if (!pos_ptr) {
        gy = (n_tokens_imm + split_tokens - 1) / split_tokens;
        if (gy > max_splits) gy = max_splits;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(n_kv, gy, n_cols);
    cfg.blockDim = dim3(ATTN_THREADS);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const bool narrow = n_head / n_kv <= 4;
    auto kern = head_dim == 64 ? (narrow ? attn_decode_kernel<64, 4> : attn_decode_kernel<64, MAX_GROUP>)
                               : (narrow ? attn_decode_kernel<128, 4> : attn_decode_kernel<128, MAX_GROUP>);
    B2_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, q, kc, vc, out, n_head, n_kv, n_ctx, pos_ptr, n_tokens_imm, gy, partials, counters,
                                     col_slot, (long long)slot_stride, split_tokens));
    return 0;
}

}  // namespace b2

// -------------------------------------------------------------------------------------------------
// b2_attn_decode_static: the same split-sequence attention, for callers that must not allocate.
//
// Some hosts of this library cannot call cudaMalloc when they need attention: code running inside
// a CUDA-graph capture or a latency-critical serving loop where an allocation would synchronise
// the device, a sandboxed plugin that was handed a fixed memory budget at start-up, or an
// embedder whose allocator owns all of HBM and gives none to third-party libraries.  For them the
// split-merge scratch lives in the library image itself: a file-scope __device__ array of a fixed,
// modest size that every GPU loading libb2.so carries until the process exits.
//
// The array's name is only a host-side handle; the device address is looked up once per device with
// cudaGetSymbolAddress (with that device current) and cached.  That look-up happens on the FIRST call
// on a device, so make that first call outside of any stream capture.  The entry point calls no
// cudaMalloc / cudaFree and no allocating library routine.
//
// Limits (rejected with B2_ERANGE beyond them): n_head <= 64, head_dim <= 128, n_tokens <= 4096.
// All calls on a device share the one array: the partial records are fully rewritten by every call
// that uses them and the arrival counters are cleared on the caller's stream before each launch, but
// two calls whose kernels overlap on the GPU (different streams, no ordering) race on it.
// -------------------------------------------------------------------------------------------------
namespace b2 {
namespace {
constexpr int STATIC_MAX_HEADS = 64, STATIC_MAX_HD = 128, STATIC_MAX_TOKENS = 4096;
constexpr int STATIC_MAX_SPLITS = STATIC_MAX_TOKENS / ATTN_SPLIT_TOKENS;
constexpr size_t STATIC_PART_BYTES = ((size_t)STATIC_MAX_HEADS * STATIC_MAX_SPLITS * (STATIC_MAX_HD + 2) * sizeof(float) + 255) & ~(size_t)255;
constexpr size_t STATIC_SCRATCH_BYTES = STATIC_PART_BYTES + 256;   // + arrival counters (one per kv head)
static __device__ __align__(256) unsigned char g_attn_static_scratch[STATIC_SCRATCH_BYTES];

void* g_static_ptr[MAX_DEVICES];   // device address of the array, per device, filled on first use
std::mutex g_static_mu;
}  // namespace
}  // namespace b2

extern "C" int b2_attn_decode_static(const float* q_dev, const void* k_cache_dev, const void* v_cache_dev, float* out_dev, int n_head,
                                     int n_kv, int head_dim, int n_ctx, int n_tokens, void* stream) {
    using namespace b2;
    if (!q_dev || !k_cache_dev || !v_cache_dev || !out_dev) {
        set_error("b2_attn_decode_static: NULL pointer");
        return B2_EINVAL;
    }
    if (n_head > STATIC_MAX_HEADS || head_dim > STATIC_MAX_HD || n_tokens > STATIC_MAX_TOKENS) {
        set_error("b2_attn_decode_static: n_head=%d head_dim=%d n_tokens=%d exceed the built-in scratch (max %d heads, head_dim %d, %d tokens); "
                  "use b2_attn_decode with caller-provided scratch",
                  n_head, head_dim, n_tokens, STATIC_MAX_HEADS, STATIC_MAX_HD, STATIC_MAX_TOKENS);
        return B2_ERANGE;
    }
    // every precondition of the launch is checked here, BEFORE the array is touched: a negative or zero size would
    // otherwise wrap the counter offset below and the memset would clear device globals in front of the array
    if (n_head < 1 || n_kv < 1 || n_head % n_kv || (head_dim != 64 && head_dim != 128) || n_tokens < 1 || n_ctx < 1 || n_tokens > n_ctx) {
        set_error("b2_attn_decode_static: n_head=%d n_kv=%d head_dim=%d n_ctx=%d n_tokens=%d: need n_head >= 1, n_kv >= 1 dividing n_head, "
                  "head_dim 64 or 128, 1 <= n_tokens <= n_ctx",
                  n_head, n_kv, head_dim, n_ctx, n_tokens);
        return B2_EINVAL;
    }
    int dev = 0;
    B2_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) {
        set_error("b2_attn_decode_static: device ordinal %d not supported", dev);
        return B2_ERANGE;
    }
    void* scratch;
    {
        // first call on this device: resolve the symbol's device address (the current device is `dev`)
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (!g_static_ptr[dev]) {
            void* p = nullptr;
            B2_CUDA_CHECK(cudaGetSymbolAddress(&p, g_attn_static_scratch));
            g_static_ptr[dev] = p;
        }
        scratch = g_static_ptr[dev];
    }
    const int max_splits = attn_max_splits(n_tokens);
    // same layout launch_attn_decode derives from (n_head, head_dim, max_splits): partial records, then counters
    size_t part = (size_t)n_head * max_splits * ((size_t)head_dim + 2) * sizeof(float);
    part = (part + 255) & ~(size_t)255;
    if (max_splits < 1 || max_splits > STATIC_MAX_SPLITS || part + 256 > STATIC_SCRATCH_BYTES) {
        set_error("b2_attn_decode_static: internal layout (%zu + 256 bytes, %d splits) does not fit the built-in scratch", part, max_splits);
        return B2_ERANGE;
    }
    // the counters must start at zero even if an earlier call on this array was interrupted
    B2_CUDA_CHECK(cudaMemsetAsync((unsigned char*)scratch + part, 0, 256, (cudaStream_t)stream));
    return launch_attn_decode(q_dev, (const __half*)k_cache_dev, (const __half*)v_cache_dev, out_dev, n_head, n_kv, head_dim, n_ctx,
                              nullptr, n_tokens, max_splits, scratch, (cudaStream_t)stream, false);
}
