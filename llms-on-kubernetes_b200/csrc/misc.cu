// misc.cu -- the small kernels around the mat-vecs: embedding-row lookup,
// greedy sampling (K10, row a16) with the device-resident loop bookkeeping, and stand-alone RMSNorm
// (K4, row a11) and RoPE (K5, row a13) for callers of the C ABI that want the operators unfused.
#include "dequant.cuh"
#include "device_utils.cuh"
#include "ops.h"

namespace b2 {

// ---- embedding: x = dequantise(table[*token]) -----------------------------------------------------
__global__ void embed_kernel(int type, const uint8_t* __restrict__ table, int64_t row_bytes, int d_model, int vocab,
                             const int32_t* __restrict__ token_ptr, float4* __restrict__ x) {
    pdl_launch_dependents();
    pdl_wait();
    int tok = *token_ptr;
    if (tok < 0 || tok >= vocab) tok = 0;   // callers validate on the host; never index out of the table
    const uint8_t* row = table + (size_t)tok * row_bytes;
    for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < d_model / 4; g += gridDim.x * blockDim.x) x[g] = dq4(type, row, g);
}

int launch_embed(int type, const void* table, int64_t row_bytes, int d_model, int vocab, const int32_t* token_ptr, float* x,
                 cudaStream_t st, bool pdl) {
    cudaLaunchConfig_t cfg{};
    const int threads = 128;
    cfg.gridDim = dim3((d_model / 4 + threads - 1) / threads);
    cfg.blockDim = dim3(threads);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    B2_CUDA_CHECK(cudaLaunchKernelEx(&cfg, embed_kernel, type, (const uint8_t*)table, row_bytes, d_model, vocab, token_ptr, (float4*)x));
    return 0;
}

// ---- greedy sampling -----------------------------------------------------------------------------
// Scratch: ARGMAX_BLOCKS {value, index} partials + one arrival counter.  The last block to arrive
// reduces the partials, writes the token, does the decode loop's bookkeeping and resets the counter.
struct ArgmaxScratch {
    float val[ARGMAX_BLOCKS];
    int idx[ARGMAX_BLOCKS];
    unsigned int counter;
};
size_t argmax_scratch_bytes() { return (sizeof(ArgmaxScratch) + 255) & ~(size_t)255; }

__device__ __forceinline__ void argmax_merge(float& v, int& i, float ov, int oi) {
    // larger value wins; equal values keep the lower index (the oracle scans upwards with '>')
    if (ov > v || (ov == v && oi < i)) {
        v = ov;
        i = oi;
    }
}

__global__ void __launch_bounds__(256) argmax_kernel(const float* __restrict__ logits, int64_t n, int32_t* __restrict__ token_out,
                                                     ArgmaxScratch* __restrict__ sc, int32_t* pos_ptr, int32_t* step_ptr,
                                                     int32_t* step_out, int step_cap) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float s_v[8];
    __shared__ int s_i[8];
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + tid; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = logits[i];
        if (v > bv) { bv = v; bi = (int)i; }   // i increases, so the first maximum is kept
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) argmax_merge(bv, bi, __shfl_xor_sync(0xffffffffu, bv, o), __shfl_xor_sync(0xffffffffu, bi, o));
    if (lane == 0) { s_v[warp] = bv; s_i[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < 8; w++) argmax_merge(bv, bi, s_v[w], s_i[w]);
        sc->val[blockIdx.x] = bv;
        sc->idx[blockIdx.x] = bi;
        __threadfence();
        const unsigned int arrived = atomicAdd(&sc->counter, 1u) + 1u;
        s_last = arrived == gridDim.x;
    }
    __syncthreads();
    if (!s_last || tid != 0) return;
    __threadfence();
    for (int b = 0; b < (int)gridDim.x; b++) argmax_merge(bv, bi, ((volatile float*)sc->val)[b], ((volatile int*)sc->idx)[b]);
    if (bi == 0x7fffffff) bi = 0;   // all -inf / NaN
    *token_out = bi;
    if (step_out) {
        const int s = *step_ptr;
        if (s < step_cap) step_out[s] = bi;
        *step_ptr = s + 1;
    }
    if (pos_ptr) *pos_ptr = *pos_ptr + 1;
    sc->counter = 0u;
}

int launch_argmax(const float* logits, int64_t n, int32_t* token_out, void* scratch, int32_t* pos_ptr, int32_t* step_ptr,
                  int32_t* step_out, int step_cap, cudaStream_t st, bool pdl) {
    if (n <= 0 || n > 0x7ffffffe) {
        set_error("argmax: n=%lld out of range", (long long)n);
        return B2_ERANGE;
    }
    cudaLaunchConfig_t cfg{};
    int64_t blocks = (n + 256 * 4 - 1) / (256 * 4);
    if (blocks > ARGMAX_BLOCKS) blocks = ARGMAX_BLOCKS;
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    B2_CUDA_CHECK(cudaLaunchKernelEx(&cfg, argmax_kernel, logits, n, token_out, (ArgmaxScratch*)scratch, pos_ptr, step_ptr, step_out, step_cap));
    return 0;
}

// ---- stand-alone RMSNorm: one CTA per row; ggml accumulates the squares in double ------------------
__global__ void __launch_bounds__(256) rmsnorm_kernel(const float* __restrict__ x, const float* __restrict__ w, float* __restrict__ y,
                                                      int64_t n, float eps) {
    __shared__ double s_red[8];
    const float* xr = x + blockIdx.x * n;
    float* yr = y + blockIdx.x * n;
    double ss = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) ss += (double)(xr[i] * xr[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = ss;
    __syncthreads();
    double tot = 0.0;
    for (int k = 0; k < 8; k++) tot += s_red[k];
    const float mean = (float)(tot / (double)n);
    const float scale = 1.0f / sqrtf(mean + eps);
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) yr[i] = (xr[i] * scale) * w[i];
}

int launch_rmsnorm(const float* x, const float* w, float* y, int64_t n, int64_t rows, float eps, cudaStream_t st) {
    if (n <= 0 || rows <= 0 || rows > 0x7fffffff) {
        set_error("rmsnorm: bad shape n=%lld rows=%lld", (long long)n, (long long)rows);
        return B2_EINVAL;
    }
    rmsnorm_kernel<<<(unsigned)rows, 256, 0, st>>>(x, w, y, n, eps);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// ---- stand-alone RoPE: theta_i = pos * theta_scale^i by repeated multiplication, as ggml builds its
// cache (ggml_rope_cache_init); the model path uses a host-built table instead (model.cu) ------------
__global__ void rope_kernel(float* __restrict__ x, int n_heads, int head_dim, int n_dims, int mode, int pos, float theta_scale,
                            const float* __restrict__ freq_factors) {
    const int half = n_dims / 2;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_heads * half) return;
    const int h = idx / half, i = idx - h * half;
    float theta = (float)pos;
    for (int k = 0; k < i; k++) theta *= theta_scale;
    if (freq_factors) theta = theta / freq_factors[i];
    const float c = cosf(theta), s = sinf(theta);
    float* v = x + (size_t)h * head_dim;
    const int ia = mode == 0 ? 2 * i : i, ib = mode == 0 ? 2 * i + 1 : i + half;
    const float x0 = v[ia], x1 = v[ib];
    v[ia] = x0 * c - x1 * s;
    v[ib] = x0 * s + x1 * c;
}

int launch_rope(float* x, int n_heads, int head_dim, int n_dims, int mode, int pos, float freq_base, const float* freq_factors,
                cudaStream_t st) {
    if (n_heads <= 0 || head_dim <= 0 || n_dims <= 0 || n_dims > head_dim || (n_dims & 1) || (mode != 0 && mode != 2) || pos < 0) {
        set_error("rope: bad arguments (n_heads=%d head_dim=%d n_dims=%d mode=%d pos=%d)", n_heads, head_dim, n_dims, mode, pos);
        return B2_EINVAL;
    }
    const float theta_scale = powf(freq_base, -2.0f / (float)n_dims);
    const int total = n_heads * (n_dims / 2);
    rope_kernel<<<(total + 127) / 128, 128, 0, st>>>(x, n_heads, head_dim, n_dims, mode, pos, theta_scale, freq_factors);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b2

namespace b2 {

int current_device() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    return dev;
}

int device_sm_count() {
    static std::mutex mu;
    static int sms[MAX_DEVICES] = {};
    const int dev = current_device();
    if (dev < 0) {
        set_error("no current CUDA device");
        return 0;
    }
    std::lock_guard<std::mutex> l(mu);
    if (dev < MAX_DEVICES && sms[dev] > 0) return sms[dev];
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
        set_error("cannot read the SM count of device %d", dev);
        return 0;
    }
    if (dev < MAX_DEVICES) sms[dev] = n;
    return n;
}

}  // namespace b2
