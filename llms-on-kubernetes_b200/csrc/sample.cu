// sample.cu -- device-side token sampling: temperature / top-k / min-p / top-p on a logits row, one
// Philox draw per (seed, draw index), one CTA per row.
//
// Why on the device: llama-server's defaults (temperature 0.8, top-k 40, top-p 0.95, min-p 0.05 -- what a request without
// sampling fields gets, e.g. the README's curl and OpenWebUI) are not greedy, and on the host every such token costs a
// 513 KB logits row over PCIe plus a partial sort of 128 K floats on the scheduler thread, for every sequence of a batch.
// Here the row never leaves HBM and a sequence that samples stays on the device-resident decode loop.
//
// The filter chain is the one the server's host sampler applies (csrc/server/engine.h: Engine::sample), in its order:
//   1. the k most probable tokens (ties: lower id first); k = top_k, at most SAMPLE_MAX_K (top_k <= 0 means SAMPLE_MAX_K)
//   2. p_i = exp((l_i - l_max) / temperature), normalised over those k
//   3. min-p: keep the leading tokens with p_i >= min_p * p_0
//   4. top-p: of those, the shortest prefix whose mass reaches top_p
//   5. draw u in [0, 1) from Philox4x32-10(key = seed, counter = draw index), pick the first token whose cumulative mass
//      reaches u * (kept mass)
// temperature <= 0 or top_k == 1 is greedy: the first index of the maximum.
// Top-k selection is exact: a 4-pass radix select over order-preserving keys finds the k-th largest logit, elements above
// it and the lowest-indexed ties on it are gathered in index order, then sorted (bitonic, value descending, id ascending).
#include <math_constants.h>

#include "common.h"
#include "ops.h"

namespace b2 {

namespace {

constexpr int ST = 1024;   // threads per row

__device__ __forceinline__ uint32_t order_key(float v) {
    uint32_t u = __float_as_uint(v);
    if (v != v) return 0u;                                   // NaN sorts below everything
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);       // larger float <=> larger key
}

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
// Philox4x32-10: first output word for (seed, index)
__device__ __forceinline__ uint32_t philox_u32(uint64_t seed, uint64_t index) {
    uint32_t c[4] = {(uint32_t)index, (uint32_t)(index >> 32), 0u, 0u};
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; r++) {
        philox_round(c, k0, k1);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c[0];
}

struct SampleSmem {
    uint32_t hist[256];
    uint32_t scan[ST];        // per-thread counts -> exclusive prefix
    float val[SAMPLE_MAX_K];
    int idx[SAMPLE_MAX_K];
    double cum[SAMPLE_MAX_K];
    uint32_t prefix, remaining, n_above;
    int keep;
    double total;
};

// exclusive prefix sum of one value per thread (ST threads); returns the total
__device__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* buf, int tid, uint32_t* excl) {
    buf[tid] = v;
    __syncthreads();
    for (int o = 1; o < ST; o <<= 1) {
        const uint32_t add = tid >= o ? buf[tid - o] : 0u;
        __syncthreads();
        buf[tid] += add;
        __syncthreads();
    }
    *excl = buf[tid] - v;
    const uint32_t total = buf[ST - 1];
    __syncthreads();
    return total;
}

__global__ void __launch_bounds__(ST) sample_rows_kernel(const float* __restrict__ logits, int64_t vocab, int64_t row_stride, const b2_sampling* __restrict__ params,
                                                        int64_t param_stride, const uint64_t* __restrict__ draw, int32_t* __restrict__ out, int out_only_sampled,
                                                        const int32_t* __restrict__ draw_i32, int32_t* __restrict__ echo_out,
                                                        const int32_t* __restrict__ echo_step, int echo_cap) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    SampleSmem& S = *(SampleSmem*)smem_raw;
    const int row = blockIdx.x, tid = threadIdx.x;
    const b2_sampling p = params[(int64_t)row * param_stride];
    const float* l = logits + (int64_t)row * row_stride;
    const bool greedy = !(p.temperature > 0.0f) || p.top_k == 1;
    if (greedy && out_only_sampled) return;   // the caller's argmax already wrote this row
    int k = p.top_k <= 0 || p.top_k > SAMPLE_MAX_K ? SAMPLE_MAX_K : p.top_k;
    if (greedy) k = 1;
    if ((int64_t)k > vocab) k = (int)vocab;
    // contiguous segment per thread: prefix sums over threads are then in index order (ties keep the lowest ids)
    const int64_t seg = (vocab + ST - 1) / ST;
    const int64_t i0 = (int64_t)tid * seg, i1 = i0 + seg < vocab ? i0 + seg : vocab;

    // ---- radix select: the key of the k-th largest logit ----
    if (tid == 0) { S.prefix = 0u; S.remaining = (uint32_t)k; }
    __syncthreads();
    for (int pass = 0; pass < 4; pass++) {
        const int shift = 24 - 8 * pass;
        const uint32_t mask = pass == 0 ? 0u : 0xffffffffu << (shift + 8);
        if (tid < 256) S.hist[tid] = 0u;
        __syncthreads();
        const uint32_t prefix = S.prefix;
        for (int64_t i = i0; i < i1; i++) {
            const uint32_t key = order_key(l[i]);
            if ((key & mask) == prefix) atomicAdd(&S.hist[(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (tid == 0) {
            // walk the buckets from the top until the k-th largest falls into one
            uint32_t rem = S.remaining;
            int b = 255;
            for (; b > 0; b--) {
                if (S.hist[b] >= rem) break;
                rem -= S.hist[b];
            }
            S.prefix = prefix | ((uint32_t)b << shift);
            S.remaining = rem;
        }
        __syncthreads();
    }
    const uint32_t tkey = S.prefix;          // key of the k-th largest
    const uint32_t n_tie_take = S.remaining; // how many elements equal to it belong to the top k (lowest ids first)

    // ---- gather in index order: everything above the threshold, then the first ties ----
    uint32_t c_above = 0, c_tie = 0;
    for (int64_t i = i0; i < i1; i++) {
        const uint32_t key = order_key(l[i]);
        c_above += key > tkey;
        c_tie += key == tkey;
    }
    uint32_t e_above, e_tie;
    const uint32_t n_above = block_exclusive_scan(c_above, S.scan, tid, &e_above);
    block_exclusive_scan(c_tie, S.scan, tid, &e_tie);
    for (int64_t i = i0; i < i1; i++) {
        const float v = l[i];
        const uint32_t key = order_key(v);
        if (key > tkey) {
            S.val[e_above] = v;
            S.idx[e_above] = (int)i;
            e_above++;
        } else if (key == tkey) {
            if (e_tie < n_tie_take) {
                S.val[n_above + e_tie] = v;
                S.idx[n_above + e_tie] = (int)i;
            }
            e_tie++;
        }
    }
    // pad to a power of two for the sort
    int n2 = 1;
    while (n2 < k) n2 <<= 1;
    for (int i = k + tid; i < n2; i += ST) { S.val[i] = -CUDART_INF_F; S.idx[i] = 0x7fffffff; }
    __syncthreads();
    // ---- bitonic sort: value descending, id ascending ----
    for (int size = 2; size <= n2; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < n2 / 2; t += ST) {
                const int lo = 2 * t - (t & (stride - 1));   // index of the lower element of the pair
                const int hi = lo + stride;
                const bool desc = (lo & size) == 0;
                const float va = S.val[lo], vb = S.val[hi];
                const int ia = S.idx[lo], ib = S.idx[hi];
                const bool a_first = va > vb || (va == vb && ia < ib);   // a belongs before b in the final order
                if (a_first != desc) {
                    S.val[lo] = vb; S.val[hi] = va;
                    S.idx[lo] = ib; S.idx[hi] = ia;
                }
            }
            __syncthreads();
        }
    }
    if (greedy) {
        if (tid == 0) {
            out[row] = S.idx[0];
            if (echo_out && echo_step) {
                const int at = *echo_step - 1;
                if (at >= 0 && at < echo_cap) echo_out[at] = S.idx[0];
            }
        }
        return;
    }
    // ---- probabilities, min-p, top-p, draw ----
    const float mx = S.val[0];
    const double inv_t = 1.0 / (double)p.temperature;
    for (int i = tid; i < k; i += ST) S.cum[i] = exp((double)(S.val[i] - mx) * inv_t);
    __syncthreads();
    if (tid == 0) {
        // k <= 1024: a serial pass in double keeps the order of the additions fixed (the same kept set as the host sampler)
        double sum = 0.0;
        for (int i = 0; i < k; i++) sum += S.cum[i];
        const double p0 = S.cum[0] / sum;
        int keep = k;
        if (p.min_p > 0.0f) {
            keep = 0;
            while (keep < k && S.cum[keep] / sum >= (double)p.min_p * p0) keep++;
            if (keep < 1) keep = 1;
        }
        if (p.top_p < 1.0f) {
            double acc = 0.0;
            int n = 0;
            while (n < keep) {
                acc += S.cum[n++] / sum;
                if (acc >= (double)p.top_p) break;
            }
            keep = n > 1 ? n : 1;
        }
        double tot = 0.0;
        for (int i = 0; i < keep; i++) tot += S.cum[i];
        const uint32_t x = philox_u32(p.seed, draw_i32 ? (uint64_t)(uint32_t)*draw_i32 : draw ? draw[row] : (uint64_t)row);
        double r = ((double)x + 0.5) * (1.0 / 4294967296.0) * tot;
        int pick = keep - 1;
        for (int i = 0; i < keep; i++) {
            r -= S.cum[i];
            if (r <= 0.0) { pick = i; break; }
        }
        out[row] = S.idx[pick];
        if (echo_out && echo_step) {
            const int at = *echo_step - 1;
            if (at >= 0 && at < echo_cap) echo_out[at] = S.idx[pick];
        }
    }
}

}  // namespace

int launch_sample_rows(const float* logits, int64_t vocab, int64_t row_stride, int rows, const b2_sampling* params_dev, int64_t param_stride,
                       const uint64_t* draw_dev, int32_t* out_dev, bool only_sampled, cudaStream_t st, const int32_t* draw_i32, int32_t* echo_out,
                       const int32_t* echo_step, int echo_cap) {
    if (!logits || !params_dev || !out_dev || rows <= 0 || vocab <= 0 || vocab > (int64_t)1 << 30) {
        set_error("sample: bad arguments (rows=%d vocab=%lld)", rows, (long long)vocab);
        return B2_EINVAL;
    }
    static PerDeviceSize configured;
    const size_t smem = sizeof(SampleSmem);
    const int dev = current_device();
    if (configured.below(dev, smem)) {
        B2_CUDA_CHECK(cudaFuncSetAttribute(sample_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured.set(dev, smem);
    }
    sample_rows_kernel<<<rows, ST, smem, st>>>(logits, vocab, row_stride, params_dev, param_stride, draw_dev, out_dev, only_sampled ? 1 : 0, draw_i32, echo_out,
                                               echo_step, echo_cap);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b2
