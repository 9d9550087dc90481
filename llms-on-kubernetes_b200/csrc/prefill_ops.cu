// prefill_ops.cu -- the elementwise kernels between the prompt-phase mat-muls (tc_gemm.cu) and attention
// (prefill_attn.cu), each over all T prompt rows at once: embedding rows (row a10), bias + RoPE + KV-cache store
// (rows a13, a15), greedy sampling over rows of logits (row a16).  Same formulas as the fused epilogues of the decode kernels
// (gemv_device.cuh / gemm.cu), so a prompt and a token-by-token replay agree to rounding.
#include <cuda_fp16.h>

#include "common.h"
#include "dequant.cuh"
#include "gemv.cuh"
#include "ops.h"

namespace b2 {
namespace {

__global__ void embed_rows_kernel(int type, const uint8_t* __restrict__ table, int64_t row_bytes, int d_model, int vocab,
                                  const int32_t* __restrict__ tokens, float* __restrict__ x) {
    int tok = tokens[blockIdx.y];
    if (tok < 0 || tok >= vocab) tok = 0;   // validated on the host
    const uint8_t* row = table + (size_t)tok * row_bytes;
    float4* dst = (float4*)(x + (size_t)blockIdx.y * d_model);
    for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < d_model / 4; g += gridDim.x * blockDim.x) dst[g] = dq4(type, row, g);
}

// qkv: [T][nq + 2 nk] mat-mul output.  One thread per element pair: Q pairs are rotated into q, K pairs rotated into
// the fp16 cache, V pairs stored to the fp16 cache.
__global__ void __launch_bounds__(256) qkv_post_kernel(PrefillQkvArgs a) {
    const int t = blockIdx.y;
    const int pi = blockIdx.x * blockDim.x + threadIdx.x;
    const int nq2 = a.nq / 2, nk2 = a.nk / 2;
    if (pi >= nq2 + 2 * nk2) return;
    const size_t row_off = (size_t)t * (a.nq + 2 * a.nk);
    // element `col` of this row of the mat-mul output: the buffer, or the sum of its deferred split-K slices
    auto val = [&](int col) {
        if (a.parts_splits <= 1) return a.qkv[row_off + col];
        float v = 0.0f;
        for (int s = 0; s < a.parts_splits; s++) v += a.parts_ws[s * a.parts_slice + row_off + col];
        return v;
    };
    const int pos = a.col_pos ? a.col_pos[t] : a.pos0 + t;
    // cache row of this position: through the page table (slot = the column's, or the prompt's one sequence), else the
    // contiguous window of the slot
    size_t slot_off = a.col_slot ? (size_t)a.col_slot[t] * a.slot_stride : 0;
    int row = pos;
    if (a.kv_table) {
        const int slot = a.col_slot ? a.col_slot[t] : a.slot;
        row = a.kv_table[slot * a.kv_pps + (pos >> 8)] * 256 + (pos & 255);
        slot_off = 0;
    }
    const int hd = a.head_dim, hp = hd / 2;
    if (pi < nq2 + nk2) {
        const bool is_q = pi < nq2;
        const int lp = is_q ? pi : pi - nq2;             // pair index inside Q or K
        const int head = lp / hp, i = lp - head * hp;    // pair i of the head
        const int ea = a.neox ? head * hd + i : head * hd + 2 * i;
        const int eb = a.neox ? ea + hp : ea + 1;
        const int src = is_q ? 0 : a.nq;
        const float* bias = is_q ? a.bq : a.bk;
        float va = val(src + ea), vb = val(src + eb);
        if (bias) {
            va += bias[ea];
            vb += bias[eb];
        }
        float oa = va, ob = vb;
        if (i < a.rope_half) {
            const float2 cs = a.rope_tab[(size_t)pos * a.rope_half + i];
            oa = va * cs.x - vb * cs.y;
            ob = va * cs.y + vb * cs.x;
        }
        if (is_q) {
            float* dst = a.q + (size_t)t * a.nq;
            dst[ea] = oa;
            dst[eb] = ob;
        } else {
            __half* dst = a.k_cache + slot_off + ((size_t)head * a.n_ctx + row) * hd;
            dst[ea - head * hd] = __float2half_rn(oa);
            dst[eb - head * hd] = __float2half_rn(ob);
        }
    } else {
        const int lp = pi - nq2 - nk2;
        const int ea = 2 * lp, head = ea / hd;
        float va = val(a.nq + a.nk + ea), vb = val(a.nq + a.nk + ea + 1);
        if (a.bv) {
            va += a.bv[ea];
            vb += a.bv[ea + 1];
        }
        __half* dst = a.v_cache + slot_off + ((size_t)head * a.n_ctx + row) * hd + (ea - head * hd);
        *(__half2*)dst = __floats2half2_rn(va, vb);
    }
}

// greedy token of every row of logits[rows][n]: one CTA per row; the largest value wins, equal values keep the lower
// index (the oracle scans upwards with '>')
__global__ void __launch_bounds__(1024) argmax_rows_kernel(const float* __restrict__ logits, int64_t n, int32_t* __restrict__ out) {
    const float* row = logits + (size_t)blockIdx.x * n;
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float v = row[i];
        if (v > bv) {   // ascending i per thread: strict '>' keeps the first maximum
            bv = v;
            bi = (int)i;
        }
    }
    auto better = [](float v, int i, float ov, int oi) { return ov > v || (ov == v && oi < i); };
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (better(bv, bi, ov, oi)) {
            bv = ov;
            bi = oi;
        }
    }
    __shared__ float sv[32];
    __shared__ int si[32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        sv[warp] = bv;
        si[warp] = bi;
    }
    __syncthreads();
    if (warp == 0) {
        bv = lane < (int)(blockDim.x >> 5) ? sv[lane] : -INFINITY;
        bi = lane < (int)(blockDim.x >> 5) ? si[lane] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (better(bv, bi, ov, oi)) {
                bv = ov;
                bi = oi;
            }
        }
        if (lane == 0) out[blockIdx.x] = bi == 0x7fffffff ? 0 : bi;
    }
}

}  // namespace

int launch_argmax_rows(const float* logits, int64_t n, int rows, int32_t* out, cudaStream_t st) {
    argmax_rows_kernel<<<rows, 1024, 0, st>>>(logits, n, out);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int launch_embed_rows(int type, const void* table, int64_t row_bytes, int d_model, int vocab, const int32_t* tokens, int T, float* x, cudaStream_t st) {
    const int threads = 128;
    const dim3 grid((d_model / 4 + threads - 1) / threads, T);
    embed_rows_kernel<<<grid, threads, 0, st>>>(type, (const uint8_t*)table, row_bytes, d_model, vocab, tokens, x);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int launch_prefill_qkv_post(const PrefillQkvArgs& a, int T, cudaStream_t st) {
    if (a.nq % 2 || a.nk % 2 || a.head_dim % 2 || T <= 0) {
        set_error("prefill_qkv_post: bad shape");
        return B2_EINVAL;
    }
    const int pairs = a.nq / 2 + a.nk;
    const dim3 grid((pairs + 255) / 256, T);
    qkv_post_kernel<<<grid, 256, 0, st>>>(a);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b2
