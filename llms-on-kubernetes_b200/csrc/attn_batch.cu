// attn_batch.cu -- decode attention of a batch of columns on the tensor cores (n_tokens = B).
//
// The persistent kernel's attention work item (step_attn.cuh: scores and P.V as mma.sync m16n8k16 with the GQA group's
// query heads as rows, K / V rows read straight from the fp16 cache in fragment order, softmax by two warps per head)
// as a stand-alone launch: one CTA of 256 threads per (column, kv head x batch of <= 4 query heads, chunk of <= 256
// tokens).  Every column has its own position and sequence slot; sequences longer than one chunk merge their chunk
// records by log-sum-exp in the last CTA to arrive, through the same scratch layout as attn.cu's batched launch
// (partial records of `cols_cap` columns, then the arrival counters).  Same numerics as the step kernel's stage, which
// the tests pin against the oracle at every sequence length.
#include "ops.h"
#include "step_attn.cuh"

namespace b2 {

using namespace stepattn;

namespace {

constexpr int AB_CHUNK = A_TOK;   // tokens per work item

template <int HD>
__global__ void __launch_bounds__(NCT) attn_batch_kernel(const float* __restrict__ q, const __half* __restrict__ kc, const __half* __restrict__ vc,
                                                         float* __restrict__ out, int n_head, int n_kv, int n_ctx, const int* __restrict__ col_pos,
                                                         const int* __restrict__ col_slot, long long slot_stride, int max_chunks,
                                                         float* __restrict__ partials, unsigned int* __restrict__ counters,
                                                         const int* __restrict__ kv_table, int kv_pps) {
    __shared__ float s_scores[A_MAXG * A_TOK], s_prob[A_MAXG * A_TOK], s_ml[2 * A_MAXG];
    __shared__ int s_flag;
    __shared__ StepOp s_op;   // only the attention fields are filled
    const int col = blockIdx.z, chunk = blockIdx.y;
    const int T = col_pos[col] + 1;
    const int n_chunks = min(max_chunks, (T + AB_CHUNK - 1) / AB_CHUNK);
    if (chunk >= n_chunks) return;
    const int gq = n_head / n_kv, nb = (gq + A_MAXG - 1) / A_MAXG;
    const int kvh = blockIdx.x / nb, h0 = (blockIdx.x % nb) * A_MAXG;
    const int g = min(A_MAXG, gq - h0);
    const int t0 = chunk * AB_CHUNK, n = min(AB_CHUNK, T - t0);
    if (threadIdx.x == 0) {
        const size_t qdim = (size_t)n_head * HD;
        // the work item addresses rows as base + (kvh * n_ctx + t0 + i) * HD.  Paged cache: the chunk [t0, t0 + 256) is one
        // page; shift the base so that t0 lands on the page's first token (n_ctx = tokens in the whole pool)
        long long slot_off = col_slot ? (long long)col_slot[col] * slot_stride : 0;
        if (kv_table) {
            const int slot = col_slot ? col_slot[col] : 0;
            slot_off = ((long long)kv_table[slot * kv_pps + chunk] * AB_CHUNK - t0) * HD;
        }
        s_op.a_q = q + (size_t)col * qdim;
        s_op.a_k = kc + slot_off;
        s_op.a_v = vc + slot_off;
        s_op.a_out = out + (size_t)col * qdim;
        s_op.a_n_head = n_head;
        s_op.a_n_kv = n_kv;
        s_op.a_hd = HD;
        s_op.a_n_ctx = n_ctx;
        s_op.a_max_chunks = max_chunks;
        s_op.a_partials = partials + (size_t)col * n_head * max_chunks * (HD + 2);
        s_op.a_counters = counters + (size_t)col * n_head;
    }
    __syncthreads();
    const AttnSmem sm{s_scores, s_prob, s_ml, &s_flag, nullptr, nullptr, nullptr};
    attn_item<HD>(s_op, kvh, h0, g, chunk, n_chunks, t0, n, sm, (int)threadIdx.x, false, nullptr);
}

}  // namespace

bool attn_batch_supported(int head_dim, int n_ctx, int max_chunks) {
    return (head_dim == 64 || head_dim == 128) && (long long)max_chunks * AB_CHUNK >= n_ctx;
}

int launch_attn_batch(const float* q, const __half* kc, const __half* vc, float* out, int n_head, int n_kv, int head_dim, int n_ctx,
                      const int* col_pos, const int* col_slot, int64_t slot_stride, int n_cols, int max_chunks, void* scratch, int cols_cap,
                      cudaStream_t st, const int* kv_table, int kv_pps) {
    if (!q || !kc || !vc || !out || !col_pos || !scratch || n_cols < 1 || cols_cap < n_cols || n_kv <= 0 || n_head % n_kv ||
        (head_dim != 64 && head_dim != 128)) {
        set_error("attn_batch: bad arguments (heads %d/%d, head_dim %d, n_ctx %d, %d chunks, %d columns)", n_head, n_kv, head_dim, n_ctx, max_chunks,
                  n_cols);
        return B2_EINVAL;
    }
    // scratch layout of attn_scratch_bytes_batch(n_head, head_dim, max_chunks, cols_cap)
    size_t part = (size_t)cols_cap * n_head * max_chunks * ((size_t)head_dim + 2) * sizeof(float);
    part = (part + 255) & ~(size_t)255;
    float* partials = (float*)scratch;
    unsigned int* counters = (unsigned int*)((uint8_t*)scratch + part);
    const int gq = n_head / n_kv, nb = (gq + A_MAXG - 1) / A_MAXG;
    const dim3 grid(n_kv * nb, max_chunks, n_cols);
    if (head_dim == 128)
        attn_batch_kernel<128><<<grid, NCT, 0, st>>>(q, kc, vc, out, n_head, n_kv, n_ctx, col_pos, col_slot, (long long)slot_stride, max_chunks, partials,
                                                     counters, kv_table, kv_pps);
    else
        attn_batch_kernel<64><<<grid, NCT, 0, st>>>(q, kc, vc, out, n_head, n_kv, n_ctx, col_pos, col_slot, (long long)slot_stride, max_chunks, partials,
                                                    counters, kv_table, kv_pps);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b2
