// step_attn.cuh -- single-query attention as a stage of the persistent decode kernel (step.cu).
//
// Work item = (kv head, batch of <= 4 query heads of its group, chunk of n <= 256 tokens); the chunk length is
// chosen per step so that the items cover the grid once.  The 256 compute threads of a CTA run one item:
//   scores : tensor cores (mma.sync m16n8k16, fp16 x fp16 -> fp32): rows = the query heads, columns = 8 tokens per
//            warp and step; a lane reads one quarter of its token's K row with 16-byte loads
//   softmax: the two warps that own a head in the PV phase compute its max and sum from shared memory
//   P.V    : tensor cores again: rows = heads, k = 16 tokens, columns = 8 dims; a warp owns 8 or 16 dims and walks all
//            tokens, V rows reach the B fragments through movmatrix.trans (no shared-memory staging, no reduction)
// Same numerics as attn.cu: q and the probabilities are rounded to fp16, fp32 accumulation; a single chunk
// normalises before rounding (the oracle's order), several chunks keep exp(s - m_chunk) rounded to fp16 and
// merge by log-sum-exp in the last CTA of the (kv head, batch) to arrive.
#pragma once
#include "gemv_device.cuh"
#include "step.cuh"

namespace b2 {
namespace stepattn {

using gemvdev::NCT;
using gemvdev::NCW;

constexpr int A_MAXG = 4;   // query heads per item; larger GQA groups run as several batches
constexpr int A_TOK = STEP_ATTN_MAX_CHUNK;

struct AttnSmem {
    float* scores;   // [A_MAXG][A_TOK]
    float* prob;     // [A_MAXG][A_TOK]
    float* ml;       // [2][A_MAXG] max, sum per head
    int* flag;
    // merge inside a cluster (rank 0 receives): records {max, sum, out[HD]} of ranks 1.., the barrier their bytes complete
    // and the parity of its next phase
    float* inbox;    // [A_MAXCL - 1][A_MAXG][128 + 2]
    uint64_t* abar;
    int* apar;
};
constexpr int A_MAXCL = 4;   // largest cluster the in-cluster merge is laid out for
constexpr size_t ATTN_SMEM_BYTES = (2 * A_MAXG * A_TOK + 2 * A_MAXG + 4 + (A_MAXCL - 1) * A_MAXG * (128 + 2)) * sizeof(float);

// tokens per work item for a sequence of T tokens: one wave of items over the grid
__host__ __device__ inline int attn_chunk_tokens(int T, int items_per_chunk, int grid, int min_chunk) {
    int c = (int)(((long long)T * items_per_chunk + grid - 1) / grid);
    c = (c + 31) & ~31;
    if (c < min_chunk) c = min_chunk;
    if (c > STEP_ATTN_MAX_CHUNK) c = STEP_ATTN_MAX_CHUNK;
    return c;
}

// the words of V row t (zero beyond the chunk) this lane feeds to its warp's TPWV output tiles: dims vcol .. vcol + 2 TPWV
template <int HD, int TPWV>
__device__ __forceinline__ void load_v_words(const __half* vbase, int t, int n, int vcol, uint32_t (&w)[TPWV]) {
    if (TPWV == 2) {
        const uint2 v = t < n ? *(const uint2*)(vbase + (size_t)t * HD + vcol) : make_uint2(0u, 0u);
        w[0] = v.x;
        w[TPWV - 1] = v.y;
    } else {
        w[0] = t < n ? *(const uint32_t*)(vbase + (size_t)t * HD + vcol) : 0u;
    }
}

// heads [h0, h0 + g) of kv head kvh, tokens [t0, t0 + n)
template <int HD>
__device__ void attn_item(const StepOp& op, int kvh, int h0, int g, int chunk, int n_chunks, int t0, int n, const AttnSmem& sm, int tid,
                          bool merge_in_cluster, unsigned long long* tr) {
    constexpr int CPH = NCT / A_MAXG;             // PV phase: threads per head (two warps)
    constexpr int EPT = HD / CPH;                 // columns per thread
    static_assert(EPT == 1 || EPT == 2, "head_dim 64 or 128");
    const int gq = op.a_n_head / op.a_n_kv;
    const int warp = tid >> 5, lane = tid & 31;
    const __half* kbase = op.a_k + ((size_t)kvh * op.a_n_ctx + t0) * HD;
    const __half* vbase = op.a_v + ((size_t)kvh * op.a_n_ctx + t0) * HD;
    // softmax / merge role: head hh and columns c0 (+1)
    const int hh = tid / CPH, cp = tid % CPH, c0 = cp * EPT;

    constexpr int TPWV = HD / 8 / NCW;   // P.V output tiles (8 dims) per warp
    constexpr int V_PRE_KS = 2;          // 16-token steps of V fetched ahead of the score phase (a 32-token chunk whole)
    const int tk = lane >> 2, c = lane & 3;
    const int vcol = (HD / 4) * c + 2 * TPWV * warp;   // first dim of this lane's V words / outputs

    if (tr && tid == 0) tr[4] = trace_ns();
    uint32_t vpre[V_PRE_KS][2][TPWV];

    // ---- scores on the tensor cores: S[head][token] = q[head] . k[token] ----
    // mma.sync m16n8k16 (fp16 operands, fp32 accumulate): rows = the item's query heads (rows >= g are zero),
    // columns = 8 tokens, k = 16 dims per step.  ggml rounds q to fp16 for this product and K is stored fp16, so the
    // operands are exactly the oracle's.  The dims are permuted so that lane (tk = lane / 4, c = lane % 4) feeds all
    // HD / 16 steps from one contiguous quarter of its token's row: dims [DPT c, DPT c + DPT).
    {
        constexpr int NMMA = HD / 16, DPT = HD / 4;
        const float scale = 1.0f / sqrtf((float)HD);
        uint32_t qa[NMMA][2];   // a0, a2 of every step: head tk's dims DPT c + 4 j + {0,1}, {2,3} as half2
        if (tk < g) {
            const float* qp = op.a_q + ((size_t)(kvh * gq + h0 + tk)) * HD + DPT * c;
#pragma unroll
            for (int j = 0; j < NMMA; j++) {
                const float4 q4 = *(const float4*)(qp + 4 * j);
                const __half2 lo = __floats2half2_rn(q4.x, q4.y), hi = __floats2half2_rn(q4.z, q4.w);
                qa[j][0] = *(const uint32_t*)&lo;
                qa[j][1] = *(const uint32_t*)&hi;
            }
        } else {
#pragma unroll
            for (int j = 0; j < NMMA; j++) qa[j][0] = qa[j][1] = 0u;
        }
        if (tr && tid == 0) tr[5] = trace_after(qa[0][0]);   // q arrived
        // the V words of the first 32 tokens are requested once q is here, so that they travel while the scores are
        // computed without standing between the score phase and its own operands
#pragma unroll
        for (int ks = 0; ks < V_PRE_KS; ks++) {
            load_v_words<HD, TPWV>(vbase, 16 * ks + tk, n, vcol, vpre[ks][0]);
            load_v_words<HD, TPWV>(vbase, 16 * ks + 8 + tk, n, vcol, vpre[ks][1]);
        }
        // token groups of 8, dealt over the warps
        for (int tg = warp; tg * 8 < n; tg += NCW) {
            const int t = tg * 8 + tk;
            uint32_t kw[DPT / 2];   // this lane's quarter of the token's K row
#pragma unroll
            for (int i = 0; i < DPT / 8; i++) {
                const uint4 v = t < n ? *(const uint4*)(kbase + (size_t)t * HD + DPT * c + 8 * i) : make_uint4(0, 0, 0, 0);
                kw[4 * i] = v.x; kw[4 * i + 1] = v.y; kw[4 * i + 2] = v.z; kw[4 * i + 3] = v.w;
            }
            if (tr && tid == 0 && tg == 0) tr[6] = trace_after(kw[0]);   // first K rows arrived
            float s4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
            for (int j = 0; j < NMMA; j++)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%5}, {%7,%8}, {%0,%1,%2,%3};"
                             : "+f"(s4[0]), "+f"(s4[1]), "+f"(s4[2]), "+f"(s4[3])
                             : "r"(qa[j][0]), "r"(0u), "r"(qa[j][1]), "r"(kw[2 * j]), "r"(kw[2 * j + 1]));
            // lane (tk, c) holds head tk's scores of tokens 8 tg + 2c, + 1
            if (tk < g) {
                const int ts = tg * 8 + 2 * c;
                if (ts < n) sm.scores[tk * A_TOK + ts] = s4[0] * scale;
                if (ts + 1 < n) sm.scores[tk * A_TOK + ts + 1] = s4[1] * scale;
            }
        }
    }
    if (tr && tid == 0) tr[7] = trace_ns();   // this warp's scores done
    named_bar_sync(1, NCT);
    if (tr && tid == 0) tr[1] = trace_ns();

    // ---- softmax: warps 2h and 2h+1 own head h (both compute the same max and sum) ----
    const bool exact = n_chunks == 1;
    if (hh < g) {
        const float* sc = sm.scores + hh * A_TOK;
        float pmax = -INFINITY, psum = 0.0f;
        for (int t = lane; t < n; t += 32) pmax = fmaxf(pmax, sc[t]);
        pmax = warp_max(pmax);
        for (int t = lane; t < n; t += 32) psum += expf(sc[t] - pmax);
        psum = warp_sum(psum);
        const float inv = 1.0f / psum;
        for (int t = (warp & 1) * 32 + lane; t < n; t += 64) {
            const float e = expf(sc[t] - pmax);
            sm.prob[hh * A_TOK + t] = __half2float(__float2half_rn(exact ? e * inv : e));
        }
        if (tid % CPH == 0) {
            sm.ml[hh] = pmax;
            sm.ml[A_MAXG + hh] = psum;
        }
    }
    named_bar_sync(1, NCT);

    // ---- P.V on the tensor cores: out[head][dim] = sum_t p[head][t] v[t][dim] ----
    // mma.sync m16n8k16: rows = heads, k = 16 tokens per step, columns = 8 dims.  Warp w owns TPWV output tiles; the dims
    // are permuted so that tile j's column 2c' + e is dim DPT c' + 2 j + e: lane (tk, c) then reads its tiles' values of
    // token row tk as one contiguous word (pair) at dim DPT c + 2 j, in the natural layout of an 8 x 8 b16 matrix, and
    // movmatrix.trans turns that into the B fragment (tokens 2c, 2c+1 of column tk).  No cross-warp reduction: every warp
    // walks all tokens for its own dims, and lane (tk, c) ends up with out[head tk][DPT c + 2 j + {0, 1}].
    float o4[TPWV][4];
#pragma unroll
    for (int i = 0; i < TPWV; i++) o4[i][0] = o4[i][1] = o4[i][2] = o4[i][3] = 0.0f;
    {
        const int n_ks = (n + 15) >> 4;
        for (int ks = 0; ks < n_ks; ks++) {
            uint32_t w0[TPWV], w1[TPWV];
            static_assert(V_PRE_KS == 2, "the two prefetched steps are selected with constant indices (registers)");
            if (ks == 0) {
#pragma unroll
                for (int i = 0; i < TPWV; i++) { w0[i] = vpre[0][0][i]; w1[i] = vpre[0][1][i]; }
            } else if (ks == 1) {
#pragma unroll
                for (int i = 0; i < TPWV; i++) { w0[i] = vpre[1][0][i]; w1[i] = vpre[1][1][i]; }
            } else {
                load_v_words<HD, TPWV>(vbase, 16 * ks + tk, n, vcol, w0);
                load_v_words<HD, TPWV>(vbase, 16 * ks + 8 + tk, n, vcol, w1);
            }
            // A: head tk's probabilities of tokens 16 ks + 2c, + 1 (a0) and + 8, + 9 (a2); fp16 values already
            uint32_t a0 = 0u, a2 = 0u;
            if (tk < g) {
                const int ta = 16 * ks + 2 * c;
                const float2 pa = *(const float2*)(sm.prob + tk * A_TOK + ta), pb = *(const float2*)(sm.prob + tk * A_TOK + ta + 8);
                const __half2 ha = __floats2half2_rn(ta < n ? pa.x : 0.0f, ta + 1 < n ? pa.y : 0.0f);
                const __half2 hb = __floats2half2_rn(ta + 8 < n ? pb.x : 0.0f, ta + 9 < n ? pb.y : 0.0f);
                a0 = *(const uint32_t*)&ha;
                a2 = *(const uint32_t*)&hb;
            }
#pragma unroll
            for (int i = 0; i < TPWV; i++) {
                uint32_t b0, b1;
                asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(b0) : "r"(w0[i]));
                asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(b1) : "r"(w1[i]));
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%5}, {%7,%8}, {%0,%1,%2,%3};"
                             : "+f"(o4[i][0]), "+f"(o4[i][1]), "+f"(o4[i][2]), "+f"(o4[i][3])
                             : "r"(a0), "r"(0u), "r"(a2), "r"(b0), "r"(b1));
            }
        }
    }
    const size_t stride = (size_t)HD + 2;
    if (!exact && merge_in_cluster) {
        // ---- the chunks of this (kv head, batch) sit in the CTAs of one cluster: rank `chunk` sends its record to rank 0
        // through distributed shared memory (st.async: the bytes complete rank 0's barrier), rank 0 merges by
        // log-sum-exp.  No global-memory round trip (record stores, fence, arrival atomic, record loads). ----
        constexpr uint32_t REC = (uint32_t)(HD + 2) * 4;          // bytes of one head's record
        if (chunk != 0) {
            if (tk < g) {
                float* rec = sm.inbox + ((size_t)(chunk - 1) * A_MAXG + tk) * (HD + 2);   // same offset in rank 0's shared memory
                const uint32_t bar = dsmem_addr(sm.abar, 0u);
#pragma unroll
                for (int i = 0; i < TPWV; i++)
                    dsmem_st_async_v2(dsmem_addr(rec + 2 + vcol + 2 * i, 0u), __float_as_uint(o4[i][0]), __float_as_uint(o4[i][1]), bar);
                if (warp == 0 && c == 0) dsmem_st_async_v2(dsmem_addr(rec, 0u), __float_as_uint(sm.ml[tk]), __float_as_uint(sm.ml[A_MAXG + tk]), bar);
            }
        } else {
            const uint32_t parity = (uint32_t)*sm.apar & 1u;
            if (tid == 0) mbar_arrive_expect_tx(sm.abar, (uint32_t)(n_chunks - 1) * (uint32_t)g * REC);
            mbar_wait(sm.abar, parity);
            if (tk < g) {
                float m = sm.ml[tk];
                for (int s = 0; s + 1 < n_chunks; s++) m = fmaxf(m, sm.inbox[((size_t)s * A_MAXG + tk) * (HD + 2)]);
                const float w0 = expf(sm.ml[tk] - m);
                float den = sm.ml[A_MAXG + tk] * w0;
                float num[TPWV][2];
#pragma unroll
                for (int i = 0; i < TPWV; i++) { num[i][0] = o4[i][0] * w0; num[i][1] = o4[i][1] * w0; }
                for (int s = 0; s + 1 < n_chunks; s++) {
                    const float* rec = sm.inbox + ((size_t)s * A_MAXG + tk) * (HD + 2);
                    const float ws = expf(rec[0] - m);
                    den += rec[1] * ws;
#pragma unroll
                    for (int i = 0; i < TPWV; i++) {
                        const float2 v = *(const float2*)(rec + 2 + vcol + 2 * i);
                        num[i][0] += v.x * ws;
                        num[i][1] += v.y * ws;
                    }
                }
                float* o = op.a_out + (size_t)(kvh * gq + h0 + tk) * HD;
#pragma unroll
                for (int i = 0; i < TPWV; i++) *(float2*)(o + vcol + 2 * i) = make_float2(num[i][0] / den, num[i][1] / den);
            }
        }
        named_bar_sync(1, NCT);
        if (chunk == 0 && tid == 0) *sm.apar ^= 1;   // the barrier's next phase
        if (tr && tid == 0) tr[2] = trace_ns();
        return;
    }
    if (tk < g) {
        const int head = kvh * gq + h0 + tk;
        // one chunk: the result; several: this chunk's record {max, sum, out[HD]}
        float* o = exact ? op.a_out + (size_t)head * HD : op.a_partials + ((size_t)head * op.a_max_chunks + chunk) * stride + 2;
#pragma unroll
        for (int i = 0; i < TPWV; i++) *(float2*)(o + vcol + 2 * i) = make_float2(o4[i][0], o4[i][1]);
        if (!exact && warp == 0 && c == 0) {
            o[-2] = sm.ml[tk];
            o[-1] = sm.ml[A_MAXG + tk];
        }
    }
    if (exact) {
        named_bar_sync(1, NCT);   // the next item reuses the shared buffers
        if (tr && tid == 0) tr[2] = trace_ns();
        return;
    }
    // the last chunk of the (kv head, batch) to arrive merges the records by log-sum-exp
    __threadfence();
    named_bar_sync(1, NCT);
    if (tr && tid == 0) tr[2] = trace_ns();
    if (tid == 0) {
        // one arrival counter per (kv head, head batch): at most n_head of them
        unsigned int* ctr = &op.a_counters[kvh * ((gq + A_MAXG - 1) / A_MAXG) + h0 / A_MAXG];
        const unsigned int arrived = atomicAdd(ctr, 1u) + 1u;
        const int last = arrived == (unsigned int)n_chunks;
        *sm.flag = last;
        if (last) *ctr = 0u;
    }
    named_bar_sync(1, NCT);
    if (*sm.flag && hh < g) {
        __threadfence();
        const int head = kvh * gq + h0 + hh;
        const float* base = op.a_partials + ((size_t)head * op.a_max_chunks) * stride;
        float m = -INFINITY;
        for (int s = 0; s < n_chunks; s++) m = fmaxf(m, __ldcg(base + s * stride));
        float den = 0.0f, n0 = 0.0f, n1 = 0.0f;
        for (int s = 0; s < n_chunks; s++) {
            const float w = expf(__ldcg(base + s * stride) - m);
            den += __ldcg(base + s * stride + 1) * w;
            n0 += __ldcg(base + s * stride + 2 + c0) * w;
            if (EPT == 2) n1 += __ldcg(base + s * stride + 2 + c0 + 1) * w;
        }
        float* o = op.a_out + (size_t)head * HD + c0;
        o[0] = n0 / den;
        if (EPT == 2) o[1] = n1 / den;
    }
    named_bar_sync(1, NCT);
}

}  // namespace stepattn
}  // namespace b2
