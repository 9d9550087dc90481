// prefill_attn.cu -- causal attention of a whole prompt chunk over the fp16 KV cache (SURVEY 8a row a14 with
// n_tokens = T: ggml's soft_max + two mul_mats per head, restated as one streaming pass).
//
// One CTA = one query head x 64 consecutive prompt tokens, four warps of 16 query rows.  Keys and values of the
// head's kv head stream through shared memory in blocks of 64 tokens (cp.async, double buffered, 16-byte chunks
// XOR-swizzled by the key index so ldmatrix reads are conflict-free).  Per block and warp:
//   S = Q K^T   mma.sync m16n8k16 f16 (Q rounded to fp16 as on the decode path), fp32 scores, * 1/sqrt(head_dim)
//   causal mask (key position <= query position), running row max / sum (fp32), P = exp(S - max) rounded to fp16
//   O = O * exp(max_old - max_new) + P V   (V through ldmatrix.trans)
// and O / sum at the end.  Same numerics as the chunked decode attention (attn.cu, step_attn.cuh): probabilities
// are rounded to fp16 relative to the running maximum, the normaliser sums the unrounded fp32 exponentials.
// Query blocks are issued longest first (the last block of the prompt sees the most keys).
#include <cuda_fp16.h>

#include "common.h"
#include "device_utils.cuh"
#include "ops.h"

namespace b2 {
namespace {

constexpr int PA_BQ = 64, PA_BK = 64, PA_THREADS = 128;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;   // 0: the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_f16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *(const uint32_t*)&h;
}

// q: [T][n_head * HD] fp32 (rotated); kc, vc: this slot's and layer's cache [n_kv][n_ctx][HD]; out: [T][n_head * HD]
template <int HD>
__global__ void __launch_bounds__(PA_THREADS, HD == 128 ? 3 : 4)
prefill_attn_kernel(const float* __restrict__ q, const __half* __restrict__ kc, const __half* __restrict__ vc, float* __restrict__ out, int n_head,
                    int n_kv, int n_ctx, int T, int pos0, float scale, const int* __restrict__ pages) {
    constexpr int CH = HD / 8;                    // 16-byte chunks per row
    constexpr int TILE = PA_BK * HD * 2;          // bytes of one K or V block
    extern __shared__ __align__(128) uint8_t pa_smem[];
    const uint32_t sbase = smem_u32(pa_smem);     // [2 buffers][K | V]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
    const int qb = gridDim.x - 1 - blockIdx.x, head = blockIdx.y;
    const int kvh = head / (n_head / n_kv);
    const int qdim = n_head * HD;
    const __half* kh = kc + (size_t)kvh * n_ctx * HD;
    const __half* vh = vc + (size_t)kvh * n_ctx * HD;
    const int t_last = min(T, qb * PA_BQ + PA_BQ) - 1;
    const int kv_end = pos0 + t_last + 1;         // keys [0, kv_end) exist
    const int n_kb = (kv_end + PA_BK - 1) / PA_BK;

    auto load_block = [&](int kb, int buf) {
        const uint32_t kdst = sbase + buf * 2 * TILE, vdst = kdst + TILE;
#pragma unroll
        for (int i = 0; i < PA_BK * CH / PA_THREADS; i++) {
            const int idx = tid + i * PA_THREADS;
            const int key = idx / CH, ch = idx % CH;
            const int kabs = kb * PA_BK + key;
            const bool ok = kabs < kv_end;
            // paged cache: position kabs is row pages[kabs / 256] * 256 + kabs % 256 (a 64-key block never straddles a page)
            const int krow = !ok ? 0 : pages ? pages[kabs >> 8] * 256 + (kabs & 255) : kabs;
            const size_t src = (size_t)krow * HD + ch * 8;
            const uint32_t off = (uint32_t)(key * HD * 2 + ((ch ^ (key & 7)) << 4));
            cp_async16(kdst + off, kh + src, ok);
            cp_async16(vdst + off, vh + src, ok);
        }
    };

    // this warp's 16 query rows as A fragments (fp16), rows g and g + 8
    const int t0 = qb * PA_BQ + warp * 16 + g, t1 = t0 + 8;
    uint32_t qa[HD / 16][4];
    {
        const float* q0 = q + (size_t)min(t0, T - 1) * qdim + head * HD;
        const float* q1 = q + (size_t)min(t1, T - 1) * qdim + head * HD;
#pragma unroll
        for (int ks = 0; ks < HD / 16; ks++) {
            const float2 a = *(const float2*)(q0 + 16 * ks + 2 * c), b = *(const float2*)(q1 + 16 * ks + 2 * c);
            const float2 a8 = *(const float2*)(q0 + 16 * ks + 8 + 2 * c), b8 = *(const float2*)(q1 + 16 * ks + 8 + 2 * c);
            qa[ks][0] = pack_h2(a.x, a.y);
            qa[ks][1] = pack_h2(b.x, b.y);
            qa[ks][2] = pack_h2(a8.x, a8.y);
            qa[ks][3] = pack_h2(b8.x, b8.y);
        }
    }
    const int p0 = pos0 + min(t0, T - 1), p1 = pos0 + min(t1, T - 1);   // key positions <= p are visible
    const int p_warp = pos0 + min(qb * PA_BQ + warp * 16 + 15, T - 1);

    float o[HD / 8][4];
#pragma unroll
    for (int j = 0; j < HD / 8; j++)
#pragma unroll
        for (int e = 0; e < 4; e++) o[j][e] = 0.0f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.0f, l1 = 0.0f;

    load_block(0, 0);
    cp_async_commit();
    for (int kb = 0; kb < n_kb; kb++) {
        const int buf = kb & 1;
        if (kb + 1 < n_kb) load_block(kb + 1, buf ^ 1);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();
        if (kb * PA_BK <= p_warp) {
            const uint32_t ks_base = sbase + buf * 2 * TILE, vs_base = ks_base + TILE;
            // ---- S = Q K^T ----
            float s[PA_BK / 8][4];
#pragma unroll
            for (int j = 0; j < PA_BK / 8; j++) {
#pragma unroll
                for (int e = 0; e < 4; e++) s[j][e] = 0.0f;
                const int key = 8 * j + (lane & 7);
#pragma unroll
                for (int ks = 0; ks < HD / 16; ks += 2) {
                    const int ch = 2 * ks + (lane >> 3);
                    uint32_t b0, b1, b2, b3;
                    ldsm_x4(ks_base + (uint32_t)(key * HD * 2 + ((ch ^ (key & 7)) << 4)), b0, b1, b2, b3);
                    mma_f16(s[j], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], b0, b1);
                    mma_f16(s[j], qa[ks + 1][0], qa[ks + 1][1], qa[ks + 1][2], qa[ks + 1][3], b2, b3);
                }
            }
            // ---- scale, causal mask, running max ----
            float bm0 = -INFINITY, bm1 = -INFINITY;
#pragma unroll
            for (int j = 0; j < PA_BK / 8; j++) {
                const int k0 = kb * PA_BK + 8 * j + 2 * c;
                s[j][0] = k0 <= p0 ? s[j][0] * scale : -INFINITY;
                s[j][1] = k0 + 1 <= p0 ? s[j][1] * scale : -INFINITY;
                s[j][2] = k0 <= p1 ? s[j][2] * scale : -INFINITY;
                s[j][3] = k0 + 1 <= p1 ? s[j][3] * scale : -INFINITY;
                bm0 = fmaxf(bm0, fmaxf(s[j][0], s[j][1]));
                bm1 = fmaxf(bm1, fmaxf(s[j][2], s[j][3]));
            }
            bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 1));
            bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 2));
            bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 1));
            bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 2));
            const float mn0 = fmaxf(m0, bm0), mn1 = fmaxf(m1, bm1);   // finite from block 0 on: key 0 is visible to every row
            const float r0 = expf(m0 - mn0), r1 = expf(m1 - mn1);     // exp(-inf) = 0 on the first block
            m0 = mn0;
            m1 = mn1;
            float bs0 = 0.0f, bs1 = 0.0f;
            uint32_t pa[PA_BK / 8][2];
#pragma unroll
            for (int j = 0; j < PA_BK / 8; j++) {
                const float e0 = expf(s[j][0] - mn0), e1 = expf(s[j][1] - mn0), e2 = expf(s[j][2] - mn1), e3 = expf(s[j][3] - mn1);
                bs0 += e0 + e1;
                bs1 += e2 + e3;
                pa[j][0] = pack_h2(e0, e1);
                pa[j][1] = pack_h2(e2, e3);
            }
            l0 = l0 * r0 + bs0;
            l1 = l1 * r1 + bs1;
#pragma unroll
            for (int j = 0; j < HD / 8; j++) {
                o[j][0] *= r0;
                o[j][1] *= r0;
                o[j][2] *= r1;
                o[j][3] *= r1;
            }
            // ---- O += P V ----
#pragma unroll
            for (int kk = 0; kk < PA_BK / 16; kk++) {
                const int key = 16 * kk + ((lane >> 3) & 1) * 8 + (lane & 7);
#pragma unroll
                for (int jd = 0; jd < HD / 8; jd += 2) {
                    const int ch = jd + (lane >> 4);
                    uint32_t b0, b1, b2, b3;
                    ldsm_x4_t(vs_base + (uint32_t)(key * HD * 2 + ((ch ^ (key & 7)) << 4)), b0, b1, b2, b3);
                    mma_f16(o[jd], pa[2 * kk][0], pa[2 * kk][1], pa[2 * kk + 1][0], pa[2 * kk + 1][1], b0, b1);
                    mma_f16(o[jd + 1], pa[2 * kk][0], pa[2 * kk][1], pa[2 * kk + 1][0], pa[2 * kk + 1][1], b2, b3);
                }
            }
        }
        __syncthreads();
    }
    // ---- normalise and store ----
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.0f / l0, i1 = 1.0f / l1;
#pragma unroll
    for (int j = 0; j < HD / 8; j++) {
        if (t0 < T) *(float2*)(out + (size_t)t0 * qdim + head * HD + 8 * j + 2 * c) = make_float2(o[j][0] * i0, o[j][1] * i0);
        if (t1 < T) *(float2*)(out + (size_t)t1 * qdim + head * HD + 8 * j + 2 * c) = make_float2(o[j][2] * i1, o[j][3] * i1);
    }
}

}  // namespace

int launch_prefill_attn(const float* q, const __half* kc, const __half* vc, float* out, int n_head, int n_kv, int head_dim, int n_ctx, int T, int pos0,
                        cudaStream_t st, const int* pages) {
    if (!q || !kc || !vc || !out || T <= 0 || pos0 < 0 || (!pages && pos0 + T > n_ctx) || n_kv <= 0 || n_head % n_kv) {
        set_error("prefill_attn: bad arguments (T=%d pos0=%d n_ctx=%d heads %d/%d)", T, pos0, n_ctx, n_head, n_kv);
        return B2_EINVAL;
    }
    const dim3 grid((T + PA_BQ - 1) / PA_BQ, n_head);
    const float scale = 1.0f / sqrtf((float)head_dim);
    if (head_dim == 128) {
        constexpr int smem = 4 * PA_BK * 128 * 2;
        B2_CUDA_CHECK(cudaFuncSetAttribute(prefill_attn_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        prefill_attn_kernel<128><<<grid, PA_THREADS, smem, st>>>(q, kc, vc, out, n_head, n_kv, n_ctx, T, pos0, scale, pages);
    } else if (head_dim == 64) {
        constexpr int smem = 4 * PA_BK * 64 * 2;
        prefill_attn_kernel<64><<<grid, PA_THREADS, smem, st>>>(q, kc, vc, out, n_head, n_kv, n_ctx, T, pos0, scale, pages);
    } else {
        set_error("prefill_attn: head_dim %d is not built (64, 128)", head_dim);
        return B2_EUNSUP;
    }
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b2
