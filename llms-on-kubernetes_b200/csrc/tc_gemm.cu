// tc_gemm.cu -- prompt-phase mat-mul: TMA -> shared memory -> tcgen05.mma -> tensor memory -> epilogue.
//
// Replaces ggml_compute_forward_mul_mat for n_tokens >> 1 (the chart's llama-server prompt
// path).  One persistent CTA per SM, six warps with fixed roles:
//   warp 0      TMA producer: per K step one 128 x 64 tile of the dequantised weights and one BN x 64 tile of
//               the activations, both fp16, 128-byte swizzle, into a ring of STAGES shared-memory slots
//   warp 1      allocates tensor memory; one lane issues tcgen05.mma (M = 128 weight rows, N = BN tokens, K = 16)
//               four times per slot and commits the slot back to the producer; two accumulators of BN columns
//               alternate between tiles so the epilogue of tile i overlaps the MMAs of tile i + 1
//   warps 2..5  epilogue: tcgen05.ld 32 lanes x 32 columns, (+ bias, + residual), coalesced fp32 stores --
//               lanes are weight rows, so for one token the 32 lanes of a warp write 128 contiguous bytes
// Tiles are handed out m-major (all token tiles of one 128-row weight tile run at the same time on neighbouring
// CTAs), so a weight tile comes out of HBM once and the activations stay in L2.
//
// Arithmetic: fp16 x fp16 products (exact in fp32) accumulated in fp32.  The block values are dequantised exactly as
// ggml does and rounded once to fp16 (2^-11 relative); the activations go through the reference's own Q8_K / Q8_0
// quantisation first (rows_to_f16_kernel) and d * q is rounded to fp16, so the mat-mul sees the quantised activations
// the reference's integer dot products see.  fp16 rather than the bf16 the task statement names: same tensor-core
// rate, three more mantissa bits -- measured on the tiny models the logit error against the oracle halves
// (DESIGN.md); values are saturated to +-65504 instead of overflowing.
#include "tc_gemm.cuh"

#include <cuda.h>

#include <algorithm>

#include "common.h"
#include "dequant.cuh"
#include "device_utils.cuh"
#include "gemv.cuh"
#include "tc_ptx.cuh"

namespace b2 {
namespace {

using namespace tcptx;

template <int BN>
struct TcGeo {
    static constexpr int STAGES = BN == 256 ? 4 : 6;
    static constexpr int A_BYTES = TC_BM * TC_BK * 2;
    static constexpr int B_BYTES = BN * TC_BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int N_BARS = 2 * STAGES + 4;
    static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /* 1024-byte alignment of the swizzled tiles */ + 8 * N_BARS + 16;
    static constexpr int TMEM_COLS = 2 * BN;   // two accumulators; a power of two >= 32
    static constexpr int THREADS = 192;
    // instruction descriptor (kind::f16): fp32 accumulator (bit 4), fp16 A and B (format 0 at bits 7 and 10), both
    // K-major, N >> 3 at bit 17, M >> 4 at bit 24
    static constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
};

template <int BN>
__global__ void __launch_bounds__(TcGeo<BN>::THREADS, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x, const TcGemmArgs a) {
    using G = TcGeo<BN>;
    extern __shared__ uint8_t tc_smem[];
    const uint32_t tiles0 = (smem_u32(tc_smem) + 1023u) & ~1023u;
    const uint32_t bars = tiles0 + G::STAGES * G::STAGE_BYTES;
    auto full = [&](int s) { return bars + 8u * s; };
    auto empty = [&](int s) { return bars + 8u * (G::STAGES + s); };
    auto acc_full = [&](int i) { return bars + 8u * (2 * G::STAGES + i); };
    auto acc_empty = [&](int i) { return bars + 8u * (2 * G::STAGES + 2 + i); };
    const uint32_t tmem_slot = bars + 8u * G::N_BARS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < G::STAGES; s++) {
            bar_init(full(s), 1);
            bar_init(empty(s), 1);
        }
        for (int i = 0; i < 2; i++) {
            bar_init(acc_full(i), 1);
            bar_init(acc_empty(i), 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" :: "l"(&map_w) : "memory");
        asm volatile("prefetch.tensormap [%0];" :: "l"(&map_x) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(tmem_slot), "n"(G::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    const int tiles_m = (a.N + TC_BM - 1) / TC_BM;
    const int tiles_t = (a.T + BN - 1) / BN;
    // split-K: `splits` CTAs share an output tile, each over n_kb consecutive K steps, partial sums to a.ws
    const int splits = a.splits > 1 ? a.splits : 1;
    const int n_tiles = tiles_m * tiles_t * splits;
    const int n_kb = a.K / TC_BK / splits;

    if (warp == 0) {
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const int ot = tile / splits, k0 = (tile % splits) * n_kb;
                const int m0 = (ot / tiles_t) * TC_BM, t0 = (ot % tiles_t) * BN;
                for (int kb = k0; kb < k0 + n_kb; kb++) {
                    bar_wait(empty(s), ph ^ 1);
                    bar_expect_tx(full(s), G::STAGE_BYTES);
                    const uint32_t sa = tiles0 + s * G::STAGE_BYTES;
                    tma_load_2d(sa, &map_w, full(s), kb * TC_BK, m0);
                    tma_load_2d(sa + G::A_BYTES, &map_x, full(s), kb * TC_BK, t0);
                    if (++s == G::STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            int j = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, j++) {
                const int as = j & 1;
                const uint32_t aph = (j >> 1) & 1;
                bar_wait(acc_empty(as), aph ^ 1);   // the epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
                for (int kb = 0; kb < n_kb; kb++) {
                    bar_wait(full(s), ph);
                    tc_fence_after();
                    const uint32_t sa = tiles0 + s * G::STAGE_BYTES;
                    const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + G::A_BYTES);
#pragma unroll
                    for (int k = 0; k < TC_BK / 16; k++)   // 16 bf16 = 32 bytes further along the swizzled row
                        umma_f16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), G::IDESC, (kb | k) != 0);
                    umma_commit(empty(s));
                    if (++s == G::STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
                umma_commit(acc_full(as));
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3;   // the quarter of the 128 tensor-memory lanes this warp may read
        int j = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, j++) {
            const int ot = tile / splits, ks = tile % splits;
            const int m0 = (ot / tiles_t) * TC_BM, t0 = (ot % tiles_t) * BN;
            const int as = j & 1;
            const uint32_t aph = (j >> 1) & 1;
            bar_wait(acc_full(as), aph);
            tc_fence_after();
            const int n = m0 + q * 32 + lane;
            const bool row_ok = n < a.N;
            const float bias = (a.bias && row_ok && splits == 1) ? a.bias[n] : 0.0f;
            // split-K partial sums go to slice ks of the workspace ([splits][T][N]); splitk_reduce_kernel applies the epilogue
            float* const obase = splits > 1 ? a.ws + (size_t)ks * a.T * a.N : a.out;
            const int64_t ldo = splits > 1 ? (int64_t)a.N : a.ldo;
            const bool add = splits == 1 && a.epi == TC_EPI_ADD;
#pragma unroll 1
            for (int c = 0; c < BN / 32; c++) {
                const int tc0 = t0 + c * 32;
                if (tc0 >= a.T) break;
                uint32_t v[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + c * 32), v);
                if (row_ok) {
                    float* p = obase + (int64_t)tc0 * ldo + n;
                    const int nt = min(32, a.T - tc0);
                    if (add) {
#pragma unroll
                        for (int i = 0; i < 32; i++)
                            if (i < nt) p[(int64_t)i * ldo] += __uint_as_float(v[i]) + bias;
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; i++)
                            if (i < nt) p[(int64_t)i * ldo] = __uint_as_float(v[i]) + bias;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) bar_arrive(acc_empty(as));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(G::TMEM_COLS) : "memory");
    }
}

// out[t][n] (+)= sum over the split-K slices of ws[s][t][n] (+ bias[n]), in slice order (deterministic)
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ ws, int splits, int T, int N, float* __restrict__ out,
                                                            int64_t ldo, const float* __restrict__ bias, int epi) {
    const int64_t n4 = (int64_t)T * N / 4;
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < n4; g += (int64_t)gridDim.x * blockDim.x) {
        const int64_t e = g * 4;
        const int t = (int)(e / N), n = (int)(e - (int64_t)t * N);
        float4 acc = *(const float4*)(ws + e);
        for (int s = 1; s < splits; s++) {
            const float4 v = *(const float4*)(ws + (size_t)s * T * N + e);
            acc.x += v.x;
            acc.y += v.y;
            acc.z += v.z;
            acc.w += v.w;
        }
        if (bias) {
            const float4 b = *(const float4*)(bias + n);
            acc.x += b.x;
            acc.y += b.y;
            acc.z += b.z;
            acc.w += b.w;
        }
        float4* o = (float4*)(out + (int64_t)t * ldo + n);
        if (epi == TC_EPI_ADD) {
            const float4 r = *o;
            acc.x += r.x;
            acc.y += r.y;
            acc.z += r.z;
            acc.w += r.w;
        }
        *o = acc;
    }
}

// ---- elementwise producers of the fp16 operands ----

__device__ __forceinline__ float clamp_h(float v) { return fminf(fmaxf(v, -65504.0f), 65504.0f); }   // saturate instead of inf
__device__ __forceinline__ uint32_t pack_h2s(float a, float b) {
    const __half2 h = __floats2half2_rn(clamp_h(a), clamp_h(b));
    return *(const uint32_t*)&h;
}

// eight consecutive values as fp16; perm: elements 1 and 2 of each group of four swapped (the K order tcq_gemm.cu's fp16
// conversion produces its weight pairs in)
__device__ __forceinline__ uint4 store8(const float (&v)[8], int perm) {
    if (perm) return make_uint4(pack_h2s(v[0], v[2]), pack_h2s(v[1], v[3]), pack_h2s(v[4], v[6]), pack_h2s(v[5], v[7]));
    return make_uint4(pack_h2s(v[0], v[1]), pack_h2s(v[2], v[3]), pack_h2s(v[4], v[5]), pack_h2s(v[6], v[7]));
}

template <int TYPE>
__global__ void dequant_f16_kernel(const uint8_t* __restrict__ src, uint4* __restrict__ dst, int64_t n8) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < n8; g += (int64_t)gridDim.x * blockDim.x) {
        const float4 lo = dq4(TYPE, src, 2 * g), hi = dq4(TYPE, src, 2 * g + 1);
        dst[g] = make_uint4(pack_h2s(lo.x, lo.y), pack_h2s(lo.z, lo.w), pack_h2s(hi.x, hi.y), pack_h2s(hi.z, hi.w));
    }
}

// The reference's activation quantisation REPRODUCED and folded back, 8 consecutive elements per lane of a warp that
// covers a 256-element span: the values are quantised exactly as ggml's quantize_row_q8_K (FMT = ACT_Q8K: one
// 256-element block per warp, iscale = -127 / signed max, nearest-int) or quantize_row_q8_0 (ACT_Q80: 32-element
// blocks = 4 lanes, d = amax / 127 stored as fp16, roundf) would, and replaced by d * q.  d * q has at most 8 + 11
// significant bits, so the only new rounding is the caller's final one to fp16 (2^-11 relative): the mat-mul then
// sees the same quantised activations as the reference's integer dot products.  FMT < 0: values unchanged.
// Warp-collective: every lane calls it (lanes past the row end with zeros).
template <int FMT>
__device__ __forceinline__ void act_requantise8(float (&v)[8]) {
    if (FMT == ACT_Q8K) {
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
        const float wmax = warp_max(amax);
        // the FIRST element of the block attaining the maximum magnitude gives the sign of the scale
        int first = 8;
#pragma unroll
        for (int i = 7; i >= 0; i--)
            if (fabsf(v[i]) == wmax) first = i;
        const uint32_t owners = __ballot_sync(0xffffffffu, first < 8);
        float mval = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++)
            if (i == first) mval = v[i];
        mval = __shfl_sync(0xffffffffu, mval, owners ? __ffs(owners) - 1 : 0);
        if (wmax != 0.0f) {
            const float iscale = -127.0f / mval;
            const float d = 1.0f / iscale;
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = d * (float)min(127, __float2int_rn(iscale * v[i]));
        }
    } else if (FMT == ACT_Q80) {
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        const float d = amax / 127.0f;
        const float id = d != 0.0f ? 1.0f / d : 0.0f;
        const float d16 = __half2float(__float2half_rn(d));
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = d16 * (float)(int)roundf(v[i] * id);
    }
}

// 8 consecutive floats of a mat-mul output: from the output buffer, or summed over the split-K slices in slice order
__device__ __forceinline__ void load8_parts(const float* __restrict__ direct, const TcParts& pp, size_t off, float (&v)[8]) {
    if (pp.splits > 1) {
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = 0.0f;
        for (int s = 0; s < pp.splits; s++) {
            const float4 p0 = *(const float4*)(pp.ws + s * pp.slice + off), p1 = *(const float4*)(pp.ws + s * pp.slice + off + 4);
            v[0] += p0.x; v[1] += p0.y; v[2] += p0.z; v[3] += p0.w; v[4] += p1.x; v[5] += p1.y; v[6] += p1.z; v[7] += p1.w;
        }
    } else {
        const float4 p0 = *(const float4*)(direct + off), p1 = *(const float4*)(direct + off + 4);
        v[0] = p0.x; v[1] = p0.y; v[2] = p0.z; v[3] = p0.w; v[4] = p1.x; v[5] = p1.y; v[6] = p1.z; v[7] = p1.w;
    }
}

// One CTA per row.  (0) a pending split-K residual: x[row] += sum of the slices of the previous mat-mul (its deferred
// epilogue; row stride n in the slices); (1) optional RMS norm (ggml's: fp32 squares summed in double, (x * scale) * w);
// (2) the activation quantisation of the mat-mul that follows (act_requantise8) and the rounding to fp16.
template <int FMT>
__global__ void __launch_bounds__(256) rows_to_f16_kernel(float* __restrict__ x, int64_t ldx, const float* __restrict__ w, float eps, int n,
                                                          __half* __restrict__ y, const TcParts pend, int perm) {
    __shared__ double ss_warp[8];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float* xr = x + (int64_t)blockIdx.x * ldx;
    __half* yr = y + (int64_t)blockIdx.x * n;
    constexpr int RC = 4;   // chunks of 2048 elements a thread keeps in registers: rows up to 8192 make ONE pass over global memory
    if (n <= RC * 2048) {
        // thread tid owns elements tid * 8 + 2048 k ..+8 = lane `lane` of span warp + 8 k: the mapping of all three phases
        float v[RC][8];
        double ss = 0.0;
#pragma unroll
        for (int k = 0; k < RC; k++) {
            const int i0 = tid * 8 + k * 2048;
            if (i0 < n) {
                const float4 p0 = *(const float4*)(xr + i0), p1 = *(const float4*)(xr + i0 + 4);
                v[k][0] = p0.x; v[k][1] = p0.y; v[k][2] = p0.z; v[k][3] = p0.w; v[k][4] = p1.x; v[k][5] = p1.y; v[k][6] = p1.z; v[k][7] = p1.w;
                if (pend.splits > 1) {
                    float d[8];
                    load8_parts(nullptr, pend, (size_t)blockIdx.x * n + i0, d);
#pragma unroll
                    for (int i = 0; i < 8; i++) v[k][i] += d[i];
                    *(float4*)(xr + i0) = make_float4(v[k][0], v[k][1], v[k][2], v[k][3]);
                    *(float4*)(xr + i0 + 4) = make_float4(v[k][4], v[k][5], v[k][6], v[k][7]);
                }
#pragma unroll
                for (int i = 0; i < 8; i++) ss += (double)(v[k][i] * v[k][i]);
            } else {
#pragma unroll
                for (int i = 0; i < 8; i++) v[k][i] = 0.0f;
            }
        }
        float scale = 1.0f;
        if (w) {
            // (the double sum is reduced in another order than the re-reading path below: within fp32 rounding of the mean)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            if (lane == 0) ss_warp[warp] = ss;
            __syncthreads();
            double tot = 0.0;
#pragma unroll
            for (int i = 0; i < 8; i++) tot += ss_warp[i];
            scale = 1.0f / sqrtf((float)(tot / (double)n) + eps);
        }
#pragma unroll
        for (int k = 0; k < RC; k++) {
            const int i0 = tid * 8 + k * 2048;
            if (k * 2048 >= n) break;                 // whole spans only beyond this point (warp-uniform)
            const bool in = i0 < n;
            if (in && w) {
                const float4 w0 = *(const float4*)(w + i0), w1 = *(const float4*)(w + i0 + 4);
                const float g[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int i = 0; i < 8; i++) v[k][i] = (v[k][i] * scale) * g[i];
            }
            if ((warp + 8 * k) * 256 < n) {           // the warp's span exists (its tail lanes may be past the row end)
                act_requantise8<FMT>(v[k]);
                if (in) *(uint4*)(yr + i0) = store8(v[k], perm);
            }
        }
        return;
    }
    if (pend.splits > 1) {
        for (int i0 = tid * 8; i0 < n; i0 += 256 * 8) {
            float v[8];
            load8_parts(nullptr, pend, (size_t)blockIdx.x * n + i0, v);
            float4 p0 = *(const float4*)(xr + i0), p1 = *(const float4*)(xr + i0 + 4);
            p0.x += v[0]; p0.y += v[1]; p0.z += v[2]; p0.w += v[3]; p1.x += v[4]; p1.y += v[5]; p1.z += v[6]; p1.w += v[7];
            *(float4*)(xr + i0) = p0;
            *(float4*)(xr + i0 + 4) = p1;
        }
        __syncthreads();   // the row is re-read below by other threads of this CTA
    }
    float scale = 1.0f;
    if (w) {
        double ss = 0.0;
        for (int i0 = tid * 8; i0 < n; i0 += 256 * 8) {
            const float4 p0 = *(const float4*)(xr + i0), p1 = *(const float4*)(xr + i0 + 4);
            const float t[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
            for (int i = 0; i < 8; i++) ss += (double)(t[i] * t[i]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if (lane == 0) ss_warp[warp] = ss;
        __syncthreads();
        double tot = 0.0;
#pragma unroll
        for (int i = 0; i < 8; i++) tot += ss_warp[i];
        scale = 1.0f / sqrtf((float)(tot / (double)n) + eps);
    }
    // a warp per 256-element span, 8 consecutive elements per lane
    for (int s = warp; s < (n + 255) / 256; s += 8) {
        const int i0 = (s << 8) + lane * 8;
        const bool in = i0 < n;
        float v[8];
        if (in) {
            const float4 p0 = *(const float4*)(xr + i0), p1 = *(const float4*)(xr + i0 + 4);
            v[0] = p0.x; v[1] = p0.y; v[2] = p0.z; v[3] = p0.w; v[4] = p1.x; v[5] = p1.y; v[6] = p1.z; v[7] = p1.w;
            if (w) {
                const float4 w0 = *(const float4*)(w + i0), w1 = *(const float4*)(w + i0 + 4);
                const float g[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int i = 0; i < 8; i++) v[i] = (v[i] * scale) * g[i];
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = 0.0f;
        }
        act_requantise8<FMT>(v);
        if (in) *(uint4*)(yr + i0) = store8(v, perm);
    }
}

// One CTA per row of the stacked gate | up mat-mul output (gu[T][2 d_ff], or its split-K slices): h = silu(gate) * up
// (the formula of the decode kernels' fused epilogue), quantised for the down projection and rounded to fp16; the
// last row's h also in fp32 (h_last, b2_ctx_get_activation)
template <int FMT>
__global__ void __launch_bounds__(256) swiglu_to_f16_kernel(const float* __restrict__ gu, const TcParts parts, int d_ff, int T,
                                                            __half* __restrict__ y, float* __restrict__ h_last, int perm) {
    const int t = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t row = (size_t)t * 2 * d_ff;
    // (the spans are independent: gridDim.y CTAs share a row, so that a 32-row decode batch is more than 32 CTAs)
    for (int s = blockIdx.y * 8 + warp; s < (d_ff + 255) / 256; s += 8 * gridDim.y) {
        const int i0 = (s << 8) + lane * 8;
        const bool in = i0 < d_ff;
        float v[8];
        if (in) {
            float g[8], u[8];
            load8_parts(gu, parts, row + i0, g);
            load8_parts(gu, parts, row + d_ff + i0, u);
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = (g[i] / (1.0f + expf(-g[i]))) * u[i];
            if (h_last && t == T - 1) {
                *(float4*)(h_last + i0) = make_float4(v[0], v[1], v[2], v[3]);
                *(float4*)(h_last + i0 + 4) = make_float4(v[4], v[5], v[6], v[7]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = 0.0f;
        }
        act_requantise8<FMT>(v);
        if (in) *(uint4*)(y + (size_t)t * d_ff + i0) = store8(v, perm);
    }
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (libb2.so does not link libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int encode_fn(EncodeTiledFn* out) {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        B2_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (q != cudaDriverEntryPointSuccess || !p) {
            set_error("tc_gemm: the driver does not export cuTensorMapEncodeTiled");
            return B2_ECUDA;
        }
        fn = (EncodeTiledFn)p;
    }
    *out = fn;
    return 0;
}

// [rows][K] bf16 row-major, boxes of box_rows x 64 elements, 128-byte swizzle, zero fill past the last row
int encode_f16_rows(CUtensorMap* map, const void* base, int64_t rows, int64_t K, int box_rows) {
    EncodeTiledFn fn;
    int rc;
    if ((rc = encode_fn(&fn))) return rc;
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    const cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("tc_gemm: cuTensorMapEncodeTiled failed (%d) for a [%lld][%lld] bf16 matrix", (int)r, (long long)rows, (long long)K);
        return B2_ECUDA;
    }
    return 0;
}

int sm_count() {
    const int n = device_sm_count();   // per device
    return n > 0 ? n : 148;
}

template <int BN>
int launch_tc(const TcGemmArgs& a, cudaStream_t st) {
    using G = TcGeo<BN>;
    CUtensorMap mw, mx;
    int rc;
    if ((rc = encode_f16_rows(&mw, a.w, a.N, a.K, TC_BM))) return rc;
    if ((rc = encode_f16_rows(&mx, a.x, a.T, a.K, BN))) return rc;
    static PerDeviceSize attr_set;   // per device: a second GPU in the same process needs its own opt-in
    const int dev = current_device();
    if (attr_set.below(dev, G::SMEM)) {
        B2_CUDA_CHECK(cudaFuncSetAttribute(tc_gemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM));
        attr_set.set(dev, G::SMEM);
    }
    const int tiles = ((a.N + TC_BM - 1) / TC_BM) * ((a.T + BN - 1) / BN) * (a.splits > 1 ? a.splits : 1);
    const int grid = tiles < sm_count() ? tiles : sm_count();
    tc_gemm_kernel<BN><<<grid, G::THREADS, G::SMEM, st>>>(mw, mx, a);
    B2_CUDA_CHECK(cudaGetLastError());
    if (a.splits > 1 && !a.defer_reduce) {
        const int64_t n4 = (int64_t)a.T * a.N / 4;
        const int blocks = (int)std::min<int64_t>((n4 + 255) / 256, 148 * 8);
        splitk_reduce_kernel<<<blocks, 256, 0, st>>>(a.ws, a.splits, a.T, a.N, a.out, a.ldo, a.bias, a.epi);
        B2_CUDA_CHECK(cudaGetLastError());
    }
    return 0;
}

}  // namespace

// token tile: 256 columns per tile halve the shared-memory traffic per flop; up to 128 rows fit the 128-column tile
static int tc_tile_cols(int T) { return T > 128 ? 256 : 128; }

int tc_gemm_splits(int N, int K, int T) {
    // fewer output tiles than SMs (short prompts, decode batches, narrow matrices): cut K until every SM streams a
    // share of the matrix; each split keeps at least 8 K steps
    const int bn = tc_tile_cols(T);
    const int tiles = ((N + TC_BM - 1) / TC_BM) * ((T + bn - 1) / bn);
    int s = 1;
    while (s < 8 && tiles * s < sm_count() && (K / TC_BK) % (2 * s) == 0 && K / TC_BK / (2 * s) >= 8) s *= 2;
    return s;
}

size_t tc_gemm_ws_bytes(int N, int K, int T) {
    const int s = tc_gemm_splits(N, K, T);
    return s > 1 ? (size_t)s * T * N * sizeof(float) : 0;
}

int launch_tc_gemm(const TcGemmArgs& a_in, cudaStream_t st, TcParts* deferred) {
    TcGemmArgs a = a_in;
    a.splits = 1;
    a.defer_reduce = deferred != nullptr;
    if (deferred) *deferred = TcParts{nullptr, 1, 0};
    if (a.ws && a.N % 4 == 0 && a.ldo % 4 == 0) {
        const int s = tc_gemm_splits(a.N, a.K, a.T);
        if (s > 1 && (size_t)s * a.T * a.N * sizeof(float) <= a.ws_bytes) a.splits = s;
    }
    if (!a.w || !a.x || !a.out || a.N <= 0 || a.K <= 0 || a.T <= 0 || a.K % TC_BK || ((uintptr_t)a.w & 15) || ((uintptr_t)a.x & 15)) {
        set_error("tc_gemm: bad arguments (N=%d K=%d T=%d; K must be a multiple of %d, operands 16-byte aligned)", a.N, a.K, a.T, TC_BK);
        return B2_EINVAL;
    }
    if (deferred && a.splits > 1) *deferred = TcParts{a.ws, a.splits, (size_t)a.T * a.N};
    return tc_tile_cols(a.T) == 256 ? launch_tc<256>(a, st) : launch_tc<128>(a, st);
}

int launch_dequant_f16(int type, const void* src, __half* dst, int64_t n, cudaStream_t st) {
    const TypeTraits* t = type_traits(type);
    if (!t || !src || !dst || n <= 0 || n % t->block_elems || n % 8 || ((uintptr_t)dst & 15)) {
        set_error("dequant_f16: bad arguments (type %d, n=%lld)", type, (long long)n);
        return t ? B2_EINVAL : B2_EUNSUP;
    }
    const int threads = 256;
    const int64_t n8 = n / 8;
    int64_t blocks = (n8 + threads - 1) / threads;
    if (blocks > 148 * 16) blocks = 148 * 16;
    const uint8_t* s = (const uint8_t*)src;
    uint4* d = (uint4*)dst;
    switch (type) {
        case B2_TYPE_F32: dequant_f16_kernel<B2_TYPE_F32><<<(int)blocks, threads, 0, st>>>(s, d, n8); break;
        case B2_TYPE_F16: dequant_f16_kernel<B2_TYPE_F16><<<(int)blocks, threads, 0, st>>>(s, d, n8); break;
        case B2_TYPE_BF16: dequant_f16_kernel<B2_TYPE_BF16><<<(int)blocks, threads, 0, st>>>(s, d, n8); break;
        case B2_TYPE_Q8_0: dequant_f16_kernel<B2_TYPE_Q8_0><<<(int)blocks, threads, 0, st>>>(s, d, n8); break;
        case B2_TYPE_Q4_K: dequant_f16_kernel<B2_TYPE_Q4_K><<<(int)blocks, threads, 0, st>>>(s, d, n8); break;
        case B2_TYPE_Q5_K: dequant_f16_kernel<B2_TYPE_Q5_K><<<(int)blocks, threads, 0, st>>>(s, d, n8); break;
        case B2_TYPE_Q6_K: dequant_f16_kernel<B2_TYPE_Q6_K><<<(int)blocks, threads, 0, st>>>(s, d, n8); break;
        default: set_error("dequant_f16: unsupported ggml type %d", type); return B2_EUNSUP;
    }
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int launch_rows_to_f16(float* x, int64_t ldx, const float* norm_w, float eps, int n, int rows, int act_fmt, __half* y, cudaStream_t st, TcParts pend, bool perm) {
    if (!x || !y || n <= 0 || rows <= 0 || n % 8 || ldx % 4 || ((uintptr_t)x & 15) || ((uintptr_t)y & 15) || (norm_w && ((uintptr_t)norm_w & 15))) {
        set_error("rows_to_f16: bad arguments (n=%d rows=%d)", n, rows);
        return B2_EINVAL;
    }
    if ((act_fmt == ACT_Q8K && n % 256) || (act_fmt == ACT_Q80 && n % 32)) {
        set_error("rows_to_f16: row length %d is not a whole number of activation blocks", n);
        return B2_EINVAL;
    }
    if (act_fmt == ACT_Q8K) rows_to_f16_kernel<ACT_Q8K><<<rows, 256, 0, st>>>(x, ldx, norm_w, eps, n, y, pend, perm ? 1 : 0);
    else if (act_fmt == ACT_Q80) rows_to_f16_kernel<ACT_Q80><<<rows, 256, 0, st>>>(x, ldx, norm_w, eps, n, y, pend, perm ? 1 : 0);
    else rows_to_f16_kernel<-1><<<rows, 256, 0, st>>>(x, ldx, norm_w, eps, n, y, pend, perm ? 1 : 0);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int launch_swiglu_to_f16(const float* gu, TcParts parts, int d_ff, int T, int act_fmt, __half* y, float* h_last, cudaStream_t st, bool perm) {
    if (!gu || !y || d_ff <= 0 || T <= 0 || d_ff % 8 || (act_fmt == ACT_Q8K && d_ff % 256) || (act_fmt == ACT_Q80 && d_ff % 32)) {
        set_error("swiglu_to_f16: bad arguments (d_ff=%d T=%d)", d_ff, T);
        return B2_EINVAL;
    }
    // CTAs per row: enough to fill the GPU with a narrow batch, never fewer than one span per warp
    const int spans = (d_ff + 255) / 256;
    int per_row = std::max(1, std::min((spans + 7) / 8, (2 * 148 + T - 1) / T));
    const dim3 grid(T, per_row);
    if (act_fmt == ACT_Q8K) swiglu_to_f16_kernel<ACT_Q8K><<<grid, 256, 0, st>>>(gu, parts, d_ff, T, y, h_last, perm ? 1 : 0);
    else if (act_fmt == ACT_Q80) swiglu_to_f16_kernel<ACT_Q80><<<grid, 256, 0, st>>>(gu, parts, d_ff, T, y, h_last, perm ? 1 : 0);
    else swiglu_to_f16_kernel<-1><<<grid, 256, 0, st>>>(gu, parts, d_ff, T, y, h_last, perm ? 1 : 0);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b2
