// gemm.cu -- batched quantised mat-mul on the tensor cores (mma.sync int8), see gemm.cuh.
//
// Pipeline of one launch (same roles as gemv.cu):
//   producer warp : per (tile of 16 rows, K-chunk of 2048): 16 bulk copies of row segments + one bulk copy
//                   of the chunk's quantised activations into the stage
//   decoder warp  : unpacks the 6-bit Q4_K scale headers of the stage once
//   compute warps : 8 (8 columns) or 16 in two groups that split the columns (16 / 32 columns); warp w of a group
//                   owns super-block w of the chunk: nibbles -> u8 A fragments (16 rows x 32 k, k permuted so that
//                   one 8-byte load feeds both halves of the fragment), activations -> s8 B fragments (32 k x 8
//                   columns), one MMA per sub-block and column group, the s32 results scaled per sub-block with
//                   the 6-bit scales; the "mins" term is two more MMAs against the 32-element sums of the
//                   activations (split into 7 low bits and the rest so that they fit int8)
//   epilogue warp : adds the warps' shares and runs the fused epilogue per (row pair, column); K / V rows go to the
//                   paged cache through the page table
// Q5_K adds the fifth bit to the A fragments, Q6_K uses m16n8k16 s8 x s8 with one MMA per 16-element scale group,
// Q8_0 (Q8_0 activations, K-chunks of 1024) one MMA per 32-element block.  Serves decode batches below 24 columns
// (and every batch <= 32 when the model's fp16 copy is not resident); wider batches and prompts: tc_gemm.cu.
#include "gemm.cuh"

#include <cstdlib>

#include "common.h"
#include "device_utils.cuh"
#include "gemv_device.cuh"

namespace b2 {

using gemvdev::block_bytes_of;
using gemvdev::decode_k4_header;
using gemvdev::pair_rows;

namespace {

constexpr int NCW = 8, NCT = NCW * 32;
constexpr int ROW_SLOT = GEMM_CHUNK_BLOCKS * 210 + 16;      // largest row segment of a chunk + alignment slack
constexpr int W_STAGE = GEMM_TILE_ROWS * ROW_SLOT;           // 27136
constexpr int DEC_ROW = GEMM_CHUNK_BLOCKS * 32 + 16;         // decoded Q4_K headers of one row (padded against bank conflicts)
constexpr int DEC_BYTES = GEMM_TILE_ROWS * DEC_ROW;

// FMT = activation format: ACT_Q8K (K-quant weights: chunks of 8 super-blocks) or ACT_Q80 (Q8_0 weights: 32 blocks of 32)
template <int NG, int FMT>
struct Geo {
    static constexpr int NB = 8 * NG;
    // 16 / 32 columns (one CTA per SM): two groups of 8 compute warps, each group takes half of the columns
    static constexpr int WG = NG >= 2 ? 2 : 1;
    static constexpr int NGL = NG / WG;            // column groups per compute warp
    static constexpr int NCWT = NCW * WG;          // compute warps
    static constexpr int THREADS = (NCWT + 3) * 32;   // + producer, header decoder, epilogue warps
    static constexpr int CB = FMT == ACT_Q8K ? GEMM_CHUNK_BLOCKS : GEMM80_CHUNK_BLOCKS;   // weight blocks per chunk and row
    static constexpr int BLK = FMT == ACT_Q8K ? 256 : 32;                                   // elements per weight block
    static constexpr int ACT_BYTES = FMT == ACT_Q8K ? (int)((size_t)NB * GEMM_Q_PITCH + (size_t)GEMM_CHUNK_BLOCKS * NB * 20)
                                                    : (int)((size_t)NB * GEMM80_Q_PITCH + (size_t)GEMM80_CHUNK_BLOCKS * NB * 4);
    static constexpr int STAGE_BYTES = (W_STAGE + ACT_BYTES + 127) & ~127;
    static constexpr int STAGES = NG == 2 ? 3 : 2;   // NG 1: two CTAs per SM; NG 2, 4: one CTA per SM
    static constexpr int RED_FLOATS = NCW * GEMM_TILE_ROWS * NB;   // one buffer of partial sums
    static constexpr int RED_BUFS = NG == 4 ? 1 : 2;
};

// ---- activation quantisation: one CTA per column ----------------------------------------------------------
// Q8_K exactly as the mat-vec prologue (quantize_row_q8_K_ref: iscale = -127 / max with the sign of the first
// largest element, nearest-int), written in the chunked layout of gemm.cuh.
__global__ void __launch_bounds__(256) act_quant_kernel(const float* __restrict__ x, int64_t x_stride, const float* __restrict__ norm_w,
                                                        float eps, int K, int cols, int nb_cols, uint8_t* __restrict__ act) {
    __shared__ double ss_warp[8];
    const int col = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const size_t chunk_bytes = gemm_act_chunk_bytes(nb_cols);
    const int nspans = K >> 8;
    const bool live = col < cols;
    const float* xc = x + (size_t)col * x_stride;
    float scale = 1.0f;
    if (live && norm_w) {
        // ggml rms_norm: fp32 squares accumulated in double, mean, 1/sqrtf(mean + eps)
        double ss = 0.0;
        for (int i0 = tid * 8; i0 < K; i0 += 256 * 8) {
            const float4 p0 = *(const float4*)(xc + i0), p1 = *(const float4*)(xc + i0 + 4);
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
        for (int w = 0; w < 8; w++) tot += ss_warp[w];
        scale = 1.0f / sqrtf((float)(tot / (double)K) + eps);
    }
    for (int s = warp; s < nspans; s += 8) {
        const int i0 = (s << 8) + lane * 8;
        float v[8];
        if (live) {
            const float4 p0 = *(const float4*)(xc + i0), p1 = *(const float4*)(xc + i0 + 4);
            v[0] = p0.x; v[1] = p0.y; v[2] = p0.z; v[3] = p0.w; v[4] = p1.x; v[5] = p1.y; v[6] = p1.z; v[7] = p1.w;
            if (norm_w) {
                const float4 w0 = *(const float4*)(norm_w + i0), w1 = *(const float4*)(norm_w + i0 + 4);
                const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int i = 0; i < 8; i++) v[i] = (v[i] * scale) * w[i];
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = 0.0f;
        }
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
        const float wmax = warp_max(amax);
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
        int q[8];
        float d = 0.0f;
        if (wmax != 0.0f) {
            const float iscale = -127.0f / mval;
#pragma unroll
            for (int i = 0; i < 8; i++) q[i] = min(127, __float2int_rn(iscale * v[i]));
            d = 1.0f / iscale;
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) q[i] = 0;
        }
        int s8 = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) s8 += q[i];
        int s32 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);
        s32 += __shfl_xor_sync(0xffffffffu, s32, 2);
        const uint32_t pk0 = (uint32_t)(q[0] & 0xff) | ((uint32_t)(q[1] & 0xff) << 8) | ((uint32_t)(q[2] & 0xff) << 16) | ((uint32_t)(q[3] & 0xff) << 24);
        const uint32_t pk1 = (uint32_t)(q[4] & 0xff) | ((uint32_t)(q[5] & 0xff) << 8) | ((uint32_t)(q[6] & 0xff) << 16) | ((uint32_t)(q[7] & 0xff) << 24);
        const int chunk = s / GEMM_CHUNK_BLOCKS, kb = s % GEMM_CHUNK_BLOCKS;
        uint8_t* base = act + (size_t)chunk * chunk_bytes;
        *(uint2*)(base + (size_t)col * GEMM_Q_PITCH + kb * 256 + lane * 8) = make_uint2(pk0, pk1);
        uint8_t* tail = base + (size_t)nb_cols * GEMM_Q_PITCH;
        if (lane == 0) ((float*)tail)[kb * nb_cols + col] = d;
        if ((lane & 3) == 0) {
            uint8_t* bl = tail + (size_t)GEMM_CHUNK_BLOCKS * nb_cols * 4;
            uint8_t* bh = bl + (size_t)GEMM_CHUNK_BLOCKS * nb_cols * 8;
            const int sub = lane >> 2;
            bl[(kb * nb_cols + col) * 8 + sub] = (uint8_t)(s32 & 127);
            bh[(kb * nb_cols + col) * 8 + sub] = (uint8_t)(int8_t)(s32 >> 7);
        }
    }
}

// Q8_0 activations (quantize_row_q8_0: d = amax / 127 per 32 elements, q = roundf(x / d), d kept fp16-rounded)
__global__ void __launch_bounds__(256) act_quant80_kernel(const float* __restrict__ x, int64_t x_stride, const float* __restrict__ norm_w,
                                                          float eps, int K, int cols, int nb_cols, uint8_t* __restrict__ act) {
    __shared__ double ss_warp[8];
    const int col = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const size_t chunk_bytes = gemm80_act_chunk_bytes(nb_cols);
    const bool live = col < cols;
    const float* xc = x + (size_t)col * x_stride;
    float scale = 1.0f;
    if (live && norm_w) {
        double ss = 0.0;
        for (int i0 = tid * 8; i0 < K; i0 += 256 * 8) {
            const float4 p0 = *(const float4*)(xc + i0), p1 = *(const float4*)(xc + i0 + 4);
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
        for (int w = 0; w < 8; w++) tot += ss_warp[w];
        scale = 1.0f / sqrtf((float)(tot / (double)K) + eps);
    }
    // 8 consecutive elements per lane, 4 lanes per 32-element block (K is a multiple of 32)
    for (int w0 = warp * 256; w0 < K; w0 += 256 * 8) {   // warp-uniform bound: the shuffles below need the whole warp
        const int i0 = w0 + lane * 8;
        const bool in_range = i0 < K;
        float v[8];
        if (live && in_range) {
            const float4 p0 = *(const float4*)(xc + i0), p1 = *(const float4*)(xc + i0 + 4);
            v[0] = p0.x; v[1] = p0.y; v[2] = p0.z; v[3] = p0.w; v[4] = p1.x; v[5] = p1.y; v[6] = p1.z; v[7] = p1.w;
            if (norm_w) {
                const float4 w0 = *(const float4*)(norm_w + i0), w1 = *(const float4*)(norm_w + i0 + 4);
                const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int i = 0; i < 8; i++) v[i] = (v[i] * scale) * w[i];
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = 0.0f;
        }
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        float d = amax / 127.0f;
        const float id = d != 0.0f ? 1.0f / d : 0.0f;
        int q[8];
#pragma unroll
        for (int i = 0; i < 8; i++) q[i] = (int)roundf(v[i] * id);
        d = __half2float(__float2half_rn(d));
        const uint32_t pk0 = (uint32_t)(q[0] & 0xff) | ((uint32_t)(q[1] & 0xff) << 8) | ((uint32_t)(q[2] & 0xff) << 16) | ((uint32_t)(q[3] & 0xff) << 24);
        const uint32_t pk1 = (uint32_t)(q[4] & 0xff) | ((uint32_t)(q[5] & 0xff) << 8) | ((uint32_t)(q[6] & 0xff) << 16) | ((uint32_t)(q[7] & 0xff) << 24);
        if (in_range) {
            const int chunk = i0 / GEMM80_CHUNK_K, kin = i0 % GEMM80_CHUNK_K;
            uint8_t* base = act + (size_t)chunk * chunk_bytes;
            *(uint2*)(base + (size_t)col * GEMM80_Q_PITCH + kin) = make_uint2(pk0, pk1);
            if ((lane & 3) == 0) ((float*)(base + (size_t)nb_cols * GEMM80_Q_PITCH))[(kin >> 5) * nb_cols + col] = d;
        }
    }
}

// ---- tensor-core primitives ---------------------------------------------------------------------------------
// D(16x8, s32) = A(16x32, u8, row) . B(32x8, s8, col).  Lane (g = lane/4, c = lane%4) holds
//   a0 = A[g][4c..4c+3], a1 = A[g+8][4c..], a2 = A[g][16+4c..], a3 = A[g+8][16+4c..]
//   b0 = B[4c..4c+3][g], b1 = B[16+4c..][g];  d0 = D[g][2c], d1 = D[g][2c+1], d2 = D[g+8][2c], d3 = D[g+8][2c+1]
__device__ __forceinline__ void mma_u8s8_k32(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}
// D(16x8) = A(16x16, s8) . B(16x8, s8): a0 = A[g][4c..], a1 = A[g+8][4c..], b0 = B[4c..][g]
__device__ __forceinline__ void mma_s8s8_k16(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%7,%7,%7};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a0), "r"(a1), "r"(b0), "r"(0));
}

// 4 bytes from shared memory at a 2-byte aligned address
__device__ __forceinline__ uint32_t lds32_a2(const uint8_t* p) {
    const uint32_t a = smem_u32(p);
    const uint32_t* w = (const uint32_t*)(p - (a & 3u));
    return __funnelshift_r(w[0], w[1], (a & 2u) << 3);
}

// source pointer of tile slot `slot` (0..15: pair slot/2, row slot&1) or nullptr if it does not exist
__device__ __forceinline__ const uint8_t* slot_row_ptr16(const GemvSeg& s, int ptile, int slot, int row_bytes) {
    const int p = ptile * GEMM_TILE_PAIRS + (slot >> 1);
    if (p >= s.n_pairs) return nullptr;
    int ra, rb;
    pair_rows(s, p, &ra, &rb);
    const int r = (slot & 1) ? rb : ra;
    if (r < 0) return nullptr;
    const uint8_t* w = (s.pairing == PAIR_SPLIT && (slot & 1)) ? s.wb : s.wa;
    return w + (size_t)r * row_bytes;
}

// views into one pipeline stage's activation chunk
template <int NB>
struct ActChunk {
    const int8_t* q;      // [NB][GEMM_Q_PITCH]
    const float* d;       // [8][NB]
    const uint8_t* bl;    // [8][NB][8]
    const uint8_t* bh;    // [8][NB][8]
    __device__ explicit ActChunk(const uint8_t* base) {
        q = (const int8_t*)base;
        d = (const float*)(base + (size_t)NB * GEMM_Q_PITCH);
        bl = (const uint8_t*)(d + GEMM_CHUNK_BLOCKS * NB);
        bh = bl + GEMM_CHUNK_BLOCKS * NB * 8;
    }
};

// Q4_K super-block kb of rows g and g+8 of the stage against NG groups of 8 columns
// NGL column groups starting at group ng0 of the NB columns
template <int NGL, int NB>
__device__ __forceinline__ void q4k_block_mma(const uint8_t* blk0, const uint8_t* blk1, const uint8_t* rec0, const uint8_t* rec1, int kb,
                                              const ActChunk<NB>& x, int ng0, int g, int c, float (&acc)[NGL][4]) {
    constexpr int NG = NGL;
    const uint4 w0 = *(const uint4*)(rec0 + 16), w1 = *(const uint4*)(rec1 + 16);   // (sc[2p], sc[2p+1], m[2p], m[2p+1]) per word
    const uint32_t ws0[4] = {w0.x, w0.y, w0.z, w0.w}, ws1[4] = {w1.x, w1.y, w1.z, w1.w};
    int ilo[NG][4], ihi[NG][4];
#pragma unroll
    for (int n = 0; n < NG; n++)
#pragma unroll
        for (int j = 0; j < 4; j++) ilo[n][j] = ihi[n][j] = 0;
#pragma unroll
    for (int p = 0; p < 4; p++) {
        // 8 bytes of qs: the low nibbles are elements 8c..8c+7 of sub-block 2p, the high nibbles of sub-block 2p+1
        const uint2 qa = *(const uint2*)(blk0 + 16 + 32 * p + 8 * c), qb = *(const uint2*)(blk1 + 16 + 32 * p + 8 * c);
        const int s0l = ws0[p] & 0xff, s0h = (ws0[p] >> 8) & 0xff, s1l = ws1[p] & 0xff, s1h = (ws1[p] >> 8) & 0xff;
#pragma unroll
        for (int n = 0; n < NG; n++) {
            const int8_t* xq = x.q + (size_t)((ng0 + n) * 8 + g) * GEMM_Q_PITCH + kb * 256 + 64 * p + 8 * c;
            const uint2 xl = *(const uint2*)xq, xh = *(const uint2*)(xq + 32);
            int dl[4], dh[4];
            mma_u8s8_k32(dl, qa.x & 0x0F0F0F0Fu, qb.x & 0x0F0F0F0Fu, qa.y & 0x0F0F0F0Fu, qb.y & 0x0F0F0F0Fu, xl.x, xl.y);
            mma_u8s8_k32(dh, qa.x & 0xF0F0F0F0u, qb.x & 0xF0F0F0F0u, qa.y & 0xF0F0F0F0u, qb.y & 0xF0F0F0F0u, xh.x, xh.y);
            ilo[n][0] += s0l * dl[0]; ilo[n][1] += s0l * dl[1]; ilo[n][2] += s1l * dl[2]; ilo[n][3] += s1l * dl[3];
            ihi[n][0] += s0h * dh[0]; ihi[n][1] += s0h * dh[1]; ihi[n][2] += s1h * dh[2]; ihi[n][3] += s1h * dh[3];
        }
    }
    // mins: M[16 rows][8 subs] . S[8 subs][8 cols] with S = 32-element sums split as 128 * bh + bl
    const uint32_t m0 = c == 0 ? __byte_perm(ws0[0], ws0[1], 0x7632) : (c == 1 ? __byte_perm(ws0[2], ws0[3], 0x7632) : 0u);
    const uint32_t m1 = c == 0 ? __byte_perm(ws1[0], ws1[1], 0x7632) : (c == 1 ? __byte_perm(ws1[2], ws1[3], 0x7632) : 0u);
    const float2 dd0 = *(const float2*)rec0, dd1 = *(const float2*)rec1;   // (d, dmin) of rows g, g+8
#pragma unroll
    for (int n = 0; n < NG; n++) {
        const int col = (ng0 + n) * 8 + g;
        const uint32_t bl = c < 2 ? *(const uint32_t*)(x.bl + (kb * NB + col) * 8 + 4 * c) : 0u;
        const uint32_t bh = c < 2 ? *(const uint32_t*)(x.bh + (kb * NB + col) * 8 + 4 * c) : 0u;
        int cl[4], ch[4];
        mma_u8s8_k32(cl, m0, m1, 0u, 0u, bl, 0u);
        mma_u8s8_k32(ch, m0, m1, 0u, 0u, bh, 0u);
        const float2 dx = *(const float2*)(x.d + kb * NB + (ng0 + n) * 8 + 2 * c);   // columns 2c, 2c+1
        // the high-nibble dots are exact multiples of 16
        acc[n][0] += dx.x * (dd0.x * (float)(ilo[n][0] + (ihi[n][0] >> 4)) - dd0.y * (float)(cl[0] + 128 * ch[0]));
        acc[n][1] += dx.y * (dd0.x * (float)(ilo[n][1] + (ihi[n][1] >> 4)) - dd0.y * (float)(cl[1] + 128 * ch[1]));
        acc[n][2] += dx.x * (dd1.x * (float)(ilo[n][2] + (ihi[n][2] >> 4)) - dd1.y * (float)(cl[2] + 128 * ch[2]));
        acc[n][3] += dx.y * (dd1.x * (float)(ilo[n][3] + (ihi[n][3] >> 4)) - dd1.y * (float)(cl[3] + 128 * ch[3]));
    }
}

// Q5_K super-block: as Q4_K with the fifth bit of element j of sub-blocks 2p / 2p+1 in bits 2p / 2p+1 of qh[j]
// (block = d, dmin, 12 scale bytes, qh[32], qs[128]).  The values reach 31, so the high nibbles are shifted down
// instead of being multiplied in place.
template <int NGL, int NB>
__device__ __forceinline__ void q5k_block_mma(const uint8_t* blk0, const uint8_t* blk1, const uint8_t* rec0, const uint8_t* rec1, int kb,
                                              const ActChunk<NB>& x, int ng0, int g, int c, float (&acc)[NGL][4]) {
    const uint4 w0 = *(const uint4*)(rec0 + 16), w1 = *(const uint4*)(rec1 + 16);
    const uint32_t ws0[4] = {w0.x, w0.y, w0.z, w0.w}, ws1[4] = {w1.x, w1.y, w1.z, w1.w};
    const uint2 ha = *(const uint2*)(blk0 + 16 + 8 * c), hb = *(const uint2*)(blk1 + 16 + 8 * c);   // qh[8c .. 8c+7] of rows g, g+8
    int ilo[NGL][4], ihi[NGL][4];
#pragma unroll
    for (int n = 0; n < NGL; n++)
#pragma unroll
        for (int j = 0; j < 4; j++) ilo[n][j] = ihi[n][j] = 0;
#pragma unroll
    for (int p = 0; p < 4; p++) {
        const uint2 qa = *(const uint2*)(blk0 + 48 + 32 * p + 8 * c), qb = *(const uint2*)(blk1 + 48 + 32 * p + 8 * c);
        // fifth bits moved to bit 4 of every byte: bit 2p for the low sub-block, bit 2p+1 for the high one
        auto lo5 = [p](uint32_t q, uint32_t h) { return (q & 0x0F0F0F0Fu) | ((p < 2 ? h << (4 - 2 * p) : h >> (2 * p - 4)) & 0x10101010u); };
        auto hi5 = [p](uint32_t q, uint32_t h) { return ((q >> 4) & 0x0F0F0F0Fu) | ((p < 2 ? h << (3 - 2 * p) : h >> (2 * p - 3)) & 0x10101010u); };
        const uint32_t al0 = lo5(qa.x, ha.x), al1 = lo5(qb.x, hb.x), al2 = lo5(qa.y, ha.y), al3 = lo5(qb.y, hb.y);
        const uint32_t ah0 = hi5(qa.x, ha.x), ah1 = hi5(qb.x, hb.x), ah2 = hi5(qa.y, ha.y), ah3 = hi5(qb.y, hb.y);
        const int s0l = ws0[p] & 0xff, s0h = (ws0[p] >> 8) & 0xff, s1l = ws1[p] & 0xff, s1h = (ws1[p] >> 8) & 0xff;
#pragma unroll
        for (int n = 0; n < NGL; n++) {
            const int8_t* xq = x.q + (size_t)((ng0 + n) * 8 + g) * GEMM_Q_PITCH + kb * 256 + 64 * p + 8 * c;
            const uint2 xl = *(const uint2*)xq, xh = *(const uint2*)(xq + 32);
            int dl[4], dh[4];
            mma_u8s8_k32(dl, al0, al1, al2, al3, xl.x, xl.y);
            mma_u8s8_k32(dh, ah0, ah1, ah2, ah3, xh.x, xh.y);
            ilo[n][0] += s0l * dl[0]; ilo[n][1] += s0l * dl[1]; ilo[n][2] += s1l * dl[2]; ilo[n][3] += s1l * dl[3];
            ihi[n][0] += s0h * dh[0]; ihi[n][1] += s0h * dh[1]; ihi[n][2] += s1h * dh[2]; ihi[n][3] += s1h * dh[3];
        }
    }
    const uint32_t m0 = c == 0 ? __byte_perm(ws0[0], ws0[1], 0x7632) : (c == 1 ? __byte_perm(ws0[2], ws0[3], 0x7632) : 0u);
    const uint32_t m1 = c == 0 ? __byte_perm(ws1[0], ws1[1], 0x7632) : (c == 1 ? __byte_perm(ws1[2], ws1[3], 0x7632) : 0u);
    const float2 dd0 = *(const float2*)rec0, dd1 = *(const float2*)rec1;
#pragma unroll
    for (int n = 0; n < NGL; n++) {
        const int col = (ng0 + n) * 8 + g;
        const uint32_t bl = c < 2 ? *(const uint32_t*)(x.bl + (kb * NB + col) * 8 + 4 * c) : 0u;
        const uint32_t bh = c < 2 ? *(const uint32_t*)(x.bh + (kb * NB + col) * 8 + 4 * c) : 0u;
        int cl[4], ch[4];
        mma_u8s8_k32(cl, m0, m1, 0u, 0u, bl, 0u);
        mma_u8s8_k32(ch, m0, m1, 0u, 0u, bh, 0u);
        const float2 dx = *(const float2*)(x.d + kb * NB + (ng0 + n) * 8 + 2 * c);
        acc[n][0] += dx.x * (dd0.x * (float)(ilo[n][0] + ihi[n][0]) - dd0.y * (float)(cl[0] + 128 * ch[0]));
        acc[n][1] += dx.y * (dd0.x * (float)(ilo[n][1] + ihi[n][1]) - dd0.y * (float)(cl[1] + 128 * ch[1]));
        acc[n][2] += dx.x * (dd1.x * (float)(ilo[n][2] + ihi[n][2]) - dd1.y * (float)(cl[2] + 128 * ch[2]));
        acc[n][3] += dx.y * (dd1.x * (float)(ilo[n][3] + ihi[n][3]) - dd1.y * (float)(cl[3] + 128 * ch[3]));
    }
}

// Q6_K super-block: 16 scale groups of 16 elements, one m16n8k16 MMA each; weights as signed q - 32
template <int NGL, int NB>
__device__ __forceinline__ void q6k_block_mma(const uint8_t* blk0, const uint8_t* blk1, int kb, const ActChunk<NB>& x, int ng0, int g, int c,
                                              float (&acc)[NGL][4]) {
    constexpr int NG = NGL;
    uint32_t sw0[4], sw1[4];   // the 16 int8 scales of each row
#pragma unroll
    for (int i = 0; i < 4; i++) {
        sw0[i] = lds32_a2(blk0 + 192 + 4 * i);
        sw1[i] = lds32_a2(blk1 + 192 + 4 * i);
    }
    int isum[NG][4];
#pragma unroll
    for (int n = 0; n < NG; n++)
#pragma unroll
        for (int j = 0; j < 4; j++) isum[n][j] = 0;
#pragma unroll
    for (int h = 0; h < 2; h++)
#pragma unroll
        for (int is = 0; is < 2; is++) {
            const int lq = 16 * is + 4 * c;
            const uint32_t la0 = lds32_a2(blk0 + 64 * h + lq), lb0 = lds32_a2(blk0 + 64 * h + 32 + lq), qh0 = lds32_a2(blk0 + 128 + 32 * h + lq);
            const uint32_t la1 = lds32_a2(blk1 + 64 * h + lq), lb1 = lds32_a2(blk1 + 64 * h + 32 + lq), qh1 = lds32_a2(blk1 + 128 + 32 * h + lq);
#pragma unroll
            for (int j = 0; j < 4; j++) {
                // v = nibble | two high bits << 4, then v - 32 as int8: t = v ^ 0x20; t + 6 * (t & 0x20) sign-extends bit 5
                const uint32_t n0 = ((j & 1) ? lb0 : la0) >> (j >= 2 ? 4 : 0), n1 = ((j & 1) ? lb1 : la1) >> (j >= 2 ? 4 : 0);
                const uint32_t h0 = j == 0 ? qh0 << 4 : (j == 1 ? qh0 << 2 : (j == 2 ? qh0 : qh0 >> 2));
                const uint32_t h1 = j == 0 ? qh1 << 4 : (j == 1 ? qh1 << 2 : (j == 2 ? qh1 : qh1 >> 2));
                const uint32_t t0 = ((n0 & 0x0F0F0F0Fu) | (h0 & 0x30303030u)) ^ 0x20202020u;
                const uint32_t t1 = ((n1 & 0x0F0F0F0Fu) | (h1 & 0x30303030u)) ^ 0x20202020u;
                const uint32_t a0 = t0 + 6u * (t0 & 0x20202020u), a1 = t1 + 6u * (t1 & 0x20202020u);
                const int si = 8 * h + 2 * j + is;
                const int sc0 = (int)(int8_t)(sw0[si >> 2] >> (8 * (si & 3))), sc1 = (int)(int8_t)(sw1[si >> 2] >> (8 * (si & 3)));
                const int kofs = 128 * h + 32 * j + 16 * is + 4 * c;
#pragma unroll
                for (int n = 0; n < NG; n++) {
                    const uint32_t xw = *(const uint32_t*)(x.q + (size_t)((ng0 + n) * 8 + g) * GEMM_Q_PITCH + kb * 256 + kofs);
                    int d4[4];
                    mma_s8s8_k16(d4, a0, a1, xw);
                    isum[n][0] += sc0 * d4[0]; isum[n][1] += sc0 * d4[1]; isum[n][2] += sc1 * d4[2]; isum[n][3] += sc1 * d4[3];
                }
            }
        }
    const float d0 = h2f_bits(*(const uint16_t*)(blk0 + 208)), d1 = h2f_bits(*(const uint16_t*)(blk1 + 208));
#pragma unroll
    for (int n = 0; n < NG; n++) {
        const float2 dx = *(const float2*)(x.d + kb * NB + (ng0 + n) * 8 + 2 * c);
        acc[n][0] += dx.x * (d0 * (float)isum[n][0]);
        acc[n][1] += dx.y * (d0 * (float)isum[n][1]);
        acc[n][2] += dx.x * (d1 * (float)isum[n][2]);
        acc[n][3] += dx.y * (d1 * (float)isum[n][3]);
    }
}

// Q8_0: one 34-byte block (fp16 d, 32 int8) per row and MMA; blk0 / blk1 = rows g and g+8 (2-byte aligned)
template <int NB>
struct ActChunk80 {
    const int8_t* q;   // [NB][GEMM80_Q_PITCH]
    const float* d;    // [32][NB]
    __device__ explicit ActChunk80(const uint8_t* base) {
        q = (const int8_t*)base;
        d = (const float*)(base + (size_t)NB * GEMM80_Q_PITCH);
    }
};
// D = A(16x32 s8) . B(32x8 s8)
__device__ __forceinline__ void mma_s8s8_k32(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}
template <int NGL, int NB>
__device__ __forceinline__ void q80_block_mma(const uint8_t* blk0, const uint8_t* blk1, int kb, const ActChunk80<NB>& x, int ng0, int g, int c,
                                              float (&acc)[NGL][4]) {
    constexpr int NG = NGL;
    // bytes 8c .. 8c+7 of the block's 32 values (k permuted as in the K-quant path)
    const uint32_t a0 = lds32_a2(blk0 + 2 + 8 * c), a2 = lds32_a2(blk0 + 6 + 8 * c);
    const uint32_t a1 = lds32_a2(blk1 + 2 + 8 * c), a3 = lds32_a2(blk1 + 6 + 8 * c);
    const float d0 = h2f_bits(*(const uint16_t*)blk0), d1 = h2f_bits(*(const uint16_t*)blk1);
#pragma unroll
    for (int n = 0; n < NG; n++) {
        const uint2 xb = *(const uint2*)(x.q + (size_t)((ng0 + n) * 8 + g) * GEMM80_Q_PITCH + kb * 32 + 8 * c);
        int d4[4];
        mma_s8s8_k32(d4, a0, a1, a2, a3, xb.x, xb.y);
        const float2 dx = *(const float2*)(x.d + kb * NB + (ng0 + n) * 8 + 2 * c);
        // ggml: sumi * (d_w * d_x)
        acc[n][0] += (float)d4[0] * (d0 * dx.x);
        acc[n][1] += (float)d4[1] * (d0 * dx.y);
        acc[n][2] += (float)d4[2] * (d1 * dx.x);
        acc[n][3] += (float)d4[3] * (d1 * dx.y);
    }
}

// cache row of (slot, kv head, position): through the page table when there is one
__device__ __forceinline__ __half* kv_row(const GemmArgs& a, const GemvSeg& s, int slot, int head, int pos) {
    if (a.kv_table) {
        const size_t tok = (size_t)a.kv_table[slot * a.kv_pps + (pos >> 8)] * 256 + (pos & 255);
        return s.cache + ((size_t)head * s.n_ctx + tok) * s.head_dim;
    }
    return s.cache + (size_t)slot * a.slot_stride + ((size_t)head * s.n_ctx + pos) * s.head_dim;
}

// fused epilogue of one (row pair, column): the mat-vec's, with per-column position, slot and output row
__device__ __forceinline__ void epilogue_col(const GemmArgs& a, const GemvSeg& s, int p, int ra, int rb, float va, float vb, int j, int pos,
                                             int slot) {
    if (s.bias) {
        va += s.bias[ra];
        if (rb >= 0) vb += (s.pairing == PAIR_SPLIT ? s.bias_b : s.bias)[rb];
    }
    float* out = s.out ? s.out + (size_t)j * a.out_stride : nullptr;
    switch (s.epi) {
        case EPI_STORE:
            out[ra] = va;
            if (rb >= 0) out[rb] = vb;
            break;
        case EPI_RESID:
            out[ra] += va;
            if (rb >= 0) out[rb] += vb;
            break;
        case EPI_SWIGLU:
            out[p] = (va / (1.0f + expf(-va))) * vb;
            break;
        case EPI_ROPE_Q:
        case EPI_ROPE_K: {
            const int hd = s.head_dim;
            const int head = ra / hd;
            const int i = s.pairing == PAIR_NEOX ? ra - head * hd : (ra - head * hd) >> 1;
            float oa = va, ob = vb;
            if (i < a.rope_half) {
                const float2 cs = a.rope_tab[(size_t)pos * a.rope_half + i];
                oa = va * cs.x - vb * cs.y;
                ob = va * cs.y + vb * cs.x;
            }
            if (s.epi == EPI_ROPE_Q) {
                out[ra] = oa;
                out[rb] = ob;
            } else {
                __half* dst = kv_row(a, s, slot, head, pos);
                dst[ra - head * hd] = __float2half_rn(oa);
                dst[rb - head * hd] = __float2half_rn(ob);
            }
            break;
        }
        case EPI_V: {
            const int hd = s.head_dim;
            const int head = ra / hd;
            __half* dst = kv_row(a, s, slot, head, pos);
            dst[ra - head * hd] = __float2half_rn(va);
            if (rb >= 0) dst[rb - head * hd] = __float2half_rn(vb);
            break;
        }
    }
}

template <int NG, int FMT>
__global__ void __launch_bounds__((Geo<NG, FMT>::THREADS), NG == 1 ? 2 : 1) gemm_kernel(const __grid_constant__ GemmArgs a) {
    using G = Geo<NG, FMT>;
    constexpr int NB = G::NB;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem;                                      // STAGES x { 16 row slots | activation chunk }
    uint8_t* dec = ring + G::STAGES * G::STAGE_BYTES;          // STAGES x decoded Q4_K headers
    float* red = (float*)(dec + G::STAGES * DEC_BYTES);        // RED_BUFS x [8 warps][16 rows][NB] partial sums
    uint64_t* full = (uint64_t*)(red + G::RED_BUFS * G::RED_FLOATS);
    uint64_t* ready = full + G::STAGES;
    uint64_t* empty = ready + G::STAGES;
    uint64_t* part_full = empty + G::STAGES;
    uint64_t* part_free = part_full + 2;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < G::STAGES; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&ready[s], 1);
            mbar_init(&empty[s], G::NCWT);
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(&part_full[i], G::NCWT);
            mbar_init(&part_free[i], 1);
        }
        mbar_fence_init();
    }
    __syncthreads();

    const int K = a.K;
    const int nblk_row = K / G::BLK;
    const int n_chunks = (nblk_row + G::CB - 1) / G::CB;

    if (warp == G::NCWT) {   // producer
        int stage = 0, phase = 0, si = 0;
        const uint64_t stream_policy = l2_policy_evict_first();
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
            while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
            const GemvSeg& s = a.seg[si];
            const int bb = block_bytes_of(s.type);
            const uint8_t* rowp = lane < GEMM_TILE_ROWS ? slot_row_ptr16(s, tile - s.tile0, lane, nblk_row * bb) : nullptr;
            for (int c = 0; c < n_chunks; c++) {
                const int nb = min(G::CB, nblk_row - c * G::CB);
                const uint8_t* src = nullptr;
                uint32_t sz = 0;
                if (rowp) {
                    const uintptr_t lo = (uintptr_t)(rowp + (size_t)c * G::CB * bb);
                    const uintptr_t hi = lo + (size_t)nb * bb;
                    src = (const uint8_t*)(lo & ~(uintptr_t)15);
                    sz = (uint32_t)(((hi + 15) & ~(uintptr_t)15) - (uintptr_t)src);
                } else if (lane == GEMM_TILE_ROWS) {
                    src = a.act + (size_t)c * G::ACT_BYTES;   // this chunk's quantised activations (stay in L2: default policy)
                    sz = G::ACT_BYTES;
                }
                uint32_t total = sz;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
                if (lane == 0) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    mbar_arrive_expect_tx(&full[stage], total);
                }
                __syncwarp();
                uint8_t* st = ring + stage * G::STAGE_BYTES;
                if (lane < GEMM_TILE_ROWS) {
                    if (sz) bulk_g2s_hint(st + lane * ROW_SLOT, src, sz, &full[stage], stream_policy);
                } else if (lane == GEMM_TILE_ROWS) {
                    bulk_g2s(st + W_STAGE, src, sz, &full[stage]);
                }
                if (++stage == G::STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    if (warp == G::NCWT + 1) {   // header decoder
        int stage = 0, phase = 0, si = 0;
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
            while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
            const int type = a.seg[si].type;
            for (int c = 0; c < n_chunks; c++) {
                const int nb = min(G::CB, nblk_row - c * G::CB);
                mbar_wait(&full[stage], phase);
                if (type == B2_TYPE_Q4_K || type == B2_TYPE_Q5_K) {
                    const int bb = block_bytes_of(type);
                    for (int i = lane; i < GEMM_TILE_ROWS * nb; i += 32) {
                        const int rr = i / nb, b = i - rr * nb;
                        uint8_t* rec = dec + stage * DEC_BYTES + rr * DEC_ROW + b * 32;
                        decode_k4_header(ring + stage * G::STAGE_BYTES + rr * ROW_SLOT + b * bb, rec, rec + 16);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&ready[stage]);
                if (++stage == G::STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    if (warp == G::NCWT + 2) {   // epilogue
        // lane (p = lane / 4, cq = lane % 4) finishes row pair p of the tile for columns n * 8 + 2 cq, + 1
        const int p_in = lane >> 2, cq = lane & 3;
        int pos[NG][2], slot[NG][2];
#pragma unroll
        for (int n = 0; n < NG; n++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int j = n * 8 + 2 * cq + e;
                pos[n][e] = (j < a.cols.n && a.cols.pos) ? a.cols.pos[j] : 0;
                slot[n][e] = (j < a.cols.n && a.cols.slot) ? a.cols.slot[j] : 0;
            }
        int parity = 0, ph = 0, si = 0;
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
            while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
            const GemvSeg& s = a.seg[si];
            mbar_wait(&part_full[parity], (ph >> parity) & 1);
            const float* rp = red + (G::RED_BUFS == 2 ? parity : 0) * G::RED_FLOATS;
            float va[NG][2], vb[NG][2];
#pragma unroll
            for (int n = 0; n < NG; n++) {
                va[n][0] = va[n][1] = vb[n][0] = vb[n][1] = 0.0f;
#pragma unroll
                for (int w = 0; w < NCW; w++) {
                    const float2 ta = *(const float2*)(rp + (w * GEMM_TILE_ROWS + 2 * p_in) * NB + n * 8 + 2 * cq);
                    const float2 tb = *(const float2*)(rp + (w * GEMM_TILE_ROWS + 2 * p_in + 1) * NB + n * 8 + 2 * cq);
                    va[n][0] += ta.x; va[n][1] += ta.y;
                    vb[n][0] += tb.x; vb[n][1] += tb.y;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&part_free[parity]);
            ph ^= 1 << parity;
            if (G::RED_BUFS == 2) parity ^= 1;
            const int p = (tile - s.tile0) * GEMM_TILE_PAIRS + p_in;
            if (p < s.n_pairs) {
                int ra, rb;
                pair_rows(s, p, &ra, &rb);
#pragma unroll
                for (int n = 0; n < NG; n++)
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const int j = n * 8 + 2 * cq + e;
                        if (j < a.cols.n) epilogue_col(a, s, p, ra, rb, va[n][e], vb[n][e], j, pos[n][e], slot[n][e]);
                    }
            }
        }
        return;
    }

    // ===== compute warps: warp wb of a group owns super-block wb of every chunk; group grp its share of the columns =====
    constexpr int NGL = G::NGL;
    const int g = lane >> 2, c = lane & 3;
    const int grp = warp / NCW, wb = warp % NCW, ng0 = grp * NGL;
    int stage = 0, phase = 0, parity = 0, fph = 3, si = 0;
    for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
        while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
        const GemvSeg& s = a.seg[si];
        const int type = s.type;
        const int bb = block_bytes_of(type);
        // row segment offset inside its slot (source address mod 16; chunks are multiples of 16 bytes for all types)
        uint32_t off0 = 0, off1 = 0;
        if (type != B2_TYPE_Q4_K && type != B2_TYPE_Q5_K) {
            off0 = (uint32_t)((uintptr_t)slot_row_ptr16(s, tile - s.tile0, g, nblk_row * bb) & 15);
            off1 = (uint32_t)((uintptr_t)slot_row_ptr16(s, tile - s.tile0, g + 8, nblk_row * bb) & 15);
        }
        float acc[NGL][4];
#pragma unroll
        for (int n = 0; n < NGL; n++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[n][j] = 0.0f;
        for (int ch = 0; ch < n_chunks; ch++) {
            const int nb = min(G::CB, nblk_row - ch * G::CB);
            mbar_wait(&ready[stage], phase);
            if (wb < nb) {
                const uint8_t* st = ring + stage * G::STAGE_BYTES;
                const uint8_t* row0 = st + g * ROW_SLOT + off0 + wb * bb;
                const uint8_t* row1 = st + (g + 8) * ROW_SLOT + off1 + wb * bb;
                if (FMT == ACT_Q8K) {
                    const ActChunk<NB> x(st + W_STAGE);
                    if (type == B2_TYPE_Q4_K) {
                        const uint8_t* dr = dec + stage * DEC_BYTES;
                        q4k_block_mma<NGL, NB>(row0, row1, dr + g * DEC_ROW + wb * 32, dr + (g + 8) * DEC_ROW + wb * 32, wb, x, ng0, g, c, acc);
                    } else if (type == B2_TYPE_Q5_K) {
                        const uint8_t* dr = dec + stage * DEC_BYTES;
                        q5k_block_mma<NGL, NB>(row0, row1, dr + g * DEC_ROW + wb * 32, dr + (g + 8) * DEC_ROW + wb * 32, wb, x, ng0, g, c, acc);
                    } else {
                        q6k_block_mma<NGL, NB>(row0, row1, wb, x, ng0, g, c, acc);
                    }
                } else {
                    // 32 blocks of 32 per chunk: warp wb takes blocks wb, wb + 8, wb + 16, wb + 24
                    const ActChunk80<NB> x(st + W_STAGE);
                    for (int kb = wb; kb < nb; kb += NCW) q80_block_mma<NGL, NB>(row0 + (kb - wb) * 34, row1 + (kb - wb) * 34, kb, x, ng0, g, c, acc);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (++stage == G::STAGES) { stage = 0; phase ^= 1; }
        }
        // hand this warp's share of the tile (16 rows x its columns) to the epilogue warp
        mbar_wait(&part_free[parity], (fph >> parity) & 1);
        fph ^= 1 << parity;
        float* rp = red + (G::RED_BUFS == 2 ? parity : 0) * G::RED_FLOATS + wb * GEMM_TILE_ROWS * NB;
#pragma unroll
        for (int n = 0; n < NGL; n++) {
            *(float2*)(rp + g * NB + (ng0 + n) * 8 + 2 * c) = make_float2(acc[n][0], acc[n][1]);
            *(float2*)(rp + (g + 8) * NB + (ng0 + n) * 8 + 2 * c) = make_float2(acc[n][2], acc[n][3]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&part_full[parity]);
        if (G::RED_BUFS == 2) parity ^= 1;
    }
}

template <int NG, int FMT>
size_t gemm_smem_bytes() {
    using G = Geo<NG, FMT>;
    return (size_t)G::STAGES * (G::STAGE_BYTES + DEC_BYTES) + (size_t)G::RED_BUFS * G::RED_FLOATS * sizeof(float) +
           (3 * G::STAGES + 4) * sizeof(uint64_t);
}

template <int NG, int FMT>
int launch_gemm_ng(const GemmArgs& a, cudaStream_t st) {
    static PerDeviceSize configured;   // the opt-in shared-memory size is a per-device attribute of the kernel
    const size_t smem = gemm_smem_bytes<NG, FMT>();
    const int dev = current_device();
    if (configured.below(dev, smem)) {
        B2_CUDA_CHECK(cudaFuncSetAttribute(gemm_kernel<NG, FMT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured.set(dev, smem);
    }
    const int sms = device_sm_count();
    if (sms <= 0) return B2_ECUDA;
    const int ctas = sms * (NG == 1 ? 2 : 1);
    const int grid = a.n_tiles < ctas ? a.n_tiles : ctas;
    gemm_kernel<NG, FMT><<<grid, Geo<NG, FMT>::THREADS, smem, st>>>(a);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace

int gemm_finalize(GemmArgs& a) {
    if (a.n_seg < 1 || a.n_seg > 3) {
        set_error("gemm: n_seg=%d out of range", a.n_seg);
        return B2_EINVAL;
    }
    a.act_fmt = gemm_act_for_type(a.seg[0].type);
    if (a.act_fmt < 0) {
        set_error("gemm: ggml type %d is not computed by the batched kernel (Q4_K, Q5_K, Q6_K, Q8_0)", a.seg[0].type);
        return B2_EUNSUP;
    }
    const int kmult = a.act_fmt == ACT_Q8K ? 256 : 32;
    if (a.K <= 0 || a.K % kmult) {
        set_error("gemm: K=%d must be a positive multiple of %d", a.K, kmult);
        return B2_EINVAL;
    }
    if (a.cols.n < 1 || a.cols.n > GEMM_MAX_COLS) {
        set_error("gemm: %d columns (1..%d)", a.cols.n, GEMM_MAX_COLS);
        return B2_EINVAL;
    }
    if ((uintptr_t)a.act & 15) {
        set_error("gemm: activation buffer must be 16-byte aligned");
        return B2_EINVAL;
    }
    int tiles = 0;
    for (int i = 0; i < a.n_seg; i++) {
        GemvSeg& s = a.seg[i];
        if (gemm_act_for_type(s.type) != a.act_fmt) {
            set_error("gemm: segment %d is ggml type %d; the batched kernel computes Q4_K / Q5_K / Q6_K (together) and Q8_0", i, s.type);
            return B2_EUNSUP;
        }
        const uintptr_t need = (s.type == B2_TYPE_Q4_K || s.type == B2_TYPE_Q5_K) ? 15 : 1;
        if (((uintptr_t)s.wa & need) || (s.pairing == PAIR_SPLIT && ((uintptr_t)s.wb & need))) {
            set_error("gemm: weight pointer of segment %d is misaligned for type %d", i, s.type);
            return B2_EINVAL;
        }
        if (s.n_pairs <= 0) {
            set_error("gemm: segment %d is empty", i);
            return B2_EINVAL;
        }
        s.tile0 = tiles;
        tiles += (s.n_pairs + GEMM_TILE_PAIRS - 1) / GEMM_TILE_PAIRS;
    }
    a.n_tiles = tiles;
    return 0;
}

int gemm_act_for_type(int type) {
    if (type == B2_TYPE_Q4_K || type == B2_TYPE_Q5_K || type == B2_TYPE_Q6_K) return ACT_Q8K;
    if (type == B2_TYPE_Q8_0) return ACT_Q80;
    return -1;
}
bool gemm_supports_type(int type) { return gemm_act_for_type(type) >= 0; }

int launch_act_quant(const float* x, int64_t x_stride, const float* norm_w, float eps, int K, int cols, uint8_t* act, cudaStream_t st,
                     int act_fmt) {
    if (K <= 0 || K % (act_fmt == ACT_Q80 ? 32 : 256) || cols < 1 || cols > GEMM_MAX_COLS) {
        set_error("act_quant: K=%d cols=%d", K, cols);
        return B2_EINVAL;
    }
    if (((uintptr_t)x & 15) || (x_stride & 3) || ((uintptr_t)norm_w & 15) || ((uintptr_t)act & 15)) {
        set_error("act_quant: pointers must be 16-byte aligned");
        return B2_EINVAL;
    }
    const int nb = gemm_nb(cols);
    if (act_fmt == ACT_Q80) act_quant80_kernel<<<nb, 256, 0, st>>>(x, x_stride, norm_w, eps, K, cols, nb, act);
    else act_quant_kernel<<<nb, 256, 0, st>>>(x, x_stride, norm_w, eps, K, cols, nb, act);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int launch_gemm(const GemmArgs& a, cudaStream_t st) {
    const int nb = gemm_nb(a.cols.n);
    if (a.act_fmt == ACT_Q80) {
        if (nb == 8) return launch_gemm_ng<1, ACT_Q80>(a, st);
        if (nb == 16) return launch_gemm_ng<2, ACT_Q80>(a, st);
        return launch_gemm_ng<4, ACT_Q80>(a, st);
    }
    if (nb == 8) return launch_gemm_ng<1, ACT_Q8K>(a, st);
    if (nb == 16) return launch_gemm_ng<2, ACT_Q8K>(a, st);
    return launch_gemm_ng<4, ACT_Q8K>(a, st);
}

}  // namespace b2
