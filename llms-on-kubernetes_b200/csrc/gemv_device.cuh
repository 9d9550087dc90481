// gemv_device.cuh -- device code of the dequant-fused mat-vec, shared by the per-launch kernel (gemv.cu) and the
// persistent per-token kernel (step.cu).  See gemv.cu for the design notes.
#pragma once
#include "gemv.cuh"

#include "common.h"
#include "device_utils.cuh"

namespace b2 {

namespace gemvdev {

constexpr int NCW = GEMV_CONSUMER_WARPS;
constexpr int NCT = NCW * 32;   // consumer threads
constexpr int TILE_ROWS = 8;
constexpr int MAX_SPANS = 128;  // K <= 32768

template <int ACT>
struct Geo {
    // blocks per K-chunk, bytes of a ring row slot (largest row segment + 16 bytes of alignment slack)
    static constexpr int CHUNK_BLOCKS = ACT == ACT_Q8K ? 16 : 64;
    static constexpr int ROW_SLOT = ACT == ACT_Q8K ? (16 * 210 + 16) : (64 * 34 + 16);
    static constexpr int STAGE_BYTES = TILE_ROWS * ROW_SLOT;
    // decoded Q4_K/Q5_K headers of one stage: 32 bytes per (row, block)
    // (row stride padded by 16 bytes so the 8 rows a warp reads at once fall into different banks)
    static constexpr int DEC_ROW = CHUNK_BLOCKS * 32 + 16;
    static constexpr int DEC_BYTES = ACT == ACT_Q8K ? TILE_ROWS * DEC_ROW : 0;
};

// The int8 activations are stored with a bank swizzle inside every 256-byte span: 16-byte granule g
// lives at granule g ^ ((g & 8) >> 2).  A warp reads four granules 64 bytes apart (Q4_K/Q5_K quarters)
// or two pairs 128 bytes apart (Q6_K quarters) in one LDS.128; unswizzled, those alias two-way.
__device__ __forceinline__ int xswz(int g) { return g ^ ((g & 8) >> 2); }

// decoded header record: {float d, float dmin, pad, pad, w[4]} with w[c] = bytes (sc[2c], sc[2c+1], m[2c], m[2c+1])
// the decoded header of one (row, block): (d, dmin) as a float2 and the four scale words; kernels choose where they live
struct DecPtr {
    const uint8_t* dd;
    const uint8_t* w;
};
__device__ __forceinline__ void decode_k4_header(const uint8_t* blk, uint8_t* dd_out, uint8_t* w_out) {
    const uint4 hdr = *(const uint4*)blk;
    const uint32_t sc_lo = hdr.y & 0x3f3f3f3fu, m_lo = hdr.z & 0x3f3f3f3fu;
    const uint32_t sc_hi = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
    const uint32_t m_hi = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
    uint4 w;
    w.x = __byte_perm(sc_lo, m_lo, 0x5410);
    w.y = __byte_perm(sc_lo, m_lo, 0x7632);
    w.z = __byte_perm(sc_hi, m_hi, 0x5410);
    w.w = __byte_perm(sc_hi, m_hi, 0x7632);
    *(float2*)dd_out = make_float2(h2f_bits(hdr.x & 0xffffu), h2f_bits(hdr.x >> 16));
    *(uint4*)w_out = w;
}

__host__ __device__ inline int block_bytes_of(int type) {
    switch (type) {
        case B2_TYPE_Q4_K: return 144;
        case B2_TYPE_Q5_K: return 176;
        case B2_TYPE_Q6_K: return 210;
        case B2_TYPE_Q8_0: return 34;
    }
    return 0;
}

// rows of the pair p of a segment; row_b < 0 when the segment has an odd row count and p is last
__device__ __forceinline__ void pair_rows(const GemvSeg& s, int p, int* ra, int* rb) {
    if (s.pairing == PAIR_ADJ) {
        *ra = 2 * p;
        *rb = 2 * p + 1 < s.n_rows ? 2 * p + 1 : -1;
    } else if (s.pairing == PAIR_SPLIT) {
        *ra = p;
        *rb = p;
    } else {
        const int half = s.head_dim >> 1;
        const int h = p / half, i = p - h * half;
        *ra = h * s.head_dim + i;
        *rb = *ra + half;
    }
}

// source pointer of tile slot `slot` (0..7: pair slot/2, row slot&1) or nullptr if it does not exist
__device__ __forceinline__ const uint8_t* slot_row_ptr(const GemvSeg& s, int ptile, int slot, int row_bytes) {
    const int p = ptile * GEMV_TILE_PAIRS + (slot >> 1);
    if (p >= s.n_pairs) return nullptr;
    int ra, rb;
    pair_rows(s, p, &ra, &rb);
    const int r = (slot & 1) ? rb : ra;
    if (r < 0) return nullptr;
    const uint8_t* w = (s.pairing == PAIR_SPLIT && (slot & 1)) ? s.wb : s.wa;
    return w + (size_t)r * row_bytes;
}

// quantised activation vector in shared memory
struct ActView {
    const int8_t* q;      // [K]
    const float* d;       // [K/256] (Q8_K) or [K/32] (Q8_0, fp16-rounded)
    const int16_t* bs16;  // [K/16] sums of 16 (Q8_K)
    const int16_t* bs32;  // [K/32] sums of 32 (Q8_K)
};

// ------------------------------------------------------------------------------------------------
// inner products: one lane = one quarter of a (row, super-block)
// ------------------------------------------------------------------------------------------------
// Q4_K quarter c: qs[32c .. 32c+31]; low nibbles are sub-block 2c, high nibbles sub-block 2c+1.
// rec = the decoded header record of this (row, block).
__device__ __forceinline__ float k4_finish(const DecPtr rec, int c, int dlo, int dhi, uint32_t bs2, float dx) {
    const float2 dd = *(const float2*)rec.dd;
    const uint32_t w = *(const uint32_t*)(rec.w + 4 * c);
    const int isum = (int)(w & 0xffu) * dlo + (int)((w >> 8) & 0xffu) * dhi;
    const int msum = (int)((w >> 16) & 0xffu) * (int)(int16_t)(bs2 & 0xffffu) + (int)(w >> 24) * ((int)bs2 >> 16);
    return dx * (dd.x * (float)isum - dd.y * (float)msum);
}

__device__ __forceinline__ float q4k_quarter(const uint8_t* blk, const DecPtr rec, int c, const uint4 (&xv)[4], uint32_t bs2, float dx) {
    const uint4 q0 = *(const uint4*)(blk + 16 + 32 * c), q1 = *(const uint4*)(blk + 32 + 32 * c);
    // four independent accumulation chains
    int l0 = dp4a_ss(q0.x & 0x0F0F0F0Fu, xv[0].x, 0), h0 = dp4a_us(q0.x & 0xF0F0F0F0u, xv[2].x, 0);
    int l1 = dp4a_ss(q1.x & 0x0F0F0F0Fu, xv[1].x, 0), h1 = dp4a_us(q1.x & 0xF0F0F0F0u, xv[3].x, 0);
    l0 = dp4a_ss(q0.y & 0x0F0F0F0Fu, xv[0].y, l0); h0 = dp4a_us(q0.y & 0xF0F0F0F0u, xv[2].y, h0);
    l1 = dp4a_ss(q1.y & 0x0F0F0F0Fu, xv[1].y, l1); h1 = dp4a_us(q1.y & 0xF0F0F0F0u, xv[3].y, h1);
    l0 = dp4a_ss(q0.z & 0x0F0F0F0Fu, xv[0].z, l0); h0 = dp4a_us(q0.z & 0xF0F0F0F0u, xv[2].z, h0);
    l1 = dp4a_ss(q1.z & 0x0F0F0F0Fu, xv[1].z, l1); h1 = dp4a_us(q1.z & 0xF0F0F0F0u, xv[3].z, h1);
    l0 = dp4a_ss(q0.w & 0x0F0F0F0Fu, xv[0].w, l0); h0 = dp4a_us(q0.w & 0xF0F0F0F0u, xv[2].w, h0);
    l1 = dp4a_ss(q1.w & 0x0F0F0F0Fu, xv[1].w, l1); h1 = dp4a_us(q1.w & 0xF0F0F0F0u, xv[3].w, h1);
    // the high-nibble dot is an exact multiple of 16
    return k4_finish(rec, c, l0 + l1, (h0 + h1) >> 4, bs2, dx);
}

// Q5_K quarter c: as Q4_K plus bit 2c / 2c+1 of qh[0..31] as the fifth bit.
__device__ __forceinline__ float q5k_quarter(const uint8_t* blk, const DecPtr rec, int c, const uint4 (&xv)[4], uint32_t bs2, float dx) {
    const uint4 h0 = *(const uint4*)(blk + 16), h1 = *(const uint4*)(blk + 32);
    const uint4 q0 = *(const uint4*)(blk + 48 + 32 * c), q1 = *(const uint4*)(blk + 64 + 32 * c);
    const int s0 = 2 * c, s1 = 2 * c + 1;
    int dlo = 0, dhi = 0;
#define B2_Q5_STEP(Q, H, XL, XH, W)                                                                  \
    dlo = dp4a_ss((Q.W & 0x0F0F0F0Fu) | (((H.W >> s0) & 0x01010101u) << 4), XL.W, dlo);              \
    dhi = dp4a_ss(((Q.W >> 4) & 0x0F0F0F0Fu) | (((H.W >> s1) & 0x01010101u) << 4), XH.W, dhi);
    B2_Q5_STEP(q0, h0, xv[0], xv[2], x) B2_Q5_STEP(q0, h0, xv[0], xv[2], y)
    B2_Q5_STEP(q0, h0, xv[0], xv[2], z) B2_Q5_STEP(q0, h0, xv[0], xv[2], w)
    B2_Q5_STEP(q1, h1, xv[1], xv[3], x) B2_Q5_STEP(q1, h1, xv[1], xv[3], y)
    B2_Q5_STEP(q1, h1, xv[1], xv[3], z) B2_Q5_STEP(q1, h1, xv[1], xv[3], w)
#undef B2_Q5_STEP
    return k4_finish(rec, c, dlo, dhi, bs2, dx);
}

// Q6_K quarter (h, p): half h, 16-column group p: 16 bytes of ql (A), 16 of ql+32 (B), 16 of qh give 64
// weights in the four scale groups 8h+2j+p, j = 0..3 (elements 128h + 32j + 16p + 0..15).  The -32
// offset is taken out through the activation's 16-element sums instead of per byte.
__device__ __forceinline__ float q6k_quarter(const uint8_t* blk, int h, int p, const uint4 (&xv)[4], const int (&bsv)[4], float dx) {
    const uint4 A = lds16_a2(blk + 64 * h + 16 * p);
    const uint4 B = lds16_a2(blk + 64 * h + 32 + 16 * p);
    const uint4 H = lds16_a2(blk + 128 + 32 * h + 16 * p);
    int d0 = 0, d1 = 0, d2 = 0, d3 = 0;
#define B2_Q6_STEP(W)                                                                      \
    d0 = dp4a_ss((A.W & 0x0F0F0F0Fu) | ((H.W << 4) & 0x30303030u), xv[0].W, d0);           \
    d1 = dp4a_ss((B.W & 0x0F0F0F0Fu) | ((H.W << 2) & 0x30303030u), xv[1].W, d1);           \
    d2 = dp4a_ss(((A.W >> 4) & 0x0F0F0F0Fu) | (H.W & 0x30303030u), xv[2].W, d2);           \
    d3 = dp4a_ss(((B.W >> 4) & 0x0F0F0F0Fu) | ((H.W >> 2) & 0x30303030u), xv[3].W, d3);
    B2_Q6_STEP(x) B2_Q6_STEP(y) B2_Q6_STEP(z) B2_Q6_STEP(w)
#undef B2_Q6_STEP
    const int8_t* sc = (const int8_t*)(blk + 192 + 8 * h + p);
    const int isum = (int)sc[0] * (d0 - 32 * bsv[0]) + (int)sc[2] * (d1 - 32 * bsv[1]) +
                     (int)sc[4] * (d2 - 32 * bsv[2]) + (int)sc[6] * (d3 - 32 * bsv[3]);
    const float d = h2f_bits(*(const uint16_t*)(blk + 208));
    return dx * (d * (float)isum);
}

// Q8_0: one 34-byte block per lane per step
__device__ __forceinline__ float q80_block(const uint8_t* blk, const uint4& x0, const uint4& x1, float dx) {
    const uint4 a = lds16_a2(blk + 2), b = lds16_a2(blk + 18);
    int s = dp4a_ss(a.x, x0.x, 0);
    s = dp4a_ss(a.y, x0.y, s);
    s = dp4a_ss(a.z, x0.z, s);
    s = dp4a_ss(a.w, x0.w, s);
    s = dp4a_ss(b.x, x1.x, s);
    s = dp4a_ss(b.y, x1.y, s);
    s = dp4a_ss(b.z, x1.z, s);
    s = dp4a_ss(b.w, x1.w, s);
    // ggml: sumi * (d_w * d_x); keep that association
    return (float)s * (h2f_bits(*(const uint16_t*)blk) * dx);
}

// one warp's share of a K-chunk for the tile's 8 rows; `row` = this lane's row segment in the ring
// DDS / WS: bytes between consecutive blocks' (d, dmin) pairs / scale words in the row's decoded-header storage
template <int ACT, int DDS = 32, int WS = 32>
__device__ __forceinline__ float dot_chunk(int type, const uint8_t* row, const DecPtr dec_row, int nb, int kb0, const ActView& x,
                                           int warp, int lane) {
    float acc = 0.0f;
    if (ACT == ACT_Q8K) {
        const int c = lane & 3;
        for (int b = warp; b < nb; b += NCW) {   // (unrolling by two was measured: no gain)
            const int kb = kb0 + b;
            const float dx = x.d[kb];
            if (type == B2_TYPE_Q6_K) {
                const int h = c >> 1, p = c & 1;
                const int8_t* xb = x.q + kb * 256;
                uint4 xv[4];
                int bsv[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    xv[j] = *(const uint4*)(xb + 16 * xswz(8 * h + p + 2 * j));
                    bsv[j] = x.bs16[kb * 16 + 8 * h + 2 * j + p];
                }
                acc += q6k_quarter(row + b * 210, h, p, xv, bsv, dx);
            } else {
                const int8_t* xb = x.q + kb * 256;
                uint4 xv[4];
                xv[0] = *(const uint4*)(xb + 16 * xswz(4 * c));       // sub-block 2c, elements 0..15
                xv[1] = *(const uint4*)(xb + 16 * xswz(4 * c + 1));   //               elements 16..31
                xv[2] = *(const uint4*)(xb + 16 * xswz(4 * c + 2));   // sub-block 2c+1
                xv[3] = *(const uint4*)(xb + 16 * xswz(4 * c + 3));
                const uint32_t bs2 = *(const uint32_t*)(x.bs32 + kb * 8 + 2 * c);
                const DecPtr rec{dec_row.dd + b * DDS, dec_row.w + b * WS};
                if (type == B2_TYPE_Q4_K) acc += q4k_quarter(row + b * 144, rec, c, xv, bs2, dx);
                else acc += q5k_quarter(row + b * 176, rec, c, xv, bs2, dx);
            }
        }
    } else {
        // 64-block chunk: warp w owns blocks [8w, 8w+8); lane (row, i) takes blocks 8w+i and 8w+i+4
        const int i = lane & 3;
        for (int b = warp * 8 + i; b < min(nb, warp * 8 + 8); b += 4) {
            const int kb = kb0 + b;
            const uint4 x0 = *(const uint4*)(x.q + kb * 32), x1 = *(const uint4*)(x.q + kb * 32 + 16);
            acc += q80_block(row + b * 34, x0, x1, x.d[kb]);
        }
    }
    return acc;
}

// ------------------------------------------------------------------------------------------------
// prologue: fp32 activation (optionally RMS-normalised) -> quantised copy in the shared memory of
// every CTA of the cluster.  Span s (256 elements) is quantised by warp (s / csize) % 8 of CTA
// s % csize; lane l holds its 8 consecutive values.
// ------------------------------------------------------------------------------------------------
struct ActSmem {
    int8_t* xq;
    float* xd;
    int16_t* bs16;
    int16_t* bs32;
    double* ss_warp;   // [NCW] per-warp sum of squares (PRO_NORM)
    uint64_t* xbar;    // completes when every span has arrived
};

// the same from the sum of the ranks' partial vectors (GemvArgs::x_parts), added in rank order
__device__ __forceinline__ void load8_parts(const float* parts, int nparts, int i0, int K, float (&v)[8]) {
    if (i0 < K) {
        float4 p0 = __ldcg((const float4*)(parts + i0)), p1 = __ldcg((const float4*)(parts + i0 + 4));
        for (int r = 1; r < nparts; r++) {
            const float4 q0 = __ldcg((const float4*)(parts + (size_t)r * K + i0)), q1 = __ldcg((const float4*)(parts + (size_t)r * K + i0 + 4));
            p0.x += q0.x; p0.y += q0.y; p0.z += q0.z; p0.w += q0.w;
            p1.x += q1.x; p1.y += q1.y; p1.z += q1.z; p1.w += q1.w;
        }
        v[0] = p0.x; v[1] = p0.y; v[2] = p0.z; v[3] = p0.w;
        v[4] = p1.x; v[5] = p1.y; v[6] = p1.z; v[7] = p1.w;
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = 0.0f;
    }
}

__device__ __forceinline__ void load8(const float* p, int i0, int K, float (&v)[8]) {
    if (i0 < K) {
        const float4 p0 = *(const float4*)(p + i0), p1 = *(const float4*)(p + i0 + 4);
        v[0] = p0.x; v[1] = p0.y; v[2] = p0.z; v[3] = p0.w;
        v[4] = p1.x; v[5] = p1.y; v[6] = p1.z; v[7] = p1.w;
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = 0.0f;
    }
}

// quantise one span held in v[] and store it into the shared memory of all `csize` CTAs
template <int ACT>
__device__ __forceinline__ void quantise_span(const float (&v)[8], int base, int lane, int K, const ActSmem& sm,
                                              const uint8_t* smem_base, uint32_t csize) {
    const int i0 = base + lane * 8;
    int q[8];
    float d = 0.0f;
    int s8 = 0, s16 = 0, s32 = 0;
    if (ACT == ACT_Q8K) {
        // the first element with the largest magnitude decides the sign of the scale
        // (strict '>' scan in quantize_row_q8_K_ref)
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
        if (wmax != 0.0f) {
            const float iscale = -127.0f / mval;
#pragma unroll
            for (int i = 0; i < 8; i++) q[i] = min(127, __float2int_rn(iscale * v[i]));
            d = 1.0f / iscale;
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) q[i] = 0;
        }
#pragma unroll
        for (int i = 0; i < 8; i++) s8 += q[i];
        s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);
        s32 = s16 + __shfl_xor_sync(0xffffffffu, s16, 2);
    } else {
        // Q8_0: 32-element blocks = 4 lanes
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        d = amax / 127.0f;
        const float id = d != 0.0f ? 1.0f / d : 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) q[i] = (int)roundf(v[i] * id);
        d = __half2float(__float2half_rn(d));
    }
    const uint32_t pk0 = (uint32_t)(q[0] & 0xff) | ((uint32_t)(q[1] & 0xff) << 8) | ((uint32_t)(q[2] & 0xff) << 16) | ((uint32_t)(q[3] & 0xff) << 24);
    const uint32_t pk1 = (uint32_t)(q[4] & 0xff) | ((uint32_t)(q[5] & 0xff) << 8) | ((uint32_t)(q[6] & 0xff) << 16) | ((uint32_t)(q[7] & 0xff) << 24);
    // the 16-bit sums travel in pairs (asynchronous stores are 4 bytes at least): sums of 16 of lanes l, l+2; sums of 32 of lanes l, l+4
    const uint32_t p16 = ((uint32_t)s16 & 0xffffu) | ((uint32_t)__shfl_down_sync(0xffffffffu, s16, 2) << 16);
    const uint32_t p32 = ((uint32_t)s32 & 0xffffu) | ((uint32_t)__shfl_down_sync(0xffffffffu, s32, 4) << 16);
    const uint32_t p32n = __shfl_up_sync(0xffffffffu, p32, 1);
    if (i0 < K) {
        const int qpos = ACT == ACT_Q8K ? base + 16 * xswz(lane >> 1) + 8 * (lane & 1) : i0;
        const uint32_t o_q = (uint32_t)((const uint8_t*)(sm.xq + qpos) - smem_base);
        const uint32_t o_d = (uint32_t)((const uint8_t*)(sm.xd + (ACT == ACT_Q8K ? (base >> 8) : (i0 >> 5))) - smem_base);
        const uint32_t o_16 = (uint32_t)((const uint8_t*)(sm.bs16 + (i0 >> 4)) - smem_base);
        const uint32_t o_32n = (uint32_t)((const uint8_t*)(sm.bs32 + ((i0 - 8) >> 5)) - smem_base);   // of the lane below (lanes 1, 9, ..)
        const uint32_t o_bar = (uint32_t)((const uint8_t*)sm.xbar - smem_base);
        // every byte is reported to the receiving CTA's barrier (activation_bytes() of them complete its phase);
        // straight-line, predicated stores: this runs on the critical path of every stage hand-off
        // (an asynchronous store costs the warp ~90 cycles of issue whatever its lanes carry: the three small items
        // of a Q8_K span share one instruction -- lanes 0, 4, .. the 16-sums, lanes 1, 9, .. the 32-sums of the
        // lane below, lane 2 the scale)
        const bool q4 = (lane & 3) == 0;
        uint32_t o_side = o_16, v_side = p16;
        bool side = q4;
        if (ACT == ACT_Q8K) {
            if ((lane & 7) == 1) { o_side = o_32n; v_side = p32n; side = true; }
            if (lane == 2) { o_side = o_d; v_side = __float_as_uint(d); side = true; }
        } else {
            o_side = o_d; v_side = __float_as_uint(d);
        }
#pragma unroll 4
        for (uint32_t r = 0; r < csize; r++) {
            const uint32_t pb = dsmem_addr(smem_base, r);
            const uint32_t bar = pb + o_bar;
            dsmem_st_async_v2(pb + o_q, pk0, pk1, bar);
            dsmem_st_async_b32_if(side, pb + o_side, v_side, bar);
        }
    }
}

// bytes one activation vector occupies in a CTA's shared memory = what its barrier waits for per phase
template <int ACT>
__device__ __forceinline__ uint32_t activation_bytes(int K) {
    return ACT == ACT_Q8K ? (uint32_t)(K + (K >> 8) * 4 + (K >> 4) * 2 + (K >> 5) * 2) : (uint32_t)(K + (K >> 5) * 4);
}

// parity: phase of xbar to wait for (0 for a kernel that builds one activation vector).
// Called by all NCT compute threads (uses named barrier 1).
// TP = false compiles the tensor-parallel input (x_parts) out: the single-GPU step kernel keeps its registers.
// DEFER (Q8_K activations only): the RMS factor r is a scalar, so it commutes with the quantiser -- q(x r w) = q(x w) and
// d(x r w) = r d(x w) up to fp32 rounding -- and with the whole mat-vec: y = r * sum_b d_w d_x' (q_w . q_x).  The spans are
// then quantised from x * w as soon as they arrive, the sum of squares is added up by the warps that own no span (or
// afterwards) while the exchange is in flight, and r goes to *scale_out for the epilogue warp to apply to the row sums.
template <int ACT, bool TP = true, bool DEFER = false>
__device__ __forceinline__ void build_activation(const GemvArgs& a, const ActSmem& sm, const uint8_t* smem_base, int tid, uint32_t rank, uint32_t csize,
                                 uint32_t parity = 0, unsigned long long* tr = nullptr, float* scale_out = nullptr) {
    const int K = a.K;
    const int lane = tid & 31, warp = tid >> 5;
    if (DEFER && ACT == ACT_Q8K && !(TP && a.x_nparts > 0) && a.prologue == PRO_NORM) {
        const int nspans = (K + 255) >> 8;
        const int first = (int)rank + (int)csize * warp, stride = (int)csize * NCW;
        // warps with a span quantise; when at least half of the warps have none, those add up the squares meanwhile
        const int owners = min(NCW, (nspans - (int)rank + (int)csize - 1) / (int)csize);   // warps 0 .. owners-1 own a span
        const bool split = owners <= NCW / 2;
        double ss = 0.0;
        auto add_squares = [&](int t0, int nt) {
#pragma unroll 4
            for (int i0 = t0 * 8; i0 < K; i0 += nt * 8) {
                float t[8];
                load8(a.x, i0, K, t);
#pragma unroll
                for (int i = 0; i < 8; i++) ss += (double)(t[i] * t[i]);
            }
        };
        if (split && warp >= owners) {
            add_squares(tid - owners * 32, (NCW - owners) * 32);
        } else {
            for (int s0 = first; s0 < nspans; s0 += stride) {
                float t[8], w[8];
                load8(a.x, (s0 << 8) + lane * 8, K, t);
                load8(a.norm_w, (s0 << 8) + lane * 8, K, w);
#pragma unroll
                for (int i = 0; i < 8; i++) t[i] *= w[i];
                quantise_span<ACT>(t, s0 << 8, lane, K, sm, smem_base, csize);
            }
            if (!split) add_squares(tid, NCT);
        }
        // the exchange's byte count does not wait for the norm: announce it, then finish the sum while the stores land
        if (tid == 0) mbar_arrive_expect_tx(sm.xbar, activation_bytes<ACT>(K));
        if (tr && tid == 0) tr[4] = trace_ns();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if (lane == 0) sm.ss_warp[warp] = ss;
        named_bar_sync(1, NCT);
        if (tid == 0) {
            double tot = 0.0;
#pragma unroll
            for (int w = 0; w < NCW; w++) tot += sm.ss_warp[w];
            const float mean = (float)(tot / (double)K);
            *scale_out = 1.0f / sqrtf(mean + a.eps);
            if (tr) tr[5] = trace_ns();
        }
        mbar_wait(sm.xbar, parity);
        if (tr && tid == 0) tr[6] = trace_ns();
        return;
    }
    // (a tensor-parallel shard reads the sum of the ranks' partial vectors instead of x)
#define LOADX(i0, KK, vv) do { if (TP && a.x_nparts > 0) load8_parts(a.x_parts, a.x_nparts, i0, KK, vv); else load8(a.x, i0, KK, vv); } while (0)
    const int nspans = (K + 255) >> 8;
    const int first = (int)rank + (int)csize * warp, stride = (int)csize * NCW;   // this warp's spans
    float scale = 1.0f;
    // the first two spans of a warp stay in registers across the norm reduction (with 8 CTAs sharing the
    // work a warp rarely owns more); further spans are re-read
    float v0[8], v1[8];
    if (a.prologue == PRO_NORM) {
        // ggml rms_norm: sum of fp32 squares accumulated in double, mean, 1/sqrtf(mean + eps); then (x*scale)*w.
        // Every CTA adds up the whole vector itself (K floats out of L2, in flight together with its own
        // spans): a round trip through the cluster for the partial sums costs more than the redundant reads.
        if (first < nspans) LOADX((first << 8) + lane * 8, K, v0);
        if (first + stride < nspans) LOADX(((first + stride) << 8) + lane * 8, K, v1);
        double ss = 0.0;
#pragma unroll 2
        for (int i0 = tid * 8; i0 < K; i0 += NCT * 8) {
            float t[8];
            LOADX(i0, K, t);
#pragma unroll
            for (int i = 0; i < 8; i++) ss += (double)(t[i] * t[i]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if (lane == 0) sm.ss_warp[warp] = ss;
        named_bar_sync(1, NCT);
        double tot = 0.0;
#pragma unroll
        for (int w = 0; w < NCW; w++) tot += sm.ss_warp[w];
        const float mean = (float)(tot / (double)K);
        scale = 1.0f / sqrtf(mean + a.eps);
        if (tr && tid == 0) {
            tr[4] = trace_after(__float_as_uint(v0[0]));
            tr[5] = trace_after(__float_as_uint(scale));
        }
    }
    if (a.prologue == PRO_NORM) {
        auto norm_quant = [&](float (&t)[8], int s) {
            float w[8];
            load8(a.norm_w, (s << 8) + lane * 8, K, w);
#pragma unroll
            for (int i = 0; i < 8; i++) t[i] = (t[i] * scale) * w[i];
            quantise_span<ACT>(t, s << 8, lane, K, sm, smem_base, csize);
        };
        if (first < nspans) norm_quant(v0, first);
        if (first + stride < nspans) norm_quant(v1, first + stride);
        for (int s = first + 2 * stride; s < nspans; s += stride) {
            float t[8];
            LOADX((s << 8) + lane * 8, K, t);
            norm_quant(t, s);
        }
    } else {
        // issue the loads of two spans before the first use
        if (first < nspans) LOADX((first << 8) + lane * 8, K, v0);
        if (first + stride < nspans) LOADX(((first + stride) << 8) + lane * 8, K, v1);
        if (tr && tid == 0) tr[4] = trace_after(__float_as_uint(v0[0]));
        if (first < nspans) quantise_span<ACT>(v0, first << 8, lane, K, sm, smem_base, csize);
        if (first + stride < nspans) quantise_span<ACT>(v1, (first + stride) << 8, lane, K, sm, smem_base, csize);
        for (int s = first + 2 * stride; s < nspans; s += stride) {
            float t[8];
            LOADX((s << 8) + lane * 8, K, t);
            quantise_span<ACT>(t, s << 8, lane, K, sm, smem_base, csize);
        }
    }
    if (tr && tid == 0) tr[6] = trace_ns();
    // the barrier's phase completes when this CTA's copy is whole: one local arrival that announces the byte count
    // (stores of faster peers may already have been counted: the transaction count is signed)
    if (tid == 0) mbar_arrive_expect_tx(sm.xbar, activation_bytes<ACT>(K));
    if (tr && tid == 0) tr[7] = trace_ns();
    mbar_wait(sm.xbar, parity);
#undef LOADX
}

// ------------------------------------------------------------------------------------------------
// epilogue for one row pair
// ------------------------------------------------------------------------------------------------
// The operands that do not depend on the dot products (bias, residual, rotation) are fetched before the
// epilogue warp waits for the partial sums, so their global-memory latency hides behind the tile's math.
struct EpiOperands {
    float bias_a, bias_b;   // 0 without bias
    float res_a, res_b;     // EPI_RESID: current value of out[ra], out[rb]
    float2 cs;              // EPI_ROPE_*: (cos, sin) of this pair
    bool rotate;
};

__device__ __forceinline__ EpiOperands epilogue_operands(const GemvArgs& a, const GemvSeg& s, int ra, int rb, int pos) {
    EpiOperands o;
    o.bias_a = o.bias_b = o.res_a = o.res_b = 0.0f;
    o.cs = make_float2(1.0f, 0.0f);
    o.rotate = false;
    if (s.bias) {
        o.bias_a = s.bias[ra];
        if (rb >= 0) o.bias_b = (s.pairing == PAIR_SPLIT ? s.bias_b : s.bias)[rb];
    }
    if (s.epi == EPI_RESID || s.epi == EPI_RESID_TP) {
        o.res_a = s.out[ra];
        if (rb >= 0) o.res_b = s.out[rb];
    } else if (s.epi == EPI_ROPE_Q || s.epi == EPI_ROPE_K) {
        const int hd = s.head_dim;
        const int head = ra / hd;
        const int i = s.pairing == PAIR_NEOX ? ra - head * hd : (ra - head * hd) >> 1;
        if (i < a.rope_half) {
            o.rotate = true;
            o.cs = a.rope_tab[(size_t)pos * a.rope_half + i];
        }
    }
    return o;
}

template <bool TP = true>
__device__ __forceinline__ void epilogue_apply(const GemvSeg& s, int p, int ra, int rb, float va, float vb, const EpiOperands& o, int pos) {
    if (s.bias) {
        va += o.bias_a;
        if (rb >= 0) vb += o.bias_b;
    }
    switch (s.epi) {
        case EPI_STORE:
            s.out[ra] = va;
            if (rb >= 0) s.out[rb] = vb;
            break;
        case EPI_RESID:
            s.out[ra] = o.res_a + va;
            if (rb >= 0) s.out[rb] = o.res_b + vb;
            break;
        case EPI_RESID_TP: {
            const float fa = o.res_a * s.res_scale + va, fb = o.res_b * s.res_scale + vb;
            if (TP && s.n_peer > 0) {
                // this rank's share of the new residual stream, delivered to every rank (the exchange stage sums them)
                for (int p = 0; p < s.n_peer; p++) {
                    s.peer_out[p][ra] = fa;
                    if (rb >= 0) s.peer_out[p][rb] = fb;
                }
            } else {
                s.out[ra] = fa;
                if (rb >= 0) s.out[rb] = fb;
            }
            break;
        }
        case EPI_SWIGLU:
            // ggml silu: x / (1 + expf(-x))
            s.out[p] = (va / (1.0f + expf(-va))) * vb;
            break;
        case EPI_ROPE_Q:
        case EPI_ROPE_K: {
            const int hd = s.head_dim;
            const int head = ra / hd;
            float oa = va, ob = vb;
            if (o.rotate) {
                oa = va * o.cs.x - vb * o.cs.y;
                ob = va * o.cs.y + vb * o.cs.x;
            }
            if (s.epi == EPI_ROPE_Q) {
                s.out[ra] = oa;
                s.out[rb] = ob;
            } else {
                __half* dst = s.cache + ((size_t)head * s.n_ctx + pos) * hd;
                dst[ra - head * hd] = __float2half_rn(oa);
                dst[rb - head * hd] = __float2half_rn(ob);
            }
            break;
        }
        case EPI_V: {
            const int hd = s.head_dim;
            const int head = ra / hd;
            __half* dst = s.cache + ((size_t)head * s.n_ctx + pos) * hd;
            dst[ra - head * hd] = __float2half_rn(va);
            if (rb >= 0) dst[rb - head * hd] = __float2half_rn(vb);
            break;
        }
    }
}

// epilogue warp, one tile: lanes 0, 2, 4, 6 own the tile's row pairs.  Fetch the operands, wait for the eight
// compute warps' shares, add them in a fixed order, release the buffer, finish the pair.
template <bool TP = true>
__device__ __forceinline__ void epilogue_tile(const GemvArgs& a, const GemvSeg& s, int ptile, int pos, const float* rp,
                                              uint64_t* part_full, uint64_t* part_free, uint32_t full_phase, int lane,
                                              const float* row_scale = nullptr) {
    const int p = ptile * GEMV_TILE_PAIRS + (lane >> 1);
    const bool owner = lane < TILE_ROWS && (lane & 1) == 0 && p < s.n_pairs;
    int ra = 0, rb = -1;
    EpiOperands o{};
    if (owner) {
        pair_rows(s, p, &ra, &rb);
        o = epilogue_operands(a, s, ra, rb, pos);
    }
    mbar_wait(part_full, full_phase);
    float v = 0.0f;
    if (lane < TILE_ROWS) {
#pragma unroll
        for (int w = 0; w < NCW; w++) v += rp[w * TILE_ROWS + lane];
        if (row_scale) v *= *row_scale;   // the RMS factor of a normalising stage whose prologue deferred it (build_activation DEFER)
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(part_free);
    const float vo = __shfl_xor_sync(0xffffffffu, v, 1);
    if (owner) epilogue_apply<TP>(s, p, ra, rb, v, vo, o, pos);
}


}  // namespace gemvdev
}  // namespace b2
