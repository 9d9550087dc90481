// gemv.cu -- dequant-fused mat-vec for decode (8a row a12 "quantised matmul", with the
// neighbouring rows a11 RMSNorm, a13 RoPE + KV write, a15 SwiGLU / residual / bias fused in).
//
// Arithmetic follows ggml's CPU mul_mat for one token: the fp32 activation row is quantised to Q8_K
// (per 256: d = 1/iscale, iscale = -127/max, q = nearest_int(iscale*x), 16-element bsums) for K-quant
// weights or to Q8_0 (per 32: d = amax/127 rounded to fp16, q = roundf(x/d)) for Q8_0 weights, and the
// dot products are integer (dp4a) with one fp32 scale per (super-)block.  See oracle/ggml_ref.c for
// the CPU restatement these kernels are tested against.
//
// Data movement (HBM-bound byte work; no tensor cores on purpose):
//   * weights stay in HBM exactly as in the GGUF.  A producer warp streams row segments into a
//     3-stage shared-memory ring with 1-D bulk async copies (cp.async.bulk, SASS UBLKCP) that
//     complete on mbarriers; 8 consumer warps read the ring with aligned 16-byte LDS.  Because the
//     copy engine moves raw bytes, the 2-byte-aligned 210-byte Q6_K and 34-byte Q8_0 blocks need no
//     repacking: consumers read them with aligned words + funnel shifts.
//   * the producer never touches activations, so under programmatic dependent launch it starts
//     filling the ring while the previous kernel in the stream is still running; consumers block in
//     griddepcontrol.wait until their input exists.
//   * each CTA rebuilds the (tiny) quantised activation vector in shared memory from the fp32 vector
//     in L2, which removes separate norm / quantise launches from the dependency chain.
//   * a tile is 4 row pairs x one K-chunk of 4096 weights; each pair is worked on by two consumer
//     warps that split the chunk's blocks between them (8 consumer warps per CTA, 2 CTAs per SM);
//     tiles are dealt round-robin over the grid.
#include "gemv.cuh"

#include "common.h"
#include "device_utils.cuh"

namespace b2 {

namespace {

constexpr int NCW = GEMV_CONSUMER_WARPS;
constexpr int NCT = NCW * 32;  // consumer threads

template <int ACT>
struct Geo {
    // blocks per K-chunk, bytes of a ring row slot (largest row segment + 16 bytes of alignment slack)
    static constexpr int CHUNK_BLOCKS = ACT == ACT_Q8K ? 16 : 64;
    static constexpr int ROW_SLOT = ACT == ACT_Q8K ? (16 * 210 + 16) : (64 * 34 + 16);
    static constexpr int STAGE_BYTES = 2 * GEMV_TILE_PAIRS * ROW_SLOT;
};

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

// ------------------------------------------------------------------------------------------------
// per-type inner products of one K-chunk of a row pair held in the ring
// ------------------------------------------------------------------------------------------------
struct ActView {
    const int8_t* q;     // [K]
    const float* d;      // [K/256] (Q8_K) or [K/32] (Q8_0, fp16-rounded)
    const int16_t* bs;   // [K/16] (Q8_K only)
};

__device__ __forceinline__ float q4k_block(const uint8_t* blk, int j, int c, const uint4& xl, const uint4& xh, int bs) {
    const uint4 hdr = *(const uint4*)blk;
    const uint4 q = *(const uint4*)(blk + 16 + 16 * j);
    int dlo = dp4a_ss(q.x & 0x0F0F0F0Fu, xl.x, 0);
    int dhi = dp4a_ss((q.x >> 4) & 0x0F0F0F0Fu, xh.x, 0);
    dlo = dp4a_ss(q.y & 0x0F0F0F0Fu, xl.y, dlo);
    dhi = dp4a_ss((q.y >> 4) & 0x0F0F0F0Fu, xh.y, dhi);
    dlo = dp4a_ss(q.z & 0x0F0F0F0Fu, xl.z, dlo);
    dhi = dp4a_ss((q.z >> 4) & 0x0F0F0F0Fu, xh.z, dhi);
    dlo = dp4a_ss(q.w & 0x0F0F0F0Fu, xl.w, dlo);
    dhi = dp4a_ss((q.w >> 4) & 0x0F0F0F0Fu, xh.w, dhi);
    // 6-bit scales/mins, four at a time (get_scale_min_k4 on packed words)
    const uint32_t sc_lo = hdr.y & 0x3f3f3f3fu, m_lo = hdr.z & 0x3f3f3f3fu;
    const uint32_t sc_hi = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
    const uint32_t m_hi = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
    const uint32_t scw = (c < 2 ? sc_lo : sc_hi) >> (16 * (c & 1));
    const int sc0 = scw & 0xff, sc1 = (scw >> 8) & 0xff;
    const int mj = ((j < 4 ? m_lo : m_hi) >> (8 * (j & 3))) & 0xff;
    const float d = h2f_bits(hdr.x & 0xffffu), dmin = h2f_bits(hdr.x >> 16);
    return d * (float)(sc0 * dlo + sc1 * dhi) - dmin * (float)(mj * bs);
}

// Q4_K: 8 lanes per 144-byte block (lane j takes qs[16j..16j+15]: low nibbles -> sub-block 2c,
// high nibbles -> sub-block 2c+1, c = j/2), 4 blocks per warp pass.
__device__ __forceinline__ void dot_chunk_q4k(const uint8_t* ra, const uint8_t* rb, int b0, int b1, int kb0, const ActView& x,
                                              int lane, float& acc0, float& acc1) {
    const int j = lane & 7, c = j >> 1, hb = j & 1;
    for (int b = b0 + (lane >> 3); b < b1; b += 4) {
        const int kb = kb0 + b;
        const int8_t* xb = x.q + kb * 256 + 64 * c + 16 * hb;
        const uint4 xl = *(const uint4*)xb, xh = *(const uint4*)(xb + 32);
        const uint32_t bsw = *(const uint32_t*)(x.bs + kb * 16 + 2 * j);
        const int bs = (int)(int16_t)(bsw & 0xffff) + (int)(int16_t)(bsw >> 16);
        const float dx = x.d[kb];
        acc0 += dx * q4k_block(ra + b * 144, j, c, xl, xh, bs);
        acc1 += dx * q4k_block(rb + b * 144, j, c, xl, xh, bs);
    }
}

__device__ __forceinline__ float q5k_block(const uint8_t* blk, int j, int c, int hb, const uint4& xl, const uint4& xh, int bs) {
    const uint4 hdr = *(const uint4*)blk;
    const uint4 qh = *(const uint4*)(blk + 16 + 16 * hb);
    const uint4 q = *(const uint4*)(blk + 48 + 16 * j);
    const int s0 = 2 * c, s1 = 2 * c + 1;
    int dlo = 0, dhi = 0;
#define B2_Q5_STEP(W)                                                                               \
    dlo = dp4a_ss((q.W & 0x0F0F0F0Fu) | (((qh.W >> s0) & 0x01010101u) << 4), xl.W, dlo);            \
    dhi = dp4a_ss(((q.W >> 4) & 0x0F0F0F0Fu) | (((qh.W >> s1) & 0x01010101u) << 4), xh.W, dhi);
    B2_Q5_STEP(x) B2_Q5_STEP(y) B2_Q5_STEP(z) B2_Q5_STEP(w)
#undef B2_Q5_STEP
    const uint32_t sc_lo = hdr.y & 0x3f3f3f3fu, m_lo = hdr.z & 0x3f3f3f3fu;
    const uint32_t sc_hi = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
    const uint32_t m_hi = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
    const uint32_t scw = (c < 2 ? sc_lo : sc_hi) >> (16 * (c & 1));
    const int sc0 = scw & 0xff, sc1 = (scw >> 8) & 0xff;
    const int mj = ((j < 4 ? m_lo : m_hi) >> (8 * (j & 3))) & 0xff;
    const float d = h2f_bits(hdr.x & 0xffffu), dmin = h2f_bits(hdr.x >> 16);
    return d * (float)(sc0 * dlo + sc1 * dhi) - dmin * (float)(mj * bs);
}

__device__ __forceinline__ void dot_chunk_q5k(const uint8_t* ra, const uint8_t* rb, int b0, int b1, int kb0, const ActView& x,
                                              int lane, float& acc0, float& acc1) {
    const int j = lane & 7, c = j >> 1, hb = j & 1;
    for (int b = b0 + (lane >> 3); b < b1; b += 4) {
        const int kb = kb0 + b;
        const int8_t* xb = x.q + kb * 256 + 64 * c + 16 * hb;
        const uint4 xl = *(const uint4*)xb, xh = *(const uint4*)(xb + 32);
        const uint32_t bsw = *(const uint32_t*)(x.bs + kb * 16 + 2 * j);
        const int bs = (int)(int16_t)(bsw & 0xffff) + (int)(int16_t)(bsw >> 16);
        const float dx = x.d[kb];
        acc0 += dx * q5k_block(ra + b * 176, j, c, hb, xl, xh, bs);
        acc1 += dx * q5k_block(rb + b * 176, j, c, hb, xl, xh, bs);
    }
}

// Q6_K: one "unit" per lane = half h, 16-column group p of a block: 16 bytes of ql (A), 16 of ql+32 (B),
// 16 of qh give 64 weights in four scale groups 8h+2j+p, j = 0..3 (elements 128h + 32j + 16p + 0..15).
// The -32 offset is taken out through the activation's 16-element sums instead of per byte.
__device__ __forceinline__ float q6k_unit(const uint8_t* blk, int h, int p, const uint4* xv, const int* bsv) {
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
    return d * (float)isum;
}

__device__ __forceinline__ void dot_chunk_q6k(const uint8_t* ra, const uint8_t* rb, int b0, int b1, int kb0, const ActView& x,
                                              int lane, float& acc0, float& acc1) {
    const int h = (lane >> 1) & 1, p = lane & 1;
    for (int b = b0 + (lane >> 2); b < b1; b += 8) {
        const int kb = kb0 + b;
        const int8_t* xb = x.q + kb * 256 + 128 * h + 16 * p;
        uint4 xv[4];
        int bsv[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            xv[j] = *(const uint4*)(xb + 32 * j);
            bsv[j] = x.bs[kb * 16 + 8 * h + 2 * j + p];
        }
        const float dx = x.d[kb];
        acc0 += dx * q6k_unit(ra + b * 210, h, p, xv, bsv);
        acc1 += dx * q6k_unit(rb + b * 210, h, p, xv, bsv);
    }
}

// Q8_0: one 34-byte block per lane.
__device__ __forceinline__ float q80_block(const uint8_t* blk, const uint4& x0, const uint4& x1) {
    const uint4 a = lds16_a2(blk + 2), b = lds16_a2(blk + 18);
    int s = dp4a_ss(a.x, x0.x, 0);
    s = dp4a_ss(a.y, x0.y, s);
    s = dp4a_ss(a.z, x0.z, s);
    s = dp4a_ss(a.w, x0.w, s);
    s = dp4a_ss(b.x, x1.x, s);
    s = dp4a_ss(b.y, x1.y, s);
    s = dp4a_ss(b.z, x1.z, s);
    s = dp4a_ss(b.w, x1.w, s);
    return (float)s * h2f_bits(*(const uint16_t*)blk);
}
__device__ __forceinline__ void dot_chunk_q80(const uint8_t* ra, const uint8_t* rb, int b0, int b1, int kb0, const ActView& x,
                                              int lane, float& acc0, float& acc1) {
    for (int b = b0 + lane; b < b1; b += 32) {
        const int kb = kb0 + b;
        const uint4 x0 = *(const uint4*)(x.q + kb * 32), x1 = *(const uint4*)(x.q + kb * 32 + 16);
        const float dx = x.d[kb];
        // ggml: sumi * (d_w * d_x); keep that association
        acc0 += q80_block(ra + b * 34, x0, x1) * dx;
        acc1 += q80_block(rb + b * 34, x0, x1) * dx;
    }
}

// ------------------------------------------------------------------------------------------------
// prologue: fp32 activation (optionally RMS-normalised) -> quantised copy in shared memory.
// Latency-bound (one small vector out of L2), so every load a warp needs is issued before the first
// use: warp w owns the 256-element spans w, w+8, ...; a lane holds 8 consecutive values of each.
// ------------------------------------------------------------------------------------------------
template <int ACT>
__device__ __forceinline__ void quantise_span(const float (&v)[8], int base, int lane, int K, int8_t* xq, float* xd, int16_t* xbs) {
    const int i0 = base + lane * 8;
    int q[8];
    if (ACT == ACT_Q8K) {
        // first element with the largest magnitude (strict '>' scan in quantize_row_q8_K_ref)
        float amax = 0.0f, mval = 0.0f;
        int midx = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float ax = fabsf(v[i]);
            if (ax > amax) { amax = ax; mval = v[i]; midx = lane * 8 + i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float oa = __shfl_xor_sync(0xffffffffu, amax, o);
            const float ov = __shfl_xor_sync(0xffffffffu, mval, o);
            const int oi = __shfl_xor_sync(0xffffffffu, midx, o);
            if (oa > amax || (oa == amax && oi < midx)) { amax = oa; mval = ov; midx = oi; }
        }
        float d = 0.0f;
        if (amax != 0.0f) {
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
        s8 += __shfl_xor_sync(0xffffffffu, s8, 1);
        if (i0 < K) {
            if ((lane & 1) == 0) xbs[(base >> 4) + (lane >> 1)] = (int16_t)s8;
            if (lane == 0) xd[base >> 8] = d;
        }
    } else {
        // Q8_0: 32-element blocks = 4 lanes
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        const float d = amax / 127.0f;
        const float id = d != 0.0f ? 1.0f / d : 0.0f;
#pragma unroll
        for (int i = 0; i < 8; i++) q[i] = (int)roundf(v[i] * id);
        if (i0 < K && (lane & 3) == 0) xd[i0 >> 5] = __half2float(__float2half_rn(d));
    }
    if (i0 < K) {
        uint2 pk;
        pk.x = (uint32_t)(q[0] & 0xff) | ((uint32_t)(q[1] & 0xff) << 8) | ((uint32_t)(q[2] & 0xff) << 16) | ((uint32_t)(q[3] & 0xff) << 24);
        pk.y = (uint32_t)(q[4] & 0xff) | ((uint32_t)(q[5] & 0xff) << 8) | ((uint32_t)(q[6] & 0xff) << 16) | ((uint32_t)(q[7] & 0xff) << 24);
        *(uint2*)(xq + i0) = pk;
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

template <int ACT>
__device__ void build_activation(const GemvArgs& a, int8_t* xq, float* xd, int16_t* xbs, double* red, int tid) {
    const int K = a.K;
    const int lane = tid & 31, warp = tid >> 5;
    const int nspans = (K + 255) >> 8;
    if (a.prologue == PRO_NORM) {
        // ggml rms_norm: sum of fp32 squares accumulated in double, mean, 1/sqrtf(mean + eps); then (x*scale)*w
        constexpr int MB = 4;   // spans per warp kept in registers (d_model <= 8192); larger vectors are re-read
        const bool resident = nspans <= NCW * MB;
        float v[MB][8], w[MB][8];
        double ss = 0.0;
        for (int s0 = 0; s0 < nspans; s0 += NCW * MB) {
#pragma unroll
            for (int m = 0; m < MB; m++) {
                const int base = (s0 + warp + m * NCW) << 8;
                load8(a.x, base < K ? base + lane * 8 : K, K, v[m]);
                if (resident) load8(a.norm_w, base < K ? base + lane * 8 : K, K, w[m]);
            }
#pragma unroll
            for (int m = 0; m < MB; m++)
#pragma unroll
                for (int i = 0; i < 8; i++) ss += (double)(v[m][i] * v[m][i]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if (lane == 0) red[warp] = ss;
        named_bar_sync(1, NCT);
        double tot = 0.0;
#pragma unroll
        for (int k = 0; k < NCW; k++) tot += red[k];
        const float mean = (float)(tot / (double)K);
        const float scale = 1.0f / sqrtf(mean + a.eps);
        for (int s0 = 0; s0 < nspans; s0 += NCW * MB) {
#pragma unroll
            for (int m = 0; m < MB; m++) {
                const int base = (s0 + warp + m * NCW) << 8;
                if (!resident) {
                    load8(a.x, base < K ? base + lane * 8 : K, K, v[m]);
                    load8(a.norm_w, base < K ? base + lane * 8 : K, K, w[m]);
                }
            }
#pragma unroll
            for (int m = 0; m < MB; m++) {
                const int base = (s0 + warp + m * NCW) << 8;
                if (base < K) {   // warp-uniform
#pragma unroll
                    for (int i = 0; i < 8; i++) v[m][i] = (v[m][i] * scale) * w[m][i];
                    quantise_span<ACT>(v[m], base, lane, K, xq, xd, xbs);
                }
            }
        }
    } else {
        constexpr int MB = 8;
        float v[MB][8];
        for (int s0 = 0; s0 < nspans; s0 += NCW * MB) {
#pragma unroll
            for (int m = 0; m < MB; m++) {
                const int base = (s0 + warp + m * NCW) << 8;
                load8(a.x, base < K ? base + lane * 8 : K, K, v[m]);
            }
#pragma unroll
            for (int m = 0; m < MB; m++) {
                const int base = (s0 + warp + m * NCW) << 8;
                if (base < K) quantise_span<ACT>(v[m], base, lane, K, xq, xd, xbs);
            }
        }
    }
    named_bar_sync(1, NCT);
}

// ------------------------------------------------------------------------------------------------
// epilogue for one row pair (lane 0 of the consumer warp)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void epilogue(const GemvArgs& a, const GemvSeg& s, int p, int ra, int rb, float va, float vb) {
    if (s.bias) {
        va += s.bias[ra];
        if (rb >= 0) vb += (s.pairing == PAIR_SPLIT ? s.bias_b : s.bias)[rb];
    }
    switch (s.epi) {
        case EPI_STORE:
            s.out[ra] = va;
            if (rb >= 0) s.out[rb] = vb;
            break;
        case EPI_RESID:
            s.out[ra] += va;
            if (rb >= 0) s.out[rb] += vb;
            break;
        case EPI_SWIGLU:
            // ggml silu: x / (1 + expf(-x))
            s.out[p] = (va / (1.0f + expf(-va))) * vb;
            break;
        case EPI_ROPE_Q:
        case EPI_ROPE_K: {
            const int pos = *a.pos_ptr;
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
            const int pos = *a.pos_ptr;
            const int hd = s.head_dim;
            const int head = ra / hd;
            __half* dst = s.cache + ((size_t)head * s.n_ctx + pos) * hd;
            dst[ra - head * hd] = __float2half_rn(va);
            if (rb >= 0) dst[rb - head * hd] = __float2half_rn(vb);
            break;
        }
    }
}

template <int ACT>
__global__ void __launch_bounds__(GEMV_THREADS, 2) gemv_kernel(const __grid_constant__ GemvArgs a) {
    using G = Geo<ACT>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem;
    uint64_t* full = (uint64_t*)(smem + GEMV_STAGES * G::STAGE_BYTES);
    uint64_t* empty = full + GEMV_STAGES;
    double* red = (double*)(empty + GEMV_STAGES);
    float* part = (float*)(red + NCW);                 // [2 parities][pairs][2 rows] partial sums of the odd warps
    int8_t* xq = (int8_t*)(part + 2 * GEMV_TILE_PAIRS * 2);
    const int K = a.K;
    float* xd = (float*)(xq + K);
    int16_t* xbs = (int16_t*)(xd + (ACT == ACT_Q8K ? K / 256 : K / 32));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < GEMV_STAGES; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NCW);
        }
        mbar_fence_init();
    }
    __syncthreads();
    // let the next kernel in the stream begin its own weight prefetch as early as possible
    pdl_launch_dependents();

    const int nblk_row = K / (ACT == ACT_Q8K ? 256 : 32);
    const int n_chunks = (nblk_row + G::CHUNK_BLOCKS - 1) / G::CHUNK_BLOCKS;

    if (warp == NCW) {
        // ===== producer: weights only, independent of the previous kernel's output =====
        int stage = 0, phase = 0;
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
            int si = 0;
            while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
            const GemvSeg& s = a.seg[si];
            const int bb = block_bytes_of(s.type);
            const uint8_t* rowp = nullptr;
            if (lane < 2 * GEMV_TILE_PAIRS) {
                const int p = (tile - s.tile0) * GEMV_TILE_PAIRS + (lane >> 1);
                if (p < s.n_pairs) {
                    int ra, rb;
                    pair_rows(s, p, &ra, &rb);
                    const int r = (lane & 1) ? rb : ra;
                    const uint8_t* w = (s.pairing == PAIR_SPLIT && (lane & 1)) ? s.wb : s.wa;
                    if (r >= 0) rowp = w + (size_t)r * nblk_row * bb;
                }
            }
            for (int c = 0; c < n_chunks; c++) {
                const int nb = min(G::CHUNK_BLOCKS, nblk_row - c * G::CHUNK_BLOCKS);
                const uint8_t* src = nullptr;
                uint32_t sz = 0;
                if (rowp) {
                    const uintptr_t lo = (uintptr_t)(rowp + (size_t)c * G::CHUNK_BLOCKS * bb);
                    const uintptr_t hi = lo + (size_t)nb * bb;
                    src = (const uint8_t*)(lo & ~(uintptr_t)15);
                    sz = (uint32_t)(((hi + 15) & ~(uintptr_t)15) - (uintptr_t)src);
                }
                uint32_t total = sz;
#pragma unroll
                for (int o = 4; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
                if (lane == 0) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    mbar_arrive_expect_tx(&full[stage], total);
                }
                __syncwarp();
                if (sz) bulk_g2s(ring + stage * G::STAGE_BYTES + lane * G::ROW_SLOT, src, sz, &full[stage]);
                if (++stage == GEMV_STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    // ===== consumers =====
    pdl_wait();  // the activation vector is written by the previous kernel
    build_activation<ACT>(a, xq, xd, xbs, red, tid);
    const ActView xv{xq, xd, xbs};

    // warps 2p and 2p+1 share row pair p of the tile and split each chunk's blocks between them
    const int pw = warp >> 1, half = warp & 1;
    int stage = 0, phase = 0, parity = 0;
    for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, parity ^= 1) {
        int si = 0;
        while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
        const GemvSeg& s = a.seg[si];
        const int bb = block_bytes_of(s.type);
        const int p = (tile - s.tile0) * GEMV_TILE_PAIRS + pw;
        const bool valid = p < s.n_pairs;
        int ra = 0, rb = -1;
        uint32_t offa = 0, offb = 0;
        if (valid) {
            pair_rows(s, p, &ra, &rb);
            // byte offset of the row segment inside its ring slot = source address mod 16
            const uint8_t* wb = s.pairing == PAIR_SPLIT ? s.wb : s.wa;
            offa = (uint32_t)((uintptr_t)(s.wa + (size_t)ra * nblk_row * bb) & 15);
            offb = (uint32_t)((uintptr_t)(wb + (size_t)(rb < 0 ? ra : rb) * nblk_row * bb) & 15);
        }
        const uint32_t chunk_step = (uint32_t)(G::CHUNK_BLOCKS * bb) & 15;  // alignment drift per chunk
        float acc0 = 0.0f, acc1 = 0.0f;
        for (int c = 0; c < n_chunks; c++) {
            const int nb = min(G::CHUNK_BLOCKS, nblk_row - c * G::CHUNK_BLOCKS);
            // split in whole warp passes (4 blocks for Q4_K/Q5_K, 8 for Q6_K, 32 for Q8_0)
            constexpr int GRAN = ACT == ACT_Q8K ? 8 : 32;
            const int h0 = min(nb, ((nb + 2 * GRAN - 1) / (2 * GRAN)) * GRAN);
            const int b0 = half ? h0 : 0, b1 = half ? nb : h0;
            mbar_wait(&full[stage], phase);
            if (valid && b0 < b1) {
                const uint8_t* slot = ring + stage * G::STAGE_BYTES + (2 * pw) * G::ROW_SLOT;
                const uint8_t* pa = slot + ((offa + c * chunk_step) & 15);
                const uint8_t* pb = slot + G::ROW_SLOT + ((offb + c * chunk_step) & 15);
                if (rb < 0) pb = pa;
                const int kb0 = c * G::CHUNK_BLOCKS;
                if (ACT == ACT_Q8K) {
                    if (s.type == B2_TYPE_Q4_K) dot_chunk_q4k(pa, pb, b0, b1, kb0, xv, lane, acc0, acc1);
                    else if (s.type == B2_TYPE_Q6_K) dot_chunk_q6k(pa, pb, b0, b1, kb0, xv, lane, acc0, acc1);
                    else dot_chunk_q5k(pa, pb, b0, b1, kb0, xv, lane, acc0, acc1);
                } else {
                    dot_chunk_q80(pa, pb, b0, b1, kb0, xv, lane, acc0, acc1);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (++stage == GEMV_STAGES) { stage = 0; phase ^= 1; }
        }
        acc0 = warp_sum(acc0);
        acc1 = warp_sum(acc1);
        float* pp = part + (parity * GEMV_TILE_PAIRS + pw) * 2;
        if (half && lane == 0) { pp[0] = acc0; pp[1] = acc1; }
        named_bar_sync(2 + pw, 64);   // the pair's two warps
        if (!half && lane == 0 && valid) epilogue(a, s, p, ra, rb, acc0 + pp[0], acc1 + pp[1]);
    }
}

}  // namespace

int gemv_act_for_type(int type) {
    switch (type) {
        case B2_TYPE_Q4_K: case B2_TYPE_Q5_K: case B2_TYPE_Q6_K: return ACT_Q8K;
        case B2_TYPE_Q8_0: return ACT_Q80;
    }
    return -1;
}

size_t gemv_smem_bytes(int act, int K) {
    const size_t stage = act == ACT_Q8K ? Geo<ACT_Q8K>::STAGE_BYTES : Geo<ACT_Q80>::STAGE_BYTES;
    size_t n = GEMV_STAGES * stage + 2 * GEMV_STAGES * sizeof(uint64_t) + NCW * sizeof(double) + 2 * GEMV_TILE_PAIRS * 2 * sizeof(float);
    n += (size_t)K;                                           // int8 activations
    n += (act == ACT_Q8K ? K / 256 : K / 32) * sizeof(float);  // block scales
    if (act == ACT_Q8K) n += (size_t)K / 16 * sizeof(int16_t);
    return (n + 127) & ~(size_t)127;
}

int gemv_finalize(GemvArgs& a, int act) {
    if (a.n_seg < 1 || a.n_seg > 3) {
        set_error("gemv: n_seg=%d out of range", a.n_seg);
        return B2_EINVAL;
    }
    const int kmult = act == ACT_Q8K ? 256 : 32;
    if (a.K <= 0 || a.K % kmult || a.K % 8) {
        set_error("gemv: K=%d is not a multiple of %d", a.K, kmult);
        return B2_EINVAL;
    }
    if (((uintptr_t)a.x & 15) || (a.prologue == PRO_NORM && ((uintptr_t)a.norm_w & 15))) {
        set_error("gemv: activation / norm weight pointers must be 16-byte aligned");
        return B2_EINVAL;
    }
    int tiles = 0;
    for (int i = 0; i < a.n_seg; i++) {
        GemvSeg& s = a.seg[i];
        if (gemv_act_for_type(s.type) != act) {
            set_error("gemv: segment %d type %d does not use activation format %d", i, s.type, act);
            return B2_EUNSUP;
        }
        const uintptr_t need = (s.type == B2_TYPE_Q4_K || s.type == B2_TYPE_Q5_K) ? 15 : 1;
        if (((uintptr_t)s.wa & need) || (s.pairing == PAIR_SPLIT && ((uintptr_t)s.wb & need))) {
            set_error("gemv: weight pointer of segment %d is misaligned for type %d", i, s.type);
            return B2_EINVAL;
        }
        if (s.pairing == PAIR_NEOX && (s.head_dim & 1)) {
            set_error("gemv: NEOX pairing needs an even head_dim");
            return B2_EINVAL;
        }
        if (s.n_pairs <= 0) {
            set_error("gemv: segment %d is empty", i);
            return B2_EINVAL;
        }
        s.tile0 = tiles;
        tiles += (s.n_pairs + GEMV_TILE_PAIRS - 1) / GEMV_TILE_PAIRS;
    }
    a.n_tiles = tiles;
    return 0;
}

static int g_num_sms = 0;

int launch_gemv(const GemvArgs& a, int act, cudaStream_t st, bool pdl) {
    if (!g_num_sms) {
        int dev = 0;
        B2_CUDA_CHECK(cudaGetDevice(&dev));
        B2_CUDA_CHECK(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const size_t smem = gemv_smem_bytes(act, a.K);
    if (smem > 227 * 1024) {
        set_error("gemv: K=%d needs %zu bytes of shared memory (limit 227 KiB)", a.K, smem);
        return B2_ERANGE;
    }
    auto kern = act == ACT_Q8K ? gemv_kernel<ACT_Q8K> : gemv_kernel<ACT_Q80>;
    static size_t configured[2] = {0, 0};
    if (smem > configured[act]) {
        B2_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[act] = smem;
    }
    int ctas_per_sm = (int)((227 * 1024) / (smem + 1024));
    if (ctas_per_sm > 2) ctas_per_sm = 2;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    int grid = g_num_sms * ctas_per_sm;
    if (grid > a.n_tiles) grid = a.n_tiles;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(GEMV_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    B2_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, a));
    return 0;
}

}  // namespace b2
