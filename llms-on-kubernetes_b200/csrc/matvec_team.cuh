// matvec_team.cuh -- the mat-vec inner loop of the persistent decode kernel, second design (step2.cu).
//
// Measured on the B200 (tools/ubench.cu): dp4a, lop3 and imad each issue at 2 warp-instructions per clock and SM, a
// 16-byte shared-memory load at one per 8 clocks, and the first design's tile loop (gemv_device.cuh: dot_chunk) spends
// ~0.14 warp-instructions per weight byte -- at ~2 issued instructions per clock and SM that is 4 TB/s, which is what
// its tile phases reach.  The limit is instruction issue, so this design cuts instructions per byte by ~3.5:
//   * work item = 8 weight rows x two 256-element blocks, done by one warp.  For Q4_K the quants go through
//     mma.sync.m16n8k32 (u8 x s8 -> s32) as the A operand -- loaded with ldmatrix, raw bytes, no nibble unpack for the
//     low halves -- against a B operand whose eight columns are 32-element sub-blocks of the int8 activation vector.
//     One MMA of the raw bytes and one of the bytes masked to their high nibbles per 32-byte chunk give every
//     sub-block's integer dot product in the lane that finishes it (layout: q4k_rows8_blocks2_mma).  The MMA unit
//     delivers for one useful column what dp4a delivers per lane, but in 1/16 of the issue slots, and the fma / alu
//     pipes are left to the scale arithmetic;
//   * the 6-bit scales and mins of a lane's two sub-blocks are unpacked by the lane itself from the block header
//     (13 integer instructions) -- no decoder warp, no second barrier per ring slot;
//   * every lane keeps one fp32 row sum over the whole K loop; the four lanes of a row meet once per tile (two
//     shuffles), not once per block.
// The other weight types run the dp4a quarter products of the first design on the same work item (lane (g, t) =
// quarter t of row g, one block after the other).
// The arithmetic per (row, sub-block pair) is the first design's (k4_finish), so the partial products are
// bit-identical; only the order in which they are added differs.
#pragma once
#include "gemv_device.cuh"

namespace b2 {
namespace teamdev {

using namespace gemvdev;

constexpr int TS_NSLOT = 8;                                            // ring allocations in flight (barrier pairs)
constexpr uint32_t TS_RING_BYTES = 5 * TILE_ROWS * (16 * 144 + 16);   // 92 800 B: five Q4_K steps, three Q6_K steps (two CTAs per SM with the K = 14336 activations)
constexpr int MT_CHUNK = 16;      // 256-element blocks of K per ring step: two per compute warp

// bytes of one 256-element block / of one row's share of a ring step (16 blocks + 16 bytes of alignment slack).
// Q4_K: 2320 = 18 x 128 + 16, so the eight rows of an ldmatrix 8x8 start in distinct 16-byte bank groups.
// (Copies of half this size -- 8 blocks per step -- were measured: the SM's copy engine takes ~46 cycles per bulk
// copy whatever its size, so 1152-byte copies cap the stream near 7 TB/s before any arithmetic.)
__host__ __device__ inline int ts_block_bytes(int type) { return type == B2_TYPE_Q8_0 ? 8 * 34 : block_bytes_of(type); }
__host__ __device__ inline uint32_t ts_pitch(int type) { return (uint32_t)(MT_CHUNK * ts_block_bytes(type) + 16); }

// activation layout classes: which 64 elements of a block a lane (b, c) owns
enum { XC_CONTIG = 0,   // Q4_K / Q5_K: elements 64c .. 64c+63 (sub-blocks 2c, 2c+1)
       XC_Q6 = 1,       // Q6_K: half h = c/2, 16-element groups 8h + p + 2j, p = c%2, j = 0..3
       XC_Q80 = 2 };    // Q8_0 weights / Q8_0 activations: elements 64c .. 64c+63 = two 32-blocks
template <int TYPE>
struct XClass {
    static constexpr int value = TYPE == B2_TYPE_Q6_K ? XC_Q6 : TYPE == B2_TYPE_Q8_0 ? XC_Q80 : XC_CONTIG;
};

// a lane's activations for one super-block: 64 int8 + what the finish needs
struct XR {
    uint4 xv[4];
    int e[4];       // CONTIG: e[0] = the two 32-element sums (int16 pair); Q6: the four 16-element sums
    float d0, d1;   // CONTIG / Q6: block scale in d0; Q80: the scales of the lane's two 32-blocks
};

template <int CLS>
__device__ __forceinline__ void load_x(XR& r, const ActView& x, int kb, int c) {
    if (CLS == XC_CONTIG) {
        const int8_t* xb = x.q + kb * 256;
#pragma unroll
        for (int j = 0; j < 4; j++) r.xv[j] = *(const uint4*)(xb + 16 * xswz(4 * c + j));
        r.e[0] = (int)*(const uint32_t*)(x.bs32 + kb * 8 + 2 * c);
        r.d0 = x.d[kb];
    } else if (CLS == XC_Q6) {
        const int h = c >> 1, p = c & 1;
        const int8_t* xb = x.q + kb * 256;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            r.xv[j] = *(const uint4*)(xb + 16 * xswz(8 * h + p + 2 * j));
            r.e[j] = x.bs16[kb * 16 + 8 * h + 2 * j + p];
        }
        r.d0 = x.d[kb];
    } else {
        const int8_t* xb = x.q + kb * 256 + 64 * c;
#pragma unroll
        for (int j = 0; j < 4; j++) r.xv[j] = *(const uint4*)(xb + 16 * j);
        r.d0 = x.d[kb * 8 + 2 * c];
        r.d1 = x.d[kb * 8 + 2 * c + 1];
    }
}

// the 6-bit (scale, min) pairs of sub-blocks 2c, 2c+1 from a raw Q4_K / Q5_K header: two bytes each, low byte = sub-block 2c
__device__ __forceinline__ void k4_scales_of_quarter(const uint4& hdr, int c, uint32_t* sc, uint32_t* mn) {
    const uint32_t sh = (uint32_t)(c & 1) << 4;
    const uint32_t y = hdr.y >> sh, z = hdr.z >> sh, w = hdr.w >> sh;
    const uint32_t sc_lo = y & 0x3f3fu, mn_lo = z & 0x3f3fu;                                  // sub-blocks 0..3
    const uint32_t sc_hi = (w & 0x0f0fu) | ((y >> 2) & 0x3030u);                             // sub-blocks 4..7
    const uint32_t mn_hi = ((w >> 4) & 0x0f0fu) | ((z >> 2) & 0x3030u);
    *sc = c < 2 ? sc_lo : sc_hi;
    *mn = c < 2 ? mn_lo : mn_hi;
}
// same value as gemvdev::k4_finish on the decoded record
__device__ __forceinline__ float k4_finish_raw(const uint4& hdr, int c, int dlo, int dhi, int bs2, float dx) {
    uint32_t sc, mn;
    k4_scales_of_quarter(hdr, c, &sc, &mn);
    const int isum = (int)(sc & 0xffu) * dlo + (int)(sc >> 8) * dhi;
    const int msum = (int)(mn & 0xffu) * (int)(int16_t)((uint32_t)bs2 & 0xffffu) + (int)(mn >> 8) * (bs2 >> 16);
    const float d = h2f_bits(hdr.x & 0xffffu), dmin = h2f_bits(hdr.x >> 16);
    return dx * (d * (float)isum - dmin * (float)msum);
}

// one lane's share of (row, block kb): `row` = the row's first byte in the ring
template <int TYPE>
__device__ __forceinline__ float row_quarter(const uint8_t* row, int kb, int c, const XR& x) {
    if (TYPE == B2_TYPE_Q4_K) {
        const uint8_t* blk = row + kb * 144;
        const uint4 hdr = *(const uint4*)blk;
        const uint4 q0 = *(const uint4*)(blk + 16 + 32 * c), q1 = *(const uint4*)(blk + 32 + 32 * c);
        int l0 = dp4a_ss(q0.x & 0x0F0F0F0Fu, x.xv[0].x, 0), h0 = dp4a_us(q0.x & 0xF0F0F0F0u, x.xv[2].x, 0);
        int l1 = dp4a_ss(q1.x & 0x0F0F0F0Fu, x.xv[1].x, 0), h1 = dp4a_us(q1.x & 0xF0F0F0F0u, x.xv[3].x, 0);
        l0 = dp4a_ss(q0.y & 0x0F0F0F0Fu, x.xv[0].y, l0); h0 = dp4a_us(q0.y & 0xF0F0F0F0u, x.xv[2].y, h0);
        l1 = dp4a_ss(q1.y & 0x0F0F0F0Fu, x.xv[1].y, l1); h1 = dp4a_us(q1.y & 0xF0F0F0F0u, x.xv[3].y, h1);
        l0 = dp4a_ss(q0.z & 0x0F0F0F0Fu, x.xv[0].z, l0); h0 = dp4a_us(q0.z & 0xF0F0F0F0u, x.xv[2].z, h0);
        l1 = dp4a_ss(q1.z & 0x0F0F0F0Fu, x.xv[1].z, l1); h1 = dp4a_us(q1.z & 0xF0F0F0F0u, x.xv[3].z, h1);
        l0 = dp4a_ss(q0.w & 0x0F0F0F0Fu, x.xv[0].w, l0); h0 = dp4a_us(q0.w & 0xF0F0F0F0u, x.xv[2].w, h0);
        l1 = dp4a_ss(q1.w & 0x0F0F0F0Fu, x.xv[1].w, l1); h1 = dp4a_us(q1.w & 0xF0F0F0F0u, x.xv[3].w, h1);
        return k4_finish_raw(hdr, c, l0 + l1, (h0 + h1) >> 4, x.e[0], x.d0);
    } else if (TYPE == B2_TYPE_Q5_K) {
        const uint8_t* blk = row + kb * 176;
        const uint4 hdr = *(const uint4*)blk;
        const uint4 h0 = *(const uint4*)(blk + 16), h1 = *(const uint4*)(blk + 32);
        const uint4 q0 = *(const uint4*)(blk + 48 + 32 * c), q1 = *(const uint4*)(blk + 64 + 32 * c);
        const int s0 = 2 * c, s1 = 2 * c + 1;
        int dlo = 0, dhi = 0;
#define B2_Q5_STEP(Q, H, XL, XH, W)                                                                  \
    dlo = dp4a_ss((Q.W & 0x0F0F0F0Fu) | (((H.W >> s0) & 0x01010101u) << 4), XL.W, dlo);              \
    dhi = dp4a_ss(((Q.W >> 4) & 0x0F0F0F0Fu) | (((H.W >> s1) & 0x01010101u) << 4), XH.W, dhi);
        B2_Q5_STEP(q0, h0, x.xv[0], x.xv[2], x) B2_Q5_STEP(q0, h0, x.xv[0], x.xv[2], y)
        B2_Q5_STEP(q0, h0, x.xv[0], x.xv[2], z) B2_Q5_STEP(q0, h0, x.xv[0], x.xv[2], w)
        B2_Q5_STEP(q1, h1, x.xv[1], x.xv[3], x) B2_Q5_STEP(q1, h1, x.xv[1], x.xv[3], y)
        B2_Q5_STEP(q1, h1, x.xv[1], x.xv[3], z) B2_Q5_STEP(q1, h1, x.xv[1], x.xv[3], w)
#undef B2_Q5_STEP
        return k4_finish_raw(hdr, c, dlo, dhi, x.e[0], x.d0);
    } else if (TYPE == B2_TYPE_Q6_K) {
        return q6k_quarter(row + kb * 210, c >> 1, c & 1, x.xv, x.e, x.d0);
    } else {
        const uint8_t* blk = row + (kb * 8 + 2 * c) * 34;
        return q80_block(blk, x.xv[0], x.xv[1], x.d0) + q80_block(blk + 34, x.xv[2], x.xv[3], x.d1);
    }
}

// ---- Q4_K: 8 rows x two blocks on the tensor cores ---------------------------------------------------------------
// A (16 x 32 bytes per MMA) = the 32-byte chunk j of block kb0 for the tile's 8 rows (A rows 0-7) and of block
// kb0 + 1 for the same 8 rows (A rows 8-15).  B's eight columns carry, for two chunks at a time (accumulator set A:
// chunks 0 and 1, set B: chunks 2 and 3), the activations that belong to them:
//     column c: chunk parity c >> 2, block (c >> 1) & 1, low / high nibbles c & 1
// and for the MMA of chunk j the columns of the other chunk parity are zero, so the two chunks of a set accumulate
// into the same registers without mixing.  Lane (g, t) provides column g and receives columns 2t, 2t+1: of weight row
// g, block kb0 + (t & 1), chunk (t >> 1) [+ 2] -- in c0, c1 when t is even (A row g), in c2, c3 when odd (A row g + 8).
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_u8s8(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
// lane-constant byte offsets into a block's 256 int8 activations (swizzled granules) and sums
struct MmaLane {
    uint32_t xo[4];   // B fragments: set A b0, b1, set B b0, b1
    uint32_t so;      // byte offset of this lane's first int16 sum pair inside the block's 8 sums of 32
    bool even_chunk;  // this lane's column belongs to the even chunk of a set (g < 4)
    bool odd_t;       // results in c2, c3
};
__device__ __forceinline__ MmaLane mma_lane(int lane) {
    const int g = lane >> 2, t = lane & 3;
    const int cp = g >> 2, bsel = (g >> 1) & 1, nh = g & 1;
    MmaLane m;
    const int sub_a = 2 * cp + nh, sub_b = 4 + 2 * cp + nh;
    m.xo[0] = (uint32_t)(bsel * 256 + 16 * xswz(2 * sub_a) + 4 * t);
    m.xo[1] = (uint32_t)(bsel * 256 + 16 * xswz(2 * sub_a + 1) + 4 * t);
    m.xo[2] = (uint32_t)(bsel * 256 + 16 * xswz(2 * sub_b) + 4 * t);
    m.xo[3] = (uint32_t)(bsel * 256 + 16 * xswz(2 * sub_b + 1) + 4 * t);
    m.so = (uint32_t)((t & 1) * 16 + 4 * (t >> 1));
    m.even_chunk = cp == 0;
    m.odd_t = (t & 1) != 0;
    return m;
}
// a lane's share of the activations of blocks kb0, kb0 + 1, ready for the eight MMAs and the finish: the B fragments
// of accumulator sets A and B, split into "column of an even chunk" / "column of an odd chunk" (one of each pair is 0),
// the sums of the lane's sub-blocks and the block scale.  Constant over a stage when a row has one K chunk.
struct XMma {
    uint32_t ea0, ea1, oa0, oa1, eb0, eb1, ob0, ob1;
    int bs_a, bs_b;
    float dx;
    bool live;   // false: this lane finishes a block beyond the end of the row (its result is dropped, whatever the stale bytes give)
};
// xq_s / bs32_s / xd_s: shared-memory addresses of the activation vector's parts; kb0: first of the warp's two blocks
// (block kb0 + 1 may lie beyond the row: n_valid = 1)
__device__ __forceinline__ void load_x_mma(XMma& x, const MmaLane& ml, uint32_t xq_s, uint32_t bs32_s, uint32_t xd_s, int kb0, int n_valid) {
    const uint32_t xb = xq_s + (uint32_t)kb0 * 256u;
    // a lane whose column belongs to the second block contributes zeros when that block does not exist
    const bool col_ok = n_valid > 1 || ml.xo[0] < 256u;
    uint32_t xa0 = 0, xa1 = 0, xb0 = 0, xb1 = 0;
    if (col_ok) {
        xa0 = lds32(xb + ml.xo[0]); xa1 = lds32(xb + ml.xo[1]);
        xb0 = lds32(xb + ml.xo[2]); xb1 = lds32(xb + ml.xo[3]);
    }
    x.ea0 = ml.even_chunk ? xa0 : 0u; x.ea1 = ml.even_chunk ? xa1 : 0u; x.oa0 = ml.even_chunk ? 0u : xa0; x.oa1 = ml.even_chunk ? 0u : xa1;
    x.eb0 = ml.even_chunk ? xb0 : 0u; x.eb1 = ml.even_chunk ? xb1 : 0u; x.ob0 = ml.even_chunk ? 0u : xb0; x.ob1 = ml.even_chunk ? 0u : xb1;
    const uint32_t blk = (uint32_t)kb0 + (ml.odd_t ? 1u : 0u);
    const uint32_t sa = bs32_s + (uint32_t)kb0 * 16u + ml.so;
    x.bs_a = (int)lds32(sa);
    x.bs_b = (int)lds32(sa + 8);
    x.live = n_valid > 1 || !ml.odd_t;
    x.dx = x.live ? __uint_as_float(lds32(xd_s + blk * 4u)) : 0.0f;
}
// lm: this lane's ldmatrix address for chunk 0; hdr: first byte of block kb0 + (t & 1) in weight row g.
// Returns this lane's share of row g's dot product.
__device__ __forceinline__ float q4k_rows8_blocks2_mma(uint32_t lm, const uint8_t* hdr, const MmaLane& ml, const XMma& x) {
    int ra[4] = {0, 0, 0, 0}, ha[4] = {0, 0, 0, 0}, rb[4] = {0, 0, 0, 0}, hb[4] = {0, 0, 0, 0};
    uint32_t a[4], h[4];
    ldmatrix_x4(lm, a);
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = a[i] & 0xF0F0F0F0u;
    mma_u8s8(ra, a, x.ea0, x.ea1);
    mma_u8s8(ha, h, x.ea0, x.ea1);
    ldmatrix_x4(lm + 64, a);
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = a[i] & 0xF0F0F0F0u;
    mma_u8s8(rb, a, x.eb0, x.eb1);
    mma_u8s8(hb, h, x.eb0, x.eb1);
    ldmatrix_x4(lm + 32, a);
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = a[i] & 0xF0F0F0F0u;
    mma_u8s8(ra, a, x.oa0, x.oa1);
    mma_u8s8(ha, h, x.oa0, x.oa1);
    ldmatrix_x4(lm + 96, a);
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = a[i] & 0xF0F0F0F0u;
    mma_u8s8(rb, a, x.ob0, x.ob1);
    mma_u8s8(hb, h, x.ob0, x.ob1);
    // raw = lo + 16 hi against the low-nibble sub-block (first column), 16 hi against the high-nibble sub-block (second)
    const int lo_a = ml.odd_t ? ra[2] - ha[2] : ra[0] - ha[0], hi_a = (ml.odd_t ? ha[3] : ha[1]) >> 4;
    const int lo_b = ml.odd_t ? rb[2] - hb[2] : rb[0] - hb[0], hi_b = (ml.odd_t ? hb[3] : hb[1]) >> 4;
    // scales and mins of sub-blocks 2c, 2c+1 (set A: c = t >> 1 < 2) and 2c+4, 2c+5 (set B) of this lane's block
    const uint4 hd = *(const uint4*)hdr;
    const uint32_t sh = (ml.so & 4u) << 2;   // 16 for the odd chunk of a set
    const uint32_t y = hd.y >> sh, z = hd.z >> sh, w = hd.w >> sh;
    const uint32_t sc_a = y & 0x3f3fu, mn_a = z & 0x3f3fu;
    const uint32_t sc_b = (w & 0x0f0fu) | ((y >> 2) & 0x3030u), mn_b = ((w >> 4) & 0x0f0fu) | ((z >> 2) & 0x3030u);
    // the integer part of ggml's block sum restricted to this lane's four sub-blocks
    const int isum = (int)(sc_a & 0xffu) * lo_a + (int)(sc_a >> 8) * hi_a + (int)(sc_b & 0xffu) * lo_b + (int)(sc_b >> 8) * hi_b;
    const int msum = (int)(mn_a & 0xffu) * (int)(int16_t)((uint32_t)x.bs_a & 0xffffu) + (int)(mn_a >> 8) * (x.bs_a >> 16) +
                     (int)(mn_b & 0xffu) * (int)(int16_t)((uint32_t)x.bs_b & 0xffffu) + (int)(mn_b >> 8) * (x.bs_b >> 16);
    const float d = h2f_bits(hd.x & 0xffffu), dmin = h2f_bits(hd.x >> 16);
    const float v = x.dx * (d * (float)isum - dmin * (float)msum);
    return x.live ? v : 0.0f;
}

// sums over the warp of two values at once: lane 0 returns the total of a, lane 16 the total of b
__device__ __forceinline__ float warp_sum_pair(float a, float b, int lane) {
    const bool hi = (lane & 16) != 0;
    const float send = hi ? a : b;
    float keep = hi ? b : a;
    keep += __shfl_xor_sync(0xffffffffu, send, 16);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) keep += __shfl_xor_sync(0xffffffffu, keep, o);
    return keep;
}

}  // namespace teamdev
}  // namespace b2
