// matvec_team.cuh -- the mat-vec inner loop of the persistent decode kernel, second design (step2.cu).
//
// What the first design (gemv_device.cuh: dot_chunk) paid for: every lane re-read its 64 int8 activations from shared
// memory for every 64 weights (64 B of x per 32 B of Q4_K), and a decoder warp rewrote the 6-bit scale headers into a
// second shared-memory table that the compute lanes read back.  ncu showed the shared-memory pipe, not HBM, saturated
// inside the tile phases (L1/TEX 38 % over a step whose tile phases are half of it, 20 % of the wavefronts bank
// conflicts).  Here the activations live in REGISTERS for a whole stage:
//   * a compute warp owns one K-slice of 8 super-blocks (2048 weights per row) for the whole stage: lane (b, c) =
//     (block b of the slice, quarter c of the block) holds its 64 int8 activations, their scale and partial sums in
//     22 registers, loaded once after the activation vector has been built;
//   * the warps of a CTA form T teams of WPT slices (K = 4096: 4 teams of 2 warps; K = 14336: one team of 7); a team
//     walks whole weight rows, so a row's 144-byte blocks are read once, by 16-byte loads that are conflict-free by
//     construction (consecutive blocks sit 36 words apart), and nothing else is read from shared memory;
//   * the 6-bit scales and mins of a lane's two sub-blocks are unpacked by the lane itself from the block header it
//     loads with the quants (13 integer instructions) -- no decoder warp, no second barrier per ring slot;
//   * the per-row sums are reduced over the 32 lanes two rows at a time (5 shuffles per pair).
// The arithmetic per (row, block quarter) is the first design's, instruction for instruction (k4_finish etc.), so the
// partial products are bit-identical; only the order in which quarters are added differs.
#pragma once
#include "gemv_device.cuh"

namespace b2 {
namespace teamdev {

using namespace gemvdev;

constexpr int TS_NSLOT = 8;                                            // ring allocations in flight (barrier pairs)
constexpr uint32_t TS_RING_BYTES = 3 * TILE_ROWS * (16 * 210 + 16);   // 81 024 B: three 8-row Q6_K steps at K = 4096
constexpr uint32_t TS_STEP_CAP = TS_RING_BYTES / 3;                    // largest allocation (one step)

// how the 8 compute warps split a row of K elements
struct TeamGeo {
    int nblk;   // 256-element blocks per row
    int S;      // 8-block slices per row
    int WPT;    // warps per team (slices worked on at the same time)
    int T;      // teams = rows in flight per step
    int nsl;    // slices per warp (1 unless a row has more than 8 slices)
};
__host__ __device__ inline TeamGeo team_geo(int K) {
    TeamGeo g;
    g.nblk = K >> 8;
    g.S = (g.nblk + 7) >> 3;
    if (g.S <= NCW) {
        g.WPT = g.S;
        g.T = NCW / g.S;          // 8, 4, 2, 2, 1, 1, 1, 1
        g.nsl = 1;
    } else {
        g.WPT = NCW;
        g.T = 1;
        g.nsl = (g.S + NCW - 1) / NCW;
    }
    return g;
}
// bytes between consecutive rows of a step in the ring (row + 16 bytes of alignment slack, 16-byte multiple)
__host__ __device__ inline uint32_t ts_pitch(int row_bytes) { return (((uint32_t)row_bytes + 15u) & ~15u) + 16u; }
// rows of a tile that share one ring allocation: 8, or fewer when 8 rows do not fit TS_STEP_CAP
__host__ __device__ inline int ts_rows_per_step(uint32_t pitch) {
    int R = TILE_ROWS;
    while (R > 1 && (uint32_t)R * pitch > TS_STEP_CAP) R >>= 1;
    return R;
}

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
