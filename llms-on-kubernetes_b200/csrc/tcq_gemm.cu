// tcq_gemm.cu -- quantised-weight mat-mul for decode batches: TMA (raw blocks) -> dequantising warps -> swizzled fp16
// tiles -> tcgen05.mma -> tensor memory (SURVEY 8a row a12, 2 <= n_tokens <= 128; north_star's "dequant fused into the
// TMA -> smem -> tcgen05 producer").
//
// One persistent CTA per SM, fifteen warps with fixed roles:
//   warp 0      weight producer: per 256-element block column one TMA box of 128 rows x (144 | 224) raw bytes of the
//               quantised tensor into a ring of 3 stages (the tensor map's element type is a byte; a box starts on a
//               16-byte boundary of the row, so a 210-byte Q6_K block sits 0 .. 14 bytes into its 224-byte slot and
//               is read with 4-byte loads and funnel shifts)
//   warp 2      activation producer: per 64-element K step one BN x 64 fp16 tile (TMA, 128-byte swizzle)
//   warps 7-14  dequantise: thread (row, half) turns 32 weights of its row per K step into fp16 -- exactly the value
//               dequant.cuh gives, rounded once -- and stores them as 16-byte chunks into the K-major 128-byte-swizzled
//               tile the UMMA descriptor expects (chunk c of row r at 128 r + 16 (c ^ (r & 7)))
//   warp 1      tcgen05.mma 128 x BN x 16, four per K step, accumulators in tensor memory (two, alternating)
//   warps 3-6   epilogue: tcgen05.ld, fp32 partial sums to the output slices
// Work is cut stream-K fashion: the (tile, block column) pairs of the whole matrix are numbered tile-major and every
// CTA takes an equal contiguous share, so each SM streams the same number of bytes whatever the matrix's shape;
// a share that ends inside a tile leaves a partial sum.  Partial sums of a tile go to consecutive slices (the k-th
// CTA to touch a tile writes slice k), slices a tile does not use are zero-filled by the CTA that finishes the tile,
// and whoever consumes the output adds the slices in order: deterministic.
//
// Dequantisation arithmetic: a 4- or 6-bit code c sits in the low bits of a 16-bit lane, i.e. it IS the fp16
// subnormal c * 2^-24; the conversion to fp32 is exact and fma(c * 2^-24, (d * sc) * 2^24, -(dmin * m)) has the exact
// product d * sc * c inside: the same single rounding as ggml's (d * sc) * c - dmin * m.
#include "tcq_gemm.cuh"

#include <cuda.h>

#include <algorithm>

#include "common.h"
#include "dequant.cuh"
#include "device_utils.cuh"
#include "tc_ptx.cuh"

namespace b2 {
namespace {

using namespace tcptx;

constexpr int TQ_BM = 128, TQ_BK = 64;
constexpr int TQ_RAW_PITCH_MAX = 224;                      // bytes per row of a raw stage: 144 (Q4_K) or 224 (Q6_K: 210 -> next multiple of 16)
constexpr int TQ_RAW_BYTES = TQ_BM * TQ_RAW_PITCH_MAX;     // 28 672
constexpr int TQ_RSTAGES = 3;
constexpr int TQ_DQ_WARPS = 8, TQ_DQ_WARP0 = 7;
constexpr int TQ_THREADS = (TQ_DQ_WARP0 + TQ_DQ_WARPS) * 32;   // 480

template <int BN>
struct TqGeo {
    static constexpr int NSLOT = BN <= 32 ? 6 : (BN == 64 ? 5 : 4);
    static constexpr int A_BYTES = TQ_BM * TQ_BK * 2;
    static constexpr int X_BYTES = BN * TQ_BK * 2;
    static constexpr int SLOT_BYTES = A_BYTES + X_BYTES;
    static constexpr int N_BARS = 2 * NSLOT + 2 * TQ_RSTAGES + 4;
    static constexpr int SMEM = NSLOT * SLOT_BYTES + TQ_RSTAGES * TQ_RAW_BYTES + 1024 + 8 * N_BARS + 16;
    static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
    static constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TQ_BM >> 4) << 24);
};

struct TcqParams {
    int n_seg;
    int seg_tile0[TCQ_MAX_SEGS + 1];   // first m-tile of every segment; [n_seg] = tiles_m
    int seg_row0[TCQ_MAX_SEGS];        // first stacked output row
    int seg_rows[TCQ_MAX_SEGS];
    int seg_type[TCQ_MAX_SEGS];
    int N, K, T, nblk, tiles_m, splits;
    float* ws;
};

// the CTA's share of the tile-major (tile, block column) sequence, one tile segment at a time
struct Walk {
    long long step, hi;
    int nblk;
    __device__ Walk(const TcqParams& p) : nblk(p.nblk) {
        const long long total = (long long)p.tiles_m * p.nblk;
        step = total * blockIdx.x / gridDim.x;
        hi = total * (blockIdx.x + 1) / gridDim.x;
    }
    __device__ bool next(int* tile, int* kb0, int* kb1) {
        if (step >= hi) return false;
        *tile = (int)(step / nblk);
        *kb0 = (int)(step - (long long)*tile * nblk);
        const long long n = min((long long)(nblk - *kb0), hi - step);
        *kb1 = *kb0 + (int)n;
        step += n;
        return true;
    }
};
// first CTA whose share contains step s: the c with floor(c total / G) <= s < floor((c + 1) total / G)
__host__ __device__ inline int cta_of_step(long long s, long long total, int G) { return (int)(((s + 1) * G - 1) / total); }

__device__ __forceinline__ int seg_of_tile(const TcqParams& p, int tile) {
    int si = 0;
    while (si + 1 < p.n_seg && tile >= p.seg_tile0[si + 1]) si++;
    return si;
}

__device__ __forceinline__ uint32_t pack_sat(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ float lo_f(uint32_t x) { return __low2float(*(const __half2*)&x); }
__device__ __forceinline__ float hi_f(uint32_t x) { return __high2float(*(const __half2*)&x); }

__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}

// ---- 32 weights of one row and one 64-element K step -> 64 bytes of the swizzled fp16 tile --------------------------------
// Q4_K, K step g of the block: bytes qs[32 g + 16 h ..+16) hold elements 16 h .. 16 h + 15 of sub-block 2 g (low nibbles)
// and of sub-block 2 g + 1 (high nibbles): chunks 2 h, 2 h + 1 and 4 + 2 h, 5 + 2 h of the row.
__device__ __forceinline__ void dequant_q4k(uint32_t raw_row, uint32_t a_row, int sw, int g, int h) {
    const uint4 hd = lds128(raw_row);   // d | dmin, scales[12]
    const __half2 dd = *(const __half2*)&hd.x;
    const float d = __low2float(dd), dmin = __high2float(dd);
    // get_scale_min_k4 for sub-blocks 2 g and 2 g + 1 from the three words of packed 6-bit scales (no byte array: registers only)
    int s0, m0, s1, m1;
    if (g < 2) {
        const int sh = 16 * g;
        s0 = (hd.y >> sh) & 63; m0 = (hd.z >> sh) & 63;
        s1 = (hd.y >> (sh + 8)) & 63; m1 = (hd.z >> (sh + 8)) & 63;
    } else {
        const int sh = 16 * (g - 2);
        s0 = ((hd.w >> sh) & 0xF) | (((hd.y >> (sh + 6)) & 3) << 4);
        m0 = ((hd.w >> (sh + 4)) & 0xF) | (((hd.z >> (sh + 6)) & 3) << 4);
        s1 = ((hd.w >> (sh + 8)) & 0xF) | (((hd.y >> (sh + 14)) & 3) << 4);
        m1 = ((hd.w >> (sh + 12)) & 0xF) | (((hd.z >> (sh + 14)) & 3) << 4);
    }
    const float ds0 = (d * (float)s0) * 16777216.0f, dm0 = -(dmin * (float)m0);
    const float ds1 = (d * (float)s1) * 16777216.0f, dm1 = -(dmin * (float)m1);
    const uint4 q = lds128(raw_row + 16 + 32 * g + 16 * h);
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    uint32_t lo[8], hi[8];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const uint32_t a = w[i] & 0x000F000Fu, b = (w[i] >> 8) & 0x000F000Fu;     // elements (0, 2), (1, 3) of the word, low nibbles
        const uint32_t c = (w[i] >> 4) & 0x000F000Fu, e = (w[i] >> 12) & 0x000F000Fu;
        lo[2 * i] = pack_sat(fmaf(lo_f(a), ds0, dm0), fmaf(lo_f(b), ds0, dm0));
        lo[2 * i + 1] = pack_sat(fmaf(hi_f(a), ds0, dm0), fmaf(hi_f(b), ds0, dm0));
        hi[2 * i] = pack_sat(fmaf(lo_f(c), ds1, dm1), fmaf(lo_f(e), ds1, dm1));
        hi[2 * i + 1] = pack_sat(fmaf(hi_f(c), ds1, dm1), fmaf(hi_f(e), ds1, dm1));
    }
    sts128(a_row + (uint32_t)(((2 * h) ^ sw) << 4), lo[0], lo[1], lo[2], lo[3]);
    sts128(a_row + (uint32_t)(((2 * h + 1) ^ sw) << 4), lo[4], lo[5], lo[6], lo[7]);
    sts128(a_row + (uint32_t)(((4 + 2 * h) ^ sw) << 4), hi[0], hi[1], hi[2], hi[3]);
    sts128(a_row + (uint32_t)(((5 + 2 * h) ^ sw) << 4), hi[4], hi[5], hi[6], hi[7]);
}

// NW 32-bit words starting at a 2-byte-aligned shared-memory address
template <int NW>
__device__ __forceinline__ void lds_words_even(uint32_t addr, uint32_t (&out)[NW]) {
    const uint32_t a = addr & ~3u, sh = (addr & 2u) << 3;
    uint32_t w[NW + 1];
#pragma unroll
    for (int i = 0; i <= NW; i++) asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w[i]) : "r"(a + 4 * i) : "memory");
#pragma unroll
    for (int i = 0; i < NW; i++) out[i] = __funnelshift_r(w[i], w[i + 1], sh);
}

// Q6_K, K step g = 2 n + u of the block: elements 128 n + 64 u + 32 h + l (l < 32) = quarter qi = 2 u + h of half n:
// ql[64 n + 32 h + l] nibble u, qh[32 n + l] bits 2 qi .. 2 qi + 1, scale[8 n + 2 qi + l / 16], minus 32.
// blk: the block's first byte in shared memory, 2-byte aligned only (210-byte blocks; a TMA box starts on a 16-byte
// boundary of the row, so the block sits 0 .. 14 bytes into its slot).
__device__ __forceinline__ void dequant_q6k(uint32_t blk, uint32_t a_row, int sw, int g, int h) {
    const int n = g >> 1, u = g & 1, qi = 2 * u + h;
    uint32_t tail;   // scales[8 n + 2 qi], [.. + 1]
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(tail) : "r"(blk + 192 + 8 * n + 2 * qi) : "memory");
    uint32_t dbits;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(dbits) : "r"(blk + 208) : "memory");
    const float d = __half2float(__ushort_as_half((unsigned short)dbits));
    const float dsA = d * (float)(int8_t)(tail & 0xff), dsB = d * (float)(int8_t)(tail >> 8);
    const float ds[2] = {dsA * 16777216.0f, dsB * 16777216.0f}, off[2] = {-32.0f * dsA, -32.0f * dsB};
    uint32_t ql[8], qh[8];
    lds_words_even<8>(blk + 64 * n + 32 * h, ql);
    lds_words_even<8>(blk + 128 + 32 * n, qh);
    uint32_t o[16];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t q = ((ql[i] >> (4 * u)) & 0x0F0F0F0Fu) | (((qh[i] >> (2 * qi)) & 0x03030303u) << 4);
        const uint32_t p01 = __byte_perm(q, 0u, 0x4140), p23 = __byte_perm(q, 0u, 0x4342);
        const float s = ds[i >> 2], f = off[i >> 2];
        o[2 * i] = pack_sat(fmaf(lo_f(p01), s, f), fmaf(hi_f(p01), s, f));
        o[2 * i + 1] = pack_sat(fmaf(lo_f(p23), s, f), fmaf(hi_f(p23), s, f));
    }
#pragma unroll
    for (int c = 0; c < 4; c++) sts128(a_row + (uint32_t)(((4 * h + c) ^ sw) << 4), o[4 * c], o[4 * c + 1], o[4 * c + 2], o[4 * c + 3]);
}

// 32 lanes x 16 consecutive fp32 columns of tensor memory
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct TcqMaps {
    CUtensorMap w[TCQ_MAX_SEGS];
    CUtensorMap x;
};

template <int BN>
__global__ void __launch_bounds__(TQ_THREADS, 1) tcq_gemm_kernel(const __grid_constant__ TcqMaps maps, const TcqParams p) {
    using G = TqGeo<BN>;
    extern __shared__ uint8_t tq_smem[];
    const uint32_t slots0 = (smem_u32(tq_smem) + 1023u) & ~1023u;
    const uint32_t raw0 = slots0 + G::NSLOT * G::SLOT_BYTES;
    const uint32_t bars = raw0 + TQ_RSTAGES * TQ_RAW_BYTES;
    auto ab_full = [&](int s) { return bars + 8u * s; };
    auto ab_empty = [&](int s) { return bars + 8u * (G::NSLOT + s); };
    auto raw_full = [&](int s) { return bars + 8u * (2 * G::NSLOT + s); };
    auto raw_empty = [&](int s) { return bars + 8u * (2 * G::NSLOT + TQ_RSTAGES + s); };
    auto acc_full = [&](int i) { return bars + 8u * (2 * G::NSLOT + 2 * TQ_RSTAGES + i); };
    auto acc_empty = [&](int i) { return bars + 8u * (2 * G::NSLOT + 2 * TQ_RSTAGES + 2 + i); };
    const uint32_t tmem_slot = bars + 8u * G::N_BARS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < G::NSLOT; s++) {
            bar_init(ab_full(s), TQ_DQ_WARPS + 1);   // the dequantising warps and the activation tile's byte count
            bar_init(ab_empty(s), 1);                // tcgen05.commit
        }
        for (int s = 0; s < TQ_RSTAGES; s++) {
            bar_init(raw_full(s), 1);
            bar_init(raw_empty(s), TQ_DQ_WARPS);
        }
        for (int i = 0; i < 2; i++) {
            bar_init(acc_full(i), 1);
            bar_init(acc_empty(i), 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (int i = 0; i < p.n_seg; i++) asm volatile("prefetch.tensormap [%0];" :: "l"(&maps.w[i]) : "memory");
        asm volatile("prefetch.tensormap [%0];" :: "l"(&maps.x) : "memory");
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

    Walk walk(p);
    int tile, kb0, kb1;
    if (warp == 0) {
        // ===== raw weight blocks =====
        if (lane == 0) {
            int rs = 0;
            uint32_t ph = 0;
            while (walk.next(&tile, &kb0, &kb1)) {
                const int si = seg_of_tile(p, tile);
                const int q4 = p.seg_type[si] == B2_TYPE_Q4_K;
                const int bb = q4 ? 144 : 210;
                const uint32_t box = (uint32_t)TQ_BM * (q4 ? 144u : 224u);
                const int m0 = (tile - p.seg_tile0[si]) * TQ_BM;
                for (int kb = kb0; kb < kb1; kb++) {
                    bar_wait(raw_empty(rs), ph ^ 1);
                    bar_expect_tx(raw_full(rs), box);
                    tma_load_2d(raw0 + rs * TQ_RAW_BYTES, &maps.w[si], raw_full(rs), (kb * bb) & ~15, m0);   // a box starts on 16 bytes
                    if (++rs == TQ_RSTAGES) { rs = 0; ph ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 2) {
        // ===== activation tiles =====
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            while (walk.next(&tile, &kb0, &kb1)) {
                for (int ks = 4 * kb0; ks < 4 * kb1; ks++) {
                    bar_wait(ab_empty(s), ph ^ 1);
                    bar_expect_tx(ab_full(s), G::X_BYTES);
                    tma_load_2d(slots0 + s * G::SLOT_BYTES + G::A_BYTES, &maps.x, ab_full(s), ks * TQ_BK, 0);
                    if (++s == G::NSLOT) { s = 0; ph ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== tensor cores =====
        if (lane == 0) {
            int s = 0, j = 0;
            uint32_t ph = 0;
            for (; walk.next(&tile, &kb0, &kb1); j++) {
                const int as = j & 1;
                bar_wait(acc_empty(as), ((j >> 1) & 1) ^ 1);   // the epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
                const int n_ks = 4 * (kb1 - kb0);
                for (int ks = 0; ks < n_ks; ks++) {
                    bar_wait(ab_full(s), ph);
                    tc_fence_after();
                    const uint32_t sa = slots0 + s * G::SLOT_BYTES;
                    const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + G::A_BYTES);
#pragma unroll
                    for (int k = 0; k < TQ_BK / 16; k++) umma_f16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), G::IDESC, (ks | k) != 0);
                    umma_commit(ab_empty(s));
                    if (++s == G::NSLOT) { s = 0; ph ^= 1; }
                }
                umma_commit(acc_full(as));
            }
        }
        __syncwarp();
    } else if (warp >= TQ_DQ_WARP0) {
        // ===== dequantisation: two threads per weight row =====
        const int td = threadIdx.x - TQ_DQ_WARP0 * 32;
        const int row = td >> 1, h = td & 1;
        const int sw = row & 7;
        int s = 0, rs = 0;
        uint32_t ph = 0, rph = 0;
        while (walk.next(&tile, &kb0, &kb1)) {
            const int q4 = p.seg_type[seg_of_tile(p, tile)] == B2_TYPE_Q4_K;
            const uint32_t pitch = q4 ? 144u : 224u;
            for (int kb = kb0; kb < kb1; kb++) {
                bar_wait(raw_full(rs), rph);
                const uint32_t raw_row = raw0 + rs * TQ_RAW_BYTES + row * pitch;
#pragma unroll
                for (int g = 0; g < 4; g++) {
                    bar_wait(ab_empty(s), ph ^ 1);
                    const uint32_t a_row = slots0 + s * G::SLOT_BYTES + row * 128;
                    if (q4) dequant_q4k(raw_row, a_row, sw, g, h);
                    else dequant_q6k(raw_row + (uint32_t)((kb * 210) & 15), a_row, sw, g, h);
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores, read by the tensor cores
                    __syncwarp();
                    if (lane == 0) bar_arrive(ab_full(s));
                    if (++s == G::NSLOT) { s = 0; ph ^= 1; }
                }
                __syncwarp();
                if (lane == 0) bar_arrive(raw_empty(rs));
                if (++rs == TQ_RSTAGES) { rs = 0; rph ^= 1; }
            }
        }
    } else {
        // ===== epilogue (warps 3..6): partial sums of the segment to its slice; the tile's last segment clears the rest =====
        const int q = warp & 3;   // the quarter of the 128 tensor-memory lanes this warp may read
        const long long total = (long long)p.tiles_m * p.nblk;
        const size_t slice = (size_t)p.T * p.N;
        for (int j = 0; walk.next(&tile, &kb0, &kb1); j++) {
            const int as = j & 1;
            const int si = seg_of_tile(p, tile);
            const int r = (tile - p.seg_tile0[si]) * TQ_BM + q * 32 + lane;   // row inside the segment
            const bool row_ok = r < p.seg_rows[si];
            const int n = p.seg_row0[si] + r;
            const int sl = (int)blockIdx.x - cta_of_step((long long)tile * p.nblk, total, (int)gridDim.x);
            bar_wait(acc_full(as), (j >> 1) & 1);
            tc_fence_after();
            float* const o = p.ws + (size_t)sl * slice + n;
#pragma unroll 1
            for (int c = 0; c < BN / 16; c++) {
                if (c * 16 >= p.T) break;
                uint32_t v[16];
                tmem_ld_32x16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + c * 16), v);
                if (row_ok) {
#pragma unroll
                    for (int i = 0; i < 16; i++)
                        if (c * 16 + i < p.T) o[(size_t)(c * 16 + i) * p.N] = __uint_as_float(v[i]);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) bar_arrive(acc_empty(as));
            if (kb1 == p.nblk && row_ok) {
                for (int z = sl + 1; z < p.splits; z++)
                    for (int t = 0; t < p.T; t++) p.ws[(size_t)z * slice + (size_t)t * p.N + n] = 0.0f;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(G::TMEM_COLS) : "memory");
    }
}

// out[t][n] (+)= sum of the slices in slice order (+ bias[n])
__global__ void __launch_bounds__(256) tcq_reduce_kernel(const float* __restrict__ ws, int splits, size_t slice, int T, int N, float* __restrict__ out,
                                                         int64_t ldo, const float* __restrict__ bias, int epi) {
    const int64_t n4 = (int64_t)T * N / 4;
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < n4; g += (int64_t)gridDim.x * blockDim.x) {
        const int64_t e = g * 4;
        const int t = (int)(e / N), n = (int)(e - (int64_t)t * N);
        float4 acc = *(const float4*)(ws + e);
        for (int s = 1; s < splits; s++) {
            const float4 v = *(const float4*)(ws + (size_t)s * slice + e);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        if (bias) {
            const float4 b = *(const float4*)(bias + n);
            acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
        }
        float4* o = (float4*)(out + (int64_t)t * ldo + n);
        if (epi == TC_EPI_ADD) {
            const float4 r = *o;
            acc.x += r.x; acc.y += r.y; acc.z += r.z; acc.w += r.w;
        }
        *o = acc;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int tq_encode_fn(EncodeTiledFn* out) {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* q = nullptr;
        cudaDriverEntryPointQueryResult r;
        B2_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r));
        if (r != cudaDriverEntryPointSuccess || !q) {
            set_error("tcq_gemm: the driver does not export cuTensorMapEncodeTiled");
            return B2_ECUDA;
        }
        fn = (EncodeTiledFn)q;
    }
    *out = fn;
    return 0;
}

// the quantised tensor as a [rows][row_bytes] byte matrix, boxes of 128 rows x box_bytes, no swizzle, zero fill outside
int encode_raw_rows(CUtensorMap* map, const void* base, int64_t rows, int64_t row_bytes, int box_bytes) {
    EncodeTiledFn fn;
    int rc;
    if ((rc = tq_encode_fn(&fn))) return rc;
    const cuuint64_t dims[2] = {(cuuint64_t)row_bytes, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)row_bytes};
    const cuuint32_t box[2] = {(cuuint32_t)box_bytes, (cuuint32_t)TQ_BM};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("tcq_gemm: cuTensorMapEncodeTiled failed (%d) for a [%lld][%lld]-byte tensor", (int)r, (long long)rows, (long long)row_bytes);
        return B2_ECUDA;
    }
    return 0;
}

int encode_x_rows(CUtensorMap* map, const void* base, int64_t rows, int64_t K, int box_rows) {
    EncodeTiledFn fn;
    int rc;
    if ((rc = tq_encode_fn(&fn))) return rc;
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    const cuuint32_t box[2] = {(cuuint32_t)TQ_BK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("tcq_gemm: cuTensorMapEncodeTiled failed (%d) for the activations [%lld][%lld]", (int)r, (long long)rows, (long long)K);
        return B2_ECUDA;
    }
    return 0;
}

int tq_sm_count() {
    const int n = device_sm_count();
    return n > 0 ? n : 148;
}

// slices needed when `tiles` tiles of `nblk` block columns are shared out over G CTAs: the most CTAs any tile meets
int tq_slices(int tiles, int nblk, int G) {
    const long long total = (long long)tiles * nblk;
    int s = 1;
    for (int t = 0; t < tiles; t++) {
        const int first = cta_of_step((long long)t * nblk, total, G), last = cta_of_step((long long)(t + 1) * nblk - 1, total, G);
        s = std::max(s, last - first + 1);
    }
    return s;
}

template <int BN>
int launch_tq(const TcqMaps& maps, const TcqParams& p, int grid, cudaStream_t st) {
    using G = TqGeo<BN>;
    static PerDeviceSize attr_set;
    const int dev = current_device();
    if (attr_set.below(dev, G::SMEM)) {
        B2_CUDA_CHECK(cudaFuncSetAttribute(tcq_gemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM));
        attr_set.set(dev, G::SMEM);
    }
    tcq_gemm_kernel<BN><<<grid, TQ_THREADS, G::SMEM, st>>>(maps, p);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

int tq_tiles(const TcqGemmArgs& a) {
    int t = 0;
    for (int i = 0; i < a.n_seg; i++) t += (a.seg[i].rows + TQ_BM - 1) / TQ_BM;
    return t;
}

}  // namespace

// (a tensor map's row stride is a multiple of 16 bytes: rows of 210-byte Q6_K blocks qualify when K is a multiple of 2048)
bool tcq_supported(int type, int K) {
    if (K <= 0 || K % 256) return false;
    return type == B2_TYPE_Q4_K || (type == B2_TYPE_Q6_K && (K / 256) % 8 == 0);
}

// CTAs for `tiles` tiles of `nblk` block columns: one per SM, fewer when a tile would otherwise be cut into more than
// TQ_MAX_SLICES partial sums (skinny matrices with a long K: the consumer reads every slice)
constexpr int TQ_MAX_SLICES = 8;
static int tq_grid(int tiles, int nblk) {
    const long long total = (long long)tiles * nblk;
    int grid = (int)std::min<long long>(tq_sm_count(), total);
    while (grid > 1 && tq_slices(tiles, nblk, grid) > TQ_MAX_SLICES) grid = grid * 7 / 8;
    return grid;
}

size_t tcq_ws_bytes(int N, int K, int T) {
    // the stacked segments round their tiles up separately: up to TCQ_MAX_SEGS - 1 tiles more than ceil(N / 128)
    const int nblk = K / 256;
    int s = 1;
    for (int extra = 0; extra < TCQ_MAX_SEGS; extra++) {
        const int tiles = (N + TQ_BM - 1) / TQ_BM + extra;
        s = std::max(s, tq_slices(tiles, nblk, tq_grid(tiles, nblk)));
    }
    return (size_t)s * T * N * sizeof(float);
}

int launch_tcq_gemm(const TcqGemmArgs& a, cudaStream_t st, TcParts* parts) {
    if (!parts || a.n_seg < 1 || a.n_seg > TCQ_MAX_SEGS || !a.x || !a.ws || a.K <= 0 || a.K % 256 || a.T < 1 || a.T > TCQ_MAX_T || ((uintptr_t)a.x & 15)) {
        set_error("tcq_gemm: bad arguments (segments %d, K=%d, T=%d)", a.n_seg, a.K, a.T);
        return B2_EINVAL;
    }
    TcqParams p{};
    TcqMaps maps;
    int rc;
    p.n_seg = a.n_seg;
    p.K = a.K;
    p.T = a.T;
    p.nblk = a.K / 256;
    int tile0 = 0, row0 = 0;
    for (int i = 0; i < a.n_seg; i++) {
        const TcqSeg& s = a.seg[i];
        if (!s.w || s.rows < 1 || !tcq_supported(s.type, a.K) || ((uintptr_t)s.w & 15) || (row0 & 3)) {
            set_error("tcq_gemm: segment %d (type %d, %d rows) is not supported", i, s.type, s.rows);
            return B2_EUNSUP;
        }
        const int bb = s.type == B2_TYPE_Q4_K ? 144 : 210;
        if ((rc = encode_raw_rows(&maps.w[i], s.w, s.rows, (int64_t)p.nblk * bb, s.type == B2_TYPE_Q4_K ? 144 : 224))) return rc;
        p.seg_tile0[i] = tile0;
        p.seg_row0[i] = row0;
        p.seg_rows[i] = s.rows;
        p.seg_type[i] = s.type;
        tile0 += (s.rows + TQ_BM - 1) / TQ_BM;
        row0 += s.rows;
    }
    for (int i = a.n_seg; i < TCQ_MAX_SEGS; i++) maps.w[i] = maps.w[0];
    p.seg_tile0[a.n_seg] = tile0;
    p.tiles_m = tile0;
    p.N = row0;
    const int bn = a.T <= 32 ? 32 : (a.T <= 64 ? 64 : 128);
    if ((rc = encode_x_rows(&maps.x, a.x, a.T, a.K, bn))) return rc;
    const int grid = tq_grid(p.tiles_m, p.nblk);
    p.splits = tq_slices(p.tiles_m, p.nblk, grid);
    if ((size_t)p.splits * a.T * p.N * sizeof(float) > a.ws_bytes) {
        set_error("tcq_gemm: workspace of %zu bytes is too small for %d slices of %d x %d", a.ws_bytes, p.splits, a.T, p.N);
        return B2_EINVAL;
    }
    p.ws = a.ws;
    *parts = TcParts{a.ws, p.splits, (size_t)a.T * p.N};
    if (bn == 32) return launch_tq<32>(maps, p, grid, st);
    if (bn == 64) return launch_tq<64>(maps, p, grid, st);
    return launch_tq<128>(maps, p, grid, st);
}

int launch_tcq_reduce(const TcParts& parts, int T, int N, float* out, int64_t ldo, const float* bias, int epi, cudaStream_t st) {
    if (!parts.ws || parts.splits < 1 || !out || N % 4 || ldo % 4) {
        set_error("tcq_reduce: bad arguments");
        return B2_EINVAL;
    }
    const int64_t n4 = (int64_t)T * N / 4;
    const int blocks = (int)std::min<int64_t>((n4 + 255) / 256, 148 * 8);
    tcq_reduce_kernel<<<blocks, 256, 0, st>>>(parts.ws, parts.splits, parts.slice, T, N, out, ldo, bias, epi);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b2
