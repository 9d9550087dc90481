// tcq_gemm.cu -- quantised-weight mat-mul for decode batches: TMA (raw blocks) -> dequantising warps -> tensor memory
// -> tcgen05.mma (SURVEY 8a row a12, 2 <= n_tokens <= 128; north_star's "dequant fused into the TMA -> smem -> tcgen05
// producer").
//
// One persistent CTA per SM, 23 warps with fixed roles:
//   warp 0      weight producer: per 256-element block column one TMA box of 128 rows x (144 | 224) raw bytes of the
//               quantised tensor into a ring of shared-memory stages (the tensor map's element type is a byte; a box
//               starts on a 16-byte boundary of the row, so a 210-byte Q6_K block sits 0 .. 14 bytes into its
//               224-byte slot and is read with 4-byte loads and funnel shifts)
//   warp 2      activation producer: per 64-element K step one BN x 64 fp16 tile (TMA, 128-byte swizzle)
//   warps 7-22  dequantise: a thread owns one weight row (lane 32 (warp % 4) + lane of tensor memory) and one of the
//               block's four K steps (warp group (warp - 7) / 4): 64 weights -> fp16, exactly the value dequant.cuh
//               gives rounded once -> 32 packed registers -> tcgen05.st into the A-operand ring in TENSOR memory.
//               The fp16 weights never touch shared memory: at 3.5 x the size of the raw blocks, writing them there
//               and having the tensor core read them back cost more shared-memory bandwidth than the SM's share of
//               the HBM stream leaves time for (first version of this kernel: 1.25 TB/s).
//   warp 1      tcgen05.mma 128 x BN x 16, A from tensor memory, B (activations) from shared memory, four per K step,
//               accumulators in tensor memory (two, alternating)
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
#include <cstdio>
#include <cstdlib>
#include <vector>

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
constexpr int TQ_DQ_WARPS = 16, TQ_DQ_WARP0 = 7;           // warp 7 + 4 g + q: K step g of the block, lane quarter (7 + q) % 4
constexpr int TQ_THREADS = (TQ_DQ_WARP0 + TQ_DQ_WARPS) * 32;   // 736
constexpr int TQ_NA = 2;                                   // A-operand ring in tensor memory: block columns of 128 columns (4 K steps)
constexpr int TQ_A_COL0 = 256;                             // columns 0 .. 255: the two accumulators; 256 .. 511: the A ring
constexpr int TQ_TMEM_COLS = 512;

// Hand-offs are per 256-element block column, not per K step: the one thread that issues the MMAs needs ~80 cycles per
// tcgen05 instruction and ~100 per barrier wait, and at 4 MMAs + 2 commits + 2 waits per K step it was the bottleneck of
// the first versions (950 cycles per K step, 1.2 TB/s).
template <int BN>
struct TqGeo {
    static constexpr int RSTAGES = BN <= 32 ? 5 : (BN == 64 ? 4 : 3);   // raw weight stages
    static constexpr int NX = BN <= 32 ? 4 : (BN == 64 ? 3 : 2);        // activation stages: BN x 256 fp16 = four swizzled BN x 64 tiles
    static constexpr int XT_BYTES = BN * TQ_BK * 2;                     // one tile
    static constexpr int X_BYTES = 4 * XT_BYTES;
    static constexpr int N_BARS = 2 * TQ_NA + 2 * NX + 2 * RSTAGES + 4;
    static constexpr int SMEM = NX * X_BYTES + RSTAGES * TQ_RAW_BYTES + 1024 + 8 * N_BARS + 16;
    static constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TQ_BM >> 4) << 24);
    // consecutive K = 16 slices go to NACC accumulators in turn and the epilogue adds them in a fixed order.  (Built on the
    // suspicion that MMAs into ONE accumulator serialise at the pipeline's latency; measured, they do not -- dependent
    // M = 128, N = 32 MMAs issue every 26 cycles, tools/ubench_umma.cu; the ~130 cycles per MMA seen before were the
    // divergent issuing thread.  Kept: it is free and shortens each accumulator's fp32 summation chain.)
    static constexpr int NACC = 128 / BN;
};
static_assert(TqGeo<32>::SMEM <= 227 * 1024 && TqGeo<64>::SMEM <= 227 * 1024 && TqGeo<128>::SMEM <= 227 * 1024, "shared memory budget");

struct TcqParams {
    int n_seg;
    int seg_tile0[TCQ_MAX_SEGS + 1];   // first m-tile of every segment; [n_seg] = tiles_m
    int seg_row0[TCQ_MAX_SEGS];        // first stacked output row
    int seg_rows[TCQ_MAX_SEGS];
    int seg_type[TCQ_MAX_SEGS];
    int N, K, T, nblk, tiles_m, splits;
    int fast;  // fp16 conversion, activations stored with the middle two of every four elements swapped (TcqGemmArgs::x_perm)
    int dbg;   // B2_TCQ_DBG (experiments): 1 = no MMAs, 2 = no dequantisation arithmetic, 4 = no tensor-memory stores
    float* ws;
    long long* trace;   // B2_TCQ_TRACE=1: clock64 stamps of CTA 0 ([0..]: MMA thread, 2 per K step; [4096..]: dequantising warp of group 0, 4 per block)
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
    __device__ __forceinline__ bool next(int& tile, int& kb0, int& kb1) {
        if (step >= hi) return false;
        tile = (int)(step / nblk);
        kb0 = (int)(step - (long long)tile * nblk);
        const long long n = min((long long)(nblk - kb0), hi - step);
        kb1 = kb0 + (int)n;
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

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}

// ---- the 64 weights of one row and one 64-element K step -> 32 registers of packed fp16 pairs (K order) -----------------
// Q4_K, K step g of the block: bytes qs[32 g ..+32) hold sub-block 2 g in their low nibbles (elements 0 .. 31 of the
// step) and sub-block 2 g + 1 in their high nibbles (elements 32 .. 63).
__device__ __forceinline__ void dequant_q4k(uint32_t blk, int g, uint32_t (&o)[32]) {
    const uint4 hd = lds128(blk);   // d | dmin, scales[12]
    const __half2 dd = *(const __half2*)&hd.x;
    const float d = __low2float(dd), dmin = __high2float(dd);
    // get_scale_min_k4 for sub-blocks 2 g and 2 g + 1 from the three words of packed 6-bit scales
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
    const uint4 qa = lds128(blk + 16 + 32 * g), qb = lds128(blk + 32 + 32 * g);
    const uint32_t w[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t a = w[i] & 0x000F000Fu, b = (w[i] >> 8) & 0x000F000Fu;     // elements (0, 2), (1, 3) of the word, low nibbles
        const uint32_t c = (w[i] >> 4) & 0x000F000Fu, e = (w[i] >> 12) & 0x000F000Fu;
        o[2 * i] = pack_sat(fmaf(lo_f(a), ds0, dm0), fmaf(lo_f(b), ds0, dm0));
        o[2 * i + 1] = pack_sat(fmaf(hi_f(a), ds0, dm0), fmaf(hi_f(b), ds0, dm0));
        o[16 + 2 * i] = pack_sat(fmaf(lo_f(c), ds1, dm1), fmaf(lo_f(e), ds1, dm1));
        o[16 + 2 * i + 1] = pack_sat(fmaf(hi_f(c), ds1, dm1), fmaf(hi_f(e), ds1, dm1));
    }
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

// Q6_K, K step g = 2 n + u of the block: element 32 j + l of the step (j < 2, l < 32) is quarter qi = 2 u + j of half n:
// ql[64 n + 32 j + l] nibble u, qh[32 n + l] bits 2 qi .. 2 qi + 1, scale[8 n + 2 qi + l / 16], minus 32.
// blk: the block's first byte in shared memory, 2-byte aligned only.
__device__ __forceinline__ void dequant_q6k(uint32_t blk, int g, uint32_t (&o)[32]) {
    const int n = g >> 1, u = g & 1;
    uint32_t dbits, sc4[1];
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(dbits) : "r"(blk + 208) : "memory");
    lds_words_even<1>(blk + 192 + 8 * n + 4 * u, sc4);   // scales[8 n + 4 u ..+4): quarters 2 u (two scales) and 2 u + 1
    const float d = __half2float(__ushort_as_half((unsigned short)dbits));
    float ds[4], off[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float t = d * (float)(int8_t)(sc4[0] >> (8 * k));
        ds[k] = t * 16777216.0f;
        off[k] = -32.0f * t;
    }
    uint32_t ql[16], qh[8];
    lds_words_even<16>(blk + 64 * n, ql);
    lds_words_even<8>(blk + 128 + 32 * n, qh);
#pragma unroll
    for (int j = 0; j < 2; j++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const uint32_t q = ((ql[8 * j + i] >> (4 * u)) & 0x0F0F0F0Fu) | (((qh[i] >> (2 * (2 * u + j))) & 0x03030303u) << 4);
            const uint32_t p01 = __byte_perm(q, 0u, 0x4140), p23 = __byte_perm(q, 0u, 0x4342);
            const float s = ds[2 * j + (i >> 2)], f = off[2 * j + (i >> 2)];
            o[16 * j + 2 * i] = pack_sat(fmaf(lo_f(p01), s, f), fmaf(hi_f(p01), s, f));
            o[16 * j + 2 * i + 1] = pack_sat(fmaf(lo_f(p23), s, f), fmaf(hi_f(p23), s, f));
        }
    }
}

// ---- the same in fp16 arithmetic (TcqGemmArgs::x_perm): 1.6 instead of 3.4 instructions per weight -----------------------
// The code bits are OR-ed into the mantissa of a half whose ulp is the code's unit -- 0x6400 (1024, ulp 1) for a low
// nibble or a byte, 0x5400 (64, ulp 1/16) for a high nibble in place -- the offset is subtracted exactly, and one HFMA2
// applies scale and minimum, both rounded to fp16 first (2^-11 relative each; the exact path rounds once at the end).
// A word's four codes come out as the pairs (0, 2) and (1, 3): the activations are stored with elements 1 and 2 of every
// group of four swapped (launch_rows_to_f16 / launch_swiglu_to_f16 `perm`), so no shuffle is needed.
__device__ __forceinline__ __half2 h2_of(uint32_t x) { return *(const __half2*)&x; }
__device__ __forceinline__ uint32_t u_of(__half2 h) { return *(const uint32_t*)&h; }
__device__ __forceinline__ __half2 h2_clamped(float v) { return __float2half2_rn(fminf(fmaxf(v, -65504.0f), 65504.0f)); }

__device__ __forceinline__ void dequant_q4k_fast(uint32_t blk, int g, uint32_t (&o)[32]) {
    const uint4 hd = lds128(blk);
    const __half2 dd = *(const __half2*)&hd.x;
    const float d = __low2float(dd), dmin = __high2float(dd);
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
    const __half2 ds0 = h2_clamped(d * (float)s0), dm0 = h2_clamped(-(dmin * (float)m0));
    const __half2 ds1 = h2_clamped(d * (float)s1), dm1 = h2_clamped(-(dmin * (float)m1));
    const __half2 k1024 = __float2half2_rn(1024.0f), k64 = __float2half2_rn(64.0f);
    const uint4 qa = lds128(blk + 16 + 32 * g), qb = lds128(blk + 32 + 32 * g);
    const uint32_t w[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t w8 = w[i] >> 8;
        const uint32_t a = (w[i] & 0x000F000Fu) | 0x64006400u, b = (w8 & 0x000F000Fu) | 0x64006400u;   // codes (0, 2), (1, 3), low nibbles
        const uint32_t c = (w[i] & 0x00F000F0u) | 0x54005400u, e = (w8 & 0x00F000F0u) | 0x54005400u;   // high nibbles in place: bits 4..7 of a half in [64, 128) count units of 16 x 1/16 = 1
        o[2 * i] = u_of(__hfma2(__hsub2(h2_of(a), k1024), ds0, dm0));
        o[2 * i + 1] = u_of(__hfma2(__hsub2(h2_of(b), k1024), ds0, dm0));
        o[16 + 2 * i] = u_of(__hfma2(__hsub2(h2_of(c), k64), ds1, dm1));
        o[16 + 2 * i + 1] = u_of(__hfma2(__hsub2(h2_of(e), k64), ds1, dm1));
    }
}

__device__ __forceinline__ void dequant_q6k_fast(uint32_t blk, int g, uint32_t (&o)[32]) {
    const int n = g >> 1, u = g & 1;
    uint32_t dbits, sc4[1];
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(dbits) : "r"(blk + 208) : "memory");
    lds_words_even<1>(blk + 192 + 8 * n + 4 * u, sc4);
    const float d = __half2float(__ushort_as_half((unsigned short)dbits));
    __half2 ds[4];
#pragma unroll
    for (int k = 0; k < 4; k++) ds[k] = h2_clamped(d * (float)(int8_t)(sc4[0] >> (8 * k)));
    const __half2 k1056 = __float2half2_rn(1056.0f);   // 1024 (the magic) + 32 (the code's bias)
    uint32_t ql[16], qh[8];
    lds_words_even<16>(blk + 64 * n, ql);
    lds_words_even<8>(blk + 128 + 32 * n, qh);
#pragma unroll
    for (int j = 0; j < 2; j++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const uint32_t q = ((ql[8 * j + i] >> (4 * u)) & 0x0F0F0F0Fu) | (((qh[i] >> (2 * (2 * u + j))) & 0x03030303u) << 4);
            const uint32_t p02 = __byte_perm(q, 0x64646464u, 0x4240), p13 = __byte_perm(q, 0x64646464u, 0x4341);
            const __half2 s = ds[2 * j + (i >> 2)];
            o[16 * j + 2 * i] = u_of(__hmul2(__hsub2(h2_of(p02), k1056), s));
            o[16 * j + 2 * i + 1] = u_of(__hmul2(__hsub2(h2_of(p13), k1056), s));
        }
    }
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

// one lane polls, the warp follows: 700 threads spinning on mbarriers starve the one thread that issues the MMAs
__device__ __forceinline__ void bar_wait_warp(uint32_t bar, uint32_t parity, int lane) {
    if (lane == 0) bar_wait(bar, parity);
    __syncwarp();
}

// one lane of a converged warp (the warp runs the surrounding loop as a whole: its variables stay warp-uniform, and the
// tcgen05 instructions get their operands from uniform registers without a per-instruction broadcast loop)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(pred));
    return pred != 0;
}

struct TcqMaps {
    CUtensorMap w[TCQ_MAX_SEGS];
    CUtensorMap x;
};

template <int BN>
__global__ void __launch_bounds__(TQ_THREADS, 1) tcq_gemm_kernel(const __grid_constant__ TcqMaps maps, const TcqParams p) {
    using G = TqGeo<BN>;
    extern __shared__ uint8_t tq_smem[];
    if (p.trace && blockIdx.x == 0 && threadIdx.x == 0) p.trace[8000] = clock64();
    const uint32_t x0 = (smem_u32(tq_smem) + 1023u) & ~1023u;
    const uint32_t raw0 = x0 + G::NX * G::X_BYTES;
    const uint32_t bars = raw0 + G::RSTAGES * TQ_RAW_BYTES;
    auto a_full = [&](int s) { return bars + 8u * s; };
    auto a_empty = [&](int s) { return bars + 8u * (TQ_NA + s); };
    auto x_full = [&](int s) { return bars + 8u * (2 * TQ_NA + s); };
    auto x_empty = [&](int s) { return bars + 8u * (2 * TQ_NA + G::NX + s); };
    constexpr int RB = 2 * TQ_NA + 2 * G::NX;
    auto raw_full = [&](int s) { return bars + 8u * (RB + s); };
    auto raw_empty = [&](int s) { return bars + 8u * (RB + G::RSTAGES + s); };
    auto acc_full = [&](int i) { return bars + 8u * (RB + 2 * G::RSTAGES + i); };
    auto acc_empty = [&](int i) { return bars + 8u * (RB + 2 * G::RSTAGES + 2 + i); };
    const uint32_t tmem_slot = bars + 8u * G::N_BARS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < TQ_NA; s++) {
            bar_init(a_full(s), TQ_DQ_WARPS);   // every dequantising warp has stored its part of the block column
            bar_init(a_empty(s), 1);            // tcgen05.commit
        }
        for (int s = 0; s < G::NX; s++) {
            bar_init(x_full(s), 1);
            bar_init(x_empty(s), 1);
        }
        for (int s = 0; s < G::RSTAGES; s++) {
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
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(tmem_slot), "n"(TQ_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    if (p.trace && blockIdx.x == 0 && threadIdx.x == 0) p.trace[8001] = clock64();
    Walk walk(p);
    int tile, kb0, kb1;
    if (warp == 0) {
        // ===== raw weight blocks =====
        if (lane == 0) {
            int rs = 0;
            uint32_t ph = 0;
            while (walk.next(tile, kb0, kb1)) {
                const int si = seg_of_tile(p, tile);
                const int q4 = p.seg_type[si] == B2_TYPE_Q4_K;
                const int bb = q4 ? 144 : 210;
                const uint32_t box = (uint32_t)TQ_BM * (q4 ? 144u : 224u);
                const int m0 = (tile - p.seg_tile0[si]) * TQ_BM;
                for (int kb = kb0; kb < kb1; kb++) {
                    bar_wait(raw_empty(rs), ph ^ 1);
                    bar_expect_tx(raw_full(rs), box);
                    tma_load_2d(raw0 + rs * TQ_RAW_BYTES, &maps.w[si], raw_full(rs), (kb * bb) & ~15, m0);   // a box starts on 16 bytes
                    if (++rs == G::RSTAGES) { rs = 0; ph ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 2) {
        // ===== activations: the block column's BN x 256 values as four swizzled BN x 64 tiles =====
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            while (walk.next(tile, kb0, kb1)) {
                for (int kb = kb0; kb < kb1; kb++) {
                    bar_wait(x_empty(s), ph ^ 1);
                    bar_expect_tx(x_full(s), G::X_BYTES);
#pragma unroll
                    for (int g = 0; g < 4; g++) tma_load_2d(x0 + s * G::X_BYTES + g * G::XT_BYTES, &maps.x, x_full(s), (4 * kb + g) * TQ_BK, 0);
                    if (++s == G::NX) { s = 0; ph ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== tensor cores: D[128 rows][BN] += A (tensor memory) . B (activation tile), 16 MMAs per block column =====
        {
            int xs = 0, j = 0, nb = 0, tn = 0;
            uint32_t xph = 0;
            for (; walk.next(tile, kb0, kb1); j++) {
                const int as = j & 1;
                bar_wait(acc_empty(as), ((j >> 1) & 1) ^ 1);   // the epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(as * 128);
                for (int kb = kb0; kb < kb1; kb++, nb++) {
                    const int s = nb & 1;
                    bar_wait(x_full(xs), xph);
                    bar_wait(a_full(s), (nb >> 1) & 1);
                    tc_fence_after();
                    const uint32_t ta = tmem_base + (uint32_t)(TQ_A_COL0 + 128 * s);
                    const uint64_t db = umma_desc_sw128(x0 + xs * G::X_BYTES);
                    if (elect_one()) {
                        if (p.trace && blockIdx.x == 0 && tn < 2000) p.trace[2 * tn] = clock64();
                        if (!(p.dbg & 1)) {
#pragma unroll
                            for (int k = 0; k < 16; k++)
                                umma_f16_ta(tmem_d + (uint32_t)((k & (G::NACC - 1)) * BN), ta + 8u * k,
                                            db + (uint64_t)((k >> 2) * (G::XT_BYTES >> 4) + 2 * (k & 3)), G::IDESC, (kb > kb0) | (k >= G::NACC));
                        }
                        const long long tm = (p.trace && blockIdx.x == 0) ? clock64() : 0;
                        umma_commit(a_empty(s));
                        umma_commit(x_empty(xs));
                        if (p.trace && blockIdx.x == 0 && tn < 2000) p.trace[2 * tn + 1] = clock64() | ((clock64() - tm) << 48);
                    }
                    tn++;
                    __syncwarp();
                    if (++xs == G::NX) { xs = 0; xph ^= 1; }
                }
                if (elect_one()) umma_commit(acc_full(as));
                __syncwarp();
            }
        }
    } else if (warp >= TQ_DQ_WARP0) {
        // ===== dequantisation: thread = weight row 32 (warp % 4) + lane, K step g of every block =====
        const int g = (warp - TQ_DQ_WARP0) >> 2;
        const int row = 32 * (warp & 3) + lane;
        const uint32_t ta = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(TQ_A_COL0 + 32 * g);
        int tb = 0;
        int rs = 0, nb = 0;
        uint32_t rph = 0;
        while (walk.next(tile, kb0, kb1)) {
            const int q4 = p.seg_type[seg_of_tile(p, tile)] == B2_TYPE_Q4_K;
            const uint32_t pitch = q4 ? 144u : 224u;
            for (int kb = kb0; kb < kb1; kb++, nb++) {
                const bool tr = p.trace && blockIdx.x == 0 && warp == TQ_DQ_WARP0 && lane == 0 && tb < 500;
                if (tr) p.trace[4096 + 5 * tb] = clock64();
                bar_wait_warp(raw_full(rs), rph, lane);
                if (tr) p.trace[4096 + 5 * tb + 1] = clock64();
                const uint32_t blk = raw0 + rs * TQ_RAW_BYTES + row * pitch;
                uint32_t o[32];
                if (p.dbg & 2) {
#pragma unroll
                    for (int i = 0; i < 32; i++) o[i] = 0u;
                } else if (p.fast) {
                    if (q4) dequant_q4k_fast(blk, g, o);
                    else dequant_q6k_fast(blk + (uint32_t)((kb * 210) & 15), g, o);
                } else if (q4) dequant_q4k(blk, g, o);
                else dequant_q6k(blk + (uint32_t)((kb * 210) & 15), g, o);
                __syncwarp();
                if (lane == 0) bar_arrive(raw_empty(rs));   // the stage's bytes are in registers
                if (++rs == G::RSTAGES) { rs = 0; rph ^= 1; }
                const int s = nb & 1;
                if (tr) p.trace[4096 + 5 * tb + 2] = clock64() + (o[0] == 0x12345678u);
                bar_wait_warp(a_empty(s), ((nb >> 1) & 1) ^ 1, lane);
                tc_fence_after();
                if (tr) p.trace[4096 + 5 * tb + 3] = clock64();
                if (!(p.dbg & 4)) tmem_st_32x32(ta + 128u * s, o);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) bar_arrive(a_full(s));
                if (tr) p.trace[4096 + 5 * tb++ + 4] = clock64();
            }
        }
    } else {
        // ===== epilogue (warps 3..6): partial sums of the segment to its slice; the tile's last segment clears the rest =====
        const int q = warp & 3;   // the quarter of the 128 tensor-memory lanes this warp may read
        const long long total = (long long)p.tiles_m * p.nblk;
        const size_t slice = (size_t)p.T * p.N;
        for (int j = 0; walk.next(tile, kb0, kb1); j++) {
            const int as = j & 1;
            const int si = seg_of_tile(p, tile);
            const int r = (tile - p.seg_tile0[si]) * TQ_BM + q * 32 + lane;   // row inside the segment
            const bool row_ok = r < p.seg_rows[si];
            const int n = p.seg_row0[si] + r;
            const int sl = (int)blockIdx.x - cta_of_step((long long)tile * p.nblk, total, (int)gridDim.x);
            bar_wait_warp(acc_full(as), (j >> 1) & 1, lane);
            tc_fence_after();
            float* const o = p.ws + (size_t)sl * slice + n;
#pragma unroll 1
            for (int c = 0; c < BN / 16; c++) {
                if (c * 16 >= p.T) break;
                uint32_t v[16];
                tmem_ld_32x16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * 128 + c * 16), v);
                float f[16];
#pragma unroll
                for (int i = 0; i < 16; i++) f[i] = __uint_as_float(v[i]);
#pragma unroll
                for (int a = 1; a < G::NACC; a++) {
                    tmem_ld_32x16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * 128 + a * BN + c * 16), v);
#pragma unroll
                    for (int i = 0; i < 16; i++) f[i] += __uint_as_float(v[i]);
                }
                if (row_ok) {
#pragma unroll
                    for (int i = 0; i < 16; i++)
                        if (c * 16 + i < p.T) o[(size_t)(c * 16 + i) * p.N] = f[i];
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
    if (p.trace && blockIdx.x == 0 && threadIdx.x == 96) p.trace[8003] = clock64();   // an epilogue warp is through
    tc_fence_before();
    __syncthreads();
    if (p.trace && blockIdx.x == 0 && threadIdx.x == 0) p.trace[8002] = clock64();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(TQ_TMEM_COLS) : "memory");
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

}  // namespace

// CTAs for `tiles` tiles of `nblk` block columns: one per SM, fewer when a tile would otherwise be cut into more than
// TQ_MAX_SLICES partial sums (skinny matrices with a long K: the consumer reads every slice)
constexpr int TQ_MAX_SLICES = 8;
static int tq_grid_for(int tiles, int nblk, int n_sms) {
    const long long total = (long long)tiles * nblk;
    int grid = (int)std::min<long long>(n_sms, total);
    while (grid > 1 && tq_slices(tiles, nblk, grid) > TQ_MAX_SLICES) grid = grid * 7 / 8;
    return grid;
}
static int tq_grid(int tiles, int nblk) { return tq_grid_for(tiles, nblk, tq_sm_count()); }

// The stream-K plan of a launch as host arithmetic (tests/test_tcq_plan.py checks it without a GPU): tiles of 128 rows per
// segment, K / 256 block columns each, shared out over `grid` CTAs; *slices_out = the most CTAs any tile meets.
int tcq_plan(const int* seg_rows, int n_seg, int K, int n_sms, int* tiles_out, int* grid_out, int* slices_out) {
    if (!seg_rows || n_seg < 1 || n_seg > TCQ_MAX_SEGS || K <= 0 || K % 256 || n_sms < 1) return B2_EINVAL;
    int tiles = 0;
    for (int i = 0; i < n_seg; i++) {
        if (seg_rows[i] < 1) return B2_EINVAL;
        tiles += (seg_rows[i] + TQ_BM - 1) / TQ_BM;
    }
    const int nblk = K / 256, grid = tq_grid_for(tiles, nblk, n_sms);
    if (tiles_out) *tiles_out = tiles;
    if (grid_out) *grid_out = grid;
    if (slices_out) *slices_out = tq_slices(tiles, nblk, grid);
    return 0;
}
// the share of CTA c: steps [lo, hi) of the tile-major sequence (step = tile * nblk + block column), as Walk computes it
void tcq_share(int tiles, int nblk, int grid, int c, long long* lo, long long* hi) {
    const long long total = (long long)tiles * nblk;
    *lo = total * c / grid;
    *hi = total * (c + 1) / grid;
}
int tcq_first_cta_of_tile(int tiles, int nblk, int grid, int tile) { return cta_of_step((long long)tile * nblk, (long long)tiles * nblk, grid); }

bool tcq_fast_conversion() { return getenv("B2_TCQ_EXACT") == nullptr; }

// (a tensor map's row stride is a multiple of 16 bytes: rows of 210-byte Q6_K blocks qualify when K is a multiple of 2048)
bool tcq_supported(int type, int K) {
    if (K <= 0 || K % 256) return false;
    return type == B2_TYPE_Q4_K || (type == B2_TYPE_Q6_K && (K / 256) % 8 == 0);
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
    p.fast = a.x_perm;
    if (const char* e = getenv("B2_TCQ_DBG")) p.dbg = atoi(e);
    *parts = TcParts{a.ws, p.splits, (size_t)a.T * p.N};
    static const bool want_trace = getenv("B2_TCQ_TRACE") != nullptr;
    if (want_trace) {
        static long long* tbuf = nullptr;
        if (!tbuf) cudaMalloc(&tbuf, 8192 * sizeof(long long));
        cudaMemsetAsync(tbuf, 0, 8192 * sizeof(long long), st);
        p.trace = tbuf;
        rc = bn == 32 ? launch_tq<32>(maps, p, grid, st) : bn == 64 ? launch_tq<64>(maps, p, grid, st) : launch_tq<128>(maps, p, grid, st);
        cudaStreamSynchronize(st);
        std::vector<long long> h(8192);
        cudaMemcpy(h.data(), tbuf, h.size() * 8, cudaMemcpyDeviceToHost);
        fprintf(stderr, "tcq trace: N=%d K=%d T=%d grid=%d slices=%d\n", p.N, p.K, p.T, grid, p.splits);
        const long long t0 = h[4096];
        fprintf(stderr, "  kernel entry %lld, set-up done %lld, epilogue warp done %lld, all warps done %lld (cycles relative to the first dequantising wait)\n",
                h[8000] - t0, h[8001] - t0, h[8003] - t0, h[8002] - t0);
        for (int b = 0; b < 40 && h[4096 + 5 * b]; b++)
            fprintf(stderr, "  dq block %2d: start %7lld raw_full +%5lld dequant +%5lld a_empty +%5lld st+arrive +%5lld\n", b, h[4096 + 5 * b] - t0,
                    h[4096 + 5 * b + 1] - h[4096 + 5 * b], h[4096 + 5 * b + 2] - h[4096 + 5 * b + 1], h[4096 + 5 * b + 3] - h[4096 + 5 * b + 2],
                    h[4096 + 5 * b + 4] - h[4096 + 5 * b + 3]);
        for (int k = 0; k < 40 && h[2 * k]; k++)
            fprintf(stderr, "  mma block %2d: ready at %7lld, issue+commit %4lld (the two commits %lld)\n", k, h[2 * k] - t0,
                    (h[2 * k + 1] & 0xffffffffffffll) - (h[2 * k] & 0xffffffffffffll), (long long)((unsigned long long)h[2 * k + 1] >> 48));
        return rc;
    }
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
