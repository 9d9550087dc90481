// step.cu -- one persistent kernel per decode step (SURVEY 8a rows a10-a16 in a single launch).
//
// Why: with one launch per stage a Llama-3.1-8B token is 163 launches; measured on the B200 each costs
// 8-9 us of fixed latency (launch ramp, activation prologue, first DRAM round trip, drain) against
// 2-10 us of HBM time, and two launches cannot be co-resident in shared memory, so weight prefetch
// under programmatic dependent launch never overlapped.  Here the same warp-specialised CTAs
// (gemv_device.cuh) stay resident for the whole token and walk a program of stages:
//   * the producer and header-decoder warps depend only on the (static) weights, so they run ahead
//     across stage boundaries: while the compute warps sit in a grid barrier or build the next
//     activation vector, the ring keeps filling with the next stage's weights -- HBM never idles for a
//     launch;
//   * a stage ends with one atomic arrival per CTA on a global counter (after the CTA's epilogue
//     stores, behind a __threadfence); the next stage's consumers poll it with ld.acquire.gpu;
//   * embedding lookup, attention and greedy sampling run as stages on the compute warps.
// All CTAs must be co-resident (they wait on each other): the grid is sized by
// cudaOccupancyMaxActiveClusters, and every wait is bounded and traps instead of hanging.
#include "step.cuh"

#include <cstdlib>

#include "dequant.cuh"
#include "gemv_device.cuh"

namespace b2 {

using namespace gemvdev;

namespace {

constexpr int W_PRODUCER = NCW, W_DECODER = NCW + 1, W_EPILOGUE = NCW + 2;

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// wait until `target` CTAs have arrived (bounded: a CTA that is not resident must not hang the GPU)
__device__ __forceinline__ void grid_wait(const unsigned int* ctr, unsigned int target) {
#pragma unroll 1
    for (uint32_t i = 0; i < (1u << 24); i++) {
        if (ld_acquire_gpu(ctr) >= target) return;
        __nanosleep(32);
    }
    asm volatile("trap;");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// profiling stamp of CTA 0 (no-op unless the program carries a trace buffer)
__device__ __forceinline__ void stamp(const StepProgram& P, int oi, int slot) {
    if (P.trace && blockIdx.x == 0) P.trace[oi * 4 + slot] = globaltimer_ns();
}

// ---- attention work item: (kv head, chunk of STEP_ATTN_CHUNK tokens), all query heads of the group ------------
// Same numerics as attn.cu: q and the probabilities are rounded to fp16, fp32 accumulation; a single
// chunk normalises before rounding (exactly the oracle's order), several chunks merge by log-sum-exp.
constexpr int A_MAXG = 4;   // query heads processed together (register budget); larger GQA groups run as two batches
struct AttnSmem {
    float* scores;   // [A_MAXG][STEP_ATTN_CHUNK]
    float* red;      // [A_MAXG][NCW]
    float* out;      // [A_MAXG][128]
    float* ml;       // [2][A_MAXG]
    int* flag;
};
constexpr size_t ATTN_SMEM_BYTES = (A_MAXG * STEP_ATTN_CHUNK + A_MAXG * NCW + A_MAXG * 128 + 2 * A_MAXG + 4) * sizeof(float);

// heads [h0, h0 + hb) of the kv head's group (hb <= A_MAXG: the register arrays below are per head)
template <int HD>
__device__ void attn_item(const StepOp& op, int kvh, int h0, int hb, int chunk, int n_chunks, int T, const AttnSmem& sm, int tid) {
    constexpr int LPT = HD / 8, TPW = 32 / LPT;
    const int gq = op.a_n_head / op.a_n_kv;
    const int g = hb;
    const int warp = tid >> 5, lane = tid & 31, sub = lane / LPT, ln = lane % LPT;
    const int t0 = chunk * STEP_ATTN_CHUNK, n = min(T - t0, STEP_ATTN_CHUNK);
    const __half* kbase = op.a_k + ((size_t)kvh * op.a_n_ctx + t0) * HD;
    const __half* vbase = op.a_v + ((size_t)kvh * op.a_n_ctx + t0) * HD;
    const float scale = 1.0f / sqrtf((float)HD);
    float qf[A_MAXG][8];
#pragma unroll
    for (int h = 0; h < A_MAXG; h++)
        if (h < g) {
            const float* qp = op.a_q + ((size_t)(kvh * gq + h0 + h)) * HD + ln * 8;
#pragma unroll
            for (int i = 0; i < 8; i++) qf[h][i] = __half2float(__float2half_rn(qp[i]));
        }
    for (int tb = warp * TPW; tb < n; tb += NCW * TPW) {
        const int t = tb + sub;
        float d[A_MAXG];
#pragma unroll
        for (int h = 0; h < A_MAXG; h++) d[h] = 0.0f;
        if (t < n) {
            const uint4 kv = *(const uint4*)(kbase + (size_t)t * HD + ln * 8);
            const __half2* k2 = (const __half2*)&kv;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float2 kf = __half22float2(k2[i]);
#pragma unroll
                for (int h = 0; h < A_MAXG; h++)
                    if (h < g) d[h] += kf.x * qf[h][2 * i] + kf.y * qf[h][2 * i + 1];
            }
        }
#pragma unroll
        for (int h = 0; h < A_MAXG; h++)
            if (h < g) {
#pragma unroll
                for (int o = LPT / 2; o > 0; o >>= 1) d[h] += __shfl_xor_sync(0xffffffffu, d[h], o);
                if (ln == 0 && t < n) sm.scores[h * STEP_ATTN_CHUNK + t] = d[h] * scale;
            }
    }
    named_bar_sync(1, NCT);
    float pmax[A_MAXG], psum[A_MAXG];
#pragma unroll
    for (int h = 0; h < A_MAXG; h++)
        if (h < g) {
            float m = -INFINITY;
            for (int t = tid; t < n; t += NCT) m = fmaxf(m, sm.scores[h * STEP_ATTN_CHUNK + t]);
            m = warp_max(m);
            if (lane == 0) sm.red[h * NCW + warp] = m;
        }
    named_bar_sync(1, NCT);
#pragma unroll
    for (int h = 0; h < A_MAXG; h++)
        if (h < g) {
            float m = sm.red[h * NCW];
#pragma unroll
            for (int w = 1; w < NCW; w++) m = fmaxf(m, sm.red[h * NCW + w]);
            pmax[h] = m;
        }
    named_bar_sync(1, NCT);
#pragma unroll
    for (int h = 0; h < A_MAXG; h++)
        if (h < g) {
            float s = 0.0f;
            for (int t = tid; t < n; t += NCT) {
                const float e = expf(sm.scores[h * STEP_ATTN_CHUNK + t] - pmax[h]);
                sm.scores[h * STEP_ATTN_CHUNK + t] = e;
                s += e;
            }
            s = warp_sum(s);
            if (lane == 0) sm.red[h * NCW + warp] = s;
        }
    named_bar_sync(1, NCT);
#pragma unroll
    for (int h = 0; h < A_MAXG; h++)
        if (h < g) {
            float s = 0.0f;
#pragma unroll
            for (int w = 0; w < NCW; w++) s += sm.red[h * NCW + w];
            psum[h] = s;
        }
    const bool exact = n_chunks == 1;
    float acc[A_MAXG][8];
#pragma unroll
    for (int h = 0; h < A_MAXG; h++)
#pragma unroll
        for (int i = 0; i < 8; i++) acc[h][i] = 0.0f;
    for (int tb = warp * TPW; tb < n; tb += NCW * TPW) {
        const int t = tb + sub;
        if (t < n) {
            const uint4 vv = *(const uint4*)(vbase + (size_t)t * HD + ln * 8);
            const __half2* v2 = (const __half2*)&vv;
            float vf[8];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float2 f = __half22float2(v2[i]);
                vf[2 * i] = f.x;
                vf[2 * i + 1] = f.y;
            }
#pragma unroll
            for (int h = 0; h < A_MAXG; h++)
                if (h < g) {
                    const float e = sm.scores[h * STEP_ATTN_CHUNK + t];
                    const float pr = __half2float(__float2half_rn(exact ? e * (1.0f / psum[h]) : e));
#pragma unroll
                    for (int i = 0; i < 8; i++) acc[h][i] += pr * vf[i];
                }
        }
    }
#pragma unroll
    for (int h = 0; h < A_MAXG; h++)
        if (h < g) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                float v = acc[h][i];
#pragma unroll
                for (int o = LPT; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                acc[h][i] = v;
            }
        }
    for (int i = tid; i < A_MAXG * 128; i += NCT) sm.out[i] = 0.0f;
    named_bar_sync(1, NCT);
    for (int w = 0; w < NCW; w++) {
        if (warp == w && sub == 0) {
#pragma unroll
            for (int h = 0; h < A_MAXG; h++)
                if (h < g) {
#pragma unroll
                    for (int i = 0; i < 8; i++) sm.out[h * 128 + ln * 8 + i] += acc[h][i];
                }
        }
        named_bar_sync(1, NCT);
    }
    if (exact) {
        for (int i = tid; i < g * HD; i += NCT) {
            const int h = i / HD, dd = i - h * HD;
            op.a_out[((size_t)(kvh * gq + h0 + h)) * HD + dd] = sm.out[h * 128 + dd];
        }
        named_bar_sync(1, NCT);
        return;
    }
    // partial record {max, sum, out[HD]} of this chunk; the last chunk of the head to arrive merges
    const size_t stride = (size_t)HD + 2;
    for (int i = tid; i < g * HD; i += NCT) {
        const int h = i / HD, dd = i - h * HD;
        op.a_partials[((size_t)(kvh * gq + h0 + h) * op.a_max_chunks + chunk) * stride + 2 + dd] = sm.out[h * 128 + dd];
    }
    if (tid == 0) {
#pragma unroll
        for (int h = 0; h < A_MAXG; h++)
            if (h < g) {
                float* rec = op.a_partials + ((size_t)(kvh * gq + h0 + h) * op.a_max_chunks + chunk) * stride;
                rec[0] = pmax[h];
                rec[1] = psum[h];
            }
    }
    __threadfence();
    named_bar_sync(1, NCT);
    if (tid == 0) {
        unsigned int* ctr = &op.a_counters[kvh * 2 + h0 / A_MAXG];   // one arrival counter per (kv head, head batch)
        const unsigned int arrived = atomicAdd(ctr, 1u) + 1u;
        *sm.flag = arrived == (unsigned int)n_chunks;
        if (*sm.flag) *ctr = 0u;
    }
    named_bar_sync(1, NCT);
    if (*sm.flag) {
        __threadfence();
        for (int i = tid; i < g * HD; i += NCT) {
            const int h = i / HD, dd = i - h * HD;
            const float* base = op.a_partials + ((size_t)(kvh * gq + h0 + h) * op.a_max_chunks) * stride;
            float m = -INFINITY;
            for (int s = 0; s < n_chunks; s++) m = fmaxf(m, __ldcg(base + s * stride));
            float den = 0.0f, num = 0.0f;
            for (int s = 0; s < n_chunks; s++) {
                const float w = expf(__ldcg(base + s * stride) - m);
                den += __ldcg(base + s * stride + 1) * w;
                num += __ldcg(base + s * stride + 2 + dd) * w;
            }
            op.a_out[((size_t)(kvh * gq + h0 + h)) * HD + dd] = num / den;
        }
    }
    named_bar_sync(1, NCT);
}

// copy a stage's mat-vec descriptor from the program (global memory) into shared memory with one warp
__device__ __forceinline__ void stage_descriptor(GemvArgs* dst, const GemvArgs* src, int lane) {
    static_assert(sizeof(GemvArgs) % 4 == 0, "descriptor copy is word-wise");
    const uint32_t* s = (const uint32_t*)src;
    uint32_t* d = (uint32_t*)dst;
    for (int i = lane; i < (int)(sizeof(GemvArgs) / 4); i += 32) d[i] = s[i];
    __syncwarp();
}

template <int ACT>
__global__ void __launch_bounds__(GEMV_THREADS, 2) step_kernel(const __grid_constant__ StepProgram P) {
    using G = Geo<ACT>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem;
    uint8_t* dec = smem + GEMV_STAGES * G::STAGE_BYTES;
    uint64_t* full = (uint64_t*)(dec + GEMV_STAGES * G::DEC_BYTES);
    uint64_t* ready = full + GEMV_STAGES;
    uint64_t* empty = ready + GEMV_STAGES;
    uint64_t* part_full = empty + GEMV_STAGES;
    uint64_t* part_free = part_full + 2;
    ActSmem sm;
    sm.xbar = part_free + 2;
    sm.ssbar = sm.xbar + 1;
    sm.ss_span = (double*)(sm.ssbar + 2);
    float* red = (float*)(sm.ss_span + MAX_SPANS);
    // the current stage's mat-vec descriptor, one private copy per role (the roles are at different stages)
    GemvArgs* slots = (GemvArgs*)(red + 2 * NCW * TILE_ROWS);
    uint8_t* xbase = (uint8_t*)(slots + 4);   // activation buffers, re-carved per mat-vec; attention aliases them

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_ctarank(), csize = cluster_nctarank();
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < GEMV_STAGES; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&ready[s], 1);
            mbar_init(&empty[s], NCW);
        }
        mbar_init(&part_full[0], NCW);
        mbar_init(&part_full[1], NCW);
        mbar_init(&part_free[0], 1);
        mbar_init(&part_free[1], 1);
        mbar_init(sm.xbar, csize * NCW);
        mbar_init(sm.ssbar, csize * NCW);
        mbar_fence_init();
    }
    cluster_sync_all();
    const unsigned int G_CTAS = gridDim.x;

    if (warp == W_PRODUCER) {
        // ===== producer: streams every mat-vec stage's weights in program order, never waits for a stage to start =====
        int stage = 0, phase = 0;
        for (int oi = 0; oi < P.n_ops; oi++) {
            const StepOp& op = P.ops[oi];
            if (op.kind != OP_GEMV) continue;
            __syncwarp();
            stage_descriptor(&slots[0], &op.g, lane);
            const GemvArgs& a = slots[0];
            const int nblk_row = a.K / (ACT == ACT_Q8K ? 256 : 32);
            const int n_chunks = (nblk_row + G::CHUNK_BLOCKS - 1) / G::CHUNK_BLOCKS;
            int si = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
                while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
                const GemvSeg& s = a.seg[si];
                const int bb = block_bytes_of(s.type);
                const uint8_t* rowp = lane < TILE_ROWS ? slot_row_ptr(s, tile - s.tile0, lane, nblk_row * bb) : nullptr;
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
        }
        return;
    }

    if (warp == W_DECODER) {
        int stage = 0, phase = 0;
        for (int oi = 0; oi < P.n_ops; oi++) {
            const StepOp& op = P.ops[oi];
            if (op.kind != OP_GEMV) continue;
            __syncwarp();
            stage_descriptor(&slots[1], &op.g, lane);
            const GemvArgs& a = slots[1];
            const int nblk_row = a.K / (ACT == ACT_Q8K ? 256 : 32);
            const int n_chunks = (nblk_row + G::CHUNK_BLOCKS - 1) / G::CHUNK_BLOCKS;
            int si = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
                while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
                const int type = a.seg[si].type;
                const int bb = block_bytes_of(type);
                for (int c = 0; c < n_chunks; c++) {
                    const int nb = min(G::CHUNK_BLOCKS, nblk_row - c * G::CHUNK_BLOCKS);
                    mbar_wait(&full[stage], phase);
                    if (ACT == ACT_Q8K && (type == B2_TYPE_Q4_K || type == B2_TYPE_Q5_K)) {
                        for (int i = lane; i < TILE_ROWS * nb; i += 32) {
                            const int rr = i / nb, b = i - rr * nb;
                            decode_k4_header(ring + stage * G::STAGE_BYTES + rr * G::ROW_SLOT + b * bb,
                                             dec + stage * G::DEC_BYTES + rr * G::DEC_ROW + b * 32);
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&ready[stage]);
                    if (++stage == GEMV_STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        return;
    }

    if (warp == W_EPILOGUE) {
        // ===== epilogue: finish row pairs; after the CTA's last tile of a stage, arrive on the grid barrier =====
        int parity = 0, ph = 0;
        for (int oi = 0; oi < P.n_ops; oi++) {
            const StepOp& op = P.ops[oi];
            if (op.kind != OP_GEMV) continue;   // the compute warps signal those stages
            // never arrive for a stage before the previous one is complete everywhere: a CTA without tiles in
            // this stage would otherwise run ahead and release other CTAs early
            if (oi > 0) {
                if (lane == 0) grid_wait(P.sync_counter, (unsigned int)oi * G_CTAS);
                __syncwarp();
            }
            stage_descriptor(&slots[2], &op.g, lane);
            const GemvArgs& a = slots[2];
            int si = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, parity ^= 1) {
                while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
                const GemvSeg& s = a.seg[si];
                mbar_wait(&part_full[parity], (ph >> parity) & 1);
                const float* rp = red + parity * NCW * TILE_ROWS;
                float v = 0.0f;
                if (lane < TILE_ROWS) {
#pragma unroll
                    for (int w = 0; w < NCW; w++) v += rp[w * TILE_ROWS + lane];
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&part_free[parity]);
                ph ^= 1 << parity;
                const float vo = __shfl_xor_sync(0xffffffffu, v, 1);
                if (lane < TILE_ROWS && (lane & 1) == 0) {
                    const int p = (tile - s.tile0) * GEMV_TILE_PAIRS + (lane >> 1);
                    if (p < s.n_pairs) {
                        int ra, rb;
                        pair_rows(s, p, &ra, &rb);
                        epilogue(a, s, p, ra, rb, v, vo);
                    }
                }
            }
            // a fence orders only the calling thread's stores: every lane that wrote results fences, then one arrives
            __threadfence();
            __syncwarp();
            if (lane == 0) {
                atomicAdd(P.sync_counter, 1u);
                stamp(P, oi, 3);
            }
        }
        return;
    }

    // ===== compute warps =====
    const int r = lane >> 2;
    int stage = 0, phase = 0, parity = 0, fph = 3;
    uint32_t xpar = 0, sspar = 0;
    for (int oi = 0; oi < P.n_ops; oi++) {
        const StepOp& op = P.ops[oi];
        // every CTA has arrived oi times per completed stage: wait for the previous stage everywhere
        if (oi > 0) {
            if (warp == 0 && lane == 0) grid_wait(P.sync_counter, (unsigned int)oi * G_CTAS);
            named_bar_sync(1, NCT);
        }
        if (tid == 0) stamp(P, oi, 0);
        if (op.kind == OP_GEMV) {
            if (warp == 0) stage_descriptor(&slots[3], &op.g, lane);
            named_bar_sync(1, NCT);
            const GemvArgs& a = slots[3];
            const int K = a.K;
            sm.xq = (int8_t*)xbase;
            sm.xd = (float*)(sm.xq + K);
            sm.bs16 = (int16_t*)(sm.xd + (ACT == ACT_Q8K ? K / 256 : K / 32));
            sm.bs32 = sm.bs16 + K / 16;
            build_activation<ACT>(a, sm, smem, tid, rank, csize, xpar, sspar);
            if (tid == 0) stamp(P, oi, 1);
            xpar ^= 1;
            if (a.prologue == PRO_NORM) sspar ^= 1;
            const ActView xv{sm.xq, sm.xd, sm.bs16, sm.bs32};
            const int nblk_row = K / (ACT == ACT_Q8K ? 256 : 32);
            const int n_chunks = (nblk_row + G::CHUNK_BLOCKS - 1) / G::CHUNK_BLOCKS;
            int si = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, parity ^= 1) {
                while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
                const GemvSeg& s = a.seg[si];
                const int type = s.type;
                const int bb = block_bytes_of(type);
                uint32_t off0 = 0;
                if (ACT != ACT_Q8K || type == B2_TYPE_Q6_K) off0 = (uint32_t)((uintptr_t)slot_row_ptr(s, tile - s.tile0, r, nblk_row * bb) & 15);
                const uint32_t chunk_step = (uint32_t)(G::CHUNK_BLOCKS * bb) & 15;
                float acc = 0.0f;
                for (int c = 0; c < n_chunks; c++) {
                    const int nb = min(G::CHUNK_BLOCKS, nblk_row - c * G::CHUNK_BLOCKS);
                    mbar_wait(&ready[stage], phase);
                    const uint8_t* row = ring + stage * G::STAGE_BYTES + r * G::ROW_SLOT + ((off0 + c * chunk_step) & 15);
                    acc += dot_chunk<ACT>(type, row, dec + stage * G::DEC_BYTES + r * G::DEC_ROW, nb, c * G::CHUNK_BLOCKS, xv, warp, lane);
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&empty[stage]);
                    if (++stage == GEMV_STAGES) { stage = 0; phase ^= 1; }
                }
                acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                mbar_wait(&part_free[parity], (fph >> parity) & 1);
                fph ^= 1 << parity;
                if ((lane & 3) == 0) red[(parity * NCW + warp) * TILE_ROWS + r] = acc;
                __syncwarp();
                if (lane == 0) mbar_arrive(&part_full[parity]);
            }
            if (tid == 0) stamp(P, oi, 2);
            // the epilogue warp arrives on the grid barrier for this stage
        } else {
            if (op.kind == OP_EMBED) {
                int tok = *op.e_token;
                if (tok < 0 || tok >= op.e_vocab) tok = 0;
                const uint8_t* row = op.e_table + (size_t)tok * op.e_row_bytes;
                for (int g4 = blockIdx.x * NCT + tid; g4 < op.e_d_model / 4; g4 += gridDim.x * NCT) ((float4*)op.e_x)[g4] = dq4(op.e_type, row, g4);
            } else if (op.kind == OP_ATTN) {
                AttnSmem as;
                as.scores = (float*)xbase;
                as.red = as.scores + A_MAXG * STEP_ATTN_CHUNK;
                as.out = as.red + A_MAXG * NCW;
                as.ml = as.out + A_MAXG * 128;
                as.flag = (int*)(as.ml + 2 * A_MAXG);
                const int T = *op.a_pos + 1;
                int n_chunks = (T + STEP_ATTN_CHUNK - 1) / STEP_ATTN_CHUNK;
                if (n_chunks > op.a_max_chunks) n_chunks = op.a_max_chunks;   // cannot happen: T <= n_ctx
                const int gq = op.a_n_head / op.a_n_kv;
                const int n_batches = (gq + A_MAXG - 1) / A_MAXG;
                // work item = (kv head, batch of <= 4 query heads, chunk of 256 tokens), dealt round-robin over the grid
                for (int item = blockIdx.x; item < op.a_n_kv * n_batches * n_chunks; item += gridDim.x) {
                    const int chunk = item % n_chunks, hb_i = (item / n_chunks) % n_batches, kvh = item / (n_chunks * n_batches);
                    const int h0 = hb_i * A_MAXG, hb = min(A_MAXG, gq - h0);
                    if (op.a_hd == 64) attn_item<64>(op, kvh, h0, hb, chunk, n_chunks, T, as, tid);
                    else attn_item<128>(op, kvh, h0, hb, chunk, n_chunks, T, as, tid);
                }
            } else if (op.kind == OP_ARGMAX) {
                float bv = -INFINITY;
                int bi = 0x7fffffff;
                for (int64_t i = (int64_t)blockIdx.x * NCT + tid; i < op.s_n; i += (int64_t)gridDim.x * NCT) {
                    const float v = op.s_logits[i];
                    if (v > bv) { bv = v; bi = (int)i; }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                    const int oi2 = __shfl_xor_sync(0xffffffffu, bi, o);
                    if (ov > bv || (ov == bv && oi2 < bi)) { bv = ov; bi = oi2; }
                }
                float* sv = (float*)xbase;
                int* sidx = (int*)(sv + NCW);
                if (lane == 0) { sv[warp] = bv; sidx[warp] = bi; }
                named_bar_sync(1, NCT);
                if (tid == 0) {
                    for (int w = 1; w < NCW; w++)
                        if (sv[w] > bv || (sv[w] == bv && sidx[w] < bi)) { bv = sv[w]; bi = sidx[w]; }
                    op.s_val[blockIdx.x] = bv;
                    op.s_idx[blockIdx.x] = bi;
                    __threadfence();
                    const unsigned int arrived = atomicAdd(op.s_counter, 1u) + 1u;
                    if (arrived == gridDim.x) {
                        __threadfence();
                        for (unsigned int b = 0; b < gridDim.x; b++) {
                            const float ov = __ldcg(op.s_val + b);
                            const int oi2 = __ldcg(op.s_idx + b);
                            if (ov > bv || (ov == bv && oi2 < bi)) { bv = ov; bi = oi2; }
                        }
                        if (bi == 0x7fffffff) bi = 0;
                        *op.s_token = bi;
                        if (op.s_out) {
                            const int s = *op.s_step;
                            if (s < op.s_cap) op.s_out[s] = bi;
                            *op.s_step = s + 1;
                        }
                        if (op.s_pos) *op.s_pos = *op.s_pos + 1;
                        *op.s_counter = 0u;
                    }
                }
            }
            __threadfence();   // every thread's stores of this stage, before the CTA's single arrival
            named_bar_sync(1, NCT);
            if (tid == 0) {
                atomicAdd(P.sync_counter, 1u);
                stamp(P, oi, 3);
            }
        }
    }
}

size_t step_smem_bytes(int act, int max_k) {
    size_t n = gemv_smem_bytes(act, max_k) + 4 * sizeof(GemvArgs);
    // attention aliases the activation buffers; make sure they are large enough
    const size_t x_bytes = (size_t)max_k + (act == ACT_Q8K ? max_k / 256 : max_k / 32) * 4 + max_k / 16 * 2 + max_k / 32 * 2;
    if (x_bytes < ATTN_SMEM_BYTES) n += ATTN_SMEM_BYTES - x_bytes;
    return (n + 127) & ~(size_t)127;
}

int g_step_sms = 0;

int step_config(int act, int max_k, size_t* smem_out, int* cluster_out, int* grid_out) {
    if (!g_step_sms) {
        int dev = 0;
        B2_CUDA_CHECK(cudaGetDevice(&dev));
        B2_CUDA_CHECK(cudaDeviceGetAttribute(&g_step_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const size_t smem = step_smem_bytes(act, max_k);
    if (smem > 227 * 1024) {
        set_error("step kernel: K=%d needs %zu bytes of shared memory", max_k, smem);
        return B2_ERANGE;
    }
    auto kern = act == ACT_Q8K ? step_kernel<ACT_Q8K> : step_kernel<ACT_Q80>;
    B2_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int ctas_per_sm = (int)((227 * 1024) / (smem + 1024));
    if (ctas_per_sm > 2) ctas_per_sm = 2;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    const int cluster = 8;
    cudaLaunchConfig_t q{};
    q.gridDim = dim3(g_step_sms * ctas_per_sm / cluster * cluster);
    q.blockDim = dim3(GEMV_THREADS);
    q.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    q.attrs = attr;
    q.numAttrs = 1;
    int n = 0;
    B2_CUDA_CHECK(cudaOccupancyMaxActiveClusters(&n, kern, &q));
    if (n < 1) {
        set_error("step kernel: no cluster of %d CTAs fits on this device", cluster);
        return B2_ECUDA;
    }
    const int cap = g_step_sms * ctas_per_sm / cluster;
    if (n > cap) n = cap;
    *smem_out = smem;
    *cluster_out = cluster;
    *grid_out = n * cluster;
    return 0;
}

}  // namespace

int step_attn_max_chunks(int n_ctx) { return (n_ctx + STEP_ATTN_CHUNK - 1) / STEP_ATTN_CHUNK; }
size_t step_attn_partials_bytes(int n_head, int head_dim, int n_ctx) {
    return (size_t)n_head * step_attn_max_chunks(n_ctx) * ((size_t)head_dim + 2) * sizeof(float);
}

int step_grid_size(int act, int max_k) {
    size_t smem;
    int cluster, grid;
    const int rc = step_config(act, max_k, &smem, &cluster, &grid);
    return rc ? rc : grid;
}

int launch_step(const StepProgram& p, int act, cudaStream_t st) {
    size_t smem;
    int cluster, grid;
    int rc = step_config(act, p.max_k, &smem, &cluster, &grid);
    if (rc) return rc;
    B2_CUDA_CHECK(cudaMemsetAsync(p.sync_counter, 0, sizeof(unsigned int), st));
    auto kern = act == ACT_Q8K ? step_kernel<ACT_Q8K> : step_kernel<ACT_Q80>;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(GEMV_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B2_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, p));
    return 0;
}

}  // namespace b2
