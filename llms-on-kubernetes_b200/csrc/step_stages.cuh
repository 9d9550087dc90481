// step_stages.cuh -- what the two persistent decode kernels (step.cu, step2.cu) share: the grid barrier, the
// descriptor staging helpers and the stages that are not mat-vecs (embedding row, attention, greedy sampling).
#pragma once
#include "dequant.cuh"
#include "gemv_device.cuh"
#include "step.cuh"
#include "step_attn.cuh"

namespace b2 {
namespace stepdev {

using namespace gemvdev;
using namespace stepattn;

constexpr int W_PRODUCER = NCW, W_DECODER = NCW + 1, W_EPILOGUE = NCW + 2;

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// ---- grid barrier -------------------------------------------------------------------------------------
// Arrivals are counted in word 0 of `sync` (stage number s is complete when it reaches s * grid).  One thread
// per CTA polls it and hands the result to the CTA's other waiting warp through shared memory.  (Tried and
// measured slower by ~1.2 us per stage: the completing CTA publishing per-group flags on separate L2 lines --
// the extra hop costs more than the polling traffic on the counter's line.)  All waits are bounded: a CTA
// that is not resident must not hang the GPU.
// one thread per CTA, after the CTA's stores of the stage are fenced
__device__ __forceinline__ void grid_arrive(unsigned int* sync, unsigned int stage_no) {
    (void)stage_no;
    atomicAdd(sync, 1u);
}
// wait until stage `stage_no` is complete everywhere, then tell the CTA's other waiters
__device__ __forceinline__ void grid_wait(const unsigned int* sync, unsigned int stage_no, volatile unsigned int* cta_stage) {
    const unsigned int target = stage_no * gridDim.x;
    // relaxed probes (no L1 invalidation each time), one acquire once the count is there
#pragma unroll 1
    for (uint32_t i = 0; i < (1u << 24); i++) {
        unsigned int v;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(sync) : "memory");
        if (v >= target) {
            (void)ld_acquire_gpu(sync);
            __threadfence_block();
            *cta_stage = stage_no;
            return;
        }
    }
    asm volatile("trap;");
}
// the CTA's other waiter (epilogue warp): the polling thread's acquire already refreshed this SM's L1
__device__ __forceinline__ void grid_wait_local(volatile unsigned int* cta_stage, unsigned int stage_no) {
#pragma unroll 1
    for (uint32_t i = 0; i < (1u << 26); i++) {
        if (*cta_stage >= stage_no) {
            __threadfence_block();
            return;
        }
        __nanosleep(20);
    }
    asm volatile("trap;");
}
// profiling stamp of CTA 0 (no-op unless the program carries a trace buffer)
__device__ __forceinline__ void stamp(const StepProgram& P, int oi, int slot) {
    if (P.trace && blockIdx.x == 0) P.trace[oi * STEP_TRACE_SLOTS + slot] = trace_ns();
}

// copy `words` 32-bit words of static program data from global into shared memory with one warp
__device__ __forceinline__ void stage_words(void* dst, const void* src, int words, int lane) {
    const uint32_t* s = (const uint32_t*)src;
    uint32_t* d = (uint32_t*)dst;
    for (int i = lane; i < words; i += 32) d[i] = s[i];
    __syncwarp();
}
static_assert(sizeof(GemvArgs) % 4 == 0 && sizeof(StepOp) % 4 == 0, "descriptor copies are word-wise");

// shared-memory carve-up behind the mat-vec kernel's own (ring, decoded headers, barriers, partial sums)
struct StepSmem {
    GemvArgs* slots;   // [3] the current mat-vec descriptor of the producer, decoder and epilogue warps
    StepOp* cslot;     // [2] the compute warps' current / next stage (double-buffered)
    int* aux;          // [0..1] per cslot: sequence length of an attention stage; [2] cta_stage; [4 + 12 cslot ..]: its AttnSplit and this CTA's first item
    uint8_t* xbase;    // activation buffers, re-carved per mat-vec; attention and argmax alias them
};
constexpr size_t STEP_EXTRA_SMEM = 3 * sizeof(GemvArgs) + 2 * sizeof(StepOp) + 28 * sizeof(int) + 16;

// ---- the stages that are not mat-vecs, run by the 256 compute threads (kept out of line: their registers
// ---- must not weigh on the mat-vec loop) --------------------------------------------------------------------
static __device__ __noinline__ void embed_stage(const StepOp& op, int tid) {
    int tok = *op.e_token;
    if (tok < 0 || tok >= op.e_vocab) tok = 0;
    const uint8_t* row = op.e_table + (size_t)tok * op.e_row_bytes;
    for (int g4 = blockIdx.x * NCT + tid; g4 < op.e_d_model / 4; g4 += gridDim.x * NCT) ((float4*)op.e_x)[g4] = dq4(op.e_type, row, g4);
}

// how an attention stage over T tokens is cut into work items (kv head, batch of <= 4 query heads, chunk),
// dealt round-robin over the grid
struct AttnSplit {
    int gq, n_batches, chunk_tokens, n_chunks, n_items;
    // cluster > 1: prefer chunks long enough that a (kv head, batch) fits the ranks of one cluster (in-cluster merge)
    __device__ AttnSplit(int n_head, int n_kv, int min_chunk, int T, int cluster = 0) {
        gq = n_head / n_kv;
        n_batches = (gq + A_MAXG - 1) / A_MAXG;
        chunk_tokens = attn_chunk_tokens(T, n_kv * n_batches, (int)gridDim.x, min_chunk);
        if (cluster > 1 && n_kv * n_batches <= (int)gridDim.x / cluster) {
            const int c = ((T + cluster - 1) / cluster + 31) & ~31;
            if (c <= STEP_ATTN_MAX_CHUNK && c > chunk_tokens) chunk_tokens = c;
        }
        n_chunks = (T + chunk_tokens - 1) / chunk_tokens;
        n_items = n_kv * n_batches * n_chunks;
    }
    // the integer divisions above cost ~1 us of dependent instructions: one thread does them while the CTA still waits
    // for the previous stage and leaves the result in shared memory
    __device__ void store(int* p) const { p[0] = gq; p[1] = n_batches; p[2] = chunk_tokens; p[3] = n_chunks; p[4] = n_items; }
    __device__ explicit AttnSplit(const int* p) : gq(p[0]), n_batches(p[1]), chunk_tokens(p[2]), n_chunks(p[3]), n_items(p[4]) {}
    __device__ void item(int i, int T, int* kvh, int* h0, int* hb, int* chunk, int* t0, int* n) const {
        *chunk = i % n_chunks;
        *kvh = i / (n_chunks * n_batches);
        *h0 = ((i / n_chunks) % n_batches) * A_MAXG;
        *hb = min(A_MAXG, gq - *h0);
        *t0 = *chunk * chunk_tokens;
        *n = min(T - *t0, chunk_tokens);
    }
};

static __device__ __noinline__ void attn_stage(const StepOp& op, uint8_t* xbase, int T, const int* split, uint64_t* abar, int* apar, int tid,
                                        unsigned long long* tr) {
    AttnSmem as;
    as.scores = (float*)xbase;
    as.prob = as.scores + A_MAXG * A_TOK;
    as.ml = as.prob + A_MAXG * A_TOK;
    as.flag = (int*)(as.ml + 2 * A_MAXG);
    as.inbox = (float*)(as.flag + 4);
    as.abar = abar;
    as.apar = apar;
    const AttnSplit sp(split);
    if (sp.n_chunks > op.a_max_chunks) asm volatile("trap;");   // scratch is sized for this bound (step_attn_max_chunks)
    if (split[11]) {
        // short sequences: cluster k holds (kv head, batch) k, its rank r chunk r; rank 0 merges (attn_item)
        if (split[5] >= 0) {
            if (op.a_hd == 64) attn_item<64>(op, split[5], split[6], split[7], split[8], sp.n_chunks, split[9], split[10], as, tid, true, tr);
            else attn_item<128>(op, split[5], split[6], split[7], split[8], sp.n_chunks, split[9], split[10], as, tid, true, tr);
        }
        return;
    }
    for (int item = blockIdx.x; item < sp.n_items; item += gridDim.x) {
        int kvh, h0, hb, chunk, t0, n;
        if (item == (int)blockIdx.x) {   // decomposed ahead of the barrier
            kvh = split[5]; h0 = split[6]; hb = split[7]; chunk = split[8]; t0 = split[9]; n = split[10];
        } else {
            sp.item(item, T, &kvh, &h0, &hb, &chunk, &t0, &n);
        }
        if (op.a_hd == 64) attn_item<64>(op, kvh, h0, hb, chunk, sp.n_chunks, t0, n, as, tid, false, tr);
        else attn_item<128>(op, kvh, h0, hb, chunk, sp.n_chunks, t0, n, as, tid, false, tr);
    }
}

// Between two tokens 4.6 GB of weights pass through L2, so whatever a stage reads besides the stream -- the KV
// cache rows of an attention stage, a norm weight vector -- would be a DRAM round trip on the stage's critical
// path, queued behind the weight prefetch.  One stage ahead (while this CTA still waits for the previous stage)
// one warp asks L2 to fetch them.  `op` is the NEXT stage's descriptor in global memory (static data).
static __device__ __noinline__ void prefetch_next_stage(const StepOp& op, int lane) {
    const int kind = op.kind;
    if (kind == OP_ATTN) {
        const int T = *op.a_pos + 1;
        const AttnSplit sp(op.a_n_head, op.a_n_kv, op.a_min_chunk, T);
        const int hd = op.a_hd;
        // the K and V rows of this CTA's items; the batches of a kv head share them (first batch fetches)
        for (int item = blockIdx.x + lane * gridDim.x; item < sp.n_items; item += 32 * gridDim.x) {
            int kvh, h0, hb, chunk, t0, n;
            sp.item(item, T, &kvh, &h0, &hb, &chunk, &t0, &n);
            if (h0 != 0) continue;
            const size_t off = ((size_t)kvh * op.a_n_ctx + t0) * hd;
            const uint32_t bytes = (uint32_t)n * hd * sizeof(__half);
            bulk_prefetch_l2(op.a_k + off, bytes);
            bulk_prefetch_l2(op.a_v + off, bytes);
        }
    } else if (kind == OP_GEMV) {
        // the norm weights: one CTA per 2 KiB piece
        if (op.g.prologue == PRO_NORM && lane == 0) {
            const int piece = 512, n_pieces = (op.g.K + piece - 1) / piece;
            for (int i = blockIdx.x; i < n_pieces; i += gridDim.x) {
                const int n = min(piece, op.g.K - i * piece);
                bulk_prefetch_l2(op.g.norm_w + (size_t)i * piece, (uint32_t)n * sizeof(float));
            }
        }
    }
}

// ---- tensor-parallel exchange at the head of a mat-vec stage (GemvArgs::x_parts) ----------------------------------------
// The grid barrier in front of the stage says: every CTA of THIS rank has stored its partial rows of the previous
// (row-split) mat-vec into every rank's inbox and fenced them at system scope.  One thread per peer then publishes this
// rank's sequence number in that peer's flag array (st.release.sys: cumulative over what the barrier made visible), and
// every CTA waits until all ranks' numbers have arrived here; build_activation then reads the inbox, summed in rank order
// (the same order on every rank: identical sums).  Peer memory is only ever written (plain stores); loads are local.
// Waits are bounded and trap.  Called by the NCT compute threads.
// (Measured and dropped: every CTA adding 1 to a per-exchange counter on every rank with red.release.sys, instead of
// one flag per rank after the local barrier -- 284 remote reductions on one address per rank and exchange: 8B at TP 2
// went from 419 to 386 tok/s.)
static __device__ __noinline__ void xchg_wait(const GemvArgs& a, int tid) {
    const unsigned int seq = *a.xg_epoch * 256u + (unsigned int)a.xg_k + 1u;
    if (blockIdx.x == 0 && tid < a.x_nparts) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.xg_peer_flags[tid]), "r"(seq) : "memory");
    }
    if (tid < a.x_nparts) {
        unsigned int v = 0;
        uint32_t spin = 0;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.xg_flags + tid) : "memory");
            if ((int)(v - seq) >= 0) break;
            if (++spin > (1u << 25)) asm volatile("trap;");
            __nanosleep(40);
        }
    }
    named_bar_sync(1, NCT);
}
// ... and the sum, every CTA a slice, is written to x: as a stage of its own (OP_XCHG) in front of the consuming
// mat-vec, so that its prologue reads one vector instead of every CTA adding up all the ranks' rows itself (at TP 8 that
// was 256 KB out of L2 per CTA and exchange).
static __device__ __noinline__ void xchg_materialise(const GemvArgs& a, int tid) {
    const int n4 = a.K / 4;
    if (a.x_part_ptr[0]) {
        // pull form: every rank left its partial row in its own memory; this CTA's slice of each is fetched over NVLink
        // (loads that bypass L1: a parity buffer is rewritten two exchanges later) and added in rank order
        for (int i = blockIdx.x * NCT + tid; i < n4; i += gridDim.x * NCT) {
            float4 v[8];
#pragma unroll
            for (int r = 0; r < 8; r++)
                if (r < a.x_nparts) asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v[r].x), "=f"(v[r].y), "=f"(v[r].z), "=f"(v[r].w)
                                                 : "l"((const float4*)a.x_part_ptr[r] + i) : "memory");
            float4 s = v[0];
#pragma unroll
            for (int r = 1; r < 8; r++)
                if (r < a.x_nparts) { s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w; }
            ((float4*)const_cast<float*>(a.x))[i] = s;
        }
        return;
    }
    const float4* in = (const float4*)a.x_parts;
    for (int i = blockIdx.x * NCT + tid; i < n4; i += gridDim.x * NCT) {
        float4 s = __ldcg(in + i);
        for (int r = 1; r < a.x_nparts; r++) {
            const float4 v = __ldcg(in + (size_t)r * n4 + i);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        ((float4*)const_cast<float*>(a.x))[i] = s;
    }
}

__device__ __forceinline__ void argmax_fold(float& bv, int& bi, float ov, int oi) {
    if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
}

// greedy sampling (first index of the maximum) + the device-resident loop's bookkeeping in the last CTA to arrive
static __device__ __noinline__ void argmax_stage(const StepOp& op, uint8_t* xbase, int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int64_t i = (int64_t)blockIdx.x * NCT + tid; i < op.s_n; i += (int64_t)gridDim.x * NCT) {
        const float v = op.s_logits[i];
        if (v > bv) { bv = v; bi = (int)i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) argmax_fold(bv, bi, __shfl_xor_sync(0xffffffffu, bv, o), __shfl_xor_sync(0xffffffffu, bi, o));
    float* sv = (float*)xbase;
    int* sidx = (int*)(sv + NCW);
    if (lane == 0) { sv[warp] = bv; sidx[warp] = bi; }
    named_bar_sync(1, NCT);
    if (warp != 0) return;
    bv = lane < NCW ? sv[lane] : -INFINITY;
    bi = lane < NCW ? sidx[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) argmax_fold(bv, bi, __shfl_xor_sync(0xffffffffu, bv, o), __shfl_xor_sync(0xffffffffu, bi, o));
    unsigned int arrived = 0;
    if (lane == 0) {
        op.s_val[blockIdx.x] = bv;
        op.s_idx[blockIdx.x] = bi;
        __threadfence();
        arrived = atomicAdd(op.s_counter, 1u) + 1u;
    }
    arrived = __shfl_sync(0xffffffffu, arrived, 0);
    if (arrived != gridDim.x) return;
    __threadfence();
    bv = -INFINITY;
    bi = 0x7fffffff;
    for (unsigned int b = lane; b < gridDim.x; b += 32) argmax_fold(bv, bi, __ldcg(op.s_val + b), __ldcg(op.s_idx + b));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) argmax_fold(bv, bi, __shfl_xor_sync(0xffffffffu, bv, o), __shfl_xor_sync(0xffffffffu, bi, o));
    if (lane == 0) {
        if (bi == 0x7fffffff) bi = 0;
        *op.s_token = bi;
        if (op.s_out) {
            const int s = *op.s_step;
            if (s < op.s_cap) op.s_out[s] = bi;
            *op.s_step = s + 1;
        }
        if (op.s_pos) *op.s_pos = *op.s_pos + 1;
        if (op.s_epoch) *op.s_epoch = *op.s_epoch + 1u;
        *op.s_counter = 0u;
    }
}
}  // namespace stepdev
}  // namespace b2
