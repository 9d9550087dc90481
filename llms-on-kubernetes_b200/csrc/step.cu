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
//   * everything a stage needs that does not depend on the previous stage's results -- its descriptor,
//     the position -- is fetched into shared memory before the wait on the grid barrier;
//   * embedding lookup, attention (step_attn.cuh) and greedy sampling run as stages on the compute warps.
// All CTAs must be co-resident (they wait on each other): the grid is sized by
// cudaOccupancyMaxActiveClusters, and every wait is bounded and traps instead of hanging.
#include "step.cuh"

#include <cstdlib>

#include "step_stages.cuh"

namespace b2 {

using namespace gemvdev;
using namespace stepattn;
using namespace stepdev;

namespace {

// ---- the weight ring of the persistent kernel ------------------------------------------------------------------
// One allocation per (tile, K-chunk): 8 row slots whose pitch depends on the weight type, placed back to back in a
// circular buffer (an allocation that would straddle the end starts at 0 again).  Up to NSLOT allocations are in
// flight -- four Q4_K tiles (18.6 KB each) or three Q6_K tiles (27 KB): with the tile loop at ~1 us per tile, three
// Q4_K tiles in flight left the compute warps waiting for data a third of the time.  Producer, decoder and compute
// warps walk the same sequence of sizes, so their offsets agree without communication; the producer also waits for
// the release of any older allocation its new one overlaps.
constexpr int NSLOT = 4;
template <int ACT>
struct Ring {
    static constexpr int CHUNK_BLOCKS = Geo<ACT>::CHUNK_BLOCKS;
    static constexpr uint32_t BYTES = ACT == ACT_Q8K ? 3 * TILE_ROWS * (16 * 210 + 16) : 4 * TILE_ROWS * (64 * 34 + 16);
    // decoded Q4_K / Q5_K headers of one slot: (d, dmin) pairs and scale words in separate, bank-padded rows
    static constexpr int DD_ROW = CHUNK_BLOCKS * 8 + 8, W_ROW = CHUNK_BLOCKS * 16 + 16;
    static constexpr int DEC_SLOT = ACT == ACT_Q8K ? TILE_ROWS * (DD_ROW + W_ROW) : 0;
    __host__ __device__ static constexpr uint32_t pitch(int bb) { return (uint32_t)(CHUNK_BLOCKS * bb + 16); }
};
struct RingPos {
    uint32_t off = 0, par = 0;   // par bit s: parity of the next use of slot s
    int slot = 0;
    template <int ACT>
    __device__ __forceinline__ uint32_t place(uint32_t sz) {
        if (off + sz > Ring<ACT>::BYTES) off = 0;
        const uint32_t o = off;
        off += sz;
        return o;
    }
    __device__ __forceinline__ uint32_t parity() const { return (par >> slot) & 1u; }
    __device__ __forceinline__ void next() {
        par ^= 1u << slot;
        slot = (slot + 1) & (NSLOT - 1);
    }
};

// a compute warp's position in the weight ring and the partial-sum hand-off, carried from tile to tile and stage to stage
struct TileWalk {
    uint8_t *ring, *dec;
    float* red;
    uint64_t *ready, *empty, *part_full, *part_free;
    RingPos rp;
    int parity, fph;
    long long cyc_ready, cyc_dot, cyc_free;   // B2_STEP_TRACE
};

// The tiles of one segment that belong to this CTA (tile, tile + grid, ... < tile_end), for a weight type known at
// compile time: nothing in the loop depends on the segment but the row-alignment offsets of the 2-byte-aligned types.
template <int ACT, int TYPE>
__device__ __forceinline__ void segment_tiles(const GemvSeg& s, int& tile, int tile_end, int nblk_row, int n_chunks, const ActView& xv,
                                              int warp, int lane, TileWalk& w, bool tracing) {
    using G = Geo<ACT>;
    using R = Ring<ACT>;
    constexpr int bb = TYPE == B2_TYPE_Q4_K ? 144 : TYPE == B2_TYPE_Q5_K ? 176 : TYPE == B2_TYPE_Q6_K ? 210 : 34;
    constexpr bool unaligned = TYPE == B2_TYPE_Q6_K || TYPE == B2_TYPE_Q8_0;
    constexpr uint32_t chunk_step = (uint32_t)(G::CHUNK_BLOCKS * bb) & 15;
    constexpr uint32_t pitch = R::pitch(bb), tile_bytes = TILE_ROWS * pitch;
    const int r = lane >> 2;
    const int step = (int)gridDim.x;
    for (; tile < tile_end; tile += step) {
        uint32_t off0 = 0;
        if (unaligned) off0 = (uint32_t)((uintptr_t)slot_row_ptr(s, tile - s.tile0, r, nblk_row * bb) & 15);
        float acc = 0.0f;
        for (int c = 0; c < n_chunks; c++) {
            const int nb = min(G::CHUNK_BLOCKS, nblk_row - c * G::CHUNK_BLOCKS);
            const uint32_t o = w.rp.place<ACT>(tile_bytes);
            const int sl = w.rp.slot;
            const long long c0 = tracing ? clock64() : 0;
            mbar_wait(&w.ready[sl], w.rp.parity());
            const long long c1 = tracing ? clock64() : 0;
            const uint8_t* row = w.ring + o + r * pitch + ((off0 + c * chunk_step) & 15);
            const uint8_t* ds = w.dec + sl * R::DEC_SLOT;
            const DecPtr drow{ds + r * R::DD_ROW, ds + TILE_ROWS * R::DD_ROW + r * R::W_ROW};
            acc += dot_chunk<ACT, 8, 16>(TYPE, row, drow, nb, c * G::CHUNK_BLOCKS, xv, warp, lane);
            __syncwarp();
            if (lane == 0) mbar_arrive(&w.empty[sl]);
            if (tracing) {
                const long long c2 = clock64();
                w.cyc_ready += c1 - c0;
                w.cyc_dot += c2 - c1 + (__float_as_uint(acc) == 0x7fc0dead ? 1 : 0);
            }
            w.rp.next();
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        const long long c3 = tracing ? clock64() : 0;
        mbar_wait(&w.part_free[w.parity], (w.fph >> w.parity) & 1);
        if (tracing) w.cyc_free += clock64() - c3;
        w.fph ^= 1 << w.parity;
        if ((lane & 3) == 0) w.red[(w.parity * NCW + warp) * TILE_ROWS + r] = acc;
        __syncwarp();
        if (lane == 0) mbar_arrive(&w.part_full[w.parity]);
        w.parity ^= 1;
    }
}

template <int ACT>
__global__ void __launch_bounds__(GEMV_THREADS, 2) step_kernel(const __grid_constant__ StepProgram P) {
    using G = Geo<ACT>;
    using R = Ring<ACT>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem;
    uint8_t* dec = smem + R::BYTES;
    uint64_t* full = (uint64_t*)(dec + NSLOT * R::DEC_SLOT);
    uint64_t* ready = full + NSLOT;
    uint64_t* empty = ready + NSLOT;
    uint64_t* part_full = empty + NSLOT;
    uint64_t* part_free = part_full + 2;
    ActSmem sm;
    sm.xbar = part_free + 2;   // 17 barriers; one pad slot keeps the 16-byte alignment below
    sm.ss_warp = (double*)(sm.xbar + 2);
    float* red = (float*)(sm.ss_warp + NCW);
    StepSmem ss;
    ss.slots = (GemvArgs*)(red + 2 * NCW * TILE_ROWS);
    ss.cslot = (StepOp*)(ss.slots + 3);
    ss.aux = (int*)(ss.cslot + 2);
    ss.xbase = (uint8_t*)(((uintptr_t)(ss.aux + 28) + 15) & ~(uintptr_t)15);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_ctarank(), csize = cluster_nctarank();
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NSLOT; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&ready[s], 1);
            mbar_init(&empty[s], NCW);
        }
        mbar_init(&part_full[0], NCW);
        mbar_init(&part_full[1], NCW);
        mbar_init(&part_free[0], 1);
        mbar_init(&part_free[1], 1);
        mbar_init(sm.xbar, 1);
        ss.aux[2] = 0;   // cta_stage: the last grid-barrier stage number this CTA has seen complete
        mbar_init(sm.xbar + 1, 1);   // attention records arriving from the cluster (step_attn.cuh); aux[3] = its next phase
        ss.aux[3] = 0;
        mbar_fence_init();
    }
    cluster_sync_all();
    volatile unsigned int* cta_stage = (volatile unsigned int*)&ss.aux[2];

    if (warp == W_PRODUCER) {
        // ===== producer: streams every mat-vec stage's weights in program order, never waits for a stage to start =====
        RingPos rp;
        // the last three allocations (they may still be in use): range, slot and the parity of that use of the slot
        uint32_t h_off[3] = {0, 0, 0}, h_end[3] = {0, 0, 0}, h_par[3] = {0, 0, 0};
        int h_slot[3] = {0, 0, 0};
        const uint64_t stream_policy = l2_policy_evict_first();
        for (int oi = 0; oi < P.n_ops; oi++) {
            const StepOp& op = P.ops[oi];
            if (op.kind != OP_GEMV) continue;
            __syncwarp();
            stage_words(&ss.slots[0], &op.g, sizeof(GemvArgs) / 4, lane);
            const GemvArgs& a = ss.slots[0];
            const int nblk_row = a.K / (ACT == ACT_Q8K ? 256 : 32);
            const int n_chunks = (nblk_row + G::CHUNK_BLOCKS - 1) / G::CHUNK_BLOCKS;
            int si = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
                while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
                const GemvSeg& s = a.seg[si];
                const int bb = block_bytes_of(s.type);
                const uint32_t pitch = R::pitch(bb), tile_bytes = TILE_ROWS * pitch;
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
                    const uint32_t o = rp.place<ACT>(tile_bytes);
                    const int sl = rp.slot;
                    const uint32_t par = rp.parity();
                    if (lane == 0) {
                        mbar_wait(&empty[sl], par ^ 1);   // the slot's previous use (four allocations ago) is released
#pragma unroll
                        for (int j = 0; j < 3; j++)       // ... and so is any younger allocation these bytes overlap
                            if (h_end[j] > h_off[j] && o < h_end[j] && h_off[j] < o + tile_bytes) mbar_wait(&empty[h_slot[j]], h_par[j]);
                        mbar_arrive_expect_tx(&full[sl], total);
                    }
                    __syncwarp();
                    if (sz) bulk_g2s_hint(ring + o + lane * pitch, src, sz, &full[sl], stream_policy);
                    h_off[2] = h_off[1]; h_end[2] = h_end[1]; h_slot[2] = h_slot[1]; h_par[2] = h_par[1];
                    h_off[1] = h_off[0]; h_end[1] = h_end[0]; h_slot[1] = h_slot[0]; h_par[1] = h_par[0];
                    h_off[0] = o; h_end[0] = o + tile_bytes; h_slot[0] = sl; h_par[0] = par;
                    rp.next();
                }
            }
        }
        return;
    }

    if (warp == W_DECODER) {
        RingPos rp;
        for (int oi = 0; oi < P.n_ops; oi++) {
            const StepOp& op = P.ops[oi];
            if (op.kind != OP_GEMV) continue;
            __syncwarp();
            stage_words(&ss.slots[1], &op.g, sizeof(GemvArgs) / 4, lane);
            const GemvArgs& a = ss.slots[1];
            const int nblk_row = a.K / (ACT == ACT_Q8K ? 256 : 32);
            const int n_chunks = (nblk_row + G::CHUNK_BLOCKS - 1) / G::CHUNK_BLOCKS;
            int si = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
                while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
                const int type = a.seg[si].type;
                const int bb = block_bytes_of(type);
                const uint32_t pitch = R::pitch(bb), tile_bytes = TILE_ROWS * pitch;
                for (int c = 0; c < n_chunks; c++) {
                    const int nb = min(G::CHUNK_BLOCKS, nblk_row - c * G::CHUNK_BLOCKS);
                    const uint32_t o = rp.place<ACT>(tile_bytes);
                    const int sl = rp.slot;
                    mbar_wait(&full[sl], rp.parity());
                    if (ACT == ACT_Q8K && (type == B2_TYPE_Q4_K || type == B2_TYPE_Q5_K)) {
                        uint8_t* ds = dec + sl * R::DEC_SLOT;
                        for (int i = lane; i < TILE_ROWS * nb; i += 32) {
                            const int rr = i / nb, b = i - rr * nb;
                            decode_k4_header(ring + o + rr * pitch + b * bb, ds + rr * R::DD_ROW + b * 8,
                                             ds + TILE_ROWS * R::DD_ROW + rr * R::W_ROW + b * 16);
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&ready[sl]);
                    rp.next();
                }
            }
        }
        return;
    }

    if (warp == W_EPILOGUE) {
        // ===== epilogue: finish row pairs; after the CTA's last tile of a stage, arrive on the grid barrier =====
        int parity = 0, ph = 0;
        int pos = 0;
        for (int oi = 0; oi < P.n_ops; oi++) {
            const StepOp& op = P.ops[oi];
            if (op.kind != OP_GEMV) continue;   // the compute warps signal those stages
            // static data first (descriptor; the position only changes in the step's last stage) ...
            stage_words(&ss.slots[2], &op.g, sizeof(GemvArgs) / 4, lane);
            const GemvArgs& a = ss.slots[2];
            if (a.pos_ptr) pos = *a.pos_ptr;
            // ... then never arrive for a stage before the previous one is complete everywhere: a CTA without
            // tiles in this stage would otherwise run ahead and release other CTAs early
            if (oi > 0) {
                if (lane == 0) grid_wait_local(cta_stage, (unsigned int)oi);
                __syncwarp();
            }
            int si = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, parity ^= 1) {
                while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
                const GemvSeg& s = a.seg[si];
                epilogue_tile(a, s, tile - s.tile0, pos, red + parity * NCW * TILE_ROWS, &part_full[parity], &part_free[parity],
                              (ph >> parity) & 1, lane);
                ph ^= 1 << parity;
            }
            // a fence orders only the calling thread's stores: every lane that wrote results fences, then one arrives
            __threadfence();
            __syncwarp();
            if (lane == 0) {
                grid_arrive(P.sync_counter, (unsigned int)oi + 1u);
                stamp(P, oi, 3);
            }
        }
        return;
    }
    // ===== compute warps =====
    const int r = lane >> 2;
    RingPos rpos;
    int parity = 0, fph = 3;
    uint32_t xpar = 0;
    for (int oi = 0; oi < P.n_ops; oi++) {
        // warp 0 stages the stage's descriptor (static) while the previous stage is still finishing elsewhere;
        // the other buffer may still be read by slower warps of this CTA
        StepOp* cs = &ss.cslot[oi & 1];
        if (warp == 0) {
            stage_words(cs, &P.ops[oi], sizeof(StepOp) / 4, lane);
            if (lane == 0 && cs->kind == OP_ATTN) {
                const int T = *cs->a_pos + 1;
                ss.aux[oi & 1] = T;
                int* pre = ss.aux + 4 + 12 * (oi & 1);
                const AttnSplit sp(cs->a_n_head, cs->a_n_kv, cs->a_min_chunk, T, (cs->a_cluster_merge && (int)csize <= A_MAXCL) ? (int)csize : 0);
                sp.store(pre);
                // a few chunks per (kv head, batch) and enough clusters: one cluster per (kv head, batch), merged through
                // distributed shared memory; otherwise items round-robin over the grid, merged through global memory
                const int groups = cs->a_n_kv * sp.n_batches;
                const bool in_cluster = cs->a_cluster_merge && sp.n_chunks > 1 && sp.n_chunks <= (int)csize && (int)csize <= A_MAXCL &&
                                        groups <= (int)(gridDim.x / csize);
                pre[11] = in_cluster;
                if (in_cluster) {
                    const int cl = (int)(blockIdx.x / csize);
                    pre[5] = -1;
                    if (cl < groups && (int)rank < sp.n_chunks) sp.item(cl * sp.n_chunks + (int)rank, T, pre + 5, pre + 6, pre + 7, pre + 8, pre + 9, pre + 10);
                } else if ((int)blockIdx.x < sp.n_items) {
                    sp.item((int)blockIdx.x, T, pre + 5, pre + 6, pre + 7, pre + 8, pre + 9, pre + 10);
                }
            }
            if (oi + 1 < P.n_ops) prefetch_next_stage(P.ops[oi + 1], lane);
            // every CTA has arrived oi times per completed stage: wait for the previous stage everywhere
            if (oi > 0 && lane == 0) grid_wait(P.sync_counter, (unsigned int)oi, cta_stage);
        }
        named_bar_sync(1, NCT);
        if (tid == 0) stamp(P, oi, 0);
        const StepOp& op = *cs;
        if (op.kind == OP_GEMV) {
            const GemvArgs& a = op.g;
            const int K = a.K;
            sm.xq = (int8_t*)ss.xbase;
            sm.xd = (float*)(sm.xq + K);
            sm.bs16 = (int16_t*)(sm.xd + (ACT == ACT_Q8K ? K / 256 : K / 32));
            sm.bs32 = sm.bs16 + K / 16;
            build_activation<ACT>(a, sm, smem, tid, rank, csize, xpar, (P.trace && blockIdx.x == 0) ? P.trace + oi * STEP_TRACE_SLOTS : nullptr);
            if (tid == 0) stamp(P, oi, 1);
            xpar ^= 1;
            const ActView xv{sm.xq, sm.xd, sm.bs16, sm.bs32};
            const int nblk_row = K / (ACT == ACT_Q8K ? 256 : 32);
            const int n_chunks = (nblk_row + G::CHUNK_BLOCKS - 1) / G::CHUNK_BLOCKS;
            // this CTA's tiles, segment by segment (a segment has one weight type: the tile loop is specialised for it)
            TileWalk tw{ring, dec, red, ready, empty, part_full, part_free, rpos, parity, fph, 0, 0, 0};
            const bool tracing = P.trace && blockIdx.x == 0 && tid == 0;
            int tile = blockIdx.x;
            for (int si = 0; si < a.n_seg; si++) {
                const GemvSeg& s = a.seg[si];
                const int tile_end = si + 1 < a.n_seg ? a.seg[si + 1].tile0 : a.n_tiles;
                if (tile >= tile_end) continue;
                if constexpr (ACT == ACT_Q8K) {
                    if (s.type == B2_TYPE_Q4_K) segment_tiles<ACT, B2_TYPE_Q4_K>(s, tile, tile_end, nblk_row, n_chunks, xv, warp, lane, tw, tracing);
                    else if (s.type == B2_TYPE_Q6_K) segment_tiles<ACT, B2_TYPE_Q6_K>(s, tile, tile_end, nblk_row, n_chunks, xv, warp, lane, tw, tracing);
                    else segment_tiles<ACT, B2_TYPE_Q5_K>(s, tile, tile_end, nblk_row, n_chunks, xv, warp, lane, tw, tracing);
                } else {
                    segment_tiles<ACT, B2_TYPE_Q8_0>(s, tile, tile_end, nblk_row, n_chunks, xv, warp, lane, tw, tracing);
                }
            }
            rpos = tw.rp; parity = tw.parity; fph = tw.fph;
            if (tracing) {
                // cycles warp 0 of CTA 0 spent waiting for data / in the dot products / waiting for the epilogue warp (marks 8..10)
                unsigned long long* tr = P.trace + oi * STEP_TRACE_SLOTS;
                tr[8] = tr[0] + (unsigned long long)(tw.cyc_ready * 1000 / 1965);
                tr[9] = tr[0] + (unsigned long long)(tw.cyc_dot * 1000 / 1965);
                tr[10] = tr[0] + (unsigned long long)(tw.cyc_free * 1000 / 1965);
            }
            if (tid == 0) stamp(P, oi, 2);
            // the epilogue warp arrives on the grid barrier for this stage
        } else {
            if (op.kind == OP_EMBED) embed_stage(op, tid);
            else if (op.kind == OP_ATTN)
                attn_stage(op, ss.xbase, ss.aux[oi & 1], ss.aux + 4 + 12 * (oi & 1), sm.xbar + 1, ss.aux + 3, tid,
                           (P.trace && blockIdx.x == 0) ? P.trace + oi * STEP_TRACE_SLOTS : nullptr);
            else if (op.kind == OP_ARGMAX) argmax_stage(op, ss.xbase, tid);
            __threadfence();   // every thread's stores of this stage, before the CTA's single arrival
            named_bar_sync(1, NCT);
            if (tid == 0) {
                grid_arrive(P.sync_counter, (unsigned int)oi + 1u);
                stamp(P, oi, 3);
            }
        }
    }
}

size_t step_smem_bytes(int act, int max_k) {
    // ring + decoded headers + 18 barrier slots + RMS partial sums + the tile's partial dot products + the descriptors ...
    size_t n = (act == ACT_Q8K ? Ring<ACT_Q8K>::BYTES + NSLOT * Ring<ACT_Q8K>::DEC_SLOT : Ring<ACT_Q80>::BYTES) + 18 * sizeof(uint64_t) +
               NCW * sizeof(double) + 2 * NCW * TILE_ROWS * sizeof(float) + STEP_EXTRA_SMEM;
    // ... + the quantised activation vector
    n += (size_t)max_k + (act == ACT_Q8K ? max_k / 256 : max_k / 32) * 4 + max_k / 16 * 2 + max_k / 32 * 2;
    // attention aliases the activation buffers; make sure they are large enough
    const size_t x_bytes = (size_t)max_k + (act == ACT_Q8K ? max_k / 256 : max_k / 32) * 4 + max_k / 16 * 2 + max_k / 32 * 2;
    if (x_bytes < ATTN_SMEM_BYTES) n += ATTN_SMEM_BYTES - x_bytes;
    return (n + 127) & ~(size_t)127;
}

int step_config(int act, int max_k, size_t* smem_out, int* cluster_out, int* grid_out) {
    const int g_step_sms = device_sm_count();
    if (g_step_sms <= 0) return B2_ECUDA;
    const size_t smem = step_smem_bytes(act, max_k);
    if (smem > 227 * 1024) {
        set_error("step kernel: K=%d needs %zu bytes of shared memory", max_k, smem);
        return B2_ERANGE;
    }
    auto kern = act == ACT_Q8K ? step_kernel<ACT_Q8K> : step_kernel<ACT_Q80>;
    B2_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int ctas_per_sm = (int)((228 * 1024) / (smem + 1024));
    if (ctas_per_sm > 2) ctas_per_sm = 2;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    // CTAs that share the work of quantising the activation vector (B2_STEP_CLUSTER: 1, 2, 4 or 8)
    // measured on Llama-3.1-8B: 4 is best (2.85 ms/token; 8: 2.92, 2: 2.90, 1: 3.09)
    int cluster = 4;
    if (const char* e = getenv("B2_STEP_CLUSTER")) {
        const int v = atoi(e);
        if (v == 1 || v == 2 || v == 4 || v == 8) cluster = v;
    }
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

int step_attn_max_chunks(int n_ctx, int grid) {
    // attn_chunk_tokens(): below the cap the items fill the grid once (at most grid + 1 chunks when a kv head is a
    // single item), at the cap a chunk is STEP_ATTN_MAX_CHUNK tokens
    const int by_cap = (n_ctx + STEP_ATTN_MAX_CHUNK - 1) / STEP_ATTN_MAX_CHUNK;
    const int by_grid = grid + 2;
    return by_cap > by_grid ? by_cap : by_grid;
}
size_t step_attn_partials_bytes(int n_head, int head_dim, int n_ctx, int grid) {
    return (size_t)n_head * step_attn_max_chunks(n_ctx, grid) * ((size_t)head_dim + 2) * sizeof(float);
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
    B2_CUDA_CHECK(cudaMemsetAsync(p.sync_counter, 0, STEP_SYNC_BYTES, st));
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
