// step2.cu -- the persistent per-token decode kernel, second design (SURVEY 8a rows a10-a16 in one launch).
//
// Same program of stages, grid barrier, attention / embedding / sampling stages and epilogues as step.cu; what changed
// is how the weight stream reaches the arithmetic (matvec_team.cuh: the first design was bound by instruction issue):
//   * tile = 8 weight rows (4 row pairs); ring allocation = one STEP = the tile's rows x 16 blocks of K, copied by the
//     producer warp with one bulk copy per row (lanes 0..7 each announce and issue their row's copy); five Q4_K steps
//     fit the 92.8 KB byte ring;
//   * compute warp w owns blocks 2w, 2w + 1 of the step for all 8 rows: Q4_K through ldmatrix + mma.sync (u8 x s8), the
//     other types through dp4a quarters; a lane keeps one fp32 row sum over the tile's whole K loop; the 6-bit scales
//     are unpacked in the lane: no decoder warp, one barrier per step;
//   * ten warps (8 compute, producer, epilogue): 320 threads x 2 CTAs per SM leave 96 registers per thread, enough to keep
//     the activation fragments of a single-chunk row in registers for the whole stage.  Measured and left out: a second
//     producer warp (the eleventh warp caps the kernel at 80 registers: 461 -> 417 tok/s) and a look-ahead warp that
//     asked L2 to fetch rows some steps ahead (cp.async.bulk.prefetch.L2: -1 .. +0.3 %; the stream is limited by how
//     fast copies are issued and slots released, not by DRAM latency -- tools/ubench_stream.cu reaches 7.2 TB/s with
//     108 KB in flight per SM, and prefetching on top of a ring that deep lowers it);
//   * a normalising stage's RMS factor is applied by the epilogue warp to the finished row sums (build_activation DEFER):
//     the quantiser does not wait for the sum of squares;
//   * TP instantiation: row-split stages leave partial rows for the other ranks, OP_XCHG stages add the ranks' rows.
// B2_STEP_V=1 selects the first design (step.cu) for comparison.
#include <cstdlib>

#include "matvec_team.cuh"
#include "step_stages.cuh"

namespace b2 {

using namespace gemvdev;
using namespace stepattn;
using namespace stepdev;
using namespace teamdev;

namespace {

// ten warps: 8 compute, producer, epilogue (320 threads x 2 CTAs per SM leave 96 registers per thread)
constexpr int W_PROD = NCW, W_EPI = NCW + 1;
constexpr int STEP2_THREADS = (NCW + 2) * 32;

struct RingPos2 {
    uint32_t off = 0, par = 0;   // par bit s: parity of the next use of slot s
    int slot = 0;
    __device__ __forceinline__ uint32_t place(uint32_t sz) {
        if (off + sz > TS_RING_BYTES) off = 0;
        const uint32_t o = off;
        off += sz;
        return o;
    }
    __device__ __forceinline__ uint32_t parity() const { return (par >> slot) & 1u; }
    __device__ __forceinline__ void next() {
        par ^= 1u << slot;
        slot = (slot + 1) & (TS_NSLOT - 1);
    }
};

struct TileWalk2 {
    uint8_t* ring;
    float* red;
    uint64_t *full, *empty, *part_full, *part_free;
    RingPos2 rp;
    int parity, fph;
    long long cyc_ready, cyc_dot, cyc_free;   // B2_STEP_TRACE
    unsigned long long* fine;                 // B2_STEP_FINE: this stage's per-step stamps, or nullptr
    int fine_k;
    bool skip_math;                           // B2_STEP_FINE=-2: release every step unread (how fast can the producer alone go?)
};

// The tiles of one segment that belong to this CTA, for a weight type known at compile time.
template <int ACT, int TYPE>
__device__ __forceinline__ void segment_tiles2(const GemvSeg& s, int& tile, int tile_end, int nblk, const ActView& xa, int warp, int lane,
                                               TileWalk2& w, bool tracing) {
    constexpr int CLS = XClass<TYPE>::value;
    constexpr bool unaligned = TYPE == B2_TYPE_Q6_K || TYPE == B2_TYPE_Q8_0;
    constexpr int bb = TYPE == B2_TYPE_Q4_K ? 144 : TYPE == B2_TYPE_Q5_K ? 176 : TYPE == B2_TYPE_Q6_K ? 210 : 8 * 34;
    constexpr uint32_t pitch = MT_CHUNK * bb + 16, step_bytes = TILE_ROWS * pitch;
    constexpr uint32_t chunk_drift = (uint32_t)(MT_CHUNK * bb) & 15;   // alignment drift of a row's data per K chunk
    const int row_bytes = nblk * bb;
    const int n_chunks = (nblk + MT_CHUNK - 1) / MT_CHUNK;
    const int g = lane >> 2, t = lane & 3;
    const int step = (int)gridDim.x;
    const MmaLane ml = mma_lane(lane);
    const uint32_t xq_s = smem_u32(xa.q), bs32_s = smem_u32(xa.bs32), xd_s = smem_u32(xa.d);
    // ldmatrix row address of this lane inside a step: row lane & 7, block 2 warp + ((lane >> 3) & 1), qs bytes 16 (lane >> 4) ..
    const uint32_t lm_off = (uint32_t)((lane & 7) * pitch + (2 * warp + ((lane >> 3) & 1)) * bb + 16 + 16 * (lane >> 4));
    const uint32_t hdr_off = (uint32_t)(g * pitch + (2 * warp + (t & 1)) * bb);
    // a row of one K chunk: the warp's two blocks, hence its activations, are the same for every tile of the stage
    XMma xm;
    if (TYPE == B2_TYPE_Q4_K && n_chunks == 1 && 2 * warp < nblk) load_x_mma(xm, ml, xq_s, bs32_s, xd_s, 2 * warp, min(2, nblk - 2 * warp));
    for (; tile < tile_end; tile += step) {
        // where the row's first byte sits inside its 16-byte-aligned ring slot (2-byte-aligned types only)
        uint32_t off_g = 0;
        if (unaligned) off_g = (uint32_t)((uintptr_t)slot_row_ptr(s, tile - s.tile0, g, row_bytes) & 15);
        float acc = 0.0f;
        for (int c = 0; c < n_chunks; c++) {
            const uint32_t o = w.rp.place(step_bytes);
            const int sl = w.rp.slot;
            const long long c0 = tracing ? clock64() : 0;
            const bool fine = tracing && w.fine && w.fine_k < STEP_FINE_STEPS;
            if (fine) w.fine[8 * w.fine_k] = (unsigned long long)clock64();
            mbar_wait(&w.full[sl], w.rp.parity());
            if (fine) w.fine[8 * w.fine_k + 1] = (unsigned long long)clock64();
            const long long c1 = tracing ? clock64() : 0;
            const int kb0 = c * MT_CHUNK + 2 * warp;
            if (kb0 < nblk && !w.skip_math) {
                const uint8_t* base = w.ring + o;
                if (TYPE == B2_TYPE_Q4_K) {
                    if (n_chunks > 1) load_x_mma(xm, ml, xq_s, bs32_s, xd_s, kb0, min(2, nblk - kb0));
                    acc += q4k_rows8_blocks2_mma(smem_u32(base) + lm_off, base + hdr_off, ml, xm);
                } else {
                    const uint8_t* row = base + g * pitch + ((off_g + c * chunk_drift) & 15);
#pragma unroll
                    for (int b = 0; b < 2; b++) {
                        if (kb0 + b < nblk) {
                            XR x;
                            load_x<CLS>(x, xa, kb0 + b, t);
                            acc += row_quarter<TYPE>(row, 2 * warp + b, t, x);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&w.empty[sl]);
            w.rp.next();
            if (tracing) {
                const long long c2 = clock64();
                w.cyc_ready += c1 - c0;
                w.cyc_dot += c2 - c1 + (__float_as_uint(acc) == 0x7fc0dead ? 1 : 0);
                if (fine) w.fine[8 * w.fine_k++ + 2] = (unsigned long long)clock64();
            }
        }
        // the four lanes of a row add up; the warp's 8 row sums go to the epilogue warp
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        const long long c3 = tracing ? clock64() : 0;
        mbar_wait(&w.part_free[w.parity], (w.fph >> w.parity) & 1);
        if (tracing) w.cyc_free += clock64() - c3;
        w.fph ^= 1 << w.parity;
        if (t == 0) w.red[(w.parity * NCW + warp) * TILE_ROWS + g] = acc;
        __syncwarp();
        if (lane == 0) mbar_arrive(&w.part_full[w.parity]);
        w.parity ^= 1;
    }
}

constexpr size_t STEP2_EXTRA_SMEM = 3 * sizeof(GemvArgs) + 2 * sizeof(StepOp) + 32 * sizeof(int) + 16;

// does this stage's prologue leave the RMS factor to the epilogue warp?  (the condition of build_activation's DEFER branch)
template <int ACT, bool TP>
__device__ __forceinline__ bool deferred_scale_of(const GemvArgs& a) {
    return ACT == ACT_Q8K && a.prologue == PRO_NORM && !(TP && a.x_nparts > 0);
}

// TP: the instantiation a tensor-parallel shard runs (partial residual rows stored to the peers, inputs summed over the
// ranks); the single-GPU instantiation carries none of it
template <int ACT, bool TP>
__global__ void __launch_bounds__(STEP2_THREADS, 2) step2_kernel(const __grid_constant__ StepProgram P) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem;
    uint64_t* full = (uint64_t*)(smem + TS_RING_BYTES);
    uint64_t* empty = full + TS_NSLOT;
    uint64_t* part_full = empty + TS_NSLOT;
    uint64_t* part_free = part_full + 2;
    ActSmem sm;
    sm.xbar = part_free + 2;   // 2 barriers: activation exchange, attention records
    sm.ss_warp = (double*)(sm.xbar + 2);
    float* red = (float*)(sm.ss_warp + NCW);
    StepSmem ss;
    ss.slots = (GemvArgs*)(red + 2 * NCW * TILE_ROWS);
    ss.cslot = (StepOp*)(ss.slots + 3);
    ss.aux = (int*)(ss.cslot + 2);
    ss.xbase = (uint8_t*)(((uintptr_t)(ss.aux + 32) + 15) & ~(uintptr_t)15);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_ctarank(), csize = cluster_nctarank();
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < TS_NSLOT; s++) {
            mbar_init(&full[s], 1);
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
    float* rms_scale = (float*)&ss.aux[28];   // RMS factor of the current normalising stage (written by the prologue, applied by the epilogue warp)

    if (warp == W_PROD) {
        // ===== producer: streams every mat-vec stage's weights in program order, never waits for a stage to start.
        // (A second producer warp taking half of the rows was measured: no gain, and the tenth warp's registers are
        // worth more to the compute warps.)  The slot of a step is not waited for by itself: eight barrier slots
        // cover more bytes than the ring holds, so its previous use was released before some younger allocation
        // overwrote it (static_assert below); only the allocations the new bytes overlap are waited for.
        static_assert(TS_NSLOT * TILE_ROWS * (MT_CHUNK * 144 + 16) > TS_RING_BYTES, "a slot's previous use must be out of the ring");
        constexpr int half = 0, rows_mine = TILE_ROWS;
        RingPos2 rp;
        uint32_t my_off = 0, my_end = 0;   // lane j < TS_NSLOT: the latest allocation placed in slot j
        const uint64_t stream_policy = l2_policy_evict_first();
        GemvArgs* slot = &ss.slots[half];
        const int rslot = rows_mine * half + (lane & (rows_mine - 1));   // the row slot of the tile this lane copies
        for (int oi = 0; oi < P.n_ops; oi++) {
            const StepOp& op = P.ops[oi];
            if (op.kind != OP_GEMV) continue;
            __syncwarp();
            stage_words(slot, &op.g, sizeof(GemvArgs) / 4, lane);
            const GemvArgs& a = *slot;
            const int nblk = a.K >> 8, n_seg = a.n_seg, n_tiles = a.n_tiles;
            const int n_chunks = (nblk + MT_CHUNK - 1) / MT_CHUNK;
            unsigned long long* fine = (half == 0 && P.trace && P.fine_op == oi && blockIdx.x == 0 && lane == 0) ? P.trace + (size_t)P.n_ops * STEP_TRACE_SLOTS : nullptr;
            int fine_k = 0;
            int tile = blockIdx.x;
            for (int si = 0; si < n_seg; si++) {
                const int tile_end = si + 1 < n_seg ? a.seg[si + 1].tile0 : n_tiles;
                if (tile >= tile_end) continue;
                // the segment's descriptor in registers (the barrier operations below are memory clobbers)
                const GemvSeg s = a.seg[si];
                const int bb = ts_block_bytes(s.type);
                const uint32_t pitch = ts_pitch(s.type), step_bytes = TILE_ROWS * pitch;
                const int row_bytes = nblk * bb;
                for (; tile < tile_end; tile += gridDim.x) {
                    const uint8_t* rowp = lane < rows_mine ? slot_row_ptr(s, tile - s.tile0, rslot, row_bytes) : nullptr;
                    for (int c = 0; c < n_chunks; c++) {
                        const int nb = min(MT_CHUNK, nblk - c * MT_CHUNK);
                        const uint8_t* src = nullptr;
                        uint32_t sz = 0;
                        if (rowp) {
                            const uintptr_t lo = (uintptr_t)(rowp + (size_t)c * MT_CHUNK * bb), hi = lo + (size_t)nb * bb;
                            src = (const uint8_t*)(lo & ~(uintptr_t)15);
                            sz = (uint32_t)(((hi + 15) & ~(uintptr_t)15) - (uintptr_t)src);
                        }
                        if (fine && fine_k < STEP_FINE_STEPS) fine[8 * fine_k + 3] = (unsigned long long)clock64();
                        const uint32_t o = rp.place(step_bytes);
                        const int sl = rp.slot;
                        if (lane < TS_NSLOT) {
                            if (lane != sl && my_end > my_off && o < my_end && my_off < o + step_bytes) {
                                // a younger allocation these bytes overlap: its latest use (the parity before the next) must be released
                                mbar_wait(&empty[lane], ((rp.par >> lane) & 1u) ^ 1u);
                                my_end = my_off;
                            }
                            if (lane == sl) {
                                my_off = o;
                                my_end = o + step_bytes;
                            }
                        }
                        __syncwarp();
                        if (fine && fine_k < STEP_FINE_STEPS) fine[8 * fine_k + 4] = (unsigned long long)clock64();
                        // every copying lane announces its own bytes; the phase cannot complete before lane 0's arrival below
                        if (sz) {
                            mbar_expect_tx(&full[sl], sz);
                            bulk_g2s_hint(ring + o + rslot * pitch, src, sz, &full[sl], stream_policy);
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&full[sl]);
                        if (fine && fine_k < STEP_FINE_STEPS) fine[8 * fine_k++ + 5] = (unsigned long long)clock64();
                        rp.next();
                    }
                }
            }
        }
        return;
    }

    if (warp == W_EPI) {
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
                epilogue_tile<TP>(a, s, tile - s.tile0, pos, red + parity * NCW * TILE_ROWS, &part_full[parity], &part_free[parity],
                              (ph >> parity) & 1, lane, deferred_scale_of<ACT, TP>(a) ? rms_scale : nullptr);
                ph ^= 1 << parity;
            }
            // a fence orders only the calling thread's stores: every lane that wrote results fences, then one arrives
            // (system scope when the rows went into the other ranks' memory)
            if (TP && a.seg[0].n_peer > 0 && !a.seg[0].peer_local) __threadfence_system();
            else __threadfence();
            __syncwarp();
            if (lane == 0) {
                if (P.trace && P.fine_op == oi) P.trace[(size_t)P.n_ops * STEP_TRACE_SLOTS + 1024 + blockIdx.x] = globaltimer_ns();
                grid_arrive(P.sync_counter, (unsigned int)oi + 1u);
                stamp(P, oi, 3);
            }
        }
        return;
    }
    // ===== compute warps =====
    RingPos2 rpos;
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
        if (tid == 0 && P.trace && P.fine_op >= 0 && P.fine_op + 1 == oi) P.trace[(size_t)P.n_ops * STEP_TRACE_SLOTS + 2048 + blockIdx.x] = globaltimer_ns();
        if (tid == 0 && blockIdx.x == 0 && P.trace && P.fine_op == oi) P.trace[(size_t)P.n_ops * STEP_TRACE_SLOTS + 6] = (unsigned long long)clock64();
        const StepOp& op = *cs;
        if (op.kind == OP_GEMV) {
            const GemvArgs& a = op.g;
            const int K = a.K;
            sm.xq = (int8_t*)ss.xbase;
            sm.xd = (float*)(sm.xq + K);
            sm.bs16 = (int16_t*)(sm.xd + (ACT == ACT_Q8K ? K / 256 : K / 32));
            sm.bs32 = sm.bs16 + K / 16;
            if (TP && a.x_nparts > 0) xchg_wait(a, tid);   // tensor-parallel shard: the ranks' partial residual streams have arrived
            build_activation<ACT, TP, true>(a, sm, smem, tid, rank, csize, xpar, (P.trace && blockIdx.x == 0) ? P.trace + oi * STEP_TRACE_SLOTS : nullptr,
                                            rms_scale);
            if (TP && a.x_nparts > 0) xchg_materialise(a, tid);
            if (tid == 0) stamp(P, oi, 1);
            xpar ^= 1;
            const ActView xv{sm.xq, sm.xd, sm.bs16, sm.bs32};
            const int nblk = K >> 8;
            // this CTA's tiles, segment by segment (a segment has one weight type: the tile loop is specialised for it)
            TileWalk2 tw{ring, red, full, empty, part_full, part_free, rpos, parity, fph, 0, 0, 0,
                         (P.trace && P.fine_op == oi) ? P.trace + (size_t)P.n_ops * STEP_TRACE_SLOTS : nullptr, 0, P.fine_op == -2};
            const bool tracing = P.trace && blockIdx.x == 0 && tid == 0;
            int tile = blockIdx.x;
            for (int si = 0; si < a.n_seg; si++) {
                const GemvSeg& s = a.seg[si];
                const int tile_end = si + 1 < a.n_seg ? a.seg[si + 1].tile0 : a.n_tiles;
                if (tile >= tile_end) continue;
                if constexpr (ACT == ACT_Q8K) {
                    if (s.type == B2_TYPE_Q4_K) segment_tiles2<ACT, B2_TYPE_Q4_K>(s, tile, tile_end, nblk, xv, warp, lane, tw, tracing);
                    else if (s.type == B2_TYPE_Q6_K) segment_tiles2<ACT, B2_TYPE_Q6_K>(s, tile, tile_end, nblk, xv, warp, lane, tw, tracing);
                    else segment_tiles2<ACT, B2_TYPE_Q5_K>(s, tile, tile_end, nblk, xv, warp, lane, tw, tracing);
                } else {
                    segment_tiles2<ACT, B2_TYPE_Q8_0>(s, tile, tile_end, nblk, xv, warp, lane, tw, tracing);
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
            else if (TP && op.kind == OP_XCHG) {
                xchg_wait(op.g, tid);
                xchg_materialise(op.g, tid);
            }
            __threadfence();   // every thread's stores of this stage, before the CTA's single arrival
            named_bar_sync(1, NCT);
            if (tid == 0) {
                grid_arrive(P.sync_counter, (unsigned int)oi + 1u);
                stamp(P, oi, 3);
            }
        }
    }
}

size_t step2_smem_bytes(int act, int max_k) {
    // ring + barriers + RMS partial sums + the tile's partial dot products + the descriptors ...
    size_t n = TS_RING_BYTES + (2 * TS_NSLOT + 6) * sizeof(uint64_t) + NCW * sizeof(double) + 2 * NCW * TILE_ROWS * sizeof(float) + STEP2_EXTRA_SMEM;
    // ... + the quantised activation vector (attention aliases it: make sure it is large enough)
    size_t x_bytes = (size_t)max_k + (act == ACT_Q8K ? max_k / 256 : max_k / 32) * 4 + max_k / 16 * 2 + max_k / 32 * 2;
    if (x_bytes < ATTN_SMEM_BYTES) x_bytes = ATTN_SMEM_BYTES;
    n += x_bytes;
    return (n + 127) & ~(size_t)127;
}

using Step2Kernel = void (*)(const StepProgram);
Step2Kernel step2_entry(int act, bool tp) {
    if (tp) return act == ACT_Q8K ? step2_kernel<ACT_Q8K, true> : step2_kernel<ACT_Q80, true>;
    return act == ACT_Q8K ? step2_kernel<ACT_Q8K, false> : step2_kernel<ACT_Q80, false>;
}

int step2_config(int act, bool tp, int max_k, size_t* smem_out, int* cluster_out, int* grid_out) {
    const int sms = device_sm_count();
    if (sms <= 0) return B2_ECUDA;
    const size_t smem = step2_smem_bytes(act, max_k);
    if (smem > 227 * 1024) {
        set_error("step kernel: K=%d needs %zu bytes of shared memory", max_k, smem);
        return B2_ERANGE;
    }
    auto kern = step2_entry(act, tp);
    B2_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int ctas_per_sm = (int)((228 * 1024) / (smem + 1024));
    if (ctas_per_sm > 2) ctas_per_sm = 2;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    int cluster = 4;   // CTAs that share the work of quantising the activation vector (B2_STEP_CLUSTER: 1, 2, 4 or 8)
    if (const char* e = getenv("B2_STEP_CLUSTER")) {
        const int v = atoi(e);
        if (v == 1 || v == 2 || v == 4 || v == 8) cluster = v;
    }
    cudaLaunchConfig_t q{};
    q.gridDim = dim3(sms * ctas_per_sm / cluster * cluster);
    q.blockDim = dim3(STEP2_THREADS);
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
    const int cap = sms * ctas_per_sm / cluster;
    if (n > cap) n = cap;
    *smem_out = smem;
    *cluster_out = cluster;
    *grid_out = n * cluster;
    return 0;
}

}  // namespace

// every mat-vec K of a program must be a whole number of 256-element blocks
bool step2_supports(int act, int K) {
    (void)act;
    return K > 0 && (K & 255) == 0;
}

int step2_grid_size(int act, int max_k, bool tp) {
    size_t smem;
    int cluster, grid;
    const int rc = step2_config(act, tp, max_k, &smem, &cluster, &grid);
    return rc ? rc : grid;
}

int launch_step2(const StepProgram& p, int act, cudaStream_t st) {
    size_t smem;
    int cluster, grid;
    int rc = step2_config(act, p.tp > 0, p.max_k, &smem, &cluster, &grid);
    if (rc) return rc;
    B2_CUDA_CHECK(cudaMemsetAsync(p.sync_counter, 0, STEP_SYNC_BYTES, st));
    auto kern = step2_entry(act, p.tp > 0);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(STEP2_THREADS);
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
