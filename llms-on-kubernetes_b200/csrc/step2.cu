// step2.cu -- the persistent per-token decode kernel, second design (SURVEY 8a rows a10-a16 in one launch).
//
// Same program of stages, grid barrier, attention / embedding / sampling stages and epilogues as step.cu; what changed
// is how the weight stream reaches the arithmetic (matvec_team.cuh) and that the stream no longer stops at a hand-off:
//   * ring allocation = one STEP: R whole weight rows of a tile (8 rows, or fewer when rows are long), copied by the
//     producer warp with one bulk copy per row; up to 8 steps in flight in an 81 KB byte ring;
//   * the 8 compute warps are T teams of WPT K-slices (matvec_team.cuh); the lane's activations sit in registers for
//     the whole stage, the 6-bit scales are unpacked in the lane: no decoder warp, one barrier per step;
//   * the warp that used to decode headers now runs AHEAD of the producer through the same sequence of rows and asks
//     L2 to fetch them (cp.async.bulk.prefetch.L2), a fixed number of steps ahead: while the compute warps sit in a
//     grid barrier and the ring is full, HBM keeps streaming into L2, and the tile phase that follows is fed from L2.
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

constexpr int W_PREFETCH = W_DECODER;   // the 10th warp: L2 look-ahead

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

// per-segment geometry every role derives the same way
struct SegGeo {
    int bb, row_bytes, R, steps, Te, RW;
    uint32_t pitch, step_bytes;
};
template <int ACT>
__device__ __forceinline__ SegGeo seg_geo(int type, int K, const TeamGeo& g) {
    SegGeo s;
    s.bb = block_bytes_of(type);
    s.row_bytes = (ACT == ACT_Q8K ? (K >> 8) : (K >> 5)) * s.bb;
    s.pitch = ts_pitch(s.row_bytes);
    s.R = ts_rows_per_step(s.pitch);
    s.steps = TILE_ROWS / s.R;
    s.Te = g.T < s.R ? g.T : s.R;
    s.RW = s.R / s.Te;
    s.step_bytes = (uint32_t)s.R * s.pitch;
    return s;
}

struct TileWalk2 {
    uint8_t* ring;
    float* red;
    uint64_t *full, *empty, *part_full, *part_free;
    RingPos2 rp;
    int parity, fph;
    long long cyc_ready, cyc_dot, cyc_free;   // B2_STEP_TRACE
};

// the lane's activation registers and the layout class they were loaded for (-1: stale, a new stage began)
struct XState {
    XR x;
    int cls;
};

// The tiles of one segment that belong to this CTA, for a weight type known at compile time.
template <int ACT, int TYPE>
__device__ __forceinline__ void segment_tiles2(const GemvSeg& s, int& tile, int tile_end, int K, const TeamGeo& g, const ActView& xa, int warp,
                                               int lane, TileWalk2& w, XState& xs, bool tracing) {
    constexpr int CLS = XClass<TYPE>::value;
    constexpr bool unaligned = TYPE == B2_TYPE_Q6_K || TYPE == B2_TYPE_Q8_0;
    const SegGeo sg = seg_geo<ACT>(TYPE, K, g);
    const int team = warp / g.WPT, sl0 = warp - team * g.WPT;
    const bool active = team < sg.Te;
    const int b = lane >> 2, c = lane & 3;
    const int kb0 = sl0 * 8 + b;
    if (g.nsl == 1 && xs.cls != CLS) {
        if (active && kb0 < g.nblk) load_x<CLS>(xs.x, xa, kb0, c);
        xs.cls = CLS;
    }
    const int step = (int)gridDim.x;
    for (; tile < tile_end; tile += step) {
        const int ptile = tile - s.tile0;
        // the partial-sum buffer of this parity must have been read by the epilogue warp (two tiles ago)
        const long long c3 = tracing ? clock64() : 0;
        mbar_wait(&w.part_free[w.parity], (w.fph >> w.parity) & 1);
        if (tracing) w.cyc_free += clock64() - c3;
        w.fph ^= 1 << w.parity;
        float* red = w.red + (w.parity * NCW + warp) * TILE_ROWS;
        for (int st = 0; st < sg.steps; st++) {
            const uint32_t o = w.rp.place(sg.step_bytes);
            const int sl = w.rp.slot;
            const long long c0 = tracing ? clock64() : 0;
            mbar_wait(&w.full[sl], w.rp.parity());
            const long long c1 = tracing ? clock64() : 0;
            float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            if (active) {
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    if (i < sg.RW) {
                        const int ris = team + sg.Te * i;     // row inside the step
                        uint32_t off = 0;
                        if (unaligned) off = (uint32_t)((uintptr_t)slot_row_ptr(s, ptile, st * sg.R + ris, sg.row_bytes) & 15);
                        const uint8_t* row = w.ring + o + ris * sg.pitch + off;
                        if (g.nsl == 1) {
                            if (kb0 < g.nblk) acc[i] += row_quarter<TYPE>(row, kb0, c, xs.x);
                        } else {
                            for (int q = 0; q < g.nsl; q++) {
                                const int kb = kb0 + q * NCW * 8;
                                if (kb < g.nblk) {
                                    load_x<CLS>(xs.x, xa, kb, c);
                                    acc[i] += row_quarter<TYPE>(row, kb, c, xs.x);
                                }
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&w.empty[sl]);
            w.rp.next();
            if (active) {
                // row sums of this warp's slice: two rows per reduction
                if (sg.RW == 1) {
                    const float v = warp_sum(acc[0]);
                    if (lane == 0) red[st * sg.R + team] = v;
                } else {
                    const float v = warp_sum_pair(acc[0], acc[1], lane);
                    if ((lane & 15) == 0) red[st * sg.R + team + sg.Te * (lane >> 4)] = v;
                    if (sg.RW == 4) {
                        const float u = warp_sum_pair(acc[2], acc[3], lane);
                        if ((lane & 15) == 0) red[st * sg.R + team + sg.Te * (2 + (lane >> 4))] = u;
                    }
                }
            }
            if (tracing) {
                const long long c2 = clock64();
                w.cyc_ready += c1 - c0;
                w.cyc_dot += c2 - c1 + (__float_as_uint(acc[0]) == 0x7fc0dead ? 1 : 0);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&w.part_full[w.parity]);
        w.parity ^= 1;
    }
}

// epilogue warp, one tile: as gemvdev::epilogue_tile, but row r's shares come from the WPT warps of its team only
__device__ __forceinline__ void epilogue_tile2(const GemvArgs& a, const GemvSeg& s, int ptile, int pos, const float* rp, uint64_t* part_full,
                                               uint64_t* part_free, uint32_t full_phase, int lane, int R, int Te, int WPT) {
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
        const int w0 = ((lane & (R - 1)) % Te) * WPT;
        for (int w = 0; w < WPT; w++) v += rp[(w0 + w) * TILE_ROWS + lane];
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(part_free);
    const float vo = __shfl_xor_sync(0xffffffffu, v, 1);
    if (owner) epilogue_apply(s, p, ra, rb, v, vo, o, pos);
}

constexpr size_t STEP2_EXTRA_SMEM = 3 * sizeof(GemvArgs) + 2 * sizeof(StepOp) + 32 * sizeof(int) + 16;

template <int ACT>
__global__ void __launch_bounds__(GEMV_THREADS, 2) step2_kernel(const __grid_constant__ StepProgram P) {
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
    // aux[28]: steps the producer has issued (read by the look-ahead warp)
    volatile int* prod_steps = (volatile int*)&ss.aux[28];

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
        ss.aux[28] = 0;
        mbar_fence_init();
    }
    cluster_sync_all();
    volatile unsigned int* cta_stage = (volatile unsigned int*)&ss.aux[2];

    if (warp == W_PRODUCER || warp == W_PREFETCH) {
        // ===== producer: streams every mat-vec stage's weights in program order, never waits for a stage to start.
        // ===== look-ahead: the same walk, P.l2_ahead steps ahead of the producer, fetching into L2 only.
        const bool is_prod = warp == W_PRODUCER;
        if (!is_prod && P.l2_ahead <= 0) return;
        RingPos2 rp;
        uint32_t my_off = 0, my_end = 0;   // lane j < TS_NSLOT: the latest allocation placed in slot j
        int n_issued = 0;
        const uint64_t stream_policy = l2_policy_evict_first();
        GemvArgs* slot = &ss.slots[is_prod ? 0 : 1];
        for (int oi = 0; oi < P.n_ops; oi++) {
            const StepOp& op = P.ops[oi];
            if (op.kind != OP_GEMV) continue;
            __syncwarp();
            stage_words(slot, &op.g, sizeof(GemvArgs) / 4, lane);
            const GemvArgs& a = *slot;
            const TeamGeo g = team_geo(a.K);
            int si = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
                while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
                const GemvSeg& s = a.seg[si];
                const SegGeo sg = seg_geo<ACT>(s.type, a.K, g);
                for (int st = 0; st < sg.steps; st++) {
                    const uint8_t* src = nullptr;
                    uint32_t sz = 0;
                    if (lane < sg.R) {
                        const uint8_t* rowp = slot_row_ptr(s, tile - s.tile0, st * sg.R + lane, sg.row_bytes);
                        if (rowp) {
                            const uintptr_t lo = (uintptr_t)rowp, hi = lo + (size_t)sg.row_bytes;
                            src = (const uint8_t*)(lo & ~(uintptr_t)15);
                            sz = (uint32_t)(((hi + 15) & ~(uintptr_t)15) - (uintptr_t)src);
                        }
                    }
                    if (!is_prod) {
                        // stay at most l2_ahead steps in front of the producer
                        if (lane == 0) {
#pragma unroll 1
                            for (uint32_t spin = 0; n_issued >= *prod_steps + P.l2_ahead; spin++) {
                                __nanosleep(64);
                                if (spin > (1u << 26)) asm volatile("trap;");
                            }
                        }
                        __syncwarp();
                        if (sz) bulk_prefetch_l2(src, sz);
                        n_issued++;
                        continue;
                    }
                    uint32_t total = sz;
#pragma unroll
                    for (int o = 4; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
                    const uint32_t o = rp.place(sg.step_bytes);
                    const int sl = rp.slot;
                    const uint32_t par = rp.parity();
                    if (lane < TS_NSLOT) {
                        if (lane == sl) {
                            mbar_wait(&empty[sl], par ^ 1);   // the slot's previous use is released ...
                            my_off = o;
                            my_end = o + sg.step_bytes;
                        } else if (my_end > my_off && o < my_end && my_off < o + sg.step_bytes) {
                            // ... and so is every younger allocation these bytes overlap (its latest use: the parity before the next)
                            mbar_wait(&empty[lane], ((rp.par >> lane) & 1u) ^ 1u);
                            my_end = my_off;
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive_expect_tx(&full[sl], total);
                    __syncwarp();
                    if (sz) bulk_g2s_hint(ring + o + lane * sg.pitch, src, sz, &full[sl], stream_policy);
                    rp.next();
                    n_issued++;
                    if (lane == 0) *prod_steps = n_issued;
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
            const TeamGeo g = team_geo(a.K);
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
                const SegGeo sg = seg_geo<ACT>(s.type, a.K, g);
                epilogue_tile2(a, s, tile - s.tile0, pos, red + parity * NCW * TILE_ROWS, &part_full[parity], &part_free[parity],
                               (ph >> parity) & 1, lane, sg.R, sg.Te, g.WPT);
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
    RingPos2 rpos;
    int parity = 0, fph = 3;
    uint32_t xpar = 0;
    XState xs;
    xs.cls = -1;
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
            const TeamGeo g = team_geo(K);
            xs.cls = -1;   // new activations: the registers are stale
            // this CTA's tiles, segment by segment (a segment has one weight type: the tile loop is specialised for it)
            TileWalk2 tw{ring, red, full, empty, part_full, part_free, rpos, parity, fph, 0, 0, 0};
            const bool tracing = P.trace && blockIdx.x == 0 && tid == 0;
            int tile = blockIdx.x;
            for (int si = 0; si < a.n_seg; si++) {
                const GemvSeg& s = a.seg[si];
                const int tile_end = si + 1 < a.n_seg ? a.seg[si + 1].tile0 : a.n_tiles;
                if (tile >= tile_end) continue;
                if constexpr (ACT == ACT_Q8K) {
                    if (s.type == B2_TYPE_Q4_K) segment_tiles2<ACT, B2_TYPE_Q4_K>(s, tile, tile_end, K, g, xv, warp, lane, tw, xs, tracing);
                    else if (s.type == B2_TYPE_Q6_K) segment_tiles2<ACT, B2_TYPE_Q6_K>(s, tile, tile_end, K, g, xv, warp, lane, tw, xs, tracing);
                    else segment_tiles2<ACT, B2_TYPE_Q5_K>(s, tile, tile_end, K, g, xv, warp, lane, tw, xs, tracing);
                } else {
                    segment_tiles2<ACT, B2_TYPE_Q8_0>(s, tile, tile_end, K, g, xv, warp, lane, tw, xs, tracing);
                }
            }
            rpos = tw.rp; parity = tw.parity; fph = tw.fph;
            if (tracing) {
                // cycles warp 0 of CTA 0 spent waiting for data / in the dot products / waiting for the epilogue warp (marks 4..6)
                unsigned long long* tr = P.trace + oi * STEP_TRACE_SLOTS;
                tr[4] = tr[0] + (unsigned long long)(tw.cyc_ready * 1000 / 1965);
                tr[5] = tr[0] + (unsigned long long)(tw.cyc_dot * 1000 / 1965);
                tr[6] = tr[0] + (unsigned long long)(tw.cyc_free * 1000 / 1965);
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

size_t step2_smem_bytes(int act, int max_k) {
    // ring + barriers + RMS partial sums + the tile's partial dot products + the descriptors ...
    size_t n = TS_RING_BYTES + (2 * TS_NSLOT + 6) * sizeof(uint64_t) + NCW * sizeof(double) + 2 * NCW * TILE_ROWS * sizeof(float) + STEP2_EXTRA_SMEM;
    // ... + the quantised activation vector (attention aliases it: make sure it is large enough)
    size_t x_bytes = (size_t)max_k + (act == ACT_Q8K ? max_k / 256 : max_k / 32) * 4 + max_k / 16 * 2 + max_k / 32 * 2;
    if (x_bytes < ATTN_SMEM_BYTES) x_bytes = ATTN_SMEM_BYTES;
    n += x_bytes;
    return (n + 127) & ~(size_t)127;
}

int step2_config(int act, int max_k, size_t* smem_out, int* cluster_out, int* grid_out) {
    int dev = 0, sms = 0;
    B2_CUDA_CHECK(cudaGetDevice(&dev));
    B2_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const size_t smem = step2_smem_bytes(act, max_k);
    if (smem > 227 * 1024) {
        set_error("step kernel: K=%d needs %zu bytes of shared memory", max_k, smem);
        return B2_ERANGE;
    }
    auto kern = act == ACT_Q8K ? step2_kernel<ACT_Q8K> : step2_kernel<ACT_Q80>;
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
    const int cap = sms * ctas_per_sm / cluster;
    if (n > cap) n = cap;
    *smem_out = smem;
    *cluster_out = cluster;
    *grid_out = n * cluster;
    return 0;
}

}  // namespace

// every mat-vec K of a program must be a whole number of 256-element blocks whose longest row fits one ring step
bool step2_supports(int act, int K) {
    if (K <= 0 || (K & 255)) return false;
    const int row_bytes = act == ACT_Q8K ? (K >> 8) * 210 : (K >> 5) * 34;   // the widest type of the format
    return ts_pitch(row_bytes) <= TS_STEP_CAP;
}

int step2_grid_size(int act, int max_k) {
    size_t smem;
    int cluster, grid;
    const int rc = step2_config(act, max_k, &smem, &cluster, &grid);
    return rc ? rc : grid;
}

int launch_step2(const StepProgram& p, int act, cudaStream_t st) {
    size_t smem;
    int cluster, grid;
    int rc = step2_config(act, p.max_k, &smem, &cluster, &grid);
    if (rc) return rc;
    B2_CUDA_CHECK(cudaMemsetAsync(p.sync_counter, 0, STEP_SYNC_BYTES, st));
    auto kern = act == ACT_Q8K ? step2_kernel<ACT_Q8K> : step2_kernel<ACT_Q80>;
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
