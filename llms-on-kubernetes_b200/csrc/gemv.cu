// gemv.cu -- dequant-fused mat-vec for decode (8a row a12 "quantised matmul", with the
// neighbouring rows a11 RMSNorm, a13 RoPE + KV write, a15 SwiGLU / residual / bias fused in).
//
// Arithmetic follows ggml's CPU mul_mat for one token: the fp32 activation row is quantised to Q8_K
// (per 256: d = 1/iscale, iscale = -127/max, q = nearest_int(iscale*x), 16-element bsums) for K-quant
// weights or to Q8_0 (per 32: d = amax/127 rounded to fp16, q = roundf(x/d)) for Q8_0 weights, and the
// dot products are integer (dp4a) with one fp32 scale per (super-)block.  See oracle/ggml_ref.c for
// the CPU restatement these kernels are tested against.
//
// The kernel is HBM-bound byte work, and at B200 bandwidth it is the *instruction issue rate* that
// threatens the roofline (11.6e12 4.5-bit weights per second), so the design is about instructions
// per weight as much as about bytes:
//   * weights stay in HBM exactly as in the GGUF.  A producer warp streams row segments into a
//     3-stage shared-memory ring with 1-D bulk async copies (cp.async.bulk, SASS UBLKCP) completing
//     on mbarriers; it never touches activations, so under programmatic dependent launch it fills the
//     ring while the previous kernel in the stream is still running.  Because the copy engine moves
//     raw bytes, the 2-byte-aligned 210-byte Q6_K / 34-byte Q8_0 blocks need no repacking.
//   * a tile is 8 weight rows x one K-chunk (16 super-blocks).  Each of the 8 consumer warps takes
//     two super-blocks of the chunk for all 8 rows: 4 lanes per (row, super-block), 64 weights per
//     lane, so the 6-bit scale unpack, the fp16 scale converts and the activation loads are amortised
//     over 64 weights, and the activation bytes are broadcast-read once for 8 rows.  High nibbles are
//     multiplied in place (dp4a.u32.s32 on q & 0xF0F0F0F0, exact shift by 4 afterwards).
//   * the quantised activation vector is built ONCE PER THREAD-BLOCK CLUSTER, not once per CTA: the
//     8 CTAs of a cluster each quantise 1/8 of the 256-element spans and store the result into every
//     peer's shared memory through DSMEM, signalling per-span on the peers' mbarriers.  (Done per CTA
//     this prologue was 39% of all issued instructions.)
//   * a decoder warp unpacks the 6-bit scale/min headers of a freshly landed stage once (128 headers)
//     into ready-to-use records, instead of every lane of every compute warp doing it per quarter.
//   * per tile the compute warps' partial sums are handed through shared memory + mbarriers to an
//     epilogue warp that adds them in a fixed order and runs the fused epilogue on row pairs (RoPE
//     pairs, gate/up pairs); compute warps never meet at a CTA barrier and never wait on a global load.
#include <cstdlib>

#include "gemv_device.cuh"

namespace b2 {

using namespace gemvdev;

namespace {

// warp roles
constexpr int W_PRODUCER = NCW, W_DECODER = NCW + 1, W_EPILOGUE = NCW + 2;

template <int ACT>
__global__ void __launch_bounds__(GEMV_THREADS, 2) gemv_kernel(const __grid_constant__ GemvArgs a) {
    using G = Geo<ACT>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem;
    uint8_t* dec = smem + GEMV_STAGES * G::STAGE_BYTES;
    uint64_t* full = (uint64_t*)(dec + GEMV_STAGES * G::DEC_BYTES);   // data landed            (producer -> decoder)
    uint64_t* ready = full + GEMV_STAGES;                             // headers decoded        (decoder -> compute)
    uint64_t* empty = ready + GEMV_STAGES;                            // stage consumed         (compute -> producer)
    uint64_t* part_full = empty + GEMV_STAGES;                        // [2] partial sums written (compute -> epilogue)
    uint64_t* part_free = part_full + 2;                              // [2] partial sums read    (epilogue -> compute)
    ActSmem sm;
    sm.xbar = part_free + 2;   // 14 barriers so far; two pad slots keep the 16-byte alignment below
    sm.ss_warp = (double*)(sm.xbar + 3);
    float* red = (float*)(sm.ss_warp + NCW);   // [2 parities][NCW warps][TILE_ROWS]
    const int K = a.K;
    sm.xq = (int8_t*)(red + 2 * NCW * TILE_ROWS);
    sm.xd = (float*)(sm.xq + K);
    sm.bs16 = (int16_t*)(sm.xd + (ACT == ACT_Q8K ? K / 256 : K / 32));
    sm.bs32 = sm.bs16 + K / 16;

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
        mbar_init(sm.xbar, 1);    // one local arrival announcing the bytes the cluster will store (build_activation)
        mbar_fence_init();
    }
    // peers store into this CTA's shared memory and arrive on its barriers: they must be initialised first
    cluster_sync_all();
    // let the next kernel in the stream begin its own weight prefetch as early as possible
    pdl_launch_dependents();

    const int nblk_row = K / (ACT == ACT_Q8K ? 256 : 32);
    const int n_chunks = (nblk_row + G::CHUNK_BLOCKS - 1) / G::CHUNK_BLOCKS;

    if (warp == W_PRODUCER) {
        // ===== producer: weights only, independent of the previous kernel's output =====
        int stage = 0, phase = 0, si = 0;
        const uint64_t stream_policy = l2_policy_evict_first();
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
                if (sz) bulk_g2s_hint(ring + stage * G::STAGE_BYTES + lane * G::ROW_SLOT, src, sz, &full[stage], stream_policy);
                if (++stage == GEMV_STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    if (warp == W_DECODER) {
        // ===== decoder: unpack the Q4_K / Q5_K scale headers of each landed stage once =====
        int stage = 0, phase = 0, si = 0;
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
            while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
            const int type = a.seg[si].type;
            const int bb = block_bytes_of(type);
            for (int c = 0; c < n_chunks; c++) {
                const int nb = min(G::CHUNK_BLOCKS, nblk_row - c * G::CHUNK_BLOCKS);
                mbar_wait(&full[stage], phase);
                if (ACT == ACT_Q8K && (type == B2_TYPE_Q4_K || type == B2_TYPE_Q5_K)) {
                    // Q4_K / Q5_K rows are 16-byte aligned in their slots
                    for (int i = lane; i < TILE_ROWS * nb; i += 32) {
                        const int rr = i / nb, b = i - rr * nb;
                        uint8_t* rec = dec + stage * G::DEC_BYTES + rr * G::DEC_ROW + b * 32;   // {d, dmin, pad, scale words}
                        decode_k4_header(ring + stage * G::STAGE_BYTES + rr * G::ROW_SLOT + b * bb, rec, rec + 16);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&ready[stage]);
                if (++stage == GEMV_STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    if (warp == W_EPILOGUE) {
        // ===== epilogue: add the compute warps' shares in a fixed order, finish row pairs =====
        pdl_wait();   // the epilogue reads and writes activations of the previous kernel
        const int pos = a.pos_ptr ? *a.pos_ptr : 0;
        int parity = 0, ph = 0, si = 0;   // ph: bit p = phase of part_full[p]
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, parity ^= 1) {
            while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
            const GemvSeg& s = a.seg[si];
            epilogue_tile(a, s, tile - s.tile0, pos, red + parity * NCW * TILE_ROWS, &part_full[parity], &part_free[parity],
                          (ph >> parity) & 1, lane);
            ph ^= 1 << parity;
        }
        return;
    }

    // ===== compute warps =====
    pdl_wait();  // the activation vector is written by the previous kernel
    build_activation<ACT>(a, sm, smem, tid, rank, csize);
    const ActView xv{sm.xq, sm.xd, sm.bs16, sm.bs32};

    const int r = lane >> 2;   // this lane's row slot in the tile
    int stage = 0, phase = 0, parity = 0, fph = 3, si = 0;   // fph: bit p = phase of part_free[p]
    for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, parity ^= 1) {
        while (si + 1 < a.n_seg && tile >= a.seg[si + 1].tile0) si++;
        const GemvSeg& s = a.seg[si];
        const int type = s.type;
        const int bb = block_bytes_of(type);
        // row segment offset inside its ring slot (source address mod 16): zero for the 16-byte-multiple
        // Q4_K / Q5_K blocks, row dependent for Q6_K / Q8_0.  Rows that do not exist (last tile of a
        // segment) are computed on whatever the slot holds and ignored by the epilogue.
        uint32_t off0 = 0;
        if (ACT != ACT_Q8K || type == B2_TYPE_Q6_K) off0 = (uint32_t)((uintptr_t)slot_row_ptr(s, tile - s.tile0, r, nblk_row * bb) & 15);
        const uint32_t chunk_step = (uint32_t)(G::CHUNK_BLOCKS * bb) & 15;   // alignment drift per chunk
        float acc = 0.0f;
        for (int c = 0; c < n_chunks; c++) {
            const int nb = min(G::CHUNK_BLOCKS, nblk_row - c * G::CHUNK_BLOCKS);
            mbar_wait(&ready[stage], phase);
            const uint8_t* row = ring + stage * G::STAGE_BYTES + r * G::ROW_SLOT + ((off0 + c * chunk_step) & 15);
            const uint8_t* drow = dec + stage * G::DEC_BYTES + r * G::DEC_ROW;
            acc += dot_chunk<ACT>(type, row, DecPtr{drow, drow + 16}, nb, c * G::CHUNK_BLOCKS, xv, warp, lane);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (++stage == GEMV_STAGES) { stage = 0; phase ^= 1; }
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        // hand the 8 row sums of this warp's K share to the epilogue warp
        mbar_wait(&part_free[parity], (fph >> parity) & 1);
        fph ^= 1 << parity;
        if ((lane & 3) == 0) red[(parity * NCW + warp) * TILE_ROWS + r] = acc;
        __syncwarp();
        if (lane == 0) mbar_arrive(&part_full[parity]);
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
    const size_t decb = act == ACT_Q8K ? Geo<ACT_Q8K>::DEC_BYTES : Geo<ACT_Q80>::DEC_BYTES;
    size_t n = GEMV_STAGES * (stage + decb) + (3 * GEMV_STAGES + 7) * sizeof(uint64_t) + NCW * sizeof(double) +
               2 * NCW * TILE_ROWS * sizeof(float);
    n += (size_t)K;                                            // int8 activations
    n += (act == ACT_Q8K ? K / 256 : K / 32) * sizeof(float);  // block scales
    n += (size_t)K / 16 * sizeof(int16_t) + (size_t)K / 32 * sizeof(int16_t);
    return (n + 127) & ~(size_t)127;
}

int gemv_finalize(GemvArgs& a, int act) {
    if (a.n_seg < 1 || a.n_seg > 3) {
        set_error("gemv: n_seg=%d out of range", a.n_seg);
        return B2_EINVAL;
    }
    const int kmult = act == ACT_Q8K ? 256 : 32;
    if (a.K <= 0 || a.K % kmult || a.K % 8 || a.K > MAX_SPANS * 256) {
        set_error("gemv: K=%d must be a multiple of %d and at most %d", a.K, kmult, MAX_SPANS * 256);
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


int launch_gemv(const GemvArgs& a, int act, cudaStream_t st, bool pdl) {
    const int g_num_sms = device_sm_count();
    if (g_num_sms <= 0) return B2_ECUDA;
    const int dev = current_device();
    const size_t smem = gemv_smem_bytes(act, a.K);
    if (smem > 227 * 1024) {
        set_error("gemv: K=%d needs %zu bytes of shared memory (limit 227 KiB)", a.K, smem);
        return B2_ERANGE;
    }
    auto kern = act == ACT_Q8K ? gemv_kernel<ACT_Q8K> : gemv_kernel<ACT_Q80>;
    static PerDeviceSize configured[2];   // per activation format and device
    if (configured[act].below(dev, smem)) {
        B2_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[act].set(dev, smem);
    }
    int ctas_per_sm = (int)((227 * 1024) / (smem + 1024));
    if (ctas_per_sm > 2) ctas_per_sm = 2;
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    // The activation prologue is shared across a thread-block cluster.  A cluster must sit inside one
    // GPC, so fewer CTAs are co-resident than SMs x CTAs/SM; a grid larger than that runs a nearly
    // empty second wave (measured: SMs idle half the kernel).  Ask the driver how many clusters fit.
    static int g_cluster_size = 0;
    if (!g_cluster_size) {
        const char* e = getenv("B2_GEMV_CLUSTER");
        g_cluster_size = e ? atoi(e) : 8;
        if (g_cluster_size != 1 && g_cluster_size != 2 && g_cluster_size != 4 && g_cluster_size != 8) g_cluster_size = 8;
    }
    int cluster = g_cluster_size;
    int want = g_num_sms * ctas_per_sm;
    if (want > a.n_tiles) want = a.n_tiles;
    while (cluster > 1 && want < cluster) cluster >>= 1;
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(GEMV_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int na = 0;
    // always a cluster launch, even of one CTA: the activation exchange uses st.async, which is a cluster instruction
    // (measured: cudaErrorIllegalInstruction in a launch without the attribute)
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = cluster;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    na++;
    int max_clusters = want / cluster;
    if (cluster > 1) {
        // cached per (activation format, cluster size, shared-memory size)
        // (every GPU of a box is the same part: the cache is not keyed by device, the attribute above is)
        static int cached[2][9] = {};
        static size_t cached_smem[2][9] = {};
        if (!cached[act][cluster] || cached_smem[act][cluster] != smem) {
            cudaLaunchConfig_t q = cfg;
            q.gridDim = dim3(g_num_sms * ctas_per_sm / cluster * cluster);
            q.attrs = attr;
            q.numAttrs = na;
            int n = 0;
            B2_CUDA_CHECK(cudaOccupancyMaxActiveClusters(&n, kern, &q));
            if (n < 1) n = 1;
            cached[act][cluster] = n;
            cached_smem[act][cluster] = smem;
        }
        if (max_clusters > cached[act][cluster]) max_clusters = cached[act][cluster];
    }
    if (max_clusters < 1) max_clusters = 1;
    const int grid = max_clusters * cluster;
    cfg.gridDim = dim3(grid);
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    B2_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, a));
    return 0;
}

}  // namespace b2
