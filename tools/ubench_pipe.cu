// ubench_pipe.cu -- the decode mat-vec's data path rebuilt level by level, to see which part costs bandwidth.
// 2 CTAs per SM, two producer warps (bulk copies of 2304-byte row chunks into a ring of 8-row steps), 8 consumer warps,
// one epilogue warp; the "matrix" is a 4 GiB buffer of Q4_K-shaped rows (K = 4096), so every step is also a tile.
//   level 0: consumers wait for a step and release it
//   level 1: + ldmatrix of the warp's two blocks and the block header (no arithmetic)
//   level 2: + the 8 mma.sync and the high-nibble masks
//   level 3: + the scale arithmetic (q4k_rows8_blocks2_mma complete, activations in shared memory)
//   level 4: + the row sums handed to the epilogue warp per tile (shuffles, part_free / part_full barriers)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I llms-on-kubernetes_b200/csrc -o tools/ubench_pipe tools/ubench_pipe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "matvec_team.cuh"

using namespace b2;
using namespace b2::gemvdev;
using namespace b2::teamdev;

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int NSLOT = 4;
constexpr uint32_t PITCH = 16 * 144 + 16, STEP = 8 * PITCH, ROW = 16 * 144;

template <int LEVEL, int PROD>
__global__ void __launch_bounds__(352, 2) pipe(const uint8_t* w, size_t n_tiles, float* out, size_t wrap_tiles) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* ring = smem;
    uint64_t* full = (uint64_t*)(smem + NSLOT * STEP);
    uint64_t* empty = full + NSLOT;
    uint64_t* part_full = empty + NSLOT;
    uint64_t* part_free = part_full + 2;
    float* red = (float*)(part_free + 2);
    int8_t* xq = (int8_t*)(red + 2 * 8 * 8);
    float* xd = (float*)(xq + 4096);
    int16_t* bs32 = (int16_t*)(xd + 16);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < NSLOT; s++) { mbar_init(&full[s], PROD == 0 ? 2 : 1); mbar_init(&empty[s], 8); }
        mbar_init(&part_full[0], 8); mbar_init(&part_full[1], 8);
        mbar_init(&part_free[0], 1); mbar_init(&part_free[1], 1);
        mbar_fence_init();
    }
    for (int i = tid; i < 4096; i += blockDim.x) xq[i] = (int8_t)(i * 7 + 3);
    for (int i = tid; i < 16; i += blockDim.x) xd[i] = 0.01f;
    for (int i = tid; i < 128; i += blockDim.x) bs32[i] = (int16_t)(i - 60);
    __syncthreads();
    if (warp == 8 || warp == 9) {
        const int half = warp - 8;
        if (PROD != 0 && half == 1) return;
        uint64_t pol = l2_policy_evict_first();
        if (wrap_tiles < n_tiles) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
        int s = 0;
        uint32_t ph = 0;
        size_t tw = blockIdx.x;
        for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x, tw += gridDim.x) {
            if (tw >= wrap_tiles) tw -= wrap_tiles;
            if (lane == 0) mbar_wait(&empty[s], ph ^ 1);
            __syncwarp();
            if (PROD == 0) {
                if (lane < 4) {
                    const int r = 4 * half + lane;
                    mbar_expect_tx(&full[s], ROW);
                    bulk_g2s_hint(ring + s * STEP + r * PITCH, w + (tw * 8 + r) * (size_t)ROW, ROW, &full[s], pol);
                }
            } else if (PROD == 1) {
                if (lane < 8) {
                    mbar_expect_tx(&full[s], ROW);
                    bulk_g2s_hint(ring + s * STEP + lane * PITCH, w + (tw * 8 + lane) * (size_t)ROW, ROW, &full[s], pol);
                }
            } else if (PROD == 2) {
                if (lane == 0) {
                    mbar_expect_tx(&full[s], 8 * ROW);
#pragma unroll
                    for (int r = 0; r < 8; r++) bulk_g2s_hint(ring + s * STEP + r * PITCH, w + (tw * 8 + r) * (size_t)ROW, ROW, &full[s], pol);
                }
            } else {
                // one copy of the tile's 8 contiguous rows (no padding between the rows: ldmatrix bank conflicts)
                if (lane == 0) {
                    mbar_expect_tx(&full[s], 8 * ROW);
                    bulk_g2s_hint(ring + s * STEP, w + (tw * 8) * (size_t)ROW, 8 * ROW, &full[s], pol);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[s]);
            if (++s == NSLOT) { s = 0; ph ^= 1; }
        }
        return;
    }
    if (warp == 10) {
        if (LEVEL < 4) return;
        int parity = 0, ph = 0;
        for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x, parity ^= 1) {
            mbar_wait(&part_full[parity], (ph >> parity) & 1);
            float v = 0.0f;
            if (lane < 8)
                for (int k = 0; k < 8; k++) v += red[(parity * 8 + k) * 8 + lane];
            __syncwarp();
            if (lane == 0) mbar_arrive(&part_free[parity]);
            ph ^= 1 << parity;
            if (lane < 8) out[t * 8 + lane] = v;
        }
        return;
    }
    // consumers
    const int g = lane >> 2, t4 = lane & 3;
    const MmaLane ml = mma_lane(lane);
    const uint32_t xq_s = smem_u32(xq), bs32_s = smem_u32(bs32), xd_s = smem_u32(xd);
    const uint32_t lm_off = (uint32_t)((lane & 7) * PITCH + (2 * warp + ((lane >> 3) & 1)) * 144 + 16 + 16 * (lane >> 4));
    const uint32_t hdr_off = (uint32_t)(g * PITCH + (2 * warp + (t4 & 1)) * 144);
    int s = 0, parity = 0, fph = 3;
    uint32_t ph = 0;
    float sink = 0.0f;
    for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        mbar_wait(&full[s], ph);
        const uint8_t* base = ring + s * STEP;
        float acc = 0.0f;
        if (LEVEL == 1) {
            uint32_t a[4], x = 0;
            for (int j = 0; j < 4; j++) { ldmatrix_x4(smem_u32(base) + lm_off + 32 * j, a); x ^= a[0] ^ a[1] ^ a[2] ^ a[3]; }
            const uint4 hd = *(const uint4*)(base + hdr_off);
            acc = __uint_as_float((x ^ hd.x ^ hd.y ^ hd.z ^ hd.w) & 0x3fffffffu);
        } else if (LEVEL == 2) {
            int ra[4] = {0, 0, 0, 0}, ha[4] = {0, 0, 0, 0};
            uint32_t a[4], h[4];
            for (int j = 0; j < 4; j++) {
                ldmatrix_x4(smem_u32(base) + lm_off + 32 * j, a);
                for (int i = 0; i < 4; i++) h[i] = a[i] & 0xF0F0F0F0u;
                mma_u8s8(ra, a, ml.xo[0] + j, ml.xo[1]);
                mma_u8s8(ha, h, ml.xo[2], ml.xo[3] + j);
            }
            const uint4 hd = *(const uint4*)(base + hdr_off);
            acc = __uint_as_float((uint32_t)(ra[0] + ra[2] + ha[1] + ha[3] + hd.x + hd.y) & 0x3fffffffu);
        } else if (LEVEL >= 3) {
            XMma xm;
            load_x_mma(xm, ml, xq_s, bs32_s, xd_s, 2 * warp, 2);
            acc = q4k_rows8_blocks2_mma(smem_u32(base) + lm_off, base + hdr_off, ml, xm);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        if (++s == NSLOT) { s = 0; ph ^= 1; }
        if (LEVEL >= 4) {
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            mbar_wait(&part_free[parity], (fph >> parity) & 1);
            fph ^= 1 << parity;
            if (t4 == 0) red[(parity * 8 + warp) * 8 + g] = acc;
            __syncwarp();
            if (lane == 0) mbar_arrive(&part_full[parity]);
            parity ^= 1;
        } else {
            sink += acc;
        }
    }
    if (LEVEL < 4 && sink == 1.2345f) out[0] = sink;
}

template <int LEVEL, int PROD>
int run(const uint8_t* buf, size_t total, float* out, int sms, size_t wrap_bytes) {
    const size_t smem = NSLOT * STEP + 12 * 8 + 2 * 64 * 4 + 4096 + 64 + 256 + 128;
    CHECK(cudaFuncSetAttribute(pipe<LEVEL, PROD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0));
    CHECK(cudaEventCreate(&e1));
    const size_t n_tiles = total / (8 * ROW);
    float ms = 0;
    for (int rep = 0; rep < 2; rep++) {
        CHECK(cudaEventRecord(e0));
        pipe<LEVEL, PROD><<<sms * 2, 352, smem>>>(buf, n_tiles, out, wrap_bytes / (8 * ROW));
        CHECK(cudaEventRecord(e1));
        CHECK(cudaEventSynchronize(e1));
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
    }
    printf("level %d, producer mode %d, source wraps every %zu MB: %.2f TB/s\n", LEVEL, PROD, wrap_bytes >> 20, total / (ms * 1e-3) / 1e12);
    return 0;
}

int main() {
    const size_t total = (size_t)4 << 30;
    uint8_t* buf;
    float* out;
    CHECK(cudaMalloc(&buf, total));
    CHECK(cudaMalloc(&out, total / ROW * 4 + 64));
    // finite fp16 scales everywhere: 0x2c00 = 0.0625
    CHECK(cudaMemset(buf, 0x2c, total));
    int sms = 0;
    CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    for (size_t wrap : {total, (size_t)48 << 20}) {
        if (run<0, 0>(buf, total, out, sms, wrap)) return 1;
        if (run<0, 1>(buf, total, out, sms, wrap)) return 1;
        if (run<0, 2>(buf, total, out, sms, wrap)) return 1;
        if (run<0, 3>(buf, total, out, sms, wrap)) return 1;
        if (run<4, 0>(buf, total, out, sms, wrap)) return 1;
        if (run<4, 1>(buf, total, out, sms, wrap)) return 1;
        if (run<4, 2>(buf, total, out, sms, wrap)) return 1;
        if (run<4, 3>(buf, total, out, sms, wrap)) return 1;
    }
    return 0;
}
