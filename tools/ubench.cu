// ubench.cu -- instruction-throughput probes for the decode mat-vec on sm_100a (build: make -C tools ubench).
// Prints warp-instructions per clock per SM for the candidates of the inner loop: dp4a, lop3, imad, the legacy
// mma.sync int8 / fp16 shapes, 16-byte shared-memory loads, warp shuffles.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;

template <int MODE>
__global__ void __launch_bounds__(1024) probe(uint32_t* out, uint32_t seed, long long* cycles) {
    __shared__ __align__(16) uint32_t sm[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = i * seed;
    __syncthreads();
    uint32_t a0 = seed + threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, a4 = a0 * 11, a5 = a0 * 13, a6 = a0 * 17, a7 = a0 * 19;
    int c0 = 0, c1 = 1, c2 = 2, c3 = 3, c4 = 4, c5 = 5, c6 = 6, c7 = 7;
    float f0 = 0, f1 = 0, f2 = 0, f3 = 0, g0 = 0, g1 = 0, g2 = 0, g3 = 0;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
        if (MODE == 0) {   // 8 independent dp4a chains
            c0 = __dp4a((int)a0, (int)a1, c0); c1 = __dp4a((int)a1, (int)a2, c1); c2 = __dp4a((int)a2, (int)a3, c2); c3 = __dp4a((int)a3, (int)a4, c3);
            c4 = __dp4a((int)a4, (int)a5, c4); c5 = __dp4a((int)a5, (int)a6, c5); c6 = __dp4a((int)a6, (int)a7, c6); c7 = __dp4a((int)a7, (int)a0, c7);
        } else if (MODE == 1) {   // lop3
            a0 = (a0 & 0x0f0f0f0fu) ^ a1; a1 = (a1 & 0xf0f0f0f0u) ^ a2; a2 = (a2 & 0x0f0f0f0fu) ^ a3; a3 = (a3 & 0xf0f0f0f0u) ^ a4;
            a4 = (a4 & 0x0f0f0f0fu) ^ a5; a5 = (a5 & 0xf0f0f0f0u) ^ a6; a6 = (a6 & 0x0f0f0f0fu) ^ a7; a7 = (a7 & 0xf0f0f0f0u) ^ a0;
        } else if (MODE == 2) {   // imad
            c0 = c0 * (int)a0 + c1; c1 = c1 * (int)a1 + c2; c2 = c2 * (int)a2 + c3; c3 = c3 * (int)a3 + c4;
            c4 = c4 * (int)a4 + c5; c5 = c5 * (int)a5 + c6; c6 = c6 * (int)a6 + c7; c7 = c7 * (int)a7 + c0;
        } else if (MODE == 3) {   // mma.sync m16n8k32 u8 x s8 -> s32, two independent accumulators
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c0), "+r"(c1), "+r"(c2), "+r"(c3) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(a4), "r"(a5));
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c4), "+r"(c5), "+r"(c6), "+r"(c7) : "r"(a1), "r"(a2), "r"(a3), "r"(a0), "r"(a6), "r"(a7));
        } else if (MODE == 4) {   // mma.sync m16n8k16 f16 -> f32, two independent accumulators
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(f0), "+f"(f1), "+f"(f2), "+f"(f3) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(a4), "r"(a5));
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(g0), "+f"(g1), "+f"(g2), "+f"(g3) : "r"(a1), "r"(a2), "r"(a3), "r"(a0), "r"(a6), "r"(a7));
        } else if (MODE == 5) {   // LDS.128, conflict-free, 4 independent
            const uint4 v0 = *(const uint4*)&sm[((threadIdx.x + it) & 255) * 4];
            const uint4 v1 = *(const uint4*)&sm[1024 + ((threadIdx.x + it) & 255) * 4];
            const uint4 v2 = *(const uint4*)&sm[2048 + ((threadIdx.x + it) & 255) * 4];
            const uint4 v3 = *(const uint4*)&sm[3072 + ((threadIdx.x + it) & 255) * 4];
            a0 ^= v0.x ^ v0.w; a1 ^= v1.y ^ v1.z; a2 ^= v2.z ^ v2.x; a3 ^= v3.w ^ v3.y;
        } else if (MODE == 6) {   // shuffles
            a0 = __shfl_xor_sync(0xffffffffu, a0, 1); a1 = __shfl_xor_sync(0xffffffffu, a1, 2); a2 = __shfl_xor_sync(0xffffffffu, a2, 4);
            a3 = __shfl_xor_sync(0xffffffffu, a3, 8);
        } else if (MODE == 7) {   // the Q4_K inner pattern: lop + dp4a pairs, 4 chains
            c0 = __dp4a((int)(a0 & 0x0f0f0f0fu), (int)a4, c0); c1 = __dp4a((int)(a0 & 0xf0f0f0f0u), (int)a5, c1);
            c2 = __dp4a((int)(a1 & 0x0f0f0f0fu), (int)a6, c2); c3 = __dp4a((int)(a1 & 0xf0f0f0f0u), (int)a7, c3);
            c0 = __dp4a((int)(a2 & 0x0f0f0f0fu), (int)a4, c0); c1 = __dp4a((int)(a2 & 0xf0f0f0f0u), (int)a5, c1);
            c2 = __dp4a((int)(a3 & 0x0f0f0f0fu), (int)a6, c2); c3 = __dp4a((int)(a3 & 0xf0f0f0f0u), (int)a7, c3);
            a0 += c0; a1 += c1; a2 += c2; a3 += c3;
        } else if (MODE == 8) {   // i2f + ffma
            f0 = f0 * g0 + (float)c0; f1 = f1 * g1 + (float)c1; f2 = f2 * g2 + (float)c2; f3 = f3 * g3 + (float)c3;
            c0 += 1; c1 += 2; c2 += 3; c3 += 4;
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7 ^ (uint32_t)(c0 + c1 + c2 + c3 + c4 + c5 + c6 + c7) ^
                                                 __float_as_uint(f0 + f1 + f2 + f3 + g0 + g1 + g2 + g3);
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
int run(const char* name, int per_iter, int threads, uint32_t* out, long long* cyc) {
    int sms = 0;
    CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    probe<MODE><<<sms, threads>>>(out, 12345u, cyc);
    CHECK(cudaDeviceSynchronize());
    probe<MODE><<<sms, threads>>>(out, 12345u, cyc);
    CHECK(cudaDeviceSynchronize());
    long long h[256];
    CHECK(cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
    double avg = 0;
    for (int i = 0; i < sms; i++) avg += (double)h[i];
    avg /= sms;
    const double warp_instr = (double)ITERS * per_iter * (threads / 32);
    printf("%-44s threads/SM %4d: %.3f warp-instr/clk/SM  (%.1f lanes/clk/SM)\n", name, threads, warp_instr / avg, 32.0 * warp_instr / avg);
    return 0;
}

int main() {
    uint32_t* out;
    long long* cyc;
    CHECK(cudaMalloc(&out, 256 * 1024 * 4));
    CHECK(cudaMalloc(&cyc, 256 * 8));
    for (int threads : {256, 512, 1024}) {
        if (run<0>("dp4a (8 chains)", 8, threads, out, cyc)) return 1;
        if (run<1>("lop3 (8 chains)", 8, threads, out, cyc)) return 1;
        if (run<2>("imad (8 chains)", 8, threads, out, cyc)) return 1;
        if (run<3>("mma.sync m16n8k32 u8.s8 (2 chains)", 2, threads, out, cyc)) return 1;
        if (run<4>("mma.sync m16n8k16 f16 (2 chains)", 2, threads, out, cyc)) return 1;
        if (run<5>("lds.128 (4 loads + 8 xor per iteration)", 4, threads, out, cyc)) return 1;
        if (run<6>("shfl (4 chains)", 4, threads, out, cyc)) return 1;
        if (run<7>("Q4_K pattern: 8 lop + 8 dp4a + 4 iadd", 20, threads, out, cyc)) return 1;
        if (run<8>("i2f + ffma + iadd (4 chains)", 12, threads, out, cyc)) return 1;
    }
    return 0;
}
