// ubench_umma.cu -- cycles per tcgen05.mma (kind::f16, K = 16) as a function of M, N and where A lives, issued back to
// back by one thread into one accumulator: what does a narrow-N MMA cost the tensor pipe?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I llms-on-kubernetes_b200/csrc -o tools/ubench_umma tools/ubench_umma.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "device_utils.cuh"
#include "tc_ptx.cuh"

using namespace b2;
using namespace b2::tcptx;

__global__ void __launch_bounds__(128, 1) k(int M, int N, int a_in_tmem, int reps, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
    const uint32_t bar = base + 96 * 1024, slot = bar + 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) ((uint32_t*)(smem + (base - smem_u32(smem))))[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { bar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(slot), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(slot) : "memory");
    if (warp == 0) {
        const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint64_t da = umma_desc_sw128(base), db = umma_desc_sw128(base + 32 * 1024);
        long long t0 = 0, t1 = 0;
        if (elect_one_sync_()) {
            t0 = clock64();
            for (int r = 0; r < reps; r++) {
                if (a_in_tmem) umma_f16_ta(tmem_base, tmem_base + 256 + 8 * (r & 15), db + 2 * (r & 3), idesc, 1);
                else umma_f16(tmem_base, da + 2 * (r & 3), db + 2 * (r & 3), idesc, 1);
            }
            t1 = clock64();
            umma_commit(bar);
        }
        __syncwarp();
        bar_wait(bar, 0);
        const long long t2 = clock64();
        if (lane == 0 && t0) { out[0] = t1 - t0; out[1] = t2 - t0; }
        long long a = __shfl_sync(0xffffffffu, t1 - t0, 0);
        (void)a;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(512) : "memory");
    }
}

int main() {
    long long* d;
    cudaMalloc(&d, 64);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    const int reps = 2048;
    for (int a_tm = 0; a_tm < 2; a_tm++)
        for (int M : {64, 128})
            for (int N : {16, 32, 64, 128, 256}) {
                long long h[2] = {0, 0};
                for (int rep = 0; rep < 2; rep++) {
                    cudaMemset(d, 0, 64);
                    k<<<1, 128, 100 * 1024>>>(M, N, a_tm, reps, d);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("M=%d N=%d: %s\n", M, N, cudaGetErrorString(e)); return 1; }
                    cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
                }
                printf("A in %s, M=%3d N=%3d: issue %6.1f cycles / MMA, complete %6.1f cycles / MMA\n", a_tm ? "tmem" : "smem", M, N, (double)h[0] / reps, (double)h[1] / reps);
            }
    return 0;
}
