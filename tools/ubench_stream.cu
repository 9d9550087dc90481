// ubench_stream.cu -- how fast can one B200 stream a weight-sized buffer out of HBM, and through which path?
//   mode bulk: per CTA one thread issues cp.async.bulk (UBLKCP) copies of `csz` bytes into a shared-memory ring of
//              `depth` slots of `per_slot` copies each; a consumer warp only waits and releases (no arithmetic).
//   mode ldg : every thread reads 16-byte words with ld.global.nc (no shared memory), 8 loads in flight per thread.
// Prints TB/s over a 4 GiB buffer (larger than L2) for a sweep of copy sizes / depths / CTAs per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    for (uint32_t i = 0; i < (1u << 24) && !ok; i++)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    if (!ok) asm volatile("trap;");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}

// each CTA streams the slices b, b + grid, ... of `slice` bytes; a slot holds per_slot copies of csz bytes (slice = per_slot * csz)
// stride: distance in bytes between the sources of consecutive copies of a slot (csz = contiguous)
__global__ void __launch_bounds__(64) bulk_stream(const uint8_t* src, size_t total, uint32_t csz, int per_slot, int depth, uint32_t stride, unsigned long long* sink, int ahead) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = (uint64_t*)smem;
    uint64_t* empty = full + 16;
    uint8_t* ring = smem + 256;
    const uint32_t slot_bytes = csz * per_slot;
    if (threadIdx.x == 0) {
        for (int i = 0; i < depth; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const size_t region = (size_t)stride * per_slot;   // bytes of source one slot spans
    const size_t n_slices = total / region;
    // with stride > csz a region holds stride / csz interleaved slots
    const int inter = (int)(stride / csz);
    if (threadIdx.x < 32) {
        uint64_t pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        int s = 0;
        uint32_t ph = 0;
        for (size_t i = blockIdx.x; i < n_slices; i += gridDim.x) {
            if (ahead > 0) {
                const size_t j = i + (size_t)ahead * gridDim.x;
                if (j < n_slices)
                    for (int c = threadIdx.x; c < per_slot; c += 32)
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src + j * region + (size_t)c * stride), "r"(stride) : "memory");
            }
            for (int k = 0; k < inter; k++) {
                if (threadIdx.x == 0) {
                    mbar_wait(&empty[s], ph ^ 1);
                    mbar_expect(&full[s], slot_bytes);
                }
                __syncwarp();
                for (int c = threadIdx.x; c < per_slot; c += 32)
                    bulk_g2s(ring + (size_t)s * slot_bytes + (size_t)c * csz, src + i * region + (size_t)c * stride + (size_t)k * csz, csz, &full[s], pol);
                if (++s == depth) { s = 0; ph ^= 1; }
            }
        }
    } else {
        int s = 0;
        uint32_t ph = 0;
        unsigned long long acc = 0;
        for (size_t i = blockIdx.x; i < n_slices; i += gridDim.x) {
            for (int k = 0; k < inter; k++) {
                mbar_wait(&full[s], ph);
                acc += *(volatile uint32_t*)(ring + (size_t)s * slot_bytes + 4 * (threadIdx.x - 32));
                __syncwarp();
                if (threadIdx.x == 32) mbar_arrive(&empty[s]);
                if (++s == depth) { s = 0; ph ^= 1; }
            }
        }
        if (acc == 0x1234567) sink[0] = acc;
    }
}

__global__ void __launch_bounds__(256) ldg_stream(const uint4* src, size_t n16, unsigned long long* sink) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0;
    for (; i + 7 * stride < n16; i += 8 * stride) {
        uint4 v[8];
#pragma unroll
        for (int k = 0; k < 8; k++)
            asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[k].x), "=r"(v[k].y), "=r"(v[k].z), "=r"(v[k].w) : "l"(src + i + k * stride));
#pragma unroll
        for (int k = 0; k < 8; k++) acc ^= v[k].x ^ v[k].y ^ v[k].z ^ v[k].w;
    }
    if (acc == 0x1234567) sink[0] = acc;
}

int main() {
    const size_t total = (size_t)4 << 30;
    uint8_t* buf;
    unsigned long long* sink;
    CHECK(cudaMalloc(&buf, total));
    CHECK(cudaMalloc(&sink, 64));
    CHECK(cudaMemset(buf, 1, total));
    int sms = 0;
    CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0));
    CHECK(cudaEventCreate(&e1));
    CHECK(cudaFuncSetAttribute(bulk_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    struct Cfg { uint32_t csz; int per_slot, depth, ctas_per_sm; uint32_t stride; int ahead; };
    const Cfg cfgs[] = {
        {1152, 16, 4, 2, 2304, 0}, {18432, 1, 4, 2, 0, 0}, {18432, 1, 8, 1, 0, 0},
        {1152, 16, 1, 2, 2304, 0}, {1152, 16, 1, 2, 2304, 2}, {1152, 16, 1, 2, 2304, 4}, {1152, 16, 1, 2, 2304, 8}, {1152, 16, 1, 2, 2304, 16},
        {1152, 16, 2, 2, 2304, 0}, {1152, 16, 2, 2, 2304, 2}, {1152, 16, 2, 2, 2304, 4}, {1152, 16, 2, 2, 2304, 8},
        {1152, 16, 4, 2, 2304, 4}, {1152, 16, 4, 2, 2304, 8},
        {1152, 16, 3, 2, 2304, 0}, {1152, 16, 5, 2, 2304, 0},
    };
    for (const Cfg& c : cfgs) {
        const uint32_t stride = c.stride ? c.stride : c.csz;
        const size_t smem = 256 + (size_t)c.csz * c.per_slot * c.depth;
        if (smem > 200 * 1024) continue;
        for (int rep = 0; rep < 2; rep++) {
            CHECK(cudaEventRecord(e0));
            bulk_stream<<<sms * c.ctas_per_sm, 64, smem>>>(buf, total, c.csz, c.per_slot, c.depth, stride, sink, c.ahead);
            CHECK(cudaEventRecord(e1));
            CHECK(cudaEventSynchronize(e1));
        }
        float ms;
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
        printf("bulk  csz %6u x %2d per slot, stride %6u, depth %2d, L2 ahead %2d slices, %d CTA/SM (%6.1f KB in flight per SM): %.2f TB/s\n", c.csz, c.per_slot, stride, c.depth,
               c.ahead, c.ctas_per_sm, c.csz * c.per_slot * c.depth * c.ctas_per_sm / 1024.0, total / (ms * 1e-3) / 1e12);
    }
    for (int ctas : {2, 4, 8}) {
        for (int rep = 0; rep < 2; rep++) {
            CHECK(cudaEventRecord(e0));
            ldg_stream<<<sms * ctas, 256>>>((const uint4*)buf, total / 16, sink);
            CHECK(cudaEventRecord(e1));
            CHECK(cudaEventSynchronize(e1));
        }
        float ms;
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
        printf("ldg.128 x8 in flight, %d CTAs of 256 per SM: %.2f TB/s\n", ctas, total / (ms * 1e-3) / 1e12);
    }
    return 0;
}
