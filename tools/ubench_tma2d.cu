// ubench_tma2d.cu -- how fast does one SM's TMA unit fill shared memory with 2-D boxes of a byte matrix, as a function of
// the box shape?  (tcq_gemm.cu fetches quantised weight rows: 128 rows x 144 bytes per box turned out to be slow.)
// Every CTA (one per SM) streams its share of a 2 GiB matrix of 2304-byte rows (Q4_K, K = 4096) through a ring of 64 KiB;
// a consumer warp releases every box as soon as it has landed.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lcuda -o tools/ubench_tma2d tools/ubench_tma2d.cu
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void bar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void bar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory"); }
__device__ __forceinline__ void bar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}

// boxes of `bw` elements x `bh` rows; cols = boxes per row, n_boxes in all
__global__ void __launch_bounds__(64, 1) stream(const __grid_constant__ CUtensorMap map, int bw_elems, int bh, uint32_t box_bytes, int cols, long long n_boxes,
                                                int stages) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
    const uint32_t bars = base + stages * ((box_bytes + 1023u) & ~1023u);
    const uint32_t stride = (box_bytes + 1023u) & ~1023u;
    if (threadIdx.x == 0) {
        for (int s = 0; s < 2 * stages; s++) bar_init(bars + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const long long lo = n_boxes * blockIdx.x / gridDim.x, hi = n_boxes * (blockIdx.x + 1) / gridDim.x;
    if (threadIdx.x == 0) {
        int s = 0; uint32_t ph = 0;
        for (long long b = lo; b < hi; b++) {
            bar_wait(bars + 8 * (stages + s), ph ^ 1);
            bar_expect_tx(bars + 8 * s, box_bytes);
            tma_load_2d(base + s * stride, &map, bars + 8 * s, (int)(b % cols) * bw_elems, (int)(b / cols) * bh);
            if (++s == stages) { s = 0; ph ^= 1; }
        }
    } else if (threadIdx.x == 32) {
        int s = 0; uint32_t ph = 0;
        for (long long b = lo; b < hi; b++) {
            bar_wait(bars + 8 * s, ph);
            bar_arrive(bars + 8 * (stages + s));
            if (++s == stages) { s = 0; ph ^= 1; }
        }
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    const size_t row_bytes = 2304, rows = (size_t)(2ull << 30) / row_bytes / 128 * 128;
    uint8_t* buf;
    CHECK(cudaMalloc(&buf, rows * row_bytes));
    CHECK(cudaMemset(buf, 1, rows * row_bytes));
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    EncodeTiledFn enc = (EncodeTiledFn)fp;
    int sms = 0;
    CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    struct Shape { int esize, bw_bytes, bh; } shapes[] = {{1, 144, 128}, {1, 144, 64}, {1, 144, 32}, {1, 144, 256}, {4, 288, 64}, {4, 576, 32}, {4, 576, 64}, {4, 576, 128},
                                                           {8, 1152, 16}, {8, 1152, 32}, {8, 1152, 64}, {8, 2304, 8}, {8, 2304, 16}, {8, 2304, 32}, {1, 256, 128}, {1, 128, 128}};
    CHECK(cudaFuncSetAttribute(stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (auto& sh : shapes) {
        const CUtensorMapDataType dt = sh.esize == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : sh.esize == 4 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT64;
        CUtensorMap map;
        const cuuint64_t dims[2] = {row_bytes / sh.esize, rows};
        const cuuint64_t strides[1] = {row_bytes};
        const cuuint32_t box[2] = {(cuuint32_t)(sh.bw_bytes / sh.esize), (cuuint32_t)sh.bh};
        const cuuint32_t estr[2] = {1, 1};
        const CUresult r = enc(&map, dt, 2, buf, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("box %d x %d: encode failed (%d)\n", sh.bw_bytes, sh.bh, (int)r); continue; }
        const uint32_t box_bytes = (uint32_t)sh.bw_bytes * sh.bh;
        const int cols = (int)(row_bytes / sh.bw_bytes);
        const long long n_boxes = (long long)cols * (rows / sh.bh);
        for (int stages : {2, 4, 8}) {
            const size_t smem = (size_t)stages * ((box_bytes + 1023) & ~1023u) + 1024 + 16 * stages + 64;
            if (smem > 200 * 1024) continue;
            cudaEvent_t e0, e1;
            CHECK(cudaEventCreate(&e0)); CHECK(cudaEventCreate(&e1));
            float ms = 0;
            for (int rep = 0; rep < 2; rep++) {
                CHECK(cudaEventRecord(e0));
                stream<<<sms, 64, smem>>>(map, sh.bw_bytes / sh.esize, sh.bh, box_bytes, cols, n_boxes, stages);
                CHECK(cudaEventRecord(e1));
                CHECK(cudaEventSynchronize(e1));
                CHECK(cudaEventElapsedTime(&ms, e0, e1));
            }
            printf("box %4d bytes x %3d rows (%6u B), %d stages (%3zu KB in flight): %6.2f TB/s\n", sh.bw_bytes, sh.bh, box_bytes, stages,
                   (size_t)stages * box_bytes >> 10, (double)n_boxes * box_bytes / (ms * 1e-3) / 1e12);
        }
    }
    return 0;
}
