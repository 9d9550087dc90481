// device_utils.cuh -- sm_100a device helpers: mbarrier, 1-D bulk async copy (TMA unit), programmatic
// dependent launch, packed-int8 dot, alignment-tolerant shared-memory loads.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier (shared::cta) ---------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    // make barrier initialisation visible to the async proxy (the TMA unit completes transactions on it)
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// add to the pending transaction count without arriving (the arrival that closes the phase comes separately)
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug must not hang the GPU (a hung box is a lost box).  2^22 probes of a
// HW-sleeping try_wait is seconds; on expiry the kernel traps and the launch reports an error.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
#pragma unroll 1
    for (uint32_t i = 0; i < (1u << 22); i++)
        if (mbar_try_wait(bar, parity)) return;
    asm volatile("trap;");
}

// ---- thread-block cluster / distributed shared memory ---------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
// all threads of all CTAs of the cluster (a plain launch is a cluster of one)
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory variable in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t dsmem_addr(const void* local, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(local)), "r"(rank));
    return r;
}
__device__ __forceinline__ void dsmem_st_u32(uint32_t addr, uint32_t v) {
    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void dsmem_st_u16(uint32_t addr, uint16_t v) {
    asm volatile("st.shared::cluster.u16 [%0], %1;" ::"r"(addr), "h"(v) : "memory");
}
__device__ __forceinline__ void dsmem_st_v2(uint32_t addr, uint32_t a, uint32_t b) {
    asm volatile("st.shared::cluster.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void dsmem_st_f64(uint32_t addr, double v) {
    asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
// Asynchronous store into a peer CTA's shared memory that reports its bytes to an mbarrier there
// (complete_tx), like a bulk copy: the receiver's wait on the barrier makes the data visible, so the
// sender needs neither a cluster fence nor a separate arrival.
__device__ __forceinline__ void dsmem_st_async_b32(uint32_t addr, uint32_t v, uint32_t mbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(addr), "r"(v), "r"(mbar) : "memory");
}
__device__ __forceinline__ void dsmem_st_async_v2(uint32_t addr, uint32_t a, uint32_t b, uint32_t mbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];" ::"r"(addr), "r"(a), "r"(b), "r"(mbar)
                 : "memory");
}
// the same store under a predicate (no branch around it: the activation exchange issues sixteen per lane)
__device__ __forceinline__ void dsmem_st_async_b32_if(bool p, uint32_t addr, uint32_t v, uint32_t mbar) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %3, 0;\n\t@q st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];\n\t}" ::"r"(addr), "r"(v),
                 "r"(mbar), "r"((uint32_t)p)
                 : "memory");
}
__device__ __forceinline__ void fence_cluster() { asm volatile("fence.acq_rel.cluster;" ::: "memory"); }
// arrive on an mbarrier that lives in another CTA of the cluster, releasing this thread's prior writes
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// poll with CTA-scope probes (a cluster-scope acquire invalidates L1 on every probe), then acquire once
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
#pragma unroll 1
    for (uint32_t i = 0; i < (1u << 22); i++)
        if (mbar_try_wait(bar, parity)) {
            mbar_try_wait_cluster(bar, parity);
            return;
        }
    asm volatile("trap;");
}

// dp4a with unsigned bytes on the left, signed on the right
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// ---- 1-D bulk async copy global -> shared, completion on an mbarrier (SASS: UBLKCP) ---------------
// dst and src 16-byte aligned, bytes a multiple of 16.
// L2 policy for data that is read once (the weight stream): evict first, so that activations, KV rows and
// prefetched operands survive in L2 next to it
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void bulk_g2s_hint(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
                 : "memory");
}
// pull [src, src + bytes) into L2 ahead of use (16-byte aligned address, size a multiple of 16)
__device__ __forceinline__ void bulk_prefetch_l2(const void* src_gmem, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// ---- programmatic dependent launch ---------------------------------------------------------------
// wait: block until every grid this launch depends on has completed and its writes are visible
// (no-op for an ordinary launch).  launch_dependents: allow the next grid in the stream to start
// its pre-wait part (weight prefetch) while this one is still running.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- named barrier for a subset of the CTA's warps ----------------------------------------------
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- wall clock (ns) for the optional in-kernel traces ---------------------------------------------
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Time stamps of the in-kernel traces: the SM's cycle counter scaled to ns at the boost clock (1965 MHz).  All marks of
// a trace come from one CTA, hence one SM, so they are comparable -- and unlike %globaltimer (updated every ~32 ns, and
// slow to read: a handful of reads per stage visibly stretched the traced CTA's stages) the read costs a few cycles.
__device__ __forceinline__ unsigned long long trace_ns() { return (unsigned long long)clock64() * 1000ull / 1965ull; }
__device__ __forceinline__ unsigned long long trace_after(uint32_t dep) {
    if (dep == 0x7fc0dead) asm volatile("trap;");
    return trace_ns();
}

// same, but not before `dep` has been computed (to time the arrival of loaded data)
__device__ __forceinline__ unsigned long long globaltimer_after(uint32_t dep) {
    if (dep == 0x7fc0dead) asm volatile("trap;");   // volatile statements keep their order: the read follows the compare
    return globaltimer_ns();
}

// ---- arithmetic ---------------------------------------------------------------------------------
__device__ __forceinline__ int dp4a_ss(uint32_t a, uint32_t b, int c) { return __dp4a((int)a, (int)b, c); }

__device__ __forceinline__ float h2f_bits(uint32_t h16) { return __half2float(__ushort_as_half((unsigned short)h16)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// 16 bytes from shared memory at an address that is only 2-byte aligned: five aligned words and a
// funnel shift by 0 or 16 bits.  (Q6_K blocks are 210 bytes and Q8_0 blocks 34, so inside a row they
// sit at every even offset.)
__device__ __forceinline__ uint4 lds16_a2(const uint8_t* p) {
    const uint32_t a = smem_u32(p);
    const uint32_t sh = (a & 2u) << 3;
    const uint32_t* w = (const uint32_t*)(p - (a & 3u));
    const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = w[4];
    uint4 r;
    r.x = __funnelshift_r(w0, w1, sh);
    r.y = __funnelshift_r(w1, w2, sh);
    r.z = __funnelshift_r(w2, w3, sh);
    r.w = __funnelshift_r(w3, w4, sh);
    return r;
}
// same for global memory (used by the dequantise kernels, which read blocks in place)
__device__ __forceinline__ uint32_t ldg32_a2(const uint8_t* p) {
    const uint16_t* h = (const uint16_t*)p;
    return (uint32_t)h[0] | ((uint32_t)h[1] << 16);
}

}  // namespace b2
