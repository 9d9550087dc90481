// tp.cu -- the exchange step of tensor parallelism (the 70B model over 2/4/8 GPUs of one box).
//
// Row/column split as the scope table prescribes: attn_q/k/v by heads and ffn_gate/up by ff rows (contiguous row
// ranges of the GGUF matrices), attn_output and ffn_down along K on super-block boundaries (model.cu uploads each
// rank's slice).  Every rank therefore ends a half-layer with a partial d_model vector; this file sums them:
// ncclAllReduce(sum, fp32) over NVLink, then x += sum.  One process per GPU; the host application hands every rank the
// same 128-byte NCCL id (b2_allreduce_unique_id on rank 0, broadcast by whatever launched the ranks).
//
// libnccl is loaded with dlopen on first use: a single-GPU deployment of libb2.so neither links nor needs it.
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>
#include <mutex>
#include <vector>

#include "common.h"
#include "ops.h"

namespace b2 {
namespace {

struct Nccl {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

int nccl(Nccl** out) {
    static Nccl n;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    if (!n.lib) {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names)
            if ((n.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL))) break;
        if (!n.lib) {
            set_error("tensor parallelism needs libnccl.so.2 on the library path: %s", dlerror());
            return B2_EUNSUP;
        }
        n.GetUniqueId = (decltype(n.GetUniqueId))dlsym(n.lib, "ncclGetUniqueId");
        n.CommInitRank = (decltype(n.CommInitRank))dlsym(n.lib, "ncclCommInitRank");
        n.CommDestroy = (decltype(n.CommDestroy))dlsym(n.lib, "ncclCommDestroy");
        n.AllReduce = (decltype(n.AllReduce))dlsym(n.lib, "ncclAllReduce");
        n.GetErrorString = (decltype(n.GetErrorString))dlsym(n.lib, "ncclGetErrorString");
        if (!n.GetUniqueId || !n.CommInitRank || !n.CommDestroy || !n.AllReduce || !n.GetErrorString) {
            set_error("libnccl does not export the expected symbols");
            dlclose(n.lib);
            n.lib = nullptr;
            return B2_EUNSUP;
        }
    }
    *out = &n;
    return 0;
}

#define B2_NCCL_CHECK(N, expr)                                                                 \
    do {                                                                                       \
        ncclResult_t _r = (expr);                                                              \
        if (_r != ncclSuccess) {                                                               \
            set_error("NCCL error at %s:%d: %s", __FILE__, __LINE__, (N)->GetErrorString(_r)); \
            return B2_ECUDA;                                                                   \
        }                                                                                      \
    } while (0)

__global__ void residual_add_kernel(float4* __restrict__ x, const float4* __restrict__ part, int n4) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n4) {
        float4 a = x[i];
        const float4 b = part[i];
        a.x += b.x;
        a.y += b.y;
        a.z += b.z;
        a.w += b.w;
        x[i] = a;
    }
}

}  // namespace

int tp_unique_id(void* out, size_t cap) {
    if (!out || cap < sizeof(ncclUniqueId)) {
        set_error("b2_allreduce_unique_id: the buffer must hold %zu bytes", sizeof(ncclUniqueId));
        return B2_EINVAL;
    }
    Nccl* n;
    int rc;
    if ((rc = nccl(&n))) return rc;
    ncclUniqueId id;
    B2_NCCL_CHECK(n, n->GetUniqueId(&id));
    memcpy(out, &id, sizeof id);
    return (int)sizeof id;
}

int tp_comm_init(const void* id_bytes, size_t len, int rank, int size, void** comm_out) {
    if (!id_bytes || len < sizeof(ncclUniqueId) || !comm_out || size < 2 || rank < 0 || rank >= size) {
        set_error("b2_allreduce_init: bad arguments (rank %d of %d, id of %zu bytes)", rank, size, len);
        return B2_EINVAL;
    }
    Nccl* n;
    int rc;
    if ((rc = nccl(&n))) return rc;
    ncclUniqueId id;
    memcpy(&id, id_bytes, sizeof id);
    ncclComm_t comm = nullptr;
    B2_NCCL_CHECK(n, n->CommInitRank(&comm, size, id, rank));
    *comm_out = comm;
    return 0;
}

void tp_comm_free(void* comm) {
    Nccl* n;
    if (comm && nccl(&n) == 0) n->CommDestroy((ncclComm_t)comm);
}

// ---- one-shot all-reduce over peer memory -------------------------------------------------------
// A d_model vector is 16-32 KB: an NCCL all-reduce of that size is ~18 us of protocol latency, 160 times per token on
// the 70B model.  Here every rank owns an exchange area {two vector buffers, two rows of arrival flags} that its peers
// map through CUDA IPC (plain loads and stores over NVLink; the handles travel once through ncclAllGather).  One
// CTA per rank: copy x into my buffer of this round's parity, fence, store the round number into my flag on every
// peer, wait until every rank's flag in my area shows the round, then sum all ranks' buffers in rank order (so every
// rank computes bit-identical sums) into x.  Two buffers suffice: a rank can start round k + 2 only after every peer
// has announced round k + 1, i.e. has finished reading round k.  The round counter lives in device memory, so the
// kernel replays inside a CUDA graph.  A peer that never arrives ends in a trap (an error on the stream), not a hang.
constexpr int TP_MAX_RANKS = 8;

struct TpXchg {
    void* peers[TP_MAX_RANKS];   // every rank's exchange area as mapped here (mine included)
    int rank, size, n;           // n floats per vector
    unsigned int* round;         // device-resident round counter (mine)
};

struct TpOneShot {
    void* area = nullptr;        // my exchange area (+ the round counter behind it)
    std::vector<void*> opened;   // peers' areas mapped through IPC
    TpXchg args{};
    int rank = 0, size = 0;
};
void tp_oneshot_free(void* state);

__device__ __forceinline__ float* xchg_buf(void* base, int parity, int n) { return (float*)base + (size_t)parity * n; }
__device__ __forceinline__ unsigned int* xchg_flags(void* base, int parity, int n) {
    return (unsigned int*)((float*)base + 2 * (size_t)n) + parity * TP_MAX_RANKS;
}

__global__ void __launch_bounds__(1024) tp_oneshot_kernel(const TpXchg a, float* __restrict__ x) {
    __shared__ unsigned int s_round;
    if (threadIdx.x == 0) s_round = *a.round + 1;
    __syncthreads();
    const unsigned int round = s_round;
    const int parity = (int)(round & 1u);
    float4* mine = (float4*)xchg_buf(a.peers[a.rank], parity, a.n);
    const int n4 = a.n / 4;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) mine[i] = ((const float4*)x)[i];
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < a.size) {
        // announce: my slot in rank threadIdx.x's flag row (a remote store for the peers), then wait for that rank's
        // announcement in my own row
        unsigned int* theirs = xchg_flags(a.peers[threadIdx.x], parity, a.n) + a.rank;
        asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(theirs), "r"(round) : "memory");
        const unsigned int* here = xchg_flags(a.peers[a.rank], parity, a.n) + threadIdx.x;
        unsigned int v;
        long long spins = 0;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(here) : "memory");
            if (++spins > (1ll << 26)) __trap();   // tens of seconds: a rank of the group is gone
        } while ((int)(v - round) < 0);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        float4 acc = ((const float4*)xchg_buf(a.peers[0], parity, a.n))[i];
        for (int r = 1; r < a.size; r++) {
            const float4 v = ((const float4*)xchg_buf(a.peers[r], parity, a.n))[i];
            acc.x += v.x;
            acc.y += v.y;
            acc.z += v.z;
            acc.w += v.w;
        }
        ((float4*)x)[i] = acc;
    }
    __syncthreads();
    if (threadIdx.x == 0) *a.round = round;
}

size_t tp_xchg_bytes(int n) { return 2 * (size_t)n * sizeof(float) + 2 * TP_MAX_RANKS * sizeof(unsigned int); }

// Collective over the communicator's ranks: allocate my exchange area, swap IPC handles, map the peers'.
// *state_out owns everything (tp_oneshot_free).  An error leaves the caller on the NCCL path.
int tp_oneshot_init(void* comm, int rank, int size, int n, cudaStream_t st, void** state_out) {
    *state_out = nullptr;
    if (size > TP_MAX_RANKS || n % 4) {
        set_error("one-shot all-reduce: %d ranks / %d floats unsupported", size, n);
        return B2_EUNSUP;
    }
    Nccl* nc;
    int rc;
    if ((rc = nccl(&nc))) return rc;
    auto AllGather = (ncclResult_t(*)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t))dlsym(nc->lib, "ncclAllGather");
    if (!AllGather) {
        set_error("libnccl does not export ncclAllGather");
        return B2_EUNSUP;
    }
    TpOneShot* s = new TpOneShot();
    s->size = size;
    s->rank = rank;
    const size_t bytes = tp_xchg_bytes(n);
    B2_CUDA_CHECK(cudaMalloc(&s->area, bytes + sizeof(unsigned int)));
    B2_CUDA_CHECK(cudaMemsetAsync(s->area, 0, bytes + sizeof(unsigned int), st));
    cudaIpcMemHandle_t mine;
    B2_CUDA_CHECK(cudaIpcGetMemHandle(&mine, s->area));
    uint8_t* d_h = nullptr;
    B2_CUDA_CHECK(cudaMalloc((void**)&d_h, (size_t)(size + 1) * sizeof mine));
    B2_CUDA_CHECK(cudaMemcpyAsync(d_h + (size_t)size * sizeof mine, &mine, sizeof mine, cudaMemcpyHostToDevice, st));
    B2_NCCL_CHECK(nc, AllGather(d_h + (size_t)size * sizeof mine, d_h, sizeof mine, ncclChar, (ncclComm_t)comm, st));
    std::vector<cudaIpcMemHandle_t> all(size);
    B2_CUDA_CHECK(cudaMemcpyAsync(all.data(), d_h, (size_t)size * sizeof mine, cudaMemcpyDeviceToHost, st));
    B2_CUDA_CHECK(cudaStreamSynchronize(st));
    cudaFree(d_h);
    TpXchg a{};
    a.rank = rank;
    a.size = size;
    a.n = n;
    a.round = (unsigned int*)((uint8_t*)s->area + bytes);
    for (int r = 0; r < size; r++) {
        if (r == rank) {
            a.peers[r] = s->area;
            continue;
        }
        void* p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("one-shot all-reduce: cannot map rank %d's exchange area (%s)", r, cudaGetErrorString(e));
            tp_oneshot_free(s);
            return B2_ECUDA;
        }
        s->opened.push_back(p);
        a.peers[r] = p;
    }
    s->args = a;
    *state_out = s;
    return 0;
}

void tp_oneshot_free(void* state) {
    TpOneShot* s = (TpOneShot*)state;
    if (!s) return;
    for (void* p : s->opened) cudaIpcCloseMemHandle(p);
    if (s->area) cudaFree(s->area);
    delete s;
}

int tp_oneshot_allreduce(void* state, float* x, cudaStream_t st) {
    TpOneShot* s = (TpOneShot*)state;
    tp_oneshot_kernel<<<1, 1024, 0, st>>>(s->args, x);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

// x[0..n) = sum over the ranks of x[0..n), in place.  The row-split mat-vecs leave x = x / tp + partial on every rank
// (EPI_RESID_TP), so the sum is residual + the whole product without a separate add kernel.
int tp_allreduce_inplace(void* comm, float* x, int n, cudaStream_t st) {
    Nccl* nc;
    int rc;
    if ((rc = nccl(&nc))) return rc;
    if (!comm) {
        set_error("tensor-parallel context used before b2_allreduce_init");
        return B2_EINVAL;
    }
    B2_NCCL_CHECK(nc, nc->AllReduce(x, x, (size_t)n, ncclFloat, ncclSum, (ncclComm_t)comm, st));
    return 0;
}

// x[0..n) += sum over the ranks of part[0..n)   (part is overwritten with the sum)
int tp_allreduce_add(void* comm, float* part, float* x, int n, cudaStream_t st) {
    Nccl* nc;
    int rc;
    if ((rc = nccl(&nc))) return rc;
    if (!comm) {
        set_error("tensor-parallel context used before b2_allreduce_init");
        return B2_EINVAL;
    }
    B2_NCCL_CHECK(nc, nc->AllReduce(part, part, (size_t)n, ncclFloat, ncclSum, (ncclComm_t)comm, st));
    residual_add_kernel<<<(n / 4 + 255) / 256, 256, 0, st>>>((float4*)x, (const float4*)part, n / 4);
    B2_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // namespace b2
