// tp.cu -- the exchange step of tensor parallelism (the 70B model over 2/4/8 GPUs of one box).
//
// Row/column split as the scope table prescribes: attn_q/k/v by heads and ffn_gate/up by ff rows (contiguous row
// ranges of the GGUF matrices), attn_output and ffn_down along K on super-block boundaries (model.cu uploads each
// rank's slice).  Every rank therefore ends a half-layer with a partial d_model vector; this file sums them:
// the row-split mat-vecs leave x / tp + partial in place of x, ncclAllReduce(sum, fp32) over NVLink makes that the new
// residual stream on every rank.  One process per GPU; the host application hands every rank the
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
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
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
        n.AllGather = (decltype(n.AllGather))dlsym(n.lib, "ncclAllGather");
        n.GetErrorString = (decltype(n.GetErrorString))dlsym(n.lib, "ncclGetErrorString");
        if (!n.GetUniqueId || !n.CommInitRank || !n.CommDestroy || !n.AllReduce || !n.AllGather || !n.GetErrorString) {
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

// ---- peer memory for the exchange inside the persistent kernel (step2.cu: OP_XCHG) ----
// Every rank allocates `bytes` of device memory (zeroed), publishes its CUDA IPC handle through the communicator
// (ncclAllGather of the 64-byte handles) and maps every peer's allocation: peers[r] is rank r's buffer as seen from this
// process (peers[rank] = the local allocation).  The kernels STORE flags into a peer's buffer and, in the pull form of the exchange, LOAD partial rows from it (plain loads over NVLink, no multicast); everything else they read is their
// own (plain st.global / ld.global over NVLink: no TMA, no multicast).
int tp_peer_buffers(void* comm, int rank, int size, size_t bytes, void** local_out, void** peers, cudaStream_t st) {
    Nccl* nc;
    int rc;
    if ((rc = nccl(&nc))) return rc;
    if (!comm || size < 2 || size > 8 || rank < 0 || rank >= size) {
        set_error("tp_peer_buffers: bad group (rank %d of %d; up to 8 ranks)", rank, size);
        return B2_EINVAL;
    }
    void* local = nullptr;
    B2_CUDA_CHECK(cudaMalloc(&local, bytes));
    B2_CUDA_CHECK(cudaMemset(local, 0, bytes));
    cudaIpcMemHandle_t mine;
    B2_CUDA_CHECK(cudaIpcGetMemHandle(&mine, local));
    uint8_t *d_mine = nullptr, *d_all = nullptr;
    B2_CUDA_CHECK(cudaMalloc((void**)&d_mine, sizeof mine));
    B2_CUDA_CHECK(cudaMalloc((void**)&d_all, sizeof mine * size));
    B2_CUDA_CHECK(cudaMemcpyAsync(d_mine, &mine, sizeof mine, cudaMemcpyHostToDevice, st));
    B2_NCCL_CHECK(nc, nc->AllGather(d_mine, d_all, sizeof mine, ncclChar, (ncclComm_t)comm, st));
    std::vector<cudaIpcMemHandle_t> all(size);
    B2_CUDA_CHECK(cudaMemcpyAsync(all.data(), d_all, sizeof mine * size, cudaMemcpyDeviceToHost, st));
    B2_CUDA_CHECK(cudaStreamSynchronize(st));
    cudaFree(d_mine);
    cudaFree(d_all);
    for (int r = 0; r < size; r++) {
        if (r == rank) {
            peers[r] = local;
            continue;
        }
        void* p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            set_error("tensor parallelism: cannot map rank %d's exchange buffer (%s); B2_TP_NCCL=1 keeps the NCCL all-reduce path", r, cudaGetErrorString(e));
            return B2_ECUDA;
        }
        peers[r] = p;
    }
    *local_out = local;
    return 0;
}

void tp_peer_release(int rank, int size, void* local, void** peers) {
    for (int r = 0; r < size; r++)
        if (r != rank && peers[r]) cudaIpcCloseMemHandle(peers[r]);
    if (local) cudaFree(local);
}

// every rank has reached this point (and finished what it enqueued before): a one-float all-reduce
int tp_barrier(void* comm, float* scratch_dev, cudaStream_t st) {
    Nccl* nc;
    int rc;
    if ((rc = nccl(&nc))) return rc;
    B2_NCCL_CHECK(nc, nc->AllReduce(scratch_dev, scratch_dev, 1, ncclFloat, ncclSum, (ncclComm_t)comm, st));
    B2_CUDA_CHECK(cudaStreamSynchronize(st));
    return 0;
}

}  // namespace b2

