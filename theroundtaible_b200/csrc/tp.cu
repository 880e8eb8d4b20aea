#include "tp.cuh"

#include <dlfcn.h>

#include <cstring>
#include <string>

#include "../../include/rt_engine.h"

namespace {
typedef struct { char internal[128]; } ncclUniqueId_t;
typedef void* ncclComm_t_;
typedef int ncclResult_t_;
enum { NCCL_FLOAT32 = 7, NCCL_FLOAT16 = 6, NCCL_INT8 = 0, NCCL_SUM = 0 };

struct Api {
    ncclResult_t_ (*GetUniqueId)(ncclUniqueId_t*) = nullptr;
    ncclResult_t_ (*CommInitRank)(ncclComm_t_*, int, ncclUniqueId_t, int) = nullptr;
    ncclResult_t_ (*CommDestroy)(ncclComm_t_) = nullptr;
    ncclResult_t_ (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t_, cudaStream_t) = nullptr;
    ncclResult_t_ (*AllGather)(const void*, void*, size_t, int, ncclComm_t_, cudaStream_t) = nullptr;
    ncclResult_t_ (*GroupStart)() = nullptr;
    ncclResult_t_ (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t_) = nullptr;
};

Api& api() {
    static Api a;
    static bool loaded = false;
    if (loaded) return a;
    // a copy already mapped into the process (e.g. torch's bundled libnccl.so.2) is reused by SONAME
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) throw RtError(RT_ERR_NCCL, std::string("libnccl.so.2 not found: ") + dlerror());
    auto sym = [&](const char* n) {
        void* p = dlsym(h, n);
        if (!p) throw RtError(RT_ERR_NCCL, std::string("NCCL symbol missing: ") + n);
        return p;
    };
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(sym("ncclAllReduce"));
    a.AllGather = reinterpret_cast<decltype(a.AllGather)>(sym("ncclAllGather"));
    a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
    a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
    loaded = true;
    return a;
}

void check(ncclResult_t_ r, const char* what) {
    if (r != 0) throw RtError(RT_ERR_NCCL, std::string("NCCL error in ") + what + ": " + api().GetErrorString(r));
}
}  // namespace

struct TpComm {
    ncclComm_t_ comm = nullptr;
    int rank = 0, size = 1;
};

void tp_unique_id(void* out128) {
    ncclUniqueId_t id;
    check(api().GetUniqueId(&id), "ncclGetUniqueId");
    memcpy(out128, &id, 128);
}

TpComm* tp_init(int rank, int size, const void* id128) {
    RT_REQUIRE(id128 != nullptr, "tp_size > 1 needs rt_config.nccl_id (rt_nccl_unique_id on rank 0)");
    ncclUniqueId_t id;
    memcpy(&id, id128, 128);
    TpComm* c = new TpComm();
    c->rank = rank;
    c->size = size;
    try {
        check(api().CommInitRank(&c->comm, size, id, rank), "ncclCommInitRank");
    } catch (...) {
        delete c;
        throw;
    }
    return c;
}

void tp_destroy(TpComm* c) {
    if (!c) return;
    if (c->comm) api().CommDestroy(c->comm);
    delete c;
}

void tp_allreduce_sum_f32(TpComm* c, float* buf, size_t count, cudaStream_t st) {
    check(api().AllReduce(buf, buf, count, NCCL_FLOAT32, NCCL_SUM, c->comm, st), "ncclAllReduce");
}

void tp_allreduce_sum_f16(TpComm* c, void* buf, size_t count, cudaStream_t st) {
    check(api().AllReduce(buf, buf, count, NCCL_FLOAT16, NCCL_SUM, c->comm, st), "ncclAllReduce");
}

void tp_allgather_bytes(TpComm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st) {
    check(api().AllGather(send, recv, bytes_per_rank, NCCL_INT8, c->comm, st), "ncclAllGather");
}

void tp_allgather_rows(TpComm* c, const void* send, size_t send_stride, void* recv, size_t recv_stride, size_t bytes_per_rank, int rows,
                       cudaStream_t st) {
    check(api().GroupStart(), "ncclGroupStart");
    for (int r = 0; r < rows; ++r)
        check(api().AllGather(static_cast<const char*>(send) + r * send_stride, static_cast<char*>(recv) + r * recv_stride, bytes_per_rank,
                              NCCL_INT8, c->comm, st), "ncclAllGather");
    check(api().GroupEnd(), "ncclGroupEnd");
}

int tp_rank(const TpComm* c) { return c->rank; }
int tp_size(const TpComm* c) { return c->size; }

void tp_ipc_exchange(TpComm* c, void* local, void** out_ptrs, cudaStream_t st) {
    cudaIpcMemHandle_t mine;
    RT_CUDA(cudaIpcGetMemHandle(&mine, local));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void *dsend = nullptr, *drecv = nullptr;
    RT_CUDA(cudaMalloc(&dsend, 64));
    RT_CUDA(cudaMalloc(&drecv, (size_t)64 * c->size));
    RT_CUDA(cudaMemcpyAsync(dsend, &mine, 64, cudaMemcpyHostToDevice, st));
    tp_allgather_bytes(c, dsend, drecv, 64, st);
    std::string all((size_t)64 * c->size, '\0');
    RT_CUDA(cudaMemcpyAsync(&all[0], drecv, all.size(), cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    cudaFree(dsend);
    cudaFree(drecv);
    for (int p = 0; p < c->size; ++p) {
        if (p == c->rank) {
            out_ptrs[p] = local;
            continue;
        }
        cudaIpcMemHandle_t h;
        memcpy(&h, &all[(size_t)64 * p], 64);
        const cudaError_t er = cudaIpcOpenMemHandle(&out_ptrs[p], h, cudaIpcMemLazyEnablePeerAccess);
        if (er != cudaSuccess) {
            cudaGetLastError();
            throw RtError(RT_ERR_CUDA, std::string("cudaIpcOpenMemHandle failed (peer access between the GPUs of the tensor-parallel group is required): ") +
                                           cudaGetErrorString(er));
        }
    }
}

void tp_ipc_close(TpComm* c, void** ptrs) {
    for (int p = 0; p < c->size; ++p)
        if (p != c->rank && ptrs[p]) cudaIpcCloseMemHandle(ptrs[p]);
}
