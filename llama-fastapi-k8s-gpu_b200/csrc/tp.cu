// Tensor-parallel collectives for 70B-class models (BASELINE.json configs[4]): one process per GPU, the
// quantised weights split by columns (attn_q/k/v, ffn_gate/up, output: whole rows per rank) or by rows' K ranges
// (attn_output, ffn_down: cut at 256-element block boundaries, no re-quantisation), and an all-reduce (sum) of the hidden
// vector after attn_output and after ffn_down, over NCCL on NVLink. The reference has no counterpart: its only multi-GPU
// strategy is independent replicas (/root/reference/helm/values.yaml:17); llama.cpp itself has no NCCL path.
//
// NCCL is loaded at run time (dlopen "libnccl.so.2"): a process that already carries a copy (torch.distributed) shares
// that one, a plain serving process picks up the system library, and a single-GPU host needs no NCCL at all.
// Only this GPU's buffers are ever passed to NCCL; there is no peer-memory access in this library's own kernels.
#include "kernels.h"

#include <dlfcn.h>
#include <nccl.h>

#include <mutex>
#include <string>

namespace b2 {

namespace {
struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;
    bool ok = false;
};

NcclApi& api() {
    static NcclApi a;
    static std::once_flag once;
    std::call_once(once, []() {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { a.err = std::string("cannot load NCCL: ") + dlerror(); return; }
        auto sym = [&](const char* n) { void* p = dlsym(h, n); if (!p) a.err = std::string("NCCL symbol missing: ") + n; return p; };
        a.GetUniqueId = (decltype(a.GetUniqueId))sym("ncclGetUniqueId");
        a.CommInitRank = (decltype(a.CommInitRank))sym("ncclCommInitRank");
        a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy");
        a.AllReduce = (decltype(a.AllReduce))sym("ncclAllReduce");
        a.AllGather = (decltype(a.AllGather))sym("ncclAllGather");
        a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
        a.ok = a.err.empty();
    });
    return a;
}

bool check(ncclResult_t r, const char* what, std::string& err) {
    if (r == ncclSuccess) return true;
    err = std::string(what) + ": " + (api().GetErrorString ? api().GetErrorString(r) : "NCCL error");
    return false;
}
}  // namespace

bool tp_unique_id(void* out128, std::string& err) {
    NcclApi& a = api();
    if (!a.ok) { err = a.err; return false; }
    static_assert(sizeof(ncclUniqueId) == 128, "unique id is 128 bytes in the C ABI");
    return check(a.GetUniqueId(reinterpret_cast<ncclUniqueId*>(out128)), "ncclGetUniqueId", err);
}

bool tp_comm_init(void** comm, int rank, int world, const void* id128, std::string& err) {
    NcclApi& a = api();
    if (!a.ok) { err = a.err; return false; }
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    ncclComm_t c = nullptr;
    if (!check(a.CommInitRank(&c, world, id, rank), "ncclCommInitRank", err)) return false;
    *comm = c;
    return true;
}

void tp_comm_destroy(void* comm) {
    if (comm && api().ok) api().CommDestroy((ncclComm_t)comm);
}

bool tp_all_reduce_sum(void* comm, const float* send, float* recv, size_t n, cudaStream_t s, std::string& err) {
    return check(api().AllReduce(send, recv, n, ncclFloat32, ncclSum, (ncclComm_t)comm, s), "ncclAllReduce", err);
}

// every rank contributes n floats; recv holds world * n (in place when send == recv + rank * n)
bool tp_all_gather(void* comm, const float* send, float* recv, size_t n, cudaStream_t s, std::string& err) {
    return check(api().AllGather(send, recv, n, ncclFloat32, (ncclComm_t)comm, s), "ncclAllGather", err);
}

}  // namespace b2
