// The one exchange step of a frame: SUM of the per-GPU film sums (camera.h:55, 117-124: the
// reference adds every sample of a pixel into one framebuffer entry) onto one GPU, over NCCL / NVLink, in place on
// the contexts' device films.  Two shapes:
//   bdpt_film_reduce(ctxs, n, root)        one process driving n devices (host/camera.cpp, `bdpt --gpus N`)
//   bdpt_comm_init + bdpt_film_reduce_comm  one process per GPU (torchrun: bench.py, tools/bench_frame.py); the
//                                           128-byte unique id travels through whatever the launcher has
// libnccl.so.2 is opened at first use (dlopen), not linked: a process that already holds an NCCL (torch's bundled
// copy) gets THAT library by soname, and a single-GPU user needs none.
#include <dlfcn.h>
#include <nccl.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "ctx.h"

using namespace bdpt;

namespace {

struct Nccl {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;
};

Nccl* nccl() {
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        n.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!n.lib) { n.err = std::string("libnccl.so.2: ") + dlerror(); return; }
        auto sym = [&](const char* name) {
            void* p = dlsym(n.lib, name);
            if (!p && n.err.empty()) n.err = std::string("libnccl.so.2 lacks ") + name;
            return p;
        };
        n.GetUniqueId = reinterpret_cast<decltype(n.GetUniqueId)>(sym("ncclGetUniqueId"));
        n.CommInitRank = reinterpret_cast<decltype(n.CommInitRank)>(sym("ncclCommInitRank"));
        n.CommInitAll = reinterpret_cast<decltype(n.CommInitAll)>(sym("ncclCommInitAll"));
        n.CommDestroy = reinterpret_cast<decltype(n.CommDestroy)>(sym("ncclCommDestroy"));
        n.Reduce = reinterpret_cast<decltype(n.Reduce)>(sym("ncclReduce"));
        n.GroupStart = reinterpret_cast<decltype(n.GroupStart)>(sym("ncclGroupStart"));
        n.GroupEnd = reinterpret_cast<decltype(n.GroupEnd)>(sym("ncclGroupEnd"));
        n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
    });
    return &n;
}

int fail(bdpt_ctx* ctx, const std::string& what) {
    if (ctx) ctx->err = what;
    return BDPT_ERR_CUDA;
}

#define BDPT_NCCL(ctx, call)                                                                            \
    do {                                                                                                \
        ncclResult_t r__ = (call);                                                                      \
        if (r__ != ncclSuccess) return fail(ctx, std::string(#call) + ": " + N->GetErrorString(r__));   \
    } while (0)

// communicators of bdpt_film_reduce, one set per device list, kept for the life of the process
std::mutex g_mu;
std::map<std::vector<int>, std::vector<ncclComm_t>> g_comms;

// communicators of bdpt_comm_init: one per device of this process, kept until the next bdpt_comm_init for that device;
// a context only borrows it (a new communicator costs a rendezvous plus ~2 s of lazy channel set-up at its first
// collective: an application joins once and renders many frames, possibly through many contexts)
struct ProcComm { ncclComm_t comm = nullptr; int rank = 0, size = 1; };
std::map<int, ProcComm> g_proc_comm;

}  // namespace

void bdpt::comm_free(bdpt_ctx* ctx) { ctx->comm = nullptr; }

extern "C" {

int bdpt_comm_unique_id(void* id128) {
    if (!id128) return BDPT_ERR_ARG;
    Nccl* N = nccl();
    if (!N->err.empty()) return BDPT_ERR_STATE;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    return N->GetUniqueId(static_cast<ncclUniqueId*>(id128)) == ncclSuccess ? BDPT_OK : BDPT_ERR_CUDA;
}

int bdpt_comm_init(bdpt_ctx* ctx, int nranks, int rank, const void* id128) {
    if (!ctx || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return BDPT_ERR_ARG;
    Nccl* N = nccl();
    if (!N->err.empty()) return fail(ctx, N->err);
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    std::lock_guard<std::mutex> lock(g_mu);
    ProcComm& pc = g_proc_comm[ctx->device];
    if (pc.comm) { N->CommDestroy(pc.comm); pc.comm = nullptr; }
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof id);
    ncclComm_t c = nullptr;
    BDPT_NCCL(ctx, N->CommInitRank(&c, nranks, id, rank));
    pc.comm = c; pc.rank = rank; pc.size = nranks;
    ctx->comm = c; ctx->comm_rank = rank; ctx->comm_size = nranks;
    return BDPT_OK;
}

int bdpt_comm_attach(bdpt_ctx* ctx) {
    if (!ctx) return BDPT_ERR_ARG;
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_proc_comm.find(ctx->device);
    if (it == g_proc_comm.end() || !it->second.comm) { ctx->err = "bdpt_comm_attach: no communicator was initialised for this device"; return BDPT_ERR_STATE; }
    ctx->comm = it->second.comm; ctx->comm_rank = it->second.rank; ctx->comm_size = it->second.size;
    return BDPT_OK;
}

int bdpt_film_reduce_comm(bdpt_ctx* ctx, int root) {
    if (!ctx) return BDPT_ERR_ARG;
    if (!ctx->comm) { ctx->err = "bdpt_film_reduce_comm: no communicator (bdpt_comm_init)"; return BDPT_ERR_STATE; }
    if (!ctx->film) { ctx->err = "no film"; return BDPT_ERR_STATE; }
    if (root < 0 || root >= ctx->comm_size) { ctx->err = "root outside the communicator"; return BDPT_ERR_ARG; }
    Nccl* N = nccl();
    BDPT_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t n = size_t(ctx->film_w) * ctx->film_h * 3;
    BDPT_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    BDPT_NCCL(ctx, N->Reduce(ctx->film, ctx->film, n, ncclDouble, ncclSum, root, static_cast<ncclComm_t>(ctx->comm), ctx->stream));
    BDPT_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    BDPT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_ms = ms;
    return BDPT_OK;
}

int bdpt_film_reduce(bdpt_ctx** ctxs, int n, int root) {
    if (!ctxs || n < 1 || root < 0 || root >= n) return BDPT_ERR_ARG;
    for (int i = 0; i < n; ++i) {
        if (!ctxs[i]) return BDPT_ERR_ARG;
        if (!ctxs[i]->film) { ctxs[i]->err = "no film"; return BDPT_ERR_STATE; }
        if (ctxs[i]->film_w != ctxs[0]->film_w || ctxs[i]->film_h != ctxs[0]->film_h) { ctxs[i]->err = "films differ in size"; return BDPT_ERR_ARG; }
    }
    if (n == 1) return BDPT_OK;
    bdpt_ctx* c0 = ctxs[root];
    Nccl* N = nccl();
    if (!N->err.empty()) return fail(c0, N->err);
    std::vector<int> devs(n);
    for (int i = 0; i < n; ++i) devs[i] = ctxs[i]->device;
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_comms.find(devs);
    if (it == g_comms.end()) {
        std::vector<ncclComm_t> comms(n, nullptr);
        BDPT_NCCL(c0, N->CommInitAll(comms.data(), n, devs.data()));
        it = g_comms.emplace(devs, comms).first;
    }
    const size_t cnt = size_t(c0->film_w) * c0->film_h * 3;
    BDPT_CUDA(c0, cudaSetDevice(c0->device));
    BDPT_CUDA(c0, cudaEventRecord(c0->ev0, c0->stream));
    BDPT_NCCL(c0, N->GroupStart());
    for (int i = 0; i < n; ++i) {
        ncclResult_t r = N->Reduce(ctxs[i]->film, ctxs[i]->film, cnt, ncclDouble, ncclSum, root, it->second[i], ctxs[i]->stream);
        if (r != ncclSuccess) { N->GroupEnd(); return fail(c0, std::string("ncclReduce: ") + N->GetErrorString(r)); }
    }
    BDPT_NCCL(c0, N->GroupEnd());
    BDPT_CUDA(c0, cudaSetDevice(c0->device));
    BDPT_CUDA(c0, cudaEventRecord(c0->ev1, c0->stream));
    for (int i = 0; i < n; ++i) {
        BDPT_CUDA(ctxs[i], cudaSetDevice(ctxs[i]->device));
        BDPT_CUDA(ctxs[i], cudaStreamSynchronize(ctxs[i]->stream));
    }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c0->ev0, c0->ev1);
    c0->last_ms = ms;
    return BDPT_OK;
}

}  // extern "C"
