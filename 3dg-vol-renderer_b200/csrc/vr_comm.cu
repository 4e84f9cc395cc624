// vr_comm.cu -- the one collective of the path: SUM all-reduce of the per-Gaussian loss differences of an SFD iteration
// (inverse_integrator.h:157-198: F_diff and the base loss are sums over pixels, hence over devices and views).
//
// Every context accumulates its share of F_diff[S x N] in int64 fixed point (2^-32 units, vr_inverse.cu), so the sum over
// devices is exact and independent of the order NCCL adds in.  vr_sfd_allreduce* enqueue ONE ncclAllReduce(SUM, int64,
// S * N + 2) on the context's own stream directly behind the last reduction kernel -- no host synchronisation, no copy to
// the host -- and every device then applies the same Adam step (vr_sfd_apply with F_diff = NULL).
//
// Two ways to form the communicator, mirroring how the contexts are owned:
//   vr_comm_init_all(ctxs, n)          one process driving n devices (the C++ drop-in API, VR_DEVICES): ncclCommInitAll
//   vr_comm_unique_id / _init_rank     one process per device (bench.py under torchrun): the 128-byte id travels by whatever
//                                      the launcher offers (torch.distributed broadcast in bench.py)
// NCCL is bound at run time (dlopen "libnccl.so.2"): the library loads on a machine without it, a process that has
// another NCCL loaded (PyTorch) shares that copy instead of bringing a second one, and a build needs only <nccl.h>.
// Contexts that share one device (VR_DEVICES="0,0", used by tests) cannot form an NCCL communicator; they, and a machine
// without NCCL, take the same sum through peer copies and an integer add kernel -- bit-identical by construction.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <mutex>
#include <set>
#include <string>
#include <vector>

#include "vr_context.h"

namespace vr {

struct NcclApi {
    void* handle = nullptr;
    bool ok = false;
    std::string why;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
};

static NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) { api.why = std::string("NCCL not found: ") + (dlerror() ? dlerror() : "dlopen failed"); return; }
        auto sym = [&](const char* n) { void* p = dlsym(api.handle, n); if (!p && api.why.empty()) api.why = std::string("NCCL symbol missing: ") + n; return p; };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
        api.ok = api.why.empty();
    });
    return api;
}

static int nccl_fail(vr_context* ctx, ncclResult_t r, const char* what) {
    NcclApi& N = nccl();
    return fail(ctx, VR_ERR_CUDA, std::string(what) + ": " + (N.GetErrorString ? N.GetErrorString(r) : "NCCL error"));
}
#define VR_NCCL(ctx_, call)                                          \
    do {                                                             \
        ncclResult_t _r = (call);                                    \
        if (_r != ncclSuccess) return nccl_fail(ctx_, _r, #call);    \
    } while (0)

void comm_destroy(vr_context* ctx) {
    if (ctx && ctx->comm) {
        NcclApi& N = nccl();
        if (N.ok) N.CommDestroy((ncclComm_t)ctx->comm);
        ctx->comm = nullptr;
    }
    if (ctx) { ctx->comm_rank = 0; ctx->comm_size = 1; }
}

__global__ void add_i64_kernel(long long* __restrict__ dst, const long long* __restrict__ src, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] += src[i];
}

}  // namespace vr

using namespace vr;

extern "C" {

int vr_comm_available(void) { return nccl().ok ? 1 : 0; }

int vr_comm_version(void) {
    NcclApi& N = nccl();
    int v = 0;
    if (N.ok && N.GetVersion) N.GetVersion(&v);
    return v;
}

int vr_comm_unique_id(void* id128) {
    if (!id128) return VR_ERR_INVALID_ARGUMENT;
    NcclApi& N = nccl();
    if (!N.ok) return VR_ERR_UNSUPPORTED;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (N.GetUniqueId(&id) != ncclSuccess) return VR_ERR_CUDA;
    memcpy(id128, &id, sizeof(id));
    return VR_OK;
}

int vr_comm_init_rank(vr_context* ctx, const void* id128, int nranks, int rank) {
    if (!ctx || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_comm_init_rank: bad arguments");
    NcclApi& N = nccl();
    if (!N.ok) return fail(ctx, VR_ERR_UNSUPPORTED, N.why);
    VR_CUDA(cudaSetDevice(ctx->device));
    comm_destroy(ctx);
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    ncclComm_t c = nullptr;
    VR_NCCL(ctx, N.CommInitRank(&c, nranks, id, rank));
    ctx->comm = c; ctx->comm_rank = rank; ctx->comm_size = nranks;
    return VR_OK;
}

int vr_comm_init_all(vr_context** ctxs, int n) {
    if (!ctxs || n < 1) return VR_ERR_INVALID_ARGUMENT;
    for (int i = 0; i < n; ++i) if (!ctxs[i]) return VR_ERR_INVALID_ARGUMENT;
    for (int i = 0; i < n; ++i) comm_destroy(ctxs[i]);
    if (n == 1) return VR_OK;
    std::set<int> distinct;
    std::vector<int> devs((size_t)n);
    for (int i = 0; i < n; ++i) { devs[(size_t)i] = ctxs[i]->device; distinct.insert(ctxs[i]->device); }
    NcclApi& N = nccl();
    // contexts sharing a device, or no NCCL on this machine: vr_sfd_allreduce_all sums through peer copies instead
    if ((int)distinct.size() != n || !N.ok) return VR_OK;
    std::vector<ncclComm_t> comms((size_t)n, nullptr);
    VR_NCCL(ctxs[0], N.CommInitAll(comms.data(), n, devs.data()));
    for (int i = 0; i < n; ++i) { ctxs[i]->comm = comms[(size_t)i]; ctxs[i]->comm_rank = i; ctxs[i]->comm_size = n; }
    return VR_OK;
}

int vr_comm_destroy(vr_context* ctx) {
    if (!ctx) return VR_ERR_INVALID_ARGUMENT;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    comm_destroy(ctx);
    return VR_OK;
}

int vr_comm_size(const vr_context* ctx) { return ctx ? (ctx->comm ? ctx->comm_size : 1) : 0; }

// one rank of a multi-process communicator: in place, on the context's stream, asynchronous
int vr_sfd_allreduce(vr_context* ctx) {
    if (!ctx) return VR_ERR_INVALID_ARGUMENT;
    long long* buf = nullptr; size_t count = 0;
    int rc = sfd_reduce_buffer(ctx, &buf, &count);
    if (rc) return rc;
    if (!ctx->comm) return VR_OK;                      // no communicator: a single context, nothing to sum
    NvtxRange range("vr.sfd.allreduce");
    NcclApi& N = nccl();
    VR_CUDA(cudaSetDevice(ctx->device));
    VR_NCCL(ctx, N.AllReduce(buf, buf, count, ncclInt64, ncclSum, (ncclComm_t)ctx->comm, ctx->stream));
    return VR_OK;
}

// all contexts of one process: a single NCCL group, each member on its own stream; or peer copies (see the file header)
int vr_sfd_allreduce_all(vr_context** ctxs, int n) {
    if (!ctxs || n < 1) return VR_ERR_INVALID_ARGUMENT;
    std::vector<long long*> buf((size_t)n);
    size_t count = 0;
    for (int i = 0; i < n; ++i) {
        if (!ctxs[i]) return VR_ERR_INVALID_ARGUMENT;
        size_t c = 0;
        int rc = sfd_reduce_buffer(ctxs[i], &buf[(size_t)i], &c);
        if (rc) return rc;
        if (i && c != count) return fail(ctxs[i], VR_ERR_INVALID_ARGUMENT, "vr_sfd_allreduce_all: contexts hold different S or N");
        count = c;
    }
    if (n == 1) return VR_OK;
    NvtxRange range("vr.sfd.allreduce_all");
    bool have_comm = true;
    for (int i = 0; i < n; ++i) have_comm = have_comm && ctxs[i]->comm && ctxs[i]->comm_size == n;
    if (have_comm) {
        NcclApi& N = nccl();
        VR_NCCL(ctxs[0], N.GroupStart());
        for (int i = 0; i < n; ++i) {
            ncclResult_t r = N.AllReduce(buf[(size_t)i], buf[(size_t)i], count, ncclInt64, ncclSum, (ncclComm_t)ctxs[i]->comm, ctxs[i]->stream);
            if (r != ncclSuccess) { N.GroupEnd(); return nccl_fail(ctxs[i], r, "ncclAllReduce"); }
        }
        VR_NCCL(ctxs[0], N.GroupEnd());
        return VR_OK;
    }
    // peer path: gather every partial sum next to context 0's, add there, copy the total back (exact: integers)
    vr_context* c0 = ctxs[0];
    vr_context* ctx = c0;
    VR_CUDA(cudaSetDevice(c0->device));
    VR_CUDA(c0->comm_scratch.ensure(count * 8));
    for (int i = 1; i < n; ++i) {
        VR_CUDA(cudaSetDevice(ctxs[i]->device));
        VR_CUDA(cudaStreamSynchronize(ctxs[i]->stream));                 // partial sums of context i are complete
        VR_CUDA(cudaSetDevice(c0->device));
        VR_CUDA(cudaMemcpyPeerAsync(c0->comm_scratch.p, c0->device, buf[(size_t)i], ctxs[i]->device, count * 8, c0->stream));
        add_i64_kernel<<<(unsigned)((count + 255) / 256), 256, 0, c0->stream>>>(buf[0], c0->comm_scratch.as<long long>(), count);
    }
    VR_CUDA(cudaGetLastError());
    VR_CUDA(cudaStreamSynchronize(c0->stream));
    for (int i = 1; i < n; ++i) {
        VR_CUDA(cudaSetDevice(ctxs[i]->device));
        VR_CUDA(cudaMemcpyPeerAsync(buf[(size_t)i], ctxs[i]->device, buf[0], c0->device, count * 8, ctxs[i]->stream));
    }
    return VR_OK;
}

}  // extern "C"
