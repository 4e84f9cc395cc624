// vr_noalloc.cu -- vr_transmittance_noalloc: vr_transmittance (gmm.h:517-578 on a batch of rays) for callers that cannot
// call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the
// reference.
//
// vr_transmittance stages the rays and the results in the context's scratch buffer, which grows with cudaMalloc.  Here
// the staging area is a file-scope device array of a fixed size, part of the library's image: every GPU that loads the
// library holds its kNoAllocBytes (2 MiB) until the process exits.  The entry point calls no cudaMalloc / cudaFree and
// no allocating library routine; it never rebuilds the culling hierarchy either (vr_transmittance does when a batch's
// origins lie farther out than the hierarchy was built for): such a batch is rejected with a clear error instead.
//
// In HOST code the array's name is a host-side stand-in, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, asked once per device with that device current (noalloc_scratch below).
//
// The first call on a device is the set-up call: it resolves the address and, with CUDA's lazy module loading, is when the
// driver materialises the array on that GPU (device memory in use grows by 2 MiB, once; measured in tests/test_zz_no_alloc.py).
// Make it outside any region that must not touch the allocator.
//
// Every call on a device shares the one array and finds the previous call's contents in it.  Nothing in it has to
// start at zero: the inputs are overwritten by this call's copies before the kernel reads them, and the kernel writes
// every one of the nrays outputs before they are copied out.  Calls that overlap on the same GPU (two contexts on two
// streams, two host threads) RACE on the array: serialise them.
#include <algorithm>
#include <cmath>
#include <mutex>
#include <string>

#include "vr_context.h"

using namespace vr;

namespace {

constexpr int kNoAllocMaxRays = VR_NOALLOC_MAX_RAYS;                  // include/volrender.h
constexpr size_t kNoAllocBytes = (size_t)kNoAllocMaxRays * 32;        // per ray: o (12) + d (12) + tmax (4) + out (4)
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mutex;
void* g_noalloc_ptr[kMaxDevices] = {nullptr};

// device address of g_noalloc_scratch on the CURRENT device (the caller has made ctx->device current)
int noalloc_scratch(vr_context* ctx, float** out) {
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return fail(ctx, VR_ERR_UNSUPPORTED, "vr_transmittance_noalloc: device index above 63");
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[ctx->device]) {
        void* p = nullptr;
        VR_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(ctx, VR_ERR_CUDA, "vr_transmittance_noalloc: cudaGetSymbolAddress returned no address");
        g_noalloc_ptr[ctx->device] = p;
    }
    *out = static_cast<float*>(g_noalloc_ptr[ctx->device]);
    return VR_OK;
}

}  // namespace

extern "C" int vr_transmittance_noalloc_max_rays(void) { return kNoAllocMaxRays; }

extern "C" int vr_transmittance_noalloc(vr_context* ctx, int nrays, const float* o3, const float* d3, const float* tmax, float* out) {
    if (!ctx || nrays < 0 || (nrays > 0 && (!o3 || !d3 || !tmax || !out)))
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_transmittance_noalloc: bad arguments");
    if (nrays > kNoAllocMaxRays)
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_transmittance_noalloc: " + std::to_string(nrays) + " rays, at most " + std::to_string(kNoAllocMaxRays) +
                                                      " per call (VR_NOALLOC_MAX_RAYS: the scratch array is fixed); split the batch or use vr_transmittance");
    if (!ctx->has_gmm) return fail(ctx, VR_ERR_NO_SCENE, "vr_transmittance_noalloc: no Gaussian mixture set");
    if (nrays == 0) return VR_OK;
    if (ctx->n > 0) {   // reach of this batch's origins (vr_bvh.h BoxReach): checked, never rebuilt for
        double R2 = 0.0, X = 0.0;
        for (int i = 0; i < nrays; ++i) {
            double d2 = 0.0;
            for (int k = 0; k < 3; ++k) {
                const double o = (double)o3[3 * (size_t)i + k];
                const double d = std::max(std::fabs(o - (double)ctx->bb_lo[k]), std::fabs(o - (double)ctx->bb_hi[k]));
                d2 += d * d;
                X = std::max(X, std::fabs(o));
            }
            R2 = std::max(R2, d2);
        }
        if (!(R2 < 1e60) || !(X < 1e30)) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_transmittance_noalloc: ray origin is not finite");
        if ((float)std::sqrt(R2) > ctx->built_R || (float)X > ctx->built_X)
            return fail(ctx, VR_ERR_UNSUPPORTED, "vr_transmittance_noalloc: a ray origin lies farther from the mixture than the culling hierarchy was built for; "
                                                 "rebuilding allocates -- call vr_transmittance once with such a batch (or set the camera there) first");
    }
    VR_CUDA(cudaSetDevice(ctx->device));
    float* d = nullptr;
    int rc = noalloc_scratch(ctx, &d);
    if (rc) return rc;
    const size_t n = (size_t)nrays;
    float *d_o = d, *d_d = d + 3 * n, *d_t = d + 6 * n, *d_out = d + 7 * n;
    cudaStream_t s = ctx->stream;
    VR_CUDA(cudaMemcpyAsync(d_o, o3, n * 12, cudaMemcpyHostToDevice, s));
    VR_CUDA(cudaMemcpyAsync(d_d, d3, n * 12, cudaMemcpyHostToDevice, s));
    VR_CUDA(cudaMemcpyAsync(d_t, tmax, n * 4, cudaMemcpyHostToDevice, s));
    launch_transmittance(ctx->scene_dev(), nrays, d_o, d_d, d_t, d_out, s);   // the kernel vr_transmittance launches
    VR_CUDA(cudaGetLastError());
    VR_CUDA(cudaMemcpyAsync(out, d_out, n * 4, cudaMemcpyDeviceToHost, s));
    VR_CUDA(cudaStreamSynchronize(s));
    return VR_OK;
}
