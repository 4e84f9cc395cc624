// vr_inverse.cu -- StochasticFiniteDiffInverseIntegrator::optimize building blocks
// (inverse_integrator.h:67-241) and AdamOptimizer::step (optimizer.h:30-41).
//
// Per-pixel contributing-Gaussian sets are bitmaps written by the path kernel; the localised loss
// difference F_diff[g] = sum over pixels p in (P_base[g] U P_plus[g]) of loss_plus[p] - loss_base[p]
// (inverse_integrator.h:157-179) is accumulated with 64-bit FIXED-POINT atomics (2^-32 resolution):
// integer addition is associative, so the result is bit-identical for any thread order, any
// row partition and any number of GPUs, and the cross-GPU all-reduce of the int64 buffer is exact.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "vr_context.h"

namespace vr {

constexpr double kFixedScale = 4294967296.0;  // 2^32

struct SfdState {
    int n = 0, t = 0;
    float lr = 1e-2f, beta1 = 0.9f, beta2 = 0.999f, adam_eps = 1e-8f;
    int tw = 0, th = 0;
    // fdiff_fixed: S * n fixed-point partial sums FOLLOWED by two int64 of loss accumulators, so that one all-reduce covers both
    DevBuf params, params_plus, eps, m, v, grad, target, loss_base, loss_plus, bits_base, fdiff_fixed, signs, fdiff_d;
    std::vector<DevBuf> view_targets;   // targets of a multi-view fit kept resident (vr_sfd_set_target_view / vr_sfd_select_view)
    int view = -1;                      // -1: `target` (vr_sfd_set_target)
    const float* target_ptr() const { return view >= 0 ? view_targets[(size_t)view].as<float>() : target.as<float>(); }
    int fdiff_S = 0;
    long long* loss_acc() const { return fdiff_fixed.as<long long>() + (size_t)fdiff_S * n; }
};

// the buffer one all-reduce per iteration sums over devices (vr_comm.cu): S * n partial sums + the two loss accumulators
int sfd_reduce_buffer(vr_context* c, long long** buf, size_t* count) {
    if (!c || !c->sfd || c->sfd->fdiff_S <= 0) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_sfd_allreduce: nothing accumulated (vr_sfd_reset_fdiff / vr_sfd_accumulate first)");
    *buf = c->sfd->fdiff_fixed.as<long long>();
    *count = (size_t)c->sfd->fdiff_S * c->sfd->n + 2;
    return VR_OK;
}

void sfd_destroy(vr_context* c) {
    if (!c->sfd) return;
    SfdState* s = c->sfd;
    for (DevBuf* b : {&s->params, &s->params_plus, &s->eps, &s->m, &s->v, &s->grad, &s->target, &s->loss_base, &s->loss_plus,
                      &s->bits_base, &s->fdiff_fixed, &s->signs, &s->fdiff_d})
        b->release();
    for (DevBuf& b : s->view_targets) b.release();
    delete s;
    c->sfd = nullptr;
}

struct OwnedPixels { int W, x0, y0, win_w, rows, part_rank, part_count; };
__device__ __forceinline__ size_t owned_pixel(const OwnedPixels& o, int i) {
    const int x = o.x0 + i % o.win_w;
    const int y = o.y0 + (i / o.win_w) * o.part_count + o.part_rank;
    return (size_t)y * o.W + x;
}

// compute_pixel_losses (inverse_integrator.h:20-29): L1 over RGB, per owned pixel
__global__ void pixel_loss_kernel(OwnedPixels o, const float* __restrict__ img, const float* __restrict__ ref,
                                  float* __restrict__ loss, long long* acc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    long long fx = 0;
    if (i < o.win_w * o.rows) {
        const size_t p = owned_pixel(o, i);
        const float d0 = __fsub_rn(img[3 * p], ref[3 * p]), d1 = __fsub_rn(img[3 * p + 1], ref[3 * p + 1]),
                    d2 = __fsub_rn(img[3 * p + 2], ref[3 * p + 2]);
        const float l = __fadd_rn(__fadd_rn(fabsf(d0), fabsf(d1)), fabsf(d2));
        loss[p] = l;
        fx = __double2ll_rn((double)l * kFixedScale);
    }
    if (acc) {
        for (int s = 16; s > 0; s >>= 1) fx += __shfl_down_sync(0xffffffffu, fx, s);
        if ((threadIdx.x & 31) == 0 && fx) atomicAdd((unsigned long long*)acc, (unsigned long long)fx);
    }
}

// one thread per (owned pixel, bitmap word): union of the two sets, fixed-point accumulation
__global__ void fdiff_kernel(OwnedPixels o, int words, int n, const uint32_t* __restrict__ bits_base,
                             const uint32_t* __restrict__ bits_plus, const float* __restrict__ loss_base,
                             const float* __restrict__ loss_plus, long long* __restrict__ F) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)o.win_w * o.rows * words) return;
    const int i = (int)(idx / words), w = (int)(idx % words);
    const size_t p = owned_pixel(o, i);
    uint32_t u = bits_base[p * words + w] | bits_plus[p * words + w];
    if (!u) return;
    const long long d = __double2ll_rn((double)loss_plus[p] * kFixedScale) - __double2ll_rn((double)loss_base[p] * kFixedScale);
    if (!d) return;
    while (u) {
        const int b = __ffs(u) - 1;
        u &= u - 1;
        const int g = w * 32 + b;
        if (g < n) atomicAdd((unsigned long long*)(F + g), (unsigned long long)d);
    }
}

static void launch_fdiff(const OwnedPixels& o, int words, int n, const uint32_t* bits_base, const uint32_t* bits_plus, const float* loss_base,
                         const float* loss_plus, long long* F, cudaStream_t st) {
    const long long total = (long long)o.win_w * o.rows * words;
    if (total > 0) fdiff_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(o, words, n, bits_base, bits_plus, loss_base, loss_plus, F);
}

__global__ void perturb_kernel(int D, const float* __restrict__ params, const signed char* __restrict__ s,
                               const float* __restrict__ eps, float* __restrict__ out) {  // inverse_integrator.h:146-148
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < D) out[i] = __fadd_rn(params[i], __fmul_rn((float)s[i], eps[i]));
}

// grad[i] = (1/S) sum_s F_diff[s][i/11] * s[i] / eps[i]  (inverse_integrator.h:182-198), double accumulate
__global__ void sfd_grad_kernel(int D, int n, int S, const double* __restrict__ F, const signed char* __restrict__ signs,
                                const float* __restrict__ eps, float* __restrict__ grad) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D) return;
    const double den = (double)eps[i];
    double acc = 0.0;
    if (!(fabs(den) < 1e-12))
        for (int s = 0; s < S; ++s) acc = __dadd_rn(acc, __ddiv_rn(__dmul_rn(F[(size_t)s * n + i / 11], (double)signs[(size_t)s * D + i]), den));
    grad[i] = (float)__ddiv_rn(acc, (double)S);
}
__global__ void fixed_to_double_kernel(int n, const long long* __restrict__ f, double* __restrict__ d) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) d[i] = (double)f[i] / kFixedScale;
}

// AdamOptimizer::step (optimizer.h:30-41); a = lr*sqrt(1-b2^t)/(1-b1^t) is computed by the caller
__global__ void adam_kernel(int D, float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, float a, float b1, float b2, float eps) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= D) return;
    const float gi = g[i];
    const float mi = __fadd_rn(__fmul_rn(b1, m[i]), __fmul_rn(__fsub_rn(1.0f, b1), gi));
    const float vi = __fadd_rn(__fmul_rn(b2, v[i]), __fmul_rn(__fmul_rn(__fsub_rn(1.0f, b2), gi), gi));
    m[i] = mi; v[i] = vi;
    p[i] = __fsub_rn(p[i], __fmul_rn(a, __fdiv_rn(mi, __fadd_rn(__fsqrt_rn(vi), eps))));
}
static float adam_alpha(float lr, float b1, float b2, int t) {
    return lr * std::sqrt(1.0f - std::pow(b2, (float)t)) / (1.0f - std::pow(b1, (float)t));
}

static OwnedPixels owned_of(const vr_render_params* p) {
    OwnedPixels o;
    o.W = p->width; o.x0 = p->x0; o.y0 = p->y0; o.win_w = p->x1 - p->x0;
    const int rows = p->y1 - p->y0;
    o.rows = p->part_rank >= rows ? 0 : (rows - p->part_rank + p->part_count - 1) / p->part_count;
    o.part_rank = p->part_rank; o.part_count = p->part_count;
    return o;
}

static int set_scene_from_params(vr_context* ctx, const float* d_params, int n) {
    const size_t m = (size_t)std::max(n, 1);
    VR_CUDA(ctx->mean3.ensure(m * 12)); VR_CUDA(ctx->cov6.ensure(m * 24));
    VR_CUDA(ctx->density.ensure(m * 4)); VR_CUDA(ctx->albedo.ensure(m * 4));
    launch_unpack_params(n, d_params, ctx->mean3.as<float>(), ctx->cov6.as<float>(), ctx->density.as<float>(),
                         ctx->albedo.as<float>(), ctx->stream);
    return set_gmm_from_device(ctx, n);
}

}  // namespace vr

using namespace vr;

extern "C" {

int vr_sfd_begin(vr_context* ctx, int n, const float* params11, const float* eps11, float lr, float beta1,
                 float beta2, float adam_eps) {
    if (!ctx || n <= 0 || !params11 || !eps11) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_begin: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    if (!ctx->sfd) ctx->sfd = new SfdState();
    SfdState* s = ctx->sfd;
    const size_t D = (size_t)11 * n;
    s->n = n; s->t = 0; s->lr = lr; s->beta1 = beta1; s->beta2 = beta2; s->adam_eps = adam_eps;
    for (DevBuf* b : {&s->params, &s->params_plus, &s->eps, &s->m, &s->v, &s->grad}) VR_CUDA(b->ensure(D * 4));
    VR_CUDA(cudaMemcpyAsync(s->params.p, params11, D * 4, cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(s->eps.p, eps11, D * 4, cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaMemsetAsync(s->m.p, 0, D * 4, ctx->stream));
    VR_CUDA(cudaMemsetAsync(s->v.p, 0, D * 4, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    s->fdiff_S = 0;
    return VR_OK;
}

int vr_sfd_set_target(vr_context* ctx, int width, int height, const float* target_rgb) {
    if (!ctx || !ctx->sfd || width <= 0 || height <= 0 || !target_rgb) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_set_target: bad arguments (vr_sfd_begin first)");
    VR_CUDA(cudaSetDevice(ctx->device));
    SfdState* s = ctx->sfd;
    const size_t bytes = (size_t)3 * width * height * 4;
    VR_CUDA(s->target.ensure(bytes));
    VR_CUDA(cudaMemcpyAsync(s->target.p, target_rgb, bytes, cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    s->tw = width; s->th = height; s->view = -1;
    return VR_OK;
}

// multi-view fits: every view's target is uploaded ONCE and stays on the device; vr_sfd_select_view picks the one the next
// vr_sfd_accumulate compares against (all views share one image size)
int vr_sfd_set_target_view(vr_context* ctx, int view, int width, int height, const float* target_rgb) {
    if (!ctx || !ctx->sfd || view < 0 || view >= 4096 || width <= 0 || height <= 0 || !target_rgb)
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_set_target_view: bad arguments (vr_sfd_begin first)");
    VR_CUDA(cudaSetDevice(ctx->device));
    SfdState* s = ctx->sfd;
    if (!s->view_targets.empty() && (s->tw != width || s->th != height) && s->view >= 0)
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_set_target_view: all views must share one image size");
    if ((size_t)view >= s->view_targets.size()) s->view_targets.resize((size_t)view + 1);
    const size_t bytes = (size_t)3 * width * height * 4;
    VR_CUDA(s->view_targets[(size_t)view].ensure(bytes));
    VR_CUDA(cudaMemcpyAsync(s->view_targets[(size_t)view].p, target_rgb, bytes, cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    s->tw = width; s->th = height; s->view = view;
    return VR_OK;
}
int vr_sfd_select_view(vr_context* ctx, int view) {
    if (!ctx || !ctx->sfd || view < 0 || (size_t)view >= ctx->sfd->view_targets.size() || !ctx->sfd->view_targets[(size_t)view].p)
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_select_view: no target was set for this view");
    ctx->sfd->view = view;
    return VR_OK;
}

int vr_sfd_reset_fdiff(vr_context* ctx, int S) {
    if (!ctx || !ctx->sfd || S <= 0) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_reset_fdiff: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    SfdState* s = ctx->sfd;
    VR_CUDA(s->fdiff_fixed.ensure(((size_t)S * s->n + 2) * 8));
    VR_CUDA(cudaMemsetAsync(s->fdiff_fixed.p, 0, ((size_t)S * s->n + 2) * 8, ctx->stream));
    s->fdiff_S = S;
    return VR_OK;
}

int64_t* vr_sfd_fdiff_dev(vr_context* ctx) { return (ctx && ctx->sfd) ? ctx->sfd->fdiff_fixed.as<int64_t>() : nullptr; }

int vr_sfd_accumulate(vr_context* ctx, const vr_render_params* rp, int S, const int8_t* signs) {
    if (!ctx || !ctx->sfd || !rp || S <= 0 || !signs) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_accumulate: bad arguments");
    SfdState* s = ctx->sfd;
    if (s->fdiff_S != S) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_accumulate: call vr_sfd_reset_fdiff(S) first");
    if (s->tw != rp->width || s->th != rp->height) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_accumulate: target image size differs from the render size");
    NvtxRange range_acc("vr.sfd.accumulate");
    VR_CUDA(cudaSetDevice(ctx->device));
    const int n = s->n, D = 11 * n;
    cudaStream_t st = ctx->stream;
    vr_render_params p = *rp;
    p.integrator = VR_INTEGRATOR_MULTI_SCATTER;
    p.record = 1;
    const OwnedPixels o = owned_of(&p);
    const int npix_owned = o.win_w * o.rows;
    const size_t npix = (size_t)p.width * p.height;
    VR_CUDA(s->signs.ensure((size_t)S * D));
    VR_CUDA(cudaMemcpyAsync(s->signs.p, signs, (size_t)S * D, cudaMemcpyHostToDevice, st));
    VR_CUDA(s->loss_base.ensure(npix * 4)); VR_CUDA(s->loss_plus.ensure(npix * 4));
    // ---- base render with recording (inverse_integrator.h:119-130)
    int rc = set_scene_from_params(ctx, s->params.as<float>(), n);
    if (rc) return rc;
    rc = render_device_impl(ctx, &p, nullptr);
    if (rc) return rc;
    const int words = ctx->rec_words;
    const size_t bits_bytes = npix * (size_t)words * 4;
    VR_CUDA(s->bits_base.ensure(bits_bytes));
    VR_CUDA(cudaMemcpyAsync(s->bits_base.p, ctx->record_bits.p, bits_bytes, cudaMemcpyDeviceToDevice, st));
    if (npix_owned > 0)
        pixel_loss_kernel<<<(npix_owned + 127) / 128, 128, 0, st>>>(o, ctx->image.as<float>(), s->target_ptr(),
                                                                     s->loss_base.as<float>(), s->loss_acc());
    // ---- S perturbed renders (inverse_integrator.h:140-188)
    for (int k = 0; k < S; ++k) {
        perturb_kernel<<<(D + 255) / 256, 256, 0, st>>>(D, s->params.as<float>(), s->signs.as<signed char>() + (size_t)k * D,
                                                         s->eps.as<float>(), s->params_plus.as<float>());
        rc = set_scene_from_params(ctx, s->params_plus.as<float>(), n);
        if (rc) return rc;
        rc = render_device_impl(ctx, &p, nullptr);
        if (rc) return rc;
        if (npix_owned > 0) {
            pixel_loss_kernel<<<(npix_owned + 127) / 128, 128, 0, st>>>(o, ctx->image.as<float>(), s->target_ptr(),
                                                                         s->loss_plus.as<float>(), nullptr);
            launch_fdiff(o, words, n, s->bits_base.as<uint32_t>(), ctx->record_bits.as<uint32_t>(), s->loss_base.as<float>(),
                         s->loss_plus.as<float>(), s->fdiff_fixed.as<long long>() + (size_t)k * n, st);
        }
    }
    VR_CUDA(cudaGetLastError());   // no synchronisation: vr_sfd_allreduce* / vr_sfd_apply / vr_sfd_read follow on the same stream
    return VR_OK;
}

int vr_sfd_read(vr_context* ctx, double* F_diff, double* loss_sum) {
    if (!ctx || !ctx->sfd || ctx->sfd->fdiff_S <= 0) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_read: nothing accumulated");
    VR_CUDA(cudaSetDevice(ctx->device));
    SfdState* s = ctx->sfd;
    const size_t cnt = (size_t)s->fdiff_S * s->n;
    if (F_diff) {
        std::vector<long long> h(cnt);
        VR_CUDA(cudaMemcpyAsync(h.data(), s->fdiff_fixed.p, cnt * 8, cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        for (size_t i = 0; i < cnt; ++i) F_diff[i] = (double)h[i] / kFixedScale;
    }
    if (loss_sum) {
        long long l = 0;
        VR_CUDA(cudaMemcpyAsync(&l, s->loss_acc(), 8, cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        *loss_sum = (double)l / kFixedScale;
    }
    return VR_OK;
}

int vr_sfd_fdiff(vr_context* ctx, const vr_render_params* rp, int S, const int8_t* signs, double* F_diff,
                 double* loss_sum, int64_t* loss_pixels) {
    int rc = vr_sfd_reset_fdiff(ctx, S);
    if (rc) return rc;
    rc = vr_sfd_accumulate(ctx, rp, S, signs);
    if (rc) return rc;
    if (loss_pixels) { const OwnedPixels o = owned_of(rp); *loss_pixels = (int64_t)o.win_w * o.rows; }
    return vr_sfd_read(ctx, F_diff, loss_sum);
}

int vr_sfd_apply(vr_context* ctx, int S, const int8_t* signs, const double* F_diff, float* grad_out) {
    if (!ctx || !ctx->sfd || S <= 0 || !signs) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_apply: bad arguments");
    NvtxRange range_apply("vr.sfd.apply");
    VR_CUDA(cudaSetDevice(ctx->device));
    SfdState* s = ctx->sfd;
    const int n = s->n, D = 11 * n;
    cudaStream_t st = ctx->stream;
    VR_CUDA(s->signs.ensure((size_t)S * D));
    VR_CUDA(cudaMemcpyAsync(s->signs.p, signs, (size_t)S * D, cudaMemcpyHostToDevice, st));
    VR_CUDA(s->fdiff_d.ensure((size_t)S * n * 8));
    if (F_diff) {
        VR_CUDA(cudaMemcpyAsync(s->fdiff_d.p, F_diff, (size_t)S * n * 8, cudaMemcpyHostToDevice, st));
    } else {  // the (possibly all-reduced) device fixed-point buffer
        if (s->fdiff_S != S) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_apply: no accumulated F_diff for this S");
        fixed_to_double_kernel<<<(S * n + 255) / 256, 256, 0, st>>>(S * n, s->fdiff_fixed.as<long long>(), s->fdiff_d.as<double>());
    }
    sfd_grad_kernel<<<(D + 255) / 256, 256, 0, st>>>(D, n, S, s->fdiff_d.as<double>(), s->signs.as<signed char>(), s->eps.as<float>(),
                                                      s->grad.as<float>());
    s->t += 1;
    const float a = adam_alpha(s->lr, s->beta1, s->beta2, s->t);
    adam_kernel<<<(D + 255) / 256, 256, 0, st>>>(D, s->params.as<float>(), s->grad.as<float>(), s->m.as<float>(), s->v.as<float>(), a,
                                                  s->beta1, s->beta2, s->adam_eps);
    VR_CUDA(cudaGetLastError());
    if (grad_out) VR_CUDA(cudaMemcpyAsync(grad_out, s->grad.p, (size_t)D * 4, cudaMemcpyDeviceToHost, st));
    VR_CUDA(cudaStreamSynchronize(st));
    return VR_OK;
}

// known-answer hook of the localised loss difference (inverse_integrator.h:157-179): caller-supplied recorded sets (bitmaps, one
// row of `words` u32 per pixel, bit g = Gaussian g) and per-pixel losses -> F_diff[n] in int64 fixed point (2^-32 units)
int vr_kat_fdiff(vr_context* ctx, int npix, int n, const uint32_t* bits_base, const uint32_t* bits_plus, const float* loss_base,
                 const float* loss_plus, int64_t* F_fixed) {
    if (!ctx || npix <= 0 || n <= 0 || !bits_base || !bits_plus || !loss_base || !loss_plus || !F_fixed)
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_fdiff: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    const int words = (n + 31) / 32;
    const size_t bb = (size_t)npix * words * 4, lb = (size_t)npix * 4, fb = (size_t)n * 8;
    VR_CUDA(ctx->scratch.ensure(2 * bb + 2 * lb + fb + 64));
    char* b = ctx->scratch.as<char>();
    long long* dF = (long long*)b; uint32_t* db0 = (uint32_t*)(b + ((fb + 15) & ~(size_t)15)); uint32_t* db1 = db0 + (size_t)npix * words;
    float* dl0 = (float*)(db1 + (size_t)npix * words); float* dl1 = dl0 + npix;
    cudaStream_t st = ctx->stream;
    VR_CUDA(cudaMemcpyAsync(db0, bits_base, bb, cudaMemcpyHostToDevice, st));
    VR_CUDA(cudaMemcpyAsync(db1, bits_plus, bb, cudaMemcpyHostToDevice, st));
    VR_CUDA(cudaMemcpyAsync(dl0, loss_base, lb, cudaMemcpyHostToDevice, st));
    VR_CUDA(cudaMemcpyAsync(dl1, loss_plus, lb, cudaMemcpyHostToDevice, st));
    VR_CUDA(cudaMemsetAsync(dF, 0, fb, st));
    OwnedPixels o; o.W = npix; o.x0 = 0; o.y0 = 0; o.win_w = npix; o.rows = 1; o.part_rank = 0; o.part_count = 1;
    launch_fdiff(o, words, n, db0, db1, dl0, dl1, dF, st);
    VR_CUDA(cudaGetLastError());
    VR_CUDA(cudaMemcpyAsync(F_fixed, dF, fb, cudaMemcpyDeviceToHost, st));
    VR_CUDA(cudaStreamSynchronize(st));
    return VR_OK;
}

// checkpoint / resume of the optimiser: parameters, Adam moments and step count
int vr_sfd_get_state(vr_context* ctx, float* params11, float* m, float* v, int* t) {
    if (!ctx || !ctx->sfd || !params11 || !m || !v || !t) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_get_state: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    SfdState* s = ctx->sfd;
    const size_t bytes = (size_t)11 * s->n * 4;
    VR_CUDA(cudaMemcpyAsync(params11, s->params.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(m, s->m.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(v, s->v.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    *t = s->t;
    return VR_OK;
}

int vr_sfd_set_state(vr_context* ctx, const float* params11, const float* m, const float* v, int t) {
    if (!ctx || !ctx->sfd || !params11 || !m || !v || t < 0) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_set_state: bad arguments (vr_sfd_begin first)");
    VR_CUDA(cudaSetDevice(ctx->device));
    SfdState* s = ctx->sfd;
    const size_t bytes = (size_t)11 * s->n * 4;
    VR_CUDA(cudaMemcpyAsync(s->params.p, params11, bytes, cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(s->m.p, m, bytes, cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(s->v.p, v, bytes, cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    s->t = t;
    return VR_OK;
}

int vr_sfd_get_params(vr_context* ctx, float* params11) {
    if (!ctx || !ctx->sfd || !params11) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_sfd_get_params: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    VR_CUDA(cudaMemcpyAsync(params11, ctx->sfd->params.p, (size_t)11 * ctx->sfd->n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

int vr_adam_step(vr_context* ctx, int D, float* params, const float* grads, float* m, float* v, int* t, float lr,
                 float beta1, float beta2, float eps) {
    if (!ctx || D <= 0 || !params || !grads || !m || !v || !t) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_adam_step: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t n = (size_t)D;
    VR_CUDA(ctx->scratch.ensure(n * 16));
    float* b = ctx->scratch.as<float>();
    cudaStream_t st = ctx->stream;
    VR_CUDA(cudaMemcpyAsync(b, params, n * 4, cudaMemcpyHostToDevice, st));
    VR_CUDA(cudaMemcpyAsync(b + n, grads, n * 4, cudaMemcpyHostToDevice, st));
    VR_CUDA(cudaMemcpyAsync(b + 2 * n, m, n * 4, cudaMemcpyHostToDevice, st));
    VR_CUDA(cudaMemcpyAsync(b + 3 * n, v, n * 4, cudaMemcpyHostToDevice, st));
    *t += 1;
    adam_kernel<<<(D + 255) / 256, 256, 0, st>>>(D, b, b + n, b + 2 * n, b + 3 * n, adam_alpha(lr, beta1, beta2, *t), beta1, beta2, eps);
    VR_CUDA(cudaGetLastError());
    VR_CUDA(cudaMemcpyAsync(params, b, n * 4, cudaMemcpyDeviceToHost, st));
    VR_CUDA(cudaMemcpyAsync(m, b + 2 * n, n * 4, cudaMemcpyDeviceToHost, st));
    VR_CUDA(cudaMemcpyAsync(v, b + 3 * n, n * 4, cudaMemcpyDeviceToHost, st));
    VR_CUDA(cudaStreamSynchronize(st));
    return VR_OK;
}

}  // extern "C"
