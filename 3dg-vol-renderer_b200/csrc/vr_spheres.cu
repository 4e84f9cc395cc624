// vr_spheres.cu -- RayMarchingSpheres::render (test_integrators.h:23-135) over a SphereMixtureModel
// (smm.h:17-103): one thread per pixel, fixed-step march with analytic shadow transmittance.
// The mixture is tiny in every scene the reference ships (<= 3 spheres), so the sphere set is a
// 32-bit mask and events live in registers/local memory; more than 32 spheres is VR_ERR_UNSUPPORTED.
// Environment directions come from PCG32(derive_path_seed(x, y, step), 2): the reference uses a
// std::random_device-seeded generator there (integrator.h:13-28), which nothing can
// reproduce -- same substitution as oracle/volpath_oracle.c, so GPU and oracle agree sample by sample.
#include <algorithm>
#include <cstring>

#include "vr_context.h"

namespace vr {

constexpr int kMaxSpheres = 32;

struct SphereSet { const float* s; int n; };  // 6 floats each: cx cy cz r sigma_a sigma_s
struct SphEvents { float t[2 * kMaxSpheres]; unsigned char code[2 * kMaxSpheres]; int n; };  // code = idx*2 + entering

__device__ __forceinline__ bool sphere_hit(const SphereSet& S, int i, f3 o, f3 d, float* t_enter, float* t_exit) {  // smm.h:29-39
    const float* p = S.s + 6 * i;
    const f3 Lv = sub3(mk3(p[0], p[1], p[2]), o);
    const float tca = dot3(Lv, d);
    const float d2 = __fsub_rn(sqnorm3(Lv), __fmul_rn(tca, tca));
    const float r2 = __fmul_rn(p[3], p[3]);
    if (d2 > r2) return false;
    const float thc = __fsqrt_rn(__fsub_rn(r2, d2));
    *t_enter = __fsub_rn(tca, thc);
    *t_exit = __fadd_rn(tca, thc);
    return *t_exit >= 0.0f;
}
__device__ __forceinline__ void sph_push(SphEvents& ev, float t, int code) {  // stable insertion by t
    int j = ev.n++;
    while (j > 0 && ev.t[j - 1] > t) { ev.t[j] = ev.t[j - 1]; ev.code[j] = ev.code[j - 1]; --j; }
    ev.t[j] = t; ev.code[j] = (unsigned char)code;
}
__device__ void sphere_events(const SphereSet& S, f3 o, f3 d, SphEvents& ev) {  // smm.h:54-63
    ev.n = 0;
    for (int i = 0; i < S.n; ++i) {
        float t0, t1;
        if (sphere_hit(S, i, o, d, &t0, &t1)) {
            if (t0 >= 0.0f) sph_push(ev, t0, 2 * i + 1);
            if (t1 >= 0.0f) sph_push(ev, t1, 2 * i);
        }
    }
}
// smm.h:79-103 with the currently active spheres switched on from t = 0 (test_integrators.h:71-91)
__device__ float sphere_shadow(const SphereSet& S, f3 o, f3 d, uint32_t active0, float tmax, SphEvents& ev) {
    sphere_events(S, o, d, ev);
    uint32_t act = active0;
    float T = 1.0f, t_prev = 0.0f;
    for (int k = 0; k < ev.n; ++k) {
        if (ev.t[k] > tmax) break;
        const float dt = __fsub_rn(ev.t[k], t_prev);
        float sigma_t = 0.0f;
        for (int i = 0; i < S.n; ++i)
            if (act & (1u << i)) sigma_t = __fadd_rn(sigma_t, __fadd_rn(S.s[6 * i + 4], S.s[6 * i + 5]));
        T = __fmul_rn(T, expf(__fmul_rn(-sigma_t, dt)));
        const int idx = ev.code[k] >> 1;
        if (ev.code[k] & 1) act |= (1u << idx); else act &= ~(1u << idx);
        t_prev = ev.t[k];
    }
    return T;
}

struct SphereArgs {
    SphereSet S; SceneDev L; CameraPod cam;
    int W, H, x0, y0, win_w, rows, part_rank, part_count, env_samples;
    float step;
    float* image;
};

__global__ void __launch_bounds__(128) march_spheres_kernel(const __grid_constant__ SphereArgs A) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= A.win_w * A.rows) return;
    const int i = A.x0 + p % A.win_w;
    const int j = A.y0 + (p / A.win_w) * A.part_count + A.part_rank;
    const float u = __fdiv_rn(__fadd_rn((float)i, 0.5f), (float)A.W), v = __fdiv_rn(__fadd_rn((float)j, 0.5f), (float)A.H);
    const RayOD ray = camera_ray(A.cam, u, v);
    float* dst = A.image + 3 * ((size_t)j * A.W + i);
    SphEvents ev, sev;
    sphere_events(A.S, ray.o, ray.d, ev);
    if (ev.n == 0) { dst[0] = A.L.env[0]; dst[1] = A.L.env[1]; dst[2] = A.L.env[2]; return; }
    float rad[3] = {0.f, 0.f, 0.f};
    uint32_t active = 0;
    int cur = 0, stepi = 0;
    float t = 0.0f, T = 1.0f;
    const float t_end = ev.t[ev.n - 1];
    const float four_pi = (float)(4.0 * VR_PI_D), inv4pi = (float)(1.0 / (4.0 * VR_PI_D));
    while (t < t_end) {
        while (cur < ev.n && ev.t[cur] <= t) {
            const int idx = ev.code[cur] >> 1;
            if (ev.code[cur] & 1) active |= (1u << idx); else active &= ~(1u << idx);
            ++cur;
        }
        const f3 pos = add3(ray.o, scale3(t, ray.d));
        float sa = 0.0f, ss = 0.0f;
        for (int k = 0; k < A.S.n; ++k)
            if (active & (1u << k)) { sa = __fadd_rn(sa, A.S.s[6 * k + 4]); ss = __fadd_rn(ss, A.S.s[6 * k + 5]); }
        if (ss > 0.0f) {
            float Li[3] = {0.f, 0.f, 0.f};
            for (int l = 0; l < A.L.nl; ++l) {
                const f3 lp = mk3(A.L.lpos[3 * l], A.L.lpos[3 * l + 1], A.L.lpos[3 * l + 2]);
                const f3 to = sub3(lp, pos);
                const f3 wi = normalized3(to);
                const float dist = __fsqrt_rn(sqnorm3(to));
                const float Tr = sphere_shadow(A.S, pos, normalized3(wi), active, dist, sev);
                const float d2 = __fmul_rn(dist, dist);
                for (int c = 0; c < 3; ++c) Li[c] = __fadd_rn(Li[c], __fdiv_rn(__fmul_rn(Tr, A.L.lint[3 * l + c]), d2));
            }
            float Le[3] = {0.f, 0.f, 0.f};
            Pcg rng; rng.seed(derive_path_seed(i, j, stepi), 2);
            for (int e = 0; e < A.env_samples; ++e) {
                const f3 wi = sample_uniform_direction(rng);
                const float Tr = sphere_shadow(A.S, pos, normalized3(wi), active, CUDART_INF_F, sev);
                for (int c = 0; c < 3; ++c) Le[c] = __fadd_rn(Le[c], __fmul_rn(Tr, A.L.env[c]));
            }
            for (int c = 0; c < 3; ++c) {
                float le = __fmul_rn(__fdiv_rn(Le[c], (float)A.env_samples), four_pi);
                rad[c] = __fadd_rn(rad[c], __fmul_rn(__fmul_rn(__fmul_rn(__fmul_rn(T, ss), __fadd_rn(Li[c], le)), A.step), inv4pi));
            }
        }
        T = __fmul_rn(T, expf(__fmul_rn(-A.step, __fadd_rn(sa, ss))));
        t = __fadd_rn(t, A.step);  // float accumulation of t, test_integrators.h:127
        ++stepi;
    }
    for (int c = 0; c < 3; ++c) dst[c] = __fadd_rn(rad[c], __fmul_rn(T, A.L.env[c]));
}

int render_spheres_impl(vr_context* ctx, const vr_render_params* p, vr_render_stats* st) {
    if (!ctx->has_smm) return fail(ctx, VR_ERR_NO_SCENE, "render: no sphere mixture set (vr_set_smm)");
    if (!(p->march_step > 0.0f) || p->env_samples < 1) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "render: march_step and env_samples must be positive");
    const int W = p->width, H = p->height;
    cudaStream_t s = ctx->stream;
    VR_CUDA(ctx->image.ensure((size_t)3 * W * H * sizeof(float)));
    if (ctx->img_w != W || ctx->img_h != H) {
        VR_CUDA(cudaMemsetAsync(ctx->image.p, 0, (size_t)3 * W * H * sizeof(float), s));
        ctx->img_w = W; ctx->img_h = H;
    }
    SphereArgs A{};
    A.S.s = ctx->spheres.as<float>(); A.S.n = ctx->n_spheres;
    A.L = ctx->scene_dev(); A.cam = ctx->cam;
    A.W = W; A.H = H; A.x0 = p->x0; A.y0 = p->y0; A.win_w = p->x1 - p->x0;
    const int rows_all = p->y1 - p->y0;
    A.rows = p->part_rank >= rows_all ? 0 : (rows_all - p->part_rank + p->part_count - 1) / p->part_count;
    A.part_rank = p->part_rank; A.part_count = p->part_count;
    A.env_samples = p->env_samples; A.step = p->march_step;
    A.image = ctx->image.as<float>();
    const int npix = A.win_w * A.rows;
    VR_CUDA(cudaEventRecord(ctx->ev[0], s));
    if (npix > 0) march_spheres_kernel<<<(npix + 127) / 128, 128, 0, s>>>(A);
    VR_CUDA(cudaGetLastError());
    VR_CUDA(cudaEventRecord(ctx->ev[1], s));
    VR_CUDA(cudaStreamSynchronize(s));
    float ms = 0.f;
    VR_CUDA(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
    if (st) { st->kernel_ms = ms; st->total_ms = ms; st->samples = (uint64_t)npix; st->kernel_launches = npix > 0; }
    return VR_OK;
}

}  // namespace vr

extern "C" int vr_set_smm(vr_context* ctx, int n, const float* center3, const float* radius, const float* sigma_a,
                          const float* sigma_s) {
    using namespace vr;
    if (!ctx || n < 0 || (n > 0 && (!center3 || !radius || !sigma_a || !sigma_s))) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_set_smm: bad arguments");
    if (n > kMaxSpheres) return fail(ctx, VR_ERR_UNSUPPORTED, "vr_set_smm: more than 32 spheres");
    VR_CUDA(cudaSetDevice(ctx->device));
    std::vector<float> packed((size_t)std::max(n, 1) * 6, 0.f);
    for (int i = 0; i < n; ++i) {
        packed[6 * i] = center3[3 * i]; packed[6 * i + 1] = center3[3 * i + 1]; packed[6 * i + 2] = center3[3 * i + 2];
        packed[6 * i + 3] = radius[i]; packed[6 * i + 4] = sigma_a[i]; packed[6 * i + 5] = sigma_s[i];
    }
    VR_CUDA(ctx->spheres.ensure(packed.size() * 4));
    VR_CUDA(cudaMemcpyAsync(ctx->spheres.p, packed.data(), packed.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->n_spheres = n; ctx->has_smm = true;
    return VR_OK;
}
