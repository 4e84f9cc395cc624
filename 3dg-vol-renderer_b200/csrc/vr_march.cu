// vr_march.cu -- the reference's validation integrators over a Gaussian mixture:
//   mode 0  TestIntegrator        integrator.h:65-94        magenta where the primary ray overlaps anything
//   mode 1  PureRayMarching       integrator.h:100-267      medium-agnostic marcher (marched shadow rays)
//   mode 2  RayMarchingGaussians  test_integrators.h:143-297  marcher with analytic (erf) transmittance
//
// One thread per pixel (they are validation tools: a 512^2 image keeps the device full, and a pixel's march is
// sequential in the reference's float arithmetic anyway).  The reference keeps a bool mask over ALL components
// and flips it while walking time-sorted event vectors.  Because events are consumed in time order, a
// component's state at time t depends only on its own interval, so no mask and no event vector are needed:
//   primary ray     active(t) = t0 <= t && !(t1 <= t)
//   secondary ray   active(t) = (pa || t0 <= t) && !(t1 <= t)
// where pa marks the components active at the scatter point, which the reference prepends as {0, entering}
// events (integrator.h:200-208, test_integrators.h:204-207); a pa component the secondary ray misses never
// exits (t0 = t1 = +inf here), exactly as its mask bit stays set in the reference.  Sums run over the active
// components in ascending file index, like the reference's mask scans.  The segment walk of mode 2 visits the
// distinct event times; the reference's zero-length segments between equal times multiply by exp(-0) = 1.
// Environment directions: PCG32(derive_path_seed(x, y, step), 2), the substitution vr_spheres.cu and the
// oracle (vo_render_march) make for the reference's random_device-seeded generator.
#include <algorithm>
#include <cstring>

#include "vr_context.h"

namespace vr {

struct MarchArgs {
    SceneDev S; CameraPod cam;
    int W, H, x0, y0, win_w, rows, part_rank, part_count, env_samples, mode, cap;
    float step;
    float* image;
    float* arena;          // [warp][2][kArenaFields * cap * 32]
    unsigned int* error;   // set when a ray overlaps more than cap components
};

enum { F_PA = F_ACTF };    // secondary store: 1 = active at the scatter point
enum { U_ORIG = U_ACT };   // both stores: file index of the slot's component

template <class St>
__device__ void sort_by_orig(St& st, int K) {  // U_ORDER = slots in ascending file index
    for (int i = 0; i < K; ++i) {
        const uint32_t key = st.U(U_ORIG, i);
        int j = i;
        while (j > 0 && st.U(U_ORIG, (int)st.U(U_ORDER, j - 1)) > key) { st.U(U_ORDER, j) = st.U(U_ORDER, j - 1); --j; }
        st.U(U_ORDER, j) = (uint32_t)i;
    }
}

// GaussianMixtureModel::evaluate_sigma (gmm.h:98-126) over the slots for which `on(slot)` holds
template <class St, class On>
__device__ __forceinline__ void eval_sigma(const SceneDev& S, St& st, int K, f3 pos, On on, float* sa, float* ss) {
    float sum = 0.0f, sum_alb = 0.0f;
    for (int j = 0; j < K; ++j) {
        const int s = (int)st.U(U_ORDER, j);
        if (!on(s)) continue;
        const GaussRec g = load_gauss(S.gauss, st.U(U_GI, s));
        const float mu = gauss_mu_t(g, pos);
        sum = __fadd_rn(sum, mu);
        sum_alb = __fadd_rn(sum_alb, __fmul_rn(mu, g.albedo));
    }
    if (sum <= 0.0f) { *sa = 0.0f; *ss = 0.0f; return; }
    const float a_mix = __fdiv_rn(sum_alb, sum);
    *ss = __fmul_rn(a_mix, sum);
    *sa = __fmul_rn(__fsub_rn(1.0f, a_mix), sum);
}

// transmittance along a secondary ray from the scatter point at primary time t_cur.
// env = false: to a light at distance dist; env = true: to the environment.
template <class St>
__device__ float secondary_transmittance(const MarchArgs& A, St& P, int Kp, float t_cur, St& Q, f3 pos, f3 dir, bool env, float dist,
                                         bool* overflow) {
    const SceneDev& S = A.S;
    int K = gather_overlaps(S, pos, dir, CUDART_INF_F, false, Q);
    if (K < 0) { *overflow = true; return 0.0f; }
    const int K_real = K;
    float t_last = 0.0f;   // the reference's events.back().t
    for (int s = 0; s < K; ++s) {
        Q.F(F_PA, s) = 0.0f;
        Q.U(U_ORIG, s) = __ldg(S.orig + Q.U(U_GI, s));
        t_last = fmaxf(t_last, Q.F(F_T1, s));
    }
    for (int k = 0; k < Kp; ++k) {
        if (!(P.F(F_T0, k) <= t_cur && !(P.F(F_T1, k) <= t_cur))) continue;
        const uint32_t gi = P.U(U_GI, k);
        int at = -1;
        for (int s = 0; s < K_real; ++s) if (Q.U(U_GI, s) == gi) { at = s; break; }
        if (at < 0) {   // active at the scatter point, but this ray's quadric misses it: on for good
            if (K >= Q.cap()) { *overflow = true; return 0.0f; }
            const GaussRec g = load_gauss(S.gauss, gi);
            const DepthConst c = depth_const(g, ray_quad(g, pos, dir));
            at = K++;
            Q.F(F_T0, at) = CUDART_INF_F; Q.F(F_T1, at) = CUDART_INF_F;
            Q.F(F_A, at) = c.A; Q.F(F_B, at) = c.B; Q.F(F_K, at) = c.K; Q.F(F_DEN, at) = c.den;
            Q.U(U_GI, at) = gi; Q.U(U_ORIG, at) = P.U(U_ORIG, k);
        }
        Q.F(F_PA, at) = 1.0f;
    }
    sort_by_orig(Q, K);
    auto on_at = [&](int s, float t) { return (Q.F(F_PA, s) != 0.0f || Q.F(F_T0, s) <= t) && !(Q.F(F_T1, s) <= t); };
    float Tr = 1.0f;
    if (A.mode == 1) {  // PureRayMarching::march_transmittance, integrator.h:105-135
        const float t_max = env ? (K_real > 0 ? t_last : 0.0f) : dist;   // env: events.back().t, 0 without real events (:235-237)
        for (float t = 0.0f; t < t_max; t = __fadd_rn(t, A.step)) {
            float sa, ss;
            eval_sigma(S, Q, K, add3(pos, scale3(t, dir)), [&](int s) { return on_at(s, t); }, &sa, &ss);
            Tr = __fmul_rn(Tr, expf(__fmul_rn(-__fadd_rn(sa, ss), A.step)));
        }
        return Tr;
    }
    // test_integrators.h:209-231 (light: up to dist, last segment not clipped) and :247-263 (environment: up to the last event)
    float t_prev = 0.0f;
    while (env || t_prev < dist) {
        float t_next = CUDART_INF_F;
        for (int s = 0; s < K_real; ++s) {
            const float a = Q.F(F_T0, s), b = Q.F(F_T1, s);
            if (a > t_prev && a < t_next) t_next = a;
            if (b > t_prev && b < t_next) t_next = b;
        }
        if (t_next == CUDART_INF_F) {
            if (env) break;
            t_next = dist;
        }
        float sum = 0.0f;
        for (int j = 0; j < K; ++j) {
            const int s = (int)Q.U(U_ORDER, j);
            if (on_at(s, t_prev)) sum = __fadd_rn(sum, depth_between(hit_const(Q, s), t_prev, t_next));
        }
        Tr = __fmul_rn(Tr, expf(-sum));
        t_prev = t_next;
    }
    return Tr;
}

__global__ void __launch_bounds__(128) march_gaussians_kernel(const __grid_constant__ MarchArgs A) {
    const SceneDev& S = A.S;
    const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t store_floats = (size_t)kArenaFields * (size_t)A.cap * 32;
    ArenaStore P, Q;
    P.base = A.arena + (2 * warp) * store_floats + (threadIdx.x & 31); P.capacity = A.cap;
    Q.base = A.arena + (2 * warp + 1) * store_floats + (threadIdx.x & 31); Q.capacity = A.cap;
    const int npix = A.win_w * A.rows;
    const float four_pi = (float)(4.0 * VR_PI_D), inv4pi = (float)(1.0 / (4.0 * VR_PI_D));
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += gridDim.x * blockDim.x) {
        const int x = A.x0 + p % A.win_w;
        const int y = A.y0 + (p / A.win_w) * A.part_count + A.part_rank;
        const float u = __fdiv_rn(__fadd_rn((float)x, 0.5f), (float)A.W), v = __fdiv_rn(__fadd_rn((float)y, 0.5f), (float)A.H);
        const RayOD ray = camera_ray(A.cam, u, v);
        float* dst = A.image + 3 * ((size_t)y * A.W + x);
        const int K = gather_overlaps(S, ray.o, ray.d, CUDART_INF_F, false, P);
        if (A.mode == 0) {
            const bool hit = K != 0;
            dst[0] = hit ? 1.0f : S.env[0]; dst[1] = hit ? 0.0f : S.env[1]; dst[2] = hit ? 1.0f : S.env[2];
            continue;
        }
        if (K < 0) { atomicExch(A.error, 1u); dst[0] = dst[1] = dst[2] = 0.0f; continue; }
        if (K == 0) { dst[0] = S.env[0]; dst[1] = S.env[1]; dst[2] = S.env[2]; continue; }
        float t_end = 0.0f;
        for (int s = 0; s < K; ++s) { P.U(U_ORIG, s) = __ldg(S.orig + P.U(U_GI, s)); t_end = fmaxf(t_end, P.F(F_T1, s)); }
        sort_by_orig(P, K);
        float T = 1.0f, L[3] = {0.f, 0.f, 0.f};
        bool overflow = false;
        int stepi = 0;
        for (float t = 0.0f; t < t_end; t = __fadd_rn(t, A.step), ++stepi) {
            auto on = [&](int s) { return P.F(F_T0, s) <= t && !(P.F(F_T1, s) <= t); };
            const f3 pos = add3(ray.o, scale3(t, ray.d));
            float sa, ss;
            eval_sigma(S, P, K, pos, on, &sa, &ss);
            if (ss > 0.0f) {
                float Li[3] = {0.f, 0.f, 0.f}, Le[3] = {0.f, 0.f, 0.f};
                for (int l = 0; l < S.nl; ++l) {
                    const f3 lp = mk3(__ldg(S.lpos + 3 * l), __ldg(S.lpos + 3 * l + 1), __ldg(S.lpos + 3 * l + 2));
                    const f3 to = sub3(lp, pos);
                    const f3 wi = normalized3(to);
                    const float dist = __fsqrt_rn(sqnorm3(to));
                    const float Tr = secondary_transmittance(A, P, K, t, Q, pos, normalized3(wi), false, dist, &overflow);
                    const float d2 = __fmul_rn(dist, dist);
                    for (int c = 0; c < 3; ++c) Li[c] = __fadd_rn(Li[c], __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * l + c)), d2));
                }
                Pcg rng; rng.seed(derive_path_seed(x, y, stepi), 2);
                for (int e = 0; e < A.env_samples; ++e) {
                    const f3 wi = sample_uniform_direction(rng);
                    const float Tr = secondary_transmittance(A, P, K, t, Q, pos, normalized3(wi), true, CUDART_INF_F, &overflow);
                    for (int c = 0; c < 3; ++c) Le[c] = __fadd_rn(Le[c], __fmul_rn(Tr, S.env[c]));
                }
                for (int c = 0; c < 3; ++c) {
                    const float le = __fmul_rn(__fdiv_rn(Le[c], (float)A.env_samples), four_pi);
                    L[c] = __fadd_rn(L[c], __fmul_rn(__fmul_rn(__fmul_rn(__fmul_rn(T, ss), __fadd_rn(Li[c], le)), A.step), inv4pi));
                }
            }
            if (A.mode == 1) {
                T = __fmul_rn(T, expf(__fmul_rn(-__fadd_rn(sa, ss), A.step)));
            } else {   // transmittance_over_segment(ray, t, t + step, active), gmm.h:146-157
                const float t1 = __fadd_rn(t, A.step);
                float sum = 0.0f;
                for (int j = 0; j < K; ++j) {
                    const int s = (int)P.U(U_ORDER, j);
                    if (on(s)) sum = __fadd_rn(sum, depth_between(hit_const(P, s), t, t1));
                }
                T = __fmul_rn(T, expf(-sum));
            }
            if (overflow) break;
        }
        if (overflow) atomicExch(A.error, 1u);
        for (int c = 0; c < 3; ++c) dst[c] = __fadd_rn(L[c], __fmul_rn(T, S.env[c]));
    }
}

int render_march_impl(vr_context* ctx, const vr_render_params* p, vr_render_stats* st) {
    if (!ctx->has_gmm) return fail(ctx, VR_ERR_NO_SCENE, "render: no Gaussian mixture set (vr_set_gmm)");
    const int mode = p->integrator == VR_INTEGRATOR_TEST ? 0 : (p->integrator == VR_INTEGRATOR_PURE_MARCH ? 1 : 2);
    if (mode != 0 && (!(p->march_step > 0.0f) || p->env_samples < 1))
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "render: march_step and env_samples must be positive");
    const int W = p->width, H = p->height;
    cudaStream_t s = ctx->stream;
    VR_CUDA(ctx->image.ensure((size_t)3 * W * H * sizeof(float)));
    if (ctx->img_w != W || ctx->img_h != H) {
        VR_CUDA(cudaMemsetAsync(ctx->image.p, 0, (size_t)3 * W * H * sizeof(float), s));
        ctx->img_w = W; ctx->img_h = H;
    }
    MarchArgs A{};
    A.S = ctx->scene_dev(); A.cam = ctx->cam;
    A.W = W; A.H = H; A.x0 = p->x0; A.y0 = p->y0; A.win_w = p->x1 - p->x0;
    const int rows_all = p->y1 - p->y0;
    A.rows = p->part_rank >= rows_all ? 0 : (rows_all - p->part_rank + p->part_count - 1) / p->part_count;
    A.part_rank = p->part_rank; A.part_count = p->part_count;
    A.env_samples = p->env_samples; A.step = p->march_step; A.mode = mode;
    A.cap = ctx->hit_capacity > 0 ? ctx->hit_capacity : std::max(16, std::min(ctx->n, 1024));
    A.image = ctx->image.as<float>();
    const int npix = A.win_w * A.rows;
    const int grid = std::max(1, std::min((npix + 127) / 128, ctx->num_sms * 4));
    const size_t warps = (size_t)grid * 4;
    VR_CUDA(ctx->arena.ensure(warps * 2 * (size_t)kArenaFields * (size_t)A.cap * 32 * sizeof(float)));
    A.arena = ctx->arena.as<float>();
    VR_CUDA(ctx->counters.ensure(64 * sizeof(unsigned int)));
    A.error = ctx->counters.as<unsigned int>();
    VR_CUDA(cudaMemsetAsync(A.error, 0, sizeof(unsigned int), s));
    VR_CUDA(cudaEventRecord(ctx->ev[0], s));
    if (npix > 0) march_gaussians_kernel<<<grid, 128, 0, s>>>(A);
    VR_CUDA(cudaGetLastError());
    VR_CUDA(cudaEventRecord(ctx->ev[1], s));
    unsigned int err = 0;
    VR_CUDA(cudaMemcpyAsync(&err, A.error, sizeof(err), cudaMemcpyDeviceToHost, s));
    VR_CUDA(cudaStreamSynchronize(s));
    float ms = 0.f;
    VR_CUDA(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
    if (st) { st->kernel_ms = ms; st->total_ms = ms; st->samples = (uint64_t)npix; st->kernel_launches = npix > 0; }
    if (err) return fail(ctx, VR_ERR_HIT_OVERFLOW, "a ray overlapped more components than hit_capacity; raise it with vr_set_option(\"hit_capacity\")");
    return VR_OK;
}

}  // namespace vr
