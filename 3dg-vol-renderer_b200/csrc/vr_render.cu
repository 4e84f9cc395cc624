// vr_render.cu -- Integrator::render behind the C-ABI: pass scheduling, kernel launches, copies.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "vr_context.h"

using namespace vr;

namespace vr {
int render_spheres_impl(vr_context* ctx, const vr_render_params* p, vr_render_stats* st);  // vr_spheres.cu
int render_march_impl(vr_context* ctx, const vr_render_params* p, vr_render_stats* st);    // vr_march.cu

// reach of a batch of ray origins (vr_transmittance, vr_kat_gather): farthest origin from the mixture's box, largest coordinate
static int ensure_origin_reach(vr_context* ctx, int nrays, const float* o3) {
    if (!ctx->has_gmm || ctx->n <= 0) return VR_OK;
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
    if (!(R2 < 1e60) || !(X < 1e30)) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "ray origin is not finite");
    return ensure_reach(ctx, (float)std::sqrt(R2), (float)X);
}

static int owned_rows(const vr_render_params* p) {
    const int rows = p->y1 - p->y0;
    if (p->part_rank >= rows) return 0;
    return (rows - p->part_rank + p->part_count - 1) / p->part_count;
}

static int validate(vr_context* ctx, const vr_render_params* p) {
    if (!ctx || !p) return VR_ERR_INVALID_ARGUMENT;
    if (p->width <= 0 || p->height <= 0 || p->spp <= 0) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "render: width, height and spp must be positive");
    if (p->width > 65535 || p->height > 65535) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "render: image side above 65535 (path seed packs x,y in 16 bits, rng.h:52-57)");
    if (p->x0 < 0 || p->y0 < 0 || p->x1 > p->width || p->y1 > p->height || p->x0 >= p->x1 || p->y0 >= p->y1)
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "render: bad pixel window");
    if (p->part_count < 1 || p->part_rank < 0 || p->part_rank >= p->part_count)
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "render: bad row partition");
    if (p->solver < 0 || p->solver > 4) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "render: unknown solver");
    if (!ctx->has_camera) return fail(ctx, VR_ERR_NO_SCENE, "render: no camera set");
    return VR_OK;
}

int render_device_impl(vr_context* ctx, const vr_render_params* p, vr_render_stats* st) {
    int rc = validate(ctx, p);
    if (rc) return rc;
    if (st) memset(st, 0, sizeof(*st));
    VR_CUDA(cudaSetDevice(ctx->device));
    if (p->integrator == VR_INTEGRATOR_MARCH_SPHERES) return render_spheres_impl(ctx, p, st);
    if (p->integrator == VR_INTEGRATOR_TEST || p->integrator == VR_INTEGRATOR_PURE_MARCH || p->integrator == VR_INTEGRATOR_MARCH_GAUSSIANS)
        return render_march_impl(ctx, p, st);
    if (p->integrator != VR_INTEGRATOR_FREE_FLIGHT_SS && p->integrator != VR_INTEGRATOR_MULTI_SCATTER)
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "render: unknown integrator");
    if (!ctx->has_gmm) return fail(ctx, VR_ERR_NO_SCENE, "render: no Gaussian mixture set (vr_set_gmm)");
    rc = ensure_camera_reach(ctx);   // a camera farther out than the culling boxes were inflated for: rebuild (vr_bvh.h BoxReach)
    if (rc) return rc;
    NvtxRange range_render("vr.render");
    const int W = p->width, H = p->height;
    cudaStream_t s = ctx->stream;
    VR_CUDA(ctx->image.ensure((size_t)3 * W * H * sizeof(float)));
    if (ctx->img_w != W || ctx->img_h != H) {
        VR_CUDA(cudaMemsetAsync(ctx->image.p, 0, (size_t)3 * W * H * sizeof(float), s));
        ctx->img_w = W; ctx->img_h = H;
    }
    // the per-light grids cost ~1 ms per 1000 components to build on the host: worth it from ~16 M samples on
    const unsigned long long render_samples = (unsigned long long)(p->x1 - p->x0) * (unsigned long long)owned_rows(p) * (unsigned long long)p->spp;
    const bool use_grids = ctx->engine == 0 && ctx->use_light_grid && (double)render_samples >= ctx->light_grid_min_samples;
    ctx->lg_speculate = use_grids;   // the next vr_set_gmm* builds the grids next to the hierarchy
    if (use_grids) { rc = ensure_light_grids(ctx); if (rc) return rc; }
    // camera rays of a pinhole camera take their candidates from a direction grid around the pinhole (wave_regen_cam_kernel)
    const bool want_cam_grid = ctx->engine == 0 && ctx->use_camera_grid && ctx->trav_version != 1 && ctx->cam.kind == VR_CAMERA_PINHOLE &&
                               (double)render_samples >= ctx->camera_grid_min_samples;
    if (want_cam_grid) { rc = ensure_camera_grid(ctx); if (rc) return rc; }
    // bounce rays take the candidates of their first stretch from a uniform grid over the mixture (vr_grid3d.cu)
    const bool want_grid3d = ctx->engine == 0 && ctx->use_grid3d && ctx->trav_version == 2 && p->integrator == VR_INTEGRATOR_MULTI_SCATTER &&
                             (double)render_samples >= ctx->grid3d_min_samples;
    if (want_grid3d) { rc = ensure_grid3d(ctx); if (rc) return rc; }
    RenderArgs A{};
    A.scene = ctx->scene_dev();
    if (!use_grids) A.scene.lg_mask = 0;
    if (want_grid3d && ctx->g3_valid && ctx->g3_ok) A.scene.g3_res = ctx->g3_res;
    if (want_cam_grid && ctx->cg_valid && ctx->cg_ok) A.scene.cg_res = ctx->lg_res;
    A.cam = ctx->cam;
    A.W = W; A.H = H; A.spp = p->spp;
    A.nsq = (int)std::sqrt((double)p->spp);  // integrator.h:562
    if (A.nsq < 1) A.nsq = 1;
    A.integrator = p->integrator;
    A.trav_version = ctx->trav_version;
    A.post_split = ctx->post_split;
    A.post_defer = ctx->post_defer;
    A.post_pair = ctx->post_pair;
    A.solver = p->solver;
    A.min_bounces = p->min_bounces;
    A.x0 = p->x0; A.y0 = p->y0; A.win_w = p->x1 - p->x0;
    A.part_rank = p->part_rank; A.part_count = p->part_count;
    A.stats = ctx->stats.as<unsigned long long>();
    A.counter = ctx->counters.as<unsigned int>();
    A.overflow_count = ctx->counters.as<unsigned int>() + 1;
    // ---- recording bitmaps (RECORD_PIXEL_GAUSSIANS, integrator.h:540-545)
    ctx->rec_valid = false;
    if (p->record && p->integrator == VR_INTEGRATOR_MULTI_SCATTER) {
        const int words = (std::max(ctx->n, 1) + 31) / 32;
        const size_t bytes = (size_t)W * H * words * 4;
        VR_CUDA(ctx->record_bits.ensure(bytes));
        VR_CUDA(cudaMemsetAsync(ctx->record_bits.p, 0, bytes, s));
        A.record_bits = ctx->record_bits.as<uint32_t>();
        A.record_words = words;
        ctx->rec_w = W; ctx->rec_h = H; ctx->rec_words = words; ctx->rec_n = ctx->n;
    }
    const unsigned long long npix = (unsigned long long)A.win_w * (unsigned long long)owned_rows(p);
    ctx->samples_valid = 0;
    VR_CUDA(cudaMemsetAsync(ctx->stats.p, 0, sizeof(unsigned long long) * kNumStats, s));
    VR_CUDA(cudaEventRecord(ctx->ev[0], s));
    double kernel_ms = 0.0;
    unsigned launches = 0;
    unsigned long long overflow_total = 0;
    if (npix > 0) {
        double cap = std::min(ctx->max_pass_samples, 4.0e9);
        unsigned long long pix_per_pass = std::max<unsigned long long>(1, (unsigned long long)(cap / p->spp));
        pix_per_pass = std::min(pix_per_pass, npix);
        if (pix_per_pass * (unsigned long long)p->spp > 0xfffffff0ULL) return fail(ctx, VR_ERR_UNSUPPORTED, "render: spp too large for one pass");
        VR_CUDA(ctx->sampleL.ensure(pix_per_pass * (size_t)p->spp * sizeof(float4)));
        A.sampleL = ctx->sampleL.as<float4>();
        const int grid_main = ctx->num_sms * ctx->blocks_per_sm_main;
        for (unsigned long long p0 = 0; p0 < npix; p0 += pix_per_pass) {
            const unsigned int pass_pixels = (unsigned int)std::min(pix_per_pass, npix - p0);
            A.pixel_base = (unsigned int)p0;
            A.work_count = pass_pixels * (unsigned int)p->spp;
            A.worklist = nullptr;
            // overflow samples are a fraction of a per cent even in heavy overlap (0.13 % at C3): 16 M entries (64 MB) bound the
            // list; a pass that needs more fails loudly below (shrink max_pass_samples) instead of dropping samples
            A.overflow_cap = (unsigned int)std::min<unsigned long long>(A.work_count, 1ull << 24);
            VR_CUDA(ctx->overflow_list.ensure((size_t)A.overflow_cap * 4));
            A.overflow_list = ctx->overflow_list.as<unsigned int>();
            VR_CUDA(cudaMemsetAsync(ctx->counters.p, 0, sizeof(unsigned int) * 4, s));
            VR_CUDA(cudaEventRecord(ctx->ev[2], s));
            if (ctx->engine == 0) {
                NvtxRange range_pass("vr.render.wavefront_pass");
                const int P = ctx->wave_slots > 0 ? ctx->wave_slots : ctx->num_sms * 32768;
                // big passes are cut into independent parts on their own streams so that kernel tails overlap
                int parts = (unsigned long long)A.work_count >= 4ULL * (unsigned long long)std::min(P, ctx->num_sms * 24576) ? ctx->wave_parts : 1;
                parts = std::max(1, std::min(parts, kMaxWaveParts));
                VR_CUDA(ctx->wavepool.ensure(wave_pool_bytes(P / parts) * parts));
                for (int i = 0; i < parts; ++i)
                    if (!ctx->wave_streams[i]) VR_CUDA(cudaStreamCreateWithFlags(&ctx->wave_streams[i], cudaStreamNonBlocking));
                if (run_wavefront_pass(A, ctx->wavepool.p, P, parts, ctx->num_sms, ctx->h_wave_counters, s, ctx->wave_streams, &launches, ctx->wave_graphs) < 0) {
                    VR_CUDA(cudaGetLastError());
                    return fail(ctx, VR_ERR_CUDA, "wavefront pass failed");
                }
            } else {
                launch_path_kernel(A, grid_main, false, s); ++launches;
            }
            VR_CUDA(cudaGetLastError());
            VR_CUDA(cudaMemcpyAsync(ctx->h_counters, ctx->counters.p, sizeof(unsigned int) * 4, cudaMemcpyDeviceToHost, s));
            VR_CUDA(cudaStreamSynchronize(s));
            const unsigned int n_over = ctx->h_counters[1];
            if (n_over > A.overflow_cap)
                return fail(ctx, VR_ERR_HIT_OVERFLOW, "more than 16 M samples of one pass overflowed the per-slot overlap records; lower max_pass_samples");
            if (n_over > 0) {  // samples whose rays overlapped more than kLocalCap components: re-trace with the arena store
                NvtxRange range_over("vr.render.overflow_retrace");
                overflow_total += n_over;
                const int cap_hits = ctx->hit_capacity > 0 ? ctx->hit_capacity : std::max(kLocalCap * 2, std::min(ctx->n, 4096));
                int grid_arena = ctx->num_sms * ctx->blocks_per_sm_arena;
                grid_arena = (int)std::min<unsigned long long>((unsigned long long)grid_arena, (n_over + kBlockThreads - 1) / kBlockThreads);
                const size_t warps = (size_t)grid_arena * (kBlockThreads / 32);
                VR_CUDA(ctx->arena.ensure(warps * kArenaFields * (size_t)cap_hits * 32 * sizeof(float)));
                RenderArgs B = A;
                B.arena = ctx->arena.as<float>(); B.arena_cap = cap_hits;
                B.worklist = ctx->overflow_list.as<unsigned int>();
                B.work_count = n_over;
                B.counter = ctx->counters.as<unsigned int>() + 2;
                launch_path_kernel(B, grid_arena, true, s); ++launches;
                VR_CUDA(cudaGetLastError());
            }
            VR_CUDA(cudaEventRecord(ctx->ev[3], s));
            { NvtxRange r("vr.render.pixel_reduce"); launch_pixel_reduce(A, pass_pixels, ctx->image.as<float>(), s); ++launches; }
            VR_CUDA(cudaGetLastError());
            VR_CUDA(cudaEventSynchronize(ctx->ev[3]));
            float ms = 0.f;
            VR_CUDA(cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]));
            kernel_ms += ms;
        }
    }
    VR_CUDA(cudaMemcpyAsync(ctx->h_stats, ctx->stats.p, sizeof(unsigned long long) * kNumStats, cudaMemcpyDeviceToHost, s));
    VR_CUDA(cudaEventRecord(ctx->ev[1], s));
    VR_CUDA(cudaStreamSynchronize(s));
    float total = 0.f;
    VR_CUDA(cudaEventElapsedTime(&total, ctx->ev[0], ctx->ev[1]));
    if (A.record_bits) ctx->rec_valid = true;
    // per-sample radiance of a render that fitted one pass stays readable (vr_get_samples)
    if (npix > 0 && npix * (unsigned long long)p->spp <= (unsigned long long)std::min(ctx->max_pass_samples, 4.0e9))
        ctx->samples_valid = npix * (unsigned long long)p->spp;
    if (st) {
        st->kernel_ms = kernel_ms; st->total_ms = total;
        st->samples = npix * (unsigned long long)p->spp;
        st->scatter_rays = ctx->h_stats[ST_SCATTER_RAYS]; st->shadow_rays = ctx->h_stats[ST_SHADOW_RAYS];
        st->hits = ctx->h_stats[ST_HITS]; st->overflow_samples = overflow_total;
        st->kernel_launches = launches;
        st->node_visits = ctx->h_stats[ST_NODES]; st->quadric_tests = ctx->h_stats[ST_QUADRICS]; st->erf_evals = ctx->h_stats[ST_ERF];
        st->mu_evals = ctx->h_stats[ST_MU]; st->analytic_solves = ctx->h_stats[ST_SA]; st->iterative_solves = ctx->h_stats[ST_ITER];
    }
    if (ctx->h_stats[ST_ERRORS])
        return fail(ctx, VR_ERR_HIT_OVERFLOW, "a ray overlapped more components than hit_capacity; raise it with vr_set_option(\"hit_capacity\")");
    return VR_OK;
}
}  // namespace vr

extern "C" {

int vr_render_device(vr_context* ctx, const vr_render_params* p, vr_render_stats* stats) {
    return render_device_impl(ctx, p, stats);
}

const float* vr_image_dev(vr_context* ctx) { return ctx ? ctx->image.as<float>() : nullptr; }

int vr_render(vr_context* ctx, const vr_render_params* p, float* out_rgb, vr_render_stats* stats) {
    if (!out_rgb) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_render: out_rgb is null");
    int rc = render_device_impl(ctx, p, stats);
    if (rc) return rc;
    const int rows = owned_rows(p);
    if (rows > 0) {
        // owned pixels only: a strided 2-D copy (row pitch = part_count image rows)
        const size_t row_bytes = (size_t)3 * p->width * sizeof(float);
        const size_t off = ((size_t)(p->y0 + p->part_rank) * p->width + p->x0) * 3;
        VR_CUDA(cudaEventRecord(ctx->ev[2], ctx->stream));
        VR_CUDA(cudaMemcpy2DAsync(out_rgb + off, row_bytes * p->part_count, ctx->image.as<float>() + off,
                                  row_bytes * p->part_count, (size_t)3 * (p->x1 - p->x0) * sizeof(float), (size_t)rows,
                                  cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaEventRecord(ctx->ev[3], ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        float ms = 0.f;
        VR_CUDA(cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]));
        if (stats) stats->total_ms += ms;
    }
    return VR_OK;
}

int vr_transmittance(vr_context* ctx, int nrays, const float* o3, const float* d3, const float* tmax, float* out) {
    if (!ctx || nrays < 0 || (nrays > 0 && (!o3 || !d3 || !tmax || !out))) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_transmittance: bad arguments");
    if (!ctx->has_gmm) return fail(ctx, VR_ERR_NO_SCENE, "vr_transmittance: no Gaussian mixture set");
    if (nrays == 0) return VR_OK;
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t n = (size_t)nrays;
    {   // reach of this batch's origins over the mixture (vr_bvh.h BoxReach)
        int rc = ensure_origin_reach(ctx, nrays, o3);
        if (rc) return rc;
    }
    VR_CUDA(ctx->scratch.ensure(n * 32));
    float* d = ctx->scratch.as<float>();
    float *d_o = d, *d_d = d + 3 * n, *d_t = d + 6 * n, *d_out = d + 7 * n;
    cudaStream_t s = ctx->stream;
    VR_CUDA(cudaMemcpyAsync(d_o, o3, n * 12, cudaMemcpyHostToDevice, s));
    VR_CUDA(cudaMemcpyAsync(d_d, d3, n * 12, cudaMemcpyHostToDevice, s));
    VR_CUDA(cudaMemcpyAsync(d_t, tmax, n * 4, cudaMemcpyHostToDevice, s));
    launch_transmittance(ctx->scene_dev(), nrays, d_o, d_d, d_t, d_out, s);
    VR_CUDA(cudaGetLastError());
    VR_CUDA(cudaMemcpyAsync(out, d_out, n * 4, cudaMemcpyDeviceToHost, s));
    VR_CUDA(cudaStreamSynchronize(s));
    return VR_OK;
}

// per-sample radiance of the last free-flight render (single pass): float4 per sample = (r, g, b, bounce index at the end of
// the path), ordered [owned pixel in row-major window order][si] -- what integrator.h:702 sums per pixel
int vr_get_samples(vr_context* ctx, float* out4, uint64_t count) {
    if (!ctx || !out4) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_get_samples: bad arguments");
    if (!ctx->samples_valid || count > ctx->samples_valid)
        return fail(ctx, VR_ERR_NO_SCENE, "vr_get_samples: the last render left no per-sample buffer of that size (multi-pass render?)");
    VR_CUDA(cudaSetDevice(ctx->device));
    VR_CUDA(cudaMemcpyAsync(out4, ctx->sampleL.p, (size_t)count * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

// per-pixel recorded sets -> counts (+ ascending original indices)
int vr_get_recorded(vr_context* ctx, int* counts, uint32_t* idx, int cap) {
    if (!ctx || !counts) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_get_recorded: bad arguments");
    if (!ctx->rec_valid) return fail(ctx, VR_ERR_NO_SCENE, "vr_get_recorded: the last render did not record");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t npix = (size_t)ctx->rec_w * ctx->rec_h, words = (size_t)ctx->rec_words;
    std::vector<uint32_t> bits(npix * words);
    VR_CUDA(cudaMemcpyAsync(bits.data(), ctx->record_bits.p, bits.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    for (size_t p = 0; p < npix; ++p) {
        int c = 0;
        for (size_t w = 0; w < words; ++w) {
            uint32_t v = bits[p * words + w];
            while (v) {
                const int b = __builtin_ctz(v);
                v &= v - 1;
                if (idx && c < cap) idx[p * (size_t)cap + c] = (uint32_t)(w * 32 + b);
                ++c;
            }
        }
        counts[p] = c;
    }
    return VR_OK;
}

// ---- known-answer hooks ------------------------------------------------------------------
#define VR_UP(dst, src, bytes) VR_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream))
#define VR_DOWN(dst, src, bytes) VR_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream))

int vr_kat_pcg(vr_context* ctx, int n, const int32_t* xysi3, uint64_t* seed, uint32_t* u32x4, float* uniformx4) {
    if (!ctx || n <= 0 || !xysi3 || !seed || !u32x4 || !uniformx4) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_pcg: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t m = (size_t)n;
    VR_CUDA(ctx->scratch.ensure(m * (12 + 8 + 16 + 16) + 64));
    char* b = ctx->scratch.as<char>();
    uint64_t* d_seed = (uint64_t*)b; int32_t* d_in = (int32_t*)(b + m * 8);
    uint32_t* d_u = (uint32_t*)(b + m * 20); float* d_f = (float*)(b + m * 36);
    VR_UP(d_in, xysi3, m * 12);
    launch_kat_pcg(n, d_in, d_seed, d_u, d_f, ctx->stream);
    VR_DOWN(seed, d_seed, m * 8); VR_DOWN(u32x4, d_u, m * 16); VR_DOWN(uniformx4, d_f, m * 16);
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

int vr_kat_gaussian(vr_context* ctx, int n, const float* o3, const float* d3, const int32_t* gidx, const float* frac,
                    int32_t* hit, float* t0, float* t1, float* tau, float* mu_mid, int32_t* ok, float* t_solved) {
    if (!ctx || n <= 0 || !o3 || !d3 || !gidx || !frac || !hit || !t0 || !t1 || !tau || !mu_mid || !ok || !t_solved)
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_gaussian: bad arguments");
    if (!ctx->has_gmm || ctx->n == 0) return fail(ctx, VR_ERR_NO_SCENE, "vr_kat_gaussian: no Gaussian mixture set");
    for (int i = 0; i < n; ++i) if (gidx[i] < 0 || gidx[i] >= ctx->n) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_gaussian: index out of range");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t m = (size_t)n;
    VR_CUDA(ctx->scratch.ensure(m * 4 * 15));
    float* b = ctx->scratch.as<float>();
    float *d_o = b, *d_d = b + 3 * m, *d_fr = b + 6 * m, *d_t0 = b + 7 * m, *d_t1 = b + 8 * m, *d_tau = b + 9 * m,
          *d_mu = b + 10 * m, *d_ts = b + 11 * m;
    int32_t *d_g = (int32_t*)(b + 12 * m), *d_hit = (int32_t*)(b + 13 * m), *d_ok = (int32_t*)(b + 14 * m);
    VR_UP(d_o, o3, m * 12); VR_UP(d_d, d3, m * 12); VR_UP(d_fr, frac, m * 4); VR_UP(d_g, gidx, m * 4);
    launch_kat_gaussian(ctx->scene_dev(), ctx->leaf_of_orig.as<uint32_t>(), n, d_o, d_d, d_g, d_fr, d_hit, d_t0, d_t1, d_tau,
                        d_mu, d_ok, d_ts, ctx->stream);
    VR_DOWN(hit, d_hit, m * 4); VR_DOWN(t0, d_t0, m * 4); VR_DOWN(t1, d_t1, m * 4); VR_DOWN(tau, d_tau, m * 4);
    VR_DOWN(mu_mid, d_mu, m * 4); VR_DOWN(ok, d_ok, m * 4); VR_DOWN(t_solved, d_ts, m * 4);
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

int vr_kat_camera_rays(vr_context* ctx, int n, const float* uv2, float* o3, float* d3) {
    if (!ctx || n <= 0 || !uv2 || !o3 || !d3) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_camera_rays: bad arguments");
    if (!ctx->has_camera) return fail(ctx, VR_ERR_NO_SCENE, "vr_kat_camera_rays: no camera set");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t m = (size_t)n;
    VR_CUDA(ctx->scratch.ensure(m * 32));
    float* b = ctx->scratch.as<float>();
    VR_UP(b, uv2, m * 8);
    launch_kat_camera(ctx->cam, n, b, b + 2 * m, b + 5 * m, ctx->stream);
    VR_DOWN(o3, b + 2 * m, m * 12); VR_DOWN(d3, b + 5 * m, m * 12);
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

int vr_kat_events(vr_context* ctx, const float* o3, const float* d3, int cap, float* t, int32_t* entering,
                  uint32_t* index, int* n_events) {
    if (!ctx || !o3 || !d3 || cap <= 0 || !t || !entering || !index || !n_events) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_events: bad arguments");
    if (!ctx->has_gmm) return fail(ctx, VR_ERR_NO_SCENE, "vr_kat_events: no Gaussian mixture set");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t hits = (size_t)cap / 2 + 1;
    const size_t arena_f = (size_t)kArenaFields * hits * 32;
    VR_CUDA(ctx->scratch.ensure((arena_f + 8 + 3 * (size_t)cap + 4) * 4));
    float* b = ctx->scratch.as<float>();
    float* d_o = b + arena_f; float* d_d = d_o + 3; float* d_t = d_d + 5;
    int32_t* d_e = (int32_t*)(d_t + cap); uint32_t* d_i = (uint32_t*)(d_e + cap); int* d_n = (int*)(d_i + cap);
    VR_UP(d_o, o3, 12); VR_UP(d_d, d3, 12);
    launch_kat_events(ctx->scene_dev(), d_o, d_d, (int)hits, b, d_t, d_e, d_i, d_n, ctx->stream);
    VR_DOWN(n_events, d_n, 4);
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    if (*n_events < 0) return fail(ctx, VR_ERR_HIT_OVERFLOW, "vr_kat_events: more events than cap");
    const size_t ne = (size_t)std::min(*n_events, cap);
    if (ne) { VR_DOWN(t, d_t, ne * 4); VR_DOWN(entering, d_e, ne * 4); VR_DOWN(index, d_i, ne * 4); }
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

int vr_kat_gather(vr_context* ctx, int n, const float* o3, const float* d3, int cap, int32_t* counts, uint32_t* idx) {
    if (!ctx || n <= 0 || !o3 || !d3 || cap <= 0 || !counts || !idx) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_gather: bad arguments");
    if (!ctx->has_gmm) return fail(ctx, VR_ERR_NO_SCENE, "vr_kat_gather: no Gaussian mixture set");
    if (n > (1 << 22)) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_gather: at most 2^22 rays per call");
    VR_CUDA(cudaSetDevice(ctx->device));
    int rc = ensure_origin_reach(ctx, n, o3);
    if (rc) return rc;
    const size_t m = (size_t)n;
    VR_CUDA(ctx->wavepool.ensure(wave_pool_bytes(n)));
    VR_CUDA(ctx->scratch.ensure(m * (24 + 16 + 4 + 4 * (size_t)cap) + 256));
    char* b = ctx->scratch.as<char>();
    float4* d_samp = (float4*)b; float* d_o = (float*)(b + m * 16); float* d_d = d_o + 3 * m;
    int* d_cnt = (int*)(d_d + 3 * m); uint32_t* d_idx = (uint32_t*)(d_cnt + m);
    VR_UP(d_o, o3, m * 12); VR_UP(d_d, d3, m * 12);
    VR_CUDA(cudaMemsetAsync(d_idx, 0xff, m * 4 * (size_t)cap, ctx->stream));
    if (run_wave_gather_kat(ctx->scene_dev(), n, d_o, d_d, ctx->wavepool.p, d_samp, ctx->stats.as<unsigned long long>(), cap, d_cnt, d_idx,
                            ctx->num_sms, ctx->stream) < 0)
        return cuda_fail(ctx, cudaGetLastError(), "vr_kat_gather");
    VR_DOWN(counts, d_cnt, m * 4); VR_DOWN(idx, d_idx, m * 4 * (size_t)cap);
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    VR_CUDA(cudaGetLastError());
    return VR_OK;
}

int vr_kat_libm(vr_context* ctx, int n, const float* x, float* sin_out, float* cos_out, float* acos_out) {
    if (!ctx || n <= 0 || !x || !sin_out || !cos_out || !acos_out) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_libm: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t m = (size_t)n;
    VR_CUDA(ctx->scratch.ensure(m * 16));
    float* b = ctx->scratch.as<float>();
    VR_UP(b, x, m * 4);
    launch_kat_libm(n, b, b + m, b + 2 * m, b + 3 * m, ctx->stream);
    VR_DOWN(sin_out, b + m, m * 4); VR_DOWN(cos_out, b + 2 * m, m * 4); VR_DOWN(acos_out, b + 3 * m, m * 4);
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

int vr_kat_erfinv(vr_context* ctx, int n, const double* x, double* y) {
    if (!ctx || n <= 0 || !x || !y) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_kat_erfinv: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t m = (size_t)n;
    VR_CUDA(ctx->scratch.ensure(m * 16));
    double* b = ctx->scratch.as<double>();
    VR_UP(b, x, m * 8);
    launch_kat_erfinv(n, b, b + m, ctx->stream);
    VR_DOWN(y, b + m, m * 8);
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

}  // extern "C"
