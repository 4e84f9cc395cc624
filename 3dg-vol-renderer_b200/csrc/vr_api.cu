// vr_api.cu -- C-ABI of libvolrender_b200.so: context, scene upload, Integrator::render.
// There is deliberately NO CPU code path: without a CUDA device vr_create fails with
// VR_ERR_NO_DEVICE and nothing else can be called.
#include <algorithm>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "vr_bvh.h"
#include "vr_lightgrid.h"
#include "vr_context.h"

using namespace vr;

vr::SceneDev vr_context::scene_dev() const {
    SceneDev S{};
    S.gauss = gauss_leaf.as<float4>();
    S.nodes = nodes.as<float4>();
    S.qnodes = qnodes.as<uint4>();
    for (int k = 0; k < 3; ++k) { S.qorg[k] = qorg[k]; S.qcell[k] = qcell[k]; }
    S.lg_start = lg_start.as<uint32_t>(); S.lg_gi = lg_gi.as<uint32_t>(); S.lg_dmin = lg_dmin.as<float>();
    S.lg_res = lg_res; S.lg_mask = lg_valid ? lg_mask : 0u;
    S.cg_start = cg_start.as<uint32_t>(); S.cg_gi = cg_gi.as<uint32_t>(); S.cg_dmin = cg_dmin.as<float>();
    S.cg_res = 0;   // switched on per render (vr_render.cu) when the grid is valid for this camera
    S.g3_start = g3_start.as<uint32_t>(); S.g3_gi = g3_gi.as<uint32_t>();
    for (int k = 0; k < 3; ++k) { S.g3_org[k] = g3_org[k]; S.g3_inv[k] = g3_inv[k]; S.g3_cell[k] = g3_cell[k]; }
    S.g3_res = 0;   // switched on per render
    S.orig = perm.as<uint32_t>();
    S.n = n;
    S.nl = nl;
    S.lpos = lpos.as<float>();
    S.lint = lint.as<float>();
    S.env[0] = env[0]; S.env[1] = env[1]; S.env[2] = env[2];
    return S;
}

// speculative light-grid build (vr_context.h): inputs are copies, so the job outlives any change of the context's vectors
struct vr_context::LightGridJob {
    std::vector<float> lo, hi, lpos;
    int n = 0, nl = 0, res = 0;
    vr::LightGridHost grid;
    std::thread worker;
};

namespace vr {
void cancel_light_grid_job(vr_context* c) {
    if (!c || !c->lg_job) return;
    if (c->lg_job->worker.joinable()) c->lg_job->worker.join();
    delete c->lg_job;
    c->lg_job = nullptr;
}
__global__ void remap_entries_kernel(uint32_t* __restrict__ gi, size_t count, const uint32_t* __restrict__ leaf_of_orig) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) gi[i] = __ldg(leaf_of_orig + gi[i]);
}
int fail(vr_context* c, int code, const std::string& msg) {
    if (c) c->err = msg;
    return code;
}
int cuda_fail(vr_context* c, cudaError_t e, const char* what) {
    std::string m = std::string(what) + ": " + cudaGetErrorString(e);
    cudaGetLastError();
    return fail(c, e == cudaErrorMemoryAllocation ? VR_ERR_OUT_OF_MEMORY : VR_ERR_CUDA, m);
}
}  // namespace vr

extern "C" {

const char* vr_version(void) { return "volrender_b200 0.1 (sm_100a)"; }

int vr_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int vr_create(vr_context** out, int device_id) {
    if (!out) return VR_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    int ndev = vr_device_count();
    if (ndev <= 0 || device_id < 0 || device_id >= ndev) return VR_ERR_NO_DEVICE;
    vr_context* ctx = new vr_context();
    ctx->device = device_id;
    *out = ctx;  // returned even on failure so that vr_last_error can be read; vr_destroy is still required
    VR_CUDA(cudaSetDevice(device_id));
    cudaDeviceProp prop;
    VR_CUDA(cudaGetDeviceProperties(&prop, device_id));
    if (prop.major < 10) return fail(ctx, VR_ERR_UNSUPPORTED, "device is not sm_100 class: kernels are built for sm_100a only");
    ctx->num_sms = prop.multiProcessorCount;
    VR_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    for (auto& e : ctx->ev) VR_CUDA(cudaEventCreate(&e));
    VR_CUDA(cudaMallocHost((void**)&ctx->h_stats, sizeof(unsigned long long) * kNumStats));
    VR_CUDA(cudaMallocHost((void**)&ctx->h_counters, sizeof(unsigned int) * 32));
    VR_CUDA(cudaMallocHost((void**)&ctx->h_wave_counters, sizeof(unsigned int) * kMaxWaveParts * kWaveCounters));
    VR_CUDA(cudaMallocHost((void**)&ctx->h_build_meta, sizeof(unsigned int) * 32));
    ctx->wave_graphs = wave_graph_cache_create();
    VR_CUDA(ctx->stats.ensure(sizeof(unsigned long long) * kNumStats));
    VR_CUDA(ctx->counters.ensure(sizeof(unsigned int) * 4));
    ctx->blocks_per_sm_main = std::max(1, path_kernel_blocks_per_sm(false));
    ctx->blocks_per_sm_arena = std::max(1, std::min(2, path_kernel_blocks_per_sm(true)));
    return VR_OK;
}

void vr_destroy(vr_context* ctx) {
    if (!ctx) return;
    vr::cancel_light_grid_job(ctx);
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    sfd_destroy(ctx);
    comm_destroy(ctx);
    ctx->comm_scratch.release();
    for (DevBuf* b : {&ctx->mean3, &ctx->cov6, &ctx->density, &ctx->albedo, &ctx->gauss_orig, &ctx->gauss_leaf, &ctx->nodes,
                      &ctx->perm, &ctx->leaf_of_orig, &ctx->box_lo, &ctx->box_hi, &ctx->lam, &ctx->qnodes, &ctx->lg_start, &ctx->lg_gi,
                      &ctx->lg_dmin, &ctx->lg_scratch, &ctx->cg_start, &ctx->cg_gi, &ctx->cg_dmin, &ctx->g3_start, &ctx->g3_gi, &ctx->g3_cursor, &ctx->spheres, &ctx->lpos, &ctx->lint,
                      &ctx->image, &ctx->sampleL, &ctx->counters, &ctx->stats, &ctx->overflow_list, &ctx->arena,
                      &ctx->record_bits, &ctx->scratch, &ctx->wavepool})
        b->release();
    if (ctx->h_stats) cudaFreeHost(ctx->h_stats);
    if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
    if (ctx->h_wave_counters) cudaFreeHost(ctx->h_wave_counters);
    if (ctx->h_build_meta) cudaFreeHost(ctx->h_build_meta);
    for (DevBuf* b : {&ctx->lbvh_keys, &ctx->lbvh_tree, &ctx->lbvh_tmp}) b->release();
    vr::wave_graph_cache_destroy(ctx->wave_graphs);
    for (cudaStream_t ws : ctx->wave_streams) if (ws) cudaStreamDestroy(ws);
    for (auto& e : ctx->ev) if (e) cudaEventDestroy(e);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* vr_last_error(const vr_context* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
void* vr_stream(vr_context* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int vr_set_option(vr_context* ctx, const char* name, double value) {
    if (!ctx || !name) return VR_ERR_INVALID_ARGUMENT;
    if (!strcmp(name, "hit_capacity")) { ctx->hit_capacity = (int)value; ctx->arena.release(); return VR_OK; }
    if (!strcmp(name, "max_pass_samples")) { ctx->max_pass_samples = std::max(1.0, value); return VR_OK; }
    if (!strcmp(name, "blocks_per_sm")) { ctx->blocks_per_sm_main = std::max(1, (int)value); return VR_OK; }
    if (!strcmp(name, "engine")) { ctx->engine = (int)value; return VR_OK; }
    if (!strcmp(name, "post_split")) { ctx->post_split = (int)value != 0; return VR_OK; }
    if (!strcmp(name, "post_pair")) { ctx->post_pair = (int)value != 0; return VR_OK; }
    if (!strcmp(name, "post_defer")) { ctx->post_defer = std::max(0, std::min(2, (int)value)); return VR_OK; }
    if (!strcmp(name, "trav")) { const int v = (int)value; ctx->trav_version = (v == 1 || v == 3) ? v : 2; return VR_OK; }   // 3: trav2 except the light-grid walk
    if (!strcmp(name, "hierarchy")) { ctx->hierarchy_builder = (int)value != 0; return VR_OK; }   // 0 device LBVH, 1 host binned SAH; next vr_set_gmm*
    if (!strcmp(name, "box_kappa")) { ctx->box_kappa = std::max(0.0, value); return VR_OK; }   // takes effect at the next vr_set_gmm*
    if (!strcmp(name, "light_grid")) { vr::cancel_light_grid_job(ctx); ctx->use_light_grid = (int)value != 0; ctx->lg_valid = false; return VR_OK; }
    if (!strcmp(name, "light_grid_min_samples")) { ctx->light_grid_min_samples = value; return VR_OK; }
    if (!strcmp(name, "grid3d")) { ctx->use_grid3d = (int)value != 0; return VR_OK; }
    if (!strcmp(name, "grid3d_min_samples")) { ctx->grid3d_min_samples = value; return VR_OK; }
    if (!strcmp(name, "grid3d_res")) { ctx->g3_res_option = std::max(0, (int)value); ctx->g3_valid = false; return VR_OK; }
    if (!strcmp(name, "camera_grid")) { ctx->use_camera_grid = (int)value != 0; return VR_OK; }
    if (!strcmp(name, "camera_grid_min_samples")) { ctx->camera_grid_min_samples = value; return VR_OK; }
    if (!strcmp(name, "light_grid_res")) { ctx->cg_valid = false; vr::cancel_light_grid_job(ctx); ctx->lg_res = std::min(1024, std::max(8, (int)value)); ctx->lg_valid = false; return VR_OK; }
    if (!strcmp(name, "wave_parts")) { ctx->wave_parts = std::max(1, std::min(kMaxWaveParts, (int)value)); ctx->wavepool.release(); return VR_OK; }
    if (!strcmp(name, "wave_slots")) { ctx->wave_slots = std::max(0, (int)value); ctx->wavepool.release(); return VR_OK; }
    return fail(ctx, VR_ERR_INVALID_ARGUMENT, std::string("unknown option ") + name);
}

// ------------------------------------------------------------------------------------------
// scene
// ------------------------------------------------------------------------------------------
int vr_set_gmm(vr_context* ctx, int n, const float* mean3, const float* cov6, const float* density,
               const float* albedo) {
    if (!ctx || n < 0 || (n > 0 && (!mean3 || !cov6 || !density || !albedo))) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_set_gmm: bad arguments");
    if (n >= (1 << 28)) return fail(ctx, VR_ERR_UNSUPPORTED, "vr_set_gmm: more than 2^28 Gaussians");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t m = (size_t)std::max(n, 1);
    VR_CUDA(ctx->mean3.ensure(m * 12)); VR_CUDA(ctx->cov6.ensure(m * 24));
    VR_CUDA(ctx->density.ensure(m * 4)); VR_CUDA(ctx->albedo.ensure(m * 4));
    if (n > 0) {
        VR_CUDA(cudaMemcpyAsync(ctx->mean3.p, mean3, (size_t)n * 12, cudaMemcpyHostToDevice, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(ctx->cov6.p, cov6, (size_t)n * 24, cudaMemcpyHostToDevice, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(ctx->density.p, density, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(ctx->albedo.p, albedo, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    return set_gmm_from_device(ctx, n);
}

int vr_set_gmm_params(vr_context* ctx, int n, const float* params11) {
    if (!ctx || n < 0 || (n > 0 && !params11)) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_set_gmm_params: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    const size_t m = (size_t)std::max(n, 1);
    VR_CUDA(ctx->mean3.ensure(m * 12)); VR_CUDA(ctx->cov6.ensure(m * 24));
    VR_CUDA(ctx->density.ensure(m * 4)); VR_CUDA(ctx->albedo.ensure(m * 4));
    VR_CUDA(ctx->scratch.ensure(m * 44));
    if (n > 0) {
        VR_CUDA(cudaMemcpyAsync(ctx->scratch.p, params11, (size_t)n * 44, cudaMemcpyHostToDevice, ctx->stream));
        launch_unpack_params(n, ctx->scratch.as<float>(), ctx->mean3.as<float>(), ctx->cov6.as<float>(),
                             ctx->density.as<float>(), ctx->albedo.as<float>(), ctx->stream);
    }
    return set_gmm_from_device(ctx, n);
}

int vr_unpack_params(vr_context* ctx, int n, const float* params11, float* mean3, float* cov6, float* density,
                     float* albedo) {
    if (!ctx || n <= 0 || !params11 || !mean3 || !cov6 || !density || !albedo) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_unpack_params: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    VR_CUDA(ctx->scratch.ensure((size_t)n * (44 + 12 + 24 + 8)));
    float* d = ctx->scratch.as<float>();
    float *dm = d + (size_t)11 * n, *dc = dm + (size_t)3 * n, *dd = dc + (size_t)6 * n, *da = dd + n;
    VR_CUDA(cudaMemcpyAsync(d, params11, (size_t)n * 44, cudaMemcpyHostToDevice, ctx->stream));
    launch_unpack_params(n, d, dm, dc, dd, da, ctx->stream);
    VR_CUDA(cudaMemcpyAsync(mean3, dm, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(cov6, dc, (size_t)n * 24, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(density, dd, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(albedo, da, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

int vr_set_lights(vr_context* ctx, int nl, const float* pos3, const float* intensity3, const float* env_rgb) {
    if (!ctx || nl < 0 || (nl > 0 && (!pos3 || !intensity3))) return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_set_lights: bad arguments");
    VR_CUDA(cudaSetDevice(ctx->device));
    VR_CUDA(ctx->lpos.ensure((size_t)std::max(nl, 1) * 12));
    VR_CUDA(ctx->lint.ensure((size_t)std::max(nl, 1) * 12));
    if (nl > 0) {
        VR_CUDA(cudaMemcpyAsync(ctx->lpos.p, pos3, (size_t)nl * 12, cudaMemcpyHostToDevice, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(ctx->lint.p, intensity3, (size_t)nl * 12, cudaMemcpyHostToDevice, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    // the direction grids depend on the light POSITIONS only: setting the same lights again (the C++ Integrator::render does
    // so on every call) keeps the grids and a build in flight
    const bool same_positions = nl == ctx->nl && ctx->h_lpos.size() == (size_t)3 * std::max(nl, 0) &&
                                (nl == 0 || memcmp(ctx->h_lpos.data(), pos3, sizeof(float) * 3 * (size_t)nl) == 0);
    if (!same_positions) {
        vr::cancel_light_grid_job(ctx);   // built for the previous lights
        ctx->h_lpos.assign(pos3, pos3 + (size_t)3 * std::max(nl, 0));
        ctx->lg_valid = false;
    }
    ctx->nl = nl;
    if (env_rgb) { ctx->env[0] = env_rgb[0]; ctx->env[1] = env_rgb[1]; ctx->env[2] = env_rgb[2]; }
    return VR_OK;
}

int vr_set_camera(vr_context* ctx, int kind, const float* pos3, const float* view3, const float* right3,
                  const float* up3, const float* pinhole3) {
    if (!ctx || !pos3 || !view3 || !right3 || !up3 || (kind != VR_CAMERA_PINHOLE && kind != VR_CAMERA_ORTHOGRAPHIC) ||
        (kind == VR_CAMERA_PINHOLE && !pinhole3))
        return fail(ctx, VR_ERR_INVALID_ARGUMENT, "vr_set_camera: bad arguments");
    ctx->cam.kind = kind;
    for (int k = 0; k < 3; ++k) {
        ctx->cam.pos[k] = pos3[k]; ctx->cam.view[k] = view3[k]; ctx->cam.right[k] = right3[k]; ctx->cam.up[k] = up3[k];
        ctx->cam.pinhole[k] = pinhole3 ? pinhole3[k] : 0.0f;
    }
    ctx->has_camera = true;
    return VR_OK;
}

}  // extern "C"

namespace vr {
// Sigma^-1 / norm / boxes on the device, hierarchy on the host (vr_bvh.cpp), permutation on the device
int set_gmm_from_device(vr_context* ctx, int n, const float* reach_RX) {
    static const bool time_setup = getenv("VR_TIME_SETUP") != nullptr;   // stderr: host time of each phase of the scene setup
    const auto ts0 = std::chrono::steady_clock::now();
    double ts_prev = 0.0;
    auto lap = [&](const char* what) {
        if (!time_setup) return;
        const double t = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - ts0).count();
        fprintf(stderr, "[setup] %-28s %7.3f ms\n", what, t - ts_prev);
        ts_prev = t;
    };
    NvtxRange range_setup("vr.set_gmm");
    cancel_light_grid_job(ctx);
    ctx->has_gmm = false;
    ctx->rec_valid = false;
    ctx->lg_valid = false;
    ctx->cg_valid = false;
    ctx->g3_valid = false;
    ctx->h_box_lo.clear(); ctx->h_box_hi.clear(); ctx->h_leaf_of_orig.clear();
    const size_t m = (size_t)std::max(n, 1);
    VR_CUDA(ctx->gauss_orig.ensure(m * 48)); VR_CUDA(ctx->gauss_leaf.ensure(m * 48));
    VR_CUDA(ctx->perm.ensure(m * 4)); VR_CUDA(ctx->leaf_of_orig.ensure(m * 4));
    VR_CUDA(ctx->box_lo.ensure(m * 12)); VR_CUDA(ctx->box_hi.ensure(m * 12)); VR_CUDA(ctx->lam.ensure(m * 4));
    VR_CUDA(ctx->nodes.ensure(64)); VR_CUDA(ctx->qnodes.ensure(32));
    ctx->n = n; ctx->n_nodes = 0;
    ctx->built_R = ctx->built_X = 0.f;
    if (n > 0) {
        launch_prepare_gaussians(n, ctx->mean3.as<float>(), ctx->cov6.as<float>(), ctx->density.as<float>(),
                                 ctx->albedo.as<float>(), ctx->gauss_orig.as<float4>(), ctx->box_lo.as<float>(),
                                 ctx->box_hi.as<float>(), ctx->lam.as<float>(), ctx->stream);
        if (ctx->hierarchy_builder == 0) {
            // everything on the device (vr_lbvh.cu): reach + slack of the boxes, Morton order, radix tree, both node formats
            int rc = build_hierarchy_device(ctx, n, reach_RX);
            if (rc) return rc;
            lap("device hierarchy build");
            launch_permute_gaussians(n, ctx->gauss_orig.as<float4>(), ctx->perm.as<uint32_t>(), ctx->gauss_leaf.as<float4>(),
                                     ctx->leaf_of_orig.as<uint32_t>(), ctx->stream);
            VR_CUDA(cudaGetLastError());
            ctx->h_box_lo.clear(); ctx->h_box_hi.clear(); ctx->h_leaf_of_orig.clear();   // (host copies, if ever needed, are fetched lazily)
            ctx->has_gmm = true;
            return VR_OK;
        }
        std::vector<float> lo((size_t)3 * n), hi((size_t)3 * n), lam((size_t)n);
        VR_CUDA(cudaMemcpyAsync(lo.data(), ctx->box_lo.p, lo.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(hi.data(), ctx->box_hi.p, hi.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(lam.data(), ctx->lam.p, lam.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        // the boxes come back TIGHT; they are inflated for the reach of the ray origins (vr_bvh.h): the camera that is set,
        // or origins within 8 units of the mixture when there is none yet, and never less than a caller-stated reach
        BoxReach reach = box_reach(n, lo.data(), hi.data(), ctx->has_camera ? ctx->cam.pos : nullptr, ctx->bb_lo, ctx->bb_hi);
        const BoxReach dflt = box_reach_of(ctx->bb_lo, ctx->bb_hi, nullptr);
        reach.R = std::max(reach.R, dflt.R); reach.X = std::max(reach.X, dflt.X);
        if (reach_RX) { reach.R = std::max(reach.R, reach_RX[0]); reach.X = std::max(reach.X, reach_RX[1]); }
        inflate_boxes(n, lo.data(), hi.data(), lam.data(), reach, ctx->box_kappa);
        ctx->built_R = reach.R; ctx->built_X = reach.X;
        lap("inverse covariances, boxes");
        if (ctx->lg_speculate && ctx->use_light_grid && ctx->nl > 0 && (int)ctx->h_lpos.size() == 3 * ctx->nl) {
            vr_context::LightGridJob* job = new vr_context::LightGridJob;
            job->lo = lo; job->hi = hi; job->lpos = ctx->h_lpos; job->n = n; job->nl = ctx->nl; job->res = ctx->lg_res;
            job->worker = std::thread([job] {
                build_light_grids(job->n, job->lo.data(), job->hi.data(), nullptr, job->nl, job->lpos.data(), job->res, &job->grid);
            });
            ctx->lg_job = job;
        }
        BvhBuildResult bvh;
        { NvtxRange r("vr.set_gmm.hierarchy_build"); build_bvh(n, lo.data(), hi.data(), &bvh); }
        lap("hierarchy build");
        if (bvh.max_depth > kStackDepth - 2)
            return fail(ctx, VR_ERR_UNSUPPORTED, "hierarchy deeper than the traversal stack");
        VR_CUDA(ctx->nodes.ensure(bvh.nodes.size() * sizeof(BvhNode)));
        VR_CUDA(cudaMemcpyAsync(ctx->nodes.p, bvh.nodes.data(), bvh.nodes.size() * sizeof(BvhNode), cudaMemcpyHostToDevice, ctx->stream));
        std::vector<BvhNodeQ> qn;
        BvhQuantGrid grid;
        quantize_bvh(bvh, &grid, &qn);
        lap("quantisation");
        for (int k = 0; k < 3; ++k) { ctx->qorg[k] = grid.org[k]; ctx->qcell[k] = grid.cell[k]; }
        VR_CUDA(ctx->qnodes.ensure(qn.size() * sizeof(BvhNodeQ)));
        VR_CUDA(cudaMemcpyAsync(ctx->qnodes.p, qn.data(), qn.size() * sizeof(BvhNodeQ), cudaMemcpyHostToDevice, ctx->stream));
        VR_CUDA(cudaMemcpyAsync(ctx->perm.p, bvh.perm.data(), (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
        launch_permute_gaussians(n, ctx->gauss_orig.as<float4>(), ctx->perm.as<uint32_t>(), ctx->gauss_leaf.as<float4>(),
                                 ctx->leaf_of_orig.as<uint32_t>(), ctx->stream);
        VR_CUDA(cudaStreamSynchronize(ctx->stream));
        VR_CUDA(cudaGetLastError());
        ctx->h_leaf_of_orig.resize((size_t)n);
        for (int i = 0; i < n; ++i) ctx->h_leaf_of_orig[bvh.perm[(size_t)i]] = (uint32_t)i;
        ctx->h_box_lo.swap(lo); ctx->h_box_hi.swap(hi);
        ctx->n_nodes = (int)bvh.nodes.size();
        ctx->bvh_depth = bvh.max_depth;
        lap("upload, permutation");
    }
    ctx->has_gmm = true;
    return VR_OK;
}

// inflated boxes (original order) and the leaf order on the host, for the light-grid builder; the device build leaves them on
// the device until a render large enough for grids asks for them
int fetch_host_boxes(vr_context* ctx) {
    const int n = ctx->n;
    if (n <= 0 || ((int)ctx->h_box_lo.size() == 3 * n && (int)ctx->h_leaf_of_orig.size() == n)) return VR_OK;
    ctx->h_box_lo.resize((size_t)3 * n); ctx->h_box_hi.resize((size_t)3 * n); ctx->h_leaf_of_orig.resize((size_t)n);
    VR_CUDA(cudaMemcpyAsync(ctx->h_box_lo.data(), ctx->box_lo.p, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(ctx->h_box_hi.data(), ctx->box_hi.p, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaMemcpyAsync(ctx->h_leaf_of_orig.data(), ctx->leaf_of_orig.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    VR_CUDA(cudaStreamSynchronize(ctx->stream));
    return VR_OK;
}

// candidate grid of the camera rays: (re)built when the mixture or the pinhole changed; device builder only
int ensure_camera_grid(vr_context* ctx) {
    if (ctx->cam.kind != VR_CAMERA_PINHOLE || ctx->hierarchy_builder != 0 || ctx->n <= 0) { ctx->cg_ok = false; return VR_OK; }
    if (ctx->cg_valid && memcmp(ctx->cg_pinhole, ctx->cam.pinhole, sizeof(float) * 3) == 0) return VR_OK;
    NvtxRange range("vr.camera_grid");
    int rc = build_camera_grid_device(ctx);
    if (rc) return rc;
    memcpy(ctx->cg_pinhole, ctx->cam.pinhole, sizeof(float) * 3);
    ctx->cg_valid = true;
    return VR_OK;
}

// uniform grid over the mixture for the bounce rays' first stretch: (re)built when the mixture changed; device builder only
int ensure_grid3d(vr_context* ctx) {
    if (ctx->hierarchy_builder != 0 || ctx->n <= 0) { ctx->g3_ok = false; return VR_OK; }
    if (ctx->g3_valid) return VR_OK;
    NvtxRange range("vr.grid3d");
    int rc = build_grid3d_device(ctx);
    if (rc) return rc;
    ctx->g3_valid = true;
    return VR_OK;
}

// A hierarchy inflated for ray origins within (R, X) must not be used by rays from farther out: the slab test could
// reject a box whose ellipsoid the reference's FP32 quadric accepts (vr_bvh.h).  Rebuild with 25 % head room instead.
int ensure_reach(vr_context* ctx, float R, float X) {
    if (!ctx->has_gmm || ctx->n <= 0) return VR_OK;
    if (R <= ctx->built_R && X <= ctx->built_X) return VR_OK;
    const float rx[2] = {std::max(R, ctx->built_R) * 1.25f, std::max(X, ctx->built_X) * 1.25f};
    return set_gmm_from_device(ctx, ctx->n, rx);
}
int ensure_camera_reach(vr_context* ctx) {
    if (!ctx->has_gmm || ctx->n <= 0 || !ctx->has_camera) return VR_OK;
    const BoxReach need = box_reach_of(ctx->bb_lo, ctx->bb_hi, ctx->cam.pos);
    return ensure_reach(ctx, need.R, need.X);
}

// per-light direction grids for the wavefront engine's point-light shadow rays (vr_lightgrid.h); (re)built when the
// mixture or the lights changed since the last render
int ensure_light_grids(vr_context* ctx) {
    if (ctx->lg_valid) return VR_OK;
    NvtxRange range_grids("vr.light_grids");
    ctx->lg_mask = 0;
    VR_CUDA(ctx->lg_start.ensure(16)); VR_CUDA(ctx->lg_gi.ensure(16)); VR_CUDA(ctx->lg_dmin.ensure(16));
    if (ctx->hierarchy_builder == 0 && ctx->use_light_grid && ctx->n > 0 && ctx->nl > 0 && (int)ctx->h_lpos.size() == 3 * ctx->nl) {
        int rcd = build_light_grids_device(ctx);   // count / scan / fill / sort on the device (vr_lightgrid_dev.cu)
        if (rcd) return rcd;
        cancel_light_grid_job(ctx);
        ctx->lg_valid = true;
        return VR_OK;
    }
    if (ctx->use_light_grid && ctx->n > 0 && ctx->nl > 0 && (int)ctx->h_lpos.size() == 3 * ctx->nl) {
        int rcb = fetch_host_boxes(ctx);
        if (rcb) return rcb;
        LightGridHost g_sync;
        const bool speculated = ctx->lg_job != nullptr && ctx->lg_job->n == ctx->n && ctx->lg_job->nl == ctx->nl && ctx->lg_job->res == ctx->lg_res;
        if (speculated) {
            if (ctx->lg_job->worker.joinable()) ctx->lg_job->worker.join();
        } else {
            build_light_grids(ctx->n, ctx->h_box_lo.data(), ctx->h_box_hi.data(), ctx->h_leaf_of_orig.data(), ctx->nl, ctx->h_lpos.data(),
                              ctx->lg_res, &g_sync);
        }
        const LightGridHost& g = speculated ? ctx->lg_job->grid : g_sync;
        if (g.mask) {
            VR_CUDA(ctx->lg_start.ensure(g.cell_start.size() * 4));
            VR_CUDA(ctx->lg_gi.ensure(std::max<size_t>(g.entry_gi.size(), 1) * 4));
            VR_CUDA(ctx->lg_dmin.ensure(std::max<size_t>(g.entry_dmin.size(), 1) * 4));
            VR_CUDA(cudaMemcpyAsync(ctx->lg_start.p, g.cell_start.data(), g.cell_start.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
            if (!g.entry_gi.empty()) {
                VR_CUDA(cudaMemcpyAsync(ctx->lg_gi.p, g.entry_gi.data(), g.entry_gi.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
                VR_CUDA(cudaMemcpyAsync(ctx->lg_dmin.p, g.entry_dmin.data(), g.entry_dmin.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
                if (speculated)   // original -> leaf-order indices
                    remap_entries_kernel<<<(unsigned int)((g.entry_gi.size() + 255) / 256), 256, 0, ctx->stream>>>(
                        ctx->lg_gi.as<uint32_t>(), g.entry_gi.size(), ctx->leaf_of_orig.as<uint32_t>());
            }
            VR_CUDA(cudaStreamSynchronize(ctx->stream));
            ctx->lg_mask = g.mask;
        }
    }
    cancel_light_grid_job(ctx);
    ctx->lg_valid = true;
    return VR_OK;
}
}  // namespace vr
