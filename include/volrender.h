/* include/volrender.h -- C-ABI of libvolrender_b200.so
 *
 * The drop-in boundary of the B200 implementation of the per-ray volume-integration hot path of
 * wantonsushi/3DG-vol-renderer.  The reference has no FFI: its boundary is a set of C++ abstract
 * classes.  The host-side mirror of those classes
 * (the headers under 3dg-vol-renderer_b200/host/include) flattens Scene/Camera into the POD arrays below and
 * calls these entry points; nothing else crosses into CUDA.
 *
 * Conventions: every function returns a vr_status (0 = ok) and never throws; a context is
 * opaque, owns all device memory of ONE device, and is driven by one caller thread; all pointer
 * arguments are HOST pointers unless the name ends in _dev.  Gaussian covariance is passed as
 * cov6 = cxx cxy cxz cyy cyz czz (the order of the "g" line of the scene format, scene.h:70-113).
 */
#ifndef VOLRENDER_H
#define VOLRENDER_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden */
#endif

typedef struct vr_context vr_context;

typedef enum vr_status {
    VR_OK = 0,
    VR_ERR_INVALID_ARGUMENT = 1,
    VR_ERR_CUDA = 2,            /* a CUDA runtime call failed; see vr_last_error */
    VR_ERR_NO_DEVICE = 3,       /* no CUDA device: the product has NO CPU fallback */
    VR_ERR_NO_SCENE = 4,        /* render/transmittance called before vr_set_gmm / vr_set_smm */
    VR_ERR_HIT_OVERFLOW = 5,    /* a ray overlapped more components than hit_capacity allows */
    VR_ERR_OUT_OF_MEMORY = 6,
    VR_ERR_UNSUPPORTED = 7
} vr_status;

/* Integrator::render implementations (integrator.h:49-57 and its subclasses) */
typedef enum vr_integrator {
    VR_INTEGRATOR_FREE_FLIGHT_SS = 0,  /* FreeFlightGaussians::render      integrator.h:300-408 */
    VR_INTEGRATOR_MULTI_SCATTER = 1,   /* MultiScatterGaussians::render    integrator.h:524-717 */
    VR_INTEGRATOR_MARCH_SPHERES = 2,   /* RayMarchingSpheres::render       test_integrators.h:23-135 */
    /* validation integrators over the Gaussian mixture: one ray per pixel, march_step / env_samples as below */
    VR_INTEGRATOR_TEST = 3,            /* TestIntegrator::render           integrator.h:65-94 */
    VR_INTEGRATOR_PURE_MARCH = 4,      /* PureRayMarching::render          integrator.h:100-267 */
    VR_INTEGRATOR_MARCH_GAUSSIANS = 5  /* RayMarchingGaussians::render     test_integrators.h:143-297 */
} vr_integrator;

/* runtime form of the #define selector distance_solvers.h:143-147 */
typedef enum vr_solver {
    VR_SOLVER_ANALYTIC_PLUS_NEWTON = 0, /* shipped default */
    VR_SOLVER_ANALYTIC_PLUS_BISECTION = 1,
    VR_SOLVER_NEWTON = 2,
    VR_SOLVER_BISECTION = 3,
    VR_SOLVER_UNIFORM = 4               /* distance_solvers.h:131-137; u is hashed from the segment (the reference's is unseeded) */
} vr_solver;

typedef enum vr_camera_kind { VR_CAMERA_PINHOLE = 0, VR_CAMERA_ORTHOGRAPHIC = 1 } vr_camera_kind;

/* Arguments of one Integrator::render call. */
typedef struct vr_render_params {
    int32_t integrator;    /* vr_integrator */
    int32_t solver;        /* vr_solver */
    int32_t width, height; /* image.get_width()/get_height() */
    int32_t spp;           /* num_samples (perfect square expected, integrator.h:562) */
    int32_t min_bounces;   /* MultiScatterGaussians min_scatter (default 5) */
    /* pixel window [x0,x1) x [y0,y1) to render (0,0,width,height = whole image) */
    int32_t x0, y0, x1, y1;
    /* row partition for multi-GPU: this context renders rows y with (y - y0) % part_count == part_rank */
    int32_t part_rank, part_count;
    int32_t record;        /* != 0: keep per-pixel contributing-Gaussian sets (integrator.h:616-644) */
    float march_step;      /* marchers: step_size (0.01) */
    int32_t env_samples;   /* marchers: env_samples (RayMarchingSpheres 5, the Gaussian marchers 20) */
    int32_t reserved[4];
} vr_render_params;

typedef struct vr_render_stats {
    double kernel_ms;        /* device time of the path kernel(s), CUDA events on the launch stream */
    double total_ms;         /* device time of the whole call incl. pixel reduction (+ copies for vr_render) */
    uint64_t samples;        /* camera paths traced = pixels rendered * spp */
    uint64_t scatter_rays;   /* free-flight rays traced (camera + bounce rays) */
    uint64_t shadow_rays;    /* transmittance (NEE) rays traced */
    uint64_t hits;           /* ray-component overlaps gathered on free-flight rays */
    uint64_t overflow_samples; /* samples re-traced by the large-capacity kernel */
    uint32_t kernel_launches;  /* kernels launched by this call */
    uint32_t reserved;
    /* EXECUTED work of the wavefront engine (what the device really did, next to the algorithmic counts of SURVEY.md 8d;
     * the overflow re-trace and the drain of a pass, < 2 % of the samples, are not counted) */
    uint64_t node_visits;      /* hierarchy slab steps (two child boxes each) + light-grid entries examined */
    uint64_t quadric_tests;    /* gaussian.h:126-164 evaluations (hits + rejected candidates) */
    uint64_t erf_evals;        /* erff evaluations (an optical_depth call of the reference = 2) */
    uint64_t mu_evals;         /* Gaussian::mu_t evaluations */
    uint64_t analytic_solves;  /* solve_for_t_given_tau calls */
    uint64_t iterative_solves; /* Newton / bisection solves */
} vr_render_stats;

/* ---- lifetime ------------------------------------------------------------------------- */
int vr_create(vr_context** out, int device_id);
void vr_destroy(vr_context* ctx);
const char* vr_last_error(const vr_context* ctx); /* never NULL; valid until the next call on ctx */
const char* vr_version(void);
int vr_device_count(void);                        /* 0 when no CUDA device is usable */
/* stream all work of ctx is launched on (cudaStream_t as void*), for event timing by the caller */
void* vr_stream(vr_context* ctx);

/* ---- scene: replaces Scene::load_GMM's in-memory result (scene.h:72-120, gaussian.h:52-109,
 *      gmm.h:75-90 BuildBVH).  Builds Sigma^-1, norm and the device BVH. ---------------------- */
int vr_set_gmm(vr_context* ctx, int n, const float* mean3, const float* cov6, const float* density,
               const float* albedo);
/* inverse path: 11 floats per Gaussian (gmm.h:583-674 apply_params_to_gmm_local) */
int vr_set_gmm_params(vr_context* ctx, int n, const float* params11);
/* SphereMixtureModel (smm.h:17-52) */
int vr_set_smm(vr_context* ctx, int n, const float* center3, const float* radius, const float* sigma_a,
               const float* sigma_s);
/* Scene::lights + env_color (scene.h:12-31) */
int vr_set_lights(vr_context* ctx, int nl, const float* pos3, const float* intensity3, const float* env_rgb);
/* Camera state after its constructor ran on the host (camera.h:15-22,38-43,57-62) */
int vr_set_camera(vr_context* ctx, int kind, const float* pos3, const float* view3, const float* right3,
                  const float* up3, const float* pinhole3);

/* ---- Integrator::render -------------------------------------------------------------------
 * vr_render: blocking; writes 3*width*height floats (row-major, 3*(y*W+x), image.h:55-60) for the
 * pixels this call owns (window + row partition); other pixels of out_rgb are left untouched.
 * vr_render_device: same work, result stays in the context's device image (vr_image_dev). */
int vr_render(vr_context* ctx, const vr_render_params* p, float* out_rgb, vr_render_stats* stats);
int vr_render_device(vr_context* ctx, const vr_render_params* p, vr_render_stats* stats);
const float* vr_image_dev(vr_context* ctx); /* device pointer, 3*W*H floats of the last render */
/* recorded per-pixel Gaussian sets of the last render with record != 0: counts[W*H] and, if
 * idx != NULL, up to cap indices per pixel (ascending original Gaussian index) */
int vr_get_recorded(vr_context* ctx, int* counts, uint32_t* idx, int cap);
/* per-SAMPLE radiance of the last vr_render* call of a free-flight integrator that fitted one pass
 * (max_pass_samples): count float4 = (r, g, b, bounce index when the path ended), ordered
 * [owned pixel, row-major in the window][si]; the si-ordered sum / spp is the pixel (integrator.h:702-706).
 * Debug / parity aid for single-sample attribution of pixel differences. */
int vr_get_samples(vr_context* ctx, float* out4, uint64_t count);

/* ---- GaussianMixtureModel::transmittance_up_to (gmm.h:517-578) on a batch of rays ----------
 * directions must be unit length (the Ray ctor normalises, ray.h:11-12) */
int vr_transmittance(vr_context* ctx, int nrays, const float* o3, const float* d3, const float* tmax, float* out);

/* ---- vr_transmittance for callers that cannot call cudaMalloc --------------------------------
 * No counterpart in the reference: added (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid allocation).  Same arguments, same kernel and same
 * results as vr_transmittance, but the rays and results are staged in a fixed file-scope device array
 * (32 bytes per ray, VR_NOALLOC_MAX_RAYS rays = 2 MiB held by every GPU that loads the library) instead
 * of the context's growing scratch buffer: the call performs no cudaMalloc / cudaFree and no allocating
 * library routine.  nrays > VR_NOALLOC_MAX_RAYS -> VR_ERR_INVALID_ARGUMENT; a batch whose origins lie
 * farther out than the culling hierarchy was built for -> VR_ERR_UNSUPPORTED (vr_transmittance would
 * rebuild, which allocates).  All calls on one GPU share the array: calls that overlap on the GPU race;
 * nothing in it needs clearing between calls (every byte read was written by the same call).
 * The first call on a device resolves the array's address (and, under lazy module loading, is when the
 * driver materialises the array there): make one warm-up call at set-up.  csrc/vr_noalloc.cu. */
#define VR_NOALLOC_MAX_RAYS 65536
int vr_transmittance_noalloc(vr_context* ctx, int nrays, const float* o3, const float* d3, const float* tmax, float* out);
int vr_transmittance_noalloc_max_rays(void);

/* ---- inverse path: inverse_integrator.h:67-241, optimizer.h:30-41 ---------------------------
 * vr_sfd_begin: uploads the target image (3*W*H floats) and the packed parameters, resets Adam.
 * vr_sfd_fdiff: for the CURRENT parameters renders the base image and S perturbed images
 *   (params + signs*eps) with recording, and returns F_diff[S*N] (inverse_integrator.h:157-179)
 *   restricted to the pixels this context owns (rp window/partition) -- partial sums that add
 *   across contexts / ranks / views.  mean_loss_sum/npix give the base L1 loss over owned pixels.
 * vr_sfd_apply: grad = mean_s F_diff*s/eps (:182-198), Adam step (optimizer.h:30-41), new params
 *   become current.  F_diff is the (all-reduced) sum over all ranks and views. */
int vr_sfd_begin(vr_context* ctx, int n, const float* params11, const float* eps11, float lr, float beta1,
                 float beta2, float adam_eps);
int vr_sfd_set_target(vr_context* ctx, int width, int height, const float* target_rgb);
/* multi-view fits: upload each view's target once, select the one the next vr_sfd_accumulate compares against */
int vr_sfd_set_target_view(vr_context* ctx, int view, int width, int height, const float* target_rgb);
int vr_sfd_select_view(vr_context* ctx, int view);
int vr_sfd_fdiff(vr_context* ctx, const vr_render_params* rp, int S, const int8_t* signs /* S*11n */,
                 double* F_diff /* S*n */, double* loss_sum, int64_t* loss_pixels);
int vr_sfd_apply(vr_context* ctx, int S, const int8_t* signs, const double* F_diff /* NULL: device buffer */,
                 float* grad_out /* 11n or NULL */);
/* multi-view / multi-GPU form of vr_sfd_fdiff: reset, then accumulate once per (camera, target)
 * view; partial sums live in a device int64 buffer of S*n FIXED-POINT values (scale 2^32) that a
 * caller may all-reduce in place (ncclAllReduce sum, int64 -- exact and order independent) before
 * vr_sfd_apply(..., F_diff = NULL, ...).  vr_sfd_read converts to double on the host. */
int vr_sfd_reset_fdiff(vr_context* ctx, int S);
int vr_sfd_accumulate(vr_context* ctx, const vr_render_params* rp, int S, const int8_t* signs);
int vr_sfd_read(vr_context* ctx, double* F_diff /* S*n or NULL */, double* loss_sum /* or NULL */);
int64_t* vr_sfd_fdiff_dev(vr_context* ctx);
/* ---- the one collective of the path: SUM all-reduce of the S*n int64 partial sums (+ the loss accumulators)
 * over NCCL, enqueued on the context's stream behind the last reduction kernel; exact and order independent (integers).
 * One process per device: rank 0 calls vr_comm_unique_id, the 128 bytes travel to every rank by the launcher's means, every
 * rank calls vr_comm_init_rank; then per iteration vr_sfd_reset_fdiff, vr_sfd_accumulate (per view), vr_sfd_allreduce,
 * vr_sfd_apply(F_diff = NULL).  One process driving several devices: vr_comm_init_all over its contexts, then
 * vr_sfd_allreduce_all (one NCCL group).  Contexts that share a device, or a machine without NCCL, get the same exact sum
 * through peer copies.  NCCL is bound at run time (libnccl.so.2); vr_comm_available() tells whether it was found. */
int vr_comm_available(void);
int vr_comm_version(void);                       /* ncclGetVersion, 0 when unavailable */
int vr_comm_unique_id(void* id128);              /* 128 bytes out */
int vr_comm_init_rank(vr_context* ctx, const void* id128, int nranks, int rank);
int vr_comm_init_all(vr_context** ctxs, int n);
int vr_comm_destroy(vr_context* ctx);
int vr_comm_size(const vr_context* ctx);         /* ranks of the context's communicator (1: none) */
int vr_sfd_allreduce(vr_context* ctx);
int vr_sfd_allreduce_all(vr_context** ctxs, int n);
int vr_sfd_get_params(vr_context* ctx, float* params11);
/* optimiser state for checkpoint / resume: current parameters, Adam moments m and v (11n floats each), step count t */
int vr_sfd_get_state(vr_context* ctx, float* params11, float* m, float* v, int* t);
int vr_sfd_set_state(vr_context* ctx, const float* params11, const float* m, const float* v, int t);
/* stand-alone AdamOptimizer::step on host arrays (state m, v, t owned by the caller) */
int vr_adam_step(vr_context* ctx, int D, float* params, const float* grads, float* m, float* v, int* t, float lr,
                 float beta1, float beta2, float eps);
/* apply_params_to_gmm_local math only: params11 -> mean3, cov6, density, albedo (device computed) */
int vr_unpack_params(vr_context* ctx, int n, const float* params11, float* mean3, float* cov6, float* density,
                     float* albedo);

/* ---- tuning ------------------------------------------------------------------------------- */
/* hit_capacity: per-ray overlap list capacity of the large-capacity kernel (default min(N, 4096)).
 * max_pass_samples: samples traced per pass (bounds the per-sample radiance buffer). */
int vr_set_option(vr_context* ctx, const char* name, double value);

/* ---- known-answer hooks: device implementations of the primitives, one thread per record ----
 * (rng.h:13-57, gaussian.h:111-297, camera.h:45-73, gmm.h:457-515) */
int vr_kat_pcg(vr_context* ctx, int n, const int32_t* xysi3, uint64_t* seed, uint32_t* u32x4, float* uniformx4);
int vr_kat_gaussian(vr_context* ctx, int n, const float* o3, const float* d3, const int32_t* gidx, const float* frac,
                    int32_t* hit, float* t0, float* t1, float* tau, float* mu_mid, int32_t* ok, float* t_solved);
int vr_kat_camera_rays(vr_context* ctx, int n, const float* uv2, float* o3, float* d3);
int vr_kat_events(vr_context* ctx, const float* o3, const float* d3, int cap, float* t, int32_t* entering,
                  uint32_t* index, int* n_events);
/* the wavefront engine's own free-flight traversal kernel (quantised hierarchy, candidate queue, batched quadric tests) on n
 * caller rays, nothing pruned: counts[n] = components overlapped (gmm.h:457-515: the quadric test decides), idx[n*cap] = their
 * original indices in gather order (first min(count, 16, cap); unused entries 0xffffffff).  Ray origins farther out than the
 * hierarchy was built for trigger the same rebuild a far camera does. */
int vr_kat_gather(vr_context* ctx, int n, const float* o3, const float* d3, int cap, int32_t* counts, uint32_t* idx);
/* the localised loss difference of the SFD estimator (inverse_integrator.h:157-179) on caller-supplied recorded sets
 * (bitmaps: (n+31)/32 words per pixel, bit g = Gaussian g) and per-pixel losses: F_fixed[n] = sum over the pixels in
 * base[g] | plus[g] of round(loss_plus * 2^32) - round(loss_base * 2^32), exact integers */
int vr_kat_fdiff(vr_context* ctx, int npix, int n, const uint32_t* bits_base, const uint32_t* bits_plus, const float* loss_base,
                 const float* loss_plus, int64_t* F_fixed);
int vr_kat_libm(vr_context* ctx, int n, const float* x, float* sin_out, float* cos_out, float* acos_out);
int vr_kat_erfinv(vr_context* ctx, int n, const double* x, double* y);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* VOLRENDER_H */
