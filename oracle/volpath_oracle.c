/* oracle/volpath_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, CPU restatement of the reference renderer's per-ray volume
 * integration path (wantonsushi/3DG-vol-renderer @ 9661b50).  Every function
 * cites the reference file:line it follows.  It is the checker the CUDA path
 * is compared against; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.  The product
 * (3dg-vol-renderer_b200/) never links, imports or calls anything in here.
 *
 * PARITY PINNED: tests/test_oracle_golden.py checks this file against golden
 * vectors emitted by the reference itself (oracle/ref_driver.cpp compiled over
 * the untouched reference headers -> oracle/_ref/ref_driver; generator:
 * oracle/make_golden.py; fixtures: tests/golden/).
 *
 * Deliberate deviations (results unaffected, see DESIGN.md "Oracle"):
 *  - BVH leaves are bounded by tight 3-sigma boxes (3*sqrt(Sigma_kk), padded)
 *    instead of the eigen-decomposition boxes of gaussian.h:304-319.  The
 *    quadric test (gaussian.h:126-164) decides every hit, so the hit SET is the
 *    same; only the pre-sort order of events can differ, which std::sort
 *    (unstable, gmm.h:511) leaves unspecified in the reference too.
 *  - events are ordered with qsort on t; ties are broken by (gather order).
 *  - the sphere / ray-marching integrators take their environment directions
 *    from a PCG stream seeded per (x, y, step) instead of the reference's
 *    std::random_device generator (integrator.h:13-28); that part
 *    of the path is therefore "parity unpinned" (statistical agreement only).
 *
 * Build: oracle/Makefile (gcc -O2 -fopenmp -ffp-contract=off -shared).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define VO_PI 3.14159265358979323846 /* std::numbers::pi (double) */
#define VO_EXPORT __attribute__((visibility("default")))

/* algorithmic-work counters: per-thread, folded into totals at the end of vo_render.
 * [0] free-flight rays V  [1] shadow rays  [2] ray-Gaussian overlaps H on free-flight rays
 * [3] optical_depth evaluations OD  [4] mu_t evaluations MU  [5] analytic solves SA
 * [6] Newton solves  [7] bisection solves  [8] camera paths */
enum { VO_C_V = 0, VO_C_SHADOW, VO_C_H, VO_C_OD, VO_C_MU, VO_C_SA, VO_C_NEWTON, VO_C_BISECT, VO_C_PATHS, VO_NCOUNT };
static __thread uint64_t vo_cnt[VO_NCOUNT];
static uint64_t vo_cnt_total[VO_NCOUNT];
VO_EXPORT void vo_counters_reset(void) { memset(vo_cnt_total, 0, sizeof(vo_cnt_total)); }
VO_EXPORT void vo_counters_get(uint64_t* out) { memcpy(out, vo_cnt_total, sizeof(vo_cnt_total)); }
static void vo_counters_fold(void) {
#pragma omp critical(vo_counters)
    { for (int i = 0; i < VO_NCOUNT; ++i) vo_cnt_total[i] += vo_cnt[i]; }
    memset(vo_cnt, 0, sizeof(vo_cnt));
}

/* Single-sample attribution: vo_render_samples can re-run a sample with its free-flight distance or
 * its sampled optical depth moved by a few ulp (vo_perturb) and reports the smallest relative margin of the sample's
 * discrete decisions (vo_min_margin): the comparison acc + seg > target (integrator.h:455), the Russian-roulette test
 * (:693) and the tolerance tests of the iterative solvers (distance_solvers.h:39,84,112).  A sample whose value moves
 * under such a perturbation is one an FP32 implementation with a different libm may legitimately disagree on. */
static __thread int vo_perturb;       /* 0 none, 1/2: t_scatter +/- 1 ulp, 3/4: target_tau * (1 +/- 2^-21), 5/6: t_scatter * (1 +/- 2^-20) */
static __thread float vo_min_margin;  /* reset per sample */
static inline void vo_margin(float lhs, float rhs) {
    float den = fabsf(rhs) > 1e-30f ? fabsf(rhs) : 1e-30f;
    float mg = fabsf(lhs - rhs) / den;
    if (mg < vo_min_margin) vo_min_margin = mg;
}

/* ------------------------------------------------------------------------- */
/* RNG -- include/rng.h:13-57                                                 */
/* ------------------------------------------------------------------------- */
VO_EXPORT uint64_t vo_splitmix64(uint64_t x) { /* rng.h:13-18 */
    x += 0x9e3779b97f4a7c15ULL;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    return x ^ (x >> 31);
}
typedef struct { uint64_t state, inc; } vo_pcg;

static inline uint32_t pcg_next(vo_pcg* r) { /* rng.h:38-44, incl. the non-standard shift-OR (F5) */
    uint64_t old = r->state;
    r->state = old * 6364136223846793005ULL + r->inc;
    uint32_t shifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    uint32_t lsh = ((0u - rot) + 1u) & 31u;
    return (shifted >> rot) | (shifted << lsh);
}
static inline void pcg_seed(vo_pcg* r, uint64_t seed_state, uint64_t seed_seq) { /* rng.h:29-35 */
    r->state = 0;
    r->inc = (seed_seq << 1) | 1;
    pcg_next(r);
    r->state += seed_state;
    pcg_next(r);
}
static inline float pcg_uniform(vo_pcg* r) { /* rng.h:47-49 */
    return (float)(pcg_next(r) >> 8) * (1.0f / 16777216.0f);
}
VO_EXPORT uint64_t vo_derive_path_seed(int x, int y, int si) { /* rng.h:52-57 */
    uint64_t seed = ((uint64_t)(int64_t)si << 32) | ((uint64_t)(int64_t)y << 16) | ((uint64_t)(int64_t)x);
    return vo_splitmix64(seed);
}
/* KAT helper: first n u32 outputs then n uniforms of the stream PCG32(derive_path_seed(x,y,si), 1) */
VO_EXPORT void vo_pcg_stream(int x, int y, int si, int n, uint32_t* out_u32, float* out_uniform) {
    vo_pcg r;
    pcg_seed(&r, vo_derive_path_seed(x, y, si), 1);
    for (int i = 0; i < n; ++i) out_u32[i] = pcg_next(&r);
    for (int i = 0; i < n; ++i) out_uniform[i] = pcg_uniform(&r);
}

/* ------------------------------------------------------------------------- */
/* small vector helpers (Eigen 3-vector semantics, float)                     */
/* ------------------------------------------------------------------------- */
typedef struct { float x, y, z; } v3;
static inline v3 v3_make(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 v3_add(v3 a, v3 b) { return v3_make(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 v3_sub(v3 a, v3 b) { return v3_make(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 v3_scale(float s, v3 a) { return v3_make(s * a.x, s * a.y, s * a.z); }
/* The vendored Eigen (3.4.90) evaluates a.dot(b) as a scalar loop of REAL fused multiply-adds:
 * r = a0*b0; r = fma(a1,b1,r); r = fma(a2,b2,r)  (Core/InnerProduct.h:117-131,229-231 ->
 * GenericPacketMath.h:1350-1354 -> numext::fma -> ::fmaf).  Verified bitwise against Eigen. */
static inline float v3_dot(v3 a, v3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
/* squaredNorm() is a redux: e0 + (e1 + e2) (Core/Redux.h:99-108, binary split), no fma */
static inline float v3_sqnorm(v3 a) { return a.x * a.x + (a.y * a.y + a.z * a.z); }
static inline v3 v3_cross(v3 a, v3 b) {
    return v3_make(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline v3 v3_normalized(v3 a) { /* Eigen normalized(): n / sqrt(squaredNorm) when > 0 */
    float z = v3_sqnorm(a);
    if (z > 0.0f) { float s = sqrtf(z); return v3_make(a.x / s, a.y / s, a.z / s); }
    return a;
}
typedef struct { v3 o, d; } vo_ray;
static inline vo_ray ray_make(v3 o, v3 d) { /* ray.h:11-12: ctor normalises */
    vo_ray r; r.o = o; r.d = v3_normalized(d); return r;
}

/* ------------------------------------------------------------------------- */
/* Gaussian -- include/gaussian.h                                             */
/* ------------------------------------------------------------------------- */
VO_EXPORT double vo_erfinv_approx(double x) { /* gaussian.h:10-25 (Winitzki, a = 0.14) */
    if (isnan(x)) return NAN;
    if (x <= -1.0) return -INFINITY;
    if (x >= 1.0) return INFINITY;
    const double a = 0.14;
    double sign = (x < 0.0) ? -1.0 : 1.0;
    double ln_term = log(1.0 - x * x);
    double first = 2.0 / (VO_PI * a) + ln_term / 2.0;
    double inside = first * first - ln_term / a;
    if (inside < 0.0) inside = 0.0;
    double result = sqrt(sqrt(inside) - first);
    return sign * result;
}

typedef struct {
    float mean[3];
    float cov[9];   /* row-major, symmetric */
    float inv[9];   /* Sigma^-1 */
    float norm;     /* (2 pi)^-1.5 det^-0.5 */
    float density, albedo;
} vo_gauss;

static inline float cof3(const float* m, int i, int j) { /* Eigen LU/InverseImpl.h:106-110 */
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return m[3 * i1 + j1] * m[3 * i2 + j2] - m[3 * i1 + j2] * m[3 * i2 + j1];
}
static void gauss_precompute(vo_gauss* g) { /* gaussian.h:52-56 (inverse, determinant, norm) */
    const float* a = g->cov;
    /* Matrix3f::inverse(): cofactors of column 0, det = sum(cof .* col0) = e0 + (e1 + e2),
     * invdet = 1/det, every entry = cofactor * invdet (Eigen LU/InverseImpl.h:112-142) */
    float c0 = cof3(a, 0, 0), c1 = cof3(a, 1, 0), c2 = cof3(a, 2, 0);
    float det_inv = c0 * a[0] + (c1 * a[3] + c2 * a[6]);
    float invdet = 1.0f / det_inv;
    float* m = g->inv;
    m[0] = c0 * invdet; m[1] = c1 * invdet; m[2] = c2 * invdet;     /* row(0) = cofactors_col0 * invdet */
    m[3] = cof3(a, 0, 1) * invdet; m[4] = cof3(a, 1, 1) * invdet; m[5] = cof3(a, 2, 1) * invdet;
    m[6] = cof3(a, 0, 2) * invdet; m[7] = cof3(a, 1, 2) * invdet; m[8] = cof3(a, 2, 2) * invdet;
    /* Matrix3f::determinant(): Eigen LU/Determinant.h:21-24,49 */
    float det = a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
    /* norm = pow(2.0f*pi, -1.5f) [double] * pow(det, -0.5f) [float]  -> float (gaussian.h:55) */
    g->norm = (float)(pow(2.0 * VO_PI, -1.5) * (double)powf(det, -0.5f));
}
static inline v3 mat_mul(const float* m, v3 v) {
    /* coefficient-based 3x3 * 3x1 product: row . vector, summed e0 + (e1 + e2) */
    return v3_make(m[0] * v.x + (m[1] * v.y + m[2] * v.z), m[3] * v.x + (m[4] * v.y + m[5] * v.z),
                   m[6] * v.x + (m[7] * v.y + m[8] * v.z));
}
static inline float gauss_mu_t(const vo_gauss* g, v3 x) { /* gaussian.h:111-117 */
    ++vo_cnt[VO_C_MU];
    v3 d = v3_sub(x, v3_make(g->mean[0], g->mean[1], g->mean[2]));
    /* -0.5f * d^T * inv_cov * d parses as ((-0.5f d^T) * M) * d */
    v3 h = v3_make(-0.5f * d.x, -0.5f * d.y, -0.5f * d.z);
    const float* m = g->inv;
    v3 hm = v3_make(h.x * m[0] + (h.y * m[3] + h.z * m[6]), h.x * m[1] + (h.y * m[4] + h.z * m[7]),
                    h.x * m[2] + (h.y * m[5] + h.z * m[8]));
    float exponent = v3_dot(hm, d);
    return g->density * (g->norm * expf(exponent));
}
static inline int gauss_intersect(const vo_gauss* g, const vo_ray* ray, float* t_enter, float* t_exit) {
    /* gaussian.h:126-164, R = 3 (gaussian.h:36) */
    v3 p = v3_sub(ray->o, v3_make(g->mean[0], g->mean[1], g->mean[2]));
    v3 Md = mat_mul(g->inv, ray->d);
    v3 Mp = mat_mul(g->inv, p);
    float A = v3_dot(ray->d, Md);
    float B = 2.0f * v3_dot(p, Md);
    float C = v3_dot(p, Mp) - (3.0f * 3.0f);
    float disc = B * B - 4.0f * A * C;
    if (disc < 0.0f) return 0;
    float sq = sqrtf(disc);
    float t0 = (-B - sq) / (2.0f * A);
    float t1 = (-B + sq) / (2.0f * A);
    if (t0 > t1) { float t = t0; t0 = t1; t1 = t; }
    if (t1 < 0.0f) return 0;
    *t_enter = (t0 >= 0.0f) ? t0 : 0.0f;
    *t_exit = t1;
    return 1;
}
static inline float gauss_optical_depth(const vo_gauss* g, const vo_ray* ray, float t0, float t1) {
    /* gaussian.h:208-231, with its float/double mix */
    ++vo_cnt[VO_C_OD];
    v3 p = v3_sub(ray->o, v3_make(g->mean[0], g->mean[1], g->mean[2]));
    v3 Md = mat_mul(g->inv, ray->d);
    v3 Mp = mat_mul(g->inv, p);
    float A = v3_dot(ray->d, Md);
    float B = 2.0f * v3_dot(p, Md);
    float C = v3_dot(p, Mp);
    float pref = (float)((double)(g->density * g->norm) * sqrt(VO_PI / (double)(2.0f * A)));
    float den = 2.0f * sqrtf(2.0f * A);
    float F1 = erff((B + 2.0f * A * t1) / den);
    float F0 = erff((B + 2.0f * A * t0) / den);
    return pref * expf(-0.5f * (C - B * B / (4.0f * A))) * (F1 - F0);
}
static inline int gauss_solve_t(const vo_gauss* g, const vo_ray* ray, float t0, float tb, float target_tau,
                                float* t_out) { /* gaussian.h:235-297 (all double) */
    ++vo_cnt[VO_C_SA];
    v3 p = v3_sub(ray->o, v3_make(g->mean[0], g->mean[1], g->mean[2]));
    v3 Md = mat_mul(g->inv, ray->d);
    v3 Mp = mat_mul(g->inv, p);
    double A = (double)v3_dot(ray->d, Md);
    if (!(A > 0.0) || !isfinite(A)) return 0;
    double B = 2.0 * (double)v3_dot(p, Md);
    double C = (double)v3_dot(p, Mp);
    double sqrtA = sqrt(A);
    double pref = (double)g->density * (double)g->norm * sqrt(VO_PI / (2.0 * A));
    double exp_factor = exp(-0.5 * (C - (B * B) / (4.0 * A)));
    double denom = pref * exp_factor;
    if (!(denom > 0.0) || !isfinite(denom)) return 0;
    double two_sqrt2_sqrtA = 2.0 * sqrt(2.0) * sqrtA;
    double erf_t0 = erf((B + 2.0 * A * (double)t0) / two_sqrt2_sqrtA);
    double target_erf = (double)target_tau / denom + erf_t0;
    const double one_eps = 1.0 - 1e-14;
    if (target_erf >= one_eps) { *t_out = tb; return 1; }
    if (target_erf <= -one_eps) { *t_out = t0; return 1; }
    if (!isfinite(target_erf)) return 0;
    if (target_erf <= -1.0 || target_erf >= 1.0) return 0;
    double arg_t = vo_erfinv_approx(target_erf);
    double numer = two_sqrt2_sqrtA * arg_t - B;
    double t_candidate = numer / (2.0 * A);
    if (!isfinite(t_candidate)) return 0;
    if (t_candidate < (double)t0 - 1e-6) t_candidate = t0;
    if (t_candidate > (double)tb + 1e-6) t_candidate = tb;
    *t_out = (float)t_candidate;
    return 1;
}

/* ------------------------------------------------------------------------- */
/* Gaussian mixture + BVH -- include/gmm.h                                    */
/* ------------------------------------------------------------------------- */
typedef struct { float bmin[3], bmax[3]; uint32_t left_first, count; } vo_node; /* gmm.h:37-43 */

typedef struct vo_gmm {
    int n;
    vo_gauss* g;
    float* gmin; /* 3n */
    float* gmax; /* 3n */
    uint32_t* indices;
    vo_node* nodes;
    int n_nodes, cap_nodes;
} vo_gmm;

static void node_bounds(vo_gmm* m, uint32_t ni) { /* gmm.h:262-274 */
    vo_node* nd = &m->nodes[ni];
    for (int k = 0; k < 3; ++k) { nd->bmin[k] = INFINITY; nd->bmax[k] = -INFINITY; }
    for (uint32_t i = 0; i < nd->count; ++i) {
        uint32_t gi = m->indices[nd->left_first + i];
        for (int k = 0; k < 3; ++k) {
            nd->bmin[k] = fminf(nd->bmin[k], m->gmin[3 * gi + k]);
            nd->bmax[k] = fmaxf(nd->bmax[k], m->gmax[3 * gi + k]);
        }
    }
}
static void subdivide(vo_gmm* m, uint32_t ni) { /* gmm.h:318-320, 408-446 (midpoint split, leaf <= 4) */
    if (m->nodes[ni].count <= 4) return;
    vo_node nd = m->nodes[ni];
    float e[3] = {nd.bmax[0] - nd.bmin[0], nd.bmax[1] - nd.bmin[1], nd.bmax[2] - nd.bmin[2]};
    int axis = (e[1] > e[0]) ? 1 : 0;
    if (e[2] > e[axis]) axis = 2;
    float split = 0.5f * (nd.bmin[axis] + nd.bmax[axis]);
    int64_t i = nd.left_first, j = (int64_t)nd.left_first + nd.count - 1;
    while (i <= j) {
        uint32_t gi = m->indices[i];
        if (m->g[gi].mean[axis] < split) ++i;
        else { uint32_t t = m->indices[i]; m->indices[i] = m->indices[j]; m->indices[j] = t; --j; }
    }
    uint32_t left_count = (uint32_t)(i - nd.left_first);
    if (left_count == 0 || left_count == nd.count) return;
    uint32_t lc = (uint32_t)m->n_nodes++, rc = (uint32_t)m->n_nodes++;
    m->nodes[lc].left_first = nd.left_first;
    m->nodes[lc].count = left_count;
    m->nodes[rc].left_first = nd.left_first + left_count;
    m->nodes[rc].count = nd.count - left_count;
    m->nodes[ni].left_first = lc;
    m->nodes[ni].count = 0;
    node_bounds(m, lc);
    node_bounds(m, rc);
    subdivide(m, lc);
    subdivide(m, rc);
}
static void gmm_build_bvh(vo_gmm* m) { /* gmm.h:231-260 */
    int n = m->n;
    for (int i = 0; i < n; ++i) {
        m->indices[i] = (uint32_t)i;
        for (int k = 0; k < 3; ++k) {
            /* tight 3-sigma extent along axis k, padded (deviation noted in the header) */
            float h = 3.0f * sqrtf(fmaxf(m->g[i].cov[4 * k], 0.0f));
            h = h * 1.001f + 1e-6f;
            m->gmin[3 * i + k] = m->g[i].mean[k] - h;
            m->gmax[3 * i + k] = m->g[i].mean[k] + h;
        }
    }
    m->n_nodes = 1;
    m->nodes[0].left_first = 0;
    m->nodes[0].count = (uint32_t)n;
    node_bounds(m, 0);
    subdivide(m, 0);
}
VO_EXPORT vo_gmm* vo_gmm_create(int n, const float* mean3, const float* cov6, const float* density,
                                const float* albedo) {
    /* scene.h:88-113: cov6 = cxx cxy cxz cyy cyz czz */
    vo_gmm* m = (vo_gmm*)calloc(1, sizeof(vo_gmm));
    m->n = n;
    m->g = (vo_gauss*)calloc((size_t)(n > 0 ? n : 1), sizeof(vo_gauss));
    m->gmin = (float*)calloc((size_t)3 * (n > 0 ? n : 1), sizeof(float));
    m->gmax = (float*)calloc((size_t)3 * (n > 0 ? n : 1), sizeof(float));
    m->indices = (uint32_t*)calloc((size_t)(n > 0 ? n : 1), sizeof(uint32_t));
    m->cap_nodes = 2 * (n > 0 ? n : 1) + 2;
    m->nodes = (vo_node*)calloc((size_t)m->cap_nodes, sizeof(vo_node));
    for (int i = 0; i < n; ++i) {
        vo_gauss* g = &m->g[i];
        memcpy(g->mean, mean3 + 3 * i, 12);
        const float* c = cov6 + 6 * i;
        float cov[9] = {c[0], c[1], c[2], c[1], c[3], c[4], c[2], c[4], c[5]};
        memcpy(g->cov, cov, 36);
        g->density = density[i];
        g->albedo = albedo[i];
        gauss_precompute(g);
    }
    if (n > 0) gmm_build_bvh(m);
    return m;
}
VO_EXPORT void vo_gmm_destroy(vo_gmm* m) {
    if (!m) return;
    free(m->g); free(m->gmin); free(m->gmax); free(m->indices); free(m->nodes); free(m);
}
VO_EXPORT int vo_gmm_num_nodes(const vo_gmm* m) { return m->n_nodes; }
/* read back the derived per-Gaussian quantities (inv 9, norm) for KATs */
VO_EXPORT void vo_gmm_get_derived(const vo_gmm* m, float* inv9, float* norm) {
    for (int i = 0; i < m->n; ++i) { memcpy(inv9 + 9 * i, m->g[i].inv, 36); norm[i] = m->g[i].norm; }
}

static float aabb_hit(const vo_ray* r, const float* bmin, const float* bmax) { /* gmm.h:48-63 */
    float rdx = 1.0f / r->d.x, rdy = 1.0f / r->d.y, rdz = 1.0f / r->d.z;
    float tx1 = (bmin[0] - r->o.x) * rdx, tx2 = (bmax[0] - r->o.x) * rdx;
    float tmin = fminf(tx1, tx2), tmax = fmaxf(tx1, tx2);
    float ty1 = (bmin[1] - r->o.y) * rdy, ty2 = (bmax[1] - r->o.y) * rdy;
    tmin = fmaxf(tmin, fminf(ty1, ty2));
    tmax = fminf(tmax, fmaxf(ty1, ty2));
    float tz1 = (bmin[2] - r->o.z) * rdz, tz2 = (bmax[2] - r->o.z) * rdz;
    tmin = fmaxf(tmin, fminf(tz1, tz2));
    tmax = fminf(tmax, fmaxf(tz1, tz2));
    if (tmax >= tmin && tmin < INFINITY && tmax > 0.0f) return tmin;
    return INFINITY;
}

typedef struct { float t; int entering; uint32_t index; uint32_t seq; } vo_event; /* smm.h:7-15 */
typedef struct { vo_event* e; int n, cap; } vo_events;
static void events_push(vo_events* ev, float t, int entering, uint32_t idx) {
    if (ev->n == ev->cap) {
        ev->cap = ev->cap ? 2 * ev->cap : 64;
        ev->e = (vo_event*)realloc(ev->e, (size_t)ev->cap * sizeof(vo_event));
    }
    vo_event x = {t, entering, idx, (uint32_t)ev->n};
    ev->e[ev->n++] = x;
}
static int event_cmp(const void* a, const void* b) {
    const vo_event* x = (const vo_event*)a; const vo_event* y = (const vo_event*)b;
    if (x->t < y->t) return -1;
    if (x->t > y->t) return 1;
    return (x->seq < y->seq) ? -1 : (x->seq > y->seq);
}
typedef struct { int* s; int cap; } vo_stack;

static void gmm_intersect_events(const vo_gmm* m, const vo_ray* ray, vo_events* ev, vo_stack* st) {
    /* gmm.h:457-515 */
    ev->n = 0;
    ++vo_cnt[VO_C_V];
    if (m->n == 0) return;
    if (st->cap < m->n_nodes + 2) { st->cap = m->n_nodes + 2; st->s = (int*)realloc(st->s, (size_t)st->cap * sizeof(int)); }
    int sp = 0;
    st->s[sp++] = 0;
    while (sp > 0) {
        int ni = st->s[--sp];
        const vo_node* nd = &m->nodes[ni];
        float tmin = aabb_hit(ray, nd->bmin, nd->bmax);
        if (tmin == INFINITY) continue;
        if (nd->count > 0) {
            for (uint32_t ii = 0; ii < nd->count; ++ii) {
                uint32_t gi = m->indices[nd->left_first + ii];
                float t0, t1;
                if (gauss_intersect(&m->g[gi], ray, &t0, &t1)) {
                    if (t0 >= 0.0f) events_push(ev, t0, 1, gi);
                    if (t1 >= 0.0f) events_push(ev, t1, 0, gi);
                    /* a grazed component: the FP32 discriminant is exactly 0 and t_enter == t_exit.  The reference sorts its events
                     * with std::sort on t alone (gmm.h:511-513), which is not stable: it may put this component's EXIT before its
                     * ENTER, and the walk (integrator.h:466-486) then never removes the component from the active list again.
                     * Which way the tie falls depends on the order of the whole event array in the reference's own traversal;
                     * this restatement (enter before exit) and the device cannot know it, so the sample is marked: margin 0. */
                    if (t0 >= 0.0f && t0 == t1) vo_min_margin = 0.0f;
                }
            }
        } else {
            int l = (int)nd->left_first, r = l + 1;
            float dl = aabb_hit(ray, m->nodes[l].bmin, m->nodes[l].bmax);
            float dr = aabb_hit(ray, m->nodes[r].bmin, m->nodes[r].bmax);
            if (dl > dr) {
                if (dl != INFINITY) st->s[sp++] = l;
                if (dr != INFINITY) st->s[sp++] = r;
            } else {
                if (dr != INFINITY) st->s[sp++] = r;
                if (dl != INFINITY) st->s[sp++] = l;
            }
        }
    }
    vo_cnt[VO_C_H] += (uint64_t)ev->n / 2;
    if (ev->n > 1) qsort(ev->e, (size_t)ev->n, sizeof(vo_event), event_cmp);
}

static float gmm_transmittance_up_to(const vo_gmm* m, const vo_ray* ray, float tmax, vo_stack* st) {
    /* gmm.h:517-578: unsorted sum of clipped optical depths, double accumulator */
    ++vo_cnt[VO_C_SHADOW];
    if (tmax <= 0.0f) return 1.0f;
    if (m->n == 0) return 1.0f;
    double sum = 0.0;
    if (st->cap < m->n_nodes + 2) { st->cap = m->n_nodes + 2; st->s = (int*)realloc(st->s, (size_t)st->cap * sizeof(int)); }
    int sp = 0;
    st->s[sp++] = 0;
    while (sp > 0) {
        int ni = st->s[--sp];
        const vo_node* nd = &m->nodes[ni];
        float tmin = aabb_hit(ray, nd->bmin, nd->bmax);
        if (tmin == INFINITY || tmin > tmax) continue;
        if (nd->count > 0) {
            for (uint32_t ii = 0; ii < nd->count; ++ii) {
                uint32_t gi = m->indices[nd->left_first + ii];
                float t0, t1;
                if (!gauss_intersect(&m->g[gi], ray, &t0, &t1)) continue;
                float a = fmaxf(0.0f, t0), b = fminf(tmax, t1);
                if (b > a) sum += (double)gauss_optical_depth(&m->g[gi], ray, a, b);
            }
        } else {
            int l = (int)nd->left_first, r = l + 1;
            float dl = aabb_hit(ray, m->nodes[l].bmin, m->nodes[l].bmax);
            float dr = aabb_hit(ray, m->nodes[r].bmin, m->nodes[r].bmax);
            if (dl == INFINITY || dl > tmax) dl = INFINITY;
            if (dr == INFINITY || dr > tmax) dr = INFINITY;
            if (dl > dr) {
                if (dl != INFINITY) st->s[sp++] = l;
                if (dr != INFINITY) st->s[sp++] = r;
            } else {
                if (dr != INFINITY) st->s[sp++] = r;
                if (dl != INFINITY) st->s[sp++] = l;
            }
        }
    }
    return expf(-(float)sum);
}

static float gmm_evaluate_albedo(const vo_gmm* m, const uint32_t* active, int na, v3 pos) { /* gmm.h:128-143 */
    float sum_mu = 0.0f, sum_mu_alb = 0.0f;
    for (int i = 0; i < na; ++i) {
        float mu = gauss_mu_t(&m->g[active[i]], pos);
        sum_mu += mu;
        sum_mu_alb += mu * m->g[active[i]].albedo;
    }
    float a = sum_mu_alb / sum_mu;
    /* std::clamp(a, 0, 1): NaN stays NaN */
    if (a < 0.0f) return 0.0f;
    if (1.0f < a) return 1.0f;
    return a;
}

/* ------------------------------------------------------------------------- */
/* Distance solvers -- include/distance_solvers.h                             */
/* ------------------------------------------------------------------------- */
static inline float clampf(float v, float lo, float hi) { return (v < lo) ? lo : ((hi < v) ? hi : v); }

static float solve_bisection(const vo_gmm* m, const vo_ray* ray, float ta, float tb, const uint32_t* act, int na,
                             float target) { /* distance_solvers.h:25-57 (15 iters, tol 1e-6) */
    ++vo_cnt[VO_C_BISECT];
    float a = ta, b = tb;
    for (int i = 0; i < 15; ++i) {
        float mid = 0.5f * (a + b);
        float tau = 0.0f;
        for (int k = 0; k < na; ++k) tau += gauss_optical_depth(&m->g[act[k]], ray, ta, mid);
        float f = tau - target;
        vo_margin(fabsf(f), 1e-6f); vo_margin(tau, target);
        if (fabsf(f) <= 1e-6f) return mid;
        if (f < 0.0f) a = mid; else b = mid;
    }
    return 0.5f * (a + b);
}
static float newton_f(const vo_gmm* m, const vo_ray* ray, float ta, float b, const uint32_t* act, int na, float tt,
                      float target) { /* distance_solvers.h:75-82 */
    float sum = 0.0f;
    float ttc = fminf(tt, b);
    for (int k = 0; k < na; ++k) sum += gauss_optical_depth(&m->g[act[k]], ray, ta, ttc);
    return sum - target;
}
static float solve_newton(const vo_gmm* m, const vo_ray* ray, float ta, float tb, const uint32_t* act, int na,
                          float target) { /* distance_solvers.h:62-127 (8 iters, tol 1e-6) */
    ++vo_cnt[VO_C_NEWTON];
    const float tol = 1e-6f;
    float a = ta, b = tb;
    float t = 0.5f * (a + b);
    for (int iter = 0; iter < 8; ++iter) {
        float f = newton_f(m, ray, ta, b, act, na, t, target);
        vo_margin(fabsf(f), tol);
        if (fabsf(f) <= tol) return clampf(t, a, b);
        float h = fmaxf(1e-5f, (b - a) * 1e-6f);
        float tp = fminf(b, t + h);
        float fp = newton_f(m, ray, ta, b, act, na, tp, target);
        float deriv = (fp - f) / (tp - t);
        if (!(deriv > 0.0f) || !isfinite(deriv) || fabsf(deriv) < 1e-12f)
            return solve_bisection(m, ray, ta, tb, act, na, target);
        float t_next = t - f / deriv;
        if (!isfinite(t_next) || t_next < a || t_next > b) return solve_bisection(m, ray, ta, tb, act, na, target);
        vo_margin(fabsf(t_next - t), tol * fmaxf(1.0f, fabsf(t)));
        if (fabsf(t_next - t) <= tol * fmaxf(1.0f, fabsf(t))) { t = t_next; return clampf(t, a, b); }
        t = t_next;
    }
    return solve_bisection(m, ray, ta, tb, act, na, target);
}
/* solver selector (runtime form of the #define block, distance_solvers.h:143-147) */
enum { VO_SOLVER_ANALYTIC_NEWTON = 0, VO_SOLVER_ANALYTIC_BISECTION = 1, VO_SOLVER_NEWTON = 2, VO_SOLVER_BISECTION = 3,
       VO_SOLVER_UNIFORM = 4 };
/* distance_solvers.h:131-137: t = ta + u (tb - ta).  The reference draws u from a random_device-seeded mt19937
 * (rng.h:6-10), i.e. it is not reproducible: PARITY UNPINNED for this mode.  Here u is a counter-based value hashed from
 * the state of the path's PCG stream right after the target draw (an integer, identical on every implementation; one
 * solve per draw), so oracle and device agree and the path's own stream is left untouched. */
static float solve_uniform(float ta, float tb, uint64_t key) {
    const uint64_t h = vo_splitmix64(key);
    const float u = (float)(uint32_t)(h >> 40) * (1.0f / 16777216.0f);
    return ta + u * (tb - ta);
}
static float solve_distance(const vo_gmm* m, const vo_ray* ray, float ta, float tb, const uint32_t* act, int na,
                            float remaining, int solver, uint64_t key) { /* distance_solvers.h:150-187 */
    if (solver == VO_SOLVER_UNIFORM) return solve_uniform(ta, tb, key);
    if ((solver == VO_SOLVER_ANALYTIC_NEWTON || solver == VO_SOLVER_ANALYTIC_BISECTION) && na == 1) {
        float t_an = 0.0f;
        if (gauss_solve_t(&m->g[act[0]], ray, ta, tb, remaining, &t_an)) return clampf(t_an, ta, tb);
    }
    if (solver == VO_SOLVER_ANALYTIC_NEWTON || solver == VO_SOLVER_NEWTON)
        return solve_newton(m, ray, ta, tb, act, na, remaining);
    return solve_bisection(m, ray, ta, tb, act, na, remaining);
}

/* ------------------------------------------------------------------------- */
/* Camera -- include/camera.h                                                 */
/* ------------------------------------------------------------------------- */
typedef struct { int kind; float pos[3], view[3], right[3], up[3], pinhole[3]; } vo_camera; /* kind 0 pinhole, 1 ortho */

/* camera.h:15-22 + 38-43.  NOTE the reference's constructor bodies use the
 * constructor PARAMETER view_dir (it shadows the member), i.e. the caller's
 * vector as passed, while the member view_dir is the normalised copy. */
VO_EXPORT void vo_camera_make(int kind, const float* position, const float* view_dir_arg, float fov, vo_camera* out) {
    v3 pos = v3_make(position[0], position[1], position[2]);
    v3 varg = v3_make(view_dir_arg[0], view_dir_arg[1], view_dir_arg[2]);
    v3 view = v3_normalized(varg);
    v3 right = v3_normalized(v3_cross(varg, v3_make(0.0f, 1.0f, 0.0f)));
    v3 up = v3_normalized(v3_cross(right, varg));
    out->kind = kind;
    out->pos[0] = pos.x; out->pos[1] = pos.y; out->pos[2] = pos.z;
    out->view[0] = view.x; out->view[1] = view.y; out->view[2] = view.z;
    out->right[0] = right.x; out->right[1] = right.y; out->right[2] = right.z;
    out->up[0] = up.x; out->up[1] = up.y; out->up[2] = up.z;
    float focal = 1.0f / tanf(0.5f * fov);
    v3 ph = v3_add(pos, v3_scale(focal, varg));
    out->pinhole[0] = ph.x; out->pinhole[1] = ph.y; out->pinhole[2] = ph.z;
}
static vo_ray camera_sample_ray(const vo_camera* c, float uu, float vv) {
    v3 pos = v3_make(c->pos[0], c->pos[1], c->pos[2]);
    v3 right = v3_make(c->right[0], c->right[1], c->right[2]);
    v3 up = v3_make(c->up[0], c->up[1], c->up[2]);
    if (c->kind == 0) { /* camera.h:45-53 */
        float u = 1.0f - uu * 2.0f;
        float v = vv * 2.0f - 1.0f;
        v3 o = v3_add(v3_add(pos, v3_scale(u, right)), v3_scale(v, up));
        v3 d = v3_sub(v3_make(c->pinhole[0], c->pinhole[1], c->pinhole[2]), o);
        return ray_make(o, v3_normalized(d)); /* normalised twice: camera.h:52 + ray.h:12 */
    } else { /* camera.h:64-73 */
        float u = uu * 2.0f - 1.0f;
        float v = 1.0f - vv * 2.0f;
        v3 o = v3_add(v3_add(pos, v3_scale(u, right)), v3_scale(v, up));
        v3 d = v3_make(c->view[0], c->view[1], c->view[2]);
        return ray_make(o, v3_normalized(d));
    }
}
VO_EXPORT void vo_camera_sample_ray(const vo_camera* c, float u, float v, float* o3, float* d3) {
    vo_ray r = camera_sample_ray(c, u, v);
    o3[0] = r.o.x; o3[1] = r.o.y; o3[2] = r.o.z; d3[0] = r.d.x; d3[1] = r.d.y; d3[2] = r.d.z;
}

/* ------------------------------------------------------------------------- */
/* KAT entry points (single-primitive calls)                                  */
/* ------------------------------------------------------------------------- */
VO_EXPORT int vo_intersect_direct(const vo_gmm* m, int gi, const float* o, const float* d, float* t0, float* t1) {
    vo_ray r; r.o = v3_make(o[0], o[1], o[2]); r.d = v3_make(d[0], d[1], d[2]);
    return gauss_intersect(&m->g[gi], &r, t0, t1);
}
VO_EXPORT float vo_optical_depth(const vo_gmm* m, int gi, const float* o, const float* d, float t0, float t1) {
    vo_ray r; r.o = v3_make(o[0], o[1], o[2]); r.d = v3_make(d[0], d[1], d[2]);
    return gauss_optical_depth(&m->g[gi], &r, t0, t1);
}
VO_EXPORT float vo_mu_t(const vo_gmm* m, int gi, const float* x) {
    return gauss_mu_t(&m->g[gi], v3_make(x[0], x[1], x[2]));
}
VO_EXPORT int vo_solve_for_t_given_tau(const vo_gmm* m, int gi, const float* o, const float* d, float t0, float tb,
                                       float tau, float* t_out) {
    vo_ray r; r.o = v3_make(o[0], o[1], o[2]); r.d = v3_make(d[0], d[1], d[2]);
    return gauss_solve_t(&m->g[gi], &r, t0, tb, tau, t_out);
}
/* batch transmittance_up_to: rays are used as given (direction must be unit) */
VO_EXPORT void vo_transmittance_batch(const vo_gmm* m, int nrays, const float* o3, const float* d3, const float* tmax,
                                      float* out) {
#pragma omp parallel
    {
        vo_stack st = {0, 0};
#pragma omp for schedule(dynamic, 64)
        for (int i = 0; i < nrays; ++i) {
            vo_ray r; r.o = v3_make(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2]);
            r.d = v3_make(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2]);
            out[i] = gmm_transmittance_up_to(m, &r, tmax[i], &st);
        }
        free(st.s);
    }
}
/* sorted event list of one ray; returns the number of events (<= cap written) */
VO_EXPORT int vo_intersect_events(const vo_gmm* m, const float* o, const float* d, int cap, float* t, int* entering,
                                  uint32_t* index) {
    vo_ray r; r.o = v3_make(o[0], o[1], o[2]); r.d = v3_make(d[0], d[1], d[2]);
    vo_events ev = {0, 0, 0}; vo_stack st = {0, 0};
    gmm_intersect_events(m, &r, &ev, &st);
    for (int i = 0; i < ev.n && i < cap; ++i) { t[i] = ev.e[i].t; entering[i] = ev.e[i].entering; index[i] = ev.e[i].index; }
    int n = ev.n;
    free(ev.e); free(st.s);
    return n;
}

/* transmittance_up_to (gmm.h:517-578) WITHOUT the hierarchy: every component is asked (the loop the reference compiles with
 * USE_BVH off, gmm.h:190-226); what a culling structure must reproduce */
VO_EXPORT void vo_transmittance_brute(const vo_gmm* m, int nrays, const float* o3, const float* d3, const float* tmax, float* out) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int r = 0; r < nrays; ++r) {
        vo_ray ray;
        ray.o = v3_make(o3[3 * r], o3[3 * r + 1], o3[3 * r + 2]);
        ray.d = v3_make(d3[3 * r], d3[3 * r + 1], d3[3 * r + 2]);
        double sum = 0.0;
        if (tmax[r] > 0.0f)
            for (int g = 0; g < m->n; ++g) {
                float t0, t1;
                if (!gauss_intersect(&m->g[g], &ray, &t0, &t1)) continue;
                float a = fmaxf(0.0f, t0), b = fminf(tmax[r], t1);
                if (b > a) sum += (double)gauss_optical_depth(&m->g[g], &ray, a, b);
            }
        out[r] = tmax[r] > 0.0f ? expf(-(float)sum) : 1.0f;
    }
}

/* every component the FP32 quadric test (gaussian.h:126-164) accepts for a batch of rays, WITHOUT any hierarchy: the set a
 * culling structure must not lose members of.  idx[r * cap + k] = component indices (ascending), counts[r] = how many. */
VO_EXPORT void vo_intersect_all(const vo_gmm* m, int nrays, const float* o3, const float* d3, int cap, int* counts, uint32_t* idx) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int r = 0; r < nrays; ++r) {
        vo_ray ray;
        ray.o = v3_make(o3[3 * r], o3[3 * r + 1], o3[3 * r + 2]);
        ray.d = v3_make(d3[3 * r], d3[3 * r + 1], d3[3 * r + 2]);
        int c = 0;
        for (int g = 0; g < m->n; ++g) {
            float t0, t1;
            if (gauss_intersect(&m->g[g], &ray, &t0, &t1)) { if (c < cap) idx[(size_t)r * cap + c] = (uint32_t)g; ++c; }
        }
        counts[r] = c;
    }
}

/* ------------------------------------------------------------------------- */
/* Forward integrators -- include/integrator.h                                */
/* ------------------------------------------------------------------------- */
typedef struct { int nl; const float* pos; const float* intensity; float env[3]; } vo_lights;

static v3 sample_uniform_direction(vo_pcg* rng) { /* integrator.h:32-44 */
    float xi1 = pcg_uniform(rng);
    float xi2 = pcg_uniform(rng);
    float theta = (float)(2.0f * VO_PI * xi1); /* 2.0f * pi(double) * float -> double -> float */
    float phi = acosf(1.0f - 2.0f * xi2);
    float x = sinf(phi) * cosf(theta);
    float y = sinf(phi) * sinf(theta);
    float z = cosf(phi);
    return v3_make(x, y, z);
}

/* per-thread scratch (the reference's thread_local vectors, integrator.h:578-586) */
typedef struct {
    vo_events ev;
    vo_stack st;
    uint32_t* active; int na, cap_active;
    int* idx_pos;    /* N */
    int* idx_epoch;  /* N */
    int epoch_counter;
    unsigned char* active_mask; /* N, single-scatter path */
    uint32_t* pix_list; int n_pix, cap_pix; /* recording */
    int last_bounce;                        /* value of `bounce` (integrator.h:590) when the last sample's path ended */
} vo_scratch;

static void active_push(vo_scratch* s, uint32_t g) {
    if (s->na == s->cap_active) {
        s->cap_active = s->cap_active ? 2 * s->cap_active : 64;
        s->active = (uint32_t*)realloc(s->active, (size_t)s->cap_active * sizeof(uint32_t));
    }
    s->active[s->na++] = g;
}
static void pix_list_add(vo_scratch* s, uint32_t g) { /* linear-search dedupe, integrator.h:625-629 */
    for (int i = 0; i < s->n_pix; ++i) if (s->pix_list[i] == g) return;
    if (s->n_pix == s->cap_pix) {
        s->cap_pix = s->cap_pix ? 2 * s->cap_pix : 64;
        s->pix_list = (uint32_t*)realloc(s->pix_list, (size_t)s->cap_pix * sizeof(uint32_t));
    }
    s->pix_list[s->n_pix++] = g;
}

/* MultiScatterGaussians::get_free_flight_distance, integrator.h:422-498 */
static float ms_free_flight(const vo_gmm* m, const vo_ray* ray, vo_scratch* s, float target_tau, int solver, uint64_t key) {
    int cur_epoch = ++s->epoch_counter;
    if (cur_epoch == 0) { memset(s->idx_epoch, 0, (size_t)m->n * sizeof(int)); cur_epoch = ++s->epoch_counter; }
    double acc_tau = 0.0;
    float t_prev = 0.0f;
    for (int ei = 0; ei < s->ev.n; ++ei) {
        float t_evt = s->ev.e[ei].t;
        double seg_tau = 0.0;
        for (int k = 0; k < s->na; ++k) seg_tau += (double)gauss_optical_depth(&m->g[s->active[k]], ray, t_prev, t_evt);
        vo_margin((float)(acc_tau + seg_tau), target_tau);
        if (acc_tau + seg_tau > (double)target_tau) {
            float remaining = (float)((double)target_tau - acc_tau);
            return solve_distance(m, ray, t_prev, t_evt, s->active, s->na, remaining, solver, key);
        }
        acc_tau += seg_tau;
        uint32_t gidx = s->ev.e[ei].index;
        if (s->ev.e[ei].entering) {
            if (s->idx_epoch[gidx] != cur_epoch) {
                s->idx_epoch[gidx] = cur_epoch;
                s->idx_pos[gidx] = s->na;
                active_push(s, gidx);
            }
        } else if (s->idx_epoch[gidx] == cur_epoch) {
            int pos = s->idx_pos[gidx];
            if (pos >= 0 && pos < s->na) {
                uint32_t last = s->active[s->na - 1];
                s->active[pos] = last;
                s->idx_pos[last] = pos;
                s->na--;
            }
            s->idx_epoch[gidx] = 0;
        }
        t_prev = t_evt;
    }
    return -1.0f;
}

/* shared NEE step: integrator.h:656-678 (MS) == :369-392 (SS).  Returns Li. */
static void nee_sample(const vo_gmm* m, const vo_lights* L, vo_pcg* rng, v3 pos, vo_stack* st, float Li[3]) {
    int is_env = (pcg_uniform(rng) < 1.0f / (float)(L->nl + 1));
    if (!is_env) {
        int li = (int)(pcg_uniform(rng) * (float)L->nl);
        v3 lp = v3_make(L->pos[3 * li], L->pos[3 * li + 1], L->pos[3 * li + 2]);
        v3 to = v3_sub(lp, pos);
        v3 wi = v3_normalized(to);
        float dist = sqrtf(v3_sqnorm(to));
        vo_ray shadow = ray_make(pos, wi);
        float Tr = gmm_transmittance_up_to(m, &shadow, dist, st);
        for (int c = 0; c < 3; ++c) Li[c] = (Tr * L->intensity[3 * li + c]) / (dist * dist);
    } else {
        v3 wi = sample_uniform_direction(rng);
        vo_ray eray = ray_make(pos, wi);
        float Tr = gmm_transmittance_up_to(m, &eray, INFINITY, st);
        float four_pi = (float)(4.0f * VO_PI);
        for (int c = 0; c < 3; ++c) Li[c] = (Tr * L->env[c]) * four_pi;
    }
}

static void scratch_init(vo_scratch* s, int n) {
    memset(s, 0, sizeof(*s));
    s->idx_pos = (int*)calloc((size_t)(n > 0 ? n : 1), sizeof(int));
    s->idx_epoch = (int*)calloc((size_t)(n > 0 ? n : 1), sizeof(int));
    s->active_mask = (unsigned char*)calloc((size_t)(n > 0 ? n : 1), 1);
}
static void scratch_free(vo_scratch* s) {
    free(s->ev.e); free(s->st.s); free(s->active); free(s->idx_pos); free(s->idx_epoch); free(s->active_mask);
    free(s->pix_list);
}

/* one multi-scatter sample: integrator.h:557-703.  record != 0 adds to s->pix_list (integrator.h:616-644) */
static void ms_sample(const vo_gmm* m, const vo_lights* L, const vo_camera* cam, int x, int y, int si, int spp, int W,
                      int H, int min_scatter, int solver, int record, vo_scratch* s, float out[3]) {
    vo_pcg rng;
    pcg_seed(&rng, vo_derive_path_seed(x, y, si), 1);
    int n = (int)sqrt((double)spp);
    int sx = si % n, sy = si / n;
    float u = ((float)x + ((float)sx + pcg_uniform(&rng)) / (float)n) / (float)W;
    float v = ((float)y + ((float)sy + pcg_uniform(&rng)) / (float)n) / (float)H;
    vo_ray ray = camera_sample_ray(cam, u, v);
    float thr[3] = {1.0f, 1.0f, 1.0f};
    float Lacc[3] = {0.0f, 0.0f, 0.0f};
    const float phase_pdf = (float)(1.0f / (4.0f * VO_PI));
    const float w_ne = (float)(L->nl + 1);
    for (int bounce = 0;; ++bounce) {
        s->last_bounce = bounce;
        gmm_intersect_events(m, &ray, &s->ev, &s->st);
        if (s->ev.n == 0) { for (int c = 0; c < 3; ++c) Lacc[c] += thr[c] * L->env[c]; break; }
        float target_tau = -logf(1.0f - pcg_uniform(&rng));
        if (vo_perturb == 3) target_tau *= 1.0f + 0x1p-21f; else if (vo_perturb == 4) target_tau *= 1.0f - 0x1p-21f;
        s->na = 0;
        float t_scatter = ms_free_flight(m, &ray, s, target_tau, solver, rng.state);
        if (t_scatter >= 0.0f && vo_perturb == 1) t_scatter = nextafterf(t_scatter, INFINITY);
        else if (t_scatter > 0.0f && vo_perturb == 2) t_scatter = nextafterf(t_scatter, 0.0f);
        else if (t_scatter > 0.0f && vo_perturb == 5) t_scatter *= 1.0f + 0x1p-20f;
        else if (t_scatter > 0.0f && vo_perturb == 6) t_scatter *= 1.0f - 0x1p-20f;
        if (record) {
            const float tol = 1e-6f;
            if (t_scatter >= 0.0f) {
                for (int i = 0; i < s->ev.n; ++i) {
                    if (s->ev.e[i].t <= t_scatter + tol) pix_list_add(s, s->ev.e[i].index); else break;
                }
            } else {
                for (int i = 0; i < s->ev.n; ++i) { if (s->ev.e[i].t < 0.0f) continue; pix_list_add(s, s->ev.e[i].index); }
            }
        }
        if (t_scatter < 0.0f) { for (int c = 0; c < 3; ++c) Lacc[c] += thr[c] * L->env[c]; break; }
        v3 pos = v3_add(ray.o, v3_scale(t_scatter, ray.d));
        float albedo = gmm_evaluate_albedo(m, s->active, s->na, pos);
        float Li[3];
        nee_sample(m, L, &rng, pos, &s->st, Li);
        for (int c = 0; c < 3; ++c) Lacc[c] += (thr[c] * (albedo * phase_pdf * w_ne)) * Li[c];
        for (int c = 0; c < 3; ++c) thr[c] *= albedo;
        if (bounce >= min_scatter) {
            float mx = fmaxf(thr[0], fmaxf(thr[1], thr[2]));
            float rr = fminf(mx, 0.9f);
            float ur = pcg_uniform(&rng);
            vo_margin(ur, rr);
            if (ur > rr) { s->last_bounce = bounce; break; }
            for (int c = 0; c < 3; ++c) thr[c] /= rr;
        }
        v3 nd = sample_uniform_direction(&rng);
        ray = ray_make(pos, nd);
    }
    out[0] = Lacc[0]; out[1] = Lacc[1]; out[2] = Lacc[2];
}

/* one single-scatter sample: integrator.h:307-401.  returns the value ADDED to pixel_L */
static void ss_sample(const vo_gmm* m, const vo_lights* L, const vo_camera* cam, int x, int y, int si, int spp, int W,
                      int H, int solver, vo_scratch* s, float out[3]) {
    vo_pcg rng;
    pcg_seed(&rng, vo_derive_path_seed(x, y, si), 1);
    int n = (int)sqrt((double)spp);
    int sx = si % n, sy = si / n;
    float u = ((float)x + ((float)sx + pcg_uniform(&rng)) / (float)n) / (float)W;
    float v = ((float)y + ((float)sy + pcg_uniform(&rng)) / (float)n) / (float)H;
    vo_ray ray = camera_sample_ray(cam, u, v);
    gmm_intersect_events(m, &ray, &s->ev, &s->st);
    if (s->ev.n == 0) { for (int c = 0; c < 3; ++c) out[c] = L->env[c]; return; }
    float target_tau = -logf(1.0f - pcg_uniform(&rng));
    if (vo_perturb == 3) target_tau *= 1.0f + 0x1p-21f; else if (vo_perturb == 4) target_tau *= 1.0f - 0x1p-21f;
    s->last_bounce = 0;
    /* active is a bool vector rescanned in ascending index order at every event (integrator.h:335,344-346) */
    float acc_tau = 0.0f, t_prev = 0.0f, t_scatter = -1.0f;
    int touched_n = 0;
    for (int ei = 0; ei < s->ev.n; ++ei) {
        float t_evt = s->ev.e[ei].t;
        s->na = 0;
        for (int i = 0; i < m->n; ++i) if (s->active_mask[i]) active_push(s, (uint32_t)i);
        float seg_tau = 0.0f;
        for (int k = 0; k < s->na; ++k) seg_tau += gauss_optical_depth(&m->g[s->active[k]], &ray, t_prev, t_evt);
        vo_margin(acc_tau + seg_tau, target_tau);
        if (acc_tau + seg_tau > target_tau) {
            t_scatter = solve_distance(m, &ray, t_prev, t_evt, s->active, s->na, target_tau - acc_tau, solver, rng.state);
            if (t_scatter >= 0.0f && vo_perturb == 1) t_scatter = nextafterf(t_scatter, INFINITY);
            else if (t_scatter > 0.0f && vo_perturb == 2) t_scatter = nextafterf(t_scatter, 0.0f);
            else if (t_scatter > 0.0f && vo_perturb == 5) t_scatter *= 1.0f + 0x1p-20f;
            else if (t_scatter > 0.0f && vo_perturb == 6) t_scatter *= 1.0f - 0x1p-20f;
            break;
        }
        acc_tau += seg_tau;
        s->active_mask[s->ev.e[ei].index] = (unsigned char)s->ev.e[ei].entering;
        t_prev = t_evt;
        ++touched_n;
    }
    for (int i = 0; i < s->ev.n; ++i) s->active_mask[s->ev.e[i].index] = 0; /* reset scratch */
    (void)touched_n;
    if (t_scatter < 0.0f) { for (int c = 0; c < 3; ++c) out[c] = L->env[c]; return; }
    v3 pos = v3_add(ray.o, v3_scale(t_scatter, ray.d));
    float albedo = gmm_evaluate_albedo(m, s->active, s->na, pos);
    float Li[3];
    nee_sample(m, L, &rng, pos, &s->st, Li);
    const float phase_pdf = (float)(1.0f / (4.0f * VO_PI));
    const float w_light = (float)(L->nl + 1);
    for (int c = 0; c < 3; ++c) out[c] = (albedo * phase_pdf * w_light) * Li[c];
}

enum { VO_INTEGRATOR_SS = 0, VO_INTEGRATOR_MS = 1 };

/* Integrator::render for the two Gaussian free-flight integrators.
 * rec_counts/rec_idx (optional, MS only): per-pixel recorded Gaussian lists, rec_idx holds
 * rec_cap entries per pixel (extra entries are dropped and the count still reports the true size).
 * (x0,y0,x1,y1) limits the pixels rendered (others are left untouched) so that bounded CPU
 * samples of a big image can be timed; pass 0,0,W,H for the full image. */
VO_EXPORT void vo_render(const vo_gmm* m, int nl, const float* lpos, const float* lint, const float* env,
                         const vo_camera* cam, int integrator, int solver, int W, int H, int spp, int min_bounces,
                         int x0, int y0, int x1, int y1, float* out_rgb, int* rec_counts, uint32_t* rec_idx, int rec_cap) {
    vo_lights L; L.nl = nl; L.pos = lpos; L.intensity = lint; L.env[0] = env[0]; L.env[1] = env[1]; L.env[2] = env[2];
#pragma omp parallel
    {
        vo_scratch s;
        scratch_init(&s, m->n);
#pragma omp for collapse(2) schedule(dynamic, 1)
        for (int y = y0; y < y1; ++y) {
            for (int x = x0; x < x1; ++x) {
                float pix[3] = {0.0f, 0.0f, 0.0f};
                s.n_pix = 0;
                vo_cnt[VO_C_PATHS] += (uint64_t)spp;
                for (int si = 0; si < spp; ++si) {
                    float Ls[3];
                    if (integrator == VO_INTEGRATOR_SS) ss_sample(m, &L, cam, x, y, si, spp, W, H, solver, &s, Ls);
                    else ms_sample(m, &L, cam, x, y, si, spp, W, H, min_bounces, solver, rec_counts != 0, &s, Ls);
                    for (int c = 0; c < 3; ++c) pix[c] += Ls[c];
                }
                size_t p = (size_t)y * W + x;
                for (int c = 0; c < 3; ++c) out_rgb[3 * p + c] = pix[c] / (float)spp; /* integrator.h:706 */
                if (rec_counts) {
                    rec_counts[p] = s.n_pix;
                    for (int i = 0; i < s.n_pix && i < rec_cap; ++i) rec_idx[p * (size_t)rec_cap + i] = s.pix_list[i];
                }
            }
        }
        scratch_free(&s);
        vo_counters_fold();
    }
}
/* per-sample radiance of one pixel (debug aid for single-sample flip attribution) */
VO_EXPORT void vo_pixel_samples(const vo_gmm* m, int nl, const float* lpos, const float* lint, const float* env,
                                const vo_camera* cam, int integrator, int solver, int W, int H, int spp, int min_bounces,
                                int x, int y, float* out_rgb_per_sample) {
    vo_lights L; L.nl = nl; L.pos = lpos; L.intensity = lint; L.env[0] = env[0]; L.env[1] = env[1]; L.env[2] = env[2];
    vo_scratch s;
    scratch_init(&s, m->n);
    for (int si = 0; si < spp; ++si) {
        if (integrator == VO_INTEGRATOR_SS) ss_sample(m, &L, cam, x, y, si, spp, W, H, solver, &s, out_rgb_per_sample + 3 * si);
        else ms_sample(m, &L, cam, x, y, si, spp, W, H, min_bounces, solver, 0, &s, out_rgb_per_sample + 3 * si);
    }
    scratch_free(&s);
}
/* per-sample output of a pixel window: out4[((y - y0) * (x1 - x0) + (x - x0)) * spp + si] = (r, g, b, bounce index at the
 * end of the path); margin (optional, same indexing, one float per sample) = smallest relative margin of the sample's
 * discrete decisions; perturb: see vo_perturb above.  Summed in si order and divided by spp the samples of a pixel give
 * vo_render's pixel bit for bit. */
VO_EXPORT void vo_render_samples(const vo_gmm* m, int nl, const float* lpos, const float* lint, const float* env,
                                 const vo_camera* cam, int integrator, int solver, int W, int H, int spp, int min_bounces,
                                 int x0, int y0, int x1, int y1, int perturb, float* out4, float* margin) {
    vo_lights L; L.nl = nl; L.pos = lpos; L.intensity = lint; L.env[0] = env[0]; L.env[1] = env[1]; L.env[2] = env[2];
#pragma omp parallel
    {
        vo_scratch s;
        scratch_init(&s, m->n);
        vo_perturb = perturb;
#pragma omp for collapse(2) schedule(dynamic, 1)
        for (int y = y0; y < y1; ++y) {
            for (int x = x0; x < x1; ++x) {
                for (int si = 0; si < spp; ++si) {
                    const size_t k = ((size_t)(y - y0) * (size_t)(x1 - x0) + (size_t)(x - x0)) * (size_t)spp + (size_t)si;
                    float Ls[3];
                    vo_min_margin = INFINITY;
                    if (integrator == VO_INTEGRATOR_SS) ss_sample(m, &L, cam, x, y, si, spp, W, H, solver, &s, Ls);
                    else ms_sample(m, &L, cam, x, y, si, spp, W, H, min_bounces, solver, 0, &s, Ls);
                    out4[4 * k] = Ls[0]; out4[4 * k + 1] = Ls[1]; out4[4 * k + 2] = Ls[2]; out4[4 * k + 3] = (float)s.last_bounce;
                    if (margin) margin[k] = vo_min_margin;
                }
            }
        }
        vo_perturb = 0;
        scratch_free(&s);
        vo_counters_fold();
    }
}
/* sinf / cosf / acosf of the host libm on a batch (the device carries ports of these, csrc/vr_math.cuh) */
VO_EXPORT void vo_libm_batch(int n, const float* x, float* s_out, float* c_out, float* a_out) {
#pragma omp parallel for
    for (int i = 0; i < n; ++i) {
        if (s_out) s_out[i] = sinf(x[i]);
        if (c_out) c_out[i] = cosf(x[i]);
        if (a_out) a_out[i] = acosf(x[i]);
    }
}
VO_EXPORT int vo_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------- */
/* Inverse path -- include/inverse_integrator.h, optimizer.h, gmm.h:583-706   */
/* ------------------------------------------------------------------------- */
VO_EXPORT void vo_pixel_losses(const float* I, const float* Iref, int npix, float* out) { /* inverse_integrator.h:20-29 */
    for (int p = 0; p < npix; ++p) {
        float d0 = I[3 * p] - Iref[3 * p], d1 = I[3 * p + 1] - Iref[3 * p + 1], d2 = I[3 * p + 2] - Iref[3 * p + 2];
        out[p] = fabsf(d0) + fabsf(d1) + fabsf(d2);
    }
}
/* F_diff[g] = sum over p in (P_base[g] U P_plus[g]) of loss_plus[p] - loss_base[p]
 * inverse_integrator.h:125-130,157-179.  Lists come as per-pixel (counts, idx[cap]) pairs. */
VO_EXPORT void vo_sfd_fdiff(int N, int npix, const int* cnt_base, const uint32_t* idx_base, const int* cnt_plus,
                            const uint32_t* idx_plus, int cap, const float* loss_base, const float* loss_plus,
                            double* F_diff) {
    double* Fp = (double*)calloc((size_t)N, sizeof(double));
    double* Fb = (double*)calloc((size_t)N, sizeof(double));
    for (int p = 0; p < npix; ++p) {
        const uint32_t* a = idx_base + (size_t)p * cap; int na = cnt_base[p] < cap ? cnt_base[p] : cap;
        const uint32_t* b = idx_plus + (size_t)p * cap; int nb = cnt_plus[p] < cap ? cnt_plus[p] : cap;
        for (int i = 0; i < na; ++i) if (a[i] < (uint32_t)N) { Fp[a[i]] += (double)loss_plus[p]; Fb[a[i]] += (double)loss_base[p]; }
        for (int j = 0; j < nb; ++j) {
            int dup = 0;
            for (int i = 0; i < na; ++i) if (a[i] == b[j]) { dup = 1; break; }
            if (!dup && b[j] < (uint32_t)N) { Fp[b[j]] += (double)loss_plus[p]; Fb[b[j]] += (double)loss_base[p]; }
        }
    }
    for (int g = 0; g < N; ++g) F_diff[g] = Fp[g] - Fb[g];
    free(Fp); free(Fb);
}
/* grad[i] += F_diff[i/11] * s[i] / eps[i]   (inverse_integrator.h:182-188); caller averages (:192-198) */
VO_EXPORT void vo_sfd_accumulate(int D, const double* F_diff, const signed char* s, const float* eps, double* grad_d) {
    for (int i = 0; i < D; ++i) {
        double denom = (double)eps[i];
        if (fabs(denom) < 1e-12) continue;
        grad_d[i] += F_diff[i / 11] * (double)s[i] / denom;
    }
}
VO_EXPORT void vo_default_eps(int D, float* eps) { /* gmm.h:677-706 */
    for (int i = 0; i < D; i += 11) {
        eps[i + 0] = eps[i + 1] = eps[i + 2] = 0.02f;
        eps[i + 3] = eps[i + 4] = eps[i + 5] = 0.10f;
        eps[i + 6] = eps[i + 7] = eps[i + 8] = 0.05f;
        eps[i + 9] = 0.25f;
        eps[i + 10] = 0.5f;
    }
}
/* AdamOptimizer::step, optimizer.h:30-41; t is the step count BEFORE the call (it is incremented first) */
VO_EXPORT void vo_adam_step(int D, float* params, const float* grads, float* m, float* v, int* t, float lr, float beta1,
                            float beta2, float eps) {
    *t += 1;
    const float a = lr * sqrtf(1.0f - powf(beta2, (float)*t)) / (1.0f - powf(beta1, (float)*t));
    for (int i = 0; i < D; ++i) {
        float g = grads[i];
        m[i] = beta1 * m[i] + (1.0f - beta1) * g;
        v[i] = beta2 * v[i] + (1.0f - beta2) * g * g;
        params[i] -= a * (m[i] / (sqrtf(v[i]) + eps));
    }
}
static float sigmoidf_safe(float x) { /* gmm.h:18-26 */
    if (x >= 0.0f) { float z = expf(-x); return 1.0f / (1.0f + z); }
    float z = expf(x);
    return z / (1.0f + z);
}
/* apply_params_to_gmm_local, gmm.h:634-674: 11 floats -> mean, cov6, density, albedo.
 * Sigma = R S S^T R^T with R = Rodrigues(rod) (Eigen AngleAxisf::toRotationMatrix). */
VO_EXPORT void vo_unpack_params(int N, const float* params, float* mean3, float* cov6, float* density, float* albedo) {
    for (int i = 0; i < N; ++i) {
        const float* p = params + 11 * i;
        mean3[3 * i] = p[0]; mean3[3 * i + 1] = p[1]; mean3[3 * i + 2] = p[2];
        float angle = sqrtf(p[3] * p[3] + p[4] * p[4] + p[5] * p[5]);
        float R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        if (angle > 1e-12f) {
            float ax = p[3] / angle, ay = p[4] / angle, az = p[5] / angle;
            float s = sinf(angle), c = cosf(angle);
            float sx = s * ax, sy = s * ay, sz = s * az;
            float cx = (1.0f - c) * ax, cy = (1.0f - c) * ay, cz = (1.0f - c) * az;
            float tmp;
            tmp = cx * ay; R[1] = tmp - sz; R[3] = tmp + sz;
            tmp = cx * az; R[2] = tmp + sy; R[6] = tmp - sy;
            tmp = cy * az; R[5] = tmp - sx; R[7] = tmp + sx;
            R[0] = cx * ax + c; R[4] = cy * ay + c; R[8] = cz * az + c;
        }
        float sc[3] = {expf(p[6]), expf(p[7]), expf(p[8])};
        /* ((R*S)*S^T)*R^T, evaluated left to right as Eigen does (gaussian.h:103) */
        float RS[9], RSS[9], C[9];
        for (int r = 0; r < 3; ++r) for (int k = 0; k < 3; ++k) RS[3 * r + k] = R[3 * r + k] * sc[k];
        for (int r = 0; r < 3; ++r) for (int k = 0; k < 3; ++k) RSS[3 * r + k] = RS[3 * r + k] * sc[k];
        for (int r = 0; r < 3; ++r)
            for (int k = 0; k < 3; ++k)
                C[3 * r + k] = RSS[3 * r] * R[3 * k] + RSS[3 * r + 1] * R[3 * k + 1] + RSS[3 * r + 2] * R[3 * k + 2];
        cov6[6 * i + 0] = C[0]; cov6[6 * i + 1] = C[1]; cov6[6 * i + 2] = C[2];
        cov6[6 * i + 3] = C[4]; cov6[6 * i + 4] = C[5]; cov6[6 * i + 5] = C[8];
        density[i] = expf(p[9]);
        float alb = sigmoidf_safe(p[10]);
        albedo[i] = alb < 0.0f ? 0.0f : (alb > 1.0f ? 1.0f : alb);
    }
}
/* like vo_gmm_create but from the 11-float parameterisation, keeping the full
 * (possibly slightly asymmetric) 3x3 product as the reference does */
VO_EXPORT vo_gmm* vo_gmm_create_from_params(int N, const float* params) {
    float* mean = (float*)malloc((size_t)N * 12);
    float* cov6 = (float*)malloc((size_t)N * 24);
    float* den = (float*)malloc((size_t)N * 4);
    float* alb = (float*)malloc((size_t)N * 4);
    vo_unpack_params(N, params, mean, cov6, den, alb);
    vo_gmm* m = vo_gmm_create(N, mean, cov6, den, alb);
    free(mean); free(cov6); free(den); free(alb);
    return m;
}

/* ------------------------------------------------------------------------- */
/* Sphere mixture + RayMarchingSpheres -- smm.h, test_integrators.h:11-136    */
/* env directions: PCG seeded per (x, y, step) [deviation, header note]       */
/* ------------------------------------------------------------------------- */
typedef struct { int n; const float* center; const float* radius; const float* sigma_a; const float* sigma_s; } vo_smm;

static int sphere_intersect(const vo_smm* s, int i, const vo_ray* ray, float* t_enter, float* t_exit) { /* smm.h:29-39 */
    v3 Lv = v3_sub(v3_make(s->center[3 * i], s->center[3 * i + 1], s->center[3 * i + 2]), ray->o);
    float tca = v3_dot(Lv, ray->d);
    float d2 = v3_sqnorm(Lv) - tca * tca;
    float r2 = s->radius[i] * s->radius[i];
    if (d2 > r2) return 0;
    float thc = sqrtf(r2 - d2);
    *t_enter = tca - thc;
    *t_exit = tca + thc;
    return *t_exit >= 0.0f;
}
static void smm_intersect_events(const vo_smm* s, const vo_ray* ray, vo_events* ev) { /* smm.h:54-63 */
    int start = ev->n; /* appends (the reference does not clear, smm.h:54) */
    for (int i = 0; i < s->n; ++i) {
        float t0, t1;
        if (sphere_intersect(s, i, ray, &t0, &t1)) {
            if (t0 >= 0.0f) events_push(ev, t0, 1, (uint32_t)i);
            if (t1 >= 0.0f) events_push(ev, t1, 0, (uint32_t)i);
        }
    }
    (void)start;
    if (ev->n > 1) qsort(ev->e, (size_t)ev->n, sizeof(vo_event), event_cmp);
}
/* smm.h:79-103 with the "currently active spheres prepended as t=0 entries" of test_integrators.h:71-91 */
static float smm_shadow_transmittance(const vo_smm* s, const vo_ray* ray, const unsigned char* active0, float tmax,
                                      vo_events* ev, unsigned char* act) {
    ev->n = 0;
    smm_intersect_events(s, ray, ev);
    /* prepended {0, true, i} events only switch sphere i on before the first real event */
    for (int i = 0; i < s->n; ++i) act[i] = active0[i];
    float T = 1.0f, t_prev = 0.0f;
    for (int k = 0; k < ev->n; ++k) {
        if (ev->e[k].t > tmax) break;
        float dt = ev->e[k].t - t_prev;
        float sigma_t = 0.0f;
        for (int i = 0; i < s->n; ++i) if (act[i]) sigma_t += s->sigma_a[i] + s->sigma_s[i];
        T *= expf(-sigma_t * dt);
        act[ev->e[k].index] = (unsigned char)ev->e[k].entering;
        t_prev = ev->e[k].t;
    }
    return T;
}
VO_EXPORT void vo_render_spheres(int n, const float* center, const float* radius, const float* sigma_a,
                                 const float* sigma_s, int nl, const float* lpos, const float* lint, const float* env,
                                 const vo_camera* cam, int W, int H, float step_size, int env_samples, float* out_rgb) {
    vo_smm S = {n, center, radius, sigma_a, sigma_s};
#pragma omp parallel
    {
        vo_events ev = {0, 0, 0}, sev = {0, 0, 0};
        unsigned char* active = (unsigned char*)calloc((size_t)(n > 0 ? n : 1), 1);
        unsigned char* act2 = (unsigned char*)calloc((size_t)(n > 0 ? n : 1), 1);
#pragma omp for collapse(2) schedule(dynamic, 1)
        for (int j = 0; j < H; ++j) {
            for (int i = 0; i < W; ++i) {
                float u = ((float)i + 0.5f) / (float)W, v = ((float)j + 0.5f) / (float)H;
                vo_ray ray = camera_sample_ray(cam, u, v);
                float rad[3] = {0, 0, 0};
                size_t p = (size_t)j * W + i;
                ev.n = 0;
                smm_intersect_events(&S, &ray, &ev);
                if (ev.n == 0) { for (int c = 0; c < 3; ++c) out_rgb[3 * p + c] = env[c]; continue; }
                memset(active, 0, (size_t)n);
                int cur = 0, stepi = 0;
                float t = 0.0f, T = 1.0f, t_end = ev.e[ev.n - 1].t;
                while (t < t_end) {
                    while (cur < ev.n && ev.e[cur].t <= t) { active[ev.e[cur].index] = (unsigned char)ev.e[cur].entering; ++cur; }
                    v3 pos = v3_add(ray.o, v3_scale(t, ray.d));
                    float sa = 0.0f, ss = 0.0f;
                    for (int k = 0; k < n; ++k) if (active[k]) { sa += sigma_a[k]; ss += sigma_s[k]; }
                    if (ss > 0.0f) {
                        float Li[3] = {0, 0, 0};
                        for (int l = 0; l < nl; ++l) {
                            v3 lp = v3_make(lpos[3 * l], lpos[3 * l + 1], lpos[3 * l + 2]);
                            v3 to = v3_sub(lp, pos);
                            v3 wi = v3_normalized(to);
                            float dist = sqrtf(v3_sqnorm(to));
                            vo_ray sh = ray_make(pos, wi);
                            float Tr = smm_shadow_transmittance(&S, &sh, active, dist, &sev, act2);
                            for (int c = 0; c < 3; ++c) Li[c] += (Tr * lint[3 * l + c]) / (dist * dist);
                        }
                        float Le[3] = {0, 0, 0};
                        vo_pcg rng;
                        pcg_seed(&rng, vo_derive_path_seed(i, j, stepi), 2);
                        for (int e = 0; e < env_samples; ++e) {
                            v3 wi = sample_uniform_direction(&rng);
                            vo_ray sh = ray_make(pos, wi);
                            float Tr = smm_shadow_transmittance(&S, &sh, active, INFINITY, &sev, act2);
                            for (int c = 0; c < 3; ++c) Le[c] += Tr * env[c];
                        }
                        float four_pi = (float)(4.0f * VO_PI);
                        float inv4pi = (float)(1.0f / (4.0f * VO_PI));
                        for (int c = 0; c < 3; ++c) {
                            Le[c] /= (float)env_samples;
                            Le[c] *= four_pi;
                            rad[c] += (((T * ss) * (Li[c] + Le[c])) * step_size) * inv4pi;
                        }
                    }
                    T *= expf(-step_size * (sa + ss));
                    t += step_size; /* float accumulation, test_integrators.h:127 */
                    ++stepi;
                }
                for (int c = 0; c < 3; ++c) out_rgb[3 * p + c] = rad[c] + T * env[c];
            }
        }
        free(ev.e); free(sev.e); free(active); free(act2);
    }
}

/* ------------------------------------------------------------------------- */
/* Validation integrators over Gaussians               */
/*   TestIntegrator        integrator.h:65-94                                */
/*   PureRayMarching       integrator.h:100-267                              */
/*   RayMarchingGaussians  test_integrators.h:143-297                        */
/* The reference draws the environment directions from a random_device-seeded */
/* generator (sample_uniform_direction_old, rng.h:6-10): PARITY UNPINNED for  */
/* that term.  Here, as in vo_render_spheres, they come from the PCG stream   */
/* seeded with (x, y, step index), sequence 2.                                */
/* ------------------------------------------------------------------------- */
static void evaluate_sigma_mask(const vo_gmm* m, const unsigned char* active, v3 pos, float* sigma_a, float* sigma_s) { /* gmm.h:98-126 */
    float sum = 0.0f, sum_alb = 0.0f;
    for (int i = 0; i < m->n; ++i) {
        if (!active[i]) continue;
        float mu = gauss_mu_t(&m->g[i], pos);
        sum += mu;
        sum_alb += mu * m->g[i].albedo;
    }
    if (sum <= 0.0f) { *sigma_s = *sigma_a = 0.0f; return; }
    float a_mix = sum_alb / sum;
    *sigma_s = a_mix * sum;
    *sigma_a = (1.0f - a_mix) * sum;
}
static float transmittance_over_segment_mask(const vo_gmm* m, const vo_ray* ray, float t0, float t1, const unsigned char* mask) { /* gmm.h:146-157 */
    float sum = 0.0f;
    for (int i = 0; i < m->n; ++i) if (mask[i]) sum += gauss_optical_depth(&m->g[i], ray, t0, t1);
    return expf(-sum);
}
/* events of `ray` with {0, entering, i} put in front for every active i (integrator.h:200-208, test_integrators.h:204-207) */
static void shadow_events(const vo_gmm* m, const vo_ray* ray, const unsigned char* active, vo_events* ev, vo_events* tmp, vo_stack* st) {
    gmm_intersect_events(m, ray, tmp, st);
    ev->n = 0;
    for (int i = m->n - 1; i >= 0; --i) if (active[i]) events_push(ev, 0.0f, 1, (uint32_t)i); /* repeated insert at begin() reverses */
    for (int k = 0; k < tmp->n; ++k) events_push(ev, tmp->e[k].t, tmp->e[k].entering, tmp->e[k].index);
}
/* PureRayMarching::march_transmittance, integrator.h:105-135 */
static float march_transmittance(const vo_gmm* m, const vo_ray* ray, float t_max, const vo_events* ev, const unsigned char* active0,
                                 float step, unsigned char* act) {
    memcpy(act, active0, (size_t)m->n);
    float T = 1.0f;
    int k = 0;
    for (float t = 0.0f; t < t_max; t += step) {
        while (k < ev->n && ev->e[k].t <= t) { act[ev->e[k].index] = (unsigned char)ev->e[k].entering; ++k; }
        v3 pos = v3_add(ray->o, v3_scale(t, ray->d));
        float sa, ss;
        evaluate_sigma_mask(m, act, pos, &sa, &ss);
        T *= expf(-(sa + ss) * step);
    }
    return T;
}
/* mode 0 TestIntegrator, 1 PureRayMarching, 2 RayMarchingGaussians */
VO_EXPORT void vo_render_march(const vo_gmm* m, int nl, const float* lpos, const float* lint, const float* env, const vo_camera* cam,
                               int W, int H, int mode, float step_size, int env_samples, float* out_rgb) {
    const int n = m->n;
#pragma omp parallel
    {
        vo_events ev = {0, 0, 0}, sev = {0, 0, 0}, tmp = {0, 0, 0};
        vo_stack st = {0, 0};
        unsigned char* active = (unsigned char*)calloc((size_t)(n > 0 ? n : 1), 1);
        unsigned char* mask = (unsigned char*)calloc((size_t)(n > 0 ? n : 1), 1);
#pragma omp for collapse(2) schedule(dynamic, 1)
        for (int y = 0; y < H; ++y) {
            for (int x = 0; x < W; ++x) {
                float u = ((float)x + 0.5f) / (float)W, v = ((float)y + 0.5f) / (float)H;
                vo_ray ray = camera_sample_ray(cam, u, v);
                float* dst = out_rgb + 3 * ((size_t)y * W + x);
                gmm_intersect_events(m, &ray, &ev, &st);
                if (mode == 0) { /* magenta where the primary ray overlaps anything */
                    dst[0] = ev.n ? 1.0f : env[0]; dst[1] = ev.n ? 0.0f : env[1]; dst[2] = ev.n ? 1.0f : env[2];
                    continue;
                }
                if (ev.n == 0) { for (int c = 0; c < 3; ++c) dst[c] = env[c]; continue; }
                const float t_end = ev.e[ev.n - 1].t;
                memset(active, 0, (size_t)n);
                if (mode == 1) for (int k = 0; k < ev.n; ++k) if (ev.e[k].t == 0.0f && ev.e[k].entering) active[ev.e[k].index] = 1;
                int cur = 0, stepi = 0;
                float T = 1.0f, L[3] = {0, 0, 0};
                const float four_pi = (float)(4.0f * VO_PI), inv4pi = (float)(1.0f / (4.0f * VO_PI));
                for (float t = 0.0f; t < t_end; t += step_size, ++stepi) {
                    while (cur < ev.n && ev.e[cur].t <= t) { active[ev.e[cur].index] = (unsigned char)ev.e[cur].entering; ++cur; }
                    v3 pos = v3_add(ray.o, v3_scale(t, ray.d));
                    float sa, ss;
                    evaluate_sigma_mask(m, active, pos, &sa, &ss);
                    if (ss > 0.0f) {
                        float Li[3] = {0, 0, 0}, Le[3] = {0, 0, 0};
                        for (int l = 0; l < nl; ++l) {
                            v3 to = v3_sub(v3_make(lpos[3 * l], lpos[3 * l + 1], lpos[3 * l + 2]), pos);
                            v3 wi = v3_normalized(to);
                            float dist = sqrtf(v3_sqnorm(to));
                            vo_ray sh = ray_make(pos, wi);
                            shadow_events(m, &sh, active, &sev, &tmp, &st);
                            float Tr = 1.0f;
                            if (mode == 1) {
                                Tr = march_transmittance(m, &sh, dist, &sev, active, step_size, mask);
                            } else { /* test_integrators.h:209-231: no clipping of the last segment to dist */
                                memcpy(mask, active, (size_t)n);
                                float t_prev = 0.0f;
                                int k = 0;
                                while (t_prev < dist) {
                                    float t_next = k < sev.n ? sev.e[k].t : dist;
                                    Tr *= transmittance_over_segment_mask(m, &sh, t_prev, t_next, mask);
                                    if (k < sev.n) { mask[sev.e[k].index] = (unsigned char)sev.e[k].entering; ++k; }
                                    t_prev = t_next;
                                }
                            }
                            for (int c = 0; c < 3; ++c) Li[c] += (Tr * lint[3 * l + c]) / (dist * dist);
                        }
                        vo_pcg rng;
                        pcg_seed(&rng, vo_derive_path_seed(x, y, stepi), 2);
                        for (int e = 0; e < env_samples; ++e) {
                            v3 wi = sample_uniform_direction(&rng);
                            vo_ray sh = ray_make(pos, wi);
                            shadow_events(m, &sh, active, &sev, &tmp, &st);
                            float Tr = 1.0f;
                            if (mode == 1) {
                                float t_env_end = sev.n ? sev.e[sev.n - 1].t : 0.0f;
                                Tr = march_transmittance(m, &sh, t_env_end, &sev, active, step_size, mask);
                            } else { /* test_integrators.h:247-263 */
                                memcpy(mask, active, (size_t)n);
                                float t_prev = 0.0f;
                                for (int k = 0; k < sev.n; ++k) {
                                    float t_next = sev.e[k].t;
                                    Tr *= transmittance_over_segment_mask(m, &sh, t_prev, t_next, mask);
                                    mask[sev.e[k].index] = (unsigned char)sev.e[k].entering;
                                    t_prev = t_next;
                                }
                            }
                            for (int c = 0; c < 3; ++c) Le[c] += Tr * env[c];
                        }
                        for (int c = 0; c < 3; ++c) {
                            Le[c] /= (float)env_samples;
                            Le[c] *= four_pi;
                            L[c] += (((T * ss) * (Li[c] + Le[c])) * step_size) * inv4pi;
                        }
                    }
                    if (mode == 1) T *= expf(-(sa + ss) * step_size);
                    else T *= transmittance_over_segment_mask(m, &ray, t, t + step_size, active);
                }
                for (int c = 0; c < 3; ++c) dst[c] = L[c] + T * env[c];
            }
        }
        free(ev.e); free(sev.e); free(tmp.e); free(st.s); free(active); free(mask);
    }
}
