// vr_device.cuh -- per-ray device routines: hierarchy traversal, overlap gathering, free-flight
// walk, distance solvers, shadow transmittance.  Included by vr_kernels.cu only.
#pragma once
#include "vr_math.cuh"

namespace vr {

struct SceneDev {
    const float4* gauss;    // 3 x float4 per Gaussian, LEAF order
    const float4* nodes;    // 4 x float4 per hierarchy node (vr_bvh.h BvhNode)
    const uint4* qnodes;    // 2 x uint4 per node: the same tree with 16-bit boxes (vr_bvh.h BvhNodeQ), wavefront kernels
    float qorg[3], qcell[3];
    // per-light direction grids (vr_lightgrid.h): cell_start[light][6*R*R+1], entries sorted by distance from the light
    const uint32_t* lg_start;
    const uint32_t* lg_gi;
    const float* lg_dmin;
    int lg_res;
    uint32_t lg_mask;       // bit l: light l has a grid
    // direction grid around the camera's pinhole (same layout, one table): candidates of the camera rays; cg_res = 0: none
    const uint32_t* cg_start;
    const uint32_t* cg_gi;
    const float* cg_dmin;
    int cg_res;
    // uniform grid over the mixture (vr_grid3d.cu): candidates of the first stretch of a bounce ray; g3_res = 0: none
    const uint32_t* g3_start;
    const uint32_t* g3_gi;
    float g3_org[3], g3_inv[3], g3_cell[3];
    int g3_res;
    const uint32_t* orig;   // leaf order -> original index
    int n;
    int nl;
    const float* lpos;      // 3*nl
    const float* lint;      // 3*nl
    float env[3];
};

constexpr int kStackDepth = 64;
// Executed-work counters (hierarchy steps, quadric tests, erf / mu_t evaluations, solves) cost registers in the hottest loops
// (-5 % at C2), so they are compiled only into the instrumented variant of the library (csrc/Makefile: libvolrender_b200_stats.so,
// -DVR_WORK_COUNTERS), which bench.py loads for one untimed render to fill roofline.executed.
#ifdef VR_WORK_COUNTERS
#define VR_COUNT(stmt) stmt
#else
#define VR_COUNT(stmt)
#endif
constexpr float kShadowTauCutoff = 120.0f;  // expf(-x) == 0 for x > 103.98: exact early-out

// ------------------------------------------------------------------------------------------
// overlap-list storage policies.  Fields per overlap: t0,t1 (clamped entry / exit), the optical
// depth constants A,B,K,den, the leaf-order Gaussian index, and scratch for the sorted walk.
// ------------------------------------------------------------------------------------------
template <int CAP>
struct LocalStore {
    float f[7][CAP];
    uint32_t u[3][CAP];
#ifdef VR_WORK_COUNTERS   // executed work of the walk / solvers on this store (bench.py roofline.executed)
    unsigned int n_erf = 0, n_mu = 0, n_sa = 0, n_it = 0;
#endif
    __device__ __forceinline__ int cap() const { return CAP; }
    __device__ __forceinline__ float& F(int field, int s) { return f[field][s]; }
    __device__ __forceinline__ uint32_t& U(int field, int s) { return u[field][s]; }
};
struct ArenaStore {   // global memory, warp-interleaved: element (field, slot) of lane l at base[(field*cap+slot)*32 + l]
    float* base;
    int capacity;
#ifdef VR_WORK_COUNTERS
    unsigned int n_erf = 0, n_mu = 0, n_sa = 0, n_it = 0;
#endif
    __device__ __forceinline__ int cap() const { return capacity; }
    __device__ __forceinline__ float& F(int field, int s) { return base[((size_t)field * capacity + s) * 32]; }
    __device__ __forceinline__ uint32_t& U(int field, int s) {
        return reinterpret_cast<uint32_t*>(base)[((size_t)(7 + field) * capacity + s) * 32];
    }
};
constexpr int kArenaFields = 10;
enum { F_T0 = 0, F_T1, F_A, F_B, F_K, F_DEN, F_ACTF };   // float fields
enum { U_GI = 0, U_ORDER, U_ACT };                        // uint fields

template <class St>
__device__ __forceinline__ DepthConst hit_const(St& st, int s) {
    DepthConst c; c.A = st.F(F_A, s); c.B = st.F(F_B, s); c.K = st.F(F_K, s); c.den = st.F(F_DEN, s);
    return c;
}

struct RaySetup { f3 idir, ood; };
__device__ __forceinline__ RaySetup ray_setup(f3 o, f3 d) {
    const float tiny = 8.271806e-25f;  // 2^-80
    RaySetup r;
    r.idir.x = 1.0f / (fabsf(d.x) > tiny ? d.x : copysignf(tiny, d.x));
    r.idir.y = 1.0f / (fabsf(d.y) > tiny ? d.y : copysignf(tiny, d.y));
    r.idir.z = 1.0f / (fabsf(d.z) > tiny ? d.z : copysignf(tiny, d.z));
    r.ood = mk3(o.x * r.idir.x, o.y * r.idir.y, o.z * r.idir.z);
    return r;
}
// slab test of the two children of a node against [0, tlimit]
__device__ __forceinline__ void node_slabs(const float4* __restrict__ nodes, int ni, const RaySetup& rs, float tlimit,
                                           bool* h0, bool* h1, float* tn0, float* tn1, int* c0, int* c1) {
    const float4 n0 = __ldg(nodes + 4 * (size_t)ni), n1 = __ldg(nodes + 4 * (size_t)ni + 1),
                 n2 = __ldg(nodes + 4 * (size_t)ni + 2), n3 = __ldg(nodes + 4 * (size_t)ni + 3);
    float ax = fmaf(n0.x, rs.idir.x, -rs.ood.x), bx = fmaf(n0.y, rs.idir.x, -rs.ood.x);
    float ay = fmaf(n0.z, rs.idir.y, -rs.ood.y), by = fmaf(n0.w, rs.idir.y, -rs.ood.y);
    float az = fmaf(n2.x, rs.idir.z, -rs.ood.z), bz = fmaf(n2.y, rs.idir.z, -rs.ood.z);
    float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
    float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tlimit));
    *h0 = tmin <= tmax; *tn0 = tmin;
    ax = fmaf(n1.x, rs.idir.x, -rs.ood.x); bx = fmaf(n1.y, rs.idir.x, -rs.ood.x);
    ay = fmaf(n1.z, rs.idir.y, -rs.ood.y); by = fmaf(n1.w, rs.idir.y, -rs.ood.y);
    az = fmaf(n2.z, rs.idir.z, -rs.ood.z); bz = fmaf(n2.w, rs.idir.z, -rs.ood.z);
    tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
    tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tlimit));
    *h1 = tmin <= tmax; *tn1 = tmin;
    *c0 = __float_as_int(n3.x); *c1 = __float_as_int(n3.y);
}

// the same test on the compact tree: coordinate = qorg + q * qcell, so the slab distance of a quantised plane is
// q * (qcell * idir) + (qorg * idir - o * idir) = q * s + b with per-ray s, b.  A 16-bit value q is turned into the
// float 2^23 + q by ONE byte permute (q lands in the low mantissa bits of 0x4B000000) and the 2^23 is folded into the
// per-ray offset: t = (2^23 + q) * s + (b - 2^23 * s).  Rounding that offset costs at most half a cell of slab
// distance, which the builder's outward rounding by two whole cells covers (vr_bvh.cpp quantize_bvh).  The permute
// selector is per ray and axis: it picks the NEAR plane (lo for a positive direction, hi for a negative one) or the
// FAR plane directly, so no min/max between the two planes of an axis is needed.
struct RaySetupQ {
    f3 s, b;                    // t = float(2^23 + q) * s + b
    unsigned int near_sel[3];   // __byte_perm selectors of the near plane; far = near ^ 0x0022
};
__device__ __forceinline__ RaySetupQ ray_setup_q(const SceneDev& S, f3 o, f3 d) {
    const RaySetup r = ray_setup(o, d);
    RaySetupQ q;
    q.s = mk3(S.qcell[0] * r.idir.x, S.qcell[1] * r.idir.y, S.qcell[2] * r.idir.z);
    const f3 b = mk3(fmaf(S.qorg[0], r.idir.x, -r.ood.x), fmaf(S.qorg[1], r.idir.y, -r.ood.y), fmaf(S.qorg[2], r.idir.z, -r.ood.z));
    q.b = mk3(fmaf(-8388608.0f, q.s.x, b.x), fmaf(-8388608.0f, q.s.y, b.y), fmaf(-8388608.0f, q.s.z, b.z));
    // selector 0x7610: bytes (w.b0, w.b1, 0x00, 0x4B) = 2^23 + lo;  0x7632: (w.b2, w.b3, 0x00, 0x4B) = 2^23 + hi
    q.near_sel[0] = q.s.x >= 0.0f ? 0x7610u : 0x7632u;
    q.near_sel[1] = q.s.y >= 0.0f ? 0x7610u : 0x7632u;
    q.near_sel[2] = q.s.z >= 0.0f ? 0x7610u : 0x7632u;
    return q;
}
__device__ __forceinline__ float q_plane(unsigned int w, unsigned int sel) { return __uint_as_float(__byte_perm(w, 0x4B000000u, sel)); }
__device__ __forceinline__ void node_slabs_q(const uint4* __restrict__ qnodes, int ni, const RaySetupQ& rs, float tlimit,
                                             bool* h0, bool* h1, float* tn0, float* tn1, int* c0, int* c1) {
    const uint4 u = __ldg(qnodes + 2 * (size_t)ni), v = __ldg(qnodes + 2 * (size_t)ni + 1);
    const unsigned int nx = rs.near_sel[0], ny = rs.near_sel[1], nz = rs.near_sel[2];
    const unsigned int fx = nx ^ 0x0022u, fy = ny ^ 0x0022u, fz = nz ^ 0x0022u;
    float tmin = fmaxf(fmaxf(fmaf(q_plane(u.x, nx), rs.s.x, rs.b.x), fmaf(q_plane(u.y, ny), rs.s.y, rs.b.y)),
                       fmaxf(fmaf(q_plane(u.z, nz), rs.s.z, rs.b.z), 0.0f));
    float tmax = fminf(fminf(fmaf(q_plane(u.x, fx), rs.s.x, rs.b.x), fmaf(q_plane(u.y, fy), rs.s.y, rs.b.y)),
                       fminf(fmaf(q_plane(u.z, fz), rs.s.z, rs.b.z), tlimit));
    *h0 = tmin <= tmax; *tn0 = tmin;
    tmin = fmaxf(fmaxf(fmaf(q_plane(u.w, nx), rs.s.x, rs.b.x), fmaf(q_plane(v.x, ny), rs.s.y, rs.b.y)),
                 fmaxf(fmaf(q_plane(v.y, nz), rs.s.z, rs.b.z), 0.0f));
    tmax = fminf(fminf(fmaf(q_plane(u.w, fx), rs.s.x, rs.b.x), fmaf(q_plane(v.x, fy), rs.s.y, rs.b.y)),
                 fminf(fmaf(q_plane(v.y, fz), rs.s.z, rs.b.z), tlimit));
    *h1 = tmin <= tmax; *tn1 = tmin;
    *c0 = (int)v.z; *c1 = (int)v.w;
}

// cell of the direction v = (ray origin - light) in the light's cube map; same formula as light_grid_cell (vr_lightgrid.h)
__device__ __forceinline__ uint32_t light_grid_cell_dev(f3 v, int res) {
    const float ax = fabsf(v.x), ay = fabsf(v.y), az = fabsf(v.z);
    int m = 0; float am = ax;
    if (ay > am) { m = 1; am = ay; }
    if (az > am) { m = 2; am = az; }
    const float vm = m == 0 ? v.x : (m == 1 ? v.y : v.z);
    const float va = m == 0 ? v.y : (m == 1 ? v.z : v.x);   // axis (m + 1) % 3
    const float vb = m == 0 ? v.z : (m == 1 ? v.x : v.y);   // axis (m + 2) % 3
    const int face = 2 * m + (vm < 0.0f ? 1 : 0);
    const float half = 0.5f * (float)res;
    int iu = (int)((va / am + 1.0f) * half), iw = (int)((vb / am + 1.0f) * half);
    iu = min(res - 1, max(0, iu)); iw = min(res - 1, max(0, iw));
    return (uint32_t)((face * res + iw) * res + iu);
}

// ------------------------------------------------------------------------------------------
// Gather the components a free-flight ray overlaps (GaussianMixtureModel::intersect_events,
// gmm.h:457-515), front to back, and stop descending behind t_bound: the exit time of a gathered
// component whose own full optical depth already exceeds the sampled target.  The collision must
// happen at or before that exit EVENT, and every event up to and including it is still gathered,
// so the reference's segment walk sees exactly the events it would act on.
// Returns the number of overlaps, or -1 when the store overflowed.
// ------------------------------------------------------------------------------------------
template <class St>
__device__ int gather_overlaps(const SceneDev& S, f3 o, f3 d, float target_tau, bool prune, St& st) {
    if (S.n == 0) return 0;
    const RaySetup rs = ray_setup(o, d);
    int stack[kStackDepth];
    int sp = 0, cur = 0, K = 0;
    float t_bound = CUDART_INF_F;
    const float need = target_tau * 1.0001f + 1e-6f;
    while (true) {
        if (cur >= 0) {
            bool h0, h1; float tn0, tn1; int c0, c1;
            node_slabs(S.nodes, cur, rs, t_bound, &h0, &h1, &tn0, &tn1, &c0, &c1);
            if (h0 && h1) {
                bool swap = tn1 < tn0;
                stack[sp++] = swap ? c0 : c1;
                cur = swap ? c1 : c0;
                continue;
            }
            if (h0) { cur = c0; continue; }
            if (h1) { cur = c1; continue; }
        } else {
            const uint32_t code = ~(uint32_t)cur;
            const uint32_t first = code >> 3, cnt = code & 7u;
            for (uint32_t i = 0; i < cnt; ++i) {
                const GaussRec g = load_gauss(S.gauss, first + i);
                const RayQuad q = ray_quad(g, o, d);
                float t0, t1;
                if (!quad_intersect(q, &t0, &t1)) continue;
                if (t0 > t_bound) continue;
                if (K >= st.cap()) return -1;
                const DepthConst c = depth_const(g, q);
                st.F(F_T0, K) = t0; st.F(F_T1, K) = t1;
                st.F(F_A, K) = c.A; st.F(F_B, K) = c.B; st.F(F_K, K) = c.K; st.F(F_DEN, K) = c.den;
                st.U(U_GI, K) = first + i;
                ++K;
                if (prune && t1 < t_bound) {
                    float full = depth_between(c, t0, t1);
                    if (full > need) t_bound = t1;
                }
            }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return K;
}

// GaussianMixtureModel::transmittance_up_to (gmm.h:517-578): exp(-sum of clipped optical depths).  The reference sums
// in double in ITS traversal order; every engine here sums the float depths in 2^-40 fixed point instead, which is
// order-independent and within 1e-12 of the double sum, so that a sample's value does not depend on which kernel
// (wavefront, drain, overflow re-trace) or which candidate order produced it.
constexpr double kTauFixed = 1099511627776.0;  // 2^40
__device__ __forceinline__ long long tau_cutoff_fixed() { return (long long)((double)kShadowTauCutoff * kTauFixed); }
static __device__ float shadow_transmittance(const SceneDev& S, f3 o, f3 d, float tmax) {
    if (tmax <= 0.0f) return 1.0f;
    if (S.n == 0) return 1.0f;
    const RaySetup rs = ray_setup(o, d);
    int stack[kStackDepth];
    int sp = 0, cur = 0;
    long long sum = 0;
    while (true) {
        if (cur >= 0) {
            bool h0, h1; float tn0, tn1; int c0, c1;
            node_slabs(S.nodes, cur, rs, tmax, &h0, &h1, &tn0, &tn1, &c0, &c1);
            if (h0 && h1) {
                bool swap = tn1 < tn0;
                stack[sp++] = swap ? c0 : c1;
                cur = swap ? c1 : c0;
                continue;
            }
            if (h0) { cur = c0; continue; }
            if (h1) { cur = c1; continue; }
        } else {
            const uint32_t code = ~(uint32_t)cur;
            const uint32_t first = code >> 3, cnt = code & 7u;
            for (uint32_t i = 0; i < cnt; ++i) {
                const GaussRec g = load_gauss(S.gauss, first + i);
                const RayQuad q = ray_quad(g, o, d);
                float t0, t1;
                if (!quad_intersect(q, &t0, &t1)) continue;
                const float a = fmaxf(0.0f, t0), b = fminf(tmax, t1);
                if (b > a) sum += __double2ll_rn((double)depth_between(depth_const(g, q), a, b) * kTauFixed);
            }
            if (sum > tau_cutoff_fixed()) return 0.0f;  // expf(-x) is exactly 0 from here on
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return sum > tau_cutoff_fixed() ? 0.0f : expf(-(float)((double)sum / kTauFixed));
}

// ------------------------------------------------------------------------------------------
// distance solvers -- include/distance_solvers.h:25-187.  act[0..na) are overlap slots, actF holds
// erf(arg(ta)) of each (bitwise what the reference recomputes inside optical_depth).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float clampf(float v, float lo, float hi) { return (v < lo) ? lo : ((hi < v) ? hi : v); }

template <class St>
__device__ __noinline__ float sum_depth_to(St& st, int na, float tt) {  // sum_k tau_k(ta, tt), float, active order
    float sum = 0.0f;
    VR_COUNT(st.n_erf += (unsigned int)na;)
    for (int k = 0; k < na; ++k) {
        const int s = (int)st.U(U_ACT, k);
        const DepthConst c = hit_const(st, s);
        sum = __fadd_rn(sum, __fmul_rn(c.K, __fsub_rn(erff_ni(erf_arg(c, tt)), st.F(F_ACTF, k))));
    }
    return sum;
}
template <class St>
__device__ __noinline__ float solve_bisection(St& st, int na, float ta, float tb, float target) {  // :25-57
    float a = ta, b = tb;
    VR_COUNT(++st.n_it;)
    for (int i = 0; i < 15; ++i) {
        float mid = __fmul_rn(0.5f, __fadd_rn(a, b));
        float f = __fsub_rn(sum_depth_to(st, na, mid), target);
        if (fabsf(f) <= 1e-6f) return mid;
        if (f < 0.0f) a = mid; else b = mid;
    }
    return __fmul_rn(0.5f, __fadd_rn(a, b));
}
template <class St>
__device__ __noinline__ float solve_newton(St& st, int na, float ta, float tb, float target) {  // :62-127
    const float tol = 1e-6f;
    const float a = ta, b = tb;
    VR_COUNT(++st.n_it;)
    float t = __fmul_rn(0.5f, __fadd_rn(a, b));
    for (int iter = 0; iter < 8; ++iter) {
        float f = __fsub_rn(sum_depth_to(st, na, fminf(t, b)), target);
        if (fabsf(f) <= tol) return clampf(t, a, b);
        float h = fmaxf(1e-5f, __fmul_rn(__fsub_rn(b, a), 1e-6f));
        float tp = fminf(b, __fadd_rn(t, h));
        float fp = __fsub_rn(sum_depth_to(st, na, fminf(tp, b)), target);
        float deriv = __fdiv_rn(__fsub_rn(fp, f), __fsub_rn(tp, t));
        if (!(deriv > 0.0f) || !isfinite(deriv) || fabsf(deriv) < 1e-12f) return solve_bisection(st, na, ta, tb, target);
        float t_next = __fsub_rn(t, __fdiv_rn(f, deriv));
        if (!isfinite(t_next) || t_next < a || t_next > b) return solve_bisection(st, na, ta, tb, target);
        if (fabsf(__fsub_rn(t_next, t)) <= __fmul_rn(tol, fmaxf(1.0f, fabsf(t)))) return clampf(t_next, a, b);
        t = t_next;
    }
    return solve_bisection(st, na, ta, tb, target);
}
// DEFER: the iterative solvers are not run here; *deferred is set instead, the segment [ta, tb] and the remaining depth
// go to *ds, and the caller hands the slot to wave_post_a_kernel<4> (or <2>), where a warp's lanes iterate together
struct DeferState { float ta, tb, remaining; };
template <bool DEFER, class St>
__device__ float solve_distance(const SceneDev& S, St& st, int na, f3 o, f3 d, float ta, float tb, float remaining,
                                int solver, uint64_t key, bool* deferred, DeferState* ds = nullptr) {  // :150-187
    if (ds) { ds->ta = ta; ds->tb = tb; ds->remaining = remaining; }   // the segment the walk ended in (also read by post_a_certified)
    if (solver == 4) {  // UNIFORM (:131-137): t = ta + u (tb - ta); the reference's u comes from a random_device-seeded
                        // generator (parity unpinned), ours is hashed from the path's PCG state after the target draw,
                        // like the oracle's solve_uniform
        const float u = __fmul_rn((float)(uint32_t)(splitmix64(key) >> 40), 1.0f / 16777216.0f);
        return __fadd_rn(ta, __fmul_rn(u, __fsub_rn(tb, ta)));
    }
    if ((solver == 0 || solver == 1) && na == 1) {
        const int s = (int)st.U(U_ACT, 0);
        const GaussRec g = load_gauss(S.gauss, st.U(U_GI, s));
        const RayQuad q = ray_quad(g, o, d);
        float t_an;
        VR_COUNT(++st.n_sa;)
        if (solve_t_given_tau(g, q, ta, tb, remaining, &t_an)) return clampf(t_an, ta, tb);
    }
    if (DEFER) { *deferred = true; return ta; }
    if (solver == 0 || solver == 2) return solve_newton(st, na, ta, tb, remaining);
    return solve_bisection(st, na, ta, tb, remaining);
}

// ------------------------------------------------------------------------------------------
// free-flight walk over the sorted events (integrator.h:422-498 multi-scatter, :333-361 single
// scatter).  Events are consumed in (t, original index, enter-before-exit) order without being
// materialised: the next event is the earlier of the next entry (overlaps sorted by entry time)
// and the earliest exit among the active components.  Returns t or -1; *na_out = active count,
// active slots in st.U(U_ACT, .) in the reference's list order.
// ------------------------------------------------------------------------------------------
template <class St>
__device__ __forceinline__ bool entry_less(const SceneDev& S, St& st, int a, int b) {
    const float ta = st.F(F_T0, a), tb = st.F(F_T0, b);
    if (ta != tb) return ta < tb;
    return __ldg(S.orig + st.U(U_GI, a)) < __ldg(S.orig + st.U(U_GI, b));
}
template <class St>
__device__ void sort_entries(const SceneDev& S, St& st, int K) {
    for (int i = 0; i < K; ++i) st.U(U_ORDER, i) = (uint32_t)i;
    // shell sort (keys are a total order, so stability is irrelevant); plain insertion sort for small K
    int gap = 1;
    while (gap < K / 3) gap = 3 * gap + 1;
    for (; gap >= 1; gap /= 3) {
        for (int i = gap; i < K; ++i) {
            const int v = (int)st.U(U_ORDER, i);
            int j = i;
            while (j >= gap && entry_less(S, st, v, (int)st.U(U_ORDER, j - gap))) {
                st.U(U_ORDER, j) = st.U(U_ORDER, j - gap);
                j -= gap;
            }
            st.U(U_ORDER, j) = (uint32_t)v;
        }
    }
}

// next event in (t, original index, enter-before-exit) order; false when the list is exhausted.
// *se = overlap slot entering (take_enter) / *kx = position in the active list leaving (!take_enter)
template <class St>
__device__ __forceinline__ bool next_event(const SceneDev& S, St& st, int K, int na, int ie, bool* take_enter_out,
                                           int* se_out, int* kx_out, float* t_evt_out) {
    int kx = -1; float tx = 0.0f;
    for (int k = 0; k < na; ++k) {  // earliest exit among the active components
        const int s = (int)st.U(U_ACT, k);
        const float t1 = st.F(F_T1, s);
        if (kx < 0 || t1 < tx) { kx = k; tx = t1; }
        else if (t1 == tx &&
                 __ldg(S.orig + st.U(U_GI, s)) < __ldg(S.orig + st.U(U_GI, (int)st.U(U_ACT, kx)))) kx = k;
    }
    const bool has_enter = ie < K;
    if (!has_enter && kx < 0) return false;
    bool take_enter = has_enter;
    int se = -1;
    if (has_enter) {
        se = (int)st.U(U_ORDER, ie);
        if (kx >= 0) {
            const float te = st.F(F_T0, se);
            if (tx < te) take_enter = false;
            else if (tx == te)
                take_enter = __ldg(S.orig + st.U(U_GI, se)) <= __ldg(S.orig + st.U(U_GI, (int)st.U(U_ACT, kx)));
        }
    }
    *take_enter_out = take_enter; *se_out = se; *kx_out = kx;
    *t_evt_out = take_enter ? st.F(F_T0, se) : tx;
    return true;
}

template <bool DEFER = false, class St>
__device__ float free_flight(const SceneDev& S, St& st, int K, f3 o, f3 d, float target_tau, int solver, bool ss,
                             int* na_out, uint64_t key, bool* deferred = nullptr, DeferState* ds = nullptr) {
    sort_entries(S, st, K);
    double acc = 0.0;
    float accf = 0.0f;
    float t_prev = 0.0f;
    int na = 0, ie = 0;
    while (true) {
        bool take_enter; int se, kx; float t_evt;
        if (!next_event(S, st, K, na, ie, &take_enter, &se, &kx, &t_evt)) break;
        // optical depth of the segment [t_prev, t_evt] over the active set
        bool hit_target;
        float remaining;
        VR_COUNT(st.n_erf += (unsigned int)na;)
        if (ss) {
            float seg = 0.0f;
            for (int k = 0; k < na; ++k) {
                const DepthConst c = hit_const(st, (int)st.U(U_ACT, k));
                const float Fn = erff_ni(erf_arg(c, t_evt));
                seg = __fadd_rn(seg, __fmul_rn(c.K, __fsub_rn(Fn, st.F(F_ACTF, k))));
                st.F(F_ACTF, k) = Fn;
            }
            hit_target = __fadd_rn(accf, seg) > target_tau;
            remaining = __fsub_rn(target_tau, accf);
            if (!hit_target) accf = __fadd_rn(accf, seg);
        } else {
            double seg = 0.0;
            for (int k = 0; k < na; ++k) {
                const DepthConst c = hit_const(st, (int)st.U(U_ACT, k));
                const float Fn = erff_ni(erf_arg(c, t_evt));
                seg = __dadd_rn(seg, (double)__fmul_rn(c.K, __fsub_rn(Fn, st.F(F_ACTF, k))));
                st.F(F_ACTF, k) = Fn;
            }
            hit_target = __dadd_rn(acc, seg) > (double)target_tau;
            remaining = (float)__dsub_rn((double)target_tau, acc);
            if (!hit_target) acc = __dadd_rn(acc, seg);
        }
        if (hit_target) {
            if (!DEFER) {   // (a deferred solve restores them itself; the analytic and the uniform one do not read them)
                VR_COUNT(st.n_erf += (unsigned int)na;)
                for (int k = 0; k < na; ++k)  // restore erf(arg(t_prev)) for the solver
                    st.F(F_ACTF, k) = erff_ni(erf_arg(hit_const(st, (int)st.U(U_ACT, k)), t_prev));
            }
            *na_out = na;
            return solve_distance<DEFER>(S, st, na, o, d, t_prev, t_evt, remaining, solver, key, deferred, ds);
        }
        if (take_enter) {
            const float F0 = erff_ni(erf_arg(hit_const(st, se), t_evt));
            VR_COUNT(++st.n_erf;)
            int pos = na;
            if (ss) {  // FreeFlightGaussians rescans a bool vector in ascending index order (integrator.h:344-346)
                const uint32_t oe = __ldg(S.orig + st.U(U_GI, se));
                while (pos > 0 && __ldg(S.orig + st.U(U_GI, (int)st.U(U_ACT, pos - 1))) > oe) {
                    st.U(U_ACT, pos) = st.U(U_ACT, pos - 1);
                    st.F(F_ACTF, pos) = st.F(F_ACTF, pos - 1);
                    --pos;
                }
            }
            st.U(U_ACT, pos) = (uint32_t)se;
            st.F(F_ACTF, pos) = F0;
            ++na; ++ie;
        } else {
            if (ss) {
                for (int k = kx; k + 1 < na; ++k) { st.U(U_ACT, k) = st.U(U_ACT, k + 1); st.F(F_ACTF, k) = st.F(F_ACTF, k + 1); }
            } else {  // swap-remove, integrator.h:476-486
                st.U(U_ACT, kx) = st.U(U_ACT, na - 1);
                st.F(F_ACTF, kx) = st.F(F_ACTF, na - 1);
            }
            --na;
        }
        t_prev = t_evt;
    }
    *na_out = 0;
    return -1.0f;
}

// GaussianMixtureModel::evaluate_albedo (gmm.h:128-143) over the active list
template <class St>
__device__ float evaluate_albedo(const SceneDev& S, St& st, int na, f3 pos) {
    float sum_mu = 0.0f, sum_mu_alb = 0.0f;
    VR_COUNT(st.n_mu += (unsigned int)na;)
    for (int k = 0; k < na; ++k) {
        const GaussRec g = load_gauss(S.gauss, st.U(U_GI, (int)st.U(U_ACT, k)));
        const float mu = gauss_mu_t(g, pos);
        sum_mu = __fadd_rn(sum_mu, mu);
        sum_mu_alb = __fadd_rn(sum_mu_alb, __fmul_rn(mu, g.albedo));
    }
    const float a = __fdiv_rn(sum_mu_alb, sum_mu);
    if (a < 0.0f) return 0.0f;
    if (1.0f < a) return 1.0f;
    return a;  // NaN stays NaN like std::clamp
}

}  // namespace vr
