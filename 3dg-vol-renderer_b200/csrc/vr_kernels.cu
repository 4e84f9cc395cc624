// vr_kernels.cu -- sm_100a kernels of the forward path.
//
//   path_kernel<Store>   persistent, one lane = one camera path at a time (MultiScatterGaussians::render
//                        integrator.h:524-717, FreeFlightGaussians::render :300-408).  Lanes whose path
//                        ended fetch the next (pixel, sample) from a global counter (warp-aggregated
//                        atomic), so path-length divergence does not idle the warp.  Each finished
//                        sample's radiance goes to a per-sample buffer; pixel_reduce_kernel then sums
//                        the spp values of a pixel IN SAMPLE ORDER, which is the reference's float
//                        accumulation order (integrator.h:702-706) -- deterministic, no atomics.
//   transmittance_kernel batch of GaussianMixtureModel::transmittance_up_to (gmm.h:517-578).
//   kat_*                known-answer hooks over the same device functions.
#include "vr_bvh.h"
#include "vr_internal.h"
#include "vr_warp.cuh"

static_assert(vr::kMaxLeafDev == vr::kMaxLeaf, "device leaf width must match the host builder");

namespace vr {

// ------------------------------------------------------------------------------------------
// one lane's path state
// ------------------------------------------------------------------------------------------
struct PathState {
    uint64_t rng_state;
    f3 o, d;
    float thr;          // albedo is scalar, so the RGB throughput is one replicated value
    float L0, L1, L2;
    unsigned int work;  // pass-local sample id
    int x, y;
    int bounce;
};

__device__ __forceinline__ void path_begin(const RenderArgs& A, unsigned int work, PathState& ps) {
    const unsigned int p = A.pixel_base + work / (unsigned int)A.spp;
    const int si = (int)(work % (unsigned int)A.spp);
    ps.work = work;
    ps.x = A.x0 + (int)(p % (unsigned int)A.win_w);
    ps.y = A.y0 + (int)(p / (unsigned int)A.win_w) * A.part_count + A.part_rank;
    Pcg rng; rng.seed(derive_path_seed(ps.x, ps.y, si), 1);
    const int n = A.nsq;
    const int sx = si % n, sy = si / n;
    // integrator.h:562-568
    const float u = __fdiv_rn(__fadd_rn((float)ps.x, __fdiv_rn(__fadd_rn((float)sx, rng.uniform()), (float)n)), (float)A.W);
    const float v = __fdiv_rn(__fadd_rn((float)ps.y, __fdiv_rn(__fadd_rn((float)sy, rng.uniform()), (float)n)), (float)A.H);
    const RayOD r = camera_ray(A.cam, u, v);
    ps.o = r.o; ps.d = r.d;
    ps.thr = 1.0f; ps.L0 = ps.L1 = ps.L2 = 0.0f;
    ps.bounce = 0;
    ps.rng_state = rng.state;
}

// next-event estimation, integrator.h:656-678 (== :369-392)
__device__ __forceinline__ void nee_sample(const SceneDev& S, Pcg& rng, f3 pos, float Li[3], unsigned int& n_shadow) {
    const bool is_env = rng.uniform() < __fdiv_rn(1.0f, (float)(S.nl + 1));
    ++n_shadow;
    if (!is_env) {
        int li = (int)__fmul_rn(rng.uniform(), (float)S.nl);
        li = min(li, S.nl - 1);
        const f3 lp = mk3(__ldg(S.lpos + 3 * li), __ldg(S.lpos + 3 * li + 1), __ldg(S.lpos + 3 * li + 2));
        const f3 to = sub3(lp, pos);
        const f3 wi = normalized3(to);
        const float dist = __fsqrt_rn(sqnorm3(to));
        const float Tr = shadow_transmittance(S, pos, normalized3(wi), dist);
        const float d2 = __fmul_rn(dist, dist);
        Li[0] = __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * li)), d2);
        Li[1] = __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * li + 1)), d2);
        Li[2] = __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * li + 2)), d2);
    } else {
        const f3 wi = sample_uniform_direction(rng);
        const float Tr = shadow_transmittance(S, pos, normalized3(wi), CUDART_INF_F);
        const float four_pi = (float)(4.0 * VR_PI_D);
        Li[0] = __fmul_rn(__fmul_rn(Tr, S.env[0]), four_pi);
        Li[1] = __fmul_rn(__fmul_rn(Tr, S.env[1]), four_pi);
        Li[2] = __fmul_rn(__fmul_rn(Tr, S.env[2]), four_pi);
    }
}

__device__ __forceinline__ void record_overlap(const RenderArgs& A, const PathState& ps, uint32_t orig) {
    uint32_t* row = A.record_bits + ((size_t)ps.y * A.W + ps.x) * (size_t)A.record_words;
    const uint32_t bit = 1u << (orig & 31u);
    if (!(row[orig >> 5] & bit)) atomicOr(row + (orig >> 5), bit);
}

// One path vertex.  Returns: 0 path continues, 1 path finished, -1 overlap store overflowed.
template <class St>
__device__ __forceinline__ int path_vertex(const RenderArgs& A, PathState& ps, St& st, unsigned int& n_scatter,
                                           unsigned int& n_shadow, unsigned int& n_hits) {
    const SceneDev& S = A.scene;
    Pcg rng; rng.state = ps.rng_state; rng.inc = 3;
    // the reference draws the target only when the ray overlaps something (integrator.h:593-600);
    // draw it on a copy so the gather can prune with it, commit the draw afterwards.
    Pcg peek = rng;
    const float target_tau = -logf(__fsub_rn(1.0f, peek.uniform()));
    ++n_scatter;
    const int K = gather_overlaps(S, ps.o, ps.d, target_tau, true, st);
    if (K < 0) return -1;
    if (K == 0) {
        ps.L0 = __fadd_rn(ps.L0, __fmul_rn(ps.thr, S.env[0]));
        ps.L1 = __fadd_rn(ps.L1, __fmul_rn(ps.thr, S.env[1]));
        ps.L2 = __fadd_rn(ps.L2, __fmul_rn(ps.thr, S.env[2]));
        return 1;
    }
    n_hits += (unsigned int)K;
    rng = peek;
    int na = 0;
    const bool ss = A.integrator == 0;
    const float t = free_flight(S, st, K, ps.o, ps.d, target_tau, A.solver, ss, &na, rng.state);
    if (A.record_bits) {  // integrator.h:616-644
        const float tlim = __fadd_rn(t, 1e-6f);
        for (int s = 0; s < K; ++s)
            if (t < 0.0f || st.F(F_T0, s) <= tlim) record_overlap(A, ps, __ldg(S.orig + st.U(U_GI, s)));
    }
    if (t < 0.0f) {
        ps.L0 = __fadd_rn(ps.L0, __fmul_rn(ps.thr, S.env[0]));
        ps.L1 = __fadd_rn(ps.L1, __fmul_rn(ps.thr, S.env[1]));
        ps.L2 = __fadd_rn(ps.L2, __fmul_rn(ps.thr, S.env[2]));
        ps.rng_state = rng.state;
        return 1;
    }
    const f3 pos = add3(ps.o, scale3(t, ps.d));
    const float albedo = evaluate_albedo(S, st, na, pos);
    float Li[3];
    nee_sample(S, rng, pos, Li, n_shadow);
    const float phase_pdf = (float)(1.0 / (4.0 * VR_PI_D));
    const float w = __fmul_rn(__fmul_rn(albedo, phase_pdf), (float)(S.nl + 1));
    if (ss) {  // integrator.h:398-401: the single-scatter estimate IS the sample
        ps.L0 = __fmul_rn(w, Li[0]); ps.L1 = __fmul_rn(w, Li[1]); ps.L2 = __fmul_rn(w, Li[2]);
        ps.rng_state = rng.state;
        return 1;
    }
    const float tw = __fmul_rn(ps.thr, w);
    ps.L0 = __fadd_rn(ps.L0, __fmul_rn(tw, Li[0]));
    ps.L1 = __fadd_rn(ps.L1, __fmul_rn(tw, Li[1]));
    ps.L2 = __fadd_rn(ps.L2, __fmul_rn(tw, Li[2]));
    ps.thr = __fmul_rn(ps.thr, albedo);
    if (ps.bounce >= A.min_bounces) {  // Russian roulette, integrator.h:691-695
        const float rr = fminf(ps.thr, 0.9f);
        if (rng.uniform() > rr) { ps.rng_state = rng.state; return 1; }
        ps.thr = __fdiv_rn(ps.thr, rr);
    }
    const f3 nd = sample_uniform_direction(rng);
    ps.o = pos;
    ps.d = normalized3(nd);
    ps.bounce++;
    ps.rng_state = rng.state;
    return 0;
}

// ------------------------------------------------------------------------------------------
// persistent path kernel with lane-level path regeneration
// ------------------------------------------------------------------------------------------
template <class St> struct StoreInit;
template <int CAP> struct StoreInit<LocalStore<CAP>> {
    static __device__ __forceinline__ void init(LocalStore<CAP>&, const RenderArgs&) {}
};
template <> struct StoreInit<ArenaStore> {
    static __device__ __forceinline__ void init(ArenaStore& st, const RenderArgs& A) {
        const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
        st.base = A.arena + warp * (size_t)kArenaFields * (size_t)A.arena_cap * 32 + (threadIdx.x & 31);
        st.capacity = A.arena_cap;
    }
};

template <class St, bool kArena>
__global__ void __launch_bounds__(kBlockThreads) path_kernel(const __grid_constant__ RenderArgs A) {
    St st;
    StoreInit<St>::init(st, A);
    PathState ps;
    bool alive = false, exhausted = false;
    unsigned int n_scatter = 0, n_shadow = 0, n_hits = 0, n_over = 0;
    const unsigned int lane = threadIdx.x & 31u;
    while (true) {
        // ---- regeneration: every idle lane takes the next sample of the pass
        const unsigned int need = __ballot_sync(0xffffffffu, !alive && !exhausted);
        if (need) {
            unsigned int base = 0;
            const int leader = __ffs(need) - 1;
            if ((int)lane == leader) base = atomicAdd(A.counter, (unsigned int)__popc(need));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!alive && !exhausted) {
                unsigned int idx = base + (unsigned int)__popc(need & ((1u << lane) - 1u));
                if (idx < A.work_count) {
                    const unsigned int work = kArena ? __ldg(A.worklist + idx) : idx;
                    path_begin(A, work, ps);
                    alive = true;
                } else {
                    exhausted = true;
                }
            }
        }
        if (!__any_sync(0xffffffffu, alive)) break;
        if (alive) {
            const int r = path_vertex(A, ps, st, n_scatter, n_shadow, n_hits);
            if (r == 1) {
                A.sampleL[ps.work] = make_float4(ps.L0, ps.L1, ps.L2, (float)ps.bounce);
                alive = false;
            } else if (r < 0) {
                if (kArena) {
                    atomicAdd(A.stats + ST_ERRORS, 1ULL);  // even the large store overflowed: reported as VR_ERR_HIT_OVERFLOW
                    A.sampleL[ps.work] = make_float4(0.f, 0.f, 0.f, -1.f);
                } else {
                    overflow_push(A, ps.work);  // re-traced from scratch by the arena kernel
                    ++n_over;
                }
                alive = false;
            }
        }
    }
    // ---- statistics: warp-shuffle reduction, one atomic per warp and counter
    for (int o = 16; o > 0; o >>= 1) {
        n_scatter += __shfl_down_sync(0xffffffffu, n_scatter, o);
        n_shadow += __shfl_down_sync(0xffffffffu, n_shadow, o);
        n_hits += __shfl_down_sync(0xffffffffu, n_hits, o);
        n_over += __shfl_down_sync(0xffffffffu, n_over, o);
    }
    if (lane == 0) {
        atomicAdd(A.stats + ST_SCATTER_RAYS, (unsigned long long)n_scatter);
        atomicAdd(A.stats + ST_SHADOW_RAYS, (unsigned long long)n_shadow);
        atomicAdd(A.stats + ST_HITS, (unsigned long long)n_hits);
        if (n_over) atomicAdd(A.stats + ST_OVERFLOW, (unsigned long long)n_over);
    }
}

// ------------------------------------------------------------------------------------------
// drain of a wavefront pass.  Once every sample has been issued, the last paths need ~100 more iterations of the
// kernel chain at a few per cent pool occupancy, each bounded by its own latency (~1.5 ms per iteration): 10 % of a
// C2 step.  The paths still alive are finished here instead, one thread per path, bounce after bounce without a
// kernel boundary.  What stays exposed is the latency of the longest path (~95 bounces at ~1 ms each: one path per
// warp was measured and is no faster), so the gain is 1 % at C2 but 25-40 % for small renders (inverse mode at low spp),
// where the drain is most of a pass.  A slot whose free-flight ray is pending holds exactly this engine's path state (origin, direction,
// throughput, radiance, bounce count, the stream state BEFORE the target draw), and both engines run the same device
// functions on it, so a sample's value does not depend on where its path was finished.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlockThreads) wave_drain_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q,
                                                                   unsigned int n_live) {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int n_scatter = 0, n_shadow = 0, n_hits = 0, n_over = 0;
    if (i < n_live) {
        const unsigned int p = Q.listTs[i];
        const float4 ro = Q.ray[kSlotF4 * (size_t)p], rd = Q.ray[kSlotF4 * (size_t)p + 1];
        const float4 s0 = Q.ps01[kSlotF4 * (size_t)p];
        const uint4 s1 = reinterpret_cast<const uint4*>(Q.ps01)[kSlotF4 * (size_t)p + 1];
        PathState ps;
        ps.rng_state = (uint64_t)s1.x | ((uint64_t)s1.y << 32);
        ps.o = mk3(ro.x, ro.y, ro.z); ps.d = mk3(rd.x, rd.y, rd.z);
        ps.L0 = s0.x; ps.L1 = s0.y; ps.L2 = s0.z; ps.thr = s0.w;
        ps.work = s1.z; ps.bounce = (int)s1.w;
        int si;
        lane_xy(A, ps.work, &ps.x, &ps.y, &si);
        LocalStore<kLocalCap> st;
        int r;
        do { r = path_vertex(A, ps, st, n_scatter, n_shadow, n_hits); } while (r == 0);
        --n_scatter;   // the pending ray was counted when the wavefront published it
        if (r == 1) {
            A.sampleL[ps.work] = make_float4(ps.L0, ps.L1, ps.L2, (float)ps.bounce);
        } else {       // more overlaps than the local store holds: the whole sample is re-traced by the arena kernel
            overflow_push(A, ps.work + A.work_base);
            ++n_over;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        n_scatter += __shfl_down_sync(0xffffffffu, n_scatter, o);
        n_shadow += __shfl_down_sync(0xffffffffu, n_shadow, o);
        n_hits += __shfl_down_sync(0xffffffffu, n_hits, o);
        n_over += __shfl_down_sync(0xffffffffu, n_over, o);
    }
    if ((threadIdx.x & 31u) == 0) {
        if (n_scatter) atomicAdd(A.stats + ST_SCATTER_RAYS, (unsigned long long)(long long)(int)n_scatter);
        if (n_shadow) atomicAdd(A.stats + ST_SHADOW_RAYS, (unsigned long long)n_shadow);
        if (n_hits) atomicAdd(A.stats + ST_HITS, (unsigned long long)n_hits);
        if (n_over) atomicAdd(A.stats + ST_OVERFLOW, (unsigned long long)n_over);
    }
}
void launch_wave_drain(const RenderArgs& a, const WavePool& q, unsigned int n_live, cudaStream_t s) {
    if (n_live) wave_drain_kernel<<<(n_live + kBlockThreads - 1) / kBlockThreads, kBlockThreads, 0, s>>>(a, q, n_live);
}

constexpr size_t kWarpKernelSmem = sizeof(WarpShared) * (kBlockThreads / 32);
int path_kernel_blocks_per_sm(bool arena) {
    int nb = 0;
    if (arena) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, path_kernel<ArenaStore, true>, kBlockThreads, 0);
    else {
        cudaFuncSetAttribute(warp_path_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWarpKernelSmem);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, warp_path_kernel, kBlockThreads, kWarpKernelSmem);
    }
    return nb;
}
void launch_path_kernel(const RenderArgs& a, int blocks, bool arena, cudaStream_t s) {
    if (arena) path_kernel<ArenaStore, true><<<blocks, kBlockThreads, 0, s>>>(a);
    else warp_path_kernel<<<blocks, kBlockThreads, kWarpKernelSmem, s>>>(a);
}

// mean of a pixel's samples in sample order (integrator.h:557,702-706)
__global__ void pixel_reduce_kernel(const __grid_constant__ RenderArgs A, unsigned int pass_pixels, float* __restrict__ image) {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pass_pixels) return;
    const float4* __restrict__ src = A.sampleL + (size_t)i * A.spp;
    float r = 0.0f, g = 0.0f, b = 0.0f;
    for (int si = 0; si < A.spp; ++si) {
        const float4 v = src[si];
        r = __fadd_rn(r, v.x); g = __fadd_rn(g, v.y); b = __fadd_rn(b, v.z);
    }
    const unsigned int p = A.pixel_base + i;
    const int x = A.x0 + (int)(p % (unsigned int)A.win_w);
    const int y = A.y0 + (int)(p / (unsigned int)A.win_w) * A.part_count + A.part_rank;
    float* dst = image + 3 * ((size_t)y * A.W + x);
    const float n = (float)A.spp;
    dst[0] = __fdiv_rn(r, n); dst[1] = __fdiv_rn(g, n); dst[2] = __fdiv_rn(b, n);
}
void launch_pixel_reduce(const RenderArgs& a, unsigned int pass_pixels, float* image, cudaStream_t s) {
    if (!pass_pixels) return;
    pixel_reduce_kernel<<<(pass_pixels + 127) / 128, 128, 0, s>>>(a, pass_pixels, image);
}

// ------------------------------------------------------------------------------------------
__global__ void transmittance_kernel(const __grid_constant__ SceneDev S, int n, const float* __restrict__ o3,
                                     const float* __restrict__ d3, const float* __restrict__ tmax,
                                     float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = shadow_transmittance(S, mk3(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2]),
                                  mk3(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2]), tmax[i]);
}
void launch_transmittance(const SceneDev& S, int n, const float* o3, const float* d3, const float* tmax, float* out,
                          cudaStream_t s) {
    if (n > 0) transmittance_kernel<<<(n + 127) / 128, 128, 0, s>>>(S, n, o3, d3, tmax, out);
}

// ------------------------------------------------------------------------------------------
// known-answer kernels (one thread per record)
// ------------------------------------------------------------------------------------------
__global__ void kat_pcg_kernel(int n, const int32_t* __restrict__ xysi, uint64_t* seed, uint32_t* u32x4, float* unix4) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t s = derive_path_seed(xysi[3 * i], xysi[3 * i + 1], xysi[3 * i + 2]);
    seed[i] = s;
    Pcg r; r.seed(s, 1);
    for (int j = 0; j < 4; ++j) u32x4[4 * i + j] = r.next_u32();
    for (int j = 0; j < 4; ++j) unix4[4 * i + j] = r.uniform();
}
void launch_kat_pcg(int n, const int32_t* xysi, uint64_t* seed, uint32_t* u32x4, float* unix4, cudaStream_t s) {
    if (n > 0) kat_pcg_kernel<<<(n + 127) / 128, 128, 0, s>>>(n, xysi, seed, u32x4, unix4);
}

__global__ void kat_gaussian_kernel(const __grid_constant__ SceneDev S, const uint32_t* __restrict__ leaf_of_orig, int n,
                                    const float* o3, const float* d3, const int32_t* gidx, const float* frac,
                                    int32_t* hit, float* t0o, float* t1o, float* tau, float* mu, int32_t* ok, float* tsol) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const f3 o = mk3(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2]), d = mk3(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2]);
    const GaussRec g = load_gauss(S.gauss, leaf_of_orig[gidx[i]]);
    const RayQuad q = ray_quad(g, o, d);
    float t0 = 0.f, t1 = 0.f;
    const bool h = quad_intersect(q, &t0, &t1);
    hit[i] = h; t0o[i] = t0; t1o[i] = t1; tau[i] = 0.f; mu[i] = 0.f; ok[i] = 0; tsol[i] = 0.f;
    if (!h) return;
    const float full = depth_between(depth_const(g, q), t0, t1);
    tau[i] = full;
    const float tm = __fmul_rn(0.5f, __fadd_rn(t0, t1));
    mu[i] = gauss_mu_t(g, add3(o, scale3(tm, d)));
    float ts = 0.f;
    ok[i] = solve_t_given_tau(g, q, t0, t1, __fmul_rn(frac[i], full), &ts);
    tsol[i] = ts;
}
void launch_kat_gaussian(const SceneDev& S, const uint32_t* leaf_of_orig, int n, const float* o3, const float* d3,
                         const int32_t* gidx, const float* frac, int32_t* hit, float* t0, float* t1, float* tau,
                         float* mu, int32_t* ok, float* tsol, cudaStream_t s) {
    if (n > 0)
        kat_gaussian_kernel<<<(n + 127) / 128, 128, 0, s>>>(S, leaf_of_orig, n, o3, d3, gidx, frac, hit, t0, t1, tau, mu,
                                                              ok, tsol);
}

__global__ void kat_camera_kernel(const __grid_constant__ CameraPod cam, int n, const float* uv, float* o3, float* d3) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const RayOD r = camera_ray(cam, uv[2 * i], uv[2 * i + 1]);
    o3[3 * i] = r.o.x; o3[3 * i + 1] = r.o.y; o3[3 * i + 2] = r.o.z;
    d3[3 * i] = r.d.x; d3[3 * i + 1] = r.d.y; d3[3 * i + 2] = r.d.z;
}
void launch_kat_camera(const CameraPod& cam, int n, const float* uv, float* o3, float* d3, cudaStream_t s) {
    if (n > 0) kat_camera_kernel<<<(n + 127) / 128, 128, 0, s>>>(cam, n, uv, o3, d3);
}

// full (unpruned) event list of one ray in the walk's order
__global__ void kat_events_kernel(const __grid_constant__ SceneDev S, const float* o3, const float* d3, int cap, float* arena,
                                  float* t, int32_t* entering, uint32_t* index, int* n_events) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    ArenaStore st; st.base = arena; st.capacity = cap;
    const f3 o = mk3(o3[0], o3[1], o3[2]), d = mk3(d3[0], d3[1], d3[2]);
    const int K = gather_overlaps(S, o, d, 0.0f, false, st);
    if (K < 0) { *n_events = -1; return; }
    sort_entries(S, st, K);
    int na = 0, ie = 0, ne = 0;
    while (true) {
        bool take_enter; int se, kx; float t_evt;
        if (!next_event(S, st, K, na, ie, &take_enter, &se, &kx, &t_evt)) break;
        const int slot = take_enter ? se : (int)st.U(U_ACT, kx);
        t[ne] = t_evt; entering[ne] = take_enter; index[ne] = S.orig[st.U(U_GI, slot)]; ++ne;
        if (take_enter) { st.U(U_ACT, na) = (uint32_t)se; ++na; ++ie; }
        else { st.U(U_ACT, kx) = st.U(U_ACT, na - 1); --na; }
    }
    *n_events = ne;
}
void launch_kat_events(const SceneDev& S, const float* o3, const float* d3, int cap, float* arena, float* t,
                       int32_t* entering, uint32_t* index, int* n_events, cudaStream_t s) {
    kat_events_kernel<<<1, 32, 0, s>>>(S, o3, d3, cap, arena, t, entering, index, n_events);
}

__global__ void kat_libm_kernel(int n, const float* x, float* so, float* co, float* ao) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    so[i] = ref_sinf(x[i]); co[i] = ref_cosf(x[i]); ao[i] = ref_acosf(x[i]);
}
void launch_kat_libm(int n, const float* x, float* s_out, float* c_out, float* a_out, cudaStream_t s) {
    if (n > 0) kat_libm_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, x, s_out, c_out, a_out);
}
__global__ void kat_erfinv_kernel(int n, const double* x, double* y) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = erfinv_approx(x[i]);
}
void launch_kat_erfinv(int n, const double* x, double* y, cudaStream_t s) {
    if (n > 0) kat_erfinv_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, x, y);
}

// ------------------------------------------------------------------------------------------
// scene preparation -- Gaussian::precompute_from_covariance (gaussian.h:52-56) on the device
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float cof3(const float* a, int i, int j) {  // Eigen LU/InverseImpl.h cofactor_3x3
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return __fsub_rn(__fmul_rn(a[3 * i1 + j1], a[3 * i2 + j2]), __fmul_rn(a[3 * i1 + j2], a[3 * i2 + j1]));
}
__global__ void prepare_gaussians_kernel(int n, const float* __restrict__ mean3, const float* __restrict__ cov6,
                                         const float* __restrict__ density, const float* __restrict__ albedo,
                                         float4* __restrict__ out, float* __restrict__ box_lo, float* __restrict__ box_hi,
                                         float* __restrict__ lam) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* c = cov6 + 6 * (size_t)i;
    const float a[9] = {c[0], c[1], c[2], c[1], c[3], c[4], c[2], c[4], c[5]};
    const float c0 = cof3(a, 0, 0), c1 = cof3(a, 1, 0), c2 = cof3(a, 2, 0);
    const float det_inv = __fadd_rn(__fmul_rn(c0, a[0]), __fadd_rn(__fmul_rn(c1, a[3]), __fmul_rn(c2, a[6])));
    const float invdet = __fdiv_rn(1.0f, det_inv);
    const float m00 = __fmul_rn(c0, invdet), m01 = __fmul_rn(c1, invdet), m02 = __fmul_rn(c2, invdet);
    const float m11 = __fmul_rn(cof3(a, 1, 1), invdet), m12 = __fmul_rn(cof3(a, 2, 1), invdet);
    const float m22 = __fmul_rn(cof3(a, 2, 2), invdet);
    // Matrix3f::determinant(), Eigen LU/Determinant.h
    const float det = __fadd_rn(
        __fsub_rn(__fmul_rn(a[0], __fsub_rn(__fmul_rn(a[4], a[8]), __fmul_rn(a[5], a[7]))),
                  __fmul_rn(a[1], __fsub_rn(__fmul_rn(a[3], a[8]), __fmul_rn(a[5], a[6])))),
        __fmul_rn(a[2], __fsub_rn(__fmul_rn(a[3], a[7]), __fmul_rn(a[4], a[6]))));
    // norm = pow(2 pi, -1.5) [double] * powf(det, -0.5f)
    const float pw = (float)(1.0 / sqrt((double)det));
    const float norm = (float)__dmul_rn(0.06349363593424097, (double)pw);
    const float mx = mean3[3 * (size_t)i], my = mean3[3 * (size_t)i + 1], mz = mean3[3 * (size_t)i + 2];
    out[3 * (size_t)i] = make_float4(mx, my, mz, m00);
    out[3 * (size_t)i + 1] = make_float4(m01, m02, m11, m12);
    out[3 * (size_t)i + 2] = make_float4(m22, density[i], norm, albedo[i]);
    const float mean[3] = {mx, my, mz};
    for (int k = 0; k < 3; ++k) {  // TIGHT 3-sigma extent; inflate_boxes() (vr_bvh.cpp) adds the slack of the FP32 quadric test
        float h = 3.0f * sqrtf(fmaxf(a[4 * k], 0.0f));
        if (!(h >= 0.0f) || !isfinite(h)) h = 3.0e38f;
        box_lo[3 * (size_t)i + k] = mean[k] - h;
        box_hi[3 * (size_t)i + k] = mean[k] + h;
    }
    // Gershgorin bound of the largest eigenvalue of Sigma^-1: scale of the quadric's rounding error (inflate_boxes)
    lam[i] = fmaxf(fmaxf(fabsf(m00) + fabsf(m01) + fabsf(m02), fabsf(m01) + fabsf(m11) + fabsf(m12)), fabsf(m02) + fabsf(m12) + fabsf(m22));
}
void launch_prepare_gaussians(int n, const float* mean3, const float* cov6, const float* density, const float* albedo,
                              float4* gauss_orig_order, float* box_lo, float* box_hi, float* lam, cudaStream_t s) {
    if (n > 0)
        prepare_gaussians_kernel<<<(n + 127) / 128, 128, 0, s>>>(n, mean3, cov6, density, albedo, gauss_orig_order, box_lo,
                                                                   box_hi, lam);
}
__global__ void permute_gaussians_kernel(int n, const float4* __restrict__ src, const uint32_t* __restrict__ perm,
                                         float4* __restrict__ dst, uint32_t* __restrict__ leaf_of_orig) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t g = perm[i];
    dst[3 * (size_t)i] = src[3 * (size_t)g];
    dst[3 * (size_t)i + 1] = src[3 * (size_t)g + 1];
    dst[3 * (size_t)i + 2] = src[3 * (size_t)g + 2];
    leaf_of_orig[g] = (uint32_t)i;
}
void launch_permute_gaussians(int n, const float4* src, const uint32_t* perm, float4* dst, uint32_t* leaf_of_orig,
                              cudaStream_t s) {
    if (n > 0) permute_gaussians_kernel<<<(n + 127) / 128, 128, 0, s>>>(n, src, perm, dst, leaf_of_orig);
}

// apply_params_to_gmm_local (gmm.h:634-674): Rodrigues -> R, S = diag(exp), Sigma = ((R S) S^T) R^T
__device__ __forceinline__ float sigmoid_safe(float x) {  // gmm.h:18-26
    if (x >= 0.0f) { const float z = expf(-x); return __fdiv_rn(1.0f, __fadd_rn(1.0f, z)); }
    const float z = expf(x);
    return __fdiv_rn(z, __fadd_rn(1.0f, z));
}
__global__ void unpack_params_kernel(int n, const float* __restrict__ params, float* mean3, float* cov6, float* density,
                                     float* albedo) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* p = params + 11 * (size_t)i;
    mean3[3 * (size_t)i] = p[0]; mean3[3 * (size_t)i + 1] = p[1]; mean3[3 * (size_t)i + 2] = p[2];
    const float angle = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(p[3], p[3]), __fmul_rn(p[4], p[4])), __fmul_rn(p[5], p[5])));
    float R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (angle > 1e-12f) {  // Eigen AngleAxisf::toRotationMatrix
        const float ax = __fdiv_rn(p[3], angle), ay = __fdiv_rn(p[4], angle), az = __fdiv_rn(p[5], angle);
        const float s = ref_sinf(angle), c = ref_cosf(angle);
        const float sx = __fmul_rn(s, ax), sy = __fmul_rn(s, ay), sz = __fmul_rn(s, az);
        const float omc = __fsub_rn(1.0f, c);
        const float cx = __fmul_rn(omc, ax), cy = __fmul_rn(omc, ay), cz = __fmul_rn(omc, az);
        float tmp;
        tmp = __fmul_rn(cx, ay); R[1] = __fsub_rn(tmp, sz); R[3] = __fadd_rn(tmp, sz);
        tmp = __fmul_rn(cx, az); R[2] = __fadd_rn(tmp, sy); R[6] = __fsub_rn(tmp, sy);
        tmp = __fmul_rn(cy, az); R[5] = __fsub_rn(tmp, sx); R[7] = __fadd_rn(tmp, sx);
        R[0] = __fadd_rn(__fmul_rn(cx, ax), c); R[4] = __fadd_rn(__fmul_rn(cy, ay), c); R[8] = __fadd_rn(__fmul_rn(cz, az), c);
    }
    const float sc[3] = {expf(p[6]), expf(p[7]), expf(p[8])};
    float RSS[9], C[9];
    for (int r = 0; r < 3; ++r)
        for (int k = 0; k < 3; ++k) RSS[3 * r + k] = __fmul_rn(__fmul_rn(R[3 * r + k], sc[k]), sc[k]);
    for (int r = 0; r < 3; ++r)
        for (int k = 0; k < 3; ++k)
            C[3 * r + k] = __fadd_rn(__fadd_rn(__fmul_rn(RSS[3 * r], R[3 * k]), __fmul_rn(RSS[3 * r + 1], R[3 * k + 1])),
                                     __fmul_rn(RSS[3 * r + 2], R[3 * k + 2]));
    float* o = cov6 + 6 * (size_t)i;
    o[0] = C[0]; o[1] = C[1]; o[2] = C[2]; o[3] = C[4]; o[4] = C[5]; o[5] = C[8];
    density[i] = expf(p[9]);
    const float alb = sigmoid_safe(p[10]);
    albedo[i] = alb < 0.0f ? 0.0f : (alb > 1.0f ? 1.0f : alb);
}
void launch_unpack_params(int n, const float* params11, float* mean3, float* cov6, float* density, float* albedo,
                          cudaStream_t s) {
    if (n > 0) unpack_params_kernel<<<(n + 127) / 128, 128, 0, s>>>(n, params11, mean3, cov6, density, albedo);
}

}  // namespace vr
