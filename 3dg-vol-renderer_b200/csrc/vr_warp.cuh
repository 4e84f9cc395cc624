// vr_warp.cuh -- warp-scheduled path kernel (the main forward kernel).
//
// One lane still owns one camera path, but the WARP decides what to execute next so that every
// piece of code runs with as many lanes as possible and the hot loop stays small enough for the
// instruction caches (profiles/r01_v0: the lane-per-path megakernel ran 4.8 of 32 lanes and sat 78%
// of its stall samples in instruction fetch).  Work kinds, each run when enough lanes want it:
//
//   node rounds   lanes whose ray is inside the hierarchy do slab tests; a lane reaching a leaf only
//                 PUSHES its Gaussians (owner lane, index) into a per-warp queue in shared memory;
//   quadric batch when >= 32 candidates are queued (or nothing else can run) the warp tests 32 of
//                 them at once, one per lane whatever path that lane owns (gaussian.h:126-164).
//                 A free-flight overlap is appended to the OWNER's list in shared memory as
//                 (t0, t1, index) and may lower the owner's pruning bound (atomicMin); a shadow-ray
//                 overlap goes to a second queue;
//   depth batch   32 queued shadow overlaps at once: exact optical depth (gaussian.h:208-231) added
//                 to the owner's 64-bit FIXED-POINT sum (associative, so deterministic);
//   post A        lanes whose free-flight ray is finished: exact depth constants of their overlaps,
//                 sorted walk, distance solve, albedo, next-event sample -> shadow ray (or escape);
//   post B        lanes whose shadow ray is finished: radiance accumulation, Russian roulette,
//                 new direction -> next free-flight ray (or the path ends);
//   regen         lanes without a path fetch the next sample (warp-aggregated atomic) and generate
//                 its camera ray.
// Lanes simply wait in their state until their kind is run; nothing is lost by waiting because the
// paths are independent.  All control flow of the scheduler is warp-uniform (ballots).
#pragma once
#include "vr_internal.h"
#include "vr_pathcommon.cuh"

namespace vr {

constexpr int kSmemCap = 8;                    // free-flight overlaps per lane kept in shared memory
constexpr int kQueueCap = 32 + 32 * kMaxLeafDev;  // candidates: < 32 left over + one leaf per lane
constexpr int kHitQueueCap = 64;               // shadow overlaps: < 32 left over + one quadric batch
constexpr unsigned kPostThreshold = 16;        // run a post kind when this many lanes wait for it

struct WarpShared {
    float hit_t0[kSmemCap][32];     // [slot][lane]: conflict-free for the owner
    float hit_t1[kSmemCap][32];
    uint32_t hit_gi[kSmemCap][32];
    float ray[8][32];               // o.xyz d.xyz lim kind   (lim: scatter = needed tau, shadow = tmax)
    int count[32];                  // overlaps appended for the lane's current ray (may exceed kSmemCap = overflow)
    int pending[32];                // candidates / shadow overlaps of the lane still queued
    unsigned int t_bound[32];       // scatter rays: pruning bound, float bits (non-negative floats order as uints)
    long long tau_fixed[32];        // shadow rays: sum of clipped optical depths, 2^-40 units (signed)
    uint32_t queue[kQueueCap];      // (owner lane << 27) | leaf-order Gaussian index
    uint32_t hq_id[kHitQueueCap];   // shadow overlaps: same packing
    float hq_a[kHitQueueCap], hq_b[kHitQueueCap];  // clipped interval
};

// overlap list of the post-A code (free_flight & co. in vr_device.cuh): everything in the owner's local memory
typedef LocalStore<kSmemCap> PostStore;

enum LaneMode { M_REGEN = 0, M_TRAV, M_WAIT, M_POSTA, M_POSTB, M_IDLE };

struct LaneState {
    // path
    uint64_t rng_state;     // scatter ray: BEFORE the target draw; shadow ray: after the NEE draws
    f3 o, d;                // current ray (scatter: path segment, shadow: NEE ray from the scatter position o)
    float thr, L0, L1, L2;
    unsigned int work;
    int bounce;
    float target_tau;       // scatter ray: sampled optical depth
    float albedo, nee_dist; // shadow ray bookkeeping (nee_dist = INF for the environment sample)
    int nee_light;          // light index, -1 = environment
    // traversal
    int mode, kind, cur, sp;
    f3 idir, ood;
    float tmax;             // shadow ray limit
};

struct WarpCounters { unsigned int scatter, shadow, hits, over; };


// start traversing the lane's current ray
__device__ __forceinline__ void begin_ray(const SceneDev& S, WarpShared& w, int lane, LaneState& ls, int kind, float lim) {
    ls.kind = kind;
    w.ray[0][lane] = ls.o.x; w.ray[1][lane] = ls.o.y; w.ray[2][lane] = ls.o.z;
    w.ray[3][lane] = ls.d.x; w.ray[4][lane] = ls.d.y; w.ray[5][lane] = ls.d.z;
    w.ray[6][lane] = lim; w.ray[7][lane] = __int_as_float(kind);
    w.count[lane] = 0; w.pending[lane] = 0;
    w.t_bound[lane] = 0x7f800000u;  // +inf
    w.tau_fixed[lane] = 0LL;
    const RaySetup rs = ray_setup(ls.o, ls.d);
    ls.idir = rs.idir; ls.ood = rs.ood;
    ls.sp = 0; ls.cur = 0;
    const bool nothing = S.n == 0 || (kind == RAY_SHADOW && !(ls.tmax > 0.0f));
    ls.mode = nothing ? (kind == RAY_SCATTER ? M_POSTA : M_POSTB) : M_TRAV;
}

__device__ __forceinline__ void finish_traversal(WarpShared& w, int lane, LaneState& ls) {
    ls.mode = w.pending[lane] > 0 ? M_WAIT : (ls.kind == RAY_SCATTER ? M_POSTA : M_POSTB);
}

// quadric batch: test up to 32 queued candidates, one per lane (whole warp; nb, nH warp-uniform)
__device__ __forceinline__ void quadric_batch(const SceneDev& S, WarpShared& w, int lane, int q_begin, int nb, int& nH) {
    bool to_hq = false;
    uint32_t e = 0; float ha = 0.f, hb = 0.f;
    if (lane < nb) {
        e = w.queue[q_begin + lane];
        const int owner = (int)(e >> 27);
        const uint32_t gi = e & 0x07ffffffu;
        const f3 o = mk3(w.ray[0][owner], w.ray[1][owner], w.ray[2][owner]);
        const f3 d = mk3(w.ray[3][owner], w.ray[4][owner], w.ray[5][owner]);
        const float lim = w.ray[6][owner];
        const int kind = __float_as_int(w.ray[7][owner]);
        const GaussRec g = load_gauss(S.gauss, gi);
        const RayQuad q = ray_quad(g, o, d);
        float t0, t1;
        bool keep_pending = false;
        if (quad_intersect(q, &t0, &t1)) {
            if (kind == RAY_SCATTER) {
                const float tb = __uint_as_float(w.t_bound[owner]);
                if (!(t0 > tb)) {
                    const int slot = atomicAdd(&w.count[owner], 1);
                    if (slot < kSmemCap) {
                        w.hit_t0[slot][owner] = t0; w.hit_t1[slot][owner] = t1; w.hit_gi[slot][owner] = gi;
                        if (t1 < tb) {
                            // pruning only: a CONSERVATIVE estimate of this component's full optical depth
                            // (fast intrinsics, 1.5x margin over the FP32 slack of the exact expression)
                            const float dn = g.density * g.norm;
                            const float inv4A = __fdividef(1.0f, 4.0f * q.A);
                            const float K = dn * sqrtf(6.2831853f * inv4A) * __expf(-0.5f * (q.C0 - q.B * q.B * inv4A));
                            const float iden = rsqrtf(8.0f * q.A);
                            const float full = K * (erff_ni((q.B + 2.0f * q.A * t1) * iden) - erff_ni((q.B + 2.0f * q.A * t0) * iden));
                            if (full > 1.5f * lim) atomicMin(&w.t_bound[owner], __float_as_uint(t1));
                        }
                    }
                }
            } else if (w.tau_fixed[owner] <= tau_cutoff_fixed()) {
                ha = fmaxf(0.0f, t0); hb = fminf(lim, t1);
                to_hq = hb > ha;
                keep_pending = to_hq;   // the overlap stays pending until the depth batch has added it
            }
        }
        if (!keep_pending) atomicSub(&w.pending[owner], 1);
    }
    const unsigned m = __ballot_sync(0xffffffffu, to_hq);
    if (to_hq) {
        const int pos = nH + __popc(m & ((1u << lane) - 1u));
        w.hq_id[pos] = e; w.hq_a[pos] = ha; w.hq_b[pos] = hb;
    }
    nH += __popc(m);
}

// depth batch: exact clipped optical depth of up to 32 shadow overlaps (gmm.h:563-571)
__device__ __forceinline__ void depth_batch(const SceneDev& S, WarpShared& w, int lane, int h_begin, int nb) {
    if (lane < nb) {
        const uint32_t e = w.hq_id[h_begin + lane];
        const int owner = (int)(e >> 27);
        const uint32_t gi = e & 0x07ffffffu;
        const f3 o = mk3(w.ray[0][owner], w.ray[1][owner], w.ray[2][owner]);
        const f3 d = mk3(w.ray[3][owner], w.ray[4][owner], w.ray[5][owner]);
        const GaussRec g = load_gauss(S.gauss, gi);
        const RayQuad q = ray_quad(g, o, d);
        const DepthConst c = depth_const(g, q);
        const float F1 = erff_ni(erf_arg(c, w.hq_b[h_begin + lane]));
        const float F0 = erff_ni(erf_arg(c, w.hq_a[h_begin + lane]));
        const float tau = __fmul_rn(c.K, __fsub_rn(F1, F0));
        atomicAdd((unsigned long long*)&w.tau_fixed[owner], (unsigned long long)__double2ll_rn((double)tau * kTauFixed));
        atomicSub(&w.pending[owner], 1);
    }
}

}  // namespace vr

namespace vr {

__device__ __forceinline__ void finish_sample(const RenderArgs& A, LaneState& ls) {
    A.sampleL[ls.work] = make_float4(ls.L0, ls.L1, ls.L2, (float)ls.bounce);
    ls.mode = M_REGEN;
}
__device__ __forceinline__ void add_env(const SceneDev& S, LaneState& ls) {
    ls.L0 = __fadd_rn(ls.L0, __fmul_rn(ls.thr, S.env[0]));
    ls.L1 = __fadd_rn(ls.L1, __fmul_rn(ls.thr, S.env[1]));
    ls.L2 = __fadd_rn(ls.L2, __fmul_rn(ls.thr, S.env[2]));
}
// start the next free-flight ray from (ls.o, ls.d): draw the target on a copy of the stream (the
// reference draws it only if the ray overlaps something, integrator.h:593-600; committed in post A)
__device__ __forceinline__ void begin_scatter_ray(const SceneDev& S, WarpShared& w, int lane, LaneState& ls, WarpCounters& cn) {
    Pcg peek; peek.state = ls.rng_state; peek.inc = 3;
    ls.target_tau = -logf(__fsub_rn(1.0f, peek.uniform()));
    ++cn.scatter;
    begin_ray(S, w, lane, ls, RAY_SCATTER, ls.target_tau);
}

// regen: generate the camera ray of sample idx (integrator.h:557-570)
static __device__ __noinline__ void lane_regen(const RenderArgs& A, WarpShared& w, int lane, LaneState& ls, unsigned int idx,
                                               WarpCounters& cn) {
    if (idx >= A.work_count) { ls.mode = M_IDLE; return; }
    ls.work = idx;
    int x, y, si;
    lane_xy(A, idx, &x, &y, &si);
    Pcg rng; rng.seed(derive_path_seed(x, y, si), 1);
    const int n = A.nsq;
    const int sx = si % n, sy = si / n;
    const float u = __fdiv_rn(__fadd_rn((float)x, __fdiv_rn(__fadd_rn((float)sx, rng.uniform()), (float)n)), (float)A.W);
    const float v = __fdiv_rn(__fadd_rn((float)y, __fdiv_rn(__fadd_rn((float)sy, rng.uniform()), (float)n)), (float)A.H);
    const RayOD r = camera_ray(A.cam, u, v);
    ls.o = r.o; ls.d = r.d;
    ls.thr = 1.0f; ls.L0 = ls.L1 = ls.L2 = 0.0f;
    ls.bounce = 0;
    ls.rng_state = rng.state;
    begin_scatter_ray(A.scene, w, lane, ls, cn);
}

// post A: the free-flight ray's overlaps are complete (integrator.h:593-678)
static __device__ __noinline__ void lane_post_a(const RenderArgs& A, WarpShared& w, int lane, LaneState& ls, WarpCounters& cn) {
    const SceneDev& S = A.scene;
    const int K = w.count[lane];
    if (K > kSmemCap) {  // more overlaps than the shared-memory list holds: the whole sample is re-traced by the arena kernel
        overflow_push(A, ls.work);
        ++cn.over;
        ls.mode = M_REGEN;
        return;
    }
    if (K == 0) { add_env(S, ls); finish_sample(A, ls); return; }
    cn.hits += (unsigned int)K;
    Pcg rng; rng.state = ls.rng_state; rng.inc = 3;
    rng.next_u32();  // commit the target draw
    PostStore st;
#pragma unroll 1
    for (int s = 0; s < K; ++s) {  // exact depth constants (gaussian.h:208-231) of the gathered overlaps
        const uint32_t gi = w.hit_gi[s][lane];
        const GaussRec g = load_gauss(S.gauss, gi);
        const DepthConst c = depth_const(g, ray_quad(g, ls.o, ls.d));
        st.F(F_T0, s) = w.hit_t0[s][lane]; st.F(F_T1, s) = w.hit_t1[s][lane];
        st.F(F_A, s) = c.A; st.F(F_B, s) = c.B; st.F(F_K, s) = c.K; st.F(F_DEN, s) = c.den;
        st.U(U_GI, s) = gi;
    }
    int na = 0;
    const bool ss = A.integrator == 0;
    const float t = free_flight(S, st, K, ls.o, ls.d, ls.target_tau, A.solver, ss, &na, rng.state);
    if (A.record_bits) {  // integrator.h:616-644
        int x, y, si;
        lane_xy(A, ls.work, &x, &y, &si);
        uint32_t* row = A.record_bits + ((size_t)y * A.W + x) * (size_t)A.record_words;
        const float tlim = __fadd_rn(t, 1e-6f);
        for (int s = 0; s < K; ++s)
            if (t < 0.0f || st.F(F_T0, s) <= tlim) {
                const uint32_t orig = __ldg(S.orig + st.U(U_GI, s));
                const uint32_t bit = 1u << (orig & 31u);
                if (!(row[orig >> 5] & bit)) atomicOr(row + (orig >> 5), bit);
            }
    }
    if (t < 0.0f) { add_env(S, ls); ls.rng_state = rng.state; finish_sample(A, ls); return; }
    const f3 pos = add3(ls.o, scale3(t, ls.d));
    ls.albedo = evaluate_albedo(S, st, na, pos);
    // next-event estimation set-up, integrator.h:656-678
    const bool is_env = rng.uniform() < __fdiv_rn(1.0f, (float)(S.nl + 1));
    ++cn.shadow;
    ls.o = pos;
    if (!is_env) {
        int li = (int)__fmul_rn(rng.uniform(), (float)S.nl);
        li = min(li, S.nl - 1);
        const f3 lp = mk3(__ldg(S.lpos + 3 * li), __ldg(S.lpos + 3 * li + 1), __ldg(S.lpos + 3 * li + 2));
        const f3 to = sub3(lp, pos);
        ls.d = normalized3(normalized3(to));
        ls.nee_dist = __fsqrt_rn(sqnorm3(to));
        ls.nee_light = li;
        ls.tmax = ls.nee_dist;
        ls.rng_state = rng.state;
    } else {
        uint64_t stt = rng.state;
        ls.d = sample_direction_ni(&stt);
        ls.rng_state = stt;
        ls.nee_dist = CUDART_INF_F;
        ls.nee_light = -1;
        ls.tmax = CUDART_INF_F;
    }
    begin_ray(S, w, lane, ls, RAY_SHADOW, ls.tmax);
}

// post B: the shadow ray's optical depth is complete (integrator.h:669-699)
static __device__ __noinline__ void lane_post_b(const RenderArgs& A, WarpShared& w, int lane, LaneState& ls, WarpCounters& cn) {
    const SceneDev& S = A.scene;
    const long long tf = w.tau_fixed[lane];
    // exp(-float(sum)) with sum the double sum of the clipped depths (gmm.h:577); exactly 0 beyond the cutoff
    const float Tr = tf > tau_cutoff_fixed() ? 0.0f : expf(-(float)((double)tf / kTauFixed));
    float Li0, Li1, Li2;
    if (ls.nee_light >= 0) {
        const int li = ls.nee_light;
        const float d2 = __fmul_rn(ls.nee_dist, ls.nee_dist);
        Li0 = __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * li)), d2);
        Li1 = __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * li + 1)), d2);
        Li2 = __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * li + 2)), d2);
    } else {
        const float four_pi = (float)(4.0 * VR_PI_D);
        Li0 = __fmul_rn(__fmul_rn(Tr, S.env[0]), four_pi);
        Li1 = __fmul_rn(__fmul_rn(Tr, S.env[1]), four_pi);
        Li2 = __fmul_rn(__fmul_rn(Tr, S.env[2]), four_pi);
    }
    const float phase_pdf = (float)(1.0 / (4.0 * VR_PI_D));
    const float wgt = __fmul_rn(__fmul_rn(ls.albedo, phase_pdf), (float)(S.nl + 1));
    if (A.integrator == 0) {  // single scatter: the estimate IS the sample (integrator.h:398-401)
        ls.L0 = __fmul_rn(wgt, Li0); ls.L1 = __fmul_rn(wgt, Li1); ls.L2 = __fmul_rn(wgt, Li2);
        finish_sample(A, ls);
        return;
    }
    const float tw = __fmul_rn(ls.thr, wgt);
    ls.L0 = __fadd_rn(ls.L0, __fmul_rn(tw, Li0));
    ls.L1 = __fadd_rn(ls.L1, __fmul_rn(tw, Li1));
    ls.L2 = __fadd_rn(ls.L2, __fmul_rn(tw, Li2));
    ls.thr = __fmul_rn(ls.thr, ls.albedo);
    if (ls.bounce >= A.min_bounces) {  // Russian roulette, integrator.h:691-695
        Pcg rng; rng.state = ls.rng_state; rng.inc = 3;
        const float rr = fminf(ls.thr, 0.9f);
        const float u = rng.uniform();
        ls.rng_state = rng.state;
        if (u > rr) { finish_sample(A, ls); return; }
        ls.thr = __fdiv_rn(ls.thr, rr);
    }
    uint64_t stt = ls.rng_state;
    ls.d = sample_direction_ni(&stt);  // ls.o is the scatter position already
    ls.rng_state = stt;
    ls.bounce++;
    begin_scatter_ray(S, w, lane, ls, cn);
}

}  // namespace vr

namespace vr {

// node rounds: slab steps for lanes inside the hierarchy; leaf lanes queue their Gaussians.  Runs until
// the candidate queue holds a full batch, a post kind has reached its threshold, or no lane traverses.
__device__ __forceinline__ void traverse_rounds(const SceneDev& S, WarpShared& w, int lane, LaneState& ls,
                                                int stack[kStackDepth], int& nQ) {
    while (true) {
        bool fin = false;
        if (ls.mode == M_TRAV && ls.cur >= 0) {
            float tlimit;
            bool stop;
            if (ls.kind == RAY_SCATTER) { tlimit = __uint_as_float(w.t_bound[lane]); stop = w.count[lane] > kSmemCap; }
            else { tlimit = ls.tmax; stop = w.tau_fixed[lane] > tau_cutoff_fixed(); }
            RaySetup rs; rs.idir = ls.idir; rs.ood = ls.ood;
            bool h0, h1; float tn0, tn1; int c0, c1;
            node_slabs(S.nodes, ls.cur, rs, tlimit, &h0, &h1, &tn0, &tn1, &c0, &c1);
            if (stop) { h0 = h1 = false; ls.sp = 0; }
            if (h0 && h1) {
                const bool swap = tn1 < tn0;
                stack[ls.sp++] = swap ? c0 : c1;
                ls.cur = swap ? c1 : c0;
            } else if (h0) ls.cur = c0;
            else if (h1) ls.cur = c1;
            else if (ls.sp > 0) ls.cur = stack[--ls.sp];
            else { finish_traversal(w, lane, ls); fin = true; }
        }
        // leaves: queue (owner, Gaussian) pairs
        const bool at_leaf = ls.mode == M_TRAV && ls.cur < 0;
        if (__ballot_sync(0xffffffffu, at_leaf)) {
            uint32_t first = 0, cnt = 0;
            if (at_leaf) {
                const uint32_t code = ~(uint32_t)ls.cur;
                first = code >> 3; cnt = code & 7u;
                w.pending[lane] += (int)cnt;
            }
#pragma unroll 1
            for (uint32_t i = 0; i < (uint32_t)kMaxLeafDev; ++i) {
                const bool has = i < cnt;
                const unsigned m = __ballot_sync(0xffffffffu, has);
                if (m == 0) break;
                if (has) w.queue[nQ + __popc(m & ((1u << lane) - 1u))] = ((uint32_t)lane << 27) | (first + i);
                nQ += __popc(m);
            }
            if (at_leaf) {
                if (ls.sp > 0) ls.cur = stack[--ls.sp];
                else { finish_traversal(w, lane, ls); fin = true; }
            }
        }
        if (nQ >= 32) break;
        if (__ballot_sync(0xffffffffu, fin)) {
            if (__popc(__ballot_sync(0xffffffffu, ls.mode == M_POSTA)) >= kPostThreshold) break;
            if (__popc(__ballot_sync(0xffffffffu, ls.mode == M_POSTB)) >= kPostThreshold) break;
            if (__ballot_sync(0xffffffffu, ls.mode == M_TRAV) == 0) break;
        }
    }
    __syncwarp();
}

__global__ void __launch_bounds__(kBlockThreads) warp_path_kernel(const __grid_constant__ RenderArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    WarpShared& w = reinterpret_cast<WarpShared*>(smem_raw)[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
    const SceneDev& S = A.scene;
    int stack[kStackDepth];
    LaneState ls;
    ls.mode = M_REGEN;
    ls.kind = RAY_SCATTER; ls.cur = 0; ls.sp = 0; ls.work = 0; ls.bounce = 0;
    WarpCounters cn = {0u, 0u, 0u, 0u};
    int nQ = 0, nH = 0;  // queued candidates / shadow overlaps (warp-uniform)
    while (true) {
        const unsigned mT = __ballot_sync(0xffffffffu, ls.mode == M_TRAV);
        // ---- depth batch: exact optical depth of queued shadow overlaps
        if (nH >= 32 || (nH > 0 && mT == 0 && nQ == 0)) {
            const int nb = min(nH, 32);
            nH -= nb;
            depth_batch(S, w, lane, nH, nb);
            __syncwarp();
            if (ls.mode == M_WAIT && w.pending[lane] == 0) ls.mode = ls.kind == RAY_SCATTER ? M_POSTA : M_POSTB;
            continue;
        }
        // ---- quadric batch: a full batch, or whatever is left when no lane can traverse
        if (nQ >= 32 || (nQ > 0 && mT == 0)) {
            const int nb = min(nQ, 32);
            nQ -= nb;
            quadric_batch(S, w, lane, nQ, nb, nH);
            __syncwarp();
            if (ls.mode == M_WAIT && w.pending[lane] == 0) ls.mode = ls.kind == RAY_SCATTER ? M_POSTA : M_POSTB;
            continue;
        }
        const unsigned mA = __ballot_sync(0xffffffffu, ls.mode == M_POSTA);
        const unsigned mB = __ballot_sync(0xffffffffu, ls.mode == M_POSTB);
        const unsigned mC = __ballot_sync(0xffffffffu, ls.mode == M_REGEN);
        const unsigned nA = __popc(mA), nB = __popc(mB), nC = __popc(mC);
        const bool force = mT == 0;  // both queues are empty here, so every non-idle lane waits for a post kind
        if (nA >= kPostThreshold || (force && nA > 0 && nA >= nB && nA >= nC)) {
            if (ls.mode == M_POSTA) lane_post_a(A, w, lane, ls, cn);
            __syncwarp();
            continue;
        }
        if (nB >= kPostThreshold || (force && nB > 0 && nB >= nC)) {
            if (ls.mode == M_POSTB) lane_post_b(A, w, lane, ls, cn);
            __syncwarp();
            continue;
        }
        if (nC >= kPostThreshold || (force && nC > 0)) {
            unsigned int base = 0;
            const int leader = __ffs(mC) - 1;
            if (lane == leader) base = atomicAdd(A.counter, nC);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (ls.mode == M_REGEN) lane_regen(A, w, lane, ls, base + (unsigned int)__popc(mC & ((1u << lane) - 1u)), cn);
            __syncwarp();
            continue;
        }
        if (force) break;  // every lane is idle
        traverse_rounds(S, w, lane, ls, stack, nQ);
    }
    for (int o = 16; o > 0; o >>= 1) {
        cn.scatter += __shfl_down_sync(0xffffffffu, cn.scatter, o);
        cn.shadow += __shfl_down_sync(0xffffffffu, cn.shadow, o);
        cn.hits += __shfl_down_sync(0xffffffffu, cn.hits, o);
        cn.over += __shfl_down_sync(0xffffffffu, cn.over, o);
    }
    if (lane == 0) {
        atomicAdd(A.stats + ST_SCATTER_RAYS, (unsigned long long)cn.scatter);
        atomicAdd(A.stats + ST_SHADOW_RAYS, (unsigned long long)cn.shadow);
        atomicAdd(A.stats + ST_HITS, (unsigned long long)cn.hits);
        if (cn.over) atomicAdd(A.stats + ST_OVERFLOW, (unsigned long long)cn.over);
    }
}

}  // namespace vr
