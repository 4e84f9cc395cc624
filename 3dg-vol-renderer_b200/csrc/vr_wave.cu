// vr_wave.cu -- wavefront path engine (the default forward engine).
//
// profiles/r01_v2: even warp-scheduled, the single big path kernel (126 KB of SASS) spent half of its
// stall samples waiting for instructions and kept only ~10 lanes per warp inside the hierarchy,
// because lanes had to wait for "their" kind of post-processing.  Here the kinds are separate
// KERNELS over a pool of P path slots in global memory (sized to stay in the 126 MB L2):
//
//   wave_regen   slots without a path take the next (pixel, sample) and generate the camera ray;
//   wave_trav    every slot's current ray (free-flight or shadow) is traversed by persistent warps
//                with a shared-memory candidate queue and batched quadric / optical-depth tests;
//                a lane whose ray is done immediately FETCHES another ray of the wave, so ~all lanes
//                stay inside the hierarchy; results go to the slot: (t0, t1, index) overlap records
//                for free-flight rays, a fixed-point optical-depth sum for shadow rays;
//   wave_post_a  slots whose free-flight ray finished: exact depth constants, sorted walk, distance
//                solve, albedo, next-event sample -> shadow ray (integrator.h:593-678);
//   wave_post_b  slots whose shadow ray finished: radiance, Russian roulette, new direction
//                (integrator.h:669-699).
// Each kernel's code is small and every warp of the chip runs the same code at the same time.
// The per-sample arithmetic is the same device code as the other engines (vr_device.cuh).
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "vr_internal.h"
#include "vr_pathcommon.cuh"

namespace vr {

// ------------------------------------------------------------------------------------------
// pool layout
// ------------------------------------------------------------------------------------------
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
size_t wave_pool_bytes(int P) {
    const size_t p = (size_t)P;
    return align256(p * 16) * 5 + align256(p * 4) + align256(p * 4 * kWaveCap) * 3 + align256(p * 8) + align256(p * 4) * 5 +
           align256(kWaveCounters * 4);
}
void wave_pool_carve(WavePool* q, int P, void* base) {
    char* b = (char*)base;
    const size_t p = (size_t)P;
    auto take = [&](size_t bytes) { char* r = b; b += align256(bytes); return (void*)r; };
    q->P = P;
    q->ray_o = (float4*)take(p * 16); q->ray_d = (float4*)take(p * 16);
    q->ps0 = (float4*)take(p * 16); q->ps1 = (uint4*)take(p * 16); q->ps2 = (float4*)take(p * 16);
    q->hit_count = (int*)take(p * 4);
    q->hit_t0 = (float*)take(p * 4 * kWaveCap); q->hit_t1 = (float*)take(p * 4 * kWaveCap);
    q->hit_gi = (uint32_t*)take(p * 4 * kWaveCap);
    q->tau_fixed = (long long*)take(p * 8);
    q->listT[0] = (unsigned int*)take(p * 4); q->listT[1] = (unsigned int*)take(p * 4);
    q->listA = (unsigned int*)take(p * 4); q->listB = (unsigned int*)take(p * 4); q->listC = (unsigned int*)take(p * 4);
    q->cnt = (unsigned int*)take(kWaveCounters * 4);
}

// ------------------------------------------------------------------------------------------
// wave_trav
// ------------------------------------------------------------------------------------------
constexpr int kTravQueueCap = 32 + 32 * kMaxLeafDev;
constexpr int kTravHitQueueCap = 64;
constexpr unsigned kFetchThreshold = 8;   // fetch new rays when this many lanes are free

struct TravShared {
    float ray[7][32];               // o.xyz d.xyz lim of the lane's current ray
    int kind[32];
    unsigned int pid[32];           // pool slot of the lane's ray
    int count[32];                  // free-flight overlaps found (may exceed kWaveCap = overflow)
    int pending[32];                // queued candidates / shadow overlaps of the lane
    unsigned int t_bound[32];       // pruning bound of a free-flight ray, float bits
    long long tau[32];              // shadow ray: fixed-point optical depth
    uint32_t queue[kTravQueueCap];  // (owner lane << 27) | leaf-order Gaussian index
    uint32_t hq_id[kTravHitQueueCap];
    float hq_a[kTravHitQueueCap], hq_b[kTravHitQueueCap];
};

enum { T_FETCH = 0, T_TRAV, T_WAIT, T_DONE };

struct TravLane {
    int mode, kind, cur, sp;
    bool has_result;
    f3 idir, ood;
    float tmax;
};

__device__ __forceinline__ void trav_quadric_batch(const SceneDev& S, const WavePool& Q, TravShared& w, int lane, int q_begin,
                                                   int nb, int& nH) {
    bool to_hq = false;
    uint32_t e = 0; float ha = 0.f, hb = 0.f;
    if (lane < nb) {
        e = w.queue[q_begin + lane];
        const int owner = (int)(e >> 27);
        const uint32_t gi = e & 0x07ffffffu;
        const f3 o = mk3(w.ray[0][owner], w.ray[1][owner], w.ray[2][owner]);
        const f3 d = mk3(w.ray[3][owner], w.ray[4][owner], w.ray[5][owner]);
        const float lim = w.ray[6][owner];
        const GaussRec g = load_gauss(S.gauss, gi);
        const RayQuad q = ray_quad(g, o, d);
        float t0, t1;
        bool keep_pending = false;
        if (quad_intersect(q, &t0, &t1)) {
            if (w.kind[owner] == RAY_SCATTER) {
                const float tb = __uint_as_float(w.t_bound[owner]);
                if (!(t0 > tb)) {
                    const int slot = atomicAdd(&w.count[owner], 1);
                    if (slot < kWaveCap) {
                        const size_t at = (size_t)slot * Q.P + w.pid[owner];
                        Q.hit_t0[at] = t0; Q.hit_t1[at] = t1; Q.hit_gi[at] = gi;
                        if (t1 < tb) {
                            // pruning only: CONSERVATIVE estimate of this component's full optical depth
                            // (fast intrinsics; 1.5x margin covers the FP32 slack of the exact expression)
                            const float dn = g.density * g.norm;
                            const float inv4A = __fdividef(1.0f, 4.0f * q.A);
                            const float K = dn * sqrtf(6.2831853f * inv4A) * __expf(-0.5f * (q.C0 - q.B * q.B * inv4A));
                            const float iden = rsqrtf(8.0f * q.A);
                            const float full = K * (erff_ni((q.B + 2.0f * q.A * t1) * iden) - erff_ni((q.B + 2.0f * q.A * t0) * iden));
                            if (full > 1.5f * lim) atomicMin(&w.t_bound[owner], __float_as_uint(t1));
                        }
                    }
                }
            } else if (w.tau[owner] <= tau_cutoff_fixed()) {
                ha = fmaxf(0.0f, t0); hb = fminf(lim, t1);
                to_hq = hb > ha;
                keep_pending = to_hq;   // stays pending until the depth batch has added it
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

// exact clipped optical depth of up to 32 shadow overlaps (gmm.h:563-571, gaussian.h:208-231)
__device__ __forceinline__ void trav_depth_batch(const SceneDev& S, TravShared& w, int lane, int h_begin, int nb) {
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
        atomicAdd((unsigned long long*)&w.tau[owner], (unsigned long long)__double2ll_rn((double)tau * kTauFixed));
        atomicSub(&w.pending[owner], 1);
    }
}

__device__ __forceinline__ void trav_finish(TravShared& w, int lane, TravLane& tl) {
    if (w.pending[lane] > 0) tl.mode = T_WAIT;
    else { tl.mode = T_FETCH; tl.has_result = true; }
}

// node rounds: slab steps for lanes inside the hierarchy; leaf lanes queue their Gaussians.  Runs until
// a full candidate batch is queued, enough lanes are free to fetch, or no lane traverses.
__device__ __forceinline__ void trav_rounds(const SceneDev& S, TravShared& w, int lane, TravLane& tl, int stack[kStackDepth],
                                            int& nQ, bool can_fetch) {
    while (true) {
        bool fin = false;
        if (tl.mode == T_TRAV && tl.cur >= 0) {
            float tlimit;
            bool stop;
            if (tl.kind == RAY_SCATTER) { tlimit = __uint_as_float(w.t_bound[lane]); stop = w.count[lane] > kWaveCap; }
            else { tlimit = tl.tmax; stop = w.tau[lane] > tau_cutoff_fixed(); }
            RaySetup rs; rs.idir = tl.idir; rs.ood = tl.ood;
            bool h0, h1; float tn0, tn1; int c0, c1;
            node_slabs(S.nodes, tl.cur, rs, tlimit, &h0, &h1, &tn0, &tn1, &c0, &c1);
            if (stop) { h0 = h1 = false; tl.sp = 0; }
            if (h0 && h1) {
                const bool swap = tn1 < tn0;
                stack[tl.sp++] = swap ? c0 : c1;
                tl.cur = swap ? c1 : c0;
            } else if (h0) tl.cur = c0;
            else if (h1) tl.cur = c1;
            else if (tl.sp > 0) tl.cur = stack[--tl.sp];
            else { trav_finish(w, lane, tl); fin = true; }
        }
        const bool at_leaf = tl.mode == T_TRAV && tl.cur < 0;
        if (__ballot_sync(0xffffffffu, at_leaf)) {
            uint32_t first = 0, cnt = 0;
            if (at_leaf) {
                const uint32_t code = ~(uint32_t)tl.cur;
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
                if (tl.sp > 0) tl.cur = stack[--tl.sp];
                else { trav_finish(w, lane, tl); fin = true; }
            }
        }
        if (nQ >= 32) break;
        if (__ballot_sync(0xffffffffu, fin)) {
            if (can_fetch && __popc(__ballot_sync(0xffffffffu, tl.mode == T_FETCH)) >= kFetchThreshold) break;
            if (__ballot_sync(0xffffffffu, tl.mode == T_TRAV) == 0) break;
        }
    }
    __syncwarp();
}

__global__ void __launch_bounds__(kBlockThreads) wave_trav_kernel(const __grid_constant__ SceneDev S, const __grid_constant__ WavePool Q,
                                                                  int cur_list) {
    __shared__ TravShared wsh[kBlockThreads / 32];
    TravShared& w = wsh[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
    const unsigned int nT = Q.cnt[WC_NT0 + cur_list];
    const unsigned int* __restrict__ list = Q.listT[cur_list];
    if (blockIdx.x == 0 && threadIdx.x == 0) Q.cnt[WC_NC] = 0;  // wave_regen (previous kernel) has consumed listC
    int stack[kStackDepth];
    TravLane tl;
    tl.mode = T_FETCH; tl.kind = RAY_SCATTER; tl.cur = 0; tl.sp = 0; tl.has_result = false; tl.tmax = 0.f;
    int nQ = 0, nH = 0;
    bool exhausted = false;  // warp-uniform: the wave has no more rays to hand out
    while (true) {
        const unsigned mT = __ballot_sync(0xffffffffu, tl.mode == T_TRAV);
        if (nH >= 32 || (nH > 0 && mT == 0 && nQ == 0)) {
            const int nb = min(nH, 32);
            nH -= nb;
            trav_depth_batch(S, w, lane, nH, nb);
            __syncwarp();
            if (tl.mode == T_WAIT && w.pending[lane] == 0) { tl.mode = T_FETCH; tl.has_result = true; }
            continue;
        }
        if (nQ >= 32 || (nQ > 0 && mT == 0)) {
            const int nb = min(nQ, 32);
            nQ -= nb;
            trav_quadric_batch(S, Q, w, lane, nQ, nb, nH);
            __syncwarp();
            if (tl.mode == T_WAIT && w.pending[lane] == 0) { tl.mode = T_FETCH; tl.has_result = true; }
            continue;
        }
        const unsigned mF = __ballot_sync(0xffffffffu, tl.mode == T_FETCH);
        if (mF && (__popc(mF) >= kFetchThreshold || mT == 0)) {
            // ---- flush finished rays: result to the slot, slot to the list of its post kernel
            const bool flush = tl.mode == T_FETCH && tl.has_result;
            const unsigned int p = w.pid[lane];
            const bool fa = flush && tl.kind == RAY_SCATTER, fb = flush && tl.kind == RAY_SHADOW;
            const unsigned ma = __ballot_sync(0xffffffffu, fa), mb = __ballot_sync(0xffffffffu, fb);
            unsigned int base_a = 0, base_b = 0;
            if (ma && lane == __ffs(ma) - 1) base_a = atomicAdd(&Q.cnt[WC_NA], (unsigned int)__popc(ma));
            if (mb && lane == __ffs(mb) - 1) base_b = atomicAdd(&Q.cnt[WC_NB], (unsigned int)__popc(mb));
            if (ma) base_a = __shfl_sync(0xffffffffu, base_a, __ffs(ma) - 1);
            if (mb) base_b = __shfl_sync(0xffffffffu, base_b, __ffs(mb) - 1);
            if (fa) { Q.hit_count[p] = w.count[lane]; Q.listA[base_a + __popc(ma & ((1u << lane) - 1u))] = p; }
            if (fb) { Q.tau_fixed[p] = w.tau[lane]; Q.listB[base_b + __popc(mb & ((1u << lane) - 1u))] = p; }
            tl.has_result = false;
            // ---- fetch the next rays of the wave
            if (!exhausted) {
                unsigned int base = 0;
                if (lane == __ffs(mF) - 1) base = atomicAdd(&Q.cnt[WC_CURSOR], (unsigned int)__popc(mF));
                base = __shfl_sync(0xffffffffu, base, __ffs(mF) - 1);
                if (tl.mode == T_FETCH) {
                    const unsigned int idx = base + (unsigned int)__popc(mF & ((1u << lane) - 1u));
                    if (idx < nT) {
                        const unsigned int np = __ldg(list + idx);
                        const float4 ro = Q.ray_o[np], rd = Q.ray_d[np];
                        const int kind = __float_as_int(rd.w);
                        w.ray[0][lane] = ro.x; w.ray[1][lane] = ro.y; w.ray[2][lane] = ro.z;
                        w.ray[3][lane] = rd.x; w.ray[4][lane] = rd.y; w.ray[5][lane] = rd.z;
                        w.ray[6][lane] = ro.w;
                        w.kind[lane] = kind; w.pid[lane] = np;
                        w.count[lane] = 0; w.pending[lane] = 0; w.t_bound[lane] = 0x7f800000u; w.tau[lane] = 0LL;
                        const RaySetup rs = ray_setup(mk3(ro.x, ro.y, ro.z), mk3(rd.x, rd.y, rd.z));
                        tl.idir = rs.idir; tl.ood = rs.ood;
                        tl.kind = kind; tl.tmax = ro.w; tl.sp = 0; tl.cur = 0;
                        const bool nothing = S.n == 0 || (kind == RAY_SHADOW && !(ro.w > 0.0f));
                        if (nothing) tl.has_result = true;   // stays in T_FETCH: flushed at the next fetch step
                        else tl.mode = T_TRAV;
                    } else {
                        tl.mode = T_DONE;
                    }
                }
                if (base + (unsigned int)__popc(mF) >= nT) exhausted = true;
            } else if (tl.mode == T_FETCH) {
                tl.mode = T_DONE;
            }
            __syncwarp();
            continue;
        }
        if (mT == 0) break;  // queues empty, nothing to flush or fetch: every lane is done
        trav_rounds(S, w, lane, tl, stack, nQ, !exhausted);
    }
}

// ------------------------------------------------------------------------------------------
// post kernels: one thread per listed slot, grid-stride in whole warps so that the aggregated
// list appends at the end of each round are executed by converged warps
// ------------------------------------------------------------------------------------------
struct WaveCounters { unsigned int scatter, shadow, hits, over; };

__device__ __forceinline__ void wave_append(unsigned int* list, unsigned int* counter, bool pred, unsigned int value) {
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    unsigned int base = 0;
    if (lane == leader) base = atomicAdd(counter, (unsigned int)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pred) list[base + __popc(m & ((1u << lane) - 1u))] = value;
}
__device__ __forceinline__ void wave_flush_counters(const RenderArgs& A, WaveCounters cn) {
    for (int o = 16; o > 0; o >>= 1) {
        cn.scatter += __shfl_down_sync(0xffffffffu, cn.scatter, o);
        cn.shadow += __shfl_down_sync(0xffffffffu, cn.shadow, o);
        cn.hits += __shfl_down_sync(0xffffffffu, cn.hits, o);
        cn.over += __shfl_down_sync(0xffffffffu, cn.over, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (cn.scatter) atomicAdd(A.stats + ST_SCATTER_RAYS, (unsigned long long)cn.scatter);
        if (cn.shadow) atomicAdd(A.stats + ST_SHADOW_RAYS, (unsigned long long)cn.shadow);
        if (cn.hits) atomicAdd(A.stats + ST_HITS, (unsigned long long)cn.hits);
        if (cn.over) atomicAdd(A.stats + ST_OVERFLOW, (unsigned long long)cn.over);
    }
}
enum { DEST_NONE = 0, DEST_TRAV = 1, DEST_REGEN = 2 };

// draw the free-flight target on a copy of the stream (committed in post A only if the ray overlaps
// something, integrator.h:593-600) and publish the slot's next free-flight ray
__device__ __forceinline__ void publish_scatter_ray(const WavePool& Q, unsigned int p, f3 o, f3 d, uint64_t rng_state, float* target) {
    Pcg peek; peek.state = rng_state; peek.inc = 3;
    const float t = -logf(__fsub_rn(1.0f, peek.uniform()));
    *target = t;
    Q.ray_o[p] = make_float4(o.x, o.y, o.z, t);
    Q.ray_d[p] = make_float4(d.x, d.y, d.z, __int_as_float(RAY_SCATTER));
}

// wave_regen: integrator.h:557-570
__global__ void __launch_bounds__(256) wave_regen_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q,
                                                          int cur_list) {
    const unsigned int nC = Q.cnt[WC_NC];
    if (blockIdx.x == 0 && threadIdx.x == 0) {  // lists consumed by the previous iteration's post kernels
        Q.cnt[WC_NT0 + (cur_list ^ 1)] = 0; Q.cnt[WC_NA] = 0; Q.cnt[WC_NB] = 0;
    }
    WaveCounters cn = {0u, 0u, 0u, 0u};
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int rounds = (nC + stride - 1) / stride;
    for (unsigned int r = 0; r < rounds; ++r) {
        const unsigned int i = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
        const bool act = i < nC;
        // sample ids: one aggregated atomic per warp
        const unsigned m = __ballot_sync(0xffffffffu, act);
        unsigned int base = 0;
        const int lane = threadIdx.x & 31;
        if (m && lane == __ffs(m) - 1) base = atomicAdd(&Q.cnt[WC_NEXT_SAMPLE], (unsigned int)__popc(m));
        if (m) base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
        const unsigned int idx = base + (unsigned int)__popc(m & ((1u << lane) - 1u));
        bool to_trav = false;
        unsigned int p = 0;
        if (act && idx < A.work_count) {
            p = Q.listC[i];
            int x, y, si;
            lane_xy(A, idx, &x, &y, &si);
            Pcg rng; rng.seed(derive_path_seed(x, y, si), 1);
            const int n = A.nsq;
            const int sx = si % n, sy = si / n;
            const float u = __fdiv_rn(__fadd_rn((float)x, __fdiv_rn(__fadd_rn((float)sx, rng.uniform()), (float)n)), (float)A.W);
            const float v = __fdiv_rn(__fadd_rn((float)y, __fdiv_rn(__fadd_rn((float)sy, rng.uniform()), (float)n)), (float)A.H);
            const RayOD ray = camera_ray(A.cam, u, v);
            float target;
            publish_scatter_ray(Q, p, ray.o, ray.d, rng.state, &target);
            Q.ps0[p] = make_float4(0.f, 0.f, 0.f, 1.0f);
            Q.ps1[p] = make_uint4((unsigned int)rng.state, (unsigned int)(rng.state >> 32), idx, 0u);
            Q.ps2[p] = make_float4(target, 0.f, 0.f, 0.f);
            ++cn.scatter;
            to_trav = true;
        }
        wave_append(Q.listT[cur_list], &Q.cnt[WC_NT0 + cur_list], to_trav, p);
    }
    wave_flush_counters(A, cn);
}

// wave_post_a: integrator.h:593-678
__global__ void __launch_bounds__(128) wave_post_a_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q,
                                                           int nxt_list) {
    const SceneDev& S = A.scene;
    const unsigned int nA = Q.cnt[WC_NA];
    WaveCounters cn = {0u, 0u, 0u, 0u};
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int rounds = (nA + stride - 1) / stride;
    for (unsigned int r = 0; r < rounds; ++r) {
        const unsigned int i = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
        int dest = DEST_NONE;
        unsigned int p = 0;
        if (i < nA) {
            p = Q.listA[i];
            const int K = Q.hit_count[p];
            float4 s0 = Q.ps0[p];
            uint4 s1 = Q.ps1[p];
            const float4 ro = Q.ray_o[p], rd = Q.ray_d[p];
            const f3 o = mk3(ro.x, ro.y, ro.z), d = mk3(rd.x, rd.y, rd.z);
            const float target_tau = ro.w;
            dest = DEST_REGEN;
            if (K > kWaveCap) {  // more overlaps than a slot holds: the whole sample is re-traced by the arena kernel
                A.overflow_list[atomicAdd(A.overflow_count, 1u)] = s1.z;
                ++cn.over;
            } else if (K == 0) {  // nothing on the ray: environment, path ends, target draw not committed
                s0.x = __fadd_rn(s0.x, __fmul_rn(s0.w, S.env[0]));
                s0.y = __fadd_rn(s0.y, __fmul_rn(s0.w, S.env[1]));
                s0.z = __fadd_rn(s0.z, __fmul_rn(s0.w, S.env[2]));
                A.sampleL[s1.z] = make_float4(s0.x, s0.y, s0.z, (float)s1.w);
            } else {
                cn.hits += (unsigned int)K;
                Pcg rng; rng.state = (uint64_t)s1.x | ((uint64_t)s1.y << 32); rng.inc = 3;
                rng.next_u32();  // commit the target draw
                LocalStore<kWaveCap> st;
#pragma unroll 1
                for (int s = 0; s < K; ++s) {  // exact depth constants (gaussian.h:208-231) of the gathered overlaps
                    const size_t at = (size_t)s * Q.P + p;
                    const uint32_t gi = Q.hit_gi[at];
                    const GaussRec g = load_gauss(S.gauss, gi);
                    const DepthConst c = depth_const(g, ray_quad(g, o, d));
                    st.F(F_T0, s) = Q.hit_t0[at]; st.F(F_T1, s) = Q.hit_t1[at];
                    st.F(F_A, s) = c.A; st.F(F_B, s) = c.B; st.F(F_K, s) = c.K; st.F(F_DEN, s) = c.den;
                    st.U(U_GI, s) = gi;
                }
                int na = 0;
                const float t = free_flight(S, st, K, o, d, target_tau, A.solver, A.integrator == 0, &na);
                if (A.record_bits) {  // integrator.h:616-644
                    int x, y, si;
                    lane_xy(A, s1.z, &x, &y, &si);
                    uint32_t* row = A.record_bits + ((size_t)y * A.W + x) * (size_t)A.record_words;
                    const float tlim = __fadd_rn(t, 1e-6f);
                    for (int s = 0; s < K; ++s)
                        if (t < 0.0f || st.F(F_T0, s) <= tlim) {
                            const uint32_t orig = __ldg(S.orig + st.U(U_GI, s));
                            const uint32_t bit = 1u << (orig & 31u);
                            if (!(row[orig >> 5] & bit)) atomicOr(row + (orig >> 5), bit);
                        }
                }
                if (t < 0.0f) {
                    s0.x = __fadd_rn(s0.x, __fmul_rn(s0.w, S.env[0]));
                    s0.y = __fadd_rn(s0.y, __fmul_rn(s0.w, S.env[1]));
                    s0.z = __fadd_rn(s0.z, __fmul_rn(s0.w, S.env[2]));
                    A.sampleL[s1.z] = make_float4(s0.x, s0.y, s0.z, (float)s1.w);
                } else {
                    const f3 pos = add3(o, scale3(t, d));
                    const float albedo = evaluate_albedo(S, st, na, pos);
                    const bool is_env = rng.uniform() < __fdiv_rn(1.0f, (float)(S.nl + 1));
                    ++cn.shadow;
                    f3 wi; float dist; int light;
                    uint64_t stt;
                    if (!is_env) {
                        int li = (int)__fmul_rn(rng.uniform(), (float)S.nl);
                        li = min(li, S.nl - 1);
                        const f3 lp = mk3(__ldg(S.lpos + 3 * li), __ldg(S.lpos + 3 * li + 1), __ldg(S.lpos + 3 * li + 2));
                        const f3 to = sub3(lp, pos);
                        wi = normalized3(normalized3(to));
                        dist = __fsqrt_rn(sqnorm3(to));
                        light = li;
                        stt = rng.state;
                    } else {
                        stt = rng.state;
                        wi = sample_direction_ni(&stt);
                        dist = CUDART_INF_F;
                        light = -1;
                    }
                    Q.ray_o[p] = make_float4(pos.x, pos.y, pos.z, dist);
                    Q.ray_d[p] = make_float4(wi.x, wi.y, wi.z, __int_as_float(RAY_SHADOW));
                    Q.ps1[p] = make_uint4((unsigned int)stt, (unsigned int)(stt >> 32), s1.z, s1.w);
                    Q.ps2[p] = make_float4(target_tau, albedo, dist, __int_as_float(light));
                    dest = DEST_TRAV;
                }
            }
        }
        wave_append(Q.listT[nxt_list], &Q.cnt[WC_NT0 + nxt_list], dest == DEST_TRAV, p);
        wave_append(Q.listC, &Q.cnt[WC_NC], dest == DEST_REGEN, p);
    }
    wave_flush_counters(A, cn);
}

// wave_post_b: integrator.h:669-699
__global__ void __launch_bounds__(256) wave_post_b_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q,
                                                           int nxt_list) {
    const SceneDev& S = A.scene;
    const unsigned int nB = Q.cnt[WC_NB];
    if (blockIdx.x == 0 && threadIdx.x == 0) Q.cnt[WC_CURSOR] = 0;  // wave_trav of this iteration is over
    WaveCounters cn = {0u, 0u, 0u, 0u};
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int rounds = (nB + stride - 1) / stride;
    for (unsigned int r = 0; r < rounds; ++r) {
        const unsigned int i = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
        int dest = DEST_NONE;
        unsigned int p = 0;
        if (i < nB) {
            p = Q.listB[i];
            float4 s0 = Q.ps0[p];
            uint4 s1 = Q.ps1[p];
            const float4 s2 = Q.ps2[p];
            const float albedo = s2.y, nee_dist = s2.z;
            const int light = __float_as_int(s2.w);
            const long long tf = Q.tau_fixed[p];
            // exp(-float(sum)), sum = double sum of the clipped depths (gmm.h:577); exactly 0 beyond the cutoff
            const float Tr = tf > tau_cutoff_fixed() ? 0.0f : expf(-(float)((double)tf / kTauFixed));
            float Li0, Li1, Li2;
            if (light >= 0) {
                const float d2 = __fmul_rn(nee_dist, nee_dist);
                Li0 = __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * light)), d2);
                Li1 = __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * light + 1)), d2);
                Li2 = __fdiv_rn(__fmul_rn(Tr, __ldg(S.lint + 3 * light + 2)), d2);
            } else {
                const float four_pi = (float)(4.0 * VR_PI_D);
                Li0 = __fmul_rn(__fmul_rn(Tr, S.env[0]), four_pi);
                Li1 = __fmul_rn(__fmul_rn(Tr, S.env[1]), four_pi);
                Li2 = __fmul_rn(__fmul_rn(Tr, S.env[2]), four_pi);
            }
            const float phase_pdf = (float)(1.0 / (4.0 * VR_PI_D));
            const float wgt = __fmul_rn(__fmul_rn(albedo, phase_pdf), (float)(S.nl + 1));
            dest = DEST_REGEN;
            if (A.integrator == 0) {  // single scatter: the estimate IS the sample (integrator.h:398-401)
                A.sampleL[s1.z] = make_float4(__fmul_rn(wgt, Li0), __fmul_rn(wgt, Li1), __fmul_rn(wgt, Li2), 0.f);
            } else {
                const float tw = __fmul_rn(s0.w, wgt);
                s0.x = __fadd_rn(s0.x, __fmul_rn(tw, Li0));
                s0.y = __fadd_rn(s0.y, __fmul_rn(tw, Li1));
                s0.z = __fadd_rn(s0.z, __fmul_rn(tw, Li2));
                s0.w = __fmul_rn(s0.w, albedo);
                uint64_t stt = (uint64_t)s1.x | ((uint64_t)s1.y << 32);
                bool alive = true;
                if ((int)s1.w >= A.min_bounces) {  // Russian roulette, integrator.h:691-695
                    Pcg rng; rng.state = stt; rng.inc = 3;
                    const float rr = fminf(s0.w, 0.9f);
                    const float u = rng.uniform();
                    stt = rng.state;
                    if (u > rr) alive = false;
                    else s0.w = __fdiv_rn(s0.w, rr);
                }
                if (!alive) {
                    A.sampleL[s1.z] = make_float4(s0.x, s0.y, s0.z, (float)s1.w);
                } else {
                    const f3 nd = sample_direction_ni(&stt);
                    const float4 ro = Q.ray_o[p];  // the scatter position is the shadow ray's origin
                    float target;
                    publish_scatter_ray(Q, p, mk3(ro.x, ro.y, ro.z), nd, stt, &target);
                    Q.ps0[p] = s0;
                    Q.ps1[p] = make_uint4((unsigned int)stt, (unsigned int)(stt >> 32), s1.z, s1.w + 1u);
                    ++cn.scatter;
                    dest = DEST_TRAV;
                }
            }
        }
        wave_append(Q.listT[nxt_list], &Q.cnt[WC_NT0 + nxt_list], dest == DEST_TRAV, p);
        wave_append(Q.listC, &Q.cnt[WC_NC], dest == DEST_REGEN, p);
    }
    wave_flush_counters(A, cn);
}

__global__ void wave_init_kernel(const __grid_constant__ WavePool Q, unsigned int n_slots) {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_slots) Q.listC[i] = i;
    if (i < kWaveCounters) Q.cnt[i] = (i == WC_NC) ? n_slots : 0u;
}

// ------------------------------------------------------------------------------------------
// host driver of one pass
// ------------------------------------------------------------------------------------------
int run_wavefront_pass(const RenderArgs& a, const WavePool& q, int num_sms, unsigned int* h_cnt, cudaStream_t s,
                       unsigned int* launches) {
    static int trav_blocks_per_sm = 0;
    if (!trav_blocks_per_sm) {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&trav_blocks_per_sm, wave_trav_kernel, kBlockThreads, 0);
        if (trav_blocks_per_sm < 1) trav_blocks_per_sm = 1;
    }
    const unsigned int slots = (unsigned int)std::min<unsigned long long>((unsigned long long)q.P, a.work_count);
    wave_init_kernel<<<(std::max(slots, (unsigned int)kWaveCounters) + 255) / 256, 256, 0, s>>>(q, slots);
    ++*launches;
    const int grid_trav = num_sms * trav_blocks_per_sm;
    const int grid_post = std::max(1, (int)std::min<unsigned int>((slots + 127) / 128, (unsigned int)num_sms * 16));
    const int grid_post256 = std::max(1, (grid_post + 1) / 2);
    int cur = 0, iters = 0;
    const int kBatch = 16;
    cudaEvent_t ev;
    if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) return -1;
    const char* prof_env = getenv("VR_WAVE_PROFILE");
    const bool prof = prof_env && prof_env[0] == '1';
    cudaEvent_t pe[5];
    double pt[4] = {0, 0, 0, 0};
    if (prof) for (auto& e : pe) cudaEventCreate(&e);
    while (true) {
        for (int b = 0; b < kBatch; ++b) {
            if (prof) cudaEventRecord(pe[0], s);
            wave_regen_kernel<<<grid_post256, 256, 0, s>>>(a, q, cur);
            if (prof) cudaEventRecord(pe[1], s);
            wave_trav_kernel<<<grid_trav, kBlockThreads, 0, s>>>(a.scene, q, cur);
            if (prof) cudaEventRecord(pe[2], s);
            wave_post_a_kernel<<<grid_post, 128, 0, s>>>(a, q, cur ^ 1);
            if (prof) cudaEventRecord(pe[3], s);
            wave_post_b_kernel<<<grid_post256, 256, 0, s>>>(a, q, cur ^ 1);
            if (prof) {
                cudaEventRecord(pe[4], s);
                cudaEventSynchronize(pe[4]);
                for (int k = 0; k < 4; ++k) { float ms = 0.f; cudaEventElapsedTime(&ms, pe[k], pe[k + 1]); pt[k] += ms; }
            }
            *launches += 4;
            cur ^= 1;
            ++iters;
        }
        if (cudaMemcpyAsync(h_cnt, q.cnt, sizeof(unsigned int) * kWaveCounters, cudaMemcpyDeviceToHost, s) != cudaSuccess ||
            cudaEventRecord(ev, s) != cudaSuccess || cudaEventSynchronize(ev) != cudaSuccess) {
            cudaEventDestroy(ev);
            return -1;
        }
        // after an iteration: rays left = listT[cur] (filled by the post kernels) + slots waiting for regeneration
        const bool more_samples = h_cnt[WC_NEXT_SAMPLE] < a.work_count;
        if (h_cnt[WC_NT0 + cur] == 0 && (h_cnt[WC_NC] == 0 || !more_samples)) break;
        if (iters > (1 << 24)) break;
    }
    cudaEventDestroy(ev);
    if (prof) {
        fprintf(stderr, "[wave] samples %u iters %d  regen %.3f ms  trav %.3f ms  post_a %.3f ms  post_b %.3f ms\n", a.work_count, iters,
                pt[0], pt[1], pt[2], pt[3]);
        for (auto& e : pe) cudaEventDestroy(e);
    }
    return iters;
}

}  // namespace vr
