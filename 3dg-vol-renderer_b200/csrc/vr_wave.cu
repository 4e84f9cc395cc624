// vr_wave.cu -- wavefront path engine (the default forward engine).
//
// profiles/r01_v2: even warp-scheduled, the single big path kernel (126 KB of SASS) spent half of its
// stall samples waiting for instructions and kept only ~10 lanes per warp inside the hierarchy,
// because lanes had to wait for "their" kind of post-processing.  Here the kinds are separate
// KERNELS over a pool of P path slots in global memory (sized to stay in the 126 MB L2):
//
//   wave_regen   slots without a path take the next (pixel, sample) and generate the camera ray;
//   wave_trav<K> every slot's current ray of kind K (free-flight / shadow) is traversed by persistent warps
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
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "vr_context.h"
#include "vr_internal.h"
#include "vr_pathcommon.cuh"

namespace vr {

// ------------------------------------------------------------------------------------------
// pool layout
// ------------------------------------------------------------------------------------------
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
size_t wave_pool_bytes(int P) {
    const size_t p = (size_t)P;
    return align256(p * 32) * 4 + align256(p * 16 * (kWaveCap - 2)) + align256(p * 4) * 5 + align256(p * 4 * (kPostClasses + 1)) +
           align256(kWaveCounters * 4);
}
void wave_pool_carve(WavePool* q, int P, void* base) {
    char* b = (char*)base;
    const size_t p = (size_t)P;
    auto take = [&](size_t bytes) { char* r = b; b += align256(bytes); return (void*)r; };
    q->P = P;
    // one 128-byte record per slot (kSlotF4 float4s): ray | ps2 | ps01 | hit01, see vr_internal.h
    float4* slab = (float4*)take(p * 128);
    q->ray = slab; q->ps2 = slab + 2; q->ps01 = slab + 4; q->hit01 = (uint4*)(slab + 6);
    q->hitx = (uint4*)take(p * 16 * (kWaveCap - 2));
    q->listTs = (unsigned int*)take(p * 4); q->listTh = (unsigned int*)take(p * 4); q->listTl = (unsigned int*)take(p * 4);
    q->listA = (unsigned int*)take(p * 4 * (kPostClasses + 1)); q->listB = (unsigned int*)take(p * 4); q->listC = (unsigned int*)take(p * 4);
    q->cnt = (unsigned int*)take(kWaveCounters * 4);
}

// ------------------------------------------------------------------------------------------
// list / counter helpers
// ------------------------------------------------------------------------------------------
struct WaveCounters { unsigned int scatter, shadow, hits, over, nodes, quads, erf, mu, sa, it; };

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
#ifdef VR_WORK_COUNTERS
        cn.nodes += __shfl_down_sync(0xffffffffu, cn.nodes, o);
        cn.quads += __shfl_down_sync(0xffffffffu, cn.quads, o);
        cn.erf += __shfl_down_sync(0xffffffffu, cn.erf, o);
        cn.mu += __shfl_down_sync(0xffffffffu, cn.mu, o);
        cn.sa += __shfl_down_sync(0xffffffffu, cn.sa, o);
        cn.it += __shfl_down_sync(0xffffffffu, cn.it, o);
#endif
    }
    if ((threadIdx.x & 31) == 0) {
#ifdef VR_WORK_COUNTERS
        if (cn.nodes) atomicAdd(A.stats + ST_NODES, (unsigned long long)cn.nodes);
        if (cn.quads) atomicAdd(A.stats + ST_QUADRICS, (unsigned long long)cn.quads);
        if (cn.erf) atomicAdd(A.stats + ST_ERF, (unsigned long long)cn.erf);
        if (cn.mu) atomicAdd(A.stats + ST_MU, (unsigned long long)cn.mu);
        if (cn.sa) atomicAdd(A.stats + ST_SA, (unsigned long long)cn.sa);
        if (cn.it) atomicAdd(A.stats + ST_ITER, (unsigned long long)cn.it);
#endif
        if (cn.scatter) atomicAdd(A.stats + ST_SCATTER_RAYS, (unsigned long long)cn.scatter);
        if (cn.shadow) atomicAdd(A.stats + ST_SHADOW_RAYS, (unsigned long long)cn.shadow);
        if (cn.hits) atomicAdd(A.stats + ST_HITS, (unsigned long long)cn.hits);
        if (cn.over) atomicAdd(A.stats + ST_OVERFLOW, (unsigned long long)cn.over);
    }
}
enum { DEST_NONE = 0, DEST_TRAV = 1, DEST_REGEN = 2, DEST_TRAV_LIGHT = 3 };
// overlap records of a slot: the count and the first two records share one 32-byte sector (85 % of the free-flight
// rays of a sparse mixture gather at most two components), the rest are 16-byte records, contiguous per slot
__device__ __forceinline__ void pool_store_hit(const WavePool& Q, unsigned int p, int k, float t0, float t1, uint32_t gi) {
    if (k < 2) {
        uint32_t* r = reinterpret_cast<uint32_t*>(Q.hit01 + kSlotF4 * (size_t)p) + 1 + 3 * k;
        r[0] = __float_as_uint(t0); r[1] = __float_as_uint(t1); r[2] = gi;
    } else {
        Q.hitx[(size_t)p * (kWaveCap - 2) + (k - 2)] = make_uint4(__float_as_uint(t0), __float_as_uint(t1), gi, 0u);
    }
}
__device__ __forceinline__ void pool_load_hit(const WavePool& Q, unsigned int p, int k, const uint4& h0, float* t0, float* t1, uint32_t* gi) {
    if (k == 0) { *t0 = __uint_as_float(h0.y); *t1 = __uint_as_float(h0.z); *gi = h0.w; }
    else if (k == 1) {
        const uint4 h1 = Q.hit01[kSlotF4 * (size_t)p + 1];
        *t0 = __uint_as_float(h1.x); *t1 = __uint_as_float(h1.y); *gi = h1.z;
    } else {
        const uint4 h = Q.hitx[(size_t)p * (kWaveCap - 2) + (k - 2)];
        *t0 = __uint_as_float(h.x); *t1 = __uint_as_float(h.y); *gi = h.z;
    }
}
__device__ __forceinline__ int post_class(int K) { return K <= 1 ? 0 : K == 2 ? 1 : K <= 4 ? 2 : K <= 8 ? 3 : 4; }


// ------------------------------------------------------------------------------------------
// wave_trav<KIND>: one kernel per ray kind (free-flight / shadow), so that neither the node loop nor
// the batches carry the other kind's code.  Leaves hold ONE Gaussian and a leaf CHILD hit by the ray
// is queued as a candidate straight from its parent's slab test (no leaf visit).
// ------------------------------------------------------------------------------------------
constexpr int kTravQueueCap = 32 + 64;    // < 32 left over + two leaf children per lane and round
constexpr int kTravHitQueueCap = 64;      // shadow overlaps: < 32 left over + one quadric batch
#ifndef VR_WAIT_S
#define VR_WAIT_S 33   // > 32 = never (no gain for free-flight rays)
#endif
#ifndef VR_WAIT_H
#define VR_WAIT_H 4
#endif
#ifndef VR_FETCH_L
#define VR_FETCH_L 8
#endif
#ifndef VR_WAIT_L
#define VR_WAIT_L 4
#endif
#ifndef VR_FETCH_S
#define VR_FETCH_S 16
#endif
#ifndef VR_FETCH_H
#define VR_FETCH_H 6
#endif
#ifdef VR_TRAV_MINBLOCKS
#define VR_TRAV_BOUNDS __launch_bounds__(kBlockThreads, VR_TRAV_MINBLOCKS)
#else
#define VR_TRAV_BOUNDS __launch_bounds__(kBlockThreads)
#endif
#ifdef VR_POSTA_MINBLOCKS
#define VR_POSTA_BOUNDS __launch_bounds__(128, VR_POSTA_MINBLOCKS)
#else
#define VR_POSTA_BOUNDS __launch_bounds__(128)
#endif
// fetch new rays when this many lanes are free: the free-flight kernel's flush step is long (escapes are finished
// there), so it waits for more lanes than the shadow kernel (sweep in profiles/r01_tuning.md)
// candidate batches normally run full (32 entries); they run early once this many lanes have finished their traversal
// and only wait for queued candidates (those lanes cannot take a new ray before their result is complete)
template <int KIND> struct WaitThreshold { static constexpr unsigned value = KIND == RAY_SCATTER ? VR_WAIT_S : (KIND == RAY_SHADOW ? VR_WAIT_H : VR_WAIT_L); };
template <int KIND> struct FetchThreshold { static constexpr unsigned value = KIND == RAY_SCATTER ? VR_FETCH_S : (KIND == RAY_SHADOW ? VR_FETCH_H : VR_FETCH_L); };

template <int KIND> struct TravShared;
template <> struct TravShared<RAY_SCATTER> {
    float ray[7][32];               // o.xyz d.xyz lim of the lane's current ray
    unsigned int pid[32];           // pool slot of the lane's ray
    int count[32];                  // overlaps found (may exceed kWaveCap = overflow)
    int pending[32];                // candidates of the lane's ray already tested (the lane counts the queued ones)
    unsigned int t_bound[32];       // pruning bound, float bits (non-negative floats order as uints)
    uint32_t queue[kTravQueueCap];  // (owner lane << 27) | leaf-order Gaussian index
};
template <> struct TravShared<RAY_SHADOW> {
    float ray[7][32];
    unsigned int pid[32];
    int pending[32];                // candidates of the lane's ray fully processed (rejected, or depth added)
    long long tau[32];              // fixed-point optical depth
    uint32_t queue[kTravQueueCap];
    uint32_t hq_id[kTravHitQueueCap];
    float hq_a[kTravHitQueueCap], hq_b[kTravHitQueueCap];
};

template <> struct TravShared<RAY_LIGHT> : TravShared<RAY_SHADOW> {};   // same per-warp state; candidates come from the light's grid

enum { T_FETCH = 0, T_TRAV, T_WAIT, T_DONE };

struct TravLane {
#ifdef VR_WORK_COUNTERS
    unsigned int n_nodes, n_quads, n_erf;   // executed work of this lane
#endif
    int mode, cur, sp;
    int pushed;   // candidates this lane queued for its current ray; w.pending[lane] counts the ones already tested
    bool has_result;
    RaySetupQ rq;
    float tmax;
};

// free-flight rays: quadric test of up to 32 candidates; overlaps go to the owner's slot as (t0, t1, index)
__device__ __forceinline__ void scatter_quadric_batch(const SceneDev& S, const WavePool& Q, TravShared<RAY_SCATTER>& w, int lane,
                                                      int q_begin, int nb, TravLane& tl) {
    if (lane < nb) {
        VR_COUNT(++tl.n_quads;)
        const uint32_t e = w.queue[q_begin + lane];
        const int owner = (int)(e >> 27);
        const uint32_t gi = e & 0x07ffffffu;
        const f3 o = mk3(w.ray[0][owner], w.ray[1][owner], w.ray[2][owner]);
        const f3 d = mk3(w.ray[3][owner], w.ray[4][owner], w.ray[5][owner]);
        const GaussRec g = load_gauss(S.gauss, gi);
        const RayQuad q = ray_quad(g, o, d);
        float t0, t1;
        if (quad_intersect(q, &t0, &t1)) {
            const float tb = __uint_as_float(w.t_bound[owner]);
            if (!(t0 > tb)) {
                const int slot = atomicAdd(&w.count[owner], 1);
                if (slot < kWaveCap) {
                    pool_store_hit(Q, w.pid[owner], slot, t0, t1, gi);
                    if (t1 < tb) {
                        VR_COUNT(tl.n_erf += 2u;)
                        // pruning only: CONSERVATIVE estimate of this component's full optical depth
                        // (fast intrinsics; 1.5x margin covers the FP32 slack of the exact expression)
                        const float dn = g.density * g.norm;
                        const float inv4A = __fdividef(1.0f, 4.0f * q.A);
                        const float K = dn * sqrtf(6.2831853f * inv4A) * __expf(-0.5f * (q.C0 - q.B * q.B * inv4A));
                        const float iden = rsqrtf(8.0f * q.A);
                        const float full = K * (erff_ni((q.B + 2.0f * q.A * t1) * iden) - erff_ni((q.B + 2.0f * q.A * t0) * iden));
                        if (full > 1.5f * w.ray[6][owner]) atomicMin(&w.t_bound[owner], __float_as_uint(t1));
                    }
                }
            }
        }
        atomicAdd(&w.pending[owner], 1);
    }
}

// shadow rays: quadric test; overlaps with a non-empty clipped interval go to the depth queue
__device__ __forceinline__ void shadow_quadric_batch(const SceneDev& S, TravShared<RAY_SHADOW>& w, int lane, int q_begin, int nb,
                                                     int& nH, TravLane& tl) {
    bool to_hq = false;
    uint32_t e = 0; float ha = 0.f, hb = 0.f;
    if (lane < nb) {
        VR_COUNT(++tl.n_quads;)
        e = w.queue[q_begin + lane];
        const int owner = (int)(e >> 27);
        const uint32_t gi = e & 0x07ffffffu;
        const f3 o = mk3(w.ray[0][owner], w.ray[1][owner], w.ray[2][owner]);
        const f3 d = mk3(w.ray[3][owner], w.ray[4][owner], w.ray[5][owner]);
        const GaussRec g = load_gauss(S.gauss, gi);
        const RayQuad q = ray_quad(g, o, d);
        float t0, t1;
        if (quad_intersect(q, &t0, &t1) && w.tau[owner] <= tau_cutoff_fixed()) {
            ha = fmaxf(0.0f, t0); hb = fminf(w.ray[6][owner], t1);
            to_hq = hb > ha;   // stays pending until the depth batch has added it
        }
        if (!to_hq) atomicAdd(&w.pending[owner], 1);
    }
    const unsigned m = __ballot_sync(0xffffffffu, to_hq);
    if (to_hq) {
        const int pos = nH + __popc(m & ((1u << lane) - 1u));
        w.hq_id[pos] = e; w.hq_a[pos] = ha; w.hq_b[pos] = hb;
    }
    nH += __popc(m);
}

// exact clipped optical depth of up to 32 shadow overlaps (gmm.h:563-571, gaussian.h:208-231)
__device__ __forceinline__ void shadow_depth_batch(const SceneDev& S, TravShared<RAY_SHADOW>& w, int lane, int h_begin, int nb, TravLane& tl) {
    if (lane < nb) {
        VR_COUNT(tl.n_erf += 2u;)
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
        atomicAdd(&w.pending[owner], 1);
    }
}

// node rounds: one slab step per traversing lane and round; hit leaf children are queued as candidates.
// Runs until a full candidate batch is queued, enough lanes are free to fetch, or no lane traverses.
template <int KIND>
__device__ __forceinline__ void trav_rounds(const SceneDev& S, TravShared<KIND>& w, int lane, TravLane& tl, int stack[kStackDepth],
                                            int& nQ, bool can_fetch) {
    while (true) {
        bool l0 = false, l1 = false, fin = false;
        uint32_t g0 = 0, g1 = 0;   // leaf-order indices of the candidates this lane queues in this round
        if (tl.mode == T_TRAV) {
            VR_COUNT(++tl.n_nodes;)
            if constexpr (KIND == RAY_LIGHT) {
                // candidates come from the cell list of the light's direction grid, nearest to the light first; the
                // list ends for this ray at the first component that lies beyond the ray's origin (vr_lightgrid.h)
                TravShared<RAY_SHADOW>& wh = reinterpret_cast<TravShared<RAY_SHADOW>&>(w);
                int pos = tl.cur;
                const int end = tl.sp;
                if (!(wh.tau[lane] > tau_cutoff_fixed()) && pos < end && __ldg(S.lg_dmin + pos) <= tl.tmax) {
                    l0 = true; g0 = __ldg(S.lg_gi + pos); ++pos;
                    if (pos < end && __ldg(S.lg_dmin + pos) <= tl.tmax) { l1 = true; g1 = __ldg(S.lg_gi + pos); ++pos; }
                }
                tl.cur = pos;
                if (!l1) { fin = true; tl.mode = T_WAIT; }
            } else {
                float tlimit;
                bool stop;
                if (KIND == RAY_SCATTER) {
                    TravShared<RAY_SCATTER>& ws = reinterpret_cast<TravShared<RAY_SCATTER>&>(w);
                    tlimit = __uint_as_float(ws.t_bound[lane]); stop = ws.count[lane] > kWaveCap;
                } else {
                    TravShared<RAY_SHADOW>& wh = reinterpret_cast<TravShared<RAY_SHADOW>&>(w);
                    tlimit = tl.tmax; stop = wh.tau[lane] > tau_cutoff_fixed();
                }
                bool h0, h1; float tn0, tn1;
                int c0, c1;
                node_slabs_q(S.qnodes, tl.cur, tl.rq, tlimit, &h0, &h1, &tn0, &tn1, &c0, &c1);
                if (stop) { h0 = h1 = false; tl.sp = 0; }
                l0 = h0 && c0 < 0 && ((~c0) & 7);   // leaf child with a Gaussian in it
                l1 = h1 && c1 < 0 && ((~c1) & 7);
                g0 = (uint32_t)(~c0) >> 3; g1 = (uint32_t)(~c1) >> 3;
                const bool i0 = h0 && c0 >= 0, i1 = h1 && c1 >= 0;
                if (i0 && i1) {
                    const bool swap = tn1 < tn0;
                    stack[tl.sp++] = swap ? c0 : c1;
                    tl.cur = swap ? c1 : c0;
                } else if (i0) tl.cur = c0;
                else if (i1) tl.cur = c1;
                else if (tl.sp > 0) tl.cur = stack[--tl.sp];
                else { fin = true; tl.mode = T_WAIT; }   // resolved below once the pushes are counted
            }
            tl.pushed += (int)l0 + (int)l1;
        }
        const unsigned m0 = __ballot_sync(0xffffffffu, l0);
        if (l0) w.queue[nQ + __popc(m0 & ((1u << lane) - 1u))] = ((uint32_t)lane << 27) | g0;
        nQ += __popc(m0);
        const unsigned m1 = __ballot_sync(0xffffffffu, l1);
        if (l1) w.queue[nQ + __popc(m1 & ((1u << lane) - 1u))] = ((uint32_t)lane << 27) | g1;
        nQ += __popc(m1);
        if (fin && w.pending[lane] == tl.pushed) { tl.mode = T_FETCH; tl.has_result = true; }
        if (nQ >= 32) break;
        if (__ballot_sync(0xffffffffu, fin)) {
            if (can_fetch && __popc(__ballot_sync(0xffffffffu, tl.mode == T_FETCH)) >= FetchThreshold<KIND>::value) break;
            if (__ballot_sync(0xffffffffu, tl.mode == T_TRAV) == 0) break;
            if (WaitThreshold<KIND>::value <= 32 && __popc(__ballot_sync(0xffffffffu, tl.mode == T_WAIT)) >= WaitThreshold<KIND>::value) break;
        }
    }
    __syncwarp();
}

template <int KIND>
__global__ void VR_TRAV_BOUNDS wave_trav_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q) {
    __shared__ TravShared<KIND> wsh[kBlockThreads / 32];
    TravShared<KIND>& w = wsh[threadIdx.x >> 5];
    const SceneDev& S = A.scene;
    const int lane = threadIdx.x & 31;
    const unsigned int nT = Q.cnt[KIND == RAY_SCATTER ? WC_NTS : (KIND == RAY_SHADOW ? WC_NTH : WC_NTL)];
    const unsigned int* __restrict__ list = KIND == RAY_SCATTER ? Q.listTs : (KIND == RAY_SHADOW ? Q.listTh : Q.listTl);
    unsigned int* cursor = &Q.cnt[KIND == RAY_SCATTER ? WC_CURSOR_S : (KIND == RAY_SHADOW ? WC_CURSOR_H : WC_CURSOR_L)];
    // the list consumed by the previous kernel of the chain (regen -> travS -> postA -> travH -> postB) is free again
    if (blockIdx.x == 0 && threadIdx.x < kPostClasses + 1) {
        if (KIND == RAY_SCATTER) { if (threadIdx.x == 0) Q.cnt[WC_NC] = 0; }
        else if (KIND == RAY_SHADOW) Q.cnt[WC_NA + threadIdx.x] = 0;
    }
    int stack[kStackDepth];
    TravLane tl;
    tl.mode = T_FETCH; tl.cur = 0; tl.sp = 0; tl.has_result = false; tl.tmax = 0.f; tl.pushed = 0;
    VR_COUNT(tl.n_nodes = tl.n_quads = tl.n_erf = 0u;)
    int nQ = 0, nH = 0;
    unsigned int n_escaped = 0;
    bool exhausted = false;  // warp-uniform: the wave has no more rays to hand out
    while (true) {
        const unsigned mT = __ballot_sync(0xffffffffu, tl.mode == T_TRAV);
        const bool drain = mT == 0 || (WaitThreshold<KIND>::value <= 32 && __popc(__ballot_sync(0xffffffffu, tl.mode == T_WAIT)) >= WaitThreshold<KIND>::value);
        if (KIND != RAY_SCATTER) {
            if (nH >= 32 || (nH > 0 && drain && nQ == 0)) {
                TravShared<RAY_SHADOW>& wh = reinterpret_cast<TravShared<RAY_SHADOW>&>(w);
                const int nb = min(nH, 32);
                nH -= nb;
                shadow_depth_batch(S, wh, lane, nH, nb, tl);
                __syncwarp();
                if (tl.mode == T_WAIT && w.pending[lane] == tl.pushed) { tl.mode = T_FETCH; tl.has_result = true; }
                continue;
            }
        }
        if (nQ >= 32 || (nQ > 0 && drain)) {
            const int nb = min(nQ, 32);
            nQ -= nb;
            if (KIND == RAY_SCATTER) scatter_quadric_batch(S, Q, reinterpret_cast<TravShared<RAY_SCATTER>&>(w), lane, nQ, nb, tl);
            else shadow_quadric_batch(S, reinterpret_cast<TravShared<RAY_SHADOW>&>(w), lane, nQ, nb, nH, tl);
            __syncwarp();
            if (tl.mode == T_WAIT && w.pending[lane] == tl.pushed) { tl.mode = T_FETCH; tl.has_result = true; }
            continue;
        }
        const unsigned mF = __ballot_sync(0xffffffffu, tl.mode == T_FETCH);
        if (mF && (__popc(mF) >= FetchThreshold<KIND>::value || mT == 0)) {
            // ---- flush finished rays: result to the slot, slot to the list of its post kernel
            const bool flush = tl.mode == T_FETCH && tl.has_result;
            const unsigned int p = w.pid[lane];
            if (KIND == RAY_SCATTER) {
                TravShared<RAY_SCATTER>& ws = reinterpret_cast<TravShared<RAY_SCATTER>&>(w);
                const int K = flush ? ws.count[lane] : 0;
                // a ray that overlapped nothing ends the path here: environment radiance, sample written,
                // slot back to the regeneration list (integrator.h:594-597; the target draw is not committed)
                const bool esc = flush && K == 0;
                if (esc) {
                    const float4 s0 = Q.ps01[kSlotF4 * (size_t)p];
                    const uint4 s1 = reinterpret_cast<uint4*>(Q.ps01)[kSlotF4 * (size_t)p + 1];
                    A.sampleL[s1.z] = make_float4(__fadd_rn(s0.x, __fmul_rn(s0.w, S.env[0])), __fadd_rn(s0.y, __fmul_rn(s0.w, S.env[1])),
                                                  __fadd_rn(s0.z, __fmul_rn(s0.w, S.env[2])), (float)s1.w);
                }
                if (flush && K > 0) reinterpret_cast<int*>(Q.hit01 + kSlotF4 * (size_t)p)[0] = K;
                wave_append(Q.listC, &Q.cnt[WC_NC], esc, p);
                const int cls = post_class(K);
#pragma unroll
                for (int c = 0; c < kPostClasses; ++c)
                    wave_append(Q.listA + (size_t)c * Q.P, &Q.cnt[WC_NA + c], flush && K > 0 && cls == c, p);
            } else {
                TravShared<RAY_SHADOW>& wh = reinterpret_cast<TravShared<RAY_SHADOW>&>(w);
                if (flush) reinterpret_cast<long long*>(Q.ps2 + kSlotF4 * (size_t)p + 1)[0] = wh.tau[lane];
                wave_append(Q.listB, &Q.cnt[WC_NB], flush, p);
            }
            tl.has_result = false;
            // ---- fetch the next rays of the wave
            if (!exhausted) {
                unsigned int base = 0;
                if (lane == __ffs(mF) - 1) base = atomicAdd(cursor, (unsigned int)__popc(mF));
                base = __shfl_sync(0xffffffffu, base, __ffs(mF) - 1);
                if (tl.mode == T_FETCH) {
                    const unsigned int idx = base + (unsigned int)__popc(mF & ((1u << lane) - 1u));
                    if (idx < nT) {
                        const unsigned int np = __ldg(list + idx);
                        const float4 ro = Q.ray[kSlotF4 * (size_t)np], rd = Q.ray[kSlotF4 * (size_t)np + 1];
                        w.ray[0][lane] = ro.x; w.ray[1][lane] = ro.y; w.ray[2][lane] = ro.z;
                        w.ray[3][lane] = rd.x; w.ray[4][lane] = rd.y; w.ray[5][lane] = rd.z;
                        w.ray[6][lane] = ro.w;
                        w.pid[lane] = np; w.pending[lane] = 0; tl.pushed = 0;
                        if (KIND == RAY_SCATTER) {
                            TravShared<RAY_SCATTER>& ws = reinterpret_cast<TravShared<RAY_SCATTER>&>(w);
                            ws.count[lane] = 0; ws.t_bound[lane] = 0x7f800000u;
                        } else {
                            reinterpret_cast<TravShared<RAY_SHADOW>&>(w).tau[lane] = 0LL;
                        }
                        if constexpr (KIND == RAY_LIGHT) {
                            const int light = (__float_as_int(rd.w) >> 8) - 1;
                            const f3 v = mk3(ro.x - __ldg(S.lpos + 3 * light), ro.y - __ldg(S.lpos + 3 * light + 1), ro.z - __ldg(S.lpos + 3 * light + 2));
                            const size_t cells = (size_t)6 * S.lg_res * S.lg_res;
                            const uint32_t* cs = S.lg_start + (size_t)light * (cells + 1) + light_grid_cell_dev(v, S.lg_res);
                            tl.cur = (int)__ldg(cs); tl.sp = (int)__ldg(cs + 1);
                            tl.tmax = __fmaf_rn(ro.w, 1.000001f, 1e-6f);   // entries farther from the light than the ray's origin cannot overlap it
                        } else {
                            tl.rq = ray_setup_q(S, mk3(ro.x, ro.y, ro.z), mk3(rd.x, rd.y, rd.z));
                            tl.tmax = ro.w; tl.sp = 0; tl.cur = 0;
                        }
                        const bool nothing = S.n == 0 || (KIND != RAY_SCATTER && !(ro.w > 0.0f));
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
        trav_rounds<KIND>(S, w, lane, tl, stack, nQ, !exhausted);
    }
    (void)n_escaped;
#ifdef VR_WORK_COUNTERS
    WaveCounters cn = {0u, 0u, 0u, 0u, tl.n_nodes, tl.n_quads, tl.n_erf, 0u, 0u, 0u};
    wave_flush_counters(A, cn);
#endif
}

// ------------------------------------------------------------------------------------------
// wave_trav2<KIND>: the same traversal with the per-round bookkeeping taken out of the node loop.
//
// SASS of wave_trav_kernel (r01): one round = one slab step per lane costs ~168 warp instructions, of which the two slab
// tests are 39; ~45 go into queueing the hit leaf children through two ballots, ~27 into the exit conditions, the rest into
// child classification and the stack.  Here a lane keeps its candidates in FOUR PRIVATE shared-memory slots (a predicated
// store each, no ballot) and the node loop is left on ONE ballot: enough lanes have stopped traversing -- finished, or their
// slots are full.  Everything warp-cooperative happens between node loops: the slots of all lanes are compacted into one
// list (one ballot per slot level per phase, not per round), the quadric tests run 32 at a time as before, every passing
// shadow overlap gets its depth in the same phase, and since a phase tests EVERYTHING that is queued, a lane that finished
// traversing is complete after it: the pending / pushed counters and their shared atomics are gone.  The pruning bound and
// the overlap count change only inside a phase, so the node loop keeps them in registers.
// Candidates, tests and sums are the same as wave_trav_kernel's (test_trav2_equals_trav: bit-identical images).
// ------------------------------------------------------------------------------------------
// leave the node loop when this many lanes do not traverse; sweep at C2 (profiles/r02/trav2_threshold_sweep.md):
// 4 -> 229, 8 -> 268, 12 -> 279, 16 -> 289, 24 -> 288 Msamples/s; with the uniform grid (lanes in a grid cell stop after a
// round or two): 12 -> 373.4, 16 -> 378.5, 20 -> 379.2, 24 -> 378.7, and 20 with a fetch threshold of 16 -> 382.5
#ifndef VR_T2_IDLE_S
#define VR_T2_IDLE_S 20
#endif
#ifndef VR_T2_IDLE_H
#define VR_T2_IDLE_H 16
#endif
#ifndef VR_T2_IDLE_L
#define VR_T2_IDLE_L 16
#endif
// finished rays are flushed (result to the slot, slot to a work list: up to six warp-aggregated appends) and their lanes
// refilled once this many wait -- or nothing traverses any more
#ifndef VR_T2_FETCH_S
#define VR_T2_FETCH_S 16   // (with the uniform grid: 8 -> 371.8, 12 -> 378.5, 16 -> 379.9)
#endif
#ifndef VR_T2_FETCH_H
#define VR_T2_FETCH_H 12
#endif
#ifndef VR_T2_FETCH_L
#define VR_T2_FETCH_L 12
#endif
constexpr int kT2Slots = 4;
// 2.3 KB per warp: the traversal kernels live on L1 hits (hierarchy nodes, component records, the per-lane stacks in local
// memory), and shared memory comes out of the same 256 KB -- at 4 KB per warp the L1 hit rate of the node loads fell from 82 to
// 72 % and that of the stacks from 52 to 19 % (profiles/r02/trav2_first_ncu.md)
// free-flight GRID kernel only.  t_cut: +(parameter at which the ray leaves its origin's grid cell) while the lane tests that
// cell's list, -(the same) once it walks the hierarchy, and at the end the ray's CERTIFICATE (see grid_done).  gsum: a lower
// bound of the optical depth the kept overlaps hold before the ray leaves the cell.
template <int KIND> struct Trav2Grid {};
template <> struct Trav2Grid<RAY_SCATTER> { float t_cut[32]; float gsum[32]; };
template <int KIND> struct Trav2Shared : Trav2Grid<KIND> {
    float ray[7][32];
    unsigned int pid[32];
    union {
        struct { int count[32]; unsigned int t_bound[32]; };   // free-flight: overlaps found, pruning bound (float bits)
        long long tau[32];                                      // shadow: fixed-point optical depth
    };
    uint32_t slot[kT2Slots][32];     // private candidate slots of the lanes
    uint32_t queue[kT2Slots * 32];   // compacted (owner << 27) | index
};
enum { T2_FETCH = 0, T2_TRAV, T2_FULL, T2_WAIT, T2_DONE, T2_GFULL, T2_GWAIT };   // G*: a GRID lane in its grid cell (slots full / list through)
struct Trav2Lane {
#ifdef VR_WORK_COUNTERS
    unsigned int n_nodes, n_quads, n_erf;
#endif
    int mode, cur, sp, nq;
    bool has_result;
    int gcell;          // GRID kernels: >= 0 while the lane tests the list of its origin's grid cell, -1 in the hierarchy
    RaySetupQ rq;
    float tmax, tlimit;
};

// all queued candidates of all lanes: compact, test 32 at a time; returns with every slot empty
template <int KIND, bool GRID = false>
__device__ __forceinline__ void trav2_phase(const SceneDev& S, const WavePool& Q, Trav2Shared<KIND>& w, int lane, Trav2Lane& tl) {
    int total = 0;
#pragma unroll
    for (int k = 0; k < kT2Slots; ++k) {
        const bool has = tl.nq > k;
        const unsigned m = __ballot_sync(0xffffffffu, has);
        if (has) w.queue[total + __popc(m & ((1u << lane) - 1u))] = ((uint32_t)lane << 27) | w.slot[k][lane];
        total += __popc(m);
    }
    tl.nq = 0;
    __syncwarp();
    for (int base = 0; base < total; base += 32) {
        const int it = base + lane;
        if (it < total) {
            VR_COUNT(++tl.n_quads;)
            const uint32_t e = w.queue[it];
            const int owner = (int)(e >> 27);
            const uint32_t gi = e & 0x07ffffffu;
            const f3 o = mk3(w.ray[0][owner], w.ray[1][owner], w.ray[2][owner]);
            const f3 d = mk3(w.ray[3][owner], w.ray[4][owner], w.ray[5][owner]);
            const GaussRec g = load_gauss(S.gauss, gi);
            const RayQuad q = ray_quad(g, o, d);
            float t0, t1;
            if (quad_intersect(q, &t0, &t1)) {
                if constexpr (KIND == RAY_SCATTER) {
                    const float tb = __uint_as_float(w.t_bound[owner]);
                    // GRID: a lane in its grid cell keeps what enters before the ray leaves that cell, a lane in the hierarchy what
                    // enters at or beyond that point: every overlap is kept exactly once
                    bool mine = true, in_cell = false;
                    float t_leave = 0.0f;
                    if constexpr (GRID) {
                        const float tc = w.t_cut[owner];
                        in_cell = __float_as_int(tc) >= 0;
                        t_leave = fabsf(tc);
                        mine = in_cell ? (t0 < t_leave) : !(t0 < t_leave);
                    }
                    if (!(t0 > tb) && mine) {
                        const int slot = atomicAdd(&w.count[owner], 1);
                        if (slot < kWaveCap) {
                            pool_store_hit(Q, w.pid[owner], slot, t0, t1, gi);
                            if (t1 < tb || in_cell) {
                                VR_COUNT(tl.n_erf += 2u;)
                                // pruning only: CONSERVATIVE estimate of this component's full optical depth (see wave_trav_kernel)
                                const float dn = g.density * g.norm;
                                const float inv4A = __fdividef(1.0f, 4.0f * q.A);
                                const float K = dn * sqrtf(6.2831853f * inv4A) * __expf(-0.5f * (q.C0 - q.B * q.B * inv4A));
                                const float iden = rsqrtf(8.0f * q.A);
                                const float F0 = erff_ni((q.B + 2.0f * q.A * t0) * iden);
                                const float full = K * (erff_ni((q.B + 2.0f * q.A * t1) * iden) - F0);
                                if (t1 < tb && full > 1.5f * w.ray[6][owner]) atomicMin(&w.t_bound[owner], __float_as_uint(t1));
                                if constexpr (GRID) {
                                    if (in_cell) {
                                        // a LOWER bound of this component's depth before the ray leaves the cell: the clipped interval can be
                                        // arbitrarily short, where the difference of two erff values (2 ulp each) cancels -- 4e-6 K covers that
                                        // absolute error eight times over; the sum is compared with 1.5 x the target like the estimate above
                                        const float part = t1 <= t_leave ? full : K * (erff_ni((q.B + 2.0f * q.A * t_leave) * iden) - F0);
                                        atomicAdd(&w.gsum[owner], part - 4e-6f * K);
                                    }
                                }
                            }
                        }
                    }
                } else {
                    // exact clipped optical depth of the overlap (gmm.h:563-571, gaussian.h:208-231), by the lane that tested it:
                    // component record and quadric coefficients are in registers (a second, compacted pass would reload both)
                    const float ha = fmaxf(0.0f, t0), hb = fminf(w.ray[6][owner], t1);
                    if (hb > ha && w.tau[owner] <= tau_cutoff_fixed()) {
                        VR_COUNT(tl.n_erf += 2u;)
                        const DepthConst c = depth_const(g, q);
                        const float F1 = erff_ni(erf_arg(c, hb));
                        const float F0 = erff_ni(erf_arg(c, ha));
                        const float tau = __fmul_rn(c.K, __fsub_rn(F1, F0));
                        atomicAdd((unsigned long long*)&w.tau[owner], (unsigned long long)__double2ll_rn((double)tau * kTauFixed));
                    }
                }
            }
        }
    }
    __syncwarp();
}

// ---- GRID lanes: the walk through the cells of the uniform grid over the mixture (vr_grid3d.cu) --------------------------
// parameter at which the ray leaves cell (ix, iy, iz), and through which axis
__device__ __forceinline__ float grid_cell_exit(const SceneDev& S, f3 o, f3 d, int ix, int iy, int iz, int* axis) {
    const float px = fmaf((float)(ix + (d.x > 0.0f ? 1 : 0)), S.g3_cell[0], S.g3_org[0]);
    const float py = fmaf((float)(iy + (d.y > 0.0f ? 1 : 0)), S.g3_cell[1], S.g3_org[1]);
    const float pz = fmaf((float)(iz + (d.z > 0.0f ? 1 : 0)), S.g3_cell[2], S.g3_org[2]);
    const float tx = d.x != 0.0f ? (px - o.x) / d.x : CUDART_INF_F;
    const float ty = d.y != 0.0f ? (py - o.y) / d.y : CUDART_INF_F;
    const float tz = d.z != 0.0f ? (pz - o.z) / d.z : CUDART_INF_F;
    int a = 0; float t = tx;
    if (ty < t) { t = ty; a = 1; }
    if (tz < t) { t = tz; a = 2; }
    *axis = a;
    return t;
}
__device__ __forceinline__ void grid_cell_list(const SceneDev& S, int ix, int iy, int iz, int* pos, int* end) {
    const uint32_t* cs = S.g3_start + ((size_t)iz * S.g3_res + iy) * S.g3_res + ix;
    *pos = (int)__ldg(cs); *end = (int)__ldg(cs + 1);
}
// the lane's ray starts with the list of its origin's cell; false: the origin is outside the grid
template <class W>
__device__ __forceinline__ bool grid_begin(const SceneDev& S, W& w, int lane, Trav2Lane& tl, f3 o, f3 d) {
    const float rx = (o.x - S.g3_org[0]) * S.g3_inv[0], ry = (o.y - S.g3_org[1]) * S.g3_inv[1], rz = (o.z - S.g3_org[2]) * S.g3_inv[2];
    const float fres = (float)S.g3_res;
    if (!(rx >= 0.0f && rx < fres && ry >= 0.0f && ry < fres && rz >= 0.0f && rz < fres)) return false;
    const int ix = min((int)rx, S.g3_res - 1), iy = min((int)ry, S.g3_res - 1), iz = min((int)rz, S.g3_res - 1);
    int axis;
    w.gsum[lane] = 0.0f;
    w.t_cut[lane] = fmaxf(grid_cell_exit(S, o, d, ix, iy, iz, &axis), 0.0f);
    grid_cell_list(S, ix, iy, iz, &tl.cur, &tl.sp);
    tl.gcell = 0;
    return true;
}
// The list of the lane's cell is through and every candidate of it has been tested: everything that ENTERS before the ray leaves
// the cell (t_leave) has been kept.  Returns the lane's new mode.  A finished ray leaves its CERTIFICATE in w.t_cut: the
// parameter up to which its overlap list is known to be complete (wave_post_a checks its result against it, see post_a_certified).
template <class W>
__device__ __forceinline__ int grid_done(const SceneDev& S, W& w, int lane, Trav2Lane& tl) {
    const float t_leave = w.t_cut[lane];
    const float tb = __uint_as_float(w.t_bound[lane]);
    tl.gcell = -1;
    // more overlaps than a slot holds: the sample is re-traced anyway
    if (w.count[lane] > kWaveCap) { w.t_cut[lane] = CUDART_INF_F; tl.has_result = true; return T2_FETCH; }
    // the pruning bound -- the exit of a component that holds more than the sampled depth, so the scatter point and the event
    // after it lie before it -- is inside the cell: everything that enters up to the bound has been kept, like in the hierarchy
    if (tb < t_leave) { w.t_cut[lane] = CUDART_INF_F; tl.has_result = true; return T2_FETCH; }
    // the kept overlaps hold more than the sampled depth before the ray leaves the cell: it scatters inside the cell.  The list
    // is complete up to t_leave only; whether that suffices depends on where the walk ends (the event AFTER the scatter point may
    // be an overlap that enters beyond t_leave), so the ray carries t_leave as its certificate
    if (w.gsum[lane] > 1.5f * w.ray[6][lane]) { tl.has_result = true; return T2_FETCH; }
    // otherwise the hierarchy, for what enters at or beyond t_leave
    w.t_cut[lane] = -t_leave;
    tl.rq = ray_setup_q(S, mk3(w.ray[0][lane], w.ray[1][lane], w.ray[2][lane]), mk3(w.ray[3][lane], w.ray[4][lane], w.ray[5][lane]));
    tl.sp = 0; tl.cur = 0;
    return T2_TRAV;
}

// GRID (free-flight rays only): a ray whose origin lies in the uniform grid over the mixture (vr_grid3d.cu) first tests the list
// of its origin's cell and keeps what enters before it leaves that cell.  In these optically thick mixtures that is nearly always
// all it needs (83 % of the bounce rays of C2 scatter within 0.008 units): the ray is finished, with a certificate, when those
// overlaps hold more than the sampled depth inside the cell; otherwise it goes on in the hierarchy for what enters later.
// (Walking further cells front to back up to the pruning bound instead was measured: exact without certificates, but slower than
// the hierarchy -- 346 vs 364 Msamples/s at C2.)
template <int KIND, bool GRID = false>
__global__ void VR_TRAV_BOUNDS wave_trav2_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q) {
    __shared__ Trav2Shared<KIND> wsh[kBlockThreads / 32];
    Trav2Shared<KIND>& w = wsh[threadIdx.x >> 5];
    const SceneDev& S = A.scene;
    const int lane = threadIdx.x & 31;
    const unsigned int nT = Q.cnt[KIND == RAY_SCATTER ? WC_NTS : (KIND == RAY_SHADOW ? WC_NTH : WC_NTL)];
    const unsigned int* __restrict__ list = KIND == RAY_SCATTER ? Q.listTs : (KIND == RAY_SHADOW ? Q.listTh : Q.listTl);
    unsigned int* cursor = &Q.cnt[KIND == RAY_SCATTER ? WC_CURSOR_S : (KIND == RAY_SHADOW ? WC_CURSOR_H : WC_CURSOR_L)];
    constexpr unsigned kIdle = KIND == RAY_SCATTER ? VR_T2_IDLE_S : (KIND == RAY_SHADOW ? VR_T2_IDLE_H : VR_T2_IDLE_L);
    constexpr int kFetch = KIND == RAY_SCATTER ? VR_T2_FETCH_S : (KIND == RAY_SHADOW ? VR_T2_FETCH_H : VR_T2_FETCH_L);
    if (blockIdx.x == 0 && threadIdx.x < kPostClasses + 1) {   // lists consumed by the previous kernel of the chain are free again
        if (KIND == RAY_SCATTER) { if (threadIdx.x == 0) Q.cnt[WC_NC] = 0; }
        else if (KIND == RAY_SHADOW) Q.cnt[WC_NA + threadIdx.x] = 0;
    }
    int stack[kStackDepth];
    Trav2Lane tl;
    tl.mode = T2_FETCH; tl.cur = 0; tl.sp = 0; tl.nq = 0; tl.has_result = false; tl.tmax = 0.f; tl.tlimit = 0.f; tl.gcell = -1;
    VR_COUNT(tl.n_nodes = tl.n_quads = tl.n_erf = 0u;)
    bool exhausted = false;  // warp-uniform: the wave has no more rays to hand out
    while (true) {
        // ---- node loop: one slab step (or two grid entries) per traversing lane and round, until kIdle of the lanes that still
        //      have a ray do not traverse (finished or slots full) -- or, near the end of the wave, none does
        if (__ballot_sync(0xffffffffu, tl.mode == T2_TRAV)) {
            const int alive = 32 - __popc(__ballot_sync(0xffffffffu, tl.mode == T2_DONE));
            const int keep_going = max(alive - (int)kIdle, 0);   // leave once at most this many lanes traverse
            while (true) {
                if (tl.mode == T2_TRAV) {
                    VR_COUNT(++tl.n_nodes;)
                    bool more = true, in_cell = false;
                    if constexpr (GRID) {
                        if (tl.gcell >= 0) {   // two entries of the cell's list per round
                            in_cell = true;
                            int pos = tl.cur;
                            const int end = tl.sp;
                            if (pos < end) w.slot[tl.nq++][lane] = __ldg(S.g3_gi + pos++);
                            if (pos < end) w.slot[tl.nq++][lane] = __ldg(S.g3_gi + pos++);
                            tl.cur = pos;
                            if (pos >= end) tl.mode = tl.nq == 0 ? grid_done(S, w, lane, tl) : T2_GWAIT;   // nothing queued: decide right away
                            else if (tl.nq > kT2Slots - 2) tl.mode = T2_GFULL;
                        }
                    }
                    if (in_cell) {
                    } else if constexpr (KIND == RAY_LIGHT) {
                        int pos = tl.cur;
                        const int end = tl.sp;
                        more = false;
                        if (pos < end && __ldg(S.lg_dmin + pos) <= tl.tmax) {
                            w.slot[tl.nq++][lane] = __ldg(S.lg_gi + pos); ++pos;
                            if (pos < end && __ldg(S.lg_dmin + pos) <= tl.tmax) { w.slot[tl.nq++][lane] = __ldg(S.lg_gi + pos); ++pos; more = true; }
                        }
                        tl.cur = pos;
                    } else {
                        bool h0, h1; float tn0, tn1;
                        int c0, c1;
                        node_slabs_q(S.qnodes, tl.cur, tl.rq, tl.tlimit, &h0, &h1, &tn0, &tn1, &c0, &c1);
                        if (h0 && c0 < 0) w.slot[tl.nq++][lane] = (uint32_t)(~c0) >> 3;   // a leaf child: ONE component (an empty
                        if (h1 && c1 < 0) w.slot[tl.nq++][lane] = (uint32_t)(~c1) >> 3;   // leaf's quantised box is unhittable)
                        const bool i0 = h0 && c0 >= 0, i1 = h1 && c1 >= 0;
                        more = true;
                        // (keeping the newest stack entry in a register and prefetching the one below was measured: 313.6 vs 317.4
                        // Msamples/s at C2 -- seven more registers cost a resident block per SM)
                        if (i0 && i1) {
                            const bool swap = tn1 < tn0;
                            stack[tl.sp++] = swap ? c0 : c1;
                            tl.cur = swap ? c1 : c0;
                        } else if (i0) tl.cur = c0;
                        else if (i1) tl.cur = c1;
                        else if (tl.sp > 0) tl.cur = stack[--tl.sp];
                        else more = false;
                    }
                    if (in_cell) {
                    } else if (!more) tl.mode = T2_WAIT;
                    else if (tl.nq > kT2Slots - 2) tl.mode = T2_FULL;   // another round could bring two more
                }
                if (__popc(__ballot_sync(0xffffffffu, tl.mode == T2_TRAV)) <= keep_going) break;
            }
        }
        // ---- phase: every queued candidate is tested; afterwards a lane that finished traversing is complete
        if (__ballot_sync(0xffffffffu, tl.nq > 0)) trav2_phase<KIND, GRID>(S, Q, w, lane, tl);
        if constexpr (GRID) {
            if (tl.mode == T2_GFULL) tl.mode = T2_TRAV;                        // more of the cell's list
            else if (tl.mode == T2_GWAIT) tl.mode = grid_done(S, w, lane, tl);   // the cell is through
        }
        if (tl.mode == T2_WAIT) { tl.mode = T2_FETCH; tl.has_result = true; }
        else if (tl.mode == T2_FULL || tl.mode == T2_TRAV) {   // the phase may have moved this lane's bound / count / depth sum
            tl.mode = T2_TRAV;
            if constexpr (KIND == RAY_SCATTER) {
                tl.tlimit = __uint_as_float(w.t_bound[lane]);
                if (w.count[lane] > kWaveCap) { tl.mode = T2_FETCH; tl.has_result = true; }     // overflow: the sample is re-traced
            } else {
                if (w.tau[lane] > tau_cutoff_fixed()) { tl.mode = T2_FETCH; tl.has_result = true; }   // transmittance is exactly 0
            }
        }
        const unsigned mT = __ballot_sync(0xffffffffu, tl.mode == T2_TRAV);
        const unsigned mF = __ballot_sync(0xffffffffu, tl.mode == T2_FETCH);
        if (mF && (__popc(mF) >= kFetch || mT == 0)) {
            // ---- flush finished rays: result to the slot, slot to the list of its post kernel
            const bool flush = tl.mode == T2_FETCH && tl.has_result;
            const unsigned int p = w.pid[lane];
            if constexpr (KIND == RAY_SCATTER) {
                const int K = flush ? w.count[lane] : 0;
                // a ray that overlapped nothing ends the path here (integrator.h:594-597; the target draw is not committed)
                const bool esc = flush && K == 0;
                if (esc) {
                    const float4 s0 = Q.ps01[kSlotF4 * (size_t)p];
                    const uint4 s1 = reinterpret_cast<uint4*>(Q.ps01)[kSlotF4 * (size_t)p + 1];
                    A.sampleL[s1.z] = make_float4(__fadd_rn(s0.x, __fmul_rn(s0.w, S.env[0])), __fadd_rn(s0.y, __fmul_rn(s0.w, S.env[1])),
                                                  __fadd_rn(s0.z, __fmul_rn(s0.w, S.env[2])), (float)s1.w);
                }
                if (flush && K > 0) reinterpret_cast<int*>(Q.hit01 + kSlotF4 * (size_t)p)[0] = K;
                if constexpr (GRID) {   // the certificate (grid_done): +inf for a list made by the hierarchy or closed by the pruning bound
                    const float tc = w.t_cut[lane];
                    if (flush && K > 0) reinterpret_cast<float*>(Q.hit01 + kSlotF4 * (size_t)p + 1)[3] = __float_as_int(tc) < 0 ? CUDART_INF_F : tc;
                }
                wave_append(Q.listC, &Q.cnt[WC_NC], esc, p);
                const int cls = post_class(K);
#pragma unroll
                for (int c = 0; c < kPostClasses; ++c)
                    wave_append(Q.listA + (size_t)c * Q.P, &Q.cnt[WC_NA + c], flush && K > 0 && cls == c, p);
            } else {
                if (flush) reinterpret_cast<long long*>(Q.ps2 + kSlotF4 * (size_t)p + 1)[0] = w.tau[lane];
                wave_append(Q.listB, &Q.cnt[WC_NB], flush, p);
            }
            tl.has_result = false;
            // ---- fetch the next rays of the wave
            if (!exhausted) {
                unsigned int base = 0;
                if (lane == __ffs(mF) - 1) base = atomicAdd(cursor, (unsigned int)__popc(mF));
                base = __shfl_sync(0xffffffffu, base, __ffs(mF) - 1);
                if (tl.mode == T2_FETCH) {
                    const unsigned int idx = base + (unsigned int)__popc(mF & ((1u << lane) - 1u));
                    if (idx < nT) {
                        const unsigned int np = __ldg(list + idx);
                        const float4 ro = Q.ray[kSlotF4 * (size_t)np], rd = Q.ray[kSlotF4 * (size_t)np + 1];
                        w.ray[0][lane] = ro.x; w.ray[1][lane] = ro.y; w.ray[2][lane] = ro.z;
                        w.ray[3][lane] = rd.x; w.ray[4][lane] = rd.y; w.ray[5][lane] = rd.z;
                        w.ray[6][lane] = ro.w;
                        w.pid[lane] = np;
                        if constexpr (KIND == RAY_SCATTER) { w.count[lane] = 0; w.t_bound[lane] = 0x7f800000u; tl.tlimit = CUDART_INF_F; }
                        else w.tau[lane] = 0LL;
                        if constexpr (KIND == RAY_LIGHT) {
                            const int light = (__float_as_int(rd.w) >> 8) - 1;
                            const f3 v = mk3(ro.x - __ldg(S.lpos + 3 * light), ro.y - __ldg(S.lpos + 3 * light + 1), ro.z - __ldg(S.lpos + 3 * light + 2));
                            const size_t cells = (size_t)6 * S.lg_res * S.lg_res;
                            const uint32_t* cs = S.lg_start + (size_t)light * (cells + 1) + light_grid_cell_dev(v, S.lg_res);
                            tl.cur = (int)__ldg(cs); tl.sp = (int)__ldg(cs + 1);
                            tl.tmax = __fmaf_rn(ro.w, 1.000001f, 1e-6f);   // entries farther from the light than the ray's origin cannot overlap it
                        } else {
                            bool in_grid = false;
                            if constexpr (GRID) {
                                // (a ray wave_post_a sent back -- its certificate did not cover its result -- takes the hierarchy)
                                if (!(__float_as_int(rd.w) & kRayNoGrid)) in_grid = grid_begin(S, w, lane, tl, mk3(ro.x, ro.y, ro.z), mk3(rd.x, rd.y, rd.z));
                                if (!in_grid) { w.t_cut[lane] = -0.0f; tl.gcell = -1; }   // the hierarchy from the start: keeps everything
                            }
                            if (!in_grid) { tl.rq = ray_setup_q(S, mk3(ro.x, ro.y, ro.z), mk3(rd.x, rd.y, rd.z)); tl.sp = 0; tl.cur = 0; }
                            tl.tmax = ro.w;
                            if constexpr (KIND == RAY_SHADOW) tl.tlimit = ro.w;
                        }
                        const bool nothing = S.n == 0 || (KIND != RAY_SCATTER && !(ro.w > 0.0f));
                        if (nothing) tl.has_result = true;   // stays in T2_FETCH: flushed at the next flush step
                        else tl.mode = T2_TRAV;
                    } else {
                        tl.mode = T2_DONE;
                    }
                }
                if (base + (unsigned int)__popc(mF) >= nT) exhausted = true;
            } else if (tl.mode == T2_FETCH) {
                tl.mode = T2_DONE;
            }
            __syncwarp();
            continue;
        }
        if (mT == 0) break;  // nothing traverses, nothing to flush or fetch: every lane is done
    }
#ifdef VR_WORK_COUNTERS
    WaveCounters cn = {0u, 0u, 0u, 0u, tl.n_nodes, tl.n_quads, tl.n_erf, 0u, 0u, 0u};
    wave_flush_counters(A, cn);
#endif
    if constexpr (GRID) {
        // the wave's list is consumed: the last block to finish empties it, because the post-A kernels of this iteration may
        // already append to it (slots sent back to the hierarchy); without the grid wave_post_a itself does that
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned int t = atomicAdd(&Q.cnt[WC_TICKET_S], 1u);
            if (t == gridDim.x - 1) { Q.cnt[WC_NTS] = 0; Q.cnt[WC_TICKET_S] = 0; __threadfence(); }
        }
    }
}

// ------------------------------------------------------------------------------------------
// post kernels: one thread per listed slot, grid-stride in whole warps so that the aggregated
// list appends at the end of each round are executed by converged warps
// ------------------------------------------------------------------------------------------
// draw the free-flight target on a copy of the stream (committed in post A only if the ray overlaps
// something, integrator.h:593-600) and publish the slot's next free-flight ray
__device__ __forceinline__ void publish_scatter_ray(const WavePool& Q, unsigned int p, f3 o, f3 d, uint64_t rng_state, float* target) {
    Pcg peek; peek.state = rng_state; peek.inc = 3;
    const float t = -logf(__fsub_rn(1.0f, peek.uniform()));
    *target = t;
    Q.ray[kSlotF4 * (size_t)p] = make_float4(o.x, o.y, o.z, t);
    Q.ray[kSlotF4 * (size_t)p + 1] = make_float4(d.x, d.y, d.z, __int_as_float(RAY_SCATTER));
}

// wave_regen: integrator.h:557-570
__global__ void __launch_bounds__(256) wave_regen_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q) {
    const unsigned int nC = Q.cnt[WC_NC];
    if (blockIdx.x == 0 && threadIdx.x == 0) Q.cnt[WC_NB] = 0;  // consumed by wave_post_b, the previous kernel of the chain
    WaveCounters cn = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
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
            const unsigned int work = wave_global_work(A, idx);
            lane_xy(A, work, &x, &y, &si);
            Pcg rng; rng.seed(derive_path_seed(x, y, si), 1);
            const int n = A.nsq;
            const int sx = si % n, sy = si / n;
            const float u = __fdiv_rn(__fadd_rn((float)x, __fdiv_rn(__fadd_rn((float)sx, rng.uniform()), (float)n)), (float)A.W);
            const float v = __fdiv_rn(__fadd_rn((float)y, __fdiv_rn(__fadd_rn((float)sy, rng.uniform()), (float)n)), (float)A.H);
            const RayOD ray = camera_ray(A.cam, u, v);
            float target;
            publish_scatter_ray(Q, p, ray.o, ray.d, rng.state, &target);
            Q.ps01[kSlotF4 * (size_t)p] = make_float4(0.f, 0.f, 0.f, 1.0f);
            reinterpret_cast<uint4*>(Q.ps01)[kSlotF4 * (size_t)p + 1] = make_uint4((unsigned int)rng.state, (unsigned int)(rng.state >> 32), work, 0u);
            ++cn.scatter;
            to_trav = true;
        }
        wave_append(Q.listTs, &Q.cnt[WC_NTS], to_trav, p);
    }
    wave_flush_counters(A, cn);
    // listC is refilled by the very next kernel (escaping rays in wave_trav), so its counter is cleared
    // here by the LAST block of this kernel to finish (every block has read nC by then)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned int t = atomicAdd(&Q.cnt[WC_TICKET], 1u);
        if (t == gridDim.x - 1) { Q.cnt[WC_NC] = 0; Q.cnt[WC_TICKET] = 0; __threadfence(); }
    }
}

// wave_regen_cam: wave_regen for a pinhole camera whose pinhole has a direction grid (vr_lightgrid_dev.cu
// build_camera_grid_device) -- the per-tile culling of the camera rays.  Every camera ray passes through the pinhole P
// (camera.h:45-53: origin on the film plane, direction towards P), so the components it can overlap beyond P are those of
// the cube-map cell of its direction, NEAREST FIRST, and those between film plane and P are in the cell of the opposite
// direction.  The lane that generates a camera ray therefore gathers its overlaps right here from that short sorted list
// (integrator.h:593, gmm.h:457-515; the quadric test and the pruning bound are wave_trav2's) instead of sending the ray
// through the hierarchy: the 32 lanes of a warp hold consecutive samples of ONE pixel, so they read the same list (one
// broadcast load per entry) and take the same branches.  A ray that overlaps nothing (background) ends its path on the
// spot and the lane takes the next sample.  Slots leave classified by overlap count, straight to wave_post_a.
// Lists average ~4 entries of 8 bytes: staging them with cp.async.bulk (16-byte granules, an mbarrier round trip) cannot
// pay; the broadcast loads hit L1.
__global__ void __launch_bounds__(256) wave_regen_cam_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q) {
    const SceneDev& S = A.scene;
    const unsigned int nC = Q.cnt[WC_NC];
    if (blockIdx.x == 0 && threadIdx.x == 0) Q.cnt[WC_NB] = 0;  // consumed by wave_post_b, the previous kernel of the chain
    WaveCounters cn = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    const int lane = threadIdx.x & 31;
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int rounds = (nC + stride - 1) / stride;
    const f3 P = mk3(A.cam.pinhole[0], A.cam.pinhole[1], A.cam.pinhole[2]);
    for (unsigned int r = 0; r < rounds; ++r) {
        const unsigned int i = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
        const unsigned int p = i < nC ? Q.listC[i] : 0u;
        bool need = i < nC;      // this lane's slot still needs a path
        int K_done = 0;          // overlaps of the path the slot got (0: none -- no samples left)
        while (__any_sync(0xffffffffu, need)) {
            // sample ids: one aggregated atomic per warp
            const unsigned m = __ballot_sync(0xffffffffu, need);
            unsigned int base = 0;
            if (lane == __ffs(m) - 1) base = atomicAdd(&Q.cnt[WC_NEXT_SAMPLE], (unsigned int)__popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            const unsigned int idx = base + (unsigned int)__popc(m & ((1u << lane) - 1u));
            if (need && idx >= A.work_count) need = false;
            if (need) {
                int x, y, si;
                const unsigned int work = wave_global_work(A, idx);
                lane_xy(A, work, &x, &y, &si);
                Pcg rng; rng.seed(derive_path_seed(x, y, si), 1);
                const int n = A.nsq;
                const int sx = si % n, sy = si / n;
                const float u = __fdiv_rn(__fadd_rn((float)x, __fdiv_rn(__fadd_rn((float)sx, rng.uniform()), (float)n)), (float)A.W);
                const float v = __fdiv_rn(__fadd_rn((float)y, __fdiv_rn(__fadd_rn((float)sy, rng.uniform()), (float)n)), (float)A.H);
                const RayOD ray = camera_ray(A.cam, u, v);
                ++cn.scatter;
                // ---- gather from the pinhole's grid
                Pcg peek; peek.state = rng.state; peek.inc = 3;
                const float target = -logf(__fsub_rn(1.0f, peek.uniform()));
                const f3 o = ray.o, d = ray.d;
                const float tP = (P.x - o.x) * d.x + (P.y - o.y) * d.y + (P.z - o.z) * d.z;   // the ray passes the pinhole at o + tP d
                const size_t cells = (size_t)6 * S.cg_res * S.cg_res;
                int K = 0;
                float tb = CUDART_INF_F;
#pragma unroll 1
                for (int side = 0; side < 2; ++side) {   // 0: between film plane and pinhole (opposite direction), 1: beyond the pinhole
                    const f3 vdir = side == 0 ? mk3(-d.x, -d.y, -d.z) : d;
                    const uint32_t cell = light_grid_cell_dev(vdir, S.cg_res);
                    (void)cells;
                    uint32_t pos = __ldg(S.cg_start + cell);
                    const uint32_t end = __ldg(S.cg_start + cell + 1);
                    for (; pos < end; ++pos) {
                        const float dm = __ldg(S.cg_dmin + pos);
                        // entries are sorted by their box's distance from the pinhole: behind it nothing farther than the film
                        // plane matters, beyond it nothing that starts behind the pruning bound (slack >> rounding of tP, dm)
                        if (side == 0 ? (dm > fmaf(tP, 1.00001f, 1e-5f)) : (tP + dm > fmaf(tb, 1.00001f, 1e-5f))) break;
                        const uint32_t gi = __ldg(S.cg_gi + pos);
                        const GaussRec g = load_gauss(S.gauss, gi);
                        const RayQuad q = ray_quad(g, o, d);
                        float t0, t1;
                        if (!quad_intersect(q, &t0, &t1) || t0 > tb) continue;
                        if (K < kWaveCap) pool_store_hit(Q, p, K, t0, t1, gi);
                        ++K;
                        if (t1 < tb) {   // pruning only: conservative estimate of this component's full optical depth (wave_trav2)
                            const float dn = g.density * g.norm;
                            const float inv4A = __fdividef(1.0f, 4.0f * q.A);
                            const float Kc = dn * sqrtf(6.2831853f * inv4A) * __expf(-0.5f * (q.C0 - q.B * q.B * inv4A));
                            const float iden = rsqrtf(8.0f * q.A);
                            const float full = Kc * (erff_ni((q.B + 2.0f * q.A * t1) * iden) - erff_ni((q.B + 2.0f * q.A * t0) * iden));
                            if (full > 1.5f * target) tb = t1;
                        }
                    }
                }
                if (K == 0) {
                    // nothing overlapped: the path ends with the environment (integrator.h:594-597) and the slot takes the next sample
                    A.sampleL[work] = make_float4(__fadd_rn(0.0f, __fmul_rn(1.0f, S.env[0])), __fadd_rn(0.0f, __fmul_rn(1.0f, S.env[1])),
                                                  __fadd_rn(0.0f, __fmul_rn(1.0f, S.env[2])), 0.0f);
                } else {
                    Q.ray[kSlotF4 * (size_t)p] = make_float4(o.x, o.y, o.z, target);
                    Q.ray[kSlotF4 * (size_t)p + 1] = make_float4(d.x, d.y, d.z, __int_as_float(RAY_SCATTER));
                    Q.ps01[kSlotF4 * (size_t)p] = make_float4(0.f, 0.f, 0.f, 1.0f);
                    reinterpret_cast<uint4*>(Q.ps01)[kSlotF4 * (size_t)p + 1] = make_uint4((unsigned int)rng.state, (unsigned int)(rng.state >> 32), work, 0u);
                    reinterpret_cast<int*>(Q.hit01 + kSlotF4 * (size_t)p)[0] = K;
                    if (S.g3_res) reinterpret_cast<float*>(Q.hit01 + kSlotF4 * (size_t)p + 1)[3] = CUDART_INF_F;   // certificate: the list is complete
                    K_done = K;
                    need = false;
                }
            }
        }
        const int cls = post_class(K_done);
#pragma unroll
        for (int c = 0; c < kPostClasses; ++c)
            wave_append(Q.listA + (size_t)c * Q.P, &Q.cnt[WC_NA + c], K_done > 0 && cls == c, p);
    }
    wave_flush_counters(A, cn);
    // listC is refilled by the next kernels (escaping rays in wave_trav, finished paths in the post kernels), so its counter is
    // cleared here by the LAST block of this kernel to finish (every block has read nC by then)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned int t = atomicAdd(&Q.cnt[WC_TICKET], 1u);
        if (t == gridDim.x - 1) { Q.cnt[WC_NC] = 0; Q.cnt[WC_TICKET] = 0; __threadfence(); }
    }
}

// wave_post_a: integrator.h:593-678.  Slots arrive binned by overlap count (post_class).
//   wave_post_a_kernel<1>      class 0 only (one overlap, the commonest case in sparse mixtures): a register-only
//       path that performs the same operations in the same order as the general walk does for K = 1.  Small code,
//       so the warps of the launch share the instruction caches (profiles/r01_v5: the one-kernel version spent 40 %
//       of its stall samples waiting for instructions).  A slot whose analytic inverse declines is handed to
//       class 1 untouched.
//   wave_post_a_kernel<0>      classes 1..4 (+ hand-overs): every class starts at a multiple of 32 in the flattened
//       index space, so a warp only ever holds slots of one class.  A slot whose walk ends in an ITERATIVE solve
//       (Newton / bisection, distance_solvers.h:15-129: ~3 % of the slots, but 40 % of this kernel's warp instructions
//       at 2-5 active lanes when solved in place, profiles/r01_v8) is left untouched and listed in class kPostClasses.
//   wave_post_a_kernel<2>      the deferred slots: the same walk again, now in warps whose lanes all iterate.
// environment / next-event part shared by both (integrator.h:652-678)
__device__ __forceinline__ int post_a_finish(const RenderArgs& A, const WavePool& Q, unsigned int p, float4 s0, uint4 s1, f3 o, f3 d,
                                             float target_tau, float t, float albedo, Pcg& rng, WaveCounters& cn) {
    const SceneDev& S = A.scene;
    if (t < 0.0f) {
        s0.x = __fadd_rn(s0.x, __fmul_rn(s0.w, S.env[0]));
        s0.y = __fadd_rn(s0.y, __fmul_rn(s0.w, S.env[1]));
        s0.z = __fadd_rn(s0.z, __fmul_rn(s0.w, S.env[2]));
        A.sampleL[s1.z] = make_float4(s0.x, s0.y, s0.z, (float)s1.w);
        return DEST_REGEN;
    }
    const f3 pos = add3(o, scale3(t, d));
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
    Q.ray[kSlotF4 * (size_t)p] = make_float4(pos.x, pos.y, pos.z, dist);
    Q.ray[kSlotF4 * (size_t)p + 1] = make_float4(wi.x, wi.y, wi.z, __int_as_float(RAY_SHADOW | ((light + 1) << 8)));
    reinterpret_cast<uint4*>(Q.ps01)[kSlotF4 * (size_t)p + 1] = make_uint4((unsigned int)stt, (unsigned int)(stt >> 32), s1.z, s1.w);
    Q.ps2[kSlotF4 * (size_t)p] = make_float4(target_tau, albedo, dist, __int_as_float(light));
    // a ray to a point light that has a direction grid takes its candidates from the grid (wave_trav_kernel<RAY_LIGHT>)
    return (light >= 0 && light < 32 && ((S.lg_mask >> light) & 1u)) ? DEST_TRAV_LIGHT : DEST_TRAV;
}

// Certificates of the uniform grid (wave_trav2_kernel<SCATTER, true>, grid_done).  A slot's overlap list is complete for what
// ENTERS before t_cert (+inf: complete as far as any walk can look).  The reference's walk (integrator.h:422-498) looks at the
// events up to and including the first one after the scatter point -- that event closes the segment [ta, tb] the solver works
// in -- and the recorded set at overlaps that enter up to t + 1e-6.  So a result computed from the certified list is the
// reference's if
//   * the closing event lies before t_cert (tb + 2e-6 < t_cert): every event the walk looked at is in the list; or
//   * one component is active at the scatter point, the analytic inverse (gaussian.h:235-297) gave t, and t + 2e-6 < t_cert: an
//     overlap the list lacks enters at t_e >= t_cert, so the reference's segment ends at min(tb, t_e) > t; it holds the target
//     (grid_done: the kept overlaps hold 1.5 x the target before t_cert), has the same start, the same active component and the
//     same remaining depth, and the analytic inverse uses tb only to clamp results beyond it.
// Anything else -- an iterative solve whose segment reaches t_cert, a ray that leaves the kept overlaps without scattering -- is
// SENT BACK: the slot returns untouched to the traversal list with kRayNoGrid set and gets its list from the hierarchy in the next
// iteration (3 % of the slots at C2).
__device__ __forceinline__ float slot_certificate(const SceneDev& S, const WavePool& Q, unsigned int p) {
    return S.g3_res ? reinterpret_cast<const float*>(Q.hit01 + kSlotF4 * (size_t)p + 1)[3] : CUDART_INF_F;
}
__device__ __forceinline__ bool before_certificate(float t, float t_cert) { return t >= 0.0f && __fadd_rn(t, 2e-6f) < t_cert; }
__device__ __forceinline__ void slot_send_back(const WavePool& Q, unsigned int p) {
    reinterpret_cast<int*>(Q.ray + kSlotF4 * (size_t)p + 1)[3] = RAY_SCATTER | kRayNoGrid;
}

// MODE: 0 general (defers iterative solves), 1 fast K = 1, 2 deferred slots (walk redone), 3 general, solves in place,
//       4 deferred slots resumed from the state MODE 0 left in the slot: the segment [ta, tb], the remaining depth and the
//       active list in the reference's order (ps2 is free between the free-flight ray and the shadow ray).
// CLS (general modes only): -1 = classes 1..4 in one launch with a 16-entry overlap store; 1..4 = that class alone with a store
// sized for it (2 / 4 / 8 / 16 entries: the store lives in local memory, 40 bytes per entry and thread)
template <int MODE, int CLS = -1>
__global__ void VR_POSTA_BOUNDS wave_post_a_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q) {
    const SceneDev& S = A.scene;
    constexpr bool FAST = MODE == 1;
    constexpr bool PAIR = MODE == 5;
    constexpr bool kOneClass = (MODE == 0 || MODE == 3) && CLS > 0;
    constexpr bool kDeferred = MODE == 2 || MODE == 4;
    constexpr int kFirst = MODE == 1 ? 0 : (kDeferred ? kPostClasses : (kOneClass ? CLS : (MODE == 6 ? 2 : 1)));              // classes [kFirst, kLast)
    constexpr int kLast = MODE == 1 ? 1 : (PAIR ? 2 : (kDeferred ? kPostClasses + 1 : (kOneClass ? CLS + 1 : kPostClasses)));
    constexpr int kStoreCap = !kOneClass ? kWaveCap : (CLS == 1 ? 2 : (CLS == 2 ? 4 : (CLS == 3 ? 8 : kWaveCap)));
    unsigned int nA[kPostClasses + 1], base[kPostClasses + 2];
    base[kFirst] = 0;
#pragma unroll
    for (int c = kFirst; c < kLast; ++c) { nA[c] = Q.cnt[WC_NA + c]; base[c + 1] = base[c] + ((nA[c] + 31u) & ~31u); }
    const unsigned int total = base[kLast];
    // (with the uniform grid the traversal kernel has emptied its list itself: slots may be sent back to it from here on)
    if ((MODE == 0 || MODE == 3 || MODE == 6) && (!kOneClass || CLS == 1) && !S.g3_res && blockIdx.x == 0 && threadIdx.x == 0) Q.cnt[WC_NTS] = 0;
    if (MODE != 1 && blockIdx.x == 0 && threadIdx.x == 0) {   // the host decides from these whether deferring pays (run_wavefront_pass)
        unsigned int sum = 0;
#pragma unroll
        for (int c = kFirst; c < kLast; ++c) sum += nA[c];
        Q.cnt[kDeferred ? WC_SUM_DEFERRED : WC_SUM_GENERAL] += sum;
    }  // consumed by wave_trav<scatter>
    WaveCounters cn = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int rounds = (total + stride - 1) / stride;
    for (unsigned int r = 0; r < rounds; ++r) {
        const unsigned int i = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
        int dest = DEST_NONE;
        bool hand_over = false, deferred = false, send_back = false;
        unsigned int p = 0;
        int cls = kFirst;
#pragma unroll
        for (int c = kFirst + 1; c < kLast; ++c) cls += (i >= base[c]) ? 1 : 0;   // (no iterations for the one-class kernels)
        unsigned int j = i, nc = nA[kFirst];
#pragma unroll
        for (int c = kFirst + 1; c < kLast; ++c) if (cls == c) { j = i - base[c]; nc = nA[c]; }
        if (i < total && j < nc) {
            p = Q.listA[(size_t)cls * Q.P + j];
            const uint4 h0 = Q.hit01[kSlotF4 * (size_t)p];   // overlap count + first overlap
            const int K = (int)h0.x;
            const float4 s0 = Q.ps01[kSlotF4 * (size_t)p];
            const uint4 s1 = reinterpret_cast<uint4*>(Q.ps01)[kSlotF4 * (size_t)p + 1];
            const float4 ro = Q.ray[kSlotF4 * (size_t)p], rd = Q.ray[kSlotF4 * (size_t)p + 1];
            const f3 o = mk3(ro.x, ro.y, ro.z), d = mk3(rd.x, rd.y, rd.z);
            const float target_tau = ro.w;
            Pcg rng; rng.state = (uint64_t)s1.x | ((uint64_t)s1.y << 32); rng.inc = 3;
            rng.next_u32();  // commit the target draw (every listed slot overlapped something)
            const float t_cert = MODE == 4 ? CUDART_INF_F : slot_certificate(S, Q, p);
            if (FAST) {
                // ---- one overlap: enter event (empty segment), exit event with the single active component,
                //      analytic inverse when the target falls inside (integrator.h:422-498, distance_solvers.h:163-171)
                float t = -1.0f, albedo = 0.0f;
                bool solved = false;
                if (K == 1 && A.solver <= 1) {
                    const uint32_t gi = h0.w;
                    const float t0 = __uint_as_float(h0.y), t1 = __uint_as_float(h0.z);
                    const GaussRec g = load_gauss(S.gauss, gi);
                    const RayQuad q = ray_quad(g, o, d);
                    const DepthConst c = depth_const(g, q);
                    const float F0 = erff_ni(erf_arg(c, t0));
                    const float F1 = erff_ni(erf_arg(c, t1));
                    const float seg = __fmul_rn(c.K, __fsub_rn(F1, F0));
                    VR_COUNT(cn.erf += 2u;)
                    if (seg > target_tau) {   // float and double comparisons agree: the conversion is exact
                        float t_an;
                        VR_COUNT(++cn.sa;)
                        if (solve_t_given_tau(g, q, t0, t1, target_tau, &t_an)) { t = clampf(t_an, t0, t1); solved = true; }
                    } else {
                        solved = true;        // the ray leaves the component: t stays -1
                    }
                    if (solved && t_cert < CUDART_INF_F && !before_certificate(t, t_cert)) { solved = false; send_back = true; }
                    if (solved) {
                        if (t >= 0.0f) {       // evaluate_albedo over the one active component (gmm.h:128-143)
                            const float mu = gauss_mu_t(g, add3(o, scale3(t, d)));
                            VR_COUNT(++cn.mu;)
                            const float a = __fdiv_rn(__fadd_rn(0.0f, __fmul_rn(mu, g.albedo)), __fadd_rn(0.0f, mu));
                            albedo = a < 0.0f ? 0.0f : (1.0f < a ? 1.0f : a);
                        }
                        if (A.record_bits && (t < 0.0f || t0 <= __fadd_rn(t, 1e-6f))) {  // integrator.h:616-644
                            int x, y, si;
                            lane_xy(A, s1.z, &x, &y, &si);
                            uint32_t* row = A.record_bits + ((size_t)y * A.W + x) * (size_t)A.record_words;
                            const uint32_t orig = __ldg(S.orig + gi);
                            const uint32_t bit = 1u << (orig & 31u);
                            if (!(row[orig >> 5] & bit)) atomicOr(row + (orig >> 5), bit);
                        }
                    }
                }
                if (solved) {
                    cn.hits += 1u;
                    dest = post_a_finish(A, Q, p, s0, s1, o, d, target_tau, t, albedo, rng, cn);
                } else if (!send_back) {
                    hand_over = true;   // untouched: the general kernel redoes it
                }
            } else if (PAIR) {
                // ---- two overlaps, register only: the general walk (free_flight, integrator.h:422-498 / :333-361) written out
                //      for K = 2.  X = the component entered first, Y = the other; the active list is {X}, {X, Y} (in the
                //      reference's list order), {the one left}, or, when they do not overlap each other, {X}, {}, {Y}.  The
                //      same operations in the same order as the general walk; a target that falls where both are active needs
                //      an iterative solve and is deferred with its state like there; anything unusual goes to the general kernel.
                if (K == 2 && A.solver <= 1 && target_tau >= 0.0f) {
                    const uint4 h1 = Q.hit01[kSlotF4 * (size_t)p + 1];
                    const bool ss = A.integrator == 0;
                    uint32_t x_gi = h0.w, y_gi = h1.z;
                    float x_t0 = __uint_as_float(h0.y), x_t1 = __uint_as_float(h0.z), y_t0 = __uint_as_float(h1.x), y_t1 = __uint_as_float(h1.y);
                    uint32_t x_or = __ldg(S.orig + x_gi), y_or = __ldg(S.orig + y_gi);
                    unsigned int x_pool = 0u, y_pool = 1u;
                    if (!((x_t0 != y_t0) ? (x_t0 < y_t0) : (x_or < y_or))) {   // entry order (sort_entries)
                        uint32_t tu; float tf;
                        tu = x_gi; x_gi = y_gi; y_gi = tu;  tu = x_or; x_or = y_or; y_or = tu;
                        tf = x_t0; x_t0 = y_t0; y_t0 = tf;  tf = x_t1; x_t1 = y_t1; y_t1 = tf;
                        x_pool = 1u; y_pool = 0u;
                    }
                    const GaussRec gx = load_gauss(S.gauss, x_gi);
                    const DepthConst cx = depth_const(gx, ray_quad(gx, o, d));
                    const GaussRec gy = load_gauss(S.gauss, y_gi);
                    const DepthConst cy = depth_const(gy, ray_quad(gy, o, d));
                    VR_COUNT(unsigned int n_erf = 0u;)
                    double acc = 0.0; float accf = 0.0f;
                    // what the walk ends with: 0 left the mixture, 1 analytic solve on `sel` in [ta, tb], 2 both active -> deferred
                    int outcome = 0; int sel = 0; float ta = 0.0f, tb = 0.0f, remaining = 0.0f;
                    // one segment [t_prev, t_evt] with one active component (constants c, erf at t_prev in *F): true = target inside
                    auto segment1 = [&](const DepthConst& c, float* F, float t_evt) -> bool {
                        const float Fn = erff_ni(erf_arg(c, t_evt));
                        VR_COUNT(++n_erf;)
                        const float term = __fmul_rn(c.K, __fsub_rn(Fn, *F));
                        *F = Fn;
                        if (ss) {
                            const float seg = __fadd_rn(0.0f, term);
                            remaining = __fsub_rn(target_tau, accf);
                            if (__fadd_rn(accf, seg) > target_tau) return true;
                            accf = __fadd_rn(accf, seg);
                        } else {
                            const double seg = __dadd_rn(0.0, (double)term);
                            remaining = (float)__dsub_rn((double)target_tau, acc);
                            if (__dadd_rn(acc, seg) > (double)target_tau) return true;
                            acc = __dadd_rn(acc, seg);
                        }
                        return false;
                    };
                    // (the empty segment before X enters cannot hold the target: 0 > target_tau is false for target_tau >= 0)
                    float xF = erff_ni(erf_arg(cx, x_t0)), yF = 0.0f;
                    VR_COUNT(++n_erf;)
                    bool take_enter = true;   // next_event: Y's entry against X's exit
                    if (x_t1 < y_t0) take_enter = false;
                    else if (x_t1 == y_t0) take_enter = y_or <= x_or;
                    if (take_enter) {
                        if (segment1(cx, &xF, y_t0)) { outcome = 1; sel = 0; ta = x_t0; tb = y_t0; }
                        else {
                            yF = erff_ni(erf_arg(cy, y_t0));
                            VR_COUNT(++n_erf;)
                            // active list: multi scatter appends (X, Y); single scatter keeps ascending original index
                            const bool x_first = !ss || x_or < y_or;
                            // earliest exit among the two, scanned in list order (ties: the lower original index)
                            const float f_t1 = x_first ? x_t1 : y_t1, s_t1 = x_first ? y_t1 : x_t1;
                            const uint32_t f_or = x_first ? x_or : y_or, s_or = x_first ? y_or : x_or;
                            const bool second_exits = s_t1 < f_t1 || (s_t1 == f_t1 && s_or < f_or);
                            const bool x_exits = x_first ? !second_exits : second_exits;
                            const float t_evt = x_exits ? x_t1 : y_t1;
                            const float Fxn = erff_ni(erf_arg(cx, t_evt)), Fyn = erff_ni(erf_arg(cy, t_evt));
                            VR_COUNT(n_erf += 2u;)
                            const float term_x = __fmul_rn(cx.K, __fsub_rn(Fxn, xF)), term_y = __fmul_rn(cy.K, __fsub_rn(Fyn, yF));
                            const float term_f = x_first ? term_x : term_y, term_s = x_first ? term_y : term_x;
                            bool hit;
                            if (ss) {
                                const float seg = __fadd_rn(__fadd_rn(0.0f, term_f), term_s);
                                remaining = __fsub_rn(target_tau, accf);
                                hit = __fadd_rn(accf, seg) > target_tau;
                                if (!hit) accf = __fadd_rn(accf, seg);
                            } else {
                                const double seg = __dadd_rn(__dadd_rn(0.0, (double)term_f), (double)term_s);
                                remaining = (float)__dsub_rn((double)target_tau, acc);
                                hit = __dadd_rn(acc, seg) > (double)target_tau;
                                if (!hit) acc = __dadd_rn(acc, seg);
                            }
                            if (hit) { outcome = 2; sel = x_first ? 0 : 1; ta = y_t0; tb = t_evt; }
                            else {
                                xF = Fxn; yF = Fyn;
                                if (x_exits) { if (segment1(cy, &yF, y_t1)) { outcome = 1; sel = 1; ta = t_evt; tb = y_t1; } }
                                else         { if (segment1(cx, &xF, x_t1)) { outcome = 1; sel = 0; ta = t_evt; tb = x_t1; } }
                            }
                        }
                    } else {
                        if (segment1(cx, &xF, x_t1)) { outcome = 1; sel = 0; ta = x_t0; tb = x_t1; }
                        else {
                            // (the empty segment [x_t1, y_t0]: acc > target_tau is false, acc + seg was not above it)
                            yF = erff_ni(erf_arg(cy, y_t0));
                            VR_COUNT(++n_erf;)
                            if (segment1(cy, &yF, y_t1)) { outcome = 1; sel = 1; ta = y_t0; tb = y_t1; }
                        }
                    }
                    float t = -1.0f;
                    bool solved = outcome == 0;
                    if (outcome == 1) {   // solve_distance with one active component: the analytic inverse (distance_solvers.h:163-171)
                        const GaussRec& g = sel == 0 ? gx : gy;
                        float t_an;
                        VR_COUNT(++cn.sa;)
                        if (solve_t_given_tau(g, ray_quad(g, o, d), ta, tb, remaining, &t_an)) { t = clampf(t_an, ta, tb); solved = true; }
                    }
                    VR_COUNT(cn.erf += n_erf;)
                    // certificate: a result needs t before it, a deferred (iterative) solve its whole segment
                    if (t_cert < CUDART_INF_F && !before_certificate(solved ? t : tb, t_cert)) {
                        send_back = true;
                    } else if (solved) {
                        cn.hits += 2u;
                        float albedo = 0.0f;
                        if (t >= 0.0f) {       // evaluate_albedo over the one active component (gmm.h:128-143)
                            const GaussRec& g = sel == 0 ? gx : gy;
                            const float mu = gauss_mu_t(g, add3(o, scale3(t, d)));
                            VR_COUNT(++cn.mu;)
                            const float a = __fdiv_rn(__fadd_rn(0.0f, __fmul_rn(mu, g.albedo)), __fadd_rn(0.0f, mu));
                            albedo = a < 0.0f ? 0.0f : (1.0f < a ? 1.0f : a);
                        }
                        if (A.record_bits) {  // integrator.h:616-644
                            int x, y, si;
                            lane_xy(A, s1.z, &x, &y, &si);
                            uint32_t* row = A.record_bits + ((size_t)y * A.W + x) * (size_t)A.record_words;
                            const float tlim = __fadd_rn(t, 1e-6f);
                            if (t < 0.0f || x_t0 <= tlim) { const uint32_t bit = 1u << (x_or & 31u); if (!(row[x_or >> 5] & bit)) atomicOr(row + (x_or >> 5), bit); }
                            if (t < 0.0f || y_t0 <= tlim) { const uint32_t bit = 1u << (y_or & 31u); if (!(row[y_or >> 5] & bit)) atomicOr(row + (y_or >> 5), bit); }
                        }
                        dest = post_a_finish(A, Q, p, s0, s1, o, d, target_tau, t, albedo, rng, cn);
                    } else {   // an iterative solve: the state wave_post_a_kernel<4> resumes from (see MODE 0 below)
                        deferred = true;
                        unsigned int order, na;
                        if (outcome == 2) { na = 2u; order = sel == 0 ? (x_pool | (y_pool << 4)) : (y_pool | (x_pool << 4)); }
                        else { na = 1u; order = sel == 0 ? x_pool : y_pool; }
                        Q.ps2[kSlotF4 * (size_t)p] = make_float4(__uint_as_float(order), __uint_as_float(0u), 0.f, 0.f);
                        Q.ps2[kSlotF4 * (size_t)p + 1] = make_float4(ta, tb, remaining, __int_as_float((int)na));
                    }
                } else {
                    hand_over = true;   // untouched: the general kernel does it
                }
            } else if (MODE == 4) {
                // ---- resume an iterative solve (distance_solvers.h:25-127) from the saved state: only the active components
                //      are loaded, in list order; constants and erf(arg(ta)) are recomputed from the same inputs, bit for bit
                const float4 d0 = Q.ps2[kSlotF4 * (size_t)p], d1 = Q.ps2[kSlotF4 * (size_t)p + 1];
                const uint64_t order = (uint64_t)__float_as_uint(d0.x) | ((uint64_t)__float_as_uint(d0.y) << 32);
                const float ta = d1.x, tb = d1.y, remaining = d1.z;
                const int na = __float_as_int(d1.w);
                cn.hits += (unsigned int)K;
                LocalStore<kWaveCap> st;
                VR_COUNT(st.n_erf += (unsigned int)na;)
#pragma unroll 1
                for (int k = 0; k < na; ++k) {
                    float ht0, ht1; uint32_t gi;
                    pool_load_hit(Q, p, (int)((order >> (4 * k)) & 15u), h0, &ht0, &ht1, &gi);
                    const GaussRec g = load_gauss(S.gauss, gi);
                    const DepthConst c = depth_const(g, ray_quad(g, o, d));
                    st.F(F_A, k) = c.A; st.F(F_B, k) = c.B; st.F(F_K, k) = c.K; st.F(F_DEN, k) = c.den;
                    st.F(F_ACTF, k) = erff_ni(erf_arg(c, ta));
                    st.U(U_GI, k) = gi; st.U(U_ACT, k) = (uint32_t)k;
                }
                const float t = (A.solver == 0 || A.solver == 2) ? solve_newton(st, na, ta, tb, remaining) : solve_bisection(st, na, ta, tb, remaining);
                if (A.record_bits) {  // integrator.h:616-644 (t >= 0 here)
                    int x, y, si;
                    lane_xy(A, s1.z, &x, &y, &si);
                    uint32_t* row = A.record_bits + ((size_t)y * A.W + x) * (size_t)A.record_words;
                    const float tlim = __fadd_rn(t, 1e-6f);
                    for (int s = 0; s < K; ++s) {
                        float ht0, ht1; uint32_t gi;
                        pool_load_hit(Q, p, s, h0, &ht0, &ht1, &gi);
                        if (t < 0.0f || ht0 <= tlim) {
                            const uint32_t orig = __ldg(S.orig + gi);
                            const uint32_t bit = 1u << (orig & 31u);
                            if (!(row[orig >> 5] & bit)) atomicOr(row + (orig >> 5), bit);
                        }
                    }
                }
                const float albedo = t >= 0.0f ? evaluate_albedo(S, st, na, add3(o, scale3(t, d))) : 0.0f;
                dest = post_a_finish(A, Q, p, s0, s1, o, d, target_tau, t, albedo, rng, cn);
                VR_COUNT(cn.erf += st.n_erf; cn.mu += st.n_mu; cn.sa += st.n_sa; cn.it += st.n_it;)
            } else if (K > kWaveCap) {  // more overlaps than a slot holds: the whole sample is re-traced by the arena kernel
                overflow_push(A, s1.z + A.work_base);
                ++cn.over;
                dest = DEST_REGEN;
            } else {
                cn.hits += (unsigned int)K;
                LocalStore<kStoreCap> st;
#pragma unroll 1
                for (int s = 0; s < K; ++s) {  // exact depth constants (gaussian.h:208-231) of the gathered overlaps
                    float ht0, ht1; uint32_t gi;
                    pool_load_hit(Q, p, s, h0, &ht0, &ht1, &gi);
                    const GaussRec g = load_gauss(S.gauss, gi);
                    const DepthConst c = depth_const(g, ray_quad(g, o, d));
                    st.F(F_T0, s) = ht0; st.F(F_T1, s) = ht1;
                    st.F(F_A, s) = c.A; st.F(F_B, s) = c.B; st.F(F_K, s) = c.K; st.F(F_DEN, s) = c.den;
                    st.U(U_GI, s) = gi;
                }
                int na = 0;
                DeferState ds;
                const float t = free_flight<MODE == 0 || MODE == 6>(S, st, K, o, d, target_tau, A.solver, A.integrator == 0, &na, rng.state, &deferred, &ds);
                bool cert_ok = true;
                if (t_cert < CUDART_INF_F)
                    cert_ok = before_certificate((!deferred && A.solver <= 1 && na == 1) || t < 0.0f ? t : ds.tb, t_cert);
                if (!cert_ok) {
                    cn.hits -= (unsigned int)K;
                    deferred = false;   // (neither finished nor deferred: the slot is untouched)
                    send_back = true;
                } else if (deferred) {
                    cn.hits -= (unsigned int)K;   // counted when the slot is finished
                    uint64_t order = 0;            // the active list in the reference's order, four bits per entry
                    for (int k = 0; k < na; ++k) order |= (uint64_t)(st.U(U_ACT, k) & 15u) << (4 * k);
                    Q.ps2[kSlotF4 * (size_t)p] = make_float4(__uint_as_float((unsigned int)order), __uint_as_float((unsigned int)(order >> 32)), 0.f, 0.f);
                    Q.ps2[kSlotF4 * (size_t)p + 1] = make_float4(ds.ta, ds.tb, ds.remaining, __int_as_float(na));
                } else {
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
                const float albedo = t >= 0.0f ? evaluate_albedo(S, st, na, add3(o, scale3(t, d))) : 0.0f;
                dest = post_a_finish(A, Q, p, s0, s1, o, d, target_tau, t, albedo, rng, cn);
                }
                VR_COUNT(cn.erf += st.n_erf; cn.mu += st.n_mu; cn.sa += st.n_sa; cn.it += st.n_it;)
            }
        }
        if (MODE != 4 && MODE != 2) {   // certificate not sufficient: back to the traversal list, through the hierarchy this time
            if (send_back) slot_send_back(Q, p);
            wave_append(Q.listTs, &Q.cnt[WC_NTS], send_back, p);
        }
        if (FAST) wave_append(Q.listA + (size_t)1 * Q.P, &Q.cnt[WC_NA + 1], hand_over, p);
        if (PAIR) wave_append(Q.listA + (size_t)2 * Q.P, &Q.cnt[WC_NA + 2], hand_over, p);
        if (MODE == 0 || MODE == 6 || PAIR) wave_append(Q.listA + (size_t)kPostClasses * Q.P, &Q.cnt[WC_NA + kPostClasses], deferred, p);
        wave_append(Q.listTh, &Q.cnt[WC_NTH], dest == DEST_TRAV, p);
        wave_append(Q.listTl, &Q.cnt[WC_NTL], dest == DEST_TRAV_LIGHT, p);
        wave_append(Q.listC, &Q.cnt[WC_NC], dest == DEST_REGEN, p);
    }
    wave_flush_counters(A, cn);
}

// wave_post_b: integrator.h:669-699
__global__ void __launch_bounds__(256) wave_post_b_kernel(const __grid_constant__ RenderArgs A, const __grid_constant__ WavePool Q) {
    const SceneDev& S = A.scene;
    const unsigned int nB = Q.cnt[WC_NB];
    if (blockIdx.x == 0 && threadIdx.x == 0) {  // both traversal kernels of this iteration are over
        Q.cnt[WC_NTH] = 0; Q.cnt[WC_NTL] = 0; Q.cnt[WC_CURSOR_S] = 0; Q.cnt[WC_CURSOR_H] = 0; Q.cnt[WC_CURSOR_L] = 0;
    }
    WaveCounters cn = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int rounds = (nB + stride - 1) / stride;
    for (unsigned int r = 0; r < rounds; ++r) {
        const unsigned int i = r * stride + blockIdx.x * blockDim.x + threadIdx.x;
        int dest = DEST_NONE;
        unsigned int p = 0;
        if (i < nB) {
            p = Q.listB[i];
            float4 s0 = Q.ps01[kSlotF4 * (size_t)p];
            uint4 s1 = reinterpret_cast<uint4*>(Q.ps01)[kSlotF4 * (size_t)p + 1];
            const float4 s2 = Q.ps2[kSlotF4 * (size_t)p];
            const float albedo = s2.y, nee_dist = s2.z;
            const int light = __float_as_int(s2.w);
            const long long tf = reinterpret_cast<const long long*>(Q.ps2 + kSlotF4 * (size_t)p + 1)[0];
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
                    const float4 ro = Q.ray[kSlotF4 * (size_t)p];  // the scatter position is the shadow ray's origin
                    float target;
                    publish_scatter_ray(Q, p, mk3(ro.x, ro.y, ro.z), nd, stt, &target);
                    Q.ps01[kSlotF4 * (size_t)p] = s0;
                    reinterpret_cast<uint4*>(Q.ps01)[kSlotF4 * (size_t)p + 1] = make_uint4((unsigned int)stt, (unsigned int)(stt >> 32), s1.z, s1.w + 1u);
                    ++cn.scatter;
                    dest = DEST_TRAV;
                }
            }
        }
        wave_append(Q.listTs, &Q.cnt[WC_NTS], dest == DEST_TRAV, p);
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
// known-answer hook: wave_trav_kernel<RAY_SCATTER> itself (quantised nodes, candidate queue, batched quadric tests) on a
// batch of caller rays with an infinite sampled depth, i.e. nothing is pruned: the overlap SET of every ray
// ------------------------------------------------------------------------------------------
__global__ void kat_gather_init_kernel(const __grid_constant__ WavePool Q, unsigned int n, const float* __restrict__ o3, const float* __restrict__ d3) {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < kWaveCounters) Q.cnt[i] = (i == WC_NTS) ? n : 0u;
    if (i >= n) return;
    Q.ray[kSlotF4 * (size_t)i] = make_float4(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2], CUDART_INF_F);
    Q.ray[kSlotF4 * (size_t)i + 1] = make_float4(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2], __int_as_float(RAY_SCATTER));
    Q.ps01[kSlotF4 * (size_t)i] = make_float4(0.f, 0.f, 0.f, 1.0f);
    reinterpret_cast<uint4*>(Q.ps01)[kSlotF4 * (size_t)i + 1] = make_uint4(0u, 0u, i, 0u);
    Q.hit01[kSlotF4 * (size_t)i] = make_uint4(0u, 0u, 0u, 0u);
    Q.listTs[i] = i;
}
__global__ void kat_gather_extract_kernel(const __grid_constant__ WavePool Q, unsigned int n, const uint32_t* __restrict__ orig, int cap,
                                          int* __restrict__ counts, uint32_t* __restrict__ idx) {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 h0 = Q.hit01[kSlotF4 * (size_t)i];
    const int K = (int)h0.x;
    counts[i] = K;
    for (int k = 0; k < K && k < kWaveCap && k < cap; ++k) {
        float t0, t1; uint32_t gi;
        pool_load_hit(Q, i, k, h0, &t0, &t1, &gi);
        idx[(size_t)i * cap + k] = __ldg(orig + gi);
    }
}
int run_wave_gather_kat(const SceneDev& S, int n, const float* d_o3, const float* d_d3, void* pool_mem, float4* sample_scratch,
                        unsigned long long* stats, int cap, int* d_counts, uint32_t* d_idx, int num_sms, cudaStream_t s) {
    WavePool Q;
    wave_pool_carve(&Q, n, pool_mem);
    RenderArgs A{};
    A.scene = S; A.scene.lg_mask = 0;
    A.sampleL = sample_scratch; A.stats = stats;
    A.W = A.H = A.spp = A.nsq = 1; A.win_w = 1; A.part_count = 1; A.work_count = (unsigned int)n;
    kat_gather_init_kernel<<<(std::max(n, (int)kWaveCounters) + 255) / 256, 256, 0, s>>>(Q, (unsigned int)n, d_o3, d_d3);
    int bps = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, wave_trav_kernel<RAY_SCATTER>, kBlockThreads, 0);
    const int grid = std::max(1, std::min(num_sms * std::max(bps, 1), (n + kBlockThreads - 1) / kBlockThreads));
    wave_trav_kernel<RAY_SCATTER><<<grid, kBlockThreads, 0, s>>>(A, Q);
    kat_gather_extract_kernel<<<(n + 255) / 256, 256, 0, s>>>(Q, (unsigned int)n, S.orig, cap, d_counts, d_idx);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// ------------------------------------------------------------------------------------------
// host driver of one pass
// ------------------------------------------------------------------------------------------
// One pass = a.work_count samples.  The pass is cut into `parts` interleaved sets of pixel rows, each with its own pool of slots, counters
// and stream; the kernel chains of the parts are independent, so the tail of one part's kernel (persistent warps waiting
// for the last rays, 9-15 % of SM cycles per kernel in profiles/r01_v7) is filled by the other part's kernels.
// A graph is reusable when every kernel argument and grid of the batch is byte-identical (repeated renders of one
// configuration: the bench loop, the S + 1 renders per view of every inverse iteration).
struct WaveGraphKey { RenderArgs a; WavePool q; int grid[6]; int defer; };
struct WaveGraphCache {
    struct Entry { bool valid; WaveGraphKey key; cudaGraphExec_t exec; } e[kMaxWaveParts][2];
};
WaveGraphCache* wave_graph_cache_create() {
    WaveGraphCache* c = new WaveGraphCache;
    memset((void*)c, 0, sizeof(*c));
    return c;
}
void wave_graph_cache_destroy(WaveGraphCache* c) {
    if (!c) return;
    for (auto& row : c->e) for (auto& en : row) if (en.exec) cudaGraphExecDestroy(en.exec);
    delete c;
}

int run_wavefront_pass(const RenderArgs& a_all, void* pool_mem, int P_total, int parts, int num_sms, unsigned int* h_cnt_all,
                       cudaStream_t s, cudaStream_t* part_streams, unsigned int* launches, WaveGraphCache* graphs) {
    // resident blocks per SM of the persistent traversal kernels, per kernel version (computed once, identical on every device)
    static std::once_flag occ_once;
    static int occ[2][3], occ_grid;
    std::call_once(occ_once, [] {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0][0], wave_trav_kernel<RAY_SCATTER>, kBlockThreads, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0][1], wave_trav_kernel<RAY_SHADOW>, kBlockThreads, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0][2], wave_trav_kernel<RAY_LIGHT>, kBlockThreads, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1][0], wave_trav2_kernel<RAY_SCATTER>, kBlockThreads, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1][1], wave_trav2_kernel<RAY_SHADOW>, kBlockThreads, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1][2], wave_trav2_kernel<RAY_LIGHT>, kBlockThreads, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_grid, wave_trav2_kernel<RAY_SCATTER, true>, kBlockThreads, 0);
        occ_grid = std::max(occ_grid, 1);
        for (auto& row : occ) for (int& v : row) v = std::max(v, 1);
    });
    const bool t2 = a_all.trav_version != 1;
    const bool post_split = a_all.post_split != 0;
    const bool pair = a_all.post_defer == 2 && a_all.post_pair != 0 && a_all.post_split == 0;
    const bool resume = a_all.post_defer == 2;   // deferred solves resumed from the state in the slot (no second walk)
    const int bps_s = occ[t2 ? 1 : 0][0], bps_h = occ[t2 ? 1 : 0][1], bps_l = occ[t2 ? 1 : 0][2];
    const char* prof_env = getenv("VR_WAVE_PROFILE");
    const bool prof = prof_env && prof_env[0] == '1';
    const unsigned int pass_pixels = a_all.work_count / (unsigned int)a_all.spp;
    parts = std::max(1, std::min(parts, kMaxWaveParts));
    if (prof || pass_pixels < (unsigned int)parts) parts = 1;
    struct Part {
        RenderArgs a; WavePool q; cudaStream_t st; cudaEvent_t ev; unsigned int* h; bool done, defer;
        int grid_s, grid_sg, grid_h, grid_l, grid_post, grid_post256;
    } part[kMaxWaveParts];
    static const bool graph_env = !(getenv("VR_WAVE_GRAPH") && getenv("VR_WAVE_GRAPH")[0] == '0');
    const bool use_graph = graph_env && !prof && graphs != nullptr;   // the profile mode synchronises inside an iteration
    static const bool timeline = getenv("VR_WAVE_TIMELINE") != nullptr;   // stderr: per batch (part, ms since pass start, live rays, samples issued)
    const auto tl0 = std::chrono::steady_clock::now();
    static const bool drain_on = !(getenv("VR_WAVE_DRAIN") && getenv("VR_WAVE_DRAIN")[0] == '0');
    static const unsigned int drain_div = getenv("VR_WAVE_DRAIN_DIV") ? (unsigned int)std::max(1, atoi(getenv("VR_WAVE_DRAIN_DIV"))) : 8u;   // hand-over below P / drain_div live paths
    static const unsigned int drain_max = getenv("VR_WAVE_DRAIN_MAX") ? (unsigned int)std::max(1, atoi(getenv("VR_WAVE_DRAIN_MAX"))) : 32768u;   // ... and below this many: per bounce the drain kernel is ~10x slower than the wave kernels, so it only gets the stragglers (profiles/r02/drain_sweep.md)
    static const int defer_env = getenv("VR_WAVE_DEFER") ? atoi(getenv("VR_WAVE_DEFER")) : -1;   // 0 / 1 force, default adaptive
    // the parts own interleaved rows of the pass (neighbouring rows cost the same, halves of an image do not: with
    // contiguous halves one part of a C2 pass ran alone for its last 8 %)
    unsigned int chunk = (unsigned int)std::max(1, a_all.win_w);
    if ((pass_pixels + chunk - 1u) / chunk < 4u * (unsigned int)parts) chunk = std::max(1u, pass_pixels / (4u * (unsigned int)parts));
    const unsigned int n_chunks = (pass_pixels + chunk - 1u) / chunk;
    const int P_part = P_total / parts;
    const size_t pool_stride = wave_pool_bytes(P_part);
    cudaEvent_t fork;
    if (cudaEventCreateWithFlags(&fork, cudaEventDisableTiming) != cudaSuccess) return -1;
    if (parts > 1) cudaEventRecord(fork, s);
    for (int i = 0; i < parts; ++i) {
        Part& pt = part[i];
        // chunks i, i + parts, ... of the pass's pixels (only the last chunk of the pass can be short)
        const unsigned int owned = (unsigned int)i < n_chunks ? (n_chunks - (unsigned int)i + (unsigned int)parts - 1u) / (unsigned int)parts : 0u;
        unsigned int part_pixels = owned * chunk;
        if (owned && (n_chunks - 1u) % (unsigned int)parts == (unsigned int)i) part_pixels -= n_chunks * chunk - pass_pixels;
        pt.a = a_all;
        pt.defer = defer_env >= 0 ? defer_env != 0 : a_all.post_defer != 0;
        pt.a.work_base = 0;
        pt.a.work_count = part_pixels * (unsigned int)a_all.spp;
        pt.a.wave_part = (unsigned int)i; pt.a.wave_parts = (unsigned int)parts; pt.a.wave_chunk = chunk;
        wave_pool_carve(&pt.q, P_part, (char*)pool_mem + pool_stride * i);
        pt.st = parts > 1 ? part_streams[i] : s;
        pt.h = h_cnt_all + kWaveCounters * i;
        pt.done = false;
        if (cudaEventCreateWithFlags(&pt.ev, cudaEventDisableTiming) != cudaSuccess) {
            for (int k = 0; k < i; ++k) cudaEventDestroy(part[k].ev);   // the events made so far
            cudaEventDestroy(fork);
            return -1;
        }
        if (parts > 1) cudaStreamWaitEvent(pt.st, fork, 0);
        const unsigned int slots = (unsigned int)std::min<unsigned long long>((unsigned long long)pt.q.P, pt.a.work_count);
        wave_init_kernel<<<(std::max(slots, (unsigned int)kWaveCounters) + 255) / 256, 256, 0, pt.st>>>(pt.q, slots);
        ++*launches;
        // persistent traversal grids, but never more warps than a wave of `slots` rays can keep busy
        const int need_blocks = (int)std::max(1u, (slots + kBlockThreads - 1) / kBlockThreads);
        pt.grid_s = std::min(num_sms * bps_s, need_blocks); pt.grid_h = std::min(num_sms * bps_h, need_blocks);
        pt.grid_l = std::min(num_sms * bps_l, need_blocks);
        pt.grid_sg = std::min(num_sms * occ_grid, need_blocks);
        pt.grid_post = std::max(1, (int)std::min<unsigned int>((slots + 127) / 128, (unsigned int)num_sms * 16));
        pt.grid_post256 = std::max(1, (pt.grid_post + 1) / 2);
    }
    int iters = 0;
    const int kBatch = 8;
    cudaEvent_t pe[7];
    double ptm[5] = {0, 0, 0, 0, 0}, pt_light = 0;
    unsigned long long cls_total[kPostClasses + 1] = {0, 0, 0, 0, 0, 0};
    if (prof) for (auto& e : pe) cudaEventCreate(&e);
    bool failed = false;
    // one iteration of a part: every kernel consumes the list the previous one filled
    // VR_WAVE_CHECK=1 (debug, with VR_WAVE_GRAPH=0): synchronise after every kernel of the chain and name the first one that fails
    static const bool check_env = getenv("VR_WAVE_CHECK") && getenv("VR_WAVE_CHECK")[0] == '1';
    auto check = [&](cudaStream_t st, const char* what) {
        if (!check_env || failed) return;
        const cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { fprintf(stderr, "[wave check] %s failed in iteration %d: %s\n", what, iters, cudaGetErrorString(e)); failed = true; }
    };
    auto enqueue_iteration = [&](Part& pt) {
        const RenderArgs& a = pt.a; const WavePool& q = pt.q; cudaStream_t st = pt.st;
        if (prof) cudaEventRecord(pe[0], st);
        if (a.scene.cg_res) wave_regen_cam_kernel<<<pt.grid_post256, 256, 0, st>>>(a, q);   // camera rays gathered from the pinhole's grid
        else wave_regen_kernel<<<pt.grid_post256, 256, 0, st>>>(a, q);
        check(st, "regen");
        if (prof) cudaEventRecord(pe[1], st);
        if (t2 && a.scene.g3_res) wave_trav2_kernel<RAY_SCATTER, true><<<pt.grid_sg, kBlockThreads, 0, st>>>(a, q);   // bounce rays start in the uniform grid
        else if (t2) wave_trav2_kernel<RAY_SCATTER><<<pt.grid_s, kBlockThreads, 0, st>>>(a, q);
        else wave_trav_kernel<RAY_SCATTER><<<pt.grid_s, kBlockThreads, 0, st>>>(a, q);
        check(st, "trav<scatter>");
        if (prof) cudaEventRecord(pe[2], st);
        wave_post_a_kernel<1><<<pt.grid_post, 128, 0, st>>>(a, q);
        check(st, "post_a<1>");
        if (post_split) {   // one launch per overlap-count class, each with an overlap store of its size
            if (pt.defer) {
                wave_post_a_kernel<0, 1><<<pt.grid_post, 128, 0, st>>>(a, q); wave_post_a_kernel<0, 2><<<pt.grid_post, 128, 0, st>>>(a, q);
                wave_post_a_kernel<0, 3><<<pt.grid_post, 128, 0, st>>>(a, q); wave_post_a_kernel<0, 4><<<pt.grid_post, 128, 0, st>>>(a, q);
                if (resume) wave_post_a_kernel<4><<<pt.grid_post, 128, 0, st>>>(a, q);
                else wave_post_a_kernel<2><<<pt.grid_post, 128, 0, st>>>(a, q);
            } else {
                wave_post_a_kernel<3, 1><<<pt.grid_post, 128, 0, st>>>(a, q); wave_post_a_kernel<3, 2><<<pt.grid_post, 128, 0, st>>>(a, q);
                wave_post_a_kernel<3, 3><<<pt.grid_post, 128, 0, st>>>(a, q); wave_post_a_kernel<3, 4><<<pt.grid_post, 128, 0, st>>>(a, q);
            }
        } else if (pt.defer && pair) {
            wave_post_a_kernel<5><<<pt.grid_post, 128, 0, st>>>(a, q);   // two overlaps, register only
            wave_post_a_kernel<6><<<pt.grid_post, 128, 0, st>>>(a, q);   // three and more, and what the first two kernels handed over
            wave_post_a_kernel<4><<<pt.grid_post, 128, 0, st>>>(a, q);
        } else if (pt.defer) {
            wave_post_a_kernel<0><<<pt.grid_post, 128, 0, st>>>(a, q);
            if (resume) wave_post_a_kernel<4><<<pt.grid_post, 128, 0, st>>>(a, q);
            else wave_post_a_kernel<2><<<pt.grid_post, 128, 0, st>>>(a, q);
        } else {
            wave_post_a_kernel<3><<<pt.grid_post, 128, 0, st>>>(a, q);
        }
        check(st, "post_a (general / pair / deferred)");
        if (prof) {
            cudaEventRecord(pe[3], st);
            unsigned int hc[kWaveCounters];
            cudaMemcpyAsync(hc, q.cnt, sizeof(hc), cudaMemcpyDeviceToHost, st);
            cudaStreamSynchronize(st);
            for (int c = 0; c <= kPostClasses; ++c) cls_total[c] += hc[WC_NA + c];
        }
        if (t2) wave_trav2_kernel<RAY_SHADOW><<<pt.grid_h, kBlockThreads, 0, st>>>(a, q);
        else wave_trav_kernel<RAY_SHADOW><<<pt.grid_h, kBlockThreads, 0, st>>>(a, q);
        check(st, "trav<shadow>");
        if (prof) cudaEventRecord(pe[6], st);
        if (a.scene.lg_mask) {
            if (t2 && a.trav_version != 3) wave_trav2_kernel<RAY_LIGHT><<<pt.grid_l, kBlockThreads, 0, st>>>(a, q);
            else wave_trav_kernel<RAY_LIGHT><<<pt.grid_l, kBlockThreads, 0, st>>>(a, q);
        }
        if (prof) cudaEventRecord(pe[4], st);
        check(st, "trav<light>");
        wave_post_b_kernel<<<pt.grid_post256, 256, 0, st>>>(a, q);
        check(st, "post_b");
        if (prof) {
            cudaEventRecord(pe[5], st);
            cudaEventSynchronize(pe[5]);
            for (int k = 0; k < 5; ++k) { float ms = 0.f; cudaEventElapsedTime(&ms, pe[k], pe[k + 1]); ptm[k] += ms; }
            { float ms = 0.f; cudaEventElapsedTime(&ms, pe[6], pe[4]); pt_light += ms; }
        }
    };
    // The batch of kBatch iterations is a fixed chain of 7-9 kernels per iteration with fixed grids and arguments: it is captured
    // once per part (and per deferral mode) into a CUDA graph and replayed, so a batch costs one launch on the host and
    // the kernels of short waves (small renders, the drain of a pass) follow each other without launch gaps.
    while (!failed) {
        bool any = false;
        NvtxRange range_batch("vr.wave.batch");
        for (int i = 0; i < parts; ++i) {
            Part& pt = part[i];
            if (pt.done) continue;
            any = true;
            const unsigned int per_iter = 6u + (pt.defer ? 1u : 0u) + (pt.defer && pair ? 1u : 0u) + (pt.a.scene.lg_mask ? 1u : 0u) + (post_split ? 3u : 0u);
            if (use_graph) {
                WaveGraphCache::Entry& en = graphs->e[i][pt.defer ? 1 : 0];
                WaveGraphKey key;
                memset((void*)&key, 0, sizeof(key));
                memcpy((void*)&key.a, (const void*)&pt.a, sizeof(RenderArgs)); memcpy((void*)&key.q, (const void*)&pt.q, sizeof(WavePool));
                key.grid[0] = pt.grid_s; key.grid[1] = pt.grid_h; key.grid[2] = pt.grid_l; key.grid[3] = pt.grid_post; key.grid[4] = pt.grid_post256; key.grid[5] = pt.grid_sg;
                key.defer = pt.defer ? 1 : 0;
                if (!en.valid || memcmp((const void*)&en.key, (const void*)&key, sizeof(key)) != 0) {
                    if (en.exec) { cudaGraphExecDestroy(en.exec); en.exec = nullptr; }
                    en.valid = false;
                    cudaGraph_t g = nullptr;
                    if (cudaStreamBeginCapture(pt.st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { failed = true; break; }
                    for (int b = 0; b < kBatch; ++b) enqueue_iteration(pt);
                    // EndCapture is called whatever happened inside the capture: a stream must never be left capturing
                    const cudaError_t ce = cudaStreamEndCapture(pt.st, &g);
                    if (ce != cudaSuccess || g == nullptr || cudaGraphInstantiate(&en.exec, g, 0) != cudaSuccess) { failed = true; en.exec = nullptr; }
                    if (g) cudaGraphDestroy(g);
                    if (failed) break;
                    memcpy((void*)&en.key, (const void*)&key, sizeof(key));
                    en.valid = true;
                }
                cudaGraphExec_t ge = en.exec;
                if (cudaGraphLaunch(ge, pt.st) != cudaSuccess) { failed = true; break; }
            } else {
                for (int b = 0; b < kBatch; ++b) enqueue_iteration(pt);
            }
            *launches += per_iter * kBatch;
        }
        iters += kBatch;
        if (!any) break;
        for (int i = 0; i < parts; ++i) {
            Part& pt = part[i];
            if (pt.done) continue;
            if (cudaMemcpyAsync(pt.h, pt.q.cnt, sizeof(unsigned int) * kWaveCounters, cudaMemcpyDeviceToHost, pt.st) != cudaSuccess ||
                cudaEventRecord(pt.ev, pt.st) != cudaSuccess) failed = true;
        }
        for (int i = 0; i < parts && !failed; ++i) {
            Part& pt = part[i];
            if (pt.done) continue;
            if (cudaEventSynchronize(pt.ev) != cudaSuccess) { failed = true; break; }
            // after wave_post_b: rays left = listTs + slots waiting for a new sample
            // deferring the iterative solves redoes those slots' walk: it pays while they are few (sparse mixtures: 8 % of the
            // general slots at C2), not when a third of all slots need one (C3, heavy overlap)
            if (defer_env < 0 && !resume && pt.defer && pt.h[WC_SUM_GENERAL] > 4096u && 5ull * pt.h[WC_SUM_DEFERRED] > pt.h[WC_SUM_GENERAL]) pt.defer = false;
            if (timeline)
                fprintf(stderr, "[timeline] part %d iter %d t %.3f ms live %u issued %u of %u\n", i, iters,
                        std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tl0).count(), pt.h[WC_NTS], std::min(pt.h[WC_NEXT_SAMPLE], pt.a.work_count), pt.a.work_count);
            const bool more_samples = pt.h[WC_NEXT_SAMPLE] < pt.a.work_count;
            if (pt.h[WC_NTS] == 0 && (pt.h[WC_NC] == 0 || !more_samples)) pt.done = true;
            // every sample is out and few paths are left: finish them in one kernel instead of ~100 more iterations
            // that are bounded by their own latency (see wave_drain_kernel)
            else if (drain_on && !more_samples && pt.h[WC_NTS] <= std::min((unsigned int)pt.q.P / drain_div, drain_max)) {
                { NvtxRange r("vr.wave.drain"); launch_wave_drain(pt.a, pt.q, pt.h[WC_NTS], pt.st); }
                ++*launches;
                pt.done = true;
                if (timeline) {
                    cudaStreamSynchronize(pt.st);
                    fprintf(stderr, "[timeline] part %d drained %u paths, t %.3f ms\n", i, pt.h[WC_NTS],
                            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tl0).count());
                }
            }
        }
        if (iters > (1 << 24)) { failed = true; break; }   // a pass that does not end is an error, not a result
    }
    for (int i = 0; i < parts; ++i) {
        if (parts > 1) { cudaEventRecord(part[i].ev, part[i].st); cudaStreamWaitEvent(s, part[i].ev, 0); }   // join
        cudaEventDestroy(part[i].ev);
    }
    cudaEventDestroy(fork);
    if (prof) {
        fprintf(stderr, "[wave] samples %u iters %d  regen %.3f  trav_scatter %.3f  post_a %.3f  trav_shadow %.3f  post_b %.3f ms\n",
                a_all.work_count, iters, ptm[0], ptm[1], ptm[2], ptm[3], ptm[4]);
        fprintf(stderr, "[wave] of trav_shadow: %.3f ms in the light-grid kernel\n", pt_light);
        fprintf(stderr, "[wave] post-A slots by overlap count: K=1 %llu  K=2 %llu  K=3..4 %llu  K=5..8 %llu  K>8 %llu  deferred (iterative solve) %llu\n", cls_total[0], cls_total[1],
                cls_total[2], cls_total[3], cls_total[4], cls_total[5]);
        for (auto& e : pe) cudaEventDestroy(e);
    }
    return failed ? -1 : iters;
}

}  // namespace vr
