// vr_internal.h -- launch-level interface between the C-ABI (vr_api.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "vr_device.cuh"

namespace vr {

// statistics counters accumulated by the path kernels (device array of kNumStats u64)
// ST_NODES .. ST_ITER: EXECUTED work of the wavefront engine (the overflow re-trace and drain kernels, < 2 % of the samples, are
// not counted): hierarchy slab steps + light-grid entries examined, quadric tests, erf evaluations, mu_t evaluations,
// analytic solves attempted, iterative (Newton / bisection) solves
enum { ST_SCATTER_RAYS = 0, ST_SHADOW_RAYS, ST_HITS, ST_OVERFLOW, ST_ERRORS, ST_NODES, ST_QUADRICS, ST_ERF, ST_MU, ST_SA, ST_ITER, kNumStats };

struct RenderArgs {
    SceneDev scene;
    CameraPod cam;
    int W, H, spp, nsq;
    int post_pair;         // wave_post_a: 1 = slots with two overlaps go through the register-only kernel <5>, the rest through <6>
    int post_defer;        // wave_post_a: iterative solves 0 in place, 1 deferred + walk redone (adaptive), 2 deferred + resumed (wave_post_a_kernel<4>)
    int post_split;        // wave_post_a: 1 = one launch per overlap-count class with a store of its size, 0 = one launch for classes 1..4
    int trav_version;      // wavefront traversal kernels: 2 = wave_trav2_kernel (private candidate slots), 1 = wave_trav_kernel
    int integrator;        // 0 single scatter, 1 multi scatter
    int solver, min_bounces;
    int x0, y0, win_w;     // owned pixel p -> x = x0 + p % win_w, y = y0 + (p / win_w) * part_count + part_rank
    int part_rank, part_count;
    unsigned int pixel_base;      // first owned pixel of this pass
    unsigned int work_base;       // added to a sample id when it is listed for the re-trace kernel (0: ids are pass-wide)
    unsigned int work_count;      // samples in this pass = pass pixels * spp (wavefront engine: of this PART of the pass)
    // wavefront engine: the parts of a pass own interleaved chunks of wave_chunk consecutive pass pixels (one window row),
    // chunk c belongs to part c % wave_parts; a part counts its samples 0..work_count-1 and wave_global_work() maps them
    // to pass-wide sample ids (what sampleL, the overflow list and lane_xy are indexed by)
    unsigned int wave_part, wave_parts, wave_chunk;
    unsigned int* counter;        // work-fetch counter of this pass
    const unsigned int* worklist; // nullptr: work ids 0..work_count-1; else ids taken from this list (re-trace kernel)
    float4* sampleL;              // per-sample radiance of this pass, [pass pixel][si]
    uint32_t* record_bits;        // per image pixel bitmap of contributing Gaussians (or nullptr)
    int record_words;
    unsigned long long* stats;
    unsigned int* overflow_list;  // samples to re-trace with the large-capacity store; overflow_cap entries
    unsigned int* overflow_count; // may run past overflow_cap: the host then fails the render instead of dropping samples
    unsigned int overflow_cap;
    float* arena;                 // re-trace kernel: kArenaFields * arena_cap * 32 floats per warp
    int arena_cap;
};

#ifdef __CUDACC__
__device__ __forceinline__ void overflow_push(const RenderArgs& A, unsigned int sample) {
    const unsigned int at = atomicAdd(A.overflow_count, 1u);
    if (at < A.overflow_cap) A.overflow_list[at] = sample;
}
#endif

// ---- wavefront engine (vr_wave.cu): a pool of P path slots whose state lives in global memory
// (sized to stay resident in the 126 MB L2), advanced one ray per iteration by four small kernels.
constexpr int kWaveCap = 16;         // free-flight overlaps kept per path slot (more = re-trace by the arena kernel)
// post-A work is binned by overlap count so that the warps of wave_post_a_kernel run slots of similar
// length: class 0: K = 1, 1: K = 2, 2: K = 3..4, 3: K = 5..8, 4: K > 8 (incl. overflow)
constexpr int kPostClasses = 5;
enum { WC_NTS = 0, WC_NTH, WC_NB, WC_NC, WC_NEXT_SAMPLE, WC_CURSOR_S, WC_CURSOR_H, WC_TICKET, WC_NTL, WC_CURSOR_L,
       WC_NA /* kPostClasses + 1 counters */,
       WC_SUM_GENERAL = WC_NA + kPostClasses + 1, WC_SUM_DEFERRED, WC_TICKET_S,   // running totals of the pass (general-kernel slots, deferred slots)
       kWaveCounters = 24 };
static_assert(WC_TICKET_S < kWaveCounters, "wave counters overflow");
constexpr int kSlotF4 = 8;           // float4s per slot record
struct WavePool {
    int P;
    // One 128-byte record per slot (kSlotF4 = 8 float4s; the four pointers below are views into it at +0, +2, +4, +6):
    // slots are visited in arbitrary order and DRAM moves 64 bytes at a time, so separate arrays cost a burst per field
    // (profiles/r01_v8: the post kernels moved 2.2x their useful bytes).  Order ray | ps2 | ps01 | hit01: the shadow
    // traversal reads ray and writes the depth sum next to it, post A reads ray + the adjacent ps01 | hit01 pair, post B
    // reads the first three sectors.
    float4* ray;          // [8p]: o.xyz, lim (scatter: sampled optical depth, shadow: tmax); [8p+1]: d.xyz, kind bits
    float4* ps2;          // [8p]: target_tau, albedo, nee_dist, nee_light bits; [8p+1]: shadow ray's optical depth sum (int64, 2^-40 units)
    float4* ps01;         // [8p]: L0 L1 L2 thr; [8p+1] (as uint4): rng state lo/hi, work (pass-local sample id), bounce
    uint4* hit01;         // [8p]: overlap count, (t0, t1, index) of overlap 0; [8p+1]: (t0, t1, index) of overlap 1, pad
    uint4* hitx;          // overlaps 2..kWaveCap-1: [p][k - 2] = (t0, t1, index, pad), contiguous per slot
    unsigned int* listTs;  // slots whose free-flight ray waits for traversal
    unsigned int* listTh;  // slots whose shadow ray waits for traversal (environment rays, lights without a grid)
    unsigned int* listTl;  // slots whose shadow ray goes to a point light that has a direction grid
    unsigned int* listA;   // [kPostClasses][P]
    unsigned int* listB;
    unsigned int* listC;
    unsigned int* cnt;    // kWaveCounters
};
size_t wave_pool_bytes(int P);
void wave_pool_carve(WavePool* q, int P, void* base);
// runs the whole pass (a.work_count samples) ; returns the number of iterations enqueued, < 0 on CUDA error
constexpr int kMaxWaveParts = 4;
// finishes the n_live paths listed in q.listTs (their free-flight rays pending) in one kernel, one thread per path (vr_kernels.cu)
void launch_wave_drain(const RenderArgs& a, const WavePool& q, unsigned int n_live, cudaStream_t s);
// instantiated CUDA graphs of the per-part iteration batches, kept between renders (a context owns one)
struct WaveGraphCache;
WaveGraphCache* wave_graph_cache_create();
void wave_graph_cache_destroy(WaveGraphCache* c);
int run_wavefront_pass(const RenderArgs& a, void* pool_mem, int P_total, int parts, int num_sms, unsigned int* h_cnt_pinned /* parts * kWaveCounters */,
                       cudaStream_t s, cudaStream_t* part_streams, unsigned int* launches, struct WaveGraphCache* graphs);

// known-answer hook over the wavefront free-flight traversal (vr_wave.cu): overlap sets of n caller rays, unpruned.
// pool_mem: wave_pool_bytes(n) bytes; sample_scratch: n float4; counts[n] = overlaps found (may exceed kWaveCap),
// idx[n][cap] = original indices of the first min(count, kWaveCap, cap) in gather order
int run_wave_gather_kat(const SceneDev& S, int n, const float* d_o3, const float* d_d3, void* pool_mem, float4* sample_scratch,
                        unsigned long long* stats, int cap, int* d_counts, uint32_t* d_idx, int num_sms, cudaStream_t s);

constexpr int kBlockThreads = 128;
constexpr int kLocalCap = 32;        // overlaps kept in local memory by the main kernel

// grid sizes chosen by occupancy query (persistent kernels)
int path_kernel_blocks_per_sm(bool arena);
void launch_path_kernel(const RenderArgs& a, int blocks, bool arena, cudaStream_t s);
// mean over spp in sample order (integrator.h:702-706), writes image[3*(y*W+x)]
void launch_pixel_reduce(const RenderArgs& a, unsigned int pass_pixels, float* image, cudaStream_t s);

void launch_transmittance(const SceneDev& S, int n, const float* o3, const float* d3, const float* tmax, float* out,
                          cudaStream_t s);

// known-answer kernels
void launch_kat_pcg(int n, const int32_t* xysi, uint64_t* seed, uint32_t* u32x4, float* unix4, cudaStream_t s);
void launch_kat_gaussian(const SceneDev& S, const uint32_t* leaf_of_orig, int n, const float* o3, const float* d3,
                         const int32_t* gidx, const float* frac, int32_t* hit, float* t0, float* t1, float* tau,
                         float* mu, int32_t* ok, float* tsol, cudaStream_t s);
void launch_kat_camera(const CameraPod& cam, int n, const float* uv, float* o3, float* d3, cudaStream_t s);
void launch_kat_events(const SceneDev& S, const float* o3, const float* d3, int cap, float* arena, float* t,
                       int32_t* entering, uint32_t* index, int* n_events, cudaStream_t s);
void launch_kat_libm(int n, const float* x, float* s_out, float* c_out, float* a_out, cudaStream_t s);
void launch_kat_erfinv(int n, const double* x, double* y, cudaStream_t s);

// scene preparation: cov6 -> Sigma^-1, norm, TIGHT 3-sigma boxes (gaussian.h:52-56,304-319) and a bound of Sigma^-1's
// largest eigenvalue per component (the boxes are inflated afterwards, vr_bvh.h inflate_boxes)
void launch_prepare_gaussians(int n, const float* mean3, const float* cov6, const float* density, const float* albedo,
                              float4* gauss_orig_order, float* box_lo, float* box_hi, float* lam, cudaStream_t s);
void launch_permute_gaussians(int n, const float4* src, const uint32_t* perm, float4* dst, uint32_t* leaf_of_orig,
                              cudaStream_t s);
// apply_params_to_gmm_local math (gmm.h:634-674)
void launch_unpack_params(int n, const float* params11, float* mean3, float* cov6, float* density, float* albedo,
                          cudaStream_t s);

}  // namespace vr
