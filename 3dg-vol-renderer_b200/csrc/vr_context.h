// vr_context.h -- the opaque context behind the C-ABI (include/volrender.h).
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

#include <string>
#include <vector>

#include "../../include/volrender.h"
#include "vr_internal.h"
#include "vr_bvh.h"

namespace vr {

// NVTX range of a host-side phase (set-up / pass / drain / reduce / collective), visible in Nsight Systems / Compute timelines;
// header-only NVTX 3: costs a predictable branch when no tool is attached
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

// grow-only device buffer
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t ensure(size_t need) {
        if (need <= bytes) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; bytes = 0; }
        cudaError_t e = cudaMalloc(&p, need);
        if (e == cudaSuccess) bytes = need;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct SfdState;  // vr_inverse.cu

}  // namespace vr

struct vr_context {
    int device = 0;
    int num_sms = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    std::string err = "";

    // ---- Gaussian mixture
    bool has_gmm = false;
    int n = 0, n_nodes = 0, bvh_depth = 0;
    vr::DevBuf mean3, cov6, density, albedo;       // inputs as given (original order)
    vr::DevBuf gauss_orig, gauss_leaf, nodes, qnodes, perm, leaf_of_orig, box_lo, box_hi, lam;
    // reach of the ray origins the culling boxes were inflated for (vr_bvh.h BoxReach) and the mixture's bounding box:
    // a camera or a ray batch that reaches farther makes the next render rebuild the hierarchy (ensure_reach)
    double box_kappa = vr::kBoxKappa;
    int hierarchy_builder = 0;                     // 0: LBVH on the device (vr_lbvh.cu), 1: binned SAH on the host (vr_bvh.cpp)
    vr::DevBuf lbvh_keys, lbvh_tree, lbvh_tmp;     // scratch of the device build
    unsigned int* h_build_meta = nullptr;          // pinned, 32 words: what the device build reports back
    float built_R = 0.f, built_X = 0.f, bb_lo[3] = {0.f, 0.f, 0.f}, bb_hi[3] = {0.f, 0.f, 0.f};
    float qorg[3] = {0.f, 0.f, 0.f}, qcell[3] = {1.f, 1.f, 1.f};   // grid of the compact tree (vr_bvh.h BvhQuantGrid)
    // host copies kept for the per-light direction grids (vr_lightgrid.h), built lazily before a render
    std::vector<float> h_box_lo, h_box_hi, h_lpos;
    std::vector<uint32_t> h_leaf_of_orig;
    vr::DevBuf lg_start, lg_gi, lg_dmin, lg_scratch;
    // direction grid around the pinhole of the camera: candidate lists of the camera rays (vr_lightgrid_dev.cu, wave_regen_kernel)
    vr::DevBuf cg_start, cg_gi, cg_dmin;
    vr::DevBuf g3_start, g3_gi, g3_cursor;         // uniform grid over the mixture (vr_grid3d.cu)
    bool g3_valid = false, g3_ok = false;          // built for the current mixture / usable
    int g3_res = 0, g3_res_option = 0;             // cells per axis (option "grid3d_res": 0 = from the component count)
    unsigned long long g3_entries = 0;
    float g3_org[3] = {0.f, 0.f, 0.f}, g3_inv[3] = {0.f, 0.f, 0.f}, g3_cell[3] = {0.f, 0.f, 0.f};
    bool use_grid3d = true;                        // option "grid3d"
    double grid3d_min_samples = 2e6;
    bool cg_valid = false, cg_ok = false;          // built for the current mixture + camera / usable (pinhole not inside a component)
    float cg_pinhole[3] = {0.f, 0.f, 0.f};
    int use_camera_grid = 1;
    double camera_grid_min_samples = 2.0e6;
    bool lg_valid = false;
    int lg_res = 256;
    uint32_t lg_mask = 0;
    int use_light_grid = 1;
    double light_grid_min_samples = 16.0e6;       // renders smaller than this skip the grids (build cost > gain)
    // speculative build: when the last render used grids, vr_set_gmm* starts their build on worker threads NEXT TO the
    // hierarchy build (boxes and lights are all it needs; entries carry original indices, mapped to leaf order by a
    // device kernel after the upload), so a set_gmm + render pair no longer pays for it in sequence
    struct LightGridJob;                          // vr_api.cu
    LightGridJob* lg_job = nullptr;
    bool lg_speculate = false;
    // ---- sphere mixture
    bool has_smm = false;
    int n_spheres = 0;
    vr::DevBuf spheres;                            // 6 floats per sphere: c3, r, sigma_a, sigma_s
    // ---- lights / camera
    int nl = 0;
    vr::DevBuf lpos, lint;
    float env[3] = {0.53f, 0.81f, 0.92f};          // scene.h:29
    bool has_camera = false;
    vr::CameraPod cam{};
    // ---- render state
    vr::DevBuf image, sampleL, counters, stats, overflow_list, arena, record_bits, scratch, wavepool;
    int img_w = 0, img_h = 0;
    int rec_w = 0, rec_h = 0, rec_words = 0, rec_n = 0;
    bool rec_valid = false;
    unsigned long long samples_valid = 0;          // samples of the last render still held in sampleL (0: none / multi-pass)
    unsigned long long* h_stats = nullptr;         // pinned
    unsigned int* h_counters = nullptr;            // pinned
    // ---- options
    int hit_capacity = 0;                          // 0 = min(n, 4096)
    double max_pass_samples = 268435456.0;         // 2^28 samples = 4 GiB of per-sample radiance
    int blocks_per_sm_main = 0, blocks_per_sm_arena = 0;
    int engine = 0;                                // 0 wavefront (vr_wave.cu), 1 warp-scheduled single kernel (vr_warp.cuh)
    int wave_slots = 0;                            // 0 = 32768 per SM (sweeps in profiles/r01_tuning.md, r01_v11/)
    int post_pair = 1;                             // slots with exactly two overlaps walked by the register-only wave_post_a_kernel<5> (needs post_defer 2, post_split 0)
    int post_defer = 2;                            // iterative free-flight solves: 0 in place, 1 deferred (walk redone), 2 deferred and resumed from saved state
    int post_split = 0;                            // wave_post_a_kernel per overlap-count class (vr_wave.cu)
    int trav_version = 2;                          // traversal kernels of the wavefront engine (vr_wave.cu): 2 wave_trav2_kernel, 1 wave_trav_kernel
    int wave_parts = 2;                            // independent parts (own pool + stream) a big pass is cut into
    cudaStream_t wave_streams[vr::kMaxWaveParts] = {nullptr, nullptr, nullptr, nullptr};
    vr::WaveGraphCache* wave_graphs = nullptr;     // CUDA graphs of the iteration batches, reused while the launch arguments stay the same
    unsigned int* h_wave_counters = nullptr;       // pinned, kMaxWaveParts * kWaveCounters
    // ---- inverse
    vr::SfdState* sfd = nullptr;
    // ---- communicator of the inverse path's one collective (vr_comm.cu); comm is an ncclComm_t
    void* comm = nullptr;
    int comm_rank = 0, comm_size = 1;
    vr::DevBuf comm_scratch;

    vr::SceneDev scene_dev() const;
};

namespace vr {
int fail(vr_context* c, int code, const std::string& msg);
int cuda_fail(vr_context* c, cudaError_t e, const char* what);
int render_device_impl(vr_context* c, const vr_render_params* p, vr_render_stats* st);
int set_gmm_from_device(vr_context* c, int n, const float* reach_RX = nullptr);   // reach_RX: minimum BoxReach {R, X} to build for
int build_hierarchy_device(vr_context* c, int n, const float* reach_RX);          // vr_lbvh.cu
int build_light_grids_device(vr_context* c);      // vr_lightgrid_dev.cu
int build_camera_grid_device(vr_context* c);      // vr_lightgrid_dev.cu
int build_grid3d_device(vr_context* c);           // vr_grid3d.cu
int ensure_grid3d(vr_context* c);
int ensure_camera_grid(vr_context* c);            // vr_api.cu
int fetch_host_boxes(vr_context* c);               // vr_api.cu: inflated boxes + leaf order on the host (light-grid builder)
int ensure_reach(vr_context* c, float R, float X);   // rebuilds the hierarchy when it was inflated for a smaller reach
int ensure_camera_reach(vr_context* c);              // ... for the camera currently set
void cancel_light_grid_job(vr_context* c);  // vr_api.cu: joins and drops a speculative grid build
int ensure_light_grids(vr_context* c);  // vr_api.cu  // mean3/cov6/density/albedo device buffers already filled
void sfd_destroy(vr_context* c);
int sfd_reduce_buffer(vr_context* c, long long** buf, size_t* count);   // vr_inverse.cu
void comm_destroy(vr_context* c);                                        // vr_comm.cu
}  // namespace vr

#define VR_CUDA(call)                                                        \
    do {                                                                     \
        cudaError_t _e = (call);                                             \
        if (_e != cudaSuccess) return vr::cuda_fail(ctx, _e, #call);         \
    } while (0)
