// vr_pathcommon.cuh -- small pieces shared by the path engines (vr_warp.cuh, vr_wave.cu)
#pragma once
#include "vr_internal.h"

namespace vr {

constexpr int kMaxLeafDev = 1;                 // == vr::kMaxLeaf of the host builder (vr_bvh.h)
enum { RAY_SCATTER = 0, RAY_SHADOW = 1, RAY_LIGHT = 2 };   // RAY_LIGHT: shadow ray to a point light with a direction grid
constexpr int kRayNoGrid = 0x100;   // flag on a RAY_SCATTER kind: gather through the hierarchy, not the uniform grid (wave_post_a sent the slot back)

__device__ __forceinline__ void lane_xy(const RenderArgs& A, unsigned int work, int* x, int* y, int* si) {
    const unsigned int p = A.pixel_base + work / (unsigned int)A.spp;
    *si = (int)(work % (unsigned int)A.spp);
    *x = A.x0 + (int)(p % (unsigned int)A.win_w);
    *y = A.y0 + (int)(p / (unsigned int)A.win_w) * A.part_count + A.part_rank;
}

// part-local sample id -> pass-wide sample id (RenderArgs::wave_chunk); the samples of a pixel stay consecutive
__device__ __forceinline__ unsigned int wave_global_work(const RenderArgs& A, unsigned int idx) {
    if (A.wave_parts <= 1u) return idx;
    const unsigned int spp = (unsigned int)A.spp;
    const unsigned int lp = idx / spp, si = idx - lp * spp;
    const unsigned int k = lp / A.wave_chunk, r = lp - k * A.wave_chunk;
    return ((k * A.wave_parts + A.wave_part) * A.wave_chunk + r) * spp + si;
}

// heavy helpers are kept out of line ONCE so the scheduler + traversal loop stays compact
static __device__ __noinline__ f3 sample_direction_ni(uint64_t* state) {
    Pcg rng; rng.state = *state; rng.inc = 3;
    const f3 d = normalized3(sample_uniform_direction(rng));  // Ray ctor normalises (ray.h:11-12)
    *state = rng.state;
    return d;
}

}  // namespace vr
