// vr_lightgrid.h -- per-light direction grids for next-event (shadow) rays to point lights.
//
// Three quarters of the shadow rays of the bundled scenes end in one of a few point lights
// (integrator.h:652-668: the environment is chosen with probability 1/(nl+1)).  Every such ray lies on a
// line through the light, so the components it can overlap are exactly those whose 3-sigma box is seen from
// the light in the ray's direction.  For each light we therefore rasterise every component's padded box into
// a cube map around the light (6 faces x R x R cells, exact perspective bounds of the box per face) and
// keep, per cell, the list of components sorted by their distance from the light.  A shadow ray from x then
// needs no hierarchy traversal: it walks the list of the cell that contains x - light up to the entries
// nearer than |x - light| and runs the reference's quadric test (gaussian.h:126-164) on each.  The grid only
// proposes candidates -- a superset of the overlapped components -- so the summed optical depth
// (gmm.h:517-578) is unchanged.  A light that sits inside or very near a component's box gets no grid and its
// rays use the hierarchy like environment rays.
#pragma once
#include <cstdint>
#include <vector>

namespace vr {

struct LightGridHost {
    int res = 0;                        // R: cells per face side
    uint32_t mask = 0;                  // bit l: light l has a grid
    std::vector<uint32_t> cell_start;   // [nl][6 * R * R + 1], offsets into entries (absolute)
    std::vector<uint32_t> entry_gi;     // leaf-order component index
    std::vector<float> entry_dmin;      // distance from the light to the component's padded box (ascending per cell)
};

// lo/hi: 3*n floats (boxes in ORIGINAL order), leaf_of_orig: n (original index -> leaf-order index),
// lpos: 3*nl floats.  Lights beyond 32 get no grid.  Deterministic.  leaf_of_orig == nullptr: entries keep ORIGINAL
// indices (the speculative build of vr_api.cu runs next to the hierarchy build and maps them on the device afterwards).
void build_light_grids(int n, const float* lo, const float* hi, const uint32_t* leaf_of_orig, int nl, const float* lpos, int res,
                       LightGridHost* out);

// cell of a direction v (from the light towards the ray origin): face = 2 * major axis + (v[major] < 0), then
// (u, w) = the other two coordinates / |v[major]| mapped from [-1, 1] to [0, R).  Host and device use the same
// formula (the device one is in vr_device.cuh); footprints are padded so that float rounding cannot matter.
inline int light_grid_cell(const double v[3], int res) {
    const double ax = v[0] < 0 ? -v[0] : v[0], ay = v[1] < 0 ? -v[1] : v[1], az = v[2] < 0 ? -v[2] : v[2];
    int m = 0; double am = ax;
    if (ay > am) { m = 1; am = ay; }
    if (az > am) { m = 2; am = az; }
    const int a = (m + 1) % 3, b = (m + 2) % 3;
    const int face = 2 * m + (v[m] < 0 ? 1 : 0);
    auto cell = [&](double c) { int i = (int)((c / am + 1.0) * 0.5 * res); return i < 0 ? 0 : (i >= res ? res - 1 : i); };
    return (face * res + cell(v[b])) * res + cell(v[a]);
}

}  // namespace vr
