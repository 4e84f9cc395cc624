// vr_lightgrid.cpp -- host builder of the per-light direction grids.  See vr_lightgrid.h.
#include "vr_lightgrid.h"

#include <algorithm>
#include <cmath>
#include <numeric>
#include <thread>

namespace vr {
namespace {

struct Rect { int face, u0, u1, w0, w1; };

// cells of face (m, negative) that the box [lo, hi] (relative to the light) projects into; false = none
bool face_rect(const double lo[3], const double hi[3], int m, bool negative, int res, Rect* r) {
    const int a = (m + 1) % 3, b = (m + 2) % 3;
    // s * x_m over the box, s = -1 for the negative face
    double xm0 = negative ? -hi[m] : lo[m];
    const double xm1 = negative ? -lo[m] : hi[m];
    if (!(xm1 > 0.0)) return false;
    // a direction belongs to this face only if |v_a| <= v_m and |v_b| <= v_m: the part of the box that can project
    // into the face has v_m >= the smallest |v_a| and |v_b| of the box
    auto min_abs = [&](int c) { return lo[c] > 0.0 ? lo[c] : (hi[c] < 0.0 ? -hi[c] : 0.0); };
    xm0 = std::max(xm0, std::max(min_abs(a), min_abs(b)));
    if (xm0 > xm1) return false;
    double ulo = -1.0, uhi = 1.0, wlo = -1.0, whi = 1.0;
    if (xm0 > 1e-9 && xm0 > 1e-6 * xm1) {   // strictly in front of the face: exact perspective bounds of the clipped box
        auto ratio = [&](int c, double* mn, double* mx) {
            *mn = std::min(lo[c] / xm0, lo[c] / xm1);
            *mx = std::max(hi[c] / xm0, hi[c] / xm1);
        };
        ratio(a, &ulo, &uhi); ratio(b, &wlo, &whi);
        if (ulo > 1.0 || uhi < -1.0 || wlo > 1.0 || whi < -1.0) return false;
    }   // else: the box reaches the light along this axis -> whole face (conservative; such lights normally get no grid)
    auto cell = [&](double c, double slack) {
        int i = (int)std::floor((c + 1.0) * 0.5 * res + slack);
        return std::min(res - 1, std::max(0, i));
    };
    r->face = 2 * m + (negative ? 1 : 0);
    r->u0 = cell(std::max(ulo, -1.0), -1e-3); r->u1 = cell(std::min(uhi, 1.0), 1e-3);
    r->w0 = cell(std::max(wlo, -1.0), -1e-3); r->w1 = cell(std::min(whi, 1.0), 1e-3);
    return true;
}

}  // namespace

// one light: its cell starts relative to its own first entry, and its entries; returns false when the light gets no grid
static bool build_one_light(int n, const float* lo, const float* hi, const uint32_t* leaf_of_orig, const float* lpos3, int res,
                            std::vector<uint32_t>* start_rel, std::vector<uint32_t>* entry_gi, std::vector<float>* entry_dmin) {
    const size_t cells = (size_t)6 * res * res;
    // absolute slack around every box (scene units): rounding of ray origins / directions / distances, which grows with the
    // magnitude of the coordinates involved (a mixture or a light far from the coordinate origin)
    double maxabs = std::max(std::max(std::fabs((double)lpos3[0]), std::fabs((double)lpos3[1])), std::fabs((double)lpos3[2]));
    for (size_t k = 0; k < (size_t)3 * n; ++k) {
        if (std::fabs(lo[k]) < 1e30f) maxabs = std::max(maxabs, std::fabs((double)lo[k]));
        if (std::fabs(hi[k]) < 1e30f) maxabs = std::max(maxabs, std::fabs((double)hi[k]));
    }
    const double pad = std::max(1e-4, 64.0 * 5.9604644775390625e-08 * maxabs);
    const double near_limit = 0.05;
    std::vector<Rect> rects;     // per component, up to 6
    std::vector<uint32_t> rect_first((size_t)n + 1);
    std::vector<float> dmin((size_t)n);
    std::vector<uint32_t> count(cells + 1), order;
    const double L[3] = {lpos3[0], lpos3[1], lpos3[2]};
    for (int g = 0; g < n; ++g) {
        double blo[3], bhi[3], d2 = 0.0;
        for (int k = 0; k < 3; ++k) {
            // NaN boxes: components whose box is not finite cannot be culled by a grid -> no grid for this light
            if (!std::isfinite(lo[3 * g + k]) || !std::isfinite(hi[3 * g + k])) return false;
            blo[k] = (double)lo[3 * g + k] - pad - L[k]; bhi[k] = (double)hi[3 * g + k] + pad - L[k];
            const double d = std::max(std::max(blo[k], -bhi[k]), 0.0);
            d2 += d * d;
        }
        rect_first[(size_t)g] = (uint32_t)rects.size();
        if (!(blo[0] <= bhi[0] && blo[1] <= bhi[1] && blo[2] <= bhi[2])) continue;   // degenerate box: never a candidate source
        const double dist = std::sqrt(d2);
        if (!(dist > near_limit)) return false;   // the light is inside or next to this component
        dmin[(size_t)g] = (float)(dist * (1.0 - 1e-6));
        for (int m = 0; m < 3; ++m)
            for (int neg = 0; neg < 2; ++neg) {
                Rect r;
                if (face_rect(blo, bhi, m, neg != 0, res, &r)) rects.push_back(r);
            }
    }
    rect_first[(size_t)n] = (uint32_t)rects.size();
    // bound the table: a light whose footprints add up to more than ~200 cells per component gets no grid
    unsigned long long area = 0;
    for (const Rect& r : rects) area += (unsigned long long)(r.u1 - r.u0 + 1) * (unsigned long long)(r.w1 - r.w0 + 1);
    if (area > 200ULL * (unsigned long long)n + (1ULL << 20)) return false;
    std::fill(count.begin(), count.end(), 0u);
    for (int g = 0; g < n; ++g)
        for (uint32_t ri = rect_first[(size_t)g]; ri < rect_first[(size_t)g + 1]; ++ri) {
            const Rect& r = rects[ri];
            for (int w = r.w0; w <= r.w1; ++w)
                for (int u = r.u0; u <= r.u1; ++u) ++count[((size_t)r.face * res + w) * res + u];
        }
    start_rel->resize(cells + 1);
    uint32_t total = 0;
    for (size_t c = 0; c < cells; ++c) { const uint32_t k = count[c]; (*start_rel)[c] = total; count[c] = total; total += k; }
    (*start_rel)[cells] = total;
    entry_gi->resize(total); entry_dmin->resize(total);
    // fill in order of increasing distance, so every cell's list comes out sorted
    order.resize((size_t)n);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return dmin[x] < dmin[y]; });
    for (uint32_t g : order)
        for (uint32_t ri = rect_first[g]; ri < rect_first[(size_t)g + 1]; ++ri) {
            const Rect& r = rects[ri];
            for (int w = r.w0; w <= r.w1; ++w)
                for (int u = r.u0; u <= r.u1; ++u) {
                    const uint32_t at = count[((size_t)r.face * res + w) * res + u]++;
                    (*entry_gi)[at] = leaf_of_orig ? leaf_of_orig[g] : g;
                    (*entry_dmin)[at] = dmin[g];
                }
        }
    return true;
}

void build_light_grids(int n, const float* lo, const float* hi, const uint32_t* leaf_of_orig, int nl, const float* lpos, int res,
                       LightGridHost* out) {
    out->res = res; out->mask = 0;
    const size_t cells = (size_t)6 * res * res;
    out->cell_start.assign((size_t)std::max(nl, 0) * (cells + 1), 0u);
    out->entry_gi.clear(); out->entry_dmin.clear();
    if (n <= 0 || nl <= 0 || res <= 0) return;
    // the lights are independent: one host thread each (lights beyond 32 get no grid), results concatenated in light order
    const int ng = std::min(nl, 32);
    std::vector<std::vector<uint32_t>> starts((size_t)ng), gis((size_t)ng);
    std::vector<std::vector<float>> dmins((size_t)ng);
    std::vector<char> ok((size_t)ng, 0);
    std::vector<std::thread> workers;
    for (int l = 0; l < ng; ++l)
        workers.emplace_back([&, l] { ok[(size_t)l] = build_one_light(n, lo, hi, leaf_of_orig, lpos + 3 * l, res, &starts[(size_t)l], &gis[(size_t)l], &dmins[(size_t)l]) ? 1 : 0; });
    for (std::thread& t : workers) t.join();
    for (int l = 0; l < nl; ++l) {
        uint32_t* start = out->cell_start.data() + (size_t)l * (cells + 1);
        const uint32_t base = (uint32_t)out->entry_gi.size();
        if (l >= ng || !ok[(size_t)l]) {   // a light without a grid has empty cells
            for (size_t c = 0; c <= cells; ++c) start[c] = base;
            continue;
        }
        for (size_t c = 0; c <= cells; ++c) start[c] = base + starts[(size_t)l][c];
        out->entry_gi.insert(out->entry_gi.end(), gis[(size_t)l].begin(), gis[(size_t)l].end());
        out->entry_dmin.insert(out->entry_dmin.end(), dmins[(size_t)l].begin(), dmins[(size_t)l].end());
        out->mask |= 1u << l;
    }
}

}  // namespace vr
