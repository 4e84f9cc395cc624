// vr_lightgrid_dev.cu -- the per-light direction grids (vr_lightgrid.h) built ON THE DEVICE.
//
// Same tables as build_light_grids() of vr_lightgrid.cpp -- per light a cube map of 6 x R x R cells, per cell the components
// whose padded box projects into it, nearest to the light first -- made by a count / scan / fill / sort chain on the
// context's stream from the inflated boxes the hierarchy build left on the device:
//   lg_rects_kernel   per component: distance of its box from the light and its footprint on each cube face (the host
//                     builder's face_rect in the same double arithmetic); per light: entries needed, "light inside or next
//                     to a component", "non-finite box"
//   (one 16-byte-per-light read-back decides which lights get a grid and sizes the entry arrays)
//   lg_count_kernel   entries per cell;  cub exclusive scan -> cell_start
//   lg_fill_kernel    entries (leaf-order index, distance) into their cells
//   lg_sort_kernel    every cell's list by (distance, index): the traversal stops at the first entry beyond the ray's origin
// The grids only PROPOSE candidates (the quadric test decides, the depth sum is order independent), so any table that holds
// a superset of the overlapped components gives the same bits: test_light_grid_is_only_a_candidate_source.
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <vector>

#include "vr_context.h"
#include "vr_lightgrid.h"

namespace vr {

struct LgRect { int face, u0, u1, w0, w1; };
struct LgLightInfo { unsigned long long area; unsigned int bad; unsigned int pad; };   // bad: bit 0 too near, bit 1 non-finite box

__device__ bool lg_face_rect(const double lo[3], const double hi[3], int m, bool negative, int res, LgRect* r) {
    const int a = (m + 1) % 3, b = (m + 2) % 3;
    double xm0 = negative ? -hi[m] : lo[m];
    const double xm1 = negative ? -lo[m] : hi[m];
    if (!(xm1 > 0.0)) return false;
    const double mina = lo[a] > 0.0 ? lo[a] : (hi[a] < 0.0 ? -hi[a] : 0.0);
    const double minb = lo[b] > 0.0 ? lo[b] : (hi[b] < 0.0 ? -hi[b] : 0.0);
    xm0 = fmax(xm0, fmax(mina, minb));
    if (xm0 > xm1) return false;
    double ulo = -1.0, uhi = 1.0, wlo = -1.0, whi = 1.0;
    if (xm0 > 1e-9 && xm0 > 1e-6 * xm1) {
        ulo = fmin(lo[a] / xm0, lo[a] / xm1); uhi = fmax(hi[a] / xm0, hi[a] / xm1);
        wlo = fmin(lo[b] / xm0, lo[b] / xm1); whi = fmax(hi[b] / xm0, hi[b] / xm1);
        if (ulo > 1.0 || uhi < -1.0 || wlo > 1.0 || whi < -1.0) return false;
    }
    r->face = 2 * m + (negative ? 1 : 0);
    r->u0 = min(res - 1, max(0, (int)floor((fmax(ulo, -1.0) + 1.0) * 0.5 * res - 1e-3)));
    r->u1 = min(res - 1, max(0, (int)floor((fmin(uhi, 1.0) + 1.0) * 0.5 * res + 1e-3)));
    r->w0 = min(res - 1, max(0, (int)floor((fmax(wlo, -1.0) + 1.0) * 0.5 * res - 1e-3)));
    r->w1 = min(res - 1, max(0, (int)floor((fmin(whi, 1.0) + 1.0) * 0.5 * res + 1e-3)));
    return true;
}

struct LgLight { double pos[3]; };

__global__ void lg_rects_kernel(int n, const float* __restrict__ lo, const float* __restrict__ hi, LgLight L, double pad, double near_limit,
                                int res, LgRect* __restrict__ rects, unsigned char* __restrict__ nrect, float* __restrict__ dmin,
                                LgLightInfo* info) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long area = 0;
    unsigned int bad = 0;
    if (g < n) {
        double blo[3], bhi[3], d2 = 0.0;
        bool finite = true;
        for (int k = 0; k < 3; ++k) {
            const float l = lo[3 * (size_t)g + k], h = hi[3 * (size_t)g + k];
            if (!isfinite(l) || !isfinite(h)) finite = false;
            blo[k] = (double)l - pad - L.pos[k]; bhi[k] = (double)h + pad - L.pos[k];
            const double d = fmax(fmax(blo[k], -bhi[k]), 0.0);
            d2 += d * d;
        }
        int cnt = 0;
        if (!finite) bad |= 2u;
        else if (blo[0] <= bhi[0] && blo[1] <= bhi[1] && blo[2] <= bhi[2]) {
            const double dist = sqrt(d2);
            if (!(dist > near_limit)) bad |= 1u;
            dmin[g] = (float)(dist * (1.0 - 1e-6));
            for (int m = 0; m < 3; ++m)
                for (int neg = 0; neg < 2; ++neg) {
                    LgRect r;
                    if (lg_face_rect(blo, bhi, m, neg != 0, res, &r)) {
                        rects[(size_t)g * 6 + cnt++] = r;
                        area += (unsigned long long)(r.u1 - r.u0 + 1) * (unsigned long long)(r.w1 - r.w0 + 1);
                    }
                }
        }
        nrect[g] = (unsigned char)cnt;
    }
    for (int o = 16; o > 0; o >>= 1) { area += __shfl_xor_sync(0xffffffffu, area, o); bad |= __shfl_xor_sync(0xffffffffu, bad, o); }
    if ((threadIdx.x & 31) == 0) {
        if (area) atomicAdd(&info->area, area);
        if (bad) atomicOr(&info->bad, bad);
    }
}

__global__ void lg_count_kernel(int n, int res, const LgRect* __restrict__ rects, const unsigned char* __restrict__ nrect, uint32_t* __restrict__ count) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    for (int k = 0; k < (int)nrect[g]; ++k) {
        const LgRect r = rects[(size_t)g * 6 + k];
        for (int w = r.w0; w <= r.w1; ++w)
            for (int u = r.u0; u <= r.u1; ++u) atomicAdd(count + ((size_t)r.face * res + w) * res + u, 1u);
    }
}

// cell_start of one light: shifted by the entries of the lights before it / set to one value (a light without a grid)
__global__ void lg_shift_kernel(size_t count, uint32_t* __restrict__ start, uint32_t base, int set) {
    const size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < count) start[c] = set ? base : start[c] + base;
}

__global__ void lg_fill_kernel(int n, int res, const LgRect* __restrict__ rects, const unsigned char* __restrict__ nrect, const float* __restrict__ dmin,
                               const uint32_t* __restrict__ leaf_of_orig, const uint32_t* __restrict__ start, uint32_t* __restrict__ cursor,
                               uint32_t* __restrict__ entry_gi, float* __restrict__ entry_dmin) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const uint32_t gi = leaf_of_orig[g];
    const float dm = dmin[g];
    for (int k = 0; k < (int)nrect[g]; ++k) {
        const LgRect r = rects[(size_t)g * 6 + k];
        for (int w = r.w0; w <= r.w1; ++w)
            for (int u = r.u0; u <= r.u1; ++u) {
                const size_t cell = ((size_t)r.face * res + w) * res + u;
                const uint32_t at = start[cell] + atomicAdd(cursor + cell, 1u);
                entry_gi[at] = gi; entry_dmin[at] = dm;
            }
    }
}

__global__ void lg_sort_kernel(size_t cells, const uint32_t* __restrict__ start, uint32_t* __restrict__ entry_gi, float* __restrict__ entry_dmin) {
    const size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cells) return;
    const uint32_t b = start[c], e = start[c + 1];
    for (uint32_t i = b + 1; i < e; ++i) {   // insertion sort by (distance, index): lists are short
        const float d = entry_dmin[i]; const uint32_t g = entry_gi[i];
        uint32_t j = i;
        while (j > b && (entry_dmin[j - 1] > d || (entry_dmin[j - 1] == d && entry_gi[j - 1] > g))) {
            entry_dmin[j] = entry_dmin[j - 1]; entry_gi[j] = entry_gi[j - 1]; --j;
        }
        entry_dmin[j] = d; entry_gi[j] = g;
    }
}

// Direction grids around `nl` points (pos3) from the device-resident boxes (ctx->box_lo/hi inflated, original order) and
// ctx->leaf_of_orig, into start / gi / dmin; *mask_out bit l: point l has a grid.  One synchronisation: the per-point sizes.
static int build_point_grids(vr_context* ctx, int nl, const float* pos3, DevBuf& start_buf, DevBuf& gi_buf, DevBuf& dmin_buf, uint32_t* mask_out) {
    const int n = ctx->n, res = ctx->lg_res;
    cudaStream_t s = ctx->stream;
    const size_t cells = (size_t)6 * res * res;
    const int ng = std::min(nl, 32);
    *mask_out = 0;
    VR_CUDA(start_buf.ensure((size_t)nl * (cells + 1) * 4));
    // scratch: rects n*6, nrect n, dmin n  per light | count / cursor cells+1 | info
    const size_t per_light = (size_t)n * (6 * sizeof(LgRect) + 4 + 4);
    VR_CUDA(ctx->lg_scratch.ensure((size_t)ng * per_light + (cells + 1) * 4 + 64 * sizeof(LgLightInfo) + 256));
    char* base = ctx->lg_scratch.as<char>();
    auto rects_of = [&](int l) { return (LgRect*)(base + (size_t)l * per_light); };
    auto dmin_of = [&](int l) { return (float*)(base + (size_t)l * per_light + (size_t)n * 6 * sizeof(LgRect)); };
    auto nrect_of = [&](int l) { return (unsigned char*)(base + (size_t)l * per_light + (size_t)n * 6 * sizeof(LgRect) + (size_t)n * 4); };
    uint32_t* count = (uint32_t*)(base + (size_t)ng * per_light);
    LgLightInfo* info = (LgLightInfo*)(count + cells + 1 + ((cells + 1) & 1));
    VR_CUDA(cudaMemsetAsync(info, 0, 64 * sizeof(LgLightInfo), s));
    // slack around every box: rounding of ray origins / directions / distances grows with the coordinates involved
    double maxabs = 0.0;
    for (int k = 0; k < 3; ++k) maxabs = std::max(maxabs, std::max(std::fabs((double)ctx->bb_lo[k]), std::fabs((double)ctx->bb_hi[k])));
    const int nb = (n + 127) / 128;
    for (int l = 0; l < ng; ++l) {
        LgLight L;
        double mabs = maxabs;
        for (int k = 0; k < 3; ++k) { L.pos[k] = (double)pos3[(size_t)3 * l + k]; mabs = std::max(mabs, std::fabs(L.pos[k])); }
        const double pad = std::max(1e-4, 64.0 * 5.9604644775390625e-08 * (mabs + 1.0));
        lg_rects_kernel<<<nb, 128, 0, s>>>(n, ctx->box_lo.as<float>(), ctx->box_hi.as<float>(), L, pad, 0.05, res, rects_of(l), nrect_of(l), dmin_of(l), info + l);
    }
    std::vector<LgLightInfo> hinfo((size_t)std::max(ng, 1));
    VR_CUDA(cudaMemcpyAsync(hinfo.data(), info, (size_t)ng * sizeof(LgLightInfo), cudaMemcpyDeviceToHost, s));
    VR_CUDA(cudaStreamSynchronize(s));
    size_t total = 0;
    std::vector<char> ok((size_t)nl, 0);
    for (int l = 0; l < ng; ++l) {
        // a light inside / next to a component, a non-finite box, or footprints beyond ~200 cells per component: no grid
        ok[(size_t)l] = hinfo[(size_t)l].bad == 0 && hinfo[(size_t)l].area <= 200ULL * (unsigned long long)n + (1ULL << 20);
        if (ok[(size_t)l]) total += (size_t)hinfo[(size_t)l].area;
    }
    if (total >= 0xfffffff0ull) { for (auto& o : ok) o = 0; total = 0; }
    VR_CUDA(gi_buf.ensure(std::max<size_t>(total, 1) * 4));
    VR_CUDA(dmin_buf.ensure(std::max<size_t>(total, 1) * 4));
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, count, count, (int)(cells + 1), s);
    VR_CUDA(ctx->lbvh_tmp.ensure(std::max<size_t>(tmp_bytes, 16)));
    uint32_t entries_before = 0;
    const unsigned grid_cells = (unsigned)((cells + 1 + 255) / 256);
    for (int l = 0; l < nl; ++l) {
        uint32_t* start = start_buf.as<uint32_t>() + (size_t)l * (cells + 1);
        if (!ok[(size_t)l]) {   // a light without a grid: every cell is empty
            lg_shift_kernel<<<grid_cells, 256, 0, s>>>(cells + 1, start, entries_before, 1);
            continue;
        }
        VR_CUDA(cudaMemsetAsync(count, 0, (cells + 1) * 4, s));
        lg_count_kernel<<<nb, 128, 0, s>>>(n, res, rects_of(l), nrect_of(l), count);
        // count[cells] stays 0, so the scan over cells + 1 items closes the table: start[cells] = entries of this light
        if (cub::DeviceScan::ExclusiveSum(ctx->lbvh_tmp.p, tmp_bytes, count, start, (int)(cells + 1), s) != cudaSuccess)
            return cuda_fail(ctx, cudaGetLastError(), "light grids: scan");
        if (entries_before) lg_shift_kernel<<<grid_cells, 256, 0, s>>>(cells + 1, start, entries_before, 0);
        VR_CUDA(cudaMemsetAsync(count, 0, (cells + 1) * 4, s));   // reused as the fill cursors
        lg_fill_kernel<<<nb, 128, 0, s>>>(n, res, rects_of(l), nrect_of(l), dmin_of(l), ctx->leaf_of_orig.as<uint32_t>(), start, count,
                                          gi_buf.as<uint32_t>(), dmin_buf.as<float>());
        lg_sort_kernel<<<(unsigned)((cells + 127) / 128), 128, 0, s>>>(cells, start, gi_buf.as<uint32_t>(), dmin_buf.as<float>());
        entries_before += (uint32_t)hinfo[(size_t)l].area;
        *mask_out |= 1u << l;
    }
    VR_CUDA(cudaGetLastError());
    return VR_OK;
}

// per-light grids: ctx->lg_start / lg_gi / lg_dmin / lg_mask
int build_light_grids_device(vr_context* ctx) {
    return build_point_grids(ctx, ctx->nl, ctx->h_lpos.data(), ctx->lg_start, ctx->lg_gi, ctx->lg_dmin, &ctx->lg_mask);
}

// The same grid around the PINHOLE of a pinhole camera (camera.h:38-53): every camera ray passes through it, so the
// components a camera ray can overlap beyond the pinhole are those of the cell of its direction, nearest first, and those it
// can overlap between film plane and pinhole are in the cell of the opposite direction (vr_wave.cu wave_regen_kernel<true>).
int build_camera_grid_device(vr_context* ctx) {
    uint32_t mask = 0;
    int rc = build_point_grids(ctx, 1, ctx->cam.pinhole, ctx->cg_start, ctx->cg_gi, ctx->cg_dmin, &mask);
    ctx->cg_ok = rc == VR_OK && mask == 1u;
    return rc;
}

}  // namespace vr
