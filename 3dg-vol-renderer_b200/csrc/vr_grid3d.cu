// vr_grid3d.cu -- a uniform grid over the mixture for the FIRST stretch of the secondary free-flight rays.
//
// In these optically thick mixtures a bounce ray (which starts inside a component) scatters again after a very short
// flight: at C2 83 % of them within 0.008 units, 97 % overlap nothing but the few components around their origin
// (profiles/r02/SUMMARY.md).  Walking the hierarchy from its root for such a ray costs ~19 node visits; the cell of a uniform
// grid that contains the origin lists the 2-5 components it can meet before it leaves that cell.
//
// The table: R^3 cells over the mixture's box; per cell the components whose culling box (the inflated 3-sigma box of the
// hierarchy build, vr_lbvh.cu inflate_kernel) touches the cell, padded by 1e-3 of a cell plus the FP32 rounding of coordinates
// of the mixture's magnitude, as leaf-order indices sorted ascending (deterministic lists).
// wave_trav2_kernel<SCATTER, GRID> (vr_wave.cu) tests the origin cell's list with the reference's quadric and keeps the overlaps
// that ENTER before the ray leaves the cell.  If they hold more than the sampled optical depth inside the cell the ray is
// finished and carries the cell's exit as a certificate; otherwise the hierarchy is walked for the overlaps that enter at or
// beyond that point, so every overlap is kept by exactly one of the two stages.  wave_post_a accepts a result computed from a
// certified list only where the reference's walk cannot have looked beyond the certificate and sends the slot back to the
// hierarchy otherwise (vr_wave.cu, "Certificates of the uniform grid").  Images and recorded sets are bit-identical with and
// without the grid: test_traversal_versions_and_builders_give_identical_images (incl. the far and x50 scenes, recorded renders,
// all solver modes), test_uniform_grid_changes_no_bit_at_full_size (C2 at full size, C3).
//   g3_count_kernel   entries per cell;  cub exclusive scan -> cell_start
//   g3_fill_kernel    leaf-order indices into their cells
//   g3_sort_kernel    every cell's list ascending
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cmath>

#include "vr_context.h"

namespace vr {

struct G3Dims { float org[3], inv[3], pad[3]; int res; };   // pad: slack of a box's cell range, in cells

__device__ __forceinline__ void g3_range(const G3Dims& G, const float* __restrict__ lo, const float* __restrict__ hi, int g, int r0[3], int r1[3], bool* finite) {
    *finite = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float a = lo[3 * (size_t)g + k], b = hi[3 * (size_t)g + k];
        if (!(fabsf(a) < 1e30f) || !(fabsf(b) < 1e30f)) *finite = false;
        // conservative: a box that comes within G.pad of a cell (1e-3 of a cell + the rounding of coordinates this far out) is listed there
        const float fa = floorf((a - G.org[k]) * G.inv[k] - G.pad[k]), fb = floorf((b - G.org[k]) * G.inv[k] + G.pad[k]);
        r0[k] = (int)fminf(fmaxf(fa, 0.0f), (float)(G.res - 1));
        r1[k] = (int)fminf(fmaxf(fb, 0.0f), (float)(G.res - 1));
    }
}

__global__ void g3_count_kernel(int n, G3Dims G, const float* __restrict__ lo, const float* __restrict__ hi, uint32_t* __restrict__ count,
                                unsigned long long* __restrict__ info) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    int r0[3], r1[3]; bool finite;
    g3_range(G, lo, hi, g, r0, r1, &finite);
    if (!finite) { atomicAdd(info + 1, 1ULL); return; }
    const unsigned long long cells = (unsigned long long)(r1[0] - r0[0] + 1) * (unsigned long long)(r1[1] - r0[1] + 1) * (unsigned long long)(r1[2] - r0[2] + 1);
    atomicAdd(info, cells);
    if (cells > 32768ULL) { atomicAdd(info + 1, 1ULL); return; }   // a component that spans a large part of the grid: no grid for this mixture
    for (int z = r0[2]; z <= r1[2]; ++z)
        for (int y = r0[1]; y <= r1[1]; ++y)
            for (int x = r0[0]; x <= r1[0]; ++x) atomicAdd(count + ((size_t)z * G.res + y) * G.res + x, 1u);
}

__global__ void g3_fill_kernel(int n, G3Dims G, const float* __restrict__ lo, const float* __restrict__ hi, const uint32_t* __restrict__ leaf_of_orig,
                               const uint32_t* __restrict__ start, uint32_t* __restrict__ cursor, uint32_t* __restrict__ gi) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    int r0[3], r1[3]; bool finite;
    g3_range(G, lo, hi, g, r0, r1, &finite);
    const uint32_t leaf = leaf_of_orig[g];
    for (int z = r0[2]; z <= r1[2]; ++z)
        for (int y = r0[1]; y <= r1[1]; ++y)
            for (int x = r0[0]; x <= r1[0]; ++x) {
                const size_t c = ((size_t)z * G.res + y) * G.res + x;
                gi[start[c] + atomicAdd(cursor + c, 1u)] = leaf;
            }
}

__global__ void g3_sort_kernel(size_t cells, const uint32_t* __restrict__ start, uint32_t* __restrict__ gi) {
    const size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cells) return;
    const uint32_t a = start[c], b = start[c + 1];
    for (uint32_t i = a + 1; i < b; ++i) {   // insertion sort: the lists are a handful of entries
        const uint32_t v = gi[i];
        uint32_t j = i;
        while (j > a && gi[j - 1] > v) { gi[j] = gi[j - 1]; --j; }
        gi[j] = v;
    }
}

// ctx->g3_start / g3_gi / g3_org / g3_inv / g3_cell / g3_res; g3_ok = false when the mixture does not suit a grid
int build_grid3d_device(vr_context* ctx) {
    ctx->g3_ok = false;
    const int n = ctx->n;
    if (n <= 0) return VR_OK;
    cudaStream_t s = ctx->stream;
    // resolution: ~26 cells per component, between 32 and 128 per axis (C2, 10k components: 64; C3, 200k: 128)
    int res = ctx->g3_res_option > 0 ? ctx->g3_res_option : (int)std::lround(std::cbrt(26.0 * (double)n));
    res = std::max(8, std::min(res, ctx->g3_res_option > 0 ? 256 : 128));
    if (ctx->g3_res_option <= 0) res = std::max(res, 32);
    G3Dims G; G.res = res;
    for (int k = 0; k < 3; ++k) {
        const double lo = (double)ctx->bb_lo[k], hi = (double)ctx->bb_hi[k];
        if (!(std::fabs(lo) < 1e30) || !(std::fabs(hi) < 1e30) || !(hi >= lo)) return VR_OK;
        const double ext = std::max(hi - lo, 1e-6) * (1.0 + 1e-5) + 1e-6;   // the mixture's box, a hair larger
        const double org = lo - 0.5 * (ext - (hi - lo));
        G.org[k] = (float)org; G.inv[k] = (float)((double)res / ext);
        // a ray's cell and the parameter at which it leaves it are computed in FP32 from coordinates of magnitude maxabs: 32 ulp of
        // those, in cells, on top of the 1e-3 (at C2: 1e-3 + 1.2e-4; for a mixture 230 units from the origin: 1e-3 + 1.4e-2)
        const double maxabs = std::max(std::fabs(lo), std::fabs(hi)) + ext;
        G.pad[k] = (float)(1e-3 + 32.0 * 5.9604644775390625e-08 * maxabs * (double)res / ext);
        ctx->g3_org[k] = G.org[k]; ctx->g3_inv[k] = G.inv[k]; ctx->g3_cell[k] = (float)(ext / (double)res);
    }
    const size_t cells = (size_t)res * res * res;
    VR_CUDA(ctx->g3_start.ensure((cells + 1) * 4));
    VR_CUDA(ctx->g3_cursor.ensure((cells + 1) * 4 + 64));
    uint32_t* count = ctx->g3_cursor.as<uint32_t>();
    unsigned long long* info = (unsigned long long*)(ctx->g3_cursor.as<char>() + (((cells + 1) * 4 + 15) & ~(size_t)15));
    VR_CUDA(cudaMemsetAsync(ctx->g3_cursor.p, 0, (cells + 1) * 4 + 64, s));
    const int nb = (n + 127) / 128;
    g3_count_kernel<<<nb, 128, 0, s>>>(n, G, ctx->box_lo.as<float>(), ctx->box_hi.as<float>(), count, info);
    unsigned long long hinfo[2] = {0, 0};
    VR_CUDA(cudaMemcpyAsync(hinfo, info, sizeof(hinfo), cudaMemcpyDeviceToHost, s));
    VR_CUDA(cudaStreamSynchronize(s));
    // non-finite or grid-spanning boxes, or lists that would not pay (more than 512 entries per component): no grid
    if (hinfo[1] != 0 || hinfo[0] > 512ULL * (unsigned long long)n + (1ULL << 20) || hinfo[0] >= 0xfffffff0ULL) return VR_OK;
    VR_CUDA(ctx->g3_gi.ensure(std::max<size_t>((size_t)hinfo[0], 1) * 4));
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, count, ctx->g3_start.as<uint32_t>(), (int)(cells + 1), s);
    VR_CUDA(ctx->lbvh_tmp.ensure(std::max<size_t>(tmp_bytes, 16)));
    // count[cells] is 0, so the scan over cells + 1 items closes the table: start[cells] = all entries
    if (cub::DeviceScan::ExclusiveSum(ctx->lbvh_tmp.p, tmp_bytes, count, ctx->g3_start.as<uint32_t>(), (int)(cells + 1), s) != cudaSuccess)
        return cuda_fail(ctx, cudaGetLastError(), "grid3d: scan");
    VR_CUDA(cudaMemsetAsync(count, 0, (cells + 1) * 4, s));   // reused as the fill cursors
    g3_fill_kernel<<<nb, 128, 0, s>>>(n, G, ctx->box_lo.as<float>(), ctx->box_hi.as<float>(), ctx->leaf_of_orig.as<uint32_t>(), ctx->g3_start.as<uint32_t>(),
                                      count, ctx->g3_gi.as<uint32_t>());
    g3_sort_kernel<<<(unsigned)((cells + 127) / 128), 128, 0, s>>>(cells, ctx->g3_start.as<uint32_t>(), ctx->g3_gi.as<uint32_t>());
    VR_CUDA(cudaGetLastError());
    ctx->g3_res = res;
    ctx->g3_entries = hinfo[0];
    ctx->g3_ok = true;
    return VR_OK;
}

}  // namespace vr
