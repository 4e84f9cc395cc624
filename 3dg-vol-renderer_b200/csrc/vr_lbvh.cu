// vr_lbvh.cu -- the culling hierarchy built ON THE DEVICE.
//
// Replaces GaussianMixtureModel::BuildBVH / SubdivideNode (gmm.h:231-260,318-446), which the reference re-runs for every
// apply_params_to_gmm_local (gmm.h:634-674), i.e. (1 + S) * views times per SFD iteration.  Like the host builder
// (vr_bvh.cpp) it produces a CULLING structure only -- the quadric test (gaussian.h:126-164) decides every overlap -- so the
// tree's shape cannot change a result, and all bit-identity tests are its acceptance tests.
//
// All of it runs on the context's stream, from the tight boxes of prepare_gaussians_kernel to the two node formats:
//   bbox_kernel        bounding box of the tight boxes (ordered-uint atomics)
//   inflate_kernel     reach of the ray origins (vr_bvh.h BoxReach: camera / caller-stated minimum) and the slack of the
//                      FP32 quadric test added to every box -- the same arithmetic as inflate_boxes()
//   morton_kernel      30-bit Morton code of the box centre | component index -> 62-bit keys (unique)
//   cub radix sort     keys; the low words are the leaf order
//   karras_kernel      binary radix tree over the sorted keys, one thread per internal node (Karras 2012)
//   fit_kernel         leaves walk up, the second arrival at a node unites its children's boxes; every node stores BOTH
//                      children's boxes (the layout the traversal kernels read, vr_bvh.h BvhNode)
//   quantize_kernel    16-bit boxes on one global grid, rounded outward by two cells (vr_bvh.h BvhNodeQ)
//   depth_kernel       deepest leaf (the traversal stack holds kStackDepth entries)
// One 64-byte read-back (bounding box, reach, grid, depth) and one synchronisation end the build.
// Tree quality: on the benchmark mixtures an LBVH costs 5-9 % more slab steps than the binned-SAH tree for rays that start
// inside the mixture and 2-9 % fewer for camera rays; it is built in ~0.2 ms instead of 5 ms of host time.
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cmath>
#include <cstring>

#include "vr_bvh.h"
#include "vr_context.h"

namespace vr {

// meta block in device memory (floats unless noted)
enum { M_BBLO = 0, M_BBHI = 3, M_R = 6, M_X = 7, M_QORG = 8, M_QCELL = 11, M_DEPTH = 14 /* int */, M_ROOTLO = 16, M_ROOTHI = 19, M_ENC = 24 /* 6 ordered uints */, kMetaWords = 32 };

__device__ __forceinline__ unsigned int f2ord(float f) { const unsigned int b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__device__ __forceinline__ float ord2f(unsigned int u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

__global__ void meta_init_kernel(unsigned int* meta) {
    const int i = threadIdx.x;
    if (i < kMetaWords) meta[i] = 0u;
    if (i < 3) { meta[M_ENC + i] = 0xffffffffu; meta[M_ENC + 3 + i] = 0u; }
}

__global__ void bbox_kernel(int n, const float* __restrict__ lo, const float* __restrict__ hi, unsigned int* meta) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float l[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, h[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
    if (i < n)
        for (int k = 0; k < 3; ++k) {   // a non-finite box (half extent 3e38) must not set the reach
            const float a = lo[3 * (size_t)i + k], b = hi[3 * (size_t)i + k];
            if (fabsf(a) < 1e30f) l[k] = a;
            if (fabsf(b) < 1e30f) h[k] = b;
        }
    for (int k = 0; k < 3; ++k) {
        for (int o = 16; o > 0; o >>= 1) {
            l[k] = fminf(l[k], __shfl_xor_sync(0xffffffffu, l[k], o));
            h[k] = fmaxf(h[k], __shfl_xor_sync(0xffffffffu, h[k], o));
        }
        if ((threadIdx.x & 31) == 0) {
            if (l[k] < CUDART_INF_F) atomicMin(meta + M_ENC + k, f2ord(l[k]));
            if (h[k] > -CUDART_INF_F) atomicMax(meta + M_ENC + 3 + k, f2ord(h[k]));
        }
    }
}

struct ReachArgs { int has_cam; float cam[3]; float min_R, min_X; double kappa; };

// BoxReach of vr_bvh.cpp (box_reach_of / box_reach) and inflate_boxes, per thread; thread 0 publishes box and reach
__global__ void inflate_kernel(int n, float* __restrict__ lo, float* __restrict__ hi, const float* __restrict__ lam, unsigned int* meta, ReachArgs ra) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float bl[3], bh[3];
    for (int k = 0; k < 3; ++k) {
        const unsigned int el = meta[M_ENC + k], eh = meta[M_ENC + 3 + k];
        bl[k] = ord2f(el); bh[k] = ord2f(eh);
        if (el == 0xffffffffu || eh == 0u || !(bl[k] <= bh[k])) { bl[k] = 0.f; bh[k] = 0.f; }
    }
    double diag2 = 0.0, maxabs = 0.0, camd2 = 0.0, camabs = 0.0;
    for (int k = 0; k < 3; ++k) {
        const double e = (double)bh[k] - (double)bl[k];
        diag2 += e * e;
        maxabs = fmax(maxabs, fmax(fabs((double)bl[k]), fabs((double)bh[k])));
        const double d = fmax(fabs((double)ra.cam[k] - (double)bl[k]), fabs((double)ra.cam[k] - (double)bh[k]));
        camd2 += d * d;
        camabs = fmax(camabs, fabs((double)ra.cam[k]));
    }
    const double diag = sqrt(diag2);
    float R = (float)(diag + 8.0), X = (float)(maxabs + 8.0);            // no camera known: origins within 8 units
    if (ra.has_cam) { R = fmaxf(R, (float)fmax(diag, sqrt(camd2) + 1.5)); X = fmaxf(X, (float)fmax(maxabs, camabs + 1.5)); }
    R = fmaxf(R, ra.min_R); X = fmaxf(X, ra.min_X);
    if (!isfinite(R)) R = 3.0e38f;
    if (!isfinite(X)) X = 3.0e38f;
    if (i == 0) {
        for (int k = 0; k < 3; ++k) { meta[M_BBLO + k] = __float_as_uint(bl[k]); meta[M_BBHI + k] = __float_as_uint(bh[k]); }
        meta[M_R] = __float_as_uint(R); meta[M_X] = __float_as_uint(X);
    }
    if (i >= n) return;
    const double eps = 5.9604644775390625e-08;   // 2^-24
    const double geom = (double)R * (2.0 * (double)R + (double)X);
    const float pad_abs = 4e-6f * fmaxf(16.0f, X);
    const double dq = ra.kappa * eps * (double)lam[i] * geom;          // inflate_boxes() of vr_bvh.cpp
    double rel = sqrt(1.0 + fmin(dq, 72.0) / 9.0) - 1.0;
    if (!(rel >= 0.003)) rel = 0.003;
    for (int k = 0; k < 3; ++k) {
        const float l = lo[3 * (size_t)i + k], h = hi[3 * (size_t)i + k];
        const float c = 0.5f * (l + h), e = 0.5f * (h - l);
        if (!(fabsf(e) < 1e30f)) continue;
        const float g = (float)((double)e * (1.0 + rel)) + pad_abs;
        lo[3 * (size_t)i + k] = c - g;
        hi[3 * (size_t)i + k] = c + g;
    }
}

__device__ __forceinline__ unsigned int expand10(unsigned int v) {
    v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu; v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__global__ void morton_kernel(int n, const float* __restrict__ lo, const float* __restrict__ hi, const unsigned int* __restrict__ meta,
                              unsigned long long* __restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned int q[3];
    for (int k = 0; k < 3; ++k) {
        const float bl = __uint_as_float(meta[M_BBLO + k]), bh = __uint_as_float(meta[M_BBHI + k]);
        const float c = 0.5f * (lo[3 * (size_t)i + k] + hi[3 * (size_t)i + k]);
        const float u = (c - bl) / fmaxf(bh - bl, 1e-30f);
        q[k] = (unsigned int)fminf(1023.0f, fmaxf(0.0f, u * 1024.0f));   // NaN -> 0
    }
    const unsigned long long code = (unsigned long long)((expand10(q[0]) << 2) | (expand10(q[1]) << 1) | expand10(q[2]));
    keys[i] = (code << 32) | (unsigned int)i;
}

// Karras 2012: internal node i covers the range of sorted keys sharing the longest common prefix around i
__device__ __forceinline__ int key_delta(const unsigned long long* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll((long long)(keys[i] ^ keys[j]));   // keys are unique (index in the low word)
}
__global__ void karras_kernel(int n, const unsigned long long* __restrict__ keys, int* __restrict__ left, int* __restrict__ right,
                              int* __restrict__ parent_node, int* __restrict__ parent_leaf, uint32_t* __restrict__ perm) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) perm[i] = (uint32_t)keys[i];
    if (i >= n - 1) return;
    const int d = (key_delta(keys, n, i, i + 1) - key_delta(keys, n, i, i - 1)) > 0 ? 1 : -1;
    const int dmin = key_delta(keys, n, i, i - d);
    int lmax = 2;
    while (key_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (key_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = key_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (key_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo_r = min(i, j), hi_r = max(i, j);
    const int lc = (lo_r == gamma) ? ~gamma : gamma;               // ~leaf position
    const int rc = (hi_r == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    left[i] = lc; right[i] = rc;
    if (lc >= 0) parent_node[lc] = i; else parent_leaf[~lc] = i;
    if (rc >= 0) parent_node[rc] = i; else parent_leaf[~rc] = i;
    if (i == 0) parent_node[0] = -1;
}

__device__ __forceinline__ void node_set_child(BvhNode* nodes, int p, int c, const float* l, const float* h, int code) {
    BvhNode& nd = nodes[p];
    if (c == 0) { nd.c0x[0] = l[0]; nd.c0x[1] = h[0]; nd.c0y[0] = l[1]; nd.c0y[1] = h[1]; nd.c0z[0] = l[2]; nd.c0z[1] = h[2]; }
    else { nd.c1x[0] = l[0]; nd.c1x[1] = h[0]; nd.c1y[0] = l[1]; nd.c1y[1] = h[1]; nd.c1z[0] = l[2]; nd.c1z[1] = h[2]; }
    nd.child[c] = code;
}
__device__ __forceinline__ void node_get_child(const BvhNode* nodes, int p, int c, float* l, float* h) {
    const volatile BvhNode& nd = nodes[p];
    if (c == 0) { l[0] = nd.c0x[0]; h[0] = nd.c0x[1]; l[1] = nd.c0y[0]; h[1] = nd.c0y[1]; l[2] = nd.c0z[0]; h[2] = nd.c0z[1]; }
    else { l[0] = nd.c1x[0]; h[0] = nd.c1x[1]; l[1] = nd.c1y[0]; h[1] = nd.c1y[1]; l[2] = nd.c1z[0]; h[2] = nd.c1z[1]; }
}
__global__ void fit_kernel(int n, const uint32_t* __restrict__ perm, const float* __restrict__ lo, const float* __restrict__ hi,
                           const int* __restrict__ left, const int* __restrict__ parent_node, const int* __restrict__ parent_leaf,
                           unsigned int* __restrict__ arrived, BvhNode* nodes, unsigned int* meta) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= n) return;
    const uint32_t g = perm[leaf];
    float cl[3], ch[3];
    for (int k = 0; k < 3; ++k) { cl[k] = lo[3 * (size_t)g + k]; ch[k] = hi[3 * (size_t)g + k]; }
    int code = ~((leaf << 3) | 1);       // one component per leaf (vr_bvh.h kMaxLeaf)
    int from = ~leaf;
    int p = parent_leaf[leaf];
    while (p >= 0) {
        const int c = (left[p] == from) ? 0 : 1;
        node_set_child(nodes, p, c, cl, ch, code);
        __threadfence();
        if (atomicAdd(arrived + p, 1u) == 0u) return;        // the sibling's subtree is not complete yet: its walker continues
        float sl[3], sh[3];
        node_get_child(nodes, p, 1 - c, sl, sh);
        for (int k = 0; k < 3; ++k) { cl[k] = fminf(cl[k], sl[k]); ch[k] = fmaxf(ch[k], sh[k]); }
        code = p; from = p;
        p = parent_node[p];
    }
    for (int k = 0; k < 3; ++k) { meta[M_ROOTLO + k] = __float_as_uint(cl[k]); meta[M_ROOTHI + k] = __float_as_uint(ch[k]); }   // box of the whole tree
}

// one component: child 0 = the leaf, child 1 = a leaf of zero components (vr_bvh.cpp does the same)
__global__ void single_leaf_kernel(const float* __restrict__ lo, const float* __restrict__ hi, BvhNode* nodes, uint32_t* perm, unsigned int* meta) {
    float l[3] = {lo[0], lo[1], lo[2]}, h[3] = {hi[0], hi[1], hi[2]};
    node_set_child(nodes, 0, 0, l, h, ~((0 << 3) | 1));
    float el[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, eh[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
    node_set_child(nodes, 0, 1, el, eh, ~0);
    nodes[0].pad[0] = nodes[0].pad[1] = 0;
    perm[0] = 0u;
    for (int k = 0; k < 3; ++k) { meta[M_ROOTLO + k] = __float_as_uint(l[k]); meta[M_ROOTHI + k] = __float_as_uint(h[k]); }
}

// vr_bvh.cpp quantize_bvh on the device: one global grid over the tree's box, child boxes rounded outward by two cells
__global__ void quantize_kernel(int n_nodes, const BvhNode* __restrict__ nodes, BvhNodeQ* __restrict__ out, unsigned int* meta) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float org[3], cell[3];
    for (int a = 0; a < 3; ++a) {
        double l = (double)__uint_as_float(meta[M_ROOTLO + a]), h = (double)__uint_as_float(meta[M_ROOTHI + a]);
        if (!(l <= h) || !isfinite(l) || !isfinite(h)) { l = 0.0; h = 0.0; }
        float c = fmaxf((float)((h - l) / 64000.0), 1e-20f);
        c = nextafterf(c, CUDART_INF_F);                      // never under-covers the extent
        cell[a] = c;
        org[a] = (float)(l - 6.0 * (double)c);
    }
    if (i == 0) for (int a = 0; a < 3; ++a) { meta[M_QORG + a] = __float_as_uint(org[a]); meta[M_QCELL + a] = __float_as_uint(cell[a]); }
    if (i >= n_nodes) return;
    BvhNodeQ q;
    for (int c = 0; c < 2; ++c) {
        float l[3], h[3];
        node_get_child(nodes, i, c, l, h);
        for (int a = 0; a < 3; ++a) {
            uint32_t ql = 0, qh = 0;                          // an empty child becomes a point outside the scene
            if (l[a] <= h[a] && isfinite(l[a]) && isfinite(h[a])) {
                const double fl = floor(((double)l[a] - (double)org[a]) / (double)cell[a]) - 2.0;
                const double fh = ceil(((double)h[a] - (double)org[a]) / (double)cell[a]) + 2.0;
                ql = (uint32_t)fmin(65535.0, fmax(0.0, fl));
                qh = (uint32_t)fmin(65535.0, fmax(0.0, fh));
            }
            q.box[3 * c + a] = ql | (qh << 16);
        }
        q.child[c] = nodes[i].child[c];
    }
    out[i] = q;
}

__global__ void depth_kernel(int n, const int* __restrict__ parent_node, const int* __restrict__ parent_leaf, unsigned int* meta) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    int depth = 0;
    if (leaf < n) for (int p = parent_leaf[leaf]; p >= 0; p = parent_node[p]) ++depth;
    for (int o = 16; o > 0; o >>= 1) depth = max(depth, __shfl_xor_sync(0xffffffffu, depth, o));
    if ((threadIdx.x & 31) == 0 && depth > 0) atomicMax((int*)meta + M_DEPTH, depth);
}

// Builds nodes / qnodes / perm of ctx from the TIGHT boxes in ctx->box_lo / box_hi and ctx->lam (all on the device); leaves the
// inflated boxes in box_lo / box_hi.  reach_RX: caller-stated minimum reach (nullptr: none).
int build_hierarchy_device(vr_context* ctx, int n, const float* reach_RX) {
    cudaStream_t s = ctx->stream;
    const int nb = (n + 127) / 128;
    const int n_nodes = std::max(n - 1, 1);
    VR_CUDA(ctx->nodes.ensure((size_t)n_nodes * sizeof(BvhNode)));
    VR_CUDA(ctx->qnodes.ensure((size_t)n_nodes * sizeof(BvhNodeQ)));
    VR_CUDA(ctx->lbvh_keys.ensure((size_t)n * 16));                               // keys in | keys out
    VR_CUDA(ctx->lbvh_tree.ensure(((size_t)n * 5 + 8) * 4 + kMetaWords * 4));    // left, right, parent_node (n-1 each), parent_leaf (n), arrived (n-1)
    unsigned long long* keys_in = ctx->lbvh_keys.as<unsigned long long>();
    unsigned long long* keys_out = keys_in + n;
    int* left = ctx->lbvh_tree.as<int>();
    int* right = left + n; int* parent_node = right + n; int* parent_leaf = parent_node + n;
    unsigned int* arrived = (unsigned int*)(parent_leaf + n);
    unsigned int* meta = arrived + n + 8;
    size_t tmp_bytes = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, keys_in, keys_out, n, 0, 62, s);
    VR_CUDA(ctx->lbvh_tmp.ensure(std::max<size_t>(tmp_bytes, 16)));

    VR_CUDA(cudaMemsetAsync(ctx->nodes.p, 0, (size_t)n_nodes * sizeof(BvhNode), s));
    meta_init_kernel<<<1, 32, 0, s>>>(meta);
    bbox_kernel<<<nb, 128, 0, s>>>(n, ctx->box_lo.as<float>(), ctx->box_hi.as<float>(), meta);
    ReachArgs ra{};
    ra.has_cam = ctx->has_camera ? 1 : 0;
    for (int k = 0; k < 3; ++k) ra.cam[k] = ctx->has_camera ? ctx->cam.pos[k] : 0.f;
    ra.min_R = reach_RX ? reach_RX[0] : 0.f; ra.min_X = reach_RX ? reach_RX[1] : 0.f;
    ra.kappa = ctx->box_kappa;
    inflate_kernel<<<nb, 128, 0, s>>>(n, ctx->box_lo.as<float>(), ctx->box_hi.as<float>(), ctx->lam.as<float>(), meta, ra);
    if (n == 1) {
        single_leaf_kernel<<<1, 1, 0, s>>>(ctx->box_lo.as<float>(), ctx->box_hi.as<float>(), ctx->nodes.as<BvhNode>(), ctx->perm.as<uint32_t>(), meta);
    } else {
        morton_kernel<<<nb, 128, 0, s>>>(n, ctx->box_lo.as<float>(), ctx->box_hi.as<float>(), meta, keys_in);
        if (cub::DeviceRadixSort::SortKeys(ctx->lbvh_tmp.p, tmp_bytes, keys_in, keys_out, n, 0, 62, s) != cudaSuccess)
            return cuda_fail(ctx, cudaGetLastError(), "hierarchy build: radix sort");
        VR_CUDA(cudaMemsetAsync(arrived, 0, (size_t)n * 4, s));
        karras_kernel<<<nb, 128, 0, s>>>(n, keys_out, left, right, parent_node, parent_leaf, ctx->perm.as<uint32_t>());
        fit_kernel<<<nb, 128, 0, s>>>(n, ctx->perm.as<uint32_t>(), ctx->box_lo.as<float>(), ctx->box_hi.as<float>(), left, parent_node, parent_leaf,
                                      arrived, ctx->nodes.as<BvhNode>(), meta);
        depth_kernel<<<nb, 128, 0, s>>>(n, parent_node, parent_leaf, meta);
    }
    quantize_kernel<<<(n_nodes + 127) / 128, 128, 0, s>>>(n_nodes, ctx->nodes.as<BvhNode>(), ctx->qnodes.as<BvhNodeQ>(), meta);
    VR_CUDA(cudaGetLastError());
    VR_CUDA(cudaMemcpyAsync(ctx->h_build_meta, meta, kMetaWords * 4, cudaMemcpyDeviceToHost, s));
    VR_CUDA(cudaStreamSynchronize(s));
    const unsigned int* hm = ctx->h_build_meta;
    auto F = [&](int i) { float f; memcpy(&f, hm + i, 4); return f; };
    for (int k = 0; k < 3; ++k) { ctx->bb_lo[k] = F(M_BBLO + k); ctx->bb_hi[k] = F(M_BBHI + k); ctx->qorg[k] = F(M_QORG + k); ctx->qcell[k] = F(M_QCELL + k); }
    ctx->built_R = F(M_R); ctx->built_X = F(M_X);
    ctx->n_nodes = n_nodes;
    ctx->bvh_depth = n == 1 ? 1 : (int)hm[M_DEPTH];
    if (ctx->bvh_depth > kStackDepth - 2) return fail(ctx, VR_ERR_UNSUPPORTED, "hierarchy deeper than the traversal stack");
    return VR_OK;
}

}  // namespace vr
