// vr_bvh.cpp -- binned-SAH builder (host).  See vr_bvh.h.
#include "vr_bvh.h"

#include <algorithm>
#include <cmath>
#include <atomic>
#include <cstdlib>
#ifdef __linux__
#include <sched.h>
#endif
#include <limits>
#include <thread>

namespace vr {
namespace {

struct Box {
    float lo[3], hi[3];
    void reset() {
        for (int k = 0; k < 3; ++k) { lo[k] = std::numeric_limits<float>::infinity(); hi[k] = -lo[k]; }
    }
    void grow(const float* l, const float* h) {
        for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], l[k]); hi[k] = std::max(hi[k], h[k]); }
    }
    void grow(const Box& b) { grow(b.lo, b.hi); }
    float half_area() const {
        float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        if (!(ex >= 0.f) || !(ey >= 0.f) || !(ez >= 0.f)) return 0.f;
        return ex * ey + ey * ez + ez * ex;
    }
};

struct Builder {
    int n;
    const float* lo;
    const float* hi;
    const float* cen;     // 3 * n centroids, 0.5f * (lo + hi)
    std::vector<uint32_t>& perm;
    std::vector<BvhNode>& nodes;
    int max_depth = 0, max_leaf = 0;

    static constexpr int kBins = 16;
    static constexpr float kTrav = 1.0f, kIsect = 2.0f;

    Box range_box(int first, int count) const {
        Box b; b.reset();
        for (int i = first; i < first + count; ++i) b.grow(lo + 3 * perm[i], hi + 3 * perm[i]);
        return b;
    }

    static int leaf_code(int first, int count) { return ~((first << 3) | count); }

    void set_child_box(BvhNode& nd, int c, const Box& b) {
        if (c == 0) {
            nd.c0x[0] = b.lo[0]; nd.c0x[1] = b.hi[0]; nd.c0y[0] = b.lo[1]; nd.c0y[1] = b.hi[1];
            nd.c0z[0] = b.lo[2]; nd.c0z[1] = b.hi[2];
        } else {
            nd.c1x[0] = b.lo[0]; nd.c1x[1] = b.hi[0]; nd.c1y[0] = b.lo[1]; nd.c1y[1] = b.hi[1];
            nd.c1z[0] = b.lo[2]; nd.c1z[1] = b.hi[2];
        }
    }

    // chooses a split of [first, first+count); returns the size of the left part (0 = make a leaf).
    // One pass over the range fills the bins of all three axes (the boxes are read once instead of three times); the
    // sweeps then pick the cheapest (axis, bin) in axis order with a strict comparison, so ties go to the lower axis / bin.
    int split(int first, int count, const Box& box) {
        Box cb; cb.reset();
        for (int i = first; i < first + count; ++i) { const float* c = cen + 3 * (size_t)perm[i]; cb.grow(c, c); }
        if (count == 2 && kMaxLeaf == 1) {
            // two components are always separated (a leaf holds one): every axis on which the centroids differ puts them
            // into the first and the last bin at the same cost, so the lowest such axis wins and the nearer one goes left
            for (int axis = 0; axis < 3; ++axis) {
                if (!(cb.hi[axis] - cb.lo[axis] > 0.f)) continue;
                if (cen[3 * (size_t)perm[first] + axis] > cen[3 * (size_t)perm[first + 1] + axis]) std::swap(perm[first], perm[first + 1]);
                return 1;
            }
            return 1;   // coinciding centroids: split by index
        }
        float cmin[3], scale[3];
        bool use[3];
        for (int axis = 0; axis < 3; ++axis) {
            const float cext = cb.hi[axis] - cb.lo[axis];
            use[axis] = cext > 0.f;
            cmin[axis] = cb.lo[axis];
            scale[axis] = use[axis] ? kBins / cext : 0.f;
        }
        Box bb[3][kBins]; int bc[3][kBins];
        for (int axis = 0; axis < 3; ++axis)
            for (int b = 0; b < kBins; ++b) { bb[axis][b].reset(); bc[axis][b] = 0; }
        for (int i = first; i < first + count; ++i) {
            const uint32_t g = perm[i];
            const float* c = cen + 3 * (size_t)g;
            for (int axis = 0; axis < 3; ++axis) {
                if (!use[axis]) continue;
                const int b = std::min(kBins - 1, std::max(0, (int)((c[axis] - cmin[axis]) * scale[axis])));
                bb[axis][b].grow(lo + 3 * (size_t)g, hi + 3 * (size_t)g); bc[axis][b]++;
            }
        }
        float best_cost = std::numeric_limits<float>::infinity();
        int best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; ++axis) {
            if (!use[axis]) continue;
            float right_area[kBins]; int right_cnt[kBins];
            Box acc; acc.reset(); int cnt = 0;
            for (int b = kBins - 1; b > 0; --b) { acc.grow(bb[axis][b]); cnt += bc[axis][b]; right_area[b] = acc.half_area(); right_cnt[b] = cnt; }
            acc.reset(); cnt = 0;
            for (int b = 0; b < kBins - 1; ++b) {
                acc.grow(bb[axis][b]); cnt += bc[axis][b];
                if (cnt == 0 || right_cnt[b + 1] == 0) continue;
                float cost = acc.half_area() * cnt + right_area[b + 1] * right_cnt[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
            }
        }
        float parent_area = box.half_area();
        if (best_axis >= 0 && count <= kMaxLeaf) {
            float split_cost = kTrav + kIsect * best_cost / std::max(parent_area, 1e-30f);
            if (split_cost >= kIsect * count) return 0;  // cheaper as a leaf
        }
        if (best_axis < 0) {
            if (count <= kMaxLeaf) return 0;
            // all centroids coincide: median split by index to bound the leaf size
            return count / 2;
        }
        const float pmin = cmin[best_axis], pscale = scale[best_axis];
        const int axis = best_axis, bin = best_bin;
        auto mid = std::stable_partition(perm.begin() + first, perm.begin() + first + count, [&](uint32_t g) {
            int b = std::min(kBins - 1, std::max(0, (int)((cen[3 * (size_t)g + axis] - pmin) * pscale)));
            return b <= bin;
        });
        int left = (int)(mid - (perm.begin() + first));
        if (left == 0 || left == count) return count <= kMaxLeaf ? 0 : count / 2;
        return left;
    }

    // a subtree left to a worker thread: nodes[ni].child[c] is patched once the subtree's place in `nodes` is known
    struct Task { int ni, c, first, count, left, depth; };
    std::vector<Task>* defer = nullptr;   // non-null: subtrees below kDeferDepth with enough work are listed, not built
    static constexpr int kDeferDepth = 4, kDeferMinCount = 512;

    // builds the subtree over [first, first+count) (count >= 2, to be split) into nodes[ni]
    void build_node(int ni, int first, int count, int left, int depth) {
        max_depth = std::max(max_depth, depth);
        int parts[2][2] = {{first, left}, {first + left, count - left}};
        for (int c = 0; c < 2; ++c) {
            int f = parts[c][0], k = parts[c][1];
            Box b = range_box(f, k);
            set_child_box(nodes[ni], c, b);
            int l = (k >= 2) ? split(f, k, b) : 0;
            if (l == 0) {
                nodes[ni].child[c] = leaf_code(f, k);
                max_leaf = std::max(max_leaf, k);
            } else if (defer && depth >= kDeferDepth && k >= kDeferMinCount) {
                defer->push_back(Task{ni, c, f, k, l, depth + 1});
            } else {
                int ci = (int)nodes.size();
                nodes.push_back(BvhNode{});
                nodes[ni].child[c] = ci;
                build_node(ci, f, k, l, depth + 1);
            }
        }
    }
};

}  // namespace

void build_bvh(int n, const float* lo, const float* hi, BvhBuildResult* out) {
    out->nodes.clear();
    out->perm.resize((size_t)std::max(n, 0));
    out->max_depth = 0; out->max_leaf = 0;
    if (n <= 0) return;
    for (int i = 0; i < n; ++i) out->perm[i] = (uint32_t)i;
    std::vector<float> cen((size_t)3 * n);
    for (size_t k = 0; k < cen.size(); ++k) cen[k] = 0.5f * (lo[k] + hi[k]);
    Builder b{n, lo, hi, cen.data(), out->perm, out->nodes};
    out->nodes.reserve((size_t)n);
    out->nodes.push_back(BvhNode{});
    Box root = b.range_box(0, n);
    int left = (n >= 2) ? b.split(0, n, root) : 0;
    if (left == 0) {
        // a single leaf: child 0 = everything (n <= kMaxLeaf here), child 1 = a leaf of ZERO Gaussians
        // (an inverted box is not enough: min/max in the slab test re-order lo and hi)
        BvhNode& nd = out->nodes[0];
        b.set_child_box(nd, 0, root);
        Box empty; empty.reset();
        b.set_child_box(nd, 1, empty);
        nd.child[0] = Builder::leaf_code(0, n);
        nd.child[1] = Builder::leaf_code(0, 0);
        b.max_leaf = n;
    } else if (n < 4096) {
        b.build_node(0, 0, n, left, 1);
    } else {
        // The top levels are built here; every subtree below them works on its own range of `perm`, so the subtrees are
        // built by worker threads into private node arrays and appended in task order.  The task list depends only on the
        // input, not on the thread count or timing: the tree is the same on every run and every machine.
        std::vector<Builder::Task> tasks;
        b.defer = &tasks;
        b.build_node(0, 0, n, left, 1);
        b.defer = nullptr;
        struct Sub { std::vector<BvhNode> nodes; int max_depth = 0, max_leaf = 0; };
        std::vector<Sub> subs(tasks.size());
        std::atomic<size_t> next{0};
        auto work = [&] {
            for (size_t t; (t = next.fetch_add(1)) < tasks.size();) {
                const Builder::Task& tk = tasks[t];
                Builder lb{n, lo, hi, cen.data(), out->perm, subs[t].nodes};
                subs[t].nodes.reserve((size_t)tk.count);
                subs[t].nodes.push_back(BvhNode{});
                lb.build_node(0, tk.first, tk.count, tk.left, tk.depth);
                subs[t].max_depth = lb.max_depth; subs[t].max_leaf = lb.max_leaf;
            }
        };
        // host cores this process may use, shared with the other ranks of the node when launched by torchrun (one process
        // per GPU: 8 ranks x 16 workers on a 16-core host only add scheduling latency to a ~1 ms build)
        unsigned hw = std::max(1u, std::thread::hardware_concurrency());
#ifdef __linux__
        cpu_set_t cpus;
        if (sched_getaffinity(0, sizeof(cpus), &cpus) == 0) hw = std::max(1u, (unsigned)CPU_COUNT(&cpus));
#endif
        if (const char* lws = std::getenv("LOCAL_WORLD_SIZE")) hw = std::max(1u, hw / (unsigned)std::max(1, std::atoi(lws)));
        const size_t nthreads = std::min<size_t>(std::min<size_t>(tasks.size(), hw), 16);
        std::vector<std::thread> pool;
        for (size_t t = 1; t < nthreads; ++t) pool.emplace_back(work);
        work();
        for (std::thread& th : pool) th.join();
        for (size_t t = 0; t < tasks.size(); ++t) {
            const int offset = (int)out->nodes.size();
            out->nodes[(size_t)tasks[t].ni].child[tasks[t].c] = offset;
            for (BvhNode nd : subs[t].nodes) {
                for (int c = 0; c < 2; ++c) if (nd.child[c] >= 0) nd.child[c] += offset;
                out->nodes.push_back(nd);
            }
            b.max_depth = std::max(b.max_depth, subs[t].max_depth);
            b.max_leaf = std::max(b.max_leaf, subs[t].max_leaf);
        }
    }
    out->max_depth = b.max_depth;
    out->max_leaf = b.max_leaf;
}

BoxReach box_reach_of(const float* bb_lo, const float* bb_hi, const float* cam_pos) {
    double diag2 = 0.0, maxabs = 0.0, camd2 = 0.0, camabs = 0.0;
    for (int k = 0; k < 3; ++k) {
        const double e = (double)bb_hi[k] - (double)bb_lo[k];
        diag2 += e * e;
        maxabs = std::max(maxabs, std::max(std::fabs((double)bb_lo[k]), std::fabs((double)bb_hi[k])));
        if (cam_pos) {
            // farthest point of the box from the camera, per axis
            const double d = std::max(std::fabs((double)cam_pos[k] - (double)bb_lo[k]), std::fabs((double)cam_pos[k] - (double)bb_hi[k]));
            camd2 += d * d;
            camabs = std::max(camabs, std::fabs((double)cam_pos[k]));
        }
    }
    BoxReach r;
    const double diag = std::sqrt(diag2);
    // ray origins: scatter positions inside the mixture (bounce and shadow rays) and points of the camera's film plane,
    // pos + u right + v up with |u|, |v| <= 1 (camera.h:45-73): sqrt(2) from pos, rounded up to 1.5
    if (cam_pos) { r.R = (float)std::max(diag, std::sqrt(camd2) + 1.5); r.X = (float)std::max(maxabs, camabs + 1.5); }
    else { r.R = (float)(diag + 8.0); r.X = (float)(maxabs + 8.0); }
    if (!std::isfinite(r.R)) r.R = 3.0e38f;
    if (!std::isfinite(r.X)) r.X = 3.0e38f;
    return r;
}

BoxReach box_reach(int n, const float* lo, const float* hi, const float* cam_pos, float* bb_lo, float* bb_hi) {
    for (int k = 0; k < 3; ++k) { bb_lo[k] = std::numeric_limits<float>::infinity(); bb_hi[k] = -bb_lo[k]; }
    for (int i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) {
            const float l = lo[3 * (size_t)i + k], h = hi[3 * (size_t)i + k];
            if (std::fabs(l) < 1e30f) bb_lo[k] = std::min(bb_lo[k], l);     // a non-finite box (h = 3e38) must not set the reach
            if (std::fabs(h) < 1e30f) bb_hi[k] = std::max(bb_hi[k], h);
        }
    for (int k = 0; k < 3; ++k) if (!(bb_lo[k] <= bb_hi[k])) { bb_lo[k] = 0.f; bb_hi[k] = 0.f; }
    return box_reach_of(bb_lo, bb_hi, cam_pos);
}

void inflate_boxes(int n, float* lo, float* hi, const float* lam, BoxReach reach, double kappa) {
    const double eps = 5.9604644775390625e-08;   // 2^-24
    const double geom = (double)reach.R * (2.0 * (double)reach.R + (double)reach.X);
    const float pad_abs = 4e-6f * std::max(16.0f, reach.X);
    for (int i = 0; i < n; ++i) {
        // the FP32 quadric accepts squared Mahalanobis radii up to 9 + dq, dq ~ kappa * eps * lam * R * (2 R + X): the box of
        // that ellipsoid is sqrt(1 + dq / 9) times the 3-sigma box (never less than +0.3 %, never more than x3)
        const double dq = kappa * eps * (double)lam[i] * geom;
        double rel = std::sqrt(1.0 + std::min(dq, 72.0) / 9.0) - 1.0;
        if (!(rel >= 0.003)) rel = 0.003;
        for (int k = 0; k < 3; ++k) {
            const float l = lo[3 * (size_t)i + k], h = hi[3 * (size_t)i + k];
            const float c = 0.5f * (l + h), e = 0.5f * (h - l);
            if (!(std::fabs(e) < 1e30f)) continue;                           // already "everything"
            const float g = (float)((double)e * (1.0 + rel)) + pad_abs;
            lo[3 * (size_t)i + k] = c - g;
            hi[3 * (size_t)i + k] = c + g;
        }
    }
}

void quantize_bvh(const BvhBuildResult& tree, BvhQuantGrid* grid, std::vector<BvhNodeQ>* out) {
    out->clear();
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    auto child_box = [](const BvhNode& nd, int c, int axis, float* l, float* h) {
        const float* p = axis == 0 ? (c == 0 ? nd.c0x : nd.c1x) : axis == 1 ? (c == 0 ? nd.c0y : nd.c1y) : (c == 0 ? nd.c0z : nd.c1z);
        *l = p[0]; *h = p[1];
    };
    for (const BvhNode& nd : tree.nodes)
        for (int c = 0; c < 2; ++c)
            for (int a = 0; a < 3; ++a) {
                float l, h;
                child_box(nd, c, a, &l, &h);
                if (!(l <= h) || !std::isfinite(l) || !std::isfinite(h)) continue;
                lo[a] = std::min(lo[a], (double)l); hi[a] = std::max(hi[a], (double)h);
            }
    const double kSteps = 64000.0;   // leaves room for the outward rounding below
    for (int a = 0; a < 3; ++a) {
        if (!(lo[a] <= hi[a])) { lo[a] = 0.0; hi[a] = 0.0; }
        grid->cell[a] = std::max((float)((hi[a] - lo[a]) / kSteps), 1e-20f);
        grid->cell[a] = std::nextafter(grid->cell[a], std::numeric_limits<float>::infinity());  // never under-covers the extent
        grid->org[a] = (float)(lo[a] - 6.0 * (double)grid->cell[a]);
    }
    out->resize(tree.nodes.size());
    for (size_t i = 0; i < tree.nodes.size(); ++i) {
        const BvhNode& nd = tree.nodes[i];
        BvhNodeQ& q = (*out)[i];
        for (int c = 0; c < 2; ++c) {
            for (int a = 0; a < 3; ++a) {
                float l, h;
                child_box(nd, c, a, &l, &h);
                uint32_t ql = 0, qh = 0;   // an empty child (zero-count leaf) becomes a point outside the scene
                if (l <= h && std::isfinite(l) && std::isfinite(h)) {
                    const double org = (double)grid->org[a], cell = (double)grid->cell[a];
                    const double fl = std::floor(((double)l - org) / cell) - 2.0, fh = std::ceil(((double)h - org) / cell) + 2.0;
                    ql = (uint32_t)std::min(65535.0, std::max(0.0, fl));
                    qh = (uint32_t)std::min(65535.0, std::max(0.0, fh));
                }
                q.box[3 * c + a] = ql | (qh << 16);
            }
            q.child[c] = nd.child[c];
        }
    }
}

}  // namespace vr
