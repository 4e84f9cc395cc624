// Host-only checks of the two culling structures (no GPU): vr_bvh.cpp and vr_lightgrid.cpp.
//   usage: test_builders <n> <seed>     prints "ok <name>" per check, exits non-zero on the first failure
// Both structures only PROPOSE candidates to the reference's quadric test, so what must hold is conservativeness:
//   * hierarchy: every component sits in exactly one leaf, every child box contains the boxes below it, and the
//     quantised box of a child contains its float box; the tree is identical run to run (threaded build);
//   * light grid: for any point x, the list of the cell that contains x - light holds every component whose padded box
//     the segment light -> x crosses, in ascending distance order.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "../vr_bvh.h"
#include "../vr_lightgrid.h"

#define CHECK(c) do { if (!(c)) { std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

static void child_box(const vr::BvhNode& nd, int c, float lo[3], float hi[3]) {
    const float* x = c == 0 ? nd.c0x : nd.c1x; const float* y = c == 0 ? nd.c0y : nd.c1y; const float* z = c == 0 ? nd.c0z : nd.c1z;
    lo[0] = x[0]; hi[0] = x[1]; lo[1] = y[0]; hi[1] = y[1]; lo[2] = z[0]; hi[2] = z[1];
}

// returns the number of components below (node ni, child c) and checks containment on the way
static long walk(const vr::BvhBuildResult& t, const std::vector<vr::BvhNodeQ>& q, const vr::BvhQuantGrid& g, const float* lo, const float* hi,
                 int ni, std::vector<int>& seen, bool* ok) {
    long total = 0;
    for (int c = 0; c < 2; ++c) {
        float bl[3], bh[3];
        child_box(t.nodes[(size_t)ni], c, bl, bh);
        const int code = t.nodes[(size_t)ni].child[c];
        if (q[(size_t)ni].child[c] != code) *ok = false;
        if (code < 0 && ((~code) & 7) == 0) continue;   // empty leaf
        for (int a = 0; a < 3; ++a) {   // quantised box contains the float box
            const uint32_t w = q[(size_t)ni].box[3 * c + a];
            const double ql = (double)g.org[a] + (double)(w & 0xffffu) * (double)g.cell[a], qh = (double)g.org[a] + (double)(w >> 16) * (double)g.cell[a];
            if (!(ql <= (double)bl[a] && qh >= (double)bh[a])) *ok = false;
        }
        if (code >= 0) {
            // the child's own two boxes lie inside this box
            for (int cc = 0; cc < 2; ++cc) {
                float l2[3], h2[3];
                child_box(t.nodes[(size_t)code], cc, l2, h2);
                const int c2 = t.nodes[(size_t)code].child[cc];
                if (c2 < 0 && ((~c2) & 7) == 0) continue;
                for (int a = 0; a < 3; ++a) if (!(l2[a] >= bl[a] && h2[a] <= bh[a])) *ok = false;
            }
            total += walk(t, q, g, lo, hi, code, seen, ok);
        } else {
            const uint32_t u = (uint32_t)~code;
            const uint32_t first = u >> 3, cnt = u & 7u;
            for (uint32_t i = 0; i < cnt; ++i) {
                const uint32_t orig = t.perm[first + i];
                ++seen[orig];
                for (int a = 0; a < 3; ++a) if (!(lo[3 * orig + a] >= bl[a] && hi[3 * orig + a] <= bh[a])) *ok = false;
            }
            total += cnt;
        }
    }
    return total;
}

int main(int argc, char** argv) {
    const int n = argc > 1 ? std::atoi(argv[1]) : 5000;
    const unsigned seed = argc > 2 ? (unsigned)std::atoi(argv[2]) : 1u;
    std::mt19937 r(seed);
    std::uniform_real_distribution<float> U(-1.f, 1.f);
    std::vector<float> lo(3 * (size_t)n), hi(3 * (size_t)n);
    for (int i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) {
            const float c = U(r) + (k == 1 ? 1.f : 0.f), h = 0.005f + 0.03f * (U(r) + 1.f) * (U(r) + 1.f);
            lo[3 * i + k] = c - h; hi[3 * i + k] = c + h;
        }
    // ---- hierarchy
    vr::BvhBuildResult t, t2;
    vr::build_bvh(n, lo.data(), hi.data(), &t);
    vr::build_bvh(n, lo.data(), hi.data(), &t2);
    CHECK(t.nodes.size() == t2.nodes.size() && t.perm == t2.perm);
    for (size_t i = 0; i < t.nodes.size(); ++i) CHECK(t.nodes[i].child[0] == t2.nodes[i].child[0] && t.nodes[i].child[1] == t2.nodes[i].child[1]);
    std::puts("ok bvh_deterministic");
    vr::BvhQuantGrid g; std::vector<vr::BvhNodeQ> q;
    vr::quantize_bvh(t, &g, &q);
    CHECK(q.size() == t.nodes.size());
    std::vector<int> seen((size_t)n, 0);
    bool ok = true;
    const long total = walk(t, q, g, lo.data(), hi.data(), 0, seen, &ok);
    CHECK(ok && total == n);
    for (int i = 0; i < n; ++i) CHECK(seen[(size_t)i] == 1);
    CHECK(t.max_leaf <= vr::kMaxLeaf || n == 1);
    std::puts("ok bvh_partition_and_containment");
    // ---- light grids
    std::vector<uint32_t> leaf_of_orig((size_t)n);
    for (int i = 0; i < n; ++i) leaf_of_orig[t.perm[(size_t)i]] = (uint32_t)i;
    const float lights[9] = {0.f, 5.f, 0.1f, -3.f, 3.f, 0.3f, 3.f, 3.f, -0.2f};
    const int res = 64;
    vr::LightGridHost lg;
    vr::build_light_grids(n, lo.data(), hi.data(), leaf_of_orig.data(), 3, lights, res, &lg);
    CHECK(lg.mask == 7u);
    const size_t cells = (size_t)6 * res * res;
    for (int l = 0; l < 3; ++l)
        for (size_t c = 0; c < cells; ++c) {
            const uint32_t a = lg.cell_start[(size_t)l * (cells + 1) + c], b = lg.cell_start[(size_t)l * (cells + 1) + c + 1];
            CHECK(a <= b && b <= lg.entry_gi.size());
            for (uint32_t e = a + 1; e < b; ++e) CHECK(lg.entry_dmin[e - 1] <= lg.entry_dmin[e]);
        }
    std::puts("ok lightgrid_sorted");
    long crossed = 0;
    for (int trial = 0; trial < 4000; ++trial) {
        const int l = trial % 3;
        const double L[3] = {lights[3 * l], lights[3 * l + 1], lights[3 * l + 2]};
        const double x[3] = {U(r) * 1.2, U(r) * 1.2 + 1.0, U(r) * 1.2};
        const double v[3] = {x[0] - L[0], x[1] - L[1], x[2] - L[2]};
        const double len = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        const int cell = vr::light_grid_cell(v, res);
        const uint32_t a = lg.cell_start[(size_t)l * (cells + 1) + (size_t)cell], b = lg.cell_start[(size_t)l * (cells + 1) + (size_t)cell + 1];
        for (int i = 0; i < n; ++i) {   // slab test of the segment L -> x against component i's box
            double t0 = 0.0, t1 = 1.0;
            bool hit = true;
            for (int k = 0; k < 3 && hit; ++k) {
                const double bl = lo[3 * i + k], bh = hi[3 * i + k];
                if (std::fabs(v[k]) < 1e-300) { hit = L[k] >= bl && L[k] <= bh; continue; }
                double ta = (bl - L[k]) / v[k], tb = (bh - L[k]) / v[k];
                if (ta > tb) std::swap(ta, tb);
                t0 = std::max(t0, ta); t1 = std::min(t1, tb);
                hit = t0 <= t1;
            }
            if (!hit) continue;
            ++crossed;
            bool listed = false;
            for (uint32_t e = a; e < b && !listed; ++e) listed = lg.entry_gi[e] == leaf_of_orig[(size_t)i] && (double)lg.entry_dmin[e] <= len;
            CHECK(listed);
        }
    }
    CHECK(crossed > 0);
    std::puts("ok lightgrid_conservative");
    std::puts("all builder checks passed");
    return 0;
}
