// vr_bvh.h -- host-side builder of the culling hierarchy over the 3-sigma ellipsoids.
//
// Replaces GaussianMixtureModel::BuildBVH / SubdivideNode (gmm.h:231-446) as a *culling* structure
// only: the quadric test (gaussian.h:126-164) still decides every overlap, so the set of
// components a ray gathers is the reference's; the hierarchy itself is free to differ.  Ours is a
// binned-SAH binary tree over tight boxes (half extent 3*sqrt(Sigma_kk), padded), flattened so
// that one 64-byte node holds BOTH children's boxes (four 128-bit loads per traversal step), with
// the Gaussians permuted into leaf order so that a leaf is a contiguous run of 48-byte records.
#pragma once
#include <cstdint>
#include <vector>

namespace vr {

struct BvhNode {       // 64 B, 16-byte aligned, read as 4 x float4
    float c0x[2], c0y[2];  // child 0: (lo.x, hi.x, lo.y, hi.y)
    float c1x[2], c1y[2];  // child 1
    float c0z[2], c1z[2];  // (lo0.z, hi0.z, lo1.z, hi1.z)
    int32_t child[2];      // >= 0: node index; < 0: leaf, ~((first << 3) | count), count 0..7
    int32_t pad[2];
};
static_assert(sizeof(BvhNode) == 64, "BvhNode must be 64 bytes");

struct BvhBuildResult {
    std::vector<BvhNode> nodes;     // nodes[0] is the root (always present when n > 0)
    std::vector<uint32_t> perm;     // leaf order -> original Gaussian index
    int max_depth = 0;
    int max_leaf = 0;
};

constexpr int kMaxLeaf = 1;         // one Gaussian per leaf: the traversal kernels queue a leaf CHILD as a candidate directly

// Compact form of the same tree for the wavefront traversal kernels (vr_wave.cu), whose bound is the L1
// data pipe (profiles/r01_v5): 32 B per node instead of 64, i.e. two 128-bit loads per step.  Child boxes
// are quantised to 16 bits per coordinate on ONE global grid (org + q * cell) and rounded OUTWARD by two
// extra cells (one for the float rounding of the slab arithmetic, one for the folded 2^23 offset, vr_device.cuh), so a quantised box always contains the float box: the hierarchy only culls and the
// quadric test still decides every overlap, hence results cannot change.
struct BvhNodeQ {
    uint32_t box[6];       // child 0: x, y, z then child 1: x, y, z; each word = q_lo | (q_hi << 16)
    int32_t child[2];      // same codes as BvhNode::child
};
static_assert(sizeof(BvhNodeQ) == 32, "BvhNodeQ must be 32 bytes");
struct BvhQuantGrid { float org[3], cell[3]; };
void quantize_bvh(const BvhBuildResult& tree, BvhQuantGrid* grid, std::vector<BvhNodeQ>* out);

// lo/hi: 3*n floats each (per-Gaussian boxes).  Deterministic.
void build_bvh(int n, const float* lo, const float* hi, BvhBuildResult* out);

// Slack of the culling boxes.  The reference decides an overlap with an FP32 quadric test (gaussian.h:126-164) whose
// discriminant B^2 - 4A(C0 - 9) cancels catastrophically: with p = o - mu, C0 = p^T M p ~ lam * R^2 (lam = largest eigenvalue
// of Sigma^-1, R = distance from the ray origin), the accepted squared Mahalanobis radius is 9 +- O(eps * lam * R^2), and
// p itself carries the rounding of coordinates of magnitude X: another O(eps * lam * R * X).  A box that is to contain every
// ray the FLOAT test accepts must therefore grow with the reach of the ray origins.  BoxReach states that reach:
//   R = the farthest a ray origin can be from a point of the mixture, X = the largest coordinate magnitude of origins and
//   mixture.  inflate_boxes grows the tight 3-sigma half extents h to h * max(1.003, sqrt(1 + dq / 9)), dq = kappa * eps * lam *
//   R * (2 R + X) the slack of the accepted squared radius, + 4e-6 * max(16, X) (the absolute term covers the slab
//   arithmetic of the traversal, vr_device.cuh).
struct BoxReach { float R, X; };
// bounding box of the tight boxes -> bb_lo/bb_hi[3]; cam_pos = nullptr: no camera known yet (the reach then assumes ray
// origins within 8 units of the mixture's bounding box)
BoxReach box_reach(int n, const float* lo, const float* hi, const float* cam_pos, float* bb_lo, float* bb_hi);
// the reach a camera at cam_pos needs over a mixture bounded by bb_lo/bb_hi (film-plane offsets of camera.h:45-73 included)
BoxReach box_reach_of(const float* bb_lo, const float* bb_hi, const float* cam_pos);
// kappa: ulps of the discriminant covered.  tools/gather_sweep.py finds no missed component on 60 000 camera and interior rays
// of the C2 mixture even without any slack, but rays from 100 units out (the quadric's error grows with the square of the
// distance) do lose grazing pseudo-overlaps below ~1; 1.5 costs < 1 % at C2 (profiles/r02/gather_sweep_kappa.log)
constexpr double kBoxKappa = 1.5;
void inflate_boxes(int n, float* lo, float* hi, const float* lam, BoxReach reach, double kappa = kBoxKappa);

}  // namespace vr
