// gmm.h -- GaussianMixtureModel of the drop-in API and the 11-float parameterisation of the
// inverse path (reference gmm.h:35-706).
//
// The reference's BVH (gmm.h:231-515) is replaced by the device hierarchy vr_set_gmm builds, so the
// host model is the public `gaussians` vector plus pack / apply.  Parameter layout per Gaussian
// (gmm.h:583-628): mean(3), Rodrigues rotation vector(3), log scale(3), log density, logit albedo.
// apply_params_to_gmm_local goes through vr_unpack_params, i.e. the SAME device arithmetic the
// renderer uses for vr_set_gmm_params, so host copy and device scene cannot drift apart.
#pragma once
#include <algorithm>
#include <cmath>
#include <stdexcept>
#include <vector>

#include "gaussian.h"
#include "smm.h"
#include "vr_session.h"

static inline float sigmoidf_safe(float x) {
    if (x >= 0.0f) return 1.0f / (1.0f + std::exp(-x));
    const float z = std::exp(x);
    return z / (1.0f + z);
}
static inline float inv_sigmoidf(float y) {
    const float yy = std::clamp(y, 1e-7f, 1.0f - 1e-7f);
    return std::log(yy / (1.0f - yy));
}

class GaussianMixtureModel {
public:
    std::vector<Gaussian> gaussians;

    GaussianMixtureModel() = default;
    explicit GaussianMixtureModel(const std::vector<Gaussian>& gs) : gaussians(gs) {}

    size_t get_num_gaussians() const { return gaussians.size(); }
    void BuildBVH() {}  // the culling hierarchy is built on vr_set_gmm; kept for source compatibility

    // flat arrays in the argument order of vr_set_gmm (cov6 = cxx cxy cxz cyy cyz czz)
    void flatten(std::vector<float>& mean3, std::vector<float>& cov6, std::vector<float>& density, std::vector<float>& albedo) const {
        const size_t n = gaussians.size();
        mean3.resize(3 * n); cov6.resize(6 * n); density.resize(n); albedo.resize(n);
        for (size_t i = 0; i < n; ++i) {
            const Gaussian& g = gaussians[i];
            const Eigen::Vector3f m = g.centroid();
            const Eigen::Matrix3f& c = g.get_covariance();
            for (int k = 0; k < 3; ++k) mean3[3 * i + k] = m[k];
            float* q = &cov6[6 * i];
            q[0] = c(0, 0); q[1] = c(0, 1); q[2] = c(0, 2); q[3] = c(1, 1); q[4] = c(1, 2); q[5] = c(2, 2);
            density[i] = g.get_density(); albedo[i] = g.get_albedo();
        }
    }

    void pack_parameters(std::vector<float>& out) const {
        out.assign(gaussians.size() * 11, 0.0f);
        for (size_t i = 0; i < gaussians.size(); ++i) {
            const Gaussian& g = gaussians[i];
            float* p = &out[11 * i];
            const Eigen::Vector3f m = g.centroid();
            Eigen::AngleAxisf aa(g.get_rotation());
            Eigen::Vector3f rod = aa.axis() * aa.angle();
            if (!std::isfinite(rod.x()) || !std::isfinite(rod.y()) || !std::isfinite(rod.z())) rod = Eigen::Vector3f::Zero();
            const Eigen::Vector3f s = g.get_scale().diagonal();
            for (int k = 0; k < 3; ++k) { p[k] = m[k]; p[3 + k] = rod[k]; p[6 + k] = std::log(std::max(s[k], 1e-12f)); }
            p[9] = std::log(std::max(g.get_density(), 1e-12f));
            p[10] = inv_sigmoidf(std::clamp(g.get_albedo(), 0.0f, 1.0f));
        }
    }
};

// gmm.h:634-674.  Throws std::runtime_error on a size mismatch like the reference.
static inline void apply_params_to_gmm_local(const std::vector<float>& params, GaussianMixtureModel& gmm) {
    const size_t n = gmm.gaussians.size();
    if (params.size() != n * 11) throw std::runtime_error("apply params size mismatch");
    if (n == 0) return;
    std::vector<float> mean(3 * n), cov6(6 * n), den(n), alb(n);
    vrhost::Session& S = vrhost::Session::get();
    S.check(0, vr_unpack_params(S.ctx(0), int(n), params.data(), mean.data(), cov6.data(), den.data(), alb.data()));
    for (size_t i = 0; i < n; ++i) {
        const float* c = &cov6[6 * i];
        Eigen::Matrix3f cov;
        cov << c[0], c[1], c[2], c[1], c[3], c[4], c[2], c[4], c[5];
        gmm.gaussians[i] = Gaussian(Eigen::Vector3f(mean[3 * i], mean[3 * i + 1], mean[3 * i + 2]), cov, den[i], alb[i]);
    }
}

// finite-difference step per parameter slot (gmm.h:677-706)
static inline std::vector<float> make_default_eps_for_params(const std::vector<float>& base_params) {
    static const float kEps[11] = {0.02f, 0.02f, 0.02f, 0.10f, 0.10f, 0.10f, 0.05f, 0.05f, 0.05f, 0.25f, 0.5f};
    std::vector<float> eps(base_params.size());
    for (size_t i = 0; i < eps.size(); ++i) eps[i] = kEps[i % 11];
    return eps;
}
