// gaussian.h -- host-side Gaussian component of the drop-in API.
//
// Keeps the reference's two constructors and public accessors (gaussian.h:85-123,302): a component
// is (mean, covariance, density, albedo, emission) or (mean, rotation R, scale S) with
// covariance = R S S^T R^T.  All ray arithmetic (intersect_direct / optical_depth /
// solve_for_t_given_tau, gaussian.h:126-297) lives on the device; the host object is only the
// container the Scene loader fills and vr_set_gmm flattens.  rotation/scale (needed by
// GaussianMixtureModel::pack_parameters) come from a symmetric eigen-decomposition like the
// reference's, with ONE deliberate difference: the eigenvector matrix is made a proper rotation
// (det = +1) so that pack -> unpack reproduces the covariance; the reference stores Eigen's raw
// eigenvectors, which lose the covariance in the axis-angle conversion whenever det = -1
// (SURVEY.md 7.3 "pack_parameters quirk").
#pragma once
#include <algorithm>
#include <cmath>

#include <Eigen/Dense>

class Gaussian {
    Eigen::Vector3f mean;
    Eigen::Matrix3f covariance;
    float density;
    float albedo;
    Eigen::Vector3f emission;
    Eigen::Matrix3f rotation;  // R
    Eigen::Matrix3f scale;     // S (diagonal)

    void factor_covariance() {
        Eigen::SelfAdjointEigenSolver<Eigen::Matrix3f> es(covariance);
        Eigen::Matrix3f V = es.eigenvectors();
        const Eigen::Vector3f lam = es.eigenvalues();
        if (V.determinant() < 0.0f) for (int r = 0; r < 3; ++r) V(r, 0) = -V(r, 0);
        rotation = V;
        scale = Eigen::Matrix3f::Zero();
        for (int k = 0; k < 3; ++k) scale(k, k) = std::sqrt(std::max(lam[k], 0.0f));
    }

public:
    static constexpr float R = 3.0f;  // support is truncated at R standard deviations (gaussian.h:36)

    Gaussian(const Eigen::Vector3f& mean_in, const Eigen::Matrix3f& covariance_in, float density_in, float albedo_in,
             const Eigen::Vector3f& emission_in = Eigen::Vector3f::Zero())
        : mean(mean_in), covariance(covariance_in), density(density_in), albedo(albedo_in), emission(emission_in) {
        factor_covariance();
    }
    Gaussian(const Eigen::Vector3f& mean_in, const Eigen::Matrix3f& rotation_in, const Eigen::Matrix3f& scale_in, float density_in,
             float albedo_in, const Eigen::Vector3f& emission_in = Eigen::Vector3f::Zero())
        : mean(mean_in), density(density_in), albedo(albedo_in), emission(emission_in), rotation(rotation_in), scale(scale_in) {
        covariance = rotation * scale * scale.transpose() * rotation.transpose();
    }

    Eigen::Vector3f centroid() const { return mean; }
    const Eigen::Matrix3f& get_covariance() const { return covariance; }
    float get_density() const { return density; }
    float get_albedo() const { return albedo; }
    const Eigen::Vector3f& get_emission() const { return emission; }
    const Eigen::Matrix3f& get_rotation() const { return rotation; }
    const Eigen::Matrix3f& get_scale() const { return scale; }

    // tight axis-aligned box of the R-sigma ellipsoid: half extent R * sqrt(Sigma_kk)
    void get_aabb(Eigen::Vector3f& lo, Eigen::Vector3f& hi) const {
        Eigen::Vector3f h(R * std::sqrt(std::max(covariance(0, 0), 0.0f)), R * std::sqrt(std::max(covariance(1, 1), 0.0f)),
                          R * std::sqrt(std::max(covariance(2, 2), 0.0f)));
        lo = mean - h; hi = mean + h;
    }
};
