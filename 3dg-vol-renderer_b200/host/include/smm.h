// smm.h -- homogeneous sphere mixture of the drop-in API (reference smm.h:7-52): containers only,
// the marcher (test_integrators.h:23-135) runs on the device (vr_set_smm + VR_INTEGRATOR_MARCH_SPHERES).
#pragma once
#include <cstddef>
#include <vector>

#include "ray.h"

struct PrimitiveHitEvent {
    float t;
    bool entering;
    size_t index;
    bool operator<(const PrimitiveHitEvent& o) const { return t < o.t; }
};

struct Sphere {
    Eigen::Vector3f center;
    float radius;
    float sigma_a;
    float sigma_s;
    Sphere(const Eigen::Vector3f& c, float r, float sa = 0.0f, float ss = 1.0f) : center(c), radius(r), sigma_a(sa), sigma_s(ss) {}
};

class SphereMixtureModel {
public:
    std::vector<Sphere> spheres;
    SphereMixtureModel() = default;
    explicit SphereMixtureModel(const std::vector<Sphere>& s) : spheres(s) {}
    size_t get_num_spheres() const { return spheres.size(); }
};
