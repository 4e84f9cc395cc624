// ray.h -- host-side Ray of the drop-in API (mirrors the reference's ray.h:5-16).
// The constructor normalises the direction, exactly like the reference's, so a camera ray is
// normalised twice (camera.h:52 + ray.h:12); the device code reproduces that double rounding.
#pragma once
#include <Eigen/Core>

struct Ray {
    Eigen::Vector3f origin;
    Eigen::Vector3f direction;
    Eigen::Vector3f throughput;

    Ray() {}
    Ray(const Eigen::Vector3f& o, const Eigen::Vector3f& d) : origin(o), direction(d.normalized()) {}
    Eigen::Vector3f operator()(float t) const { return origin + t * direction; }
};
