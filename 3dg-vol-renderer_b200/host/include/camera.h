// camera.h -- Camera / Pinhole_Camera / Orthographic_Camera of the drop-in API.
//
// Same constructors and sample_ray contract as the reference (camera.h:7-73).  The device cannot
// call the virtual, and the reference keeps its fields protected, so every camera additionally
// exposes pod(): the constructed basis as plain floats, which is what vr_set_camera
// (include/volrender.h) takes.  Constructor arithmetic follows camera.h:15-22,38-43: `right` and
// `up` are built from the view direction AS PASSED (the parameter shadows the normalised member),
// and the pinhole sits at position + view_dir_as_passed / tan(fov / 2).
#pragma once
#include <cmath>

#include "ray.h"

struct CameraPod {
    int kind;  // 0 pinhole, 1 orthographic (vr_camera_kind)
    float pos[3], view[3], right[3], up[3], pinhole[3];
};

class Camera {
protected:
    Eigen::Vector3f position;
    Eigen::Vector3f view_dir;
    Eigen::Vector3f right;
    Eigen::Vector3f up;

public:
    Camera(const Eigen::Vector3f& position_in, const Eigen::Vector3f& view_in) : position(position_in), view_dir(view_in.normalized()) {
        right = view_in.cross(Eigen::Vector3f(0.0f, 1.0f, 0.0f)).normalized();
        up = right.cross(view_in).normalized();
    }
    virtual ~Camera() = default;

    virtual Ray sample_ray(const Eigen::Vector2d& uv) const = 0;
    virtual CameraPod pod() const = 0;

protected:
    CameraPod base_pod(int kind) const {
        CameraPod p{};
        p.kind = kind;
        for (int k = 0; k < 3; ++k) { p.pos[k] = position[k]; p.view[k] = view_dir[k]; p.right[k] = right[k]; p.up[k] = up[k]; p.pinhole[k] = position[k]; }
        return p;
    }
};

class Pinhole_Camera : public Camera {
    float fov;
    float focal_length;
    Eigen::Vector3f pinhole;

public:
    Pinhole_Camera(const Eigen::Vector3f& position_in, const Eigen::Vector3f& view_in, float fov_in) : Camera(position_in, view_in), fov(fov_in) {
        focal_length = 1.0f / std::tan(0.5f * fov);
        pinhole = position_in + focal_length * view_in;
    }
    // film plane spans [-1,1]^2 around the position; u is mirrored, v points up
    Ray sample_ray(const Eigen::Vector2d& uv) const override {
        const float su = 1.0f - float(uv.x()) * 2.0f, sv = float(uv.y()) * 2.0f - 1.0f;
        const Eigen::Vector3f o = position + su * right + sv * up;
        return Ray(o, (pinhole - o).normalized());
    }
    CameraPod pod() const override {
        CameraPod p = base_pod(0);
        for (int k = 0; k < 3; ++k) p.pinhole[k] = pinhole[k];
        return p;
    }
};

class Orthographic_Camera : public Camera {
public:
    Orthographic_Camera(const Eigen::Vector3f& position_in, const Eigen::Vector3f& forward) : Camera(position_in, forward) {}
    Ray sample_ray(const Eigen::Vector2d& uv) const override {
        const float su = float(uv.x()) * 2.0f - 1.0f, sv = 1.0f - float(uv.y()) * 2.0f;
        return Ray(position + su * right + sv * up, view_dir.normalized());
    }
    CameraPod pod() const override { return base_pod(1); }
};
