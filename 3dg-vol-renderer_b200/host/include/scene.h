// scene.h -- Scene aggregate and text loaders of the drop-in API (reference scene.h:12-120).
//
// Same public aggregate (volume_type, optional gmm/smm vectors of which only element [0] is used,
// lights, env_color) and the same token formats:
//     l x y z r g b                                         point light
//     s x y z radius sigma_a sigma_s                         sphere
//     g x y z cxx cxy cxz cyy cyz czz density albedo [er eg eb]   Gaussian
// The readers use the same stream operations as the reference (operator>> per token, peek() after
// albedo), so they inherit its behaviour on malformed input as well: unknown tokens are skipped,
// and a blank between albedo and the end of line makes the optional-emission read swallow the next
// tag and stop parsing (scene.h:99-105).
// New here: load_XML for the Mitsuba subset of tests/env_one_sphere_test_ortho.xml (scene_xml.h).
#pragma once
#include <fstream>
#include <iostream>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "gmm.h"
#include "smm.h"

struct Light {
    Eigen::Vector3f position;
    Eigen::Vector3f intensity;
};

struct Scene {
    enum class VolumeType { GAUSSIANS, SPHERES, VOXELS } volume_type = VolumeType::GAUSSIANS;

    std::optional<std::vector<GaussianMixtureModel>> gmm;
    std::optional<std::vector<SphereMixtureModel>> smm;
    std::vector<Light> lights;
    Eigen::Vector3f env_color = {0.53f, 0.81f, 0.92f};

private:
    static std::ifstream open_or_throw(const std::string& filename) {
        std::ifstream f(filename);
        if (!f) throw std::runtime_error("Failed to open scene file: " + filename);
        return f;
    }
    static bool read_light(std::istream& in, std::vector<Light>& out) {
        float v[6];
        for (float& x : v) in >> x;
        out.push_back({Eigen::Vector3f(v[0], v[1], v[2]), Eigen::Vector3f(v[3], v[4], v[5])});
        return bool(in);
    }

public:
    static Scene load_SMM(const std::string& filename) {
        Scene s;
        s.volume_type = VolumeType::SPHERES;
        std::ifstream in = open_or_throw(filename);
        std::vector<Sphere> spheres;
        for (std::string tag; in >> tag;) {
            if (tag == "l") read_light(in, s.lights);
            else if (tag == "s") {
                float v[6];
                for (float& x : v) in >> x;
                spheres.emplace_back(Eigen::Vector3f(v[0], v[1], v[2]), v[3], v[4], v[5]);
            }
        }
        s.smm = std::vector<SphereMixtureModel>{SphereMixtureModel(spheres)};
        return s;
    }

    static Scene load_GMM(const std::string& filename) {
        Scene s;
        s.volume_type = VolumeType::GAUSSIANS;
        std::ifstream in = open_or_throw(filename);
        std::vector<Gaussian> gaussians;
        for (std::string tag; in >> tag;) {
            if (tag == "l") read_light(in, s.lights);
            else if (tag == "g") {
                float v[11];
                for (float& x : v) in >> x;
                Eigen::Vector3f emission(0.0f, 0.0f, 0.0f);
                const int next = in.peek();
                if (next != '\n' && next != EOF) {
                    float e[3];
                    if (in >> e[0] >> e[1] >> e[2]) emission = Eigen::Vector3f(e[0], e[1], e[2]);
                }
                Eigen::Matrix3f cov;
                cov << v[3], v[4], v[5], v[4], v[6], v[7], v[5], v[7], v[8];
                gaussians.emplace_back(Eigen::Vector3f(v[0], v[1], v[2]), cov, v[9], v[10], emission);
            }
        }
        s.gmm = std::vector<GaussianMixtureModel>{GaussianMixtureModel(gaussians)};
        return s;
    }

    // Mitsuba-subset XML (scene_xml.h); returns the scene, and through the optional out-parameters the
    // sensor the file describes
    struct XmlSensor {
        bool orthographic = true;
        Eigen::Vector3f origin{0.f, 1.f, 6.f}, target{0.f, 1.f, 0.f};
        float fov_radians = 0.78539816339f;
        int width = 512, height = 512, sample_count = 1;
    };
    static Scene load_XML(const std::string& filename, XmlSensor* sensor = nullptr);

    size_t get_num_primitives() const {
        switch (volume_type) {
            case VolumeType::SPHERES: return smm && !smm->empty() ? (*smm)[0].get_num_spheres() : 0;
            case VolumeType::GAUSSIANS: return gmm && !gmm->empty() ? (*gmm)[0].get_num_gaussians() : 0;
            default: throw std::runtime_error("Trying to get num_primitives with uninitialized medium");
        }
    }
};

#include "scene_xml.h"
