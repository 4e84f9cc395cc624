// march_integrators.h -- the ray-marching validation integrators of the reference
// (RayMarchingSpheres test_integrators.h:10-136, RayMarchingGaussians test_integrators.h:143-297,
// PureRayMarching integrator.h:100-267, TestIntegrator integrator.h:65-96).  Included by integrator.h.
//
// RayMarchingSpheres runs on the device (vr_set_smm + VR_INTEGRATOR_MARCH_SPHERES).  The reference
// seeds its environment directions from std::random_device; the device marcher uses
// the PCG stream seeded per (x, y, step, sample) instead, so images agree statistically, not bit-wise.
// The other three run on the device for Gaussian scenes (vr_march.cu, same substitution for the environment directions).
#pragma once

class RayMarchingSpheres : public Integrator {
    float step_size;
    int env_samples;

public:
    RayMarchingSpheres(const std::shared_ptr<Camera>& cam, float step_size_in = 0.01f, int env_samples_in = 5)
        : Integrator(cam), step_size(step_size_in), env_samples(env_samples_in) {}

    void render(const Scene& scene, Image& image) override {
        if (scene.volume_type != Scene::VolumeType::SPHERES || !scene.smm || scene.smm->empty())
            throw std::runtime_error("Sphere mixture model not initialized");
        const SphereMixtureModel& model = scene.smm->at(0);
        const size_t n = model.get_num_spheres();
        std::vector<float> c(3 * n + 3), r(n + 1), sa(n + 1), ss(n + 1);
        for (size_t i = 0; i < n; ++i) {
            const Sphere& s = model.spheres[i];
            for (int k = 0; k < 3; ++k) c[3 * i + k] = s.center[k];
            r[i] = s.radius; sa[i] = s.sigma_a; ss[i] = s.sigma_s;
        }
        const int W = int(image.get_width()), H = int(image.get_height());
        vrhost::Session& S = vrhost::Session::get();
        const int nd = S.num_devices();
        vrhost::for_each_device(S, [&](int i) {
            vrhost::upload_lights_and_camera(S, i, scene, *camera);
            S.check(i, vr_set_smm(S.ctx(i), int(n), c.data(), r.data(), sa.data(), ss.data()));
            vr_render_params p{};
            p.integrator = VR_INTEGRATOR_MARCH_SPHERES; p.width = W; p.height = H; p.spp = 1; p.min_bounces = 0;
            p.x1 = W; p.y1 = H; p.part_rank = i; p.part_count = nd;
            p.march_step = step_size; p.env_samples = env_samples;
            vr_render_stats st{};
            S.check(i, vr_render(S.ctx(i), &p, image.data(), &st));
        });
    }
};

// The three validation integrators over a Gaussian scene run on the device (vr_march.cu): one ray per pixel through
// the pixel centre, fixed-step march.  Sphere scenes are served by RayMarchingSpheres only.
namespace vrhost {
inline void render_validation(int kind, const char* what, const Scene& scene, const Camera& camera, Image& image, float step, int env_samples) {
    if (scene.volume_type != Scene::VolumeType::GAUSSIANS)
        throw std::runtime_error(std::string(what) + ": only Gaussian scenes run on the device (use RayMarchingSpheres for spheres); there is no CPU fallback");
    render_gaussians(kind, VR_SOLVER_ANALYTIC_PLUS_NEWTON, scene, camera, image, 1, 0, nullptr, step, env_samples);
}
}  // namespace vrhost

class RayMarchingGaussians : public Integrator {
    float step_size;
    int env_samples;
public:
    RayMarchingGaussians(const std::shared_ptr<Camera>& cam, float step_size_in = 0.01f, int env_samples_in = 20)
        : Integrator(cam), step_size(step_size_in), env_samples(env_samples_in) {}
    void render(const Scene& scene, Image& image) override {
        vrhost::render_validation(VR_INTEGRATOR_MARCH_GAUSSIANS, "RayMarchingGaussians", scene, *camera, image, step_size, env_samples);
    }
};

class PureRayMarching : public Integrator {
    float step_size;
    int env_samples;
public:
    PureRayMarching(const std::shared_ptr<Camera>& cam, float step_size_in = 0.01f, int env_samples_in = 20)
        : Integrator(cam), step_size(step_size_in), env_samples(env_samples_in) {}
    void render(const Scene& scene, Image& image) override {
        vrhost::render_validation(VR_INTEGRATOR_PURE_MARCH, "PureRayMarching", scene, *camera, image, step_size, env_samples);
    }
};

class TestIntegrator : public Integrator {
public:
    TestIntegrator(const std::shared_ptr<Camera>& cam) : Integrator(cam) {}
    void render(const Scene& scene, Image& image) override {
        vrhost::render_validation(VR_INTEGRATOR_TEST, "TestIntegrator", scene, *camera, image, 0.01f, 1);
    }
};
