// integrator.h -- Integrator hierarchy of the drop-in API (reference integrator.h:49-57 and its
// subclasses).  Constructors, defaults and render() signatures are the reference's; every body is
// "flatten Scene + Camera, call the C-ABI" (include/volrender.h).  With VR_DEVICES="0,1,.." the image
// rows are interleaved over the devices (one vr_context and one host thread per device, no data-path
// collective); each context writes only the pixels it owns straight into Image::data().
//
// The reference picks its distance solver with a #define (distance_solvers.h:143-147); here it is a
// runtime property: set_solver(), default ANALYTIC_PLUS_NEWTON (the shipped mode), and the same macros
// (-DUNIFORM, -DBISECTION, -DNEWTON, -DANALYTIC_PLUS_BISECTION) select the default when user code defines them.
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <thread>
#include <vector>

#include "camera.h"
#include "image.h"
#include "scene.h"
#include "vr_session.h"

namespace vrhost {

inline int default_solver() {
#if defined(UNIFORM)
    return VR_SOLVER_UNIFORM;
#elif defined(BISECTION)
    return VR_SOLVER_BISECTION;
#elif defined(NEWTON)
    return VR_SOLVER_NEWTON;
#elif defined(ANALYTIC_PLUS_BISECTION)
    return VR_SOLVER_ANALYTIC_PLUS_BISECTION;
#else
    return VR_SOLVER_ANALYTIC_PLUS_NEWTON;
#endif
}

struct RenderTotals {  // summed over devices; ms = slowest device
    double kernel_ms = 0.0, total_ms = 0.0;
    uint64_t samples = 0, scatter_rays = 0, shadow_rays = 0, hits = 0;
    uint32_t kernel_launches = 0;
};

inline void upload_lights_and_camera(Session& S, int i, const Scene& scene, const Camera& camera) {
    std::vector<float> lp(3 * scene.lights.size() + 3), li(3 * scene.lights.size() + 3);
    for (size_t k = 0; k < scene.lights.size(); ++k)
        for (int c = 0; c < 3; ++c) { lp[3 * k + c] = scene.lights[k].position[c]; li[3 * k + c] = scene.lights[k].intensity[c]; }
    const float env[3] = {scene.env_color[0], scene.env_color[1], scene.env_color[2]};
    S.check(i, vr_set_lights(S.ctx(i), int(scene.lights.size()), lp.data(), li.data(), env));
    const CameraPod c = camera.pod();
    S.check(i, vr_set_camera(S.ctx(i), c.kind, c.pos, c.view, c.right, c.up, c.pinhole));
}

// runs fn(device slot) on one host thread per device and rethrows the first failure
template <class F>
inline void for_each_device(Session& S, F fn) {
    const int nd = S.num_devices();
    if (nd == 1) { fn(0); return; }
    std::vector<std::thread> th;
    std::vector<std::exception_ptr> err((size_t)nd);
    for (int i = 0; i < nd; ++i)
        th.emplace_back([&, i] { try { fn(i); } catch (...) { err[(size_t)i] = std::current_exception(); } });
    for (auto& t : th) t.join();
    for (auto& e : err) if (e) std::rethrow_exception(e);
}

// One Integrator::render over Gaussians: kind = VR_INTEGRATOR_FREE_FLIGHT_SS / VR_INTEGRATOR_MULTI_SCATTER.
inline RenderTotals render_gaussians(int kind, int solver, const Scene& scene, const Camera& camera, Image& image, int spp, int min_bounces,
                                     std::vector<std::vector<uint32_t>>* per_pixel_gaussians, float march_step = 0.01f, int env_samples = 5) {
    if (scene.volume_type != Scene::VolumeType::GAUSSIANS) throw std::runtime_error("Trying to intersect with uninitialized medium");
    static const GaussianMixtureModel empty_model;
    const GaussianMixtureModel& model = scene.gmm && !scene.gmm->empty() ? scene.gmm->at(0) : empty_model;
    std::vector<float> mean, cov6, den, alb;
    model.flatten(mean, cov6, den, alb);
    const int n = int(model.get_num_gaussians());
    const int W = int(image.get_width()), H = int(image.get_height());
    Session& S = Session::get();
    const int nd = S.num_devices();
    std::vector<vr_render_stats> stats((size_t)nd);
    for_each_device(S, [&](int i) {
        upload_lights_and_camera(S, i, scene, camera);
        S.check(i, vr_set_gmm(S.ctx(i), n, mean.data(), cov6.data(), den.data(), alb.data()));
        vr_render_params p{};
        p.integrator = kind; p.solver = solver; p.width = W; p.height = H; p.spp = spp; p.min_bounces = min_bounces;
        p.x0 = 0; p.y0 = 0; p.x1 = W; p.y1 = H; p.part_rank = i; p.part_count = nd;
        p.record = per_pixel_gaussians ? 1 : 0;
        p.march_step = march_step; p.env_samples = env_samples;
        S.check(i, vr_render(S.ctx(i), &p, image.data(), &stats[(size_t)i]));
    });
    if (per_pixel_gaussians) {  // integrator.h:540-545,709-714: one de-duplicated index list per pixel
        per_pixel_gaussians->assign(size_t(W) * size_t(H), std::vector<uint32_t>());
        for (int i = 0; i < nd; ++i) {
            std::vector<int> counts(size_t(W) * H, 0);
            S.check(i, vr_get_recorded(S.ctx(i), counts.data(), nullptr, 0));
            int cap = 1;
            for (int c : counts) cap = std::max(cap, c);
            std::vector<uint32_t> idx(size_t(W) * H * cap);
            S.check(i, vr_get_recorded(S.ctx(i), counts.data(), idx.data(), cap));
            for (size_t px = 0; px < counts.size(); ++px)
                if (counts[px] > 0) {
                    auto& dst = (*per_pixel_gaussians)[px];
                    dst.insert(dst.end(), idx.begin() + px * cap, idx.begin() + px * cap + counts[px]);
                }
        }
    }
    RenderTotals t;
    for (const vr_render_stats& s : stats) {
        t.kernel_ms = std::max(t.kernel_ms, s.kernel_ms); t.total_ms = std::max(t.total_ms, s.total_ms);
        t.samples += s.samples; t.scatter_rays += s.scatter_rays; t.shadow_rays += s.shadow_rays; t.hits += s.hits;
        t.kernel_launches += s.kernel_launches;
    }
    return t;
}

}  // namespace vrhost

// ---- abstract base (integrator.h:49-57) -------------------------------------------------------
class Integrator {
protected:
    const std::shared_ptr<Camera> camera;

public:
    Integrator(const std::shared_ptr<Camera>& cam) : camera(cam) {}
    virtual ~Integrator() = default;
    virtual void render(const Scene& scene, Image& image) = 0;
};

// ---- FreeFlightGaussians: single scattering (integrator.h:273-409) ------------------------------
class FreeFlightGaussians : public Integrator {
    int num_samples;
    int solver = vrhost::default_solver();
    vrhost::RenderTotals last{};

public:
    FreeFlightGaussians(const std::shared_ptr<Camera>& cam, int num_samples_in = 256) : Integrator(cam), num_samples(num_samples_in) {}
    void render(const Scene& scene, Image& image) override {
        last = vrhost::render_gaussians(VR_INTEGRATOR_FREE_FLIGHT_SS, solver, scene, *camera, image, num_samples, 0, nullptr);
    }
    void set_num_samples(int n) { num_samples = n; }
    void set_solver(int vr_solver_value) { solver = vr_solver_value; }
    const vrhost::RenderTotals& last_stats() const { return last; }
};

// ---- MultiScatterGaussians (integrator.h:414-720) ----------------------------------------------
class MultiScatterGaussians : public Integrator {
    int num_samples;
    int min_scatter;
    int solver = vrhost::default_solver();
    vrhost::RenderTotals last{};

public:
    MultiScatterGaussians(const std::shared_ptr<Camera>& cam, int samples = 16, int min_bounces = 5)
        : Integrator(cam), num_samples(samples), min_scatter(min_bounces) {}
    void render(const Scene& scene, Image& image) override { render(scene, image, nullptr); }
    // recording overload (integrator.h:532-536; always available here, no RECORD_PIXEL_GAUSSIANS switch)
    void render(const Scene& scene, Image& image, std::vector<std::vector<uint32_t>>* per_pixel_gaussians) {
        last = vrhost::render_gaussians(VR_INTEGRATOR_MULTI_SCATTER, solver, scene, *camera, image, num_samples, min_scatter, per_pixel_gaussians);
    }
    void set_num_samples(int n) { num_samples = n; }
    int get_num_samples() const { return num_samples; }
    int get_min_bounces() const { return min_scatter; }
    void set_solver(int vr_solver_value) { solver = vr_solver_value; }
    int get_solver() const { return solver; }
    const std::shared_ptr<Camera>& get_camera() const { return camera; }
    const vrhost::RenderTotals& last_stats() const { return last; }
};
#include "march_integrators.h"
