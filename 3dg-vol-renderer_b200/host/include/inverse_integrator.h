// inverse_integrator.h -- InverseIntegrator / StochasticFiniteDiffInverseIntegrator of the drop-in
// API (reference inverse_integrator.h:34-247).
//
// optimize() keeps the reference's contract: Scene by value, false when there is nothing to optimise
// or a step fails, progress as ./sfd_output/iter_%04d.ppm + stdout lines every save_every iterations,
// a final high-spp render.  One iteration is the reference's estimator (:114-207): a recorded base
// render, S sign vectors s in {-1,+1}^D, S recorded renders of params + s*eps, per-Gaussian loss
// differences over the union of the pixels either render touched, grad = mean_s F_diff*s/eps, Adam.
// All of it runs on the device behind vr_sfd_begin / vr_sfd_accumulate / vr_sfd_apply; with several
// devices (VR_DEVICES) the image rows are split, the per-Gaussian partial sums (exact int64 fixed point)
// are summed by ONE ncclAllReduce per iteration on the devices' streams (Session::allreduce_sfd,
// csrc/vr_comm.cu; the reduction inverse_integrator.h:157-198 sums over pixels, hence over devices),
// and every device applies the same Adam step.  The SFD renders use the FORWARD integrator's camera,
// as the reference does (:119,153 go through forward_integrator->render); lights and camera are
// uploaded again at the top of every iteration, so nothing depends on what a save render left behind.
//
// Differences, all additive: the sign vectors come from a SEEDED mt19937 (SFDDConfig::seed; the
// reference seeds from std::random_device) so runs are reproducible; the optimised
// scene, which the reference discards, is available from optimized_scene() afterwards; output
// directory and the final render's spp are configurable (defaults = the reference's).
#pragma once
#include <chrono>
#include <cstdio>
#include <filesystem>
#include <iomanip>
#include <iostream>
#include <random>
#include <sstream>

#include "integrator.h"
#include "optimizer.h"

static inline void ensure_dir_exists(const std::string& path) {
    std::filesystem::path p(path);
    if (!std::filesystem::exists(p)) std::filesystem::create_directories(p);
}
static inline std::string out_filename(const std::string& dir, int it) {
    char name[32];
    std::snprintf(name, sizeof name, "/iter_%04d.ppm", it);
    return dir + name;
}
// per-pixel L1 loss (inverse_integrator.h:20-29)
static inline void compute_pixel_losses(const Image& I, const Image& Iref, std::vector<float>& out) {
    const size_t n = size_t(I.get_width()) * I.get_height();
    out.assign(n, 0.0f);
    const float *a = I.data(), *b = Iref.data();
    for (size_t p = 0; p < n; ++p) out[p] = std::abs(a[3 * p] - b[3 * p]) + std::abs(a[3 * p + 1] - b[3 * p + 1]) + std::abs(a[3 * p + 2] - b[3 * p + 2]);
}

class InverseIntegrator {
protected:
    const std::shared_ptr<Camera> camera;

public:
    InverseIntegrator(const std::shared_ptr<Camera>& cam) : camera(cam) {}
    virtual ~InverseIntegrator() = default;
    virtual bool optimize(Scene scene_initial, const Image& I_ref) = 0;
};

struct SFDDConfig {
    int max_iters = 1000;
    int save_every = 25;
    int num_stoch_samples = 4;
    float lr = 1e-2f;
    // ---- additions
    uint32_t seed = 0x5fdd2025u;           // sign-vector stream
    std::string output_dir = "./sfd_output";
    int final_samples = 16384;             // inverse_integrator.h:231
    bool verbose = true;
    // checkpoint / resume: every checkpoint_every iterations the optimiser state (iteration,
    // parameters, Adam moments, sign-vector generator) is written to checkpoint_path; a non-empty resume_from continues
    // a run from such a file and reproduces the uninterrupted run bit for bit
    std::string checkpoint_path;
    int checkpoint_every = 0;
    std::string resume_from;
};

namespace vrhost {
struct SfdCheckpoint {
    int iteration = 0, adam_t = 0;
    std::vector<float> params, m, v;
    std::string rng_state;   // std::mt19937 text form

    bool save(const std::string& path) const {
        std::ofstream f(path, std::ios::binary);
        if (!f) return false;
        const uint64_t n = params.size(), r = rng_state.size();
        f.write("VRSFD001", 8);
        f.write((const char*)&iteration, 4); f.write((const char*)&adam_t, 4);
        f.write((const char*)&n, 8); f.write((const char*)&r, 8);
        f.write((const char*)params.data(), std::streamsize(n * 4)); f.write((const char*)m.data(), std::streamsize(n * 4));
        f.write((const char*)v.data(), std::streamsize(n * 4)); f.write(rng_state.data(), std::streamsize(r));
        return bool(f);
    }
    bool load(const std::string& path) {
        std::ifstream f(path, std::ios::binary);
        char magic[8];
        uint64_t n = 0, r = 0;
        if (!f || !f.read(magic, 8) || std::string(magic, 8) != "VRSFD001") return false;
        f.read((char*)&iteration, 4); f.read((char*)&adam_t, 4); f.read((char*)&n, 8); f.read((char*)&r, 8);
        if (!f || n > (1ull << 32) || r > (1ull << 20)) return false;
        params.resize(n); m.resize(n); v.resize(n); rng_state.resize(r);
        f.read((char*)params.data(), std::streamsize(n * 4)); f.read((char*)m.data(), std::streamsize(n * 4));
        f.read((char*)v.data(), std::streamsize(n * 4)); f.read(&rng_state[0], std::streamsize(r));
        return bool(f);
    }
};
}  // namespace vrhost

class StochasticFiniteDiffInverseIntegrator : public InverseIntegrator {
    std::shared_ptr<MultiScatterGaussians> forward_integrator;
    SFDDConfig cfg;
    Scene result;
    std::vector<float> final_params;
    double last_mean_loss = 0.0;

    static double mean_loss(const Image& a, const Image& ref) {
        std::vector<float> l;
        compute_pixel_losses(a, ref, l);
        double s = 0.0;
        for (float x : l) s += x;
        return l.empty() ? 0.0 : s / double(l.size());
    }

public:
    StochasticFiniteDiffInverseIntegrator(const std::shared_ptr<Camera>& cam, const std::shared_ptr<MultiScatterGaussians>& fwd,
                                          const SFDDConfig& config = SFDDConfig())
        : InverseIntegrator(cam), forward_integrator(fwd), cfg(config) {}

    bool optimize(Scene scene_initial, const Image& I_ref) override {
        using clock = std::chrono::high_resolution_clock;
        ensure_dir_exists(cfg.output_dir);
        if (!scene_initial.gmm || scene_initial.gmm->empty()) {
            std::cerr << "Scene has no GMM." << std::endl;
            return false;
        }
        result = scene_initial;
        GaussianMixtureModel& model = result.gmm->at(0);
        std::vector<float> params;
        model.pack_parameters(params);
        if (params.empty()) {
            std::cerr << "No params to optimize." << std::endl;
            return false;
        }
        const std::vector<float> eps = make_default_eps_for_params(params);
        const int n = int(model.get_num_gaussians());
        const int D = int(params.size());
        const int W = int(I_ref.get_width()), H = int(I_ref.get_height());
        const int S = cfg.num_stoch_samples;

        vrhost::Session& dev = vrhost::Session::get();
        const int nd = dev.num_devices();
        const Camera& sfd_camera = *forward_integrator->get_camera();
        for (int i = 0; i < nd; ++i) {
            vrhost::upload_lights_and_camera(dev, i, result, sfd_camera);
            dev.check(i, vr_sfd_begin(dev.ctx(i), n, params.data(), eps.data(), cfg.lr, 0.9f, 0.999f, 1e-8f));
            dev.check(i, vr_sfd_set_target(dev.ctx(i), W, H, I_ref.data()));
        }
        std::mt19937 rng(cfg.seed);
        int first_iter = 0;
        if (!cfg.resume_from.empty()) {
            vrhost::SfdCheckpoint ck;
            if (!ck.load(cfg.resume_from) || int(ck.params.size()) != D) {
                std::cerr << "Cannot resume from " << cfg.resume_from << std::endl;
                return false;
            }
            for (int i = 0; i < nd; ++i) dev.check(i, vr_sfd_set_state(dev.ctx(i), ck.params.data(), ck.m.data(), ck.v.data(), ck.adam_t));
            std::istringstream(ck.rng_state) >> rng;
            first_iter = ck.iteration;
        }
        std::vector<int8_t> signs(size_t(S) * D);
        Image I_base(W, H);
        const auto t0 = clock::now();

        for (int iter = first_iter; iter < cfg.max_iters; ++iter) {
            for (int8_t& s : signs) s = (rng() & 1u) ? 1 : -1;
            vrhost::for_each_device(dev, [&](int i) {
                vrhost::upload_lights_and_camera(dev, i, result, sfd_camera);
                vr_render_params rp{};
                rp.integrator = VR_INTEGRATOR_MULTI_SCATTER; rp.solver = forward_integrator->get_solver();
                rp.width = W; rp.height = H; rp.spp = forward_integrator->get_num_samples(); rp.min_bounces = forward_integrator->get_min_bounces();
                rp.x1 = W; rp.y1 = H; rp.part_rank = i; rp.part_count = nd; rp.record = 1;
                dev.check(i, vr_sfd_reset_fdiff(dev.ctx(i), S));
                dev.check(i, vr_sfd_accumulate(dev.ctx(i), &rp, S, signs.data()));
            });
            dev.allreduce_sfd();   // nd > 1: one all-reduce of the S * n int64 partial sums, in place on the devices
            for (int i = 0; i < nd; ++i) dev.check(i, vr_sfd_apply(dev.ctx(i), S, signs.data(), nullptr, nullptr));
            if (cfg.checkpoint_every > 0 && !cfg.checkpoint_path.empty() && (iter + 1) % cfg.checkpoint_every == 0) {
                vrhost::SfdCheckpoint ck;
                ck.iteration = iter + 1;
                ck.params.resize(size_t(D)); ck.m.resize(size_t(D)); ck.v.resize(size_t(D));
                dev.check(0, vr_sfd_get_state(dev.ctx(0), ck.params.data(), ck.m.data(), ck.v.data(), &ck.adam_t));
                std::ostringstream os; os << rng; ck.rng_state = os.str();
                if (!ck.save(cfg.checkpoint_path)) std::cerr << "Cannot write checkpoint " << cfg.checkpoint_path << std::endl;
            }
            if (iter % cfg.save_every == 0) {
                dev.check(0, vr_sfd_get_params(dev.ctx(0), params.data()));
                apply_params_to_gmm_local(params, model);
                const std::string fn = out_filename(cfg.output_dir, iter);
                forward_integrator->render(result, I_base);
                I_base.make_PPM(fn);
                last_mean_loss = mean_loss(I_base, I_ref);
                if (cfg.verbose) {
                    const std::chrono::duration<double> el = clock::now() - t0;
                    std::cout << "[SFD] iter " << iter << " saving " << fn << std::endl;
                    std::cout << "[SFD] iter " << iter << " mean loss: " << last_mean_loss << std::endl;
                    std::cout << "[SFD] iter " << iter << " elapsed " << el.count() << "s" << std::endl;
                }
            }
        }
        dev.check(0, vr_sfd_get_params(dev.ctx(0), params.data()));
        apply_params_to_gmm_local(params, model);
        final_params = params;
        forward_integrator->set_num_samples(cfg.final_samples);
        forward_integrator->render(result, I_base);
        I_base.make_PPM(out_filename(cfg.output_dir, cfg.max_iters - 1));
        last_mean_loss = mean_loss(I_base, I_ref);
        if (cfg.verbose) {
            std::cout << "[SFD] Optimization finished." << std::endl;
            std::cout << "Final mean loss: " << last_mean_loss << std::endl;
        }
        return true;
    }

    // ---- additions: what the reference throws away
    const Scene& optimized_scene() const { return result; }
    const std::vector<float>& optimized_params() const { return final_params; }
    double final_mean_loss() const { return last_mean_loss; }
};
