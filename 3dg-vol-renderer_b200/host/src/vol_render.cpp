// vol_render -- command-line driver over the drop-in C++ API.
//
// Replaces the reference's tests/main.cpp:17-117, whose scene, camera, integrator, solver and sample
// count are compile-time edits, by runtime flags; output is the same 8-bit PPM (Image::make_PPM)
// and, optionally, the raw float image and the recorded per-pixel Gaussian lists.
//
//   vol_render --scene S.txt [--type gmm|smm|xml] [--integrator ms|ss|spheres|march|pure|test] [--solver 0..4]
//              [--spp 256] [--min-bounces 5] [--width 512 --height 512]
//              [--camera pinhole|ortho] [--pos 0,1,6] [--look 0,1,0] [--fov 0.785398]
//              [--step 0.01 --env-samples 5] [--out output.ppm] [--raw image.f32] [--record lists.u32]
//              [--repeat 1]
//   vol_render --scene S.txt --inverse target.ppm [--iters 1000 --save-every 25 --sfd-samples 4 --lr 0.01
//              --seed N --final-spp 16384 --out-dir ./sfd_output --save-scene fitted.txt
//              --checkpoint state.bin --checkpoint-every 50 --resume state.bin --save-params params.f32
//              --inverse-camera-pos x,y,z]
// Devices: VR_DEVICES="0,1,..." (default "0").
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>
#include <memory>
#include <string>

#include "integrator.h"
#include "inverse_integrator.h"

namespace {

struct Args {
    std::map<std::string, std::string> kv;
    bool has(const std::string& k) const { return kv.count(k) != 0; }
    std::string str(const std::string& k, const std::string& d) const { auto it = kv.find(k); return it == kv.end() ? d : it->second; }
    int i(const std::string& k, int d) const { return has(k) ? std::atoi(kv.at(k).c_str()) : d; }
    float f(const std::string& k, float d) const { return has(k) ? float(std::atof(kv.at(k).c_str())) : d; }
    Eigen::Vector3f v3(const std::string& k, Eigen::Vector3f d) const {
        if (!has(k)) return d;
        float v[3] = {0, 0, 0};
        vrhost::xml::floats3(kv.at(k), v);
        return Eigen::Vector3f(v[0], v[1], v[2]);
    }
};

bool ends_with(const std::string& s, const char* suf) {
    const size_t n = std::strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

void write_scene_txt(const std::string& path, const Scene& s) {
    FILE* f = std::fopen(path.c_str(), "w");
    if (!f) throw std::runtime_error("cannot write " + path);
    for (const Light& l : s.lights)
        std::fprintf(f, "l %.9g %.9g %.9g %.9g %.9g %.9g\n", l.position[0], l.position[1], l.position[2], l.intensity[0], l.intensity[1], l.intensity[2]);
    if (s.gmm && !s.gmm->empty())
        for (const Gaussian& g : s.gmm->at(0).gaussians) {
            const Eigen::Vector3f m = g.centroid();
            const Eigen::Matrix3f& c = g.get_covariance();
            std::fprintf(f, "g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g\n", m[0], m[1], m[2], c(0, 0), c(0, 1), c(0, 2), c(1, 1), c(1, 2),
                         c(2, 2), g.get_density(), g.get_albedo());
        }
    std::fclose(f);
}

int usage() {
    std::fprintf(stderr, "usage: vol_render --scene FILE [--type gmm|smm|xml] [--integrator ms|ss|spheres] [--spp N] [--width W --height H] ...\n"
                         "       (see the header of host/src/vol_render.cpp for every flag)\n");
    return 2;
}

}  // namespace

int main(int argc, char** argv) {
    Args a;
    for (int k = 1; k < argc; ++k) {
        std::string key = argv[k];
        if (key.rfind("--", 0) != 0) return usage();
        key.erase(0, 2);
        if (k + 1 < argc && std::strncmp(argv[k + 1], "--", 2) != 0) a.kv[key] = argv[++k];
        else a.kv[key] = "1";
    }
    if (!a.has("scene")) return usage();
    try {
        const std::string path = a.str("scene", "");
        std::string type = a.str("type", ends_with(path, ".xml") ? "xml" : "gmm");
        Scene::XmlSensor sensor;
        Scene scene = type == "xml" ? Scene::load_XML(path, &sensor) : type == "smm" ? Scene::load_SMM(path) : Scene::load_GMM(path);
        if (a.has("env")) scene.env_color = a.v3("env", scene.env_color);
        const bool spheres = scene.volume_type == Scene::VolumeType::SPHERES;

        int W = a.i("width", type == "xml" ? sensor.width : 512), H = a.i("height", type == "xml" ? sensor.height : 512);
        const Eigen::Vector3f pos = a.v3("pos", type == "xml" ? sensor.origin : Eigen::Vector3f(0.f, 1.f, 6.f));
        const Eigen::Vector3f look = a.v3("look", type == "xml" ? sensor.target : Eigen::Vector3f(0.f, 1.f, 0.f));
        const float fov = a.f("fov", type == "xml" ? sensor.fov_radians : 0.25f * 3.14159265358979323846f);
        const std::string cam_kind = a.str("camera", type == "xml" && sensor.orthographic ? "ortho" : "pinhole");
        const Eigen::Vector3f view = (look - pos).normalized();
        std::shared_ptr<Camera> camera;
        if (cam_kind == "ortho") camera = std::make_shared<Orthographic_Camera>(pos, view);
        else camera = std::make_shared<Pinhole_Camera>(pos, view, fov);

        const std::string integ = a.str("integrator", spheres ? "spheres" : "ms");
        const int spp = a.i("spp", 256);
        const int solver = a.i("solver", VR_SOLVER_ANALYTIC_PLUS_NEWTON);
        std::printf("scene %s: %zu primitives, %zu lights; %s camera; integrator %s; %dx%d; devices %d\n", path.c_str(), scene.get_num_primitives(),
                    scene.lights.size(), cam_kind.c_str(), integ.c_str(), W, H, vrhost::Session::get().num_devices());

        if (a.has("inverse")) {
            Image target(a.str("inverse", ""));
            auto fwd = std::make_shared<MultiScatterGaussians>(camera, spp, a.i("min-bounces", 5));
            fwd->set_solver(solver);
            SFDDConfig cfg;
            cfg.max_iters = a.i("iters", cfg.max_iters); cfg.save_every = a.i("save-every", cfg.save_every);
            cfg.num_stoch_samples = a.i("sfd-samples", cfg.num_stoch_samples); cfg.lr = a.f("lr", cfg.lr);
            cfg.seed = uint32_t(a.i("seed", int(cfg.seed))); cfg.final_samples = a.i("final-spp", cfg.final_samples);
            cfg.output_dir = a.str("out-dir", cfg.output_dir);
            cfg.checkpoint_path = a.str("checkpoint", ""); cfg.checkpoint_every = a.i("checkpoint-every", 0);
            cfg.resume_from = a.str("resume", ""); cfg.verbose = !a.has("quiet");
            // the reference's InverseIntegrator never uses its own camera member: every SFD render goes through the forward
            // integrator's (inverse_integrator.h:119,153).  --inverse-camera-pos builds it from another camera to show that.
            std::shared_ptr<Camera> inv_camera = camera;
            if (a.has("inverse-camera-pos")) {
                const Eigen::Vector3f ip = a.v3("inverse-camera-pos", pos);
                inv_camera = std::make_shared<Pinhole_Camera>(ip, (look - ip).normalized(), fov);
            }
            StochasticFiniteDiffInverseIntegrator inv(inv_camera, fwd, cfg);
            const auto t0 = std::chrono::high_resolution_clock::now();
            const bool ok = inv.optimize(scene, target);
            const std::chrono::duration<double> el = std::chrono::high_resolution_clock::now() - t0;
            std::printf("Inverse optimization time: %.6f seconds (%.3f iterations/s)\n", el.count(), cfg.max_iters / el.count());
            if (ok && a.has("save-scene")) write_scene_txt(a.str("save-scene", ""), inv.optimized_scene());
            if (ok && a.has("save-params")) {
                std::ofstream pf(a.str("save-params", ""), std::ios::binary);
                pf.write((const char*)inv.optimized_params().data(), std::streamsize(inv.optimized_params().size() * 4));
            }
            return ok ? 0 : 1;
        }

        Image image(W, H);
        std::vector<std::vector<uint32_t>> lists;
        const bool record = a.has("record");
        const int repeat = std::max(1, a.i("repeat", 1));
        double best = 1e300;
        for (int r = 0; r < repeat; ++r) {
            const auto t0 = std::chrono::high_resolution_clock::now();
            if (integ == "spheres") {
                RayMarchingSpheres rm(camera, a.f("step", 0.01f), a.i("env-samples", 5));
                rm.render(scene, image);
            } else if (integ == "ss") {
                FreeFlightGaussians ff(camera, spp);
                ff.set_solver(solver);
                ff.render(scene, image);
            } else if (integ == "ms") {
                MultiScatterGaussians ms(camera, spp, a.i("min-bounces", 5));
                ms.set_solver(solver);
                if (record) ms.render(scene, image, &lists); else ms.render(scene, image);
            } else if (integ == "march") {
                RayMarchingGaussians rm(camera, a.f("step", 0.01f), a.i("env-samples", 20));
                rm.render(scene, image);
            } else if (integ == "pure") {
                PureRayMarching pm(camera, a.f("step", 0.01f), a.i("env-samples", 20));
                pm.render(scene, image);
            } else if (integ == "test") {
                TestIntegrator ti(camera);
                ti.render(scene, image);
            } else {
                throw std::runtime_error("unknown integrator '" + integ + "'");
            }
            const std::chrono::duration<double> el = std::chrono::high_resolution_clock::now() - t0;
            best = std::min(best, el.count());
            std::printf("Render time: %.6f seconds\n", el.count());
        }
        const double samples = double(W) * H * (integ == "ss" || integ == "ms" ? spp : 1);
        std::printf("best %.6f s, %.3f Msamples/s\n", best, samples / best * 1e-6);
        image.make_PPM(a.str("out", "output.ppm"));
        if (a.has("raw")) image.write_raw(a.str("raw", ""));
        if (record) {  // u32 stream: npix, then per pixel count followed by the indices (ascending)
            FILE* f = std::fopen(a.str("record", "").c_str(), "wb");
            if (!f) throw std::runtime_error("cannot write the record file");
            const uint32_t npix = uint32_t(lists.size());
            std::fwrite(&npix, 4, 1, f);
            for (const auto& l : lists) {
                const uint32_t c = uint32_t(l.size());
                std::fwrite(&c, 4, 1, f);
                if (c) std::fwrite(l.data(), 4, c, f);
            }
            std::fclose(f);
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "vol_render: %s\n", e.what());
        return 1;
    }
    return 0;
}
