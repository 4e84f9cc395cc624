// test_host -- checks of the host-side drop-in API that need no GPU: scene parsing, camera rays,
// image I/O, parameter packing, error behaviour.  Run by tests/test_host_api.py.
//   test_host <golden scene dir> <tmp dir>
// Prints one "ok <name>" line per check and exits non-zero on the first failure.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>

#include "integrator.h"
#include "inverse_integrator.h"
#include "test_integrators.h"

#define CHECK(cond)                                                                  \
    do {                                                                             \
        if (!(cond)) { std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); std::exit(1); } \
    } while (0)

static bool close_to(float a, float b, float rtol = 1e-5f, float atol = 1e-7f) { return std::abs(a - b) <= atol + rtol * std::abs(b); }

int main(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: test_host <scene dir> <tmp dir>\n"); return 2; }
    const std::string scenes = argv[1], tmp = argv[2];

    {   // ---- scene text formats (scene.h:36-120)
        std::ofstream f(tmp + "/mini.txt");
        f << "l 0 5 0.1 50 0 0\nl -3 3 0.3 0 30 0\n# comment tokens are skipped\n"
          << "g 0.1 1.2 -0.3  0.04 0.01 0.0 0.05 0.002 0.03  7.5 0.8\n"
          << "g -0.5 0.7 0.2  0.02 0.0 0.0 0.02 0.0 0.02  3.0 0.6 1 2 3\n"
          << "g 0.5 0.9 0.0  0.01 0.0 0.0 0.01 0.0 0.01  2.0 0.5\n";
        f.close();
        Scene s = Scene::load_GMM(tmp + "/mini.txt");
        CHECK(s.volume_type == Scene::VolumeType::GAUSSIANS);
        CHECK(s.lights.size() == 2 && s.get_num_primitives() == 3);
        CHECK(s.lights[1].position == Eigen::Vector3f(-3.f, 3.f, 0.3f) && s.lights[1].intensity == Eigen::Vector3f(0.f, 30.f, 0.f));
        const Gaussian& g0 = s.gmm->at(0).gaussians[0];
        CHECK(g0.centroid() == Eigen::Vector3f(0.1f, 1.2f, -0.3f));
        CHECK(g0.get_covariance()(0, 1) == 0.01f && g0.get_covariance()(1, 0) == 0.01f && g0.get_covariance()(1, 2) == 0.002f);
        CHECK(g0.get_density() == 7.5f && g0.get_albedo() == 0.8f);
        CHECK(s.gmm->at(0).gaussians[1].get_emission() == Eigen::Vector3f(1.f, 2.f, 3.f));
        CHECK(s.env_color == Eigen::Vector3f(0.53f, 0.81f, 0.92f));
        // the reference's trailing-blank quirk (scene.h:99-105): parsing stops after that line
        std::ofstream q(tmp + "/quirk.txt");
        q << "g 0 1 0  0.01 0 0 0.01 0 0.01  2.0 0.5 \ng 1 1 0  0.01 0 0 0.01 0 0.01  2.0 0.5\n";
        q.close();
        CHECK(Scene::load_GMM(tmp + "/quirk.txt").get_num_primitives() == 1);
        bool threw = false;
        try { Scene::load_GMM(tmp + "/does_not_exist.txt"); } catch (const std::runtime_error&) { threw = true; }
        CHECK(threw);
        std::puts("ok scene_gmm_text");
    }
    {   // ---- spheres + XML (tests/env_one_sphere_test_ortho.xml subset)
        Scene s = Scene::load_SMM(scenes + "/one_sphere.txt");
        CHECK(s.volume_type == Scene::VolumeType::SPHERES && s.get_num_primitives() == 1 && s.lights.size() == 1);
        const Sphere& sp = s.smm->at(0).spheres[0];
        std::ofstream x(tmp + "/one.xml");
        x << "<?xml version=\"1.0\"?>\n<scene version=\"3.0.0\">\n <!-- c -->\n <integrator type=\"volpath\"><integer name=\"max_depth\" value=\"2\"/></integrator>\n"
          << " <sensor type=\"orthographic\"><transform name=\"to_world\"><lookat origin=\"0,1,6\" target=\"0,1,0\" up=\"0,1,0\"/></transform>\n"
          << "  <sampler type=\"independent\"><integer name=\"sample_count\" value=\"1\"/></sampler>\n"
          << "  <film type=\"hdrfilm\"><integer name=\"width\" value=\"64\"/><integer name=\"height\" value=\"48\"/><rfilter type=\"box\"/></film></sensor>\n"
          << " <emitter type=\"constant\"><rgb name=\"radiance\" value=\"0.53, 0.81, 0.92\"/></emitter>\n"
          << " <emitter type=\"point\"><point name=\"position\" x=\"0\" y=\"4\" z=\"0\"/><rgb name=\"intensity\" value=\"35.0\"/></emitter>\n"
          << " <medium id=\"vol\" type=\"homogeneous\"><rgb name=\"albedo\" value=\"0.875\"/><rgb name=\"sigma_t\" value=\"0.8\"/><float name=\"scale\" value=\"1.0\"/>"
          << "<phase name=\"phase\" type=\"isotropic\"/></medium>\n"
          << " <shape type=\"sphere\"><point name=\"center\" x=\"0\" y=\"1\" z=\"0\"/><float name=\"radius\" value=\"1\"/><bsdf type=\"null\"/>"
          << "<ref name=\"interior\" id=\"vol\"/></shape>\n</scene>\n";
        x.close();
        Scene::XmlSensor sen;
        Scene xs = Scene::load_XML(tmp + "/one.xml", &sen);
        CHECK(xs.volume_type == Scene::VolumeType::SPHERES && xs.get_num_primitives() == 1);
        const Sphere& xp = xs.smm->at(0).spheres[0];
        CHECK(xp.center == sp.center && xp.radius == sp.radius);
        CHECK(close_to(xp.sigma_a, sp.sigma_a, 1e-5f) && close_to(xp.sigma_s, sp.sigma_s, 1e-5f));   // 0.1 / 0.7 (1_spheres.txt)
        CHECK(xs.lights.size() == 1 && xs.lights[0].position == Eigen::Vector3f(0.f, 4.f, 0.f) && xs.lights[0].intensity == Eigen::Vector3f(35.f, 35.f, 35.f));
        CHECK(xs.env_color == Eigen::Vector3f(0.53f, 0.81f, 0.92f));
        CHECK(sen.orthographic && sen.width == 64 && sen.height == 48 && sen.sample_count == 1);
        CHECK(sen.origin == Eigen::Vector3f(0.f, 1.f, 6.f) && sen.target == Eigen::Vector3f(0.f, 1.f, 0.f));
        std::puts("ok scene_spheres_xml");
    }
    {   // ---- cameras (camera.h:15-73): the hand-checkable default view
        const Eigen::Vector3f pos(0.f, 1.f, 6.f), view(0.f, 0.f, -1.f);
        Pinhole_Camera pc(pos, view, 0.25f * 3.14159265358979f);
        const CameraPod p = pc.pod();
        CHECK(p.kind == 0 && p.right[0] == 1.f && p.right[1] == 0.f && p.up[1] == 1.f && p.view[2] == -1.f);
        CHECK(close_to(p.pinhole[2], 6.f - 1.f / std::tan(0.125f * 3.14159265358979f)));
        Ray c = pc.sample_ray(Eigen::Vector2d(0.5, 0.5));
        CHECK(c.origin == pos && close_to(c.direction[2], -1.f) && std::abs(c.direction[0]) < 1e-7f);
        Ray tl = pc.sample_ray(Eigen::Vector2d(0.0, 0.0));   // u is mirrored: uv (0,0) starts at +right, -up
        CHECK(close_to(tl.origin[0], 1.f) && close_to(tl.origin[1], 0.f, 1e-5f, 1e-6f) && tl.direction[0] < 0.f && tl.direction[1] > 0.f);
        CHECK(close_to(tl.direction.norm(), 1.f, 1e-6f));
        Orthographic_Camera oc(pos, view);
        Ray o = oc.sample_ray(Eigen::Vector2d(0.25, 0.75));
        CHECK(oc.pod().kind == 1 && close_to(o.origin[0], -0.5f) && close_to(o.origin[1], 0.5f) && o.direction == view);
        std::puts("ok cameras");
    }
    {   // ---- Image (image.h:9-106)
        Image im(5, 3);
        CHECK(im.get_width() == 5 && im.get_height() == 3);
        im.set_pixel(4, 2, Eigen::Vector3f(0.25f, 2.0f, -1.0f));
        im.set_pixel(1, 0, Eigen::Vector3f(1.0f, 0.5f, 0.0f));
        CHECK(im.data()[3 * (2 * 5 + 4) + 1] == 2.0f && im.get_pixel(1, 0) == Eigen::Vector3f(1.0f, 0.5f, 0.0f));
        im.make_PPM(tmp + "/t.ppm");
        Image back(tmp + "/t.ppm");
        CHECK(back.get_width() == 5 && back.get_height() == 3);
        CHECK(back.get_pixel(4, 2) == Eigen::Vector3f(63 / 255.f, 1.0f, 0.0f));   // 0.25*255 = 63.75 truncates; clamps at both ends
        CHECK(back.get_pixel(1, 0) == Eigen::Vector3f(1.0f, 127 / 255.f, 0.0f));
        const std::vector<uint8_t> rgba = im.get_rgba_buffer();
        CHECK(rgba.size() == 60 && rgba[4 * (2 * 5 + 4)] == 63 && rgba[4 * (2 * 5 + 4) + 3] == 255);
        bool threw = false;
        try { Image bad(tmp + "/mini.txt"); } catch (const std::runtime_error&) { threw = true; }
        CHECK(threw);
        std::puts("ok image");
    }
    {   // ---- 11-float parameterisation (gmm.h:583-706): pack, then the unpack formulas on the host
        Scene s = Scene::load_GMM(scenes + "/rand50.txt");
        const GaussianMixtureModel& m = s.gmm->at(0);
        std::vector<float> p;
        m.pack_parameters(p);
        CHECK(p.size() == 11 * m.get_num_gaussians());
        const std::vector<float> eps = make_default_eps_for_params(p);
        CHECK(eps[0] == 0.02f && eps[3] == 0.10f && eps[6] == 0.05f && eps[9] == 0.25f && eps[10] == 0.5f && eps[11] == 0.02f);
        for (size_t i = 0; i < m.get_num_gaussians(); ++i) {
            const Gaussian& g = m.gaussians[i];
            const float* q = &p[11 * i];
            CHECK(q[0] == g.centroid()[0] && q[2] == g.centroid()[2]);
            CHECK(close_to(std::exp(q[9]), g.get_density(), 1e-5f) && close_to(sigmoidf_safe(q[10]), g.get_albedo(), 1e-5f));
            Eigen::Vector3f rod(q[3], q[4], q[5]);
            Eigen::Matrix3f R = Eigen::Matrix3f::Identity();
            const float ang = rod.norm();
            if (ang > 1e-12f) R = Eigen::AngleAxisf(ang, rod / ang).toRotationMatrix();
            Eigen::Matrix3f S = Eigen::Matrix3f::Zero();
            for (int k = 0; k < 3; ++k) S(k, k) = std::exp(q[6 + k]);
            const Eigen::Matrix3f cov = R * S * S.transpose() * R.transpose();
            float scale = 0.f;
            for (int k = 0; k < 3; ++k) scale = std::max(scale, g.get_covariance()(k, k));
            for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) CHECK(std::abs(cov(r, c) - g.get_covariance()(r, c)) <= 2e-5f * scale);
        }
        GaussianMixtureModel wrong = m;
        std::vector<float> short_params(11);
        bool threw = false;
        try { apply_params_to_gmm_local(short_params, wrong); } catch (const std::runtime_error&) { threw = true; }
        CHECK(threw);
        std::puts("ok pack_parameters");
    }
    {   // ---- optimizer / integrator error behaviour that needs no device
        AdamOptimizer adam(4, 1e-2f);
        std::vector<float> p(4, 1.f), g(3, 1.f);
        CHECK(!adam.step(p, g) && adam.dim() == 4);
        auto cam = std::make_shared<Pinhole_Camera>(Eigen::Vector3f(0.f, 1.f, 6.f), Eigen::Vector3f(0.f, 0.f, -1.f), 0.785398f);
        Scene s = Scene::load_GMM(scenes + "/rand50.txt");
        Image im(8, 8);
        bool threw = false;
        Scene sph = Scene::load_SMM(scenes + "/one_sphere.txt");   // the Gaussian validation integrators do not take sphere scenes
        try { PureRayMarching(cam).render(sph, im); } catch (const std::runtime_error&) { threw = true; }
        CHECK(threw);
        if (vr_device_count() == 0) {   // no GPU: the product must fail loudly, not fall back
            threw = false;
            try { MultiScatterGaussians(cam, 4).render(s, im); } catch (const std::runtime_error& e) { threw = std::string(e.what()).find("no CPU fallback") != std::string::npos; }
            CHECK(threw);
            std::puts("ok no_cpu_fallback");
        }
        std::puts("ok errors");
    }
    std::puts("all host checks passed");
    return 0;
}
