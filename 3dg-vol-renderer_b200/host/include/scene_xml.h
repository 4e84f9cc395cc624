// scene_xml.h -- reader for the Mitsuba-3 subset used by tests/env_one_sphere_test_ortho.xml
// (the reference ships the XML as the external cross-check of config C0 but no
// loader for it).  Understood elements:
//   <sensor type="orthographic|perspective"> <lookat origin= target= up=/> <float name="fov" value=deg/>
//        <integer name="width|height|sample_count" value=/>
//   <emitter type="constant"> <rgb name="radiance" value="r, g, b | v"/>
//   <emitter type="point">    <point name="position" x= y= z=/> <rgb name="intensity" value=/>
//   <medium id= type="homogeneous"> <rgb name="albedo"/> <rgb name="sigma_t"/> <float name="scale"/>
//   <shape type="sphere"> <point name="center"/> <float name="radius"/> <ref name="interior" id=/>
// A sphere becomes Sphere(center, radius, sigma_a, sigma_s) with sigma_s = albedo * sigma_t * scale and
// sigma_a = (1 - albedo) * sigma_t * scale (channel 0 of the rgb values; the sphere model is grey).
// Everything else (integrator, sampler, film filter, bsdf, phase) is ignored.  Included by scene.h.
#pragma once
#include <cctype>
#include <cstdlib>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

namespace vrhost {
namespace xml {

struct Element {
    std::string name;
    std::map<std::string, std::string> attr;
    bool closing = false, self_closing = false;
    const std::string& get(const std::string& k) const {
        static const std::string empty;
        auto it = attr.find(k);
        return it == attr.end() ? empty : it->second;
    }
};

// splits the text into elements; comments, declarations and character data are dropped
inline std::vector<Element> scan(const std::string& text) {
    std::vector<Element> out;
    size_t i = 0;
    while ((i = text.find('<', i)) != std::string::npos) {
        if (text.compare(i, 4, "<!--") == 0) { const size_t e = text.find("-->", i); i = e == std::string::npos ? text.size() : e + 3; continue; }
        const size_t end = text.find('>', i);
        if (end == std::string::npos) break;
        std::string body = text.substr(i + 1, end - i - 1);
        i = end + 1;
        if (body.empty() || body[0] == '?' || body[0] == '!') continue;
        Element el;
        if (body[0] == '/') { el.closing = true; body.erase(0, 1); }
        if (!body.empty() && body.back() == '/') { el.self_closing = true; body.pop_back(); }
        size_t p = 0;
        while (p < body.size() && !std::isspace((unsigned char)body[p])) ++p;
        el.name = body.substr(0, p);
        while (p < body.size()) {
            while (p < body.size() && std::isspace((unsigned char)body[p])) ++p;
            const size_t eq = body.find('=', p);
            if (eq == std::string::npos) break;
            std::string key = body.substr(p, eq - p);
            while (!key.empty() && std::isspace((unsigned char)key.back())) key.pop_back();
            size_t q = eq + 1;
            while (q < body.size() && std::isspace((unsigned char)body[q])) ++q;
            if (q >= body.size() || (body[q] != '"' && body[q] != '\'')) break;
            const char quote = body[q];
            const size_t qe = body.find(quote, q + 1);
            if (qe == std::string::npos) break;
            el.attr[key] = body.substr(q + 1, qe - q - 1);
            p = qe + 1;
        }
        out.push_back(el);
    }
    return out;
}

// "a, b, c" or "a b c" or "v" -> up to three floats; a single value is replicated
inline void floats3(const std::string& s, float out[3]) {
    std::string t = s;
    for (char& c : t) if (c == ',') c = ' ';
    std::stringstream ss(t);
    int n = 0;
    float v;
    while (n < 3 && (ss >> v)) out[n++] = v;
    if (n == 1) out[1] = out[2] = out[0];
}
inline void xyz(const Element& e, float out[3]) {
    out[0] = float(std::atof(e.get("x").c_str())); out[1] = float(std::atof(e.get("y").c_str())); out[2] = float(std::atof(e.get("z").c_str()));
}

}  // namespace xml
}  // namespace vrhost

inline Scene Scene::load_XML(const std::string& filename, XmlSensor* sensor_out) {
    using vrhost::xml::Element;
    std::ifstream f(filename);
    if (!f) throw std::runtime_error("Failed to open scene file: " + filename);
    std::stringstream buf;
    buf << f.rdbuf();
    const std::vector<Element> els = vrhost::xml::scan(buf.str());

    struct Medium { float albedo = 1.f, sigma_t = 1.f, scale = 1.f; };
    struct Shape { float center[3] = {0, 0, 0}; float radius = 1.f; std::string interior; Medium inline_medium; bool has_inline = false; };
    std::map<std::string, Medium> media;
    std::vector<Shape> shapes;
    XmlSensor sensor;
    Scene s;
    s.volume_type = VolumeType::SPHERES;
    bool have_env = false;

    std::vector<Element> stack;   // open elements
    Medium cur_medium; std::string cur_medium_id;
    Shape cur_shape;
    float light_pos[3] = {0, 0, 0}, light_int[3] = {0, 0, 0};
    auto inside = [&](const char* name, const char* type = nullptr) {
        for (auto it = stack.rbegin(); it != stack.rend(); ++it)
            if (it->name == name && (!type || it->get("type") == type)) return true;
        return false;
    };
    for (const Element& e : els) {
        if (e.closing) {
            if (e.name == "medium") {
                if (inside("shape")) { cur_shape.inline_medium = cur_medium; cur_shape.has_inline = true; }
                else media[cur_medium_id] = cur_medium;
            } else if (e.name == "shape") {
                if (!stack.empty() && stack.back().get("type") == "sphere") shapes.push_back(cur_shape);
            } else if (e.name == "emitter") {
                if (!stack.empty() && stack.back().get("type") == "point")
                    s.lights.push_back({Eigen::Vector3f(light_pos[0], light_pos[1], light_pos[2]), Eigen::Vector3f(light_int[0], light_int[1], light_int[2])});
            }
            if (!stack.empty()) stack.pop_back();
            continue;
        }
        const std::string& nm = e.get("name");
        if (e.name == "sensor") sensor.orthographic = e.get("type") != "perspective";
        else if (e.name == "medium") { cur_medium = Medium(); cur_medium_id = e.get("id"); }
        else if (e.name == "shape") cur_shape = Shape();
        else if (e.name == "emitter") { light_pos[0] = light_pos[1] = light_pos[2] = 0; light_int[0] = light_int[1] = light_int[2] = 0; }
        else if (e.name == "lookat" && inside("sensor")) {
            float o[3] = {0, 0, 0}, t[3] = {0, 0, 0};
            vrhost::xml::floats3(e.get("origin"), o); vrhost::xml::floats3(e.get("target"), t);
            sensor.origin = Eigen::Vector3f(o[0], o[1], o[2]); sensor.target = Eigen::Vector3f(t[0], t[1], t[2]);
        } else if (e.name == "integer" && inside("sensor")) {
            const int v = std::atoi(e.get("value").c_str());
            if (nm == "width") sensor.width = v; else if (nm == "height") sensor.height = v; else if (nm == "sample_count") sensor.sample_count = v;
        } else if (e.name == "float") {
            const float v = float(std::atof(e.get("value").c_str()));
            if (nm == "fov" && inside("sensor")) sensor.fov_radians = v * 3.14159265358979323846f / 180.0f;
            else if (nm == "scale" && inside("medium")) cur_medium.scale = v;
            else if (nm == "radius" && inside("shape")) cur_shape.radius = v;
        } else if (e.name == "rgb" || e.name == "spectrum") {
            float v[3] = {0, 0, 0};
            vrhost::xml::floats3(e.get("value"), v);
            if (inside("medium")) { if (nm == "albedo") cur_medium.albedo = v[0]; else if (nm == "sigma_t") cur_medium.sigma_t = v[0]; }
            else if (inside("emitter", "constant") && nm == "radiance") { s.env_color = Eigen::Vector3f(v[0], v[1], v[2]); have_env = true; }
            else if (inside("emitter", "point") && nm == "intensity") { light_int[0] = v[0]; light_int[1] = v[1]; light_int[2] = v[2]; }
        } else if (e.name == "point") {
            if (inside("emitter", "point") && nm == "position") vrhost::xml::xyz(e, light_pos);
            else if (inside("shape") && nm == "center") vrhost::xml::xyz(e, cur_shape.center);
        } else if (e.name == "ref" && inside("shape") && nm == "interior") cur_shape.interior = e.get("id");
        if (!e.self_closing) stack.push_back(e);
    }
    if (!have_env) s.env_color = Eigen::Vector3f(0.f, 0.f, 0.f);   // Mitsuba: no constant emitter = black background

    std::vector<Sphere> spheres;
    for (const Shape& sh : shapes) {
        Medium m;
        if (sh.has_inline) m = sh.inline_medium;
        else if (media.count(sh.interior)) m = media[sh.interior];
        else throw std::runtime_error("XML scene: sphere refers to unknown medium '" + sh.interior + "'");
        const float st = m.sigma_t * m.scale;
        spheres.emplace_back(Eigen::Vector3f(sh.center[0], sh.center[1], sh.center[2]), sh.radius, (1.0f - m.albedo) * st, m.albedo * st);
    }
    s.smm = std::vector<SphereMixtureModel>{SphereMixtureModel(spheres)};
    if (sensor_out) *sensor_out = sensor;
    return s;
}
