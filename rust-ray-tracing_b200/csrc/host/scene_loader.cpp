// scene_loader.cpp — see scene_loader.hpp.
#include "scene_loader.hpp"

#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>

namespace rtbhost {

using tomls::Table;
using tomls::Value;

namespace {

const Value& index(const Table& t, const char* key, const char* where) {
    const Value* v = t.find(key);
    if (!v) throw SceneError(std::string(where) + ": missing key \"" + key + "\"");   // table["key"] panics in the reference
    return *v;
}

std::string lower(std::string s) {
    for (auto& c : s) c = (char)std::tolower((unsigned char)c);   // to_ascii_lowercase
    return s;
}

}  // namespace

double to_float(const Value& v) {
    if (v.is_float()) return v.f;
    if (v.is_integer()) return (double)v.i;
    throw SceneError(std::string("expected a TOML float or integer, got ") + v.type_name());   // to_float(..).unwrap()
}

void vec3_from_toml(const Value& v, double out[3]) {
    if (v.is_table()) {
        out[0] = to_float(index(*v.tab, "x", "vector"));
        out[1] = to_float(index(*v.tab, "y", "vector"));
        out[2] = to_float(index(*v.tab, "z", "vector"));
        return;
    }
    if (v.is_array()) {
        if (v.arr->size() < 3) throw SceneError("assertion failed: array.len() >= 3");
        for (int k = 0; k < 3; k++) out[k] = to_float((*v.arr)[k]);
        return;
    }
    throw SceneError(std::string("expected a table or an array for a vector, got ") + v.type_name());   // None.unwrap()
}

void color_from_toml(const Value& v, double out[3]) {
    if (v.is_table()) {
        const char* keys[3] = {"red", "green", "blue"};
        for (int k = 0; k < 3; k++) {
            const Value& c = index(*v.tab, keys[k], "colour");          // a missing key panics ...
            out[k] = (c.is_float() || c.is_integer()) ? to_float(c) : 0.0;   // ... a wrongly typed one is unwrap_or(0.0)
        }
        return;
    }
    if (v.is_array()) {
        if (v.arr->size() < 3) throw SceneError("assertion failed: array.len() >= 3");
        for (int k = 0; k < 3; k++) out[k] = to_float((*v.arr)[k]);
        return;
    }
    throw SceneError(std::string("expected a table or an array for a colour, got ") + v.type_name());
}

static rtb200_material load_material(const Table& t) {   // load_material_from_toml, src/materials.rs:21-33
    const Value& ty = index(t, "type", "material");
    if (!ty.is_string()) throw SceneError("material type must be a string");
    std::string kind = lower(ty.s);
    rtb200_material m;
    std::memset(&m, 0, sizeof m);
    if (kind == "lambertian") {                           // :65-68
        m.kind = RTB200_LAMBERTIAN;
        color_from_toml(index(t, "albedo", "lambertian"), m.albedo);
    } else if (kind == "metal") {                         // :99-106
        m.kind = RTB200_METAL;
        color_from_toml(index(t, "albedo", "metal"), m.albedo);
        const double fuzz = to_float(index(t, "fuzzy_factor", "metal"));
        m.param = fuzz < 1.0 ? fuzz : 1.0;                // Metal::new, src/materials.rs:79-88
    } else if (kind == "dielectric") {                    // :149-154
        m.kind = RTB200_DIELECTRIC;
        m.param = to_float(index(t, "index_of_refraction", "dielectric"));
        const Value& h = index(t, "hollow", "dielectric");
        if (!h.is_boolean()) throw SceneError("dielectric: hollow must be a boolean");   // as_bool().unwrap()
        m.hollow = h.b ? 1 : 0;
    } else {
        throw SceneError("Unknown material type " + kind + "!");   // :31
    }
    return m;
}

LoadedScene load_scene(const Table& doc) {
    LoadedScene out;
    const Value& mats = index(doc, "materials", "scene");            // scene_data["materials"].as_table().unwrap()
    if (!mats.is_table()) throw SceneError("materials must be a table");
    std::vector<std::pair<std::string, const Value*>> named;
    for (const auto& kv : mats.tab->items) named.emplace_back(kv.first, &kv.second);
    std::sort(named.begin(), named.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    for (const auto& kv : named) {
        if (!kv.second->is_table()) throw SceneError("material \"" + kv.first + "\" must be a table");
        out.material_names.push_back(kv.first);
        out.materials.push_back(load_material(*kv.second->tab));
    }
    const Value& hit = index(doc, "hitables", "scene");               // scene_data["hitables"].as_array().unwrap()
    if (!hit.is_array()) throw SceneError("hitables must be an array of tables");
    for (const Value& ov : *hit.arr) {
        if (!ov.is_table()) throw SceneError("hitables must be an array of tables");
        const Table& t = *ov.tab;
        const Value& ty = index(t, "type", "object");
        if (!ty.is_string()) throw SceneError("object type must be a string");
        std::string kind = lower(ty.s);
        if (kind != "sphere") throw SceneError("Unknown object type " + kind);   // src/objects.rs:50
        rtb200_sphere s;
        std::memset(&s, 0, sizeof s);
        vec3_from_toml(index(t, "center", "sphere"), s.center);          // :293
        s.radius = to_float(index(t, "radius", "sphere"));               // :294
        const Value& mn = index(t, "material", "sphere");
        if (!mn.is_string()) throw SceneError("sphere: material must be a string");
        auto it = std::lower_bound(out.material_names.begin(), out.material_names.end(), mn.s);
        if (it == out.material_names.end() || *it != mn.s) throw SceneError("sphere names material \"" + mn.s + "\" which does not exist");   // :296
        s.material = (int32_t)(it - out.material_names.begin());
        out.spheres.push_back(s);
    }
    return out;
}

LoadedScene load_scene_text(const std::string& toml_text) {
    Table doc;
    try {
        doc = tomls::parse(toml_text);
    } catch (const tomls::ParseError& e) {
        throw SceneError(std::string("Failed parsing scene: ") + e.what());   // src/main.rs:36-39
    }
    return load_scene(doc);
}

LoadedScene load_scene_file(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw SceneError("Can't read scene from file " + path + ": cannot open");   // src/main.rs:31-34
    std::stringstream ss;
    ss << f.rdbuf();
    try {
        return load_scene_text(ss.str());
    } catch (const SceneError& e) {
        throw SceneError(std::string(e.what()) + " (file " + path + ")");
    }
}

bool load_camera(const Table& doc, CameraDesc& cam) {
    const Value* c = doc.find("camera");
    if (!c) return false;
    if (!c->is_table()) throw SceneError("camera must be a table");
    const Table& t = *c->tab;
    const Value& w = index(t, "image_width", "camera");
    const Value& h = index(t, "image_height", "camera");
    if (!w.is_integer() || !h.is_integer() || w.i <= 0 || h.i <= 0 || w.i > 1 << 20 || h.i > 1 << 20)
        throw SceneError("camera: image_width / image_height must be positive integers");   // as_integer().unwrap().try_into().unwrap()
    cam.image_width = (int32_t)w.i; cam.image_height = (int32_t)h.i;
    const Value& fl = index(t, "focal_length", "camera");
    const Value& va = index(t, "view_angle", "camera");
    if (!fl.is_float() || !va.is_float()) throw SceneError("camera: focal_length / view_angle must be TOML floats");   // as_float().unwrap()
    cam.focal_length = fl.f; cam.view_angle = va.f;
    vec3_from_toml(index(t, "center", "camera"), cam.center);
    vec3_from_toml(index(t, "look_at", "camera"), cam.look_at);
    vec3_from_toml(index(t, "up", "camera"), cam.up);
    const Value* da = t.find("defocus_angle");
    cam.defocus_angle = da ? to_float(*da) : 0.0;
    return true;
}

std::string scene_to_json(const LoadedScene& s) {
    std::string out = "{\"spheres\": [";
    char b[512];
    for (size_t i = 0; i < s.spheres.size(); i++) {
        const rtb200_sphere& sp = s.spheres[i];
        const rtb200_material& m = s.materials[(size_t)sp.material];
        std::snprintf(b, sizeof b,
                      "%s{\"center\": [%.17g, %.17g, %.17g], \"radius\": %.17g, \"material\": \"%s\", \"kind\": %d, \"hollow\": %d, "
                      "\"albedo\": [%.17g, %.17g, %.17g], \"param\": %.17g}",
                      i ? ", " : "", sp.center[0], sp.center[1], sp.center[2], sp.radius, s.material_names[(size_t)sp.material].c_str(), m.kind,
                      m.hollow, m.albedo[0], m.albedo[1], m.albedo[2], m.param);
        out += b;
    }
    out += "], \"n_materials\": " + std::to_string(s.materials.size()) + "}";
    return out;
}

}  // namespace rtbhost
