// scene_loader.hpp — the native host's L1 layer: scene.toml -> the C-ABI scene (include/rtb200.h).
// Mirrors the reference loaders (paths relative to the reference root): `to_float` src/toml_utils.rs:2-11,
// `Vec3::from_toml` src/geometry.rs:190-211, `Color::from_toml` src/color.rs:99-123, `get_materials` /
// `load_material_from_toml` src/materials.rs:12-33 and the three `from_table` loaders (:65-68, :99-106, :149-154),
// `get_object_list` / `load_object_from_toml` / `Sphere::from_table` src/objects.rs:38-52,292-299, and
// `Camera::from_toml` src/ray_tracing.rs:64-75.  Where the reference panics, these throw SceneError with the
// reference's message where it has one.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../../../include/rtb200.h"
#include "toml_subset.hpp"

namespace rtbhost {

struct SceneError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// Owns the arrays an rtb200_scene points into.  Materials are in sorted-name order (the reference's toml::Table is
// a BTreeMap, src/materials.rs:13); spheres in file order (src/objects.rs:39) — the order the tie-break depends on.
struct LoadedScene {
    std::vector<std::string> material_names;
    std::vector<rtb200_material> materials;
    std::vector<rtb200_sphere> spheres;
    rtb200_scene abi() const {
        rtb200_scene s;
        s.spheres = spheres.data(); s.n_spheres = (int32_t)spheres.size();
        s.materials = materials.data(); s.n_materials = (int32_t)materials.size();
        return s;
    }
};

// Camera description = the arguments of Camera::new (src/ray_tracing.rs:27-35); defaults are src/main.rs:45-58.
struct CameraDesc {
    int32_t image_width = 3840, image_height = 2160;
    double focal_length = 10.0, view_angle = 30.0;
    double center[3] = {16.0, 2.0, 18.5};
    double look_at[3] = {0.0, 0.0, 0.0};
    double up[3] = {0.0, 1.0, 0.0};
    double defocus_angle = 0.0;
};

double to_float(const tomls::Value& v);                              // src/toml_utils.rs:2-11
void vec3_from_toml(const tomls::Value& v, double out[3]);           // src/geometry.rs:190-211
void color_from_toml(const tomls::Value& v, double out[3]);          // src/color.rs:99-123

// scene_data["materials"] + scene_data["hitables"] (src/main.rs:41-43).
LoadedScene load_scene(const tomls::Table& doc);
LoadedScene load_scene_text(const std::string& toml_text);
LoadedScene load_scene_file(const std::string& path);                // "Can't read scene from file ..." (src/main.rs:31-34)

// Optional [camera] table with Camera::from_toml's keys (src/ray_tracing.rs:64-75).  Returns false when the document
// has no [camera] table (the reference's main never reads one: src/main.rs:51-58).  Unlike the reference (which reads
// `defocus_angle` from the "view_angle" key, :72 — SURVEY F12) the defocus angle comes from its own key, default 0.
bool load_camera(const tomls::Table& doc, CameraDesc& cam);

// Flattened scene as JSON (cross-check against the Python host in tests).
std::string scene_to_json(const LoadedScene& s);

}  // namespace rtbhost
