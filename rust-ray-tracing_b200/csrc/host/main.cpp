// main.cpp — rtb200_cli, the native drop-in for the reference's `main` (src/main.rs:27-74): reads scene.toml, renders
// through the C ABI (include/rtb200.h) on a B200, saves output.png and prints the same summary lines.  What the
// reference hard-codes (src/main.rs:29,45-64: scene path, camera, 3840x2160, 100 spp, depth 50, output path) are the
// defaults of the options below.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "../../../include/rtb200.h"
#include "image_out.hpp"
#include "scene_loader.hpp"
#include "toml_subset.hpp"

namespace {

void usage() {
    std::puts(
        "usage: rtb200_cli [scene.toml] [options]\n"
        "  -o, --output PATH       output image (.png or .ppm)                 [output.png]\n"
        "  --width N --height N    image size                                  [3840 x 2160]\n"
        "  --spp N                 samples per pixel                           [100]\n"
        "  --depth N               max bounces                                 [50]\n"
        "  --seed N                Philox key of the sample stream             [1]\n"
        "  --device N              CUDA device                                 [0]\n"
        "  --gpus N                render on devices 0..N-1 in ONE call (rtb200_multi_render: traces sharded over the GPUs,\n"
        "                          partial framebuffers summed over NVLink peer memory by the resolve kernel)   [1]\n"
        "  --devices A,B,...       the same on the named devices (a device may repeat)\n"
        "  --sharding MODE         samples (default) | rows | RxS (R interleaved-scanline groups x S sample groups)\n"
        "  --focal-length F --view-angle DEG --defocus-angle DEG               [10, 30, 0]\n"
        "  --center X,Y,Z --look-at X,Y,Z --up X,Y,Z                           [16,2,18.5  0,0,0  0,1,0]\n"
        "  --camera-from-scene     use the scene file's [camera] table (Camera::from_toml keys)\n"
        "  --both-roots            scalar Sphere::hit rule (far root when the near one is out of range): physically right\n"
        "                          glass; NOT what the reference's packed renderer draws\n"
        "  --dump-toml             print the parsed TOML document as JSON and exit (no GPU)\n"
        "  --dump-scene            print the flattened scene as JSON and exit (no GPU)\n"
        "  --dump-camera           print the derived camera (Camera::new) as JSON and exit (no GPU)\n"
        "  --progress              redraw a progress line on stderr while rendering (the reference's progress grid)\n"
        "  --stats-json            also print the render statistics as one JSON line\n"
        "  --quiet                 no summary lines");
}

bool parse3(const char* s, double out[3]) { return std::sscanf(s, "%lf,%lf,%lf", &out[0], &out[1], &out[2]) == 3; }

std::string read_file(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw rtbhost::SceneError("Can't read scene from file " + path + ": cannot open");   // src/main.rs:31-34
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

}  // namespace

int main(int argc, char** argv) {
    std::string scene_path = "scene.toml", out_path = "output.png";   // src/main.rs:29,66
    rtbhost::CameraDesc cam;                                            // src/main.rs:45-58
    int spp = 100, depth = 50, device = 0;                              // src/main.rs:64
    unsigned long long seed = 1;
    bool dump_toml = false, dump_scene = false, dump_camera = false, stats_json = false, quiet = false, camera_from_scene = false;
    bool both_roots = false, progress = false;
    bool have_scene_arg = false;
    std::vector<int32_t> devices;                                       // --gpus / --devices: the shards of rtb200_multi
    std::string sharding = "samples";
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto need = [&](const char* name) -> const char* {
            if (i + 1 >= argc) { std::fprintf(stderr, "rtb200_cli: %s needs a value\n", name); std::exit(2); }
            return argv[++i];
        };
        if (a == "-h" || a == "--help") { usage(); return 0; }
        else if (a == "-o" || a == "--output") out_path = need("--output");
        else if (a == "--width") cam.image_width = std::atoi(need("--width"));
        else if (a == "--height") cam.image_height = std::atoi(need("--height"));
        else if (a == "--spp") spp = std::atoi(need("--spp"));
        else if (a == "--depth") depth = std::atoi(need("--depth"));
        else if (a == "--seed") seed = std::strtoull(need("--seed"), nullptr, 10);
        else if (a == "--device") device = std::atoi(need("--device"));
        else if (a == "--gpus") { const int n = std::atoi(need("--gpus")); devices.clear(); for (int g = 0; g < n; g++) devices.push_back(g); }
        else if (a == "--devices") {
            devices.clear();
            std::stringstream ss(need("--devices"));
            std::string tok;
            while (std::getline(ss, tok, ',')) if (!tok.empty()) devices.push_back(std::atoi(tok.c_str()));
        }
        else if (a == "--sharding") sharding = need("--sharding");
        else if (a == "--focal-length") cam.focal_length = std::atof(need("--focal-length"));
        else if (a == "--view-angle") cam.view_angle = std::atof(need("--view-angle"));
        else if (a == "--defocus-angle") cam.defocus_angle = std::atof(need("--defocus-angle"));
        else if (a == "--center") { if (!parse3(need("--center"), cam.center)) { std::fprintf(stderr, "rtb200_cli: --center wants X,Y,Z\n"); return 2; } }
        else if (a == "--look-at") { if (!parse3(need("--look-at"), cam.look_at)) { std::fprintf(stderr, "rtb200_cli: --look-at wants X,Y,Z\n"); return 2; } }
        else if (a == "--up") { if (!parse3(need("--up"), cam.up)) { std::fprintf(stderr, "rtb200_cli: --up wants X,Y,Z\n"); return 2; } }
        else if (a == "--camera-from-scene") camera_from_scene = true;
        else if (a == "--both-roots") both_roots = true;
        else if (a == "--progress") progress = true;
        else if (a == "--dump-toml") dump_toml = true;
        else if (a == "--dump-scene") dump_scene = true;
        else if (a == "--dump-camera") dump_camera = true;
        else if (a == "--stats-json") stats_json = true;
        else if (a == "--quiet") quiet = true;
        else if (!a.empty() && a[0] != '-' && !have_scene_arg) { scene_path = a; have_scene_arg = true; }
        else { std::fprintf(stderr, "rtb200_cli: unknown argument %s\n", a.c_str()); usage(); return 2; }
    }

    rtbhost::LoadedScene scene;
    try {
        std::string text = read_file(scene_path);
        tomls::Table doc;
        try {
            doc = tomls::parse(text);
        } catch (const tomls::ParseError& e) {
            throw rtbhost::SceneError("Failed parsing scene from file " + scene_path + ": " + e.what());   // src/main.rs:36-39
        }
        if (dump_toml) { std::puts(tomls::to_json(doc).c_str()); return 0; }
        scene = rtbhost::load_scene(doc);
        if (camera_from_scene && !rtbhost::load_camera(doc, cam)) throw rtbhost::SceneError("--camera-from-scene: the scene has no [camera] table");
    } catch (const std::exception& e) {
        std::fprintf(stderr, "rtb200_cli: %s\n", e.what());
        return 1;
    }
    if (dump_scene) { std::puts(rtbhost::scene_to_json(scene).c_str()); return 0; }

    rtb200_camera camera;
    if (rtb200_camera_init(&camera, cam.image_width, cam.image_height, cam.focal_length, cam.view_angle, cam.center, cam.look_at, cam.up,
                           cam.defocus_angle) != RTB200_OK) {
        std::fprintf(stderr, "rtb200_cli: %s\n", rtb200_last_error());
        return 1;
    }
    if (dump_camera) {
        auto v = [](const double* p) { char b[128]; std::snprintf(b, sizeof b, "[%.17g, %.17g, %.17g]", p[0], p[1], p[2]); return std::string(b); };
        std::printf("{\"image_width\": %d, \"image_height\": %d, \"center\": %s, \"viewport_upper_left_corner\": %s, \"viewport_u\": %s, "
                    "\"viewport_v\": %s, \"defocus_u\": %s, \"defocus_v\": %s}\n",
                    camera.image_width, camera.image_height, v(camera.center).c_str(), v(camera.viewport_upper_left_corner).c_str(),
                    v(camera.viewport_u).c_str(), v(camera.viewport_v).c_str(), v(camera.defocus_u).c_str(), v(camera.defocus_v).c_str());
        return 0;
    }
    if (!quiet) {   // Camera::new's report (src/ray_tracing.rs:46-50) and main's object count (src/main.rs:60)
        auto len = [](const double* p) { return std::sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]); };
        std::printf("Parameters:\n\tCamera Center:            (%g, %g, %g)\n\tViewport Height:          %g\n\tViewport Width:           %g\n"
                    "\tViewport Top Left Corner: (%g, %g, %g)\n",
                    camera.center[0], camera.center[1], camera.center[2], len(camera.viewport_v), len(camera.viewport_u),
                    camera.viewport_upper_left_corner[0], camera.viewport_upper_left_corner[1], camera.viewport_upper_left_corner[2]);
        std::printf("\t Number of objects: \t %zu\n", scene.spheres.size());
    }

    rtb200_ctx* ctx = nullptr;
    rtb200_multi* multi = nullptr;
    auto die = [&](const char* what) {
        std::fprintf(stderr, "rtb200_cli: %s: %s\n", what, rtb200_last_error());
        if (ctx) rtb200_destroy(ctx);
        if (multi) rtb200_multi_destroy(multi);
        return 1;
    };
    int row_groups = 1;
    if (!devices.empty()) {   // several GPUs behind the one render call (the reference's render owns its worker threads the same way)
        const int n = (int)devices.size();
        if (sharding == "samples") row_groups = 1;
        else if (sharding == "rows") row_groups = n;
        else {
            int r = 0, sgr = 0;
            if (std::sscanf(sharding.c_str(), "%dx%d", &r, &sgr) != 2 || r < 1 || sgr < 1 || r * sgr != n) {
                std::fprintf(stderr, "rtb200_cli: --sharding wants samples, rows or RxS with R*S = %d shards\n", n);
                return 2;
            }
            row_groups = r;
        }
        if (rtb200_multi_create(devices.data(), n, &multi) != RTB200_OK) return die("rtb200_multi_create");
        if (both_roots && rtb200_multi_set_hit_mode(multi, RTB200_HIT_BOTH_ROOTS) != RTB200_OK) return die("rtb200_multi_set_hit_mode");
    } else {
        if (rtb200_create(device, &ctx) != RTB200_OK) return die("rtb200_create");       // TileRenderer::new, src/main.rs:62
        if (both_roots && rtb200_set_hit_mode(ctx, RTB200_HIT_BOTH_ROOTS) != RTB200_OK) return die("rtb200_set_hit_mode");
    }
    if (progress) {   // src/renderer.rs:300-365 draws a tile grid; one line is enough for a render that takes milliseconds to seconds
        auto draw = [](void*, uint64_t done, uint64_t total, double secs) {
            const double frac = total ? (double)done / (double)total : 1.0;
            const double eta = frac > 0.0 ? secs * (1.0 - frac) / frac : 0.0;
            char bar[41];
            const int fill = (int)(frac * 40.0);
            for (int i = 0; i < 40; i++) bar[i] = i < fill ? '#' : '.';
            bar[40] = 0;
            std::fprintf(stderr, "\r[%s] %5.1f %%  %.2f s elapsed, ETA %.2f s ", bar, 100.0 * frac, secs, eta);
            if (done == total) std::fputc('\n', stderr);
            std::fflush(stderr);
        };
        if ((multi ? rtb200_multi_set_progress(multi, draw, nullptr, 100) : rtb200_set_progress(ctx, draw, nullptr, 100)) != RTB200_OK)
            return die("rtb200_set_progress");
    }
    rtb200_scene abi = scene.abi();
    if ((multi ? rtb200_multi_upload_scene(multi, &abi) : rtb200_upload_scene(ctx, &abi)) != RTB200_OK) return die("rtb200_upload_scene");  // Scene::from_list, src/main.rs:43
    rtb200_params params;
    std::memset(&params, 0, sizeof params);
    params.max_bounces = depth; params.samples_per_pixel = spp; params.seed = seed;
    params.sample_begin = 0; params.sample_end = spp; params.row_begin = 0; params.row_end = cam.image_height;
    const size_t n_bytes = (size_t)cam.image_width * cam.image_height * 3;
    void* pinned = nullptr;
    if (rtb200_host_alloc(n_bytes, &pinned) != RTB200_OK) return die("rtb200_host_alloc");
    rtb200_stats st;
    const int render_rc = multi ? rtb200_multi_render(multi, &camera, &params, row_groups, static_cast<uint8_t*>(pinned), nullptr, &st)
                                : rtb200_render(ctx, &camera, &params, static_cast<uint8_t*>(pinned), nullptr, &st);
    if (render_rc != RTB200_OK) {   // renderer.render(50, 100, ..), src/main.rs:64
        rtb200_host_free(pinned);
        return die("rtb200_render");
    }
    int rc = 0;
    try {
        rtbhost::save_image(out_path, static_cast<const uint8_t*>(pinned), cam.image_width, cam.image_height);   // src/main.rs:66
    } catch (const std::exception& e) {
        std::fprintf(stderr, "rtb200_cli: %s\n", e.what());
        rc = 1;
    }
    rtb200_host_free(pinned);
    if (!quiet) {   // src/main.rs:68-71
        std::printf("Image Size: %d x %d\n", cam.image_width, cam.image_height);
        std::printf("Total Pixels: %llu\n", (unsigned long long)st.pixels_rendered);
        std::printf("Time Taken: %.3f seconds\n", st.seconds);
        std::printf("Average Pixel Rate: %.2f px/s\n", st.pixels_per_second);
    }
    if (stats_json)
        std::printf("{\"seconds\": %.6f, \"pixels\": %llu, \"paths\": %llu, \"segments\": %llu, \"sphere_tests\": %llu, \"trace_ms\": %.4f, "
                    "\"resolve_ms\": %.4f, \"d2h_ms\": %.4f, \"mpaths_per_s\": %.3f, \"shards\": %d, \"row_groups\": %d}\n",
                    st.seconds, (unsigned long long)st.pixels_rendered, (unsigned long long)st.paths, (unsigned long long)st.segments,
                    (unsigned long long)st.sphere_tests, st.trace_ms, st.resolve_ms, st.d2h_ms,
                    st.trace_ms > 0 ? (double)st.paths / (st.trace_ms * 1e-3) * 1e-6 : 0.0, multi ? (int)devices.size() : 1, row_groups);
    if (ctx) rtb200_destroy(ctx);
    if (multi) rtb200_multi_destroy(multi);
    return rc;
}
