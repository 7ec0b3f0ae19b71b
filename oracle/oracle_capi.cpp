// oracle_capi.cpp — C entry points over rt_oracle.hpp for ctypes.  TEST INFRASTRUCTURE ONLY (see rt_oracle.hpp).
// Plain arrays in, plain arrays out.  Every entry point that computes takes `use_f32`: 0 = the reference-faithful
// double instantiation; 1 = MIXED, the GPU comparand (FP32 path state, closest-hit sweep in f64 on the f64 scene,
// rt_oracle.hpp); 2 = pure FP32 (kept to measure what FP32 hit arithmetic costs against the reference).  Scene/camera
// inputs are always double (the reference's types are f64) and are cast to float exactly as the CUDA upload does
// ((float)x, round to nearest).
#include "rt_oracle.hpp"

#include <algorithm>
#include <atomic>
#include <cstring>
#include <thread>

using namespace rto;

namespace {

// Hit mode of the scenes built after the call (oracle_set_both_roots): false = the reference's packed path.
static bool g_both_roots = false;

template <class R>
Scene<R> make_scene(const double* spheres /*[n][4] cx cy cz r*/, const int32_t* sphere_material, int32_t n,
                    const int32_t* mat_kind, const int32_t* mat_hollow, const double* mat_albedo /*[m][3]*/,
                    const double* mat_param, int32_t m, bool mixed = false) {
    Scene<R> s;
    s.both_roots = g_both_roots;
    s.spheres.resize(n);
    if (mixed) {
        s.exact64.resize(n);
        for (int32_t i = 0; i < n; i++)
            s.exact64[i] = Sphere<double>{{spheres[4 * i + 0], spheres[4 * i + 1], spheres[4 * i + 2]}, spheres[4 * i + 3], 0};
    }
    for (int32_t i = 0; i < n; i++) {
        s.spheres[i].center = {(R)spheres[4 * i + 0], (R)spheres[4 * i + 1], (R)spheres[4 * i + 2]};
        s.spheres[i].radius = (R)spheres[4 * i + 3];
        s.spheres[i].material = sphere_material ? sphere_material[i] : 0;
    }
    s.materials.resize(m);
    for (int32_t j = 0; j < m; j++) {
        s.materials[j].kind = mat_kind[j];
        s.materials[j].hollow = mat_hollow[j] != 0;
        s.materials[j].albedo = {(R)mat_albedo[3 * j + 0], (R)mat_albedo[3 * j + 1], (R)mat_albedo[3 * j + 2]};
        s.materials[j].param = (R)mat_param[j];
    }
    return s;
}

CameraD make_camera(int32_t w, int32_t h, const double* cam18) {
    CameraD c;
    c.width = w; c.height = h;
    auto v = [&](int k) { return V3<double>{cam18[3 * k], cam18[3 * k + 1], cam18[3 * k + 2]}; };
    c.center = v(0); c.ulc = v(1); c.viewport_u = v(2); c.viewport_v = v(3); c.defocus_u = v(4); c.defocus_v = v(5);
    return c;
}

template <class R>
void intersect_impl(const Scene<R>& scene, const double* rays /*[n][6]*/, const int32_t* enabled, int64_t n,
                    int32_t* idx_out, double* t_out) {
    for (int64_t i = 0; i < n; i++) {
        Ray<R> r{{(R)rays[6 * i], (R)rays[6 * i + 1], (R)rays[6 * i + 2]},
                 {(R)rays[6 * i + 3], (R)rays[6 * i + 4], (R)rays[6 * i + 5]}};
        SweepHit<R> h = sweep(scene, r, enabled ? enabled[i] != 0 : true);
        idx_out[i] = h.hit ? h.index : -1;
        t_out[i] = (double)h.t;
    }
}

template <class R>
void render_impl(const Scene<R>& scene, const CameraD& camd, int32_t s0, int32_t s1, int32_t spp_total,
                 int32_t depth, uint32_t k0, uint32_t k1, int32_t row0, int32_t row1, int32_t threads,
                 double* sum_out /*[H][W][3] or null*/, uint8_t* rgb8_out /*[H][W][3] or null*/, uint64_t* segments_out) {
    Camera<R> cam = camera_cast<R>(camd);
    std::atomic<int32_t> next_row(row0);
    std::atomic<uint64_t> seg_total(0);
    if (threads < 1) threads = 1;
    auto work = [&]() {
        uint64_t seg = 0;
        for (;;) {
            int32_t row = next_row.fetch_add(1);
            if (row >= row1) break;
            for (int32_t col = 0; col < cam.width; col++) {
                V3<R> sum = pixel_sum(scene, cam, col, row, s0, s1, depth, Seed{k0, k1}, &seg);
                size_t o = ((size_t)row * cam.width + col) * 3;
                if (sum_out) { sum_out[o] = (double)sum.x; sum_out[o + 1] = (double)sum.y; sum_out[o + 2] = (double)sum.z; }
                if (rgb8_out) {
                    V3<R> px = sum / (R)spp_total;                       // src/renderer.rs:161
                    rgb8_out[o] = quantise(px.x); rgb8_out[o + 1] = quantise(px.y); rgb8_out[o + 2] = quantise(px.z);
                }
            }
        }
        seg_total += seg;
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; t++) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    if (segments_out) *segments_out = seg_total.load();
}

}  // namespace

extern "C" {

// 0 (default): near root only, as the reference's live renderer; 1: the scalar Sphere::hit rule (rt_oracle.hpp sweep).
// Process-wide; the test harness sets it around a call.
void oracle_set_both_roots(int32_t on) { g_both_roots = on != 0; }

void oracle_philox4x32_10(const uint32_t* ctr4, const uint32_t* key2, uint32_t* out4) {
    Philox4 p = philox4x32_10(ctr4[0], ctr4[1], ctr4[2], ctr4[3], key2[0], key2[1]);
    std::memcpy(out4, p.v, sizeof(p.v));
}

void oracle_cossin_2pi(double v, int32_t use_f32, double* cs, double* sn) {
    if (use_f32) { float c, s; cossin_2pi<float>((float)v, c, s); *cs = c; *sn = s; }
    else cossin_2pi<double>(v, *cs, *sn);
}

void oracle_unit_vector_from(double u, double v, int32_t use_f32, double* out3) {
    if (use_f32) { V3<float> r = unit_vector_from<float>((float)u, (float)v); out3[0] = r.x; out3[1] = r.y; out3[2] = r.z; }
    else { V3<double> r = unit_vector_from<double>(u, v); out3[0] = r.x; out3[1] = r.y; out3[2] = r.z; }
}

// Camera::new (src/ray_tracing.rs:27-62).  out18 = center, ulc, viewport_u, viewport_v, defocus_u, defocus_v.
void oracle_camera_new(int32_t width, int32_t height, double focal_length, double view_angle_deg, const double* center,
                       const double* look_at, const double* up, double defocus_angle_deg, double* out18) {
    CameraD c = camera_new(width, height, focal_length, view_angle_deg, {center[0], center[1], center[2]},
                           {look_at[0], look_at[1], look_at[2]}, {up[0], up[1], up[2]}, defocus_angle_deg);
    const V3<double> v[6] = {c.center, c.ulc, c.viewport_u, c.viewport_v, c.defocus_u, c.defocus_v};
    for (int k = 0; k < 6; k++) { out18[3 * k] = v[k].x; out18[3 * k + 1] = v[k].y; out18[3 * k + 2] = v[k].z; }
}

// Camera::get_ray with explicit jitter words (rnd4 = the Philox block).  out6 = origin, direction.
void oracle_camera_get_ray(int32_t width, int32_t height, const double* cam18, int32_t col, int32_t row,
                           const uint32_t* rnd4, int32_t use_f32, double* out6) {
    CameraD cd = make_camera(width, height, cam18);
    Philox4 p{{rnd4[0], rnd4[1], rnd4[2], rnd4[3]}};
    if (use_f32) {
        Ray<float> r = camera_get_ray(camera_cast<float>(cd), col, row, p);
        out6[0] = r.o.x; out6[1] = r.o.y; out6[2] = r.o.z; out6[3] = r.d.x; out6[4] = r.d.y; out6[5] = r.d.z;
    } else {
        Ray<double> r = camera_get_ray(camera_cast<double>(cd), col, row, p);
        out6[0] = r.o.x; out6[1] = r.o.y; out6[2] = r.o.z; out6[3] = r.d.x; out6[4] = r.d.y; out6[5] = r.d.z;
    }
}

// Closest-hit sweep for a batch of rays.  idx = -1 on miss; t = +inf on miss.
void oracle_intersect(const double* spheres, int32_t n_spheres, const double* rays, const int32_t* enabled, int64_t n_rays,
                      int32_t use_f32, int32_t* idx_out, double* t_out) {
    const int32_t kind = 0, hollow = 0; const double alb[3] = {0, 0, 0}, par = 0;
    if (use_f32) intersect_impl(make_scene<float>(spheres, nullptr, n_spheres, &kind, &hollow, alb, &par, 1, use_f32 == 1), rays, enabled, n_rays, idx_out, t_out);
    else intersect_impl(make_scene<double>(spheres, nullptr, n_spheres, &kind, &hollow, alb, &par, 1), rays, enabled, n_rays, idx_out, t_out);
}

// One sphere, one ray: sweep + finalize + scatter with an explicit Philox block.
// out: [0]=hit(0/1) [1]=t [2..4]=p [5..7]=n [8]=front [9..11]=attenuation [12..14]=scattered origin [15..17]=scattered dir
void oracle_hit_and_scatter(const double* sphere4, int32_t mat_kind, int32_t mat_hollow, const double* albedo3, double param,
                            const double* ray6, const uint32_t* rnd4, int32_t use_f32, double* out18) {
    const int32_t mi = 0;
    auto run = [&](auto tag) {
        using R = decltype(tag);
        Scene<R> sc = make_scene<R>(sphere4, &mi, 1, &mat_kind, &mat_hollow, albedo3, &param, 1, sizeof(R) == 4 && use_f32 == 1);
        Ray<R> r{{(R)ray6[0], (R)ray6[1], (R)ray6[2]}, {(R)ray6[3], (R)ray6[4], (R)ray6[5]}};
        SweepHit<R> h = sweep(sc, r);
        for (int i = 0; i < 18; i++) out18[i] = 0.0;
        out18[0] = h.hit ? 1.0 : 0.0; out18[1] = (double)h.t;
        if (!h.hit) return;
        HitRecord<R> rec = finalize(sc, r, h);
        out18[2] = rec.p.x; out18[3] = rec.p.y; out18[4] = rec.p.z;
        out18[5] = rec.n.x; out18[6] = rec.n.y; out18[7] = rec.n.z;
        out18[8] = rec.front ? 1.0 : 0.0;
        V3<R> att; Ray<R> nx;
        scatter(sc.materials[0], r, rec, Philox4{{rnd4[0], rnd4[1], rnd4[2], rnd4[3]}}, att, nx);
        out18[9] = att.x; out18[10] = att.y; out18[11] = att.z;
        out18[12] = nx.o.x; out18[13] = nx.o.y; out18[14] = nx.o.z;
        out18[15] = nx.d.x; out18[16] = nx.d.y; out18[17] = nx.d.z;
    };
    if (use_f32) run(float{}); else run(double{});
}

void oracle_sky(const double* d3, int32_t use_f32, double* out3) {
    if (use_f32) { V3<float> s = sky<float>({(float)d3[0], (float)d3[1], (float)d3[2]}); out3[0] = s.x; out3[1] = s.y; out3[2] = s.z; }
    else { V3<double> s = sky<double>({d3[0], d3[1], d3[2]}); out3[0] = s.x; out3[1] = s.y; out3[2] = s.z; }
}

int32_t oracle_quantise(double c, int32_t use_f32) { return use_f32 ? quantise<float>((float)c) : quantise<double>(c); }

double oracle_reflectance(double cosine, double ior, int32_t use_f32) {
    return use_f32 ? (double)reflectance<float>((float)cosine, (float)ior) : reflectance<double>(cosine, ior);
}

// Full render of rows [row0,row1), samples [s0,s1) of every pixel, per-path semantics (v1/v3).
// sum_out: linear RGB SUM over the sample range (not the mean); rgb8_out: quantised sum/spp_total.
void oracle_render(const double* spheres, const int32_t* sphere_material, int32_t n_spheres, const int32_t* mat_kind,
                   const int32_t* mat_hollow, const double* mat_albedo, const double* mat_param, int32_t n_materials,
                   int32_t width, int32_t height, const double* cam18, int32_t s0, int32_t s1, int32_t spp_total, int32_t depth,
                   uint32_t seed_lo, uint32_t seed_hi, int32_t row0, int32_t row1, int32_t threads, int32_t use_f32,
                   double* sum_out, uint8_t* rgb8_out, uint64_t* segments_out) {
    CameraD cd = make_camera(width, height, cam18);
    if (use_f32)
        render_impl(make_scene<float>(spheres, sphere_material, n_spheres, mat_kind, mat_hollow, mat_albedo, mat_param, n_materials, use_f32 == 1),
                    cd, s0, s1, spp_total, depth, seed_lo, seed_hi, row0, row1, threads, sum_out, rgb8_out, segments_out);
    else
        render_impl(make_scene<double>(spheres, sphere_material, n_spheres, mat_kind, mat_hollow, mat_albedo, mat_param, n_materials),
                    cd, s0, s1, spp_total, depth, seed_lo, seed_hi, row0, row1, threads, sum_out, rgb8_out, segments_out);
}

}  // extern "C"
