// cpu_ref.cpp — CPU port of the reference's LIVE render path, for the reported CPU baseline and for quantifying the
// lane-bookkeeping defects of trace_vectorized2.  TEST / BENCH INFRASTRUCTURE ONLY (see rt_oracle.hpp header).
//
// What is restated, statement for statement (paths relative to the reference root):
//   TileRenderer::render          src/renderer.rs:244-387   128x128 tiles, dynamic queue, one worker per hardware thread
//   render_vectorized2            src/renderer.rs:141-176   per pixel: spp camera rays -> chunks of 4 -> trace -> /spp -> u8
//   Scene::trace_vectorized2      src/ray_tracing.rs:375-505 two ray buffers, copy-compaction, INCLUDING its defects
//                                 (sky from the primary ray directions, final buffer picked by chunk-count
//                                 parity, stale chunks; F11: padding lanes turn sky-coloured)
//   Sphere::hit_packed            src/objects.rs:249-290    4 x f64 AVX2 lanes, near root only
//   PackedHitRecords              src/objects.rs:109-201
//   materials                     src/materials.rs:53-147   scalar, per lane, dynamic dispatch
// The reference's lanes are std::simd f64x4 (src/renderer.rs:142) compiled with -Ctarget-cpu=native
// (.cargo/config.toml:3); here they are __m256d intrinsics compiled with -O3 -march=native.
// Randomness: the reference uses rand::thread_rng (ChaCha12) + rand_distr::Normal; this port uses a per-thread
// xoshiro256++ stream and the same distributions (uniform jitter, rejection-sampled unit disk, normalised Gaussian
// triple for the unit vector), so the work per event matches; the random numbers themselves cannot.
#include <immintrin.h>

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <thread>
#include <vector>

namespace cpuref {

// ---- scalar Vec3 / Color (src/geometry.rs:9-188, src/color.rs:8-165) ---------------------------------------
struct V3 { double x, y, z; };
static inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
static inline V3 operator*(double s, V3 a) { return a * s; }
static inline V3 operator/(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
static inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
static inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline double length_squared(V3 a) { return dot(a, a); }
static inline V3 unit(V3 a) { return a / std::sqrt(length_squared(a)); }
static inline bool near_zero(V3 a) { return std::fabs(a.x) < 1e-8 && std::fabs(a.y) < 1e-8 && std::fabs(a.z) < 1e-8; }
static inline V3 reflect(V3 v, V3 n) { return v - (n * (2.0 * dot(v, n))); }
static inline V3 refract(V3 v, V3 n, double ratio) {
    double cos_theta = std::fmin(dot(-v, n), 1.0);
    V3 perp = (v + (n * cos_theta)) * ratio;
    V3 par = n * (-(std::sqrt(std::fabs(1.0 - length_squared(perp)))));
    return perp + par;
}
struct Ray { V3 o, d; };

// ---- per-thread generator (stands in for thread_rng) ------------------------------------------------------
struct Rng {
    uint64_t s[4];
    explicit Rng(uint64_t seed) {
        for (int i = 0; i < 4; i++) { seed += 0x9E3779B97F4A7C15ull; uint64_t z = seed; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; s[i] = z ^ (z >> 31); }
    }
    static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    inline uint64_t next() {
        uint64_t r = rotl(s[0] + s[3], 23) + s[0], t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
        return r;
    }
    inline double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }   // gen_range(0.0..1.0)
    inline double normal() {                                                                  // rand_distr::Normal(0,1)
        for (;;) {                                                                            // Marsaglia polar method
            double u = 2.0 * uniform() - 1.0, v = 2.0 * uniform() - 1.0, q = u * u + v * v;
            if (q > 0.0 && q < 1.0) return u * std::sqrt(-2.0 * std::log(q) / q);
        }
    }
};
static inline V3 random_unit_vector(Rng& g) {                       // src/geometry.rs:139-152
    V3 v{g.normal(), g.normal(), g.normal()};
    if (near_zero(v)) return {0.0, 0.0, 1.0};
    return unit(v);
}
static inline V3 random_in_unit_disk(Rng& g) {                      // src/geometry.rs:154-168
    for (;;) {
        double x = 2.0 * g.uniform() - 1.0, y = 2.0 * g.uniform() - 1.0;
        if (x * x + y * y <= 1.0) return {x, y, 0.0};
    }
}

// ---- materials: trait object + three impls (src/materials.rs:35-147) ---------------------------------------
struct HitRecord { V3 location, normal; double t; bool front_face; };
struct HitResult { V3 attenuation; Ray scattered; };
struct Material { virtual ~Material() = default; virtual HitResult get_hit_result(const Ray& ray, const HitRecord& rec, Rng& g) const = 0; };
struct Lambertian : Material {
    V3 albedo;
    explicit Lambertian(V3 a) : albedo(a) {}
    HitResult get_hit_result(const Ray&, const HitRecord& rec, Rng& g) const override {
        V3 dir = random_unit_vector(g) + rec.normal;
        if (near_zero(dir)) dir = rec.normal;
        return {albedo, {rec.location, dir}};
    }
};
struct Metal : Material {
    V3 albedo; double fuzz;
    Metal(V3 a, double f) : albedo(a), fuzz(f < 1.0 ? f : 1.0) {}
    HitResult get_hit_result(const Ray& ray, const HitRecord& rec, Rng& g) const override {
        V3 reflected = reflect(ray.d, rec.normal) + (fuzz * random_unit_vector(g));
        return {albedo, {rec.location, reflected}};
    }
};
struct Dielectric : Material {
    double ior; bool hollow;
    Dielectric(double i, bool h) : ior(i), hollow(h) {}
    static double reflectance(double cosine, double r) {
        double q = (1.0 - r) / (1.0 + r), r0 = q * q, x = 1.0 - cosine, x2 = x * x;
        return r0 + (1.0 - r0) * (x * (x2 * x2));
    }
    HitResult get_hit_result(const Ray& ray, const HitRecord& rec, Rng& g) const override {
        double ratio = rec.front_face ? 1.0 / ior : ior;
        V3 n = hollow ? -rec.normal : rec.normal;
        double cos_theta = std::fmin(dot(-ray.d, n), 1.0);
        double sin_theta = std::sqrt(1.0 - cos_theta * cos_theta);
        bool cannot_refract = ratio * sin_theta > 1.0;
        V3 dir = (cannot_refract || reflectance(cos_theta, ratio) > g.uniform()) ? reflect(ray.d, n) : refract(ray.d, n, ratio);
        return {{1.0, 1.0, 1.0}, {rec.location, dir}};
    }
};

struct Sphere { V3 center; double radius; const Material* material; };
struct Scene { std::vector<Sphere> objects; std::vector<std::unique_ptr<Material>> materials; };
struct Camera { int width, height; V3 center, ulc, viewport_u, viewport_v, defocus_u, defocus_v; };

static inline Ray camera_get_ray(const Camera& c, int col, int row, Rng& g) {   // src/ray_tracing.rs:77-89
    double x_offset = g.uniform(), y_offset = g.uniform();
    V3 pixel_offset = c.viewport_u * (((double)col + x_offset) / (double)c.width) + c.viewport_v * (((double)row + y_offset) / (double)c.height);
    V3 pixel_center = c.ulc + pixel_offset;
    V3 disk = random_in_unit_disk(g);                                           // always drawn (:82)
    V3 origin = c.defocus_u * disk.x + c.defocus_v * disk.y + c.center;
    return {origin, unit(pixel_center - origin)};
}

// ---- 4 x f64 lanes (std::simd f64x4 in the reference) -------------------------------------------------------
typedef __m256d F4;
static inline F4 splat(double v) { return _mm256_set1_pd(v); }
static inline double lane(F4 v, int i) { alignas(32) double t[4]; _mm256_store_pd(t, v); return t[i]; }
static inline F4 set_lane(F4 v, int i, double x) { alignas(32) double t[4]; _mm256_store_pd(t, v); t[i] = x; return _mm256_load_pd(t); }
static inline F4 mask_true() { return _mm256_castsi256_pd(_mm256_set1_epi64x(-1)); }
static inline F4 mask_false() { return _mm256_setzero_pd(); }
static inline bool mask_test(F4 m, int i) { return (_mm256_movemask_pd(m) >> i) & 1; }
static inline F4 mask_set(F4 m, int i, bool v) { alignas(32) int64_t t[4]; _mm256_store_si256((__m256i*)t, _mm256_castpd_si256(m)); t[i] = v ? -1 : 0; return _mm256_castsi256_pd(_mm256_load_si256((const __m256i*)t)); }
static inline F4 select(F4 base, F4 other, F4 mask) { return _mm256_blendv_pd(base, other, mask); }     // masked_select, src/simd_util.rs:4-37

struct PV3 { F4 x, y, z; };                                                                              // PackedVec3, src/geometry.rs:217-226
static inline PV3 pv_splat(V3 v) { return {splat(v.x), splat(v.y), splat(v.z)}; }
static inline PV3 operator-(PV3 a, PV3 b) { return {_mm256_sub_pd(a.x, b.x), _mm256_sub_pd(a.y, b.y), _mm256_sub_pd(a.z, b.z)}; }
static inline PV3 operator+(PV3 a, PV3 b) { return {_mm256_add_pd(a.x, b.x), _mm256_add_pd(a.y, b.y), _mm256_add_pd(a.z, b.z)}; }
static inline PV3 operator*(PV3 a, F4 s) { return {_mm256_mul_pd(a.x, s), _mm256_mul_pd(a.y, s), _mm256_mul_pd(a.z, s)}; }
static inline PV3 operator/(PV3 a, F4 s) { return {_mm256_div_pd(a.x, s), _mm256_div_pd(a.y, s), _mm256_div_pd(a.z, s)}; }
static inline F4 pv_dot(PV3 a, PV3 b) { return _mm256_fmadd_pd(a.z, b.z, _mm256_fmadd_pd(a.y, b.y, _mm256_mul_pd(a.x, b.x))); }   // :466-468
static inline F4 pv_length_squared(PV3 a) { return pv_dot(a, a); }                                                                // :434-436
static inline V3 pv_at(PV3 a, int i) { return {lane(a.x, i), lane(a.y, i), lane(a.z, i)}; }
static inline void pv_update(PV3& a, int i, V3 v) { a.x = set_lane(a.x, i, v.x); a.y = set_lane(a.y, i, v.y); a.z = set_lane(a.z, i, v.z); }
static inline void pv_assign_masked(PV3& a, PV3 v, F4 m) { a.x = select(a.x, v.x, m); a.y = select(a.y, v.y, m); a.z = select(a.z, v.z, m); }

struct PackedRays {                                                                                      // src/ray.rs:34-153
    PV3 origins, directions; F4 enabled;
    static PackedRays zeros_enabled() { return {pv_splat({0, 0, 0}), pv_splat({0, 0, 0}), mask_true()}; }  // PackedRays::new(default, default)
    static PackedRays from_rays(const Ray* r, int n) {                                                   // FromIterator, :134-152
        PackedRays p{pv_splat({0, 0, 0}), pv_splat({0, 0, 0}), mask_false()};
        for (int i = 0; i < n; i++) p.update(i, r[i]);
        return p;
    }
    bool is_enabled(int i) const { return mask_test(enabled, i); }
    Ray at_including_disabled(int i) const { return {pv_at(origins, i), pv_at(directions, i)}; }
    void update(int i, const Ray& r) { pv_update(origins, i, r.o); pv_update(directions, i, r.d); enabled = mask_set(enabled, i, true); }
    void update_with_enable(int i, const Ray& r, bool e) { pv_update(origins, i, r.o); pv_update(directions, i, r.d); enabled = mask_set(enabled, i, e); }
    void disable(int i) { enabled = mask_set(enabled, i, false); }
    PV3 at_t(F4 t) const { return origins + directions * t; }
};

struct PackedColor { F4 r, g, b; };                                                                      // src/color.rs:174-296
static inline PackedColor pc_splat(double r, double g, double b) { return {splat(r), splat(g), splat(b)}; }

struct PackedHitRecords {                                                                                // src/objects.rs:109-201
    PV3 locations, normals; F4 t, front_face, hits; __m256i materials;
    PackedHitRecords() : locations(pv_splat({0, 0, 0})), normals(pv_splat({0, 0, 0})), t(splat(std::numeric_limits<double>::infinity())),
                         front_face(mask_false()), hits(mask_false()), materials(_mm256_setzero_si256()) {}
    inline void update(PV3 outward_normals, F4 tt, F4 valid, const Material* m) {                        // :140-155
        F4 update_mask = _mm256_and_pd(valid, _mm256_cmp_pd(tt, t, _CMP_LE_OQ));
        pv_assign_masked(normals, outward_normals, update_mask);
        t = select(t, tt, update_mask);
        hits = _mm256_or_pd(hits, update_mask);
        materials = _mm256_blendv_epi8(materials, _mm256_set1_epi64x((long long)(uintptr_t)m), _mm256_castpd_si256(update_mask));
    }
    inline void finalize(const PackedRays& rays) {                                                       // :157-162
        normals = normals / _mm256_sqrt_pd(pv_length_squared(normals));
        locations = rays.at_t(t);
        front_face = _mm256_cmp_pd(pv_dot(rays.directions, normals), splat(0.0), _CMP_LT_OQ);
        PV3 neg{_mm256_xor_pd(normals.x, splat(-0.0)), _mm256_xor_pd(normals.y, splat(-0.0)), _mm256_xor_pd(normals.z, splat(-0.0))};
        pv_assign_masked(normals, neg, _mm256_xor_pd(front_face, mask_true()));
    }
    const Material* material_at(int i) const { alignas(32) int64_t tmp[4]; _mm256_store_si256((__m256i*)tmp, materials); return (const Material*)(uintptr_t)tmp[i]; }
};

// Sphere::hit_packed, src/objects.rs:249-290 (t_range = 0.001..inf).  Returns 4 lane tests performed.
static inline void hit_packed(const Sphere& s, const PackedRays& rays, PackedHitRecords& rec) {
    PV3 center_offset = rays.origins - pv_splat(s.center);                                               // :252
    F4 a = pv_length_squared(rays.directions);                                                           // :253
    F4 inverse_a = _mm256_div_pd(splat(1.0), a);                                                         // :254
    F4 half_b = pv_dot(center_offset, rays.directions);                                                  // :255
    F4 c = _mm256_sub_pd(pv_length_squared(center_offset), splat(s.radius * s.radius));                  // :256
    F4 discriminant = _mm256_fmadd_pd(half_b, half_b, _mm256_xor_pd(_mm256_mul_pd(a, c), splat(-0.0)));  // :257
    F4 discriminant_positive = _mm256_and_pd(_mm256_cmp_pd(discriminant, splat(0.0), _CMP_GE_OQ), rays.enabled);   // :259
    if (_mm256_movemask_pd(discriminant_positive)) {                                                     // :261
        F4 neg_half_b = _mm256_xor_pd(half_b, splat(-0.0));                                              // negate_simd_float, src/simd_util.rs:173-222
        F4 sqrt_discriminant = _mm256_sqrt_pd(discriminant);
        F4 root1 = _mm256_mul_pd(_mm256_sub_pd(neg_half_b, sqrt_discriminant), inverse_a);               // :270
        F4 root2 = _mm256_mul_pd(_mm256_add_pd(neg_half_b, sqrt_discriminant), inverse_a);               // :271
        F4 inf = splat(std::numeric_limits<double>::infinity());
        F4 root1_valid = _mm256_and_pd(_mm256_cmp_pd(root1, splat(0.001), _CMP_GE_OQ), _mm256_cmp_pd(root1, inf, _CMP_LT_OQ));   // simd_inside
        F4 root2_valid = root1_valid;                                                                    // :274 (computed from root1: SURVEY F4)
        F4 root = select(root2, root1, root1_valid);                                                     // :276
        F4 valid = _mm256_and_pd(_mm256_or_pd(root1_valid, root2_valid), rays.enabled);                  // :278
        PV3 normal = rays.at_t(root) - pv_splat(s.center);                                               // :280-281
        rec.update(normal, root, valid, s.material);
    }
}

struct Counters { uint64_t lane_tests = 0, segments = 0, paths = 0; };

// Scene::trace_vectorized2, src/ray_tracing.rs:375-505 — restated with its defects.
static V3 trace_vectorized2(const Scene& scene, const std::vector<PackedRays>& rays, int depth_limit, Rng& g, Counters& cnt) {
    const size_t len = rays.size();
    std::vector<PackedRays> ray_buffers[2] = {rays, std::vector<PackedRays>(len, PackedRays::zeros_enabled())};      // :382
    std::vector<PackedColor> color[2] = {std::vector<PackedColor>(len, pc_splat(1, 1, 1)), std::vector<PackedColor>(len, pc_splat(1, 1, 1))};
    std::vector<F4> hit_sky[2] = {std::vector<F4>(len, mask_false()), std::vector<F4>(len, mask_false())};
    size_t last_active_chunk = len;                                                                      // :386
    for (int k = 0; k < depth_limit; k++) {
        if (last_active_chunk == 0) break;                                                               // :389-392
        const int sel = k % 2;
        for (size_t j = 0; j < last_active_chunk; j++) {
            PackedRays& pr = ray_buffers[sel][j];
            PackedHitRecords rec;
            for (const Sphere& s : scene.objects) hit_packed(s, pr, rec);                                // :399-401
            int en = __builtin_popcount(_mm256_movemask_pd(pr.enabled));
            cnt.segments += en; cnt.lane_tests += (uint64_t)en * scene.objects.size();
            rec.finalize(pr);                                                                            // :403
            for (int i = 0; i < 4; i++) {                                                                // :406-426
                if (mask_test(rec.hits, i)) {
                    HitRecord hr{pv_at(rec.locations, i), pv_at(rec.normals, i), lane(rec.t, i), mask_test(rec.front_face, i)};
                    Ray ray = pr.at_including_disabled(i);
                    HitResult res = rec.material_at(i)->get_hit_result(ray, hr, g);
                    color[sel][j].r = set_lane(color[sel][j].r, i, lane(color[sel][j].r, i) * res.attenuation.x);
                    color[sel][j].g = set_lane(color[sel][j].g, i, lane(color[sel][j].g, i) * res.attenuation.y);
                    color[sel][j].b = set_lane(color[sel][j].b, i, lane(color[sel][j].b, i) * res.attenuation.z);
                    pr.update(i, res.scattered);
                } else {
                    pr.disable(i);
                    hit_sky[sel][j] = mask_set(hit_sky[sel][j], i, true);
                }
            }
        }
        // shuffle (:429-483): enabled lanes to the front of the other buffer, then the disabled ones
        size_t out_chunk = 0; int out_slot = 0;
        const int nsel = 1 - sel;
        auto copy_lane = [&](size_t i, int j, bool enable) {
            ray_buffers[nsel][out_chunk].update_with_enable(out_slot, ray_buffers[sel][i].at_including_disabled(j), enable);
            color[nsel][out_chunk].r = set_lane(color[nsel][out_chunk].r, out_slot, lane(color[sel][i].r, j));
            color[nsel][out_chunk].g = set_lane(color[nsel][out_chunk].g, out_slot, lane(color[sel][i].g, j));
            color[nsel][out_chunk].b = set_lane(color[nsel][out_chunk].b, out_slot, lane(color[sel][i].b, j));
            hit_sky[nsel][out_chunk] = mask_set(hit_sky[nsel][out_chunk], out_slot, mask_test(hit_sky[sel][i], j));
            if (++out_slot >= 4) { out_slot = 0; out_chunk++; }
        };
        for (size_t i = 0; i < last_active_chunk; i++)
            for (int j = 0; j < 4; j++)
                if (ray_buffers[sel][i].is_enabled(j)) copy_lane(i, j, true);
        const size_t new_last = out_slot == 0 ? out_chunk : out_chunk + 1;                               // :463
        for (size_t i = 0; i < last_active_chunk; i++)
            for (int j = 0; j < 4; j++)
                if (!ray_buffers[sel][i].is_enabled(j)) copy_lane(i, j, false);
        last_active_chunk = new_last;
    }
    const int sel = len ? (int)((len - 1) % 2) : 0;                                                      // :486 (chunk-count parity, F7b)
    F4 sr = splat(0.0), sg = splat(0.0), sb = splat(0.0);
    for (size_t j = 0; j < len; j++) {
        F4 a = _mm256_mul_pd(_mm256_add_pd(rays[j].directions.y, splat(1.0)), splat(0.5));               // :490 — the PRIMARY rays (F7a)
        F4 w = _mm256_add_pd(_mm256_xor_pd(a, splat(-0.0)), splat(1.0));
        F4 kr = _mm256_add_pd(w, _mm256_mul_pd(splat(0.5), a)), kg = _mm256_add_pd(w, _mm256_mul_pd(splat(0.7), a)), kb = _mm256_add_pd(w, _mm256_mul_pd(splat(1.0), a));
        PackedColor& c = color[sel][j];
        c.r = select(c.r, _mm256_mul_pd(c.r, kr), hit_sky[sel][j]);
        c.g = select(c.g, _mm256_mul_pd(c.g, kg), hit_sky[sel][j]);
        c.b = select(c.b, _mm256_mul_pd(c.b, kb), hit_sky[sel][j]);
        F4 en = ray_buffers[sel][j].enabled;                                                             // :496 still enabled => black
        c.r = select(c.r, splat(0.0), en); c.g = select(c.g, splat(0.0), en); c.b = select(c.b, splat(0.0), en);
        sr = _mm256_add_pd(sr, c.r); sg = _mm256_add_pd(sg, c.g); sb = _mm256_add_pd(sb, c.b);           // :499-502
    }
    auto hsum = [](F4 v) { return ((lane(v, 0) + lane(v, 1)) + lane(v, 2)) + lane(v, 3); };
    return {hsum(sr), hsum(sg), hsum(sb)};
}

static inline uint8_t to_u8(double c) {                                                                  // src/color.rs:54-64
    double v = std::sqrt(c) * 255.999;
    if (!(v >= 0.0)) return 0;
    if (v >= 255.0) return 255;
    return (uint8_t)(int)v;
}

}  // namespace cpuref

extern "C" {

// TileRenderer::render (src/renderer.rs:244-387) over rows [row0,row1): 128x128 tiles from a shared queue, `threads`
// workers, render_vectorized2 per tile.  Returns wall seconds; counts lane tests (enabled lanes x spheres) and segments.
double cpuref_render(const double* spheres, const int32_t* sphere_material, int32_t n_spheres, const int32_t* mat_kind,
                     const int32_t* mat_hollow, const double* mat_albedo, const double* mat_param, int32_t n_materials, int32_t width,
                     int32_t height, const double* cam18, int32_t spp, int32_t depth, int32_t threads, int32_t block_size, uint64_t seed,
                     int32_t row0, int32_t row1, uint8_t* rgb8_out, uint64_t* lane_tests, uint64_t* segments) {
    using namespace cpuref;
    Scene scene;
    for (int j = 0; j < n_materials; j++) {
        V3 al{mat_albedo[3 * j], mat_albedo[3 * j + 1], mat_albedo[3 * j + 2]};
        if (mat_kind[j] == 0) scene.materials.emplace_back(new Lambertian(al));
        else if (mat_kind[j] == 1) scene.materials.emplace_back(new Metal(al, mat_param[j]));
        else scene.materials.emplace_back(new Dielectric(mat_param[j], mat_hollow[j] != 0));
    }
    for (int i = 0; i < n_spheres; i++)
        scene.objects.push_back({{spheres[4 * i], spheres[4 * i + 1], spheres[4 * i + 2]}, spheres[4 * i + 3], scene.materials[sphere_material[i]].get()});
    auto v = [&](int k) { return V3{cam18[3 * k], cam18[3 * k + 1], cam18[3 * k + 2]}; };
    Camera cam{width, height, v(0), v(1), v(2), v(3), v(4), v(5)};
    if (threads < 1) threads = (int)std::thread::hardware_concurrency();
    if (block_size < 1) block_size = 128;
    const int rows = row1 - row0;
    const int wb = (width + block_size - 1) / block_size, hb = (rows + block_size - 1) / block_size;     // :248-249
    std::atomic<int> next_tile(0);
    std::atomic<uint64_t> tot_tests(0), tot_seg(0);
    auto t0 = std::chrono::steady_clock::now();
    auto worker = [&](int tid) {
        Rng g(seed * 0x9E3779B97F4A7C15ull + (uint64_t)tid + 1);
        Counters cnt;
        std::vector<Ray> prim((size_t)spp);
        for (;;) {
            int tile = next_tile.fetch_add(1);                                                           // task_rx.recv(), :280
            if (tile >= wb * hb) break;
            int bx = tile % wb, by = tile / wb;
            int size_x = std::min(block_size, width - bx * block_size), size_y = std::min(block_size, rows - by * block_size);
            for (int j = 0; j < size_y; j++) {
                for (int i = 0; i < size_x; i++) {                                                       // render_vectorized2, :149-164
                    int col = i + bx * block_size, row = row0 + j + by * block_size;
                    for (int s = 0; s < spp; s++) prim[s] = camera_get_ray(cam, col, row, g);
                    std::vector<PackedRays> chunks;
                    for (int s = 0; s < spp; s += 4) chunks.push_back(PackedRays::from_rays(&prim[s], std::min(4, spp - s)));
                    V3 sum = trace_vectorized2(scene, chunks, depth, g, cnt);
                    cnt.paths += spp;
                    if (rgb8_out) {
                        size_t o = ((size_t)row * width + col) * 3;
                        rgb8_out[o] = to_u8(sum.x / (double)spp); rgb8_out[o + 1] = to_u8(sum.y / (double)spp); rgb8_out[o + 2] = to_u8(sum.z / (double)spp);
                    }
                }
            }
        }
        tot_tests += cnt.lane_tests; tot_seg += cnt.segments;
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; t++) pool.emplace_back(worker, t);
    worker(0);
    for (auto& t : pool) t.join();
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (lane_tests) *lane_tests = tot_tests.load();
    if (segments) *segments = tot_seg.load();
    return sec;
}

}  // extern "C"
