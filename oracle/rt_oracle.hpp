// rt_oracle.hpp — CPU restatement of the reference's render hot path.  TEST INFRASTRUCTURE ONLY.
//
// This header is the parity oracle for the rtb200 CUDA path.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may build, link or call anything under oracle/.
// The product path (rust-ray-tracing_b200/) never includes this file.
//
// PARITY STATUS: "parity unpinned" at the reference boundary.  The reference
// (semicolonTransistor/rust-ray-tracing) ships no tests, golden vectors or images, cannot be compiled
// here (nightly Rust + portable_simd, no cargo/rustc in this image) and draws all randomness from an
// OS-seeded thread_rng (src/ray_tracing.rs:78-79, src/geometry.rs:139-168, src/materials.rs:137).  What
// pins this oracle is (1) hand-derived known-answer vectors for each cited routine (tests/test_oracle_kat.py,
// SURVEY.md §8c), (2) the Random123 known-answer vectors for Philox4x32-10, and (3) the committed fixtures under
// tests/golden/ (written from this oracle by tests/golden/make_golden.py; they freeze it against silent drift and travel to
// the GPU box — they are NOT reference output, there is none).
//
// Every function cites the reference file:line it restates.  The code is templated on the scalar type:
//   Real = double : reference-faithful arithmetic (the reference is f64 throughout, src/geometry.rs:9-13)
//   Real = float  : same operation order, same FMA placement, IEEE sqrt/div, in FP32 (the north_star's GPU arithmetic).
//   Real = float with Scene::exact64 filled ("mixed", the GPU COMPARAND): the path state (rays, hit points, normals,
//                   materials, colours) is FP32, but the closest-hit sweep — Sphere::hit_packed + PackedHitRecords::update —
//                   runs in f64 on the f64 scene from the FP32-valued ray, and hands back the f64 root rounded once to
//                   FP32.  This is what the CUDA kernels do (rt_kernels.cuh test_candidate): it keeps the intersection
//                   inside the north_star's 1e-5 bar against the reference's f64 routine, which pure FP32 misses on the
//                   1000-unit ground sphere (cancellation in the discriminant).
// Compile with -ffp-contract=off so that the ONLY fused multiply-adds are the explicit std::fma calls,
// which sit exactly where the reference calls mul_add (src/geometry.rs:435,467; src/objects.rs:257).
//
// Randomness: the reference has no deterministic mode.  The build defines one: a counter-based
// Philox4x32-10 stream keyed by (seed) and indexed by (pixel, sample, dimension), consumed identically by
// this oracle and by the CUDA kernels.  Distributions match the reference's (uniform pixel jitter, uniform
// unit disk, uniform unit sphere, uniform Fresnel draw); the mapping from uniforms uses only + - * / sqrt fma
// so that CPU and GPU agree bit-for-bit in float.
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>
#include <vector>

namespace rto {

// ----------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11; Random123).
// Stands in for rand::thread_rng (not reproducible in the reference).  Known-answer vectors are the
// Random123 kat_vectors entries for philox4x32 10 rounds.
// ----------------------------------------------------------------------------------------------
struct Philox4 { uint32_t v[4]; };

inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)M0 * c0;
        uint64_t p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return Philox4{{c0, c1, c2, c3}};
}

// Stream dimensions (third counter word).  0 = camera sample, 1+k = scatter event of bounce k.
constexpr uint32_t DIM_CAMERA = 0;

// 24-bit uniform in [0,1): exactly representable in float, so float and double oracles see the same value.
// Replaces gen_range(0.0..1.0) (src/ray_tracing.rs:78-79, src/materials.rs:137).
template <class R> inline R u01(uint32_t x) { return (R)(x >> 8) * (R)(1.0 / 16777216.0); }

// cos(2*pi*v), sin(2*pi*v) for v in [0,1) from + * fma only (bit-reproducible on CPU and GPU).
// Quadrant reduction is exact (4v and 4v-q are exact); the quarter-wave polynomials are Taylor series in
// y = f*pi/2 (|error| < 1e-7), evaluated by Horner with explicit fma.  Coefficients are float literals so the
// float and double instantiations use the same real numbers.
template <class R> inline void cossin_2pi(R v, R& cs, R& sn) {
    R t = v * (R)4.0f;
    int q = (int)t;             // 0..3
    R f = t - (R)q;             // [0,1)
    R y = f * (R)1.57079637f;   // float(pi/2)
    R y2 = y * y;
    // sin(y) = y * (1 + y2*(s3 + y2*(s5 + y2*(s7 + y2*(s9 + y2*s11)))))
    R ps = (R)-2.50521084e-08f;                       // -1/11!
    ps = std::fma(ps, y2, (R)2.75573192e-06f);        //  1/9!
    ps = std::fma(ps, y2, (R)-1.98412701e-04f);       // -1/7!
    ps = std::fma(ps, y2, (R)8.33333377e-03f);        //  1/5!
    ps = std::fma(ps, y2, (R)-1.66666672e-01f);       // -1/3!
    ps = std::fma(ps, y2, (R)1.0f);
    R s = y * ps;
    // cos(y) = 1 + y2*(c2 + y2*(c4 + y2*(c6 + y2*(c8 + y2*(c10 + y2*c12)))))
    R pc = (R)2.08767570e-09f;                        //  1/12!
    pc = std::fma(pc, y2, (R)-2.75573192e-07f);       // -1/10!
    pc = std::fma(pc, y2, (R)2.48015876e-05f);        //  1/8!
    pc = std::fma(pc, y2, (R)-1.38888892e-03f);       // -1/6!
    pc = std::fma(pc, y2, (R)4.16666679e-02f);        //  1/4!
    pc = std::fma(pc, y2, (R)-0.5f);
    pc = std::fma(pc, y2, (R)1.0f);
    R c = pc;
    // rotate by the quadrant
    switch (q & 3) {
        case 0: cs = c;  sn = s;  break;
        case 1: cs = -s; sn = c;  break;
        case 2: cs = -c; sn = -s; break;
        default: cs = s; sn = -c; break;
    }
}

// ----------------------------------------------------------------------------------------------
// Vec3 — scalar algebra of src/geometry.rs:9-188.  NOTE: the scalar Vec3 never fuses (plain * and +,
// left to right); only PackedVec3 uses mul_add.  Both forms are restated separately below.
// ----------------------------------------------------------------------------------------------
template <class R> struct V3 { R x, y, z; };

template <class R> inline V3<R> operator+(V3<R> a, V3<R> b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }   // geometry.rs:38-47
template <class R> inline V3<R> operator-(V3<R> a, V3<R> b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }   // geometry.rs:49-58
template <class R> inline V3<R> operator*(V3<R> a, R s) { return {a.x * s, a.y * s, a.z * s}; }             // geometry.rs:60-69
template <class R> inline V3<R> operator*(R s, V3<R> a) { return a * s; }                                     // geometry.rs:71-76 (rhs * self)
template <class R> inline V3<R> operator/(V3<R> a, R s) { return {a.x / s, a.y / s, a.z / s}; }             // geometry.rs:78-87
template <class R> inline V3<R> operator-(V3<R> a) { return {-a.x, -a.y, -a.z}; }                             // geometry.rs:89-98
template <class R> inline R length_squared(V3<R> a) { return a.x * a.x + a.y * a.y + a.z * a.z; }            // geometry.rs:107-109
template <class R> inline R length(V3<R> a) { return std::sqrt(length_squared(a)); }                          // geometry.rs:111-113
template <class R> inline V3<R> unit(V3<R> a) { return a / length(a); }                                       // geometry.rs:119-121
template <class R> inline R dot(V3<R> a, V3<R> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }              // geometry.rs:123-125
template <class R> inline V3<R> cross(V3<R> a, V3<R> b) {                                                     // geometry.rs:127-133
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
template <class R> inline bool near_zero(V3<R> a) {                                                           // geometry.rs:134-137
    const R eps = (R)1e-8;
    return std::fabs(a.x) < eps && std::fabs(a.y) < eps && std::fabs(a.z) < eps;
}
template <class R> inline V3<R> reflect(V3<R> v, V3<R> n) {                                                   // geometry.rs:179-181
    // self - 2.0 * self.dot(normal) * normal  ==  self - (normal * (2.0 * dot))
    R s = (R)2.0 * dot(v, n);
    return v - (n * s);
}
template <class R> inline V3<R> refract(V3<R> v, V3<R> n, R ratio) {                                          // geometry.rs:183-188
    R cos_theta = std::fmin(dot(-v, n), (R)1.0);
    V3<R> perp = (v + (n * cos_theta)) * ratio;
    R s = -(std::sqrt(std::fabs((R)1.0 - length_squared(perp))));
    V3<R> par = n * s;
    return perp + par;
}

// PackedVec3 lane algebra (src/geometry.rs:432-468): the fused forms.
template <class R> inline R packed_length_squared(V3<R> a) {                                                  // geometry.rs:434-436
    return std::fma(a.z, a.z, std::fma(a.y, a.y, a.x * a.x));
}
template <class R> inline R packed_dot(V3<R> a, V3<R> b) {                                                    // geometry.rs:466-468
    return std::fma(a.z, b.z, std::fma(a.y, b.y, a.x * b.x));
}

// Deterministic replacement for Vec3::random_unit_vector (src/geometry.rs:139-152: three N(0,1) draws,
// normalised => uniform on the unit sphere).  Same distribution from two uniforms: z uniform in (-1,1],
// azimuth uniform.
template <class R> inline V3<R> unit_vector_from(R u, R v) {
    R z = (R)1.0 - (R)2.0 * u;
    R r2 = (R)1.0 - z * z;
    R r = std::sqrt(r2 > (R)0.0 ? r2 : (R)0.0);
    R cs, sn;
    cossin_2pi(v, cs, sn);
    return {r * cs, r * sn, z};
}
// Deterministic replacement for Vec3::random_in_unit_disk (src/geometry.rs:154-168: rejection sampling =>
// uniform on the unit disk).
template <class R> inline void unit_disk_from(R u, R v, R& dx, R& dy) {
    R r = std::sqrt(u);
    R cs, sn;
    cossin_2pi(v, cs, sn);
    dx = r * cs; dy = r * sn;
}

// ----------------------------------------------------------------------------------------------
// Scene description (src/objects.rs:203-214, src/materials.rs:43-118, src/ray_tracing.rs:100-104)
// ----------------------------------------------------------------------------------------------
enum MaterialKind : int32_t { LAMBERTIAN = 0, METAL = 1, DIELECTRIC = 2 };

template <class R> struct Material {
    int32_t kind;
    bool hollow;      // dielectric only (materials.rs:113)
    V3<R> albedo;     // lambertian, metal
    R param;          // metal: fuzzy_factor (already clamped to <= 1, materials.rs:79-88); dielectric: index_of_refraction
};
template <class R> struct Sphere { V3<R> center; R radius; int32_t material; };
template <class R> struct Scene {
    std::vector<Sphere<R>> spheres;
    std::vector<Material<R>> materials;
    bool both_roots = false;
    std::vector<Sphere<double>> exact64;   // mixed mode (R = float only): the scene at the reference's precision; empty = sweep in R
};

template <class R> struct Ray { V3<R> o, d; };   // src/ray.rs:5-18 — direction stored as given, never normalised

// ----------------------------------------------------------------------------------------------
// Camera (src/ray_tracing.rs:15-62).  camera_new is evaluated in double (as the reference does) and the
// resulting eight vectors are cast to Real for the per-ray arithmetic.
// ----------------------------------------------------------------------------------------------
struct CameraD {
    int32_t width, height;
    V3<double> center, ulc, viewport_u, viewport_v, defocus_u, defocus_v;
};

inline CameraD camera_new(int32_t width, int32_t height, double focal_length, double view_angle_deg,
                          V3<double> center, V3<double> look_at, V3<double> up, double defocus_angle_deg) {
    const double PI = 3.14159265358979323846;
    double aspect = (double)width / (double)height;                                                           // :28
    double viewport_height = std::tan((view_angle_deg * (PI / 180.0)) / 2.0) * focal_length * 2.0;            // :29
    double viewport_width = viewport_height * aspect;                                                         // :32
    V3<double> direction = unit(look_at - center);                                                            // :34
    V3<double> w = -direction;                                                                                // :35
    V3<double> u = unit(cross(up, w));                                                                        // :36
    V3<double> v = cross(w, u);                                                                               // :37
    V3<double> viewport_u = u * viewport_width;                                                               // :39
    V3<double> viewport_v = (-v) * viewport_height;                                                           // :40
    V3<double> ulc = ((center - (w * focal_length)) - (viewport_u / 2.0)) - (viewport_v / 2.0);               // :41
    double defocus_radius = focal_length * std::tan((defocus_angle_deg / 2.0) * (PI / 180.0));                // :43
    V3<double> defocus_u = defocus_radius * u;                                                                // :44
    V3<double> defocus_v = defocus_radius * v;                                                                // :45
    return CameraD{width, height, center, ulc, viewport_u, viewport_v, defocus_u, defocus_v};
}

template <class R> struct Camera {
    int32_t width, height;
    V3<R> center, ulc, viewport_u, viewport_v, defocus_u, defocus_v;
    bool defocus_enabled;
};
template <class R> inline V3<R> cast3(V3<double> a) { return {(R)a.x, (R)a.y, (R)a.z}; }
template <class R> inline Camera<R> camera_cast(const CameraD& c) {
    Camera<R> r;
    r.width = c.width; r.height = c.height;
    r.center = cast3<R>(c.center); r.ulc = cast3<R>(c.ulc);
    r.viewport_u = cast3<R>(c.viewport_u); r.viewport_v = cast3<R>(c.viewport_v);
    r.defocus_u = cast3<R>(c.defocus_u); r.defocus_v = cast3<R>(c.defocus_v);
    auto nz = [](V3<R> a) { return a.x != (R)0 || a.y != (R)0 || a.z != (R)0; };
    r.defocus_enabled = nz(r.defocus_u) || nz(r.defocus_v);
    return r;
}

// Camera::get_ray (src/ray_tracing.rs:77-89).  The reference always draws a disk sample; with a zero defocus
// radius the sample is multiplied by exact zeros, so skipping it is bit-identical (origin == center).
template <class R> inline Ray<R> camera_get_ray(const Camera<R>& cam, int32_t col, int32_t row, const Philox4& rnd) {
    R x_offset = u01<R>(rnd.v[0]);
    R y_offset = u01<R>(rnd.v[1]);
    R sx = ((R)col + x_offset) / (R)cam.width;
    R sy = ((R)row + y_offset) / (R)cam.height;
    V3<R> pixel_offset = (cam.viewport_u * sx) + (cam.viewport_v * sy);
    V3<R> pixel_center = cam.ulc + pixel_offset;
    V3<R> origin = cam.center;
    if (cam.defocus_enabled) {
        R dx, dy;
        unit_disk_from(u01<R>(rnd.v[2]), u01<R>(rnd.v[3]), dx, dy);
        origin = ((cam.defocus_u * dx) + (cam.defocus_v * dy)) + cam.center;
    }
    V3<R> direction = unit(pixel_center - origin);
    return Ray<R>{origin, direction};
}

// ----------------------------------------------------------------------------------------------
// Closest-hit sweep: one lane of Sphere::hit_packed (src/objects.rs:249-290) applied to every object in list
// order (src/ray_tracing.rs:399-401) with PackedHitRecords::update (src/objects.rs:140-155).
//
//  * near root only: root2_valid is computed from root1 (objects.rs:274), so valid == root1_valid & enabled and
//    root == root1 whenever valid.  A ray starting inside/on a sphere misses it.
//  * t range is 0.001..inf: inclusive start, exclusive end (src/simd_util.rs:76-129).
//  * closest-hit tie-break is `t <= best` — the LAST object in list order wins (objects.rs:141).
//  * the reference enters the root computation when ANY lane of the packet has disc >= 0 (objects.rs:259-261);
//    a lane with disc < 0 then gets sqrt(negative) = NaN -> root NaN -> outside the range -> invalid, which is
//    the same outcome as skipping that lane, so the per-lane form below is exact.
// ----------------------------------------------------------------------------------------------
template <class R> struct SweepHit { bool hit; int32_t index; R t; };

template <class R> inline R sphere_discriminant(const Sphere<R>& s, const Ray<R>& ray, R& half_b, R& a) {
    V3<R> oc = ray.o - s.center;                                   // objects.rs:252
    a = packed_length_squared(ray.d);                              // :253
    half_b = packed_dot(oc, ray.d);                                // :255
    R c = packed_length_squared(oc) - (s.radius * s.radius);       // :256  (radius.powi(2))
    return std::fma(half_b, half_b, -(a * c));                     // :257
}

// scene.both_roots (NOT the reference's live path; rtb200's RTB200_HIT_BOTH_ROOTS option): the rule of the scalar
// Sphere::hit (src/objects.rs:227-234) inside the packed arithmetic — when root1 is outside the range, root2 (:271) is
// tried.
template <class R> inline SweepHit<R> sweep_spheres(const std::vector<Sphere<R>>& spheres, bool both_roots, const Ray<R>& ray, bool enabled) {
    const R T_MIN = (R)0.001;
    const R INF = std::numeric_limits<R>::infinity();
    SweepHit<R> best{false, -1, INF};                              // PackedHitRecords::default, objects.rs:121-134
    const int32_t n = (int32_t)spheres.size();
    for (int32_t i = 0; i < n; i++) {
        R half_b, a;
        R disc = sphere_discriminant(spheres[i], ray, half_b, a);
        R inv_a = (R)1.0 / a;                                      // :254
        if (disc >= (R)0.0 && enabled) {                           // :259
            R sq = std::sqrt(disc);                                // :263
            R root1 = (-half_b - sq) * inv_a;                      // :270
            bool valid = (root1 >= T_MIN) && (root1 < INF) && enabled;   // :272-277
            if (!valid && both_roots) {                            // scalar rule, :230-234
                root1 = (-half_b + sq) * inv_a;                    // :271
                valid = (root1 >= T_MIN) && (root1 < INF) && enabled;
            }
            if (valid && root1 <= best.t) {                        // update_mask, :141
                best.hit = true; best.index = i; best.t = root1;   // :143-147
            }
        }
    }
    return best;
}

// The scene's sweep.  Mixed mode (exact64 filled): the ray's FP32 values are promoted exactly, the routine above runs in
// f64 over the f64 scene, and the winning root is rounded once to R.
template <class R> inline SweepHit<R> sweep(const Scene<R>& scene, const Ray<R>& ray, bool enabled = true) {
    if (!scene.exact64.empty()) {
        const Ray<double> r64{{(double)ray.o.x, (double)ray.o.y, (double)ray.o.z}, {(double)ray.d.x, (double)ray.d.y, (double)ray.d.z}};
        const SweepHit<double> h = sweep_spheres<double>(scene.exact64, scene.both_roots, r64, enabled);
        return SweepHit<R>{h.hit, h.index, (R)h.t};
    }
    return sweep_spheres<R>(scene.spheres, scene.both_roots, ray, enabled);
}

// PackedHitRecords::finalize (src/objects.rs:157-162) for the winning sphere.  The un-normalised normal is
// (o + d*t) - c as computed in hit_packed (objects.rs:279-280) with the winning root.
template <class R> struct HitRecord { V3<R> p, n; R t; bool front; int32_t material; };

template <class R> inline HitRecord<R> finalize(const Scene<R>& scene, const Ray<R>& ray, const SweepHit<R>& h) {
    const Sphere<R>& s = scene.spheres[h.index];
    V3<R> loc = ray.o + (ray.d * h.t);                             // PackedRays::at_t, src/ray.rs:98-100
    V3<R> nraw = loc - s.center;                                   // objects.rs:280
    R len = std::sqrt(packed_length_squared(nraw));                // geometry.rs:439-441
    V3<R> n = nraw / len;                                          // geometry.rs:444-446
    bool front = packed_dot(ray.d, n) < (R)0.0;                    // objects.rs:160
    if (!front) n = -n;                                            // :161
    return HitRecord<R>{loc, n, h.t, front, s.material};
}

// ----------------------------------------------------------------------------------------------
// Materials (src/materials.rs:53-147).  rnd = Philox block of this bounce: v[0],v[1] -> unit vector,
// v[2] -> Fresnel draw.
// ----------------------------------------------------------------------------------------------
template <class R> inline R reflectance(R cosine, R ior) {                                                    // materials.rs:121-124
    R q = ((R)1.0 - ior) / ((R)1.0 + ior);
    R r0 = q * q;
    R x = (R)1.0 - cosine;
    R x2 = x * x;
    R x4 = x2 * x2;
    R x5 = x * x4;                                                                                            // powi(5): x * ((x^2)^2)
    return r0 + ((R)1.0 - r0) * x5;
}

template <class R> inline void scatter(const Material<R>& m, const Ray<R>& ray, const HitRecord<R>& rec,
                                       const Philox4& rnd, V3<R>& attenuation, Ray<R>& out) {
    if (m.kind == LAMBERTIAN) {                                                                               // materials.rs:54-63
        V3<R> ruv = unit_vector_from(u01<R>(rnd.v[0]), u01<R>(rnd.v[1]));
        V3<R> dir = ruv + rec.n;
        if (near_zero(dir)) dir = rec.n;
        attenuation = m.albedo;
        out = Ray<R>{rec.p, dir};
    } else if (m.kind == METAL) {                                                                             // materials.rs:92-97
        V3<R> ruv = unit_vector_from(u01<R>(rnd.v[0]), u01<R>(rnd.v[1]));
        V3<R> dir = reflect(ray.d, rec.n) + (m.param * ruv);
        attenuation = m.albedo;
        out = Ray<R>{rec.p, dir};
    } else {                                                                                                  // materials.rs:128-147
        R ior = m.param;
        R ratio = rec.front ? ((R)1.0 / ior) : ior;
        V3<R> n = m.hollow ? -rec.n : rec.n;
        R cos_theta = std::fmin(dot(-ray.d, n), (R)1.0);
        R sin_theta = std::sqrt((R)1.0 - cos_theta * cos_theta);
        bool cannot_refract = ratio * sin_theta > (R)1.0;
        R u = u01<R>(rnd.v[2]);
        V3<R> dir;
        if (cannot_refract || reflectance(cos_theta, ratio) > u) dir = reflect(ray.d, n);
        else dir = refract(ray.d, n, ratio);
        attenuation = V3<R>{(R)1.0, (R)1.0, (R)1.0};
        out = Ray<R>{rec.p, dir};
    }
}

// Sky gradient (src/ray_tracing.rs:365-368 / :613-617): a = (d.y + 1) * 0.5 with the UN-normalised escaped
// direction; white*(1-a) + (0.5,0.7,1.0)*a.
template <class R> inline V3<R> sky(V3<R> d) {
    R a = (d.y + (R)1.0) * (R)0.5;
    R w = (R)1.0 - a;
    return {w + (R)0.5 * a, w + (R)0.7 * a, w + (R)1.0 * a};
}

// ----------------------------------------------------------------------------------------------
// One path: per-lane semantics of Scene::trace_vectorized (src/ray_tracing.rs:312-373) and trace_vectorized3
// (:508-628) == scalar trace_rays (:264-306) with the packed hit routine.  Returns the lane's colour
// contribution; segments (rays cast) is accumulated for the tests/s metric.
// ----------------------------------------------------------------------------------------------
struct Seed { uint32_t k0, k1; };

template <class R> inline V3<R> trace_path(const Scene<R>& scene, Ray<R> ray, int32_t depth_limit, uint32_t pixel,
                                           uint32_t sample, Seed seed, uint64_t* segments) {
    V3<R> color{(R)1.0, (R)1.0, (R)1.0};                                                                      // :319-320
    for (int32_t k = 0; k < depth_limit; k++) {                                                               // :323
        SweepHit<R> h = sweep(scene, ray);
        if (segments) (*segments)++;
        if (!h.hit) {                                                                                         // :353-357 + :365-369
            V3<R> s = sky(ray.d);
            return {color.x * s.x, color.y * s.y, color.z * s.z};
        }
        HitRecord<R> rec = finalize(scene, ray, h);                                                           // :335
        Philox4 rnd = philox4x32_10(pixel, sample, 1u + (uint32_t)k, 0u, seed.k0, seed.k1);
        V3<R> att; Ray<R> next;
        scatter(scene.materials[rec.material], ray, rec, rnd, att, next);                                     // :342
        color = {color.x * att.x, color.y * att.y, color.z * att.z};                                          // :343
        ray = next;                                                                                           // :346
    }
    return {(R)0.0, (R)0.0, (R)0.0};                                                                          // :370 still enabled => black
}

// Color::to_u8_array (src/color.rs:54-64): (sqrt(c) * 255.999) as u8 — Rust `as` truncates toward zero,
// saturates, and maps NaN to 0.
template <class R> inline uint8_t quantise(R c) {
    R v = std::sqrt(c) * (R)255.999;
    if (!(v >= (R)0.0)) return 0;      // NaN (sqrt of a negative) or negative
    if (v >= (R)255.0) return 255;
    return (uint8_t)(int)v;
}

// Pixel = sum over samples / spp (src/renderer.rs:155-163).  Samples [s0, s1) of the pixel are summed in
// sample order; `spp_total` is the divisor.
template <class R> inline V3<R> pixel_sum(const Scene<R>& scene, const Camera<R>& cam, int32_t col, int32_t row,
                                          int32_t s0, int32_t s1, int32_t depth_limit, Seed seed, uint64_t* segments) {
    V3<R> sum{(R)0.0, (R)0.0, (R)0.0};
    uint32_t pixel = (uint32_t)row * (uint32_t)cam.width + (uint32_t)col;
    for (int32_t s = s0; s < s1; s++) {
        Philox4 rnd = philox4x32_10(pixel, (uint32_t)s, DIM_CAMERA, 0u, seed.k0, seed.k1);
        Ray<R> ray = camera_get_ray(cam, col, row, rnd);
        V3<R> c = trace_path(scene, ray, depth_limit, pixel, (uint32_t)s, seed, segments);
        sum = sum + c;
    }
    return sum;
}

}  // namespace rto
