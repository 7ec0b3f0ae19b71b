// rt_device.cuh — device-side arithmetic of the path tracer: sampling, camera, hit finalisation, materials, sky,
// quantisation.  Operation order and FMA placement follow the reference (file:line cited per function; paths are
// relative to the reference repository root).  This translation unit is compiled with -fmad=false, so the ONLY
// fused multiply-adds are the explicit __fmaf_rn calls, which sit where the reference calls mul_add
// (src/geometry.rs:435,467; src/objects.rs:257).  sqrtf and '/' are IEEE (-prec-sqrt=true -prec-div=true).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

struct f3 { float x, y, z; };

__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
// Scalar Vec3 algebra (src/geometry.rs:38-137): plain mul/add, never fused.
__device__ __forceinline__ f3 add3(f3 a, f3 b) { return mk3(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z)); }
__device__ __forceinline__ f3 sub3(f3 a, f3 b) { return mk3(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)); }
__device__ __forceinline__ f3 mul3(f3 a, float s) { return mk3(__fmul_rn(a.x, s), __fmul_rn(a.y, s), __fmul_rn(a.z, s)); }
__device__ __forceinline__ f3 div3(f3 a, float s) { return mk3(__fdiv_rn(a.x, s), __fdiv_rn(a.y, s), __fdiv_rn(a.z, s)); }
__device__ __forceinline__ f3 neg3(f3 a) { return mk3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float dot3(f3 a, f3 b) {                                   // geometry.rs:123-125
    return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z));
}
__device__ __forceinline__ float len2_3(f3 a) { return dot3(a, a); }                  // geometry.rs:107-109
__device__ __forceinline__ f3 unit3(f3 a) { return div3(a, __fsqrt_rn(len2_3(a))); }  // geometry.rs:119-121
// PackedVec3 lane algebra (src/geometry.rs:434-436, :466-468): fused.
__device__ __forceinline__ float pdot3(f3 a, f3 b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, __fmul_rn(a.x, b.x))); }
__device__ __forceinline__ float plen2_3(f3 a) { return pdot3(a, a); }

// ---- Philox4x32-10: the deterministic sample stream (replaces rand::thread_rng, src/ray_tracing.rs:78-79) ------
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0;
        uint32_t n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += W0; k1 += W1;
    }
    return make_uint4(c0, c1, c2, c3);
}
#define RTB_DIM_CAMERA 0u

// 24-bit uniform in [0,1) (replaces gen_range(0.0..1.0), src/ray_tracing.rs:78-79, src/materials.rs:137).
__device__ __forceinline__ float u01(uint32_t x) { return __fmul_rn((float)(x >> 8), 5.9604644775390625e-08f); }

// cos(2*pi*v), sin(2*pi*v), v in [0,1): exact quadrant reduction + quarter-wave Taylor polynomials by Horner with
// explicit fma.  Bit-identical to rto::cossin_2pi<float> (oracle/rt_oracle.hpp).
__device__ __forceinline__ void cossin_2pi(float v, float& cs, float& sn) {
    float t = __fmul_rn(v, 4.0f);
    int q = (int)t;
    float f = __fsub_rn(t, (float)q);
    float y = __fmul_rn(f, 1.57079637f);
    float y2 = __fmul_rn(y, y);
    float ps = -2.50521084e-08f;
    ps = __fmaf_rn(ps, y2, 2.75573192e-06f);
    ps = __fmaf_rn(ps, y2, -1.98412701e-04f);
    ps = __fmaf_rn(ps, y2, 8.33333377e-03f);
    ps = __fmaf_rn(ps, y2, -1.66666672e-01f);
    ps = __fmaf_rn(ps, y2, 1.0f);
    float s = __fmul_rn(y, ps);
    float pc = 2.08767570e-09f;
    pc = __fmaf_rn(pc, y2, -2.75573192e-07f);
    pc = __fmaf_rn(pc, y2, 2.48015876e-05f);
    pc = __fmaf_rn(pc, y2, -1.38888892e-03f);
    pc = __fmaf_rn(pc, y2, 4.16666679e-02f);
    pc = __fmaf_rn(pc, y2, -0.5f);
    pc = __fmaf_rn(pc, y2, 1.0f);
    float c = pc;
    q &= 3;
    cs = (q == 0) ? c : (q == 1) ? -s : (q == 2) ? -c : s;
    sn = (q == 0) ? s : (q == 1) ? c : (q == 2) ? -s : -c;
}

// Uniform unit vector from two uniforms (stands in for Vec3::random_unit_vector, src/geometry.rs:139-152).
__device__ __forceinline__ f3 unit_vector_from(float u, float v) {
    float z = __fsub_rn(1.0f, __fmul_rn(2.0f, u));
    float r2 = __fsub_rn(1.0f, __fmul_rn(z, z));
    float r = __fsqrt_rn(r2 > 0.0f ? r2 : 0.0f);
    float cs, sn;
    cossin_2pi(v, cs, sn);
    return mk3(__fmul_rn(r, cs), __fmul_rn(r, sn), z);
}

// ---- camera (src/ray_tracing.rs:15-24, :77-89) ------------------------------------------------------------------
struct DeviceCamera {
    int32_t width, height;
    f3 center, ulc, viewport_u, viewport_v, defocus_u, defocus_v;
    int32_t defocus_enabled;
};

// Camera::get_ray (src/ray_tracing.rs:77-89).  With a zero defocus radius the disk sample multiplies exact zeros, so
// it is skipped (origin == center bit-for-bit).
__device__ __forceinline__ void camera_get_ray(const DeviceCamera& cam, int col, int row, uint4 rnd, f3& o, f3& d) {
    float x_offset = u01(rnd.x);
    float y_offset = u01(rnd.y);
    float sx = __fdiv_rn(__fadd_rn((float)col, x_offset), (float)cam.width);
    float sy = __fdiv_rn(__fadd_rn((float)row, y_offset), (float)cam.height);
    f3 pixel_offset = add3(mul3(cam.viewport_u, sx), mul3(cam.viewport_v, sy));
    f3 pixel_center = add3(cam.ulc, pixel_offset);
    o = cam.center;
    if (cam.defocus_enabled) {
        float r = __fsqrt_rn(u01(rnd.z));                       // uniform unit disk (src/geometry.rs:154-168)
        float cs, sn;
        cossin_2pi(u01(rnd.w), cs, sn);
        float ddx = __fmul_rn(r, cs), ddy = __fmul_rn(r, sn);
        o = add3(add3(mul3(cam.defocus_u, ddx), mul3(cam.defocus_v, ddy)), cam.center);
    }
    d = unit3(sub3(pixel_center, o));
}

// ---- hit record (PackedHitRecords::finalize, src/objects.rs:157-162; normal from :279-280) ------------------------
struct HitRec { f3 p, n; bool front; };

__device__ __forceinline__ HitRec finalize_hit(f3 o, f3 d, float t, f3 c) {
    HitRec h;
    h.p = add3(o, mul3(d, t));                  // PackedRays::at_t, src/ray.rs:98-100
    f3 nraw = sub3(h.p, c);                     // objects.rs:280
    float len = __fsqrt_rn(plen2_3(nraw));      // geometry.rs:439-441
    f3 n = div3(nraw, len);                     // geometry.rs:444-446
    h.front = pdot3(d, n) < 0.0f;               // objects.rs:160
    h.n = h.front ? n : neg3(n);                // :161
    return h;
}

// ---- materials (src/materials.rs:53-147) ------------------------------------------------------------------------
__device__ __forceinline__ f3 reflect3(f3 v, f3 n) {             // geometry.rs:179-181
    float s = __fmul_rn(2.0f, dot3(v, n));
    return sub3(v, mul3(n, s));
}
__device__ __forceinline__ f3 refract3(f3 v, f3 n, float ratio) { // geometry.rs:183-188
    float cos_theta = fminf(dot3(neg3(v), n), 1.0f);
    f3 perp = mul3(add3(v, mul3(n, cos_theta)), ratio);
    float s = -(__fsqrt_rn(fabsf(__fsub_rn(1.0f, len2_3(perp)))));
    return add3(perp, mul3(n, s));
}
__device__ __forceinline__ float reflectance(float cosine, float ior) {   // materials.rs:121-124
    float q = __fdiv_rn(__fsub_rn(1.0f, ior), __fadd_rn(1.0f, ior));
    float r0 = __fmul_rn(q, q);
    float x = __fsub_rn(1.0f, cosine);
    float x2 = __fmul_rn(x, x);
    float x4 = __fmul_rn(x2, x2);
    float x5 = __fmul_rn(x, x4);
    return __fadd_rn(r0, __fmul_rn(__fsub_rn(1.0f, r0), x5));
}

#define RTB_LAMBERTIAN 0
#define RTB_METAL 1
#define RTB_DIELECTRIC 2
#define RTB_HOLLOW_BIT 0x100

// Per-sphere material record on the device: {albedo r,g,b, param} + kind word (kind | hollow<<8).
// Material::get_hit_result for the three materials; rnd = this bounce's Philox block.
__device__ __forceinline__ void scatter(int kind_word, float4 mat, f3 d, const HitRec& rec, uint4 rnd, f3& att, f3& out_d) {
    int kind = kind_word & 0xff;
    if (kind == RTB_DIELECTRIC) {                                  // materials.rs:128-147
        float ior = mat.w;
        float ratio = rec.front ? __fdiv_rn(1.0f, ior) : ior;
        f3 n = (kind_word & RTB_HOLLOW_BIT) ? neg3(rec.n) : rec.n;
        float cos_theta = fminf(dot3(neg3(d), n), 1.0f);
        float sin_theta = __fsqrt_rn(__fsub_rn(1.0f, __fmul_rn(cos_theta, cos_theta)));
        bool cannot_refract = __fmul_rn(ratio, sin_theta) > 1.0f;
        float u = u01(rnd.z);
        if (cannot_refract || reflectance(cos_theta, ratio) > u) out_d = reflect3(d, n);
        else out_d = refract3(d, n, ratio);
        att = mk3(1.0f, 1.0f, 1.0f);
    } else {
        f3 ruv = unit_vector_from(u01(rnd.x), u01(rnd.y));
        if (kind == RTB_LAMBERTIAN) {                              // materials.rs:54-63
            f3 dir = add3(ruv, rec.n);
            const float eps = 1e-8f;
            if (fabsf(dir.x) < eps && fabsf(dir.y) < eps && fabsf(dir.z) < eps) dir = rec.n;   // near_zero, geometry.rs:134-137
            out_d = dir;
        } else {                                                   // materials.rs:92-97
            out_d = add3(reflect3(d, rec.n), mul3(ruv, mat.w));
        }
        att = mk3(mat.x, mat.y, mat.z);
    }
}

// Sky gradient with the un-normalised escaped direction (src/ray_tracing.rs:365-368, :613-617).
__device__ __forceinline__ f3 sky(f3 d) {
    float a = __fmul_rn(__fadd_rn(d.y, 1.0f), 0.5f);
    float w = __fsub_rn(1.0f, a);
    return mk3(__fadd_rn(w, __fmul_rn(0.5f, a)), __fadd_rn(w, __fmul_rn(0.7f, a)), __fadd_rn(w, __fmul_rn(1.0f, a)));
}

// Color::to_u8_array (src/color.rs:54-64): (sqrt(c) * 255.999) as u8 — truncating, saturating, NaN -> 0.
__device__ __forceinline__ uint32_t quantise(float c) {
    float v = __fmul_rn(__fsqrt_rn(c), 255.999f);
    if (!(v >= 0.0f)) return 0u;
    if (v >= 255.0f) return 255u;
    return (uint32_t)(int)v;
}

}  // namespace rtb
