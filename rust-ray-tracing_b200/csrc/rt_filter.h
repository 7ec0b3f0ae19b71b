// rt_filter.h — the conservative candidate filter of the closest-hit sweep.  Host + device (plain C++ / CUDA).
//
// The reference tests every (ray, sphere) pair with the discriminant of Sphere::hit_packed (src/objects.rs:252-257,
// 12 FP32 instructions in its operation order).  The sweep kernels run that exact arithmetic only on CANDIDATES; the
// all-pairs pass is this filter: 8 FP32 instructions per pair whose sign bit is clear for EVERY pair whose reference
// discriminant is >= 0 (no false negatives), and for a few near-silhouette pairs besides (false positives, removed
// by the exact test).  Results are therefore bit-identical to the reference-order arithmetic.
//
// Geometry.  With OC = o - c and D = d/|d|, the reference discriminant has the sign of
//     (OC.D)^2 - |OC|^2 + r^2                                    (r^2 = the stored float radius.powi(2))
// Expanded around the world origin this is  h^2 - (g + k)  with
//     h = c.D - o.D                      3 FFMA   (per ray: D, -(o.D))
//     g = |o|^2 - 2 c.o                  3 FFMA   (per ray: -2o, |o|^2)
//     k = |c|^2 - r^2                    per sphere, precomputed in f64 on the host (float4.w of the staged array)
//     w = g + k ; q = fma(h, h, -w)      FADD + FFMA
// The expansion cancels (|c|^2 - 2c.o + |o|^2), so q carries an absolute error of order u*S^2, S = |o| + |c| + r,
// u = 2^-24; the reference's own computed discriminant is off by up to ~24u(|OC|^2 + r^2) in the same units.  Both are
// covered by a margin M = 96u*S^2 <= 192u*|o|^2 + 192u*(|c|+r)^2, which is SEPARABLE: the ray's half is folded into
// its |o|^2 term and the sphere's half into k, so the margin costs no instruction.
//
// Error budget (all in units of u*S^2; DESIGN.md §4 has the derivation): reference discriminant 24, direction
// normalisation 20, h rounding 12, g rounding 7, k/w/q rounding 5  => 68 <= 96.
// tests/test_filter_cpu.py checks the no-false-negative property on 10^8-pair adversarial samples with this very code
// compiled for the host.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define RTB_HD __host__ __device__ __forceinline__
#else
#define RTB_HD inline
#endif

namespace rtb {

// 192 * 2^-24
#define RTB_FILTER_MARGIN 1.1444091796875e-05f

#if defined(__CUDA_ARCH__)
#define RTB_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define RTB_MUL(a, b) __fmul_rn((a), (b))
#define RTB_ADD(a, b) __fadd_rn((a), (b))
#define RTB_RSQRT(a) rsqrtf(a)
#else
#define RTB_FMA(a, b, c) fmaf((a), (b), (c))
#define RTB_MUL(a, b) ((a) * (b))
#define RTB_ADD(a, b) ((a) + (b))
#define RTB_RSQRT(a) (1.0f / sqrtf(a))
#endif

// Per-ray half of the filter.
struct FilterRay {
    float hx, hy, hz, nod;   // D = d/|d|, and -(o.D)
    float gx, gy, gz, oo;    // -2o, and |o|^2 * (1 - 192u)
};

// a = d.d as the sweep computes it (src/objects.rs:253).  A ray whose direction or origin is not finite and normal
// (the reference then never reports a hit, or the margins cannot be trusted) gets the "everything is a candidate"
// ray: q = +3e38 - k for every sphere.
RTB_HD FilterRay make_filter_ray(float ox, float oy, float oz, float dx, float dy, float dz, float a) {
    FilterRay f;
    float o2 = RTB_FMA(oz, oz, RTB_FMA(oy, oy, RTB_MUL(ox, ox)));
    if (!(a >= 1e-30f && a <= 1e30f && o2 <= 1e30f)) {
        f.hx = f.hy = f.hz = f.nod = 0.0f;
        f.gx = f.gy = f.gz = 0.0f;
        f.oo = -3.0e38f;
        return f;
    }
    float inv_len = RTB_RSQRT(a);
    f.hx = RTB_MUL(dx, inv_len); f.hy = RTB_MUL(dy, inv_len); f.hz = RTB_MUL(dz, inv_len);
    f.nod = -RTB_FMA(oz, f.hz, RTB_FMA(oy, f.hy, RTB_MUL(ox, f.hx)));
    f.gx = RTB_MUL(-2.0f, ox); f.gy = RTB_MUL(-2.0f, oy); f.gz = RTB_MUL(-2.0f, oz);
    f.oo = RTB_FMA(-RTB_FILTER_MARGIN, o2, o2);
    return f;
}

// The 8 FP32 instructions per (ray, sphere): sign bit clear <=> candidate.  s = {cx, cy, cz, k}.
RTB_HD float filter_value(const FilterRay& f, float cx, float cy, float cz, float k) {
    float h = RTB_FMA(cx, f.hx, RTB_FMA(cy, f.hy, RTB_FMA(cz, f.hz, f.nod)));
    float g = RTB_FMA(cx, f.gx, RTB_FMA(cy, f.gy, RTB_FMA(cz, f.gz, f.oo)));
    float w = RTB_ADD(g, k);
    return RTB_FMA(h, h, -w);
}

// Per-sphere half, evaluated in f64 on the host from the DEVICE values (float centre, float r^2):
// k = |c|^2 - r^2 - 192u*(|c| + r)^2, rounded to float and stepped one ulp towards -inf.
inline float filter_sphere_k(float cx, float cy, float cz, float r2) {
    double c2 = (double)cx * cx + (double)cy * cy + (double)cz * cz;
    double r = sqrt(r2 > 0.0f ? (double)r2 : 0.0);
    double b = sqrt(c2) + r;
    double k = c2 - (double)r2 - (double)RTB_FILTER_MARGIN * b * b;
    float kf = (float)k;
    return nextafterf(kf, -INFINITY);
}

}  // namespace rtb
