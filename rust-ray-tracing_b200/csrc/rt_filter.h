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
// Since round 2 the exact test behind the filters is the reference's routine in f64 on the f64 spheres (rt_kernels.cuh,
// test_candidate).  Its discriminant is exact to ~2^-53 relative, i.e. the "reference discriminant 24" term shrinks to nothing,
// while the f64 sphere differs from the FP32 one the filters see by <= u|c| in the centre and <= 2u r^2 in r^2: the true
// discriminant moves by <= 2u|OC||c| + 2u r^2 <= 4u*S^2, inside the 28 units of slack.
// tests/test_filter_cpu.py checks the no-false-negative property on adversarial samples with this very code compiled for the
// host (IEEE 1/sqrt); rtb200_filter_check (k_filter_check) does it ON THE DEVICE with the arithmetic the kernels run
// (approximate rsqrtf) against the f64 discriminant: 1e8 pairs per scene family in the GPU suite, 0 false negatives.
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

// The same, also handing back h = (c - o).D, the FP32 projection of the centre on the ray (stage 2 of the two-stage sweep
// uses its sign to drop spheres behind the origin).  |h - (c - o).d/|d|| <= 8u(|o| + |c| + r): direction normalisation u|OC|,
// three roundings each in -(o.D) and in the chain, f32 centre against the f64 centre u|c|.
RTB_HD float filter_value_h(const FilterRay& f, float cx, float cy, float cz, float k, float& h) {
    h = RTB_FMA(cx, f.hx, RTB_FMA(cy, f.hy, RTB_FMA(cz, f.hz, f.nod)));
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

// ---- stage 1 of the two-stage sweep: the slab test ------------------------------------------------------------------
// A cheaper conservative test that runs for ALL pairs in the parameter-space kernel; its candidates (a few per cent
// of the pairs) then go through filter_value above and, if they pass that too, through the exact arithmetic.
//
// Drop one coordinate axis (`up`, chosen per scene by the host: the axis along which the sphere centres spread least)
// and look at the scene from along it: the ray becomes a line in the remaining (a, b) plane, and a sphere the line of
// sight misses by more than its radius cannot be hit.  With U = (d_b, -d_a)/|(d_a, d_b)| (a unit vector orthogonal to
// the ray, with no `up` component)
//     p = (c_a - o_a) U_a + (c_b - o_b) U_b          2 FFMA   (per ray: U_a, U_b, -(o.U))
//     candidate  <=>  p^2 <= R' + M_r
// |p| is the length of a projection of the centre-to-line vector, so |p| <= dist(c, line); the reference's computed
// discriminant is >= 0 only if dist^2 <= r^2 + 24u(|OC|^2 + r^2), and p carries a rounding error of at most 16u*S
// (S = |o| + |c| + r): (dist + 16uS)^2 <= r^2 + 96u*S^2 again, split as for the filter into the sphere's half
// R' = r^2 + 192u(|c|+r)^2 (host, f64, rounded up) and the ray's half M_r = 192u|o|^2.
//
// The ray's half costs no instruction either: with R_min = min over the scene of R', scale U by
//     s = sqrt(R_min / (R_min + M_r)) * (1 - 2^-20)
// (per ray, folded into U_a, U_b and -(o.U)).  If p^2 <= R' + M_r then (s p)^2 <= (R' + M_r) R_min / (R_min + M_r) <= R'
// because R' >= R_min, so
//     q = fma(-p', p', R')                           FFMA, candidate <=> sign bit clear      (p' = the scaled p)
// is conservative for every sphere: 3 FFMA per pair.  It is looser than the unscaled test by the factor
// 1 + M_r/R_min on r^2 — 11 % for a camera ray 20 units from the origin over 0.2-radius spheres, under 1 % for bounce
// rays — and stage 2 removes what that lets through.  A scene with one tiny sphere would make R_min useless, so the
// host raises every R' to a floor of a quarter of the scene's median R' (slab_apply_floor): enlarging a sphere's R' is
// always safe.  The (1 - 2^-20) covers the rounding of s and of the scaled coefficients.
struct SlabRay {
    float ua, ub, nou;
};

// Component `axis` of a vector given as three scalars (axis is warp-uniform).
RTB_HD float axis_pick(float x, float y, float z, int axis) { return axis == 0 ? x : axis == 1 ? y : z; }

// up = the dropped axis (0, 1, 2); a = (up + 1) % 3, b = (up + 2) % 3.  a2 = d.d as the sweep computes it;
// rp_min = the scene's smallest (floored) R'.
RTB_HD SlabRay make_slab_ray(float ox, float oy, float oz, float dx, float dy, float dz, float a2, int up, float rp_min) {
    SlabRay s;
    s.ua = s.ub = s.nou = 0.0f;                               // p' = 0: q = R' >= 0, everything is a candidate
    const int ia = up == 2 ? 0 : up + 1, ib = ia == 2 ? 0 : ia + 1;
    float o2 = RTB_FMA(oz, oz, RTB_FMA(oy, oy, RTB_MUL(ox, ox)));
    if (!(a2 >= 1e-30f && a2 <= 1e30f && o2 <= 1e30f)) return s;             // degenerate ray
    float t = RTB_MUL(rp_min, RTB_FMA(RTB_FILTER_MARGIN, o2, rp_min));       // R_min (R_min + M_r)
    if (!(t >= 1e-36f && t <= 1e36f)) return s;
    float scale = RTB_MUL(RTB_MUL(rp_min, RTB_RSQRT(t)), 0.99999904632568359375f);   // sqrt(R_min/(R_min+M_r)) (1 - 2^-20)
    float inv_len = RTB_RSQRT(a2);
    float ha = RTB_MUL(axis_pick(dx, dy, dz, ia), inv_len), hb = RTB_MUL(axis_pick(dx, dy, dz, ib), inv_len);
    float h2 = RTB_FMA(hb, hb, RTB_MUL(ha, ha));
    if (h2 < 1e-15f) {                                        // ray along `up`: any direction in the (a, b) plane is orthogonal
        s.ua = scale; s.ub = 0.0f;
    } else {
        float inv = RTB_MUL(RTB_RSQRT(h2), scale);
        s.ua = RTB_MUL(hb, inv); s.ub = -RTB_MUL(ha, inv);
    }
    float oa = axis_pick(ox, oy, oz, ia), ob = axis_pick(ox, oy, oz, ib);
    s.nou = -RTB_FMA(ob, s.ub, RTB_MUL(oa, s.ua));
    return s;
}

// 3 FP32 instructions per (ray, sphere): sign bit clear <=> candidate.  Sphere = {c_a, c_b, R'}.
RTB_HD float slab_value(const SlabRay& s, float ca, float cb, float rp) {
    float p = RTB_FMA(ca, s.ua, RTB_FMA(cb, s.ub, s.nou));
    return RTB_FMA(-p, p, rp);
}

// Per-sphere half of the slab test, f64 on the host from the DEVICE values: R' = r^2 + 192u(|c| + r)^2, rounded to float
// and stepped one ulp up.
inline float slab_sphere_r(float cx, float cy, float cz, float r2) {
    double c2 = (double)cx * cx + (double)cy * cy + (double)cz * cz;
    double r = sqrt(r2 > 0.0f ? (double)r2 : 0.0);
    double b = sqrt(c2) + r;
    double rp = (double)(r2 > 0.0f ? r2 : 0.0f) + (double)RTB_FILTER_MARGIN * b * b;
    return nextafterf((float)rp, INFINITY);
}

// Host: raise every R' of the scene (stride floats apart) to the floor max(median/4, 1e-18) and return the minimum,
// the R_min of make_slab_ray.  n == 0 returns 1.
inline float slab_apply_floor(float* rp, int n, int stride) {
    if (n <= 0) return 1.0f;
    float lo = INFINITY, hi = -INFINITY;
    for (int i = 0; i < n; i++) { float v = rp[(size_t)i * stride]; lo = v < lo ? v : lo; hi = v > hi ? v : hi; }
    // median by bisection on the value range (no allocation, n is a few thousand at most)
    float a = lo, b = hi;
    for (int it = 0; it < 48 && a < b; it++) {
        float m = a + 0.5f * (b - a);
        int below = 0;
        for (int i = 0; i < n; i++) below += rp[(size_t)i * stride] <= m;
        if (2 * below >= n) b = m; else a = nextafterf(m, INFINITY);
    }
    float floor_v = 0.25f * b;
    if (!(floor_v >= 1e-18f)) floor_v = 1e-18f;
    float mn = INFINITY;
    for (int i = 0; i < n; i++) {
        float& v = rp[(size_t)i * stride];
        if (v < floor_v) v = floor_v;
        mn = v < mn ? v : mn;
    }
    return mn;
}

}  // namespace rtb
