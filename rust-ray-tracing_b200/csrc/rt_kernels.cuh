// rt_kernels.cuh — the sm_100a kernels of the render hot path.
//
//   sweep_spheres     the closest-hit sweep of RAYS rays per lane against a shared-memory sphere array; the device
//                     mirror of `for object in &self.objects { object.hit_packed(..) }` (src/ray_tracing.rs:399-401)
//                     with Sphere::hit_packed (src/objects.rs:249-290) and PackedHitRecords::update (:140-155).
//                     A warp is the reference's PackedRays<N> with N = 32*RAYS: one sphere is broadcast from shared
//                     memory per step (= Simd::splat, src/objects.rs:252-256) and tested against every lane's rays.
//   k_intersect_batch K1: the sweep alone for a batch of rays (parity + throughput probe).
//   k_trace_paths     K2: persistent path-tracing kernel — Camera::get_ray + Scene::trace_vectorized* + materials
//                     + sky + per-pixel accumulation (src/renderer.rs:141-176, src/ray_tracing.rs:77-89, :312-628,
//                     src/materials.rs:53-147).  A lane slot = one pixel's sample stream; path regeneration replaces
//                     the reference's lane compaction (src/ray_tracing.rs:429-483, :561-607).
//   k_intersect_batch_f64  K1d: the same routine in f64 over f64 spheres, every pair (the reference-precision probe).
//   k_combine         sums the K sample-split partial planes in plane order.
//   k_resolve         K3: sum/spp + Color::to_u8_array + image assembly (src/renderer.rs:161-163, :376-384;
//                     src/color.rs:54-64).
//   k_fp32_peak       K0: dependent-free FFMA chains, the measured FP32 roofline for this path.
#pragma once
#include "rt_device.cuh"
#include "rt_filter.h"

namespace rtb {

constexpr int SWEEP_BLOCK = 32;          // spheres per candidate-mask word
constexpr int LIST_CAP = 6;              // deferred candidate-mask entries per ray before a forced drain (general case)
constexpr int LIST_CAP_SMALL = 16;       // scenes of <= 512 spheres: one entry per 32-block always fits, no overflow check
constexpr int LIST_CAP_V = 8;            // parameter-space variant (two rays per lane, path state in shared memory)
#define RTB_INF __int_as_float(0x7f800000)

// ---- sphere staging ------------------------------------------------------------------------------------------
// Three device arrays per scene take part in the sweep, all in list order and padded to whole 32-blocks with never-hit
// spheres: the FILTER spheres float4 {cx, cy, cz, k} (rt_filter.h), which the one-stage sweeps stage in shared memory
// and stage 2 of the two-stage sweep reads per candidate; the stage-1 triples; and the EXACT spheres
// double4 {cx, cy, cz, r*r} at the reference's own precision (the f64 values of the scene description, not their f32
// roundings), read only by the exact candidate test below.  Shading reads the f32 centres float4 {cx, cy, cz, r*r}.
__device__ __forceinline__ void stage_spheres(float4* s_sph, const float4* __restrict__ g_flt, int begin, int count) {
    for (int i = threadIdx.x; i < count; i += blockDim.x) s_sph[i] = __ldg(&g_flt[begin + i]);
}

// One ray as the exact candidate test needs it (origin, direction as stored: never normalised, src/ray.rs:13-18).
// both_roots: warp-uniform hit mode (false = the reference's packed path: near root only; true = the scalar
// Sphere::hit rule, src/objects.rs:227-234: the far root when the near one is out of range).
// inv_a = 1 / (d.d) in f64 (src/objects.rs:253-254), evaluated ONCE per ray and sweep (the division is the one expensive
// f64 operation of the exact test) and carried with the ray: in registers, or parked in the two shared-memory words that
// hold the sweep's result between trips.
struct ExactRay { float ox, oy, oz, dx, dy, dz; double inv_a; bool both_roots; };

__device__ __forceinline__ double exact_inv_a(f3 d) {
    const double dx = (double)d.x, dy = (double)d.y, dz = (double)d.z;
    const double a = __fma_rn(dz, dz, __fma_rn(dy, dy, __dmul_rn(dx, dx)));                         // src/objects.rs:253
    return __drcp_rn(a);                                                                            // :254 (1.0 / a, correctly rounded)
}
__device__ __forceinline__ ExactRay make_exact_ray(f3 o, f3 d, bool both_roots) {
    ExactRay r;
    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.dx = d.x; r.dy = d.y; r.dz = d.z; r.both_roots = both_roots;
    r.inv_a = exact_inv_a(d);
    return r;
}
// The ray's half of the candidate filter; a = d.d as the sweep computes it (src/objects.rs:253).
__device__ __forceinline__ FilterRay make_filter(f3 o, f3 d) { return make_filter_ray(o.x, o.y, o.z, d.x, d.y, d.z, plen2_3(d)); }

// ---- the exact candidate test: Sphere::hit_packed at the reference's precision -----------------------------------------
// The reference is f64 throughout (src/geometry.rs:9-13).  The all-pairs passes of the sweep are FP32 filters with no
// false negatives; what survives them (~1 % of the pairs) is tested HERE with the reference's own arithmetic in its own
// scalar type: the FP32 ray of the path state promoted to f64 (exact), the f64 spheres of the scene description, the
// reference's operation order and mul_add sites (src/objects.rs:252-257,263,270-277), IEEE sqrt and division, and the
// `<=` closest-hit update on the f64 root (:141).  For a given (FP32-valued) ray the sweep therefore returns the object
// index of the reference's f64 routine exactly and its hit distance rounded once to FP32 (<= 6e-8 relative): the
// north_star's intersection bar (index exact, t within 1e-5 of the packed-SIMD hit routine) holds on the render path, not
// only on a probe.  An FP32 version of this test (round 1) cancelled on the 1000-unit ground sphere: 77 % of hits inside
// 1e-5, worst 6e-4.  B200 issues FP64 at half the FP32 rate and this code sees ~1 % of the pairs.
struct ExactRay64 { double ox, oy, oz, dx, dy, dz, a, inv_a; };

__device__ __forceinline__ ExactRay64 promote_ray(const ExactRay& r) {
    ExactRay64 e;
    e.ox = (double)r.ox; e.oy = (double)r.oy; e.oz = (double)r.oz;
    e.dx = (double)r.dx; e.dy = (double)r.dy; e.dz = (double)r.dz;
    e.a = __fma_rn(e.dz, e.dz, __fma_rn(e.dy, e.dy, __dmul_rn(e.dx, e.dx)));                        // src/objects.rs:253
    e.inv_a = r.inv_a;                                                                              // :254, hoisted (exact_inv_a)
    return e;
}

// The body of hit_packed for one lane (src/objects.rs:252-288) followed by PackedHitRecords::update (:140-147).
// s = {cx, cy, cz, r*r} (radius.powi(2) evaluated in f64 on the host, :256).
__device__ __forceinline__ double exact_discriminant(const ExactRay64& r, double4 s, double& hb) {
    const double ocx = __dsub_rn(r.ox, s.x), ocy = __dsub_rn(r.oy, s.y), ocz = __dsub_rn(r.oz, s.z);   // :252
    hb = __fma_rn(ocz, r.dz, __fma_rn(ocy, r.dy, __dmul_rn(ocx, r.dx)));                               // :255
    const double cc = __dsub_rn(__fma_rn(ocz, ocz, __fma_rn(ocy, ocy, __dmul_rn(ocx, ocx))), s.w);     // :256
    return __fma_rn(hb, hb, -__dmul_rn(r.a, cc));                                                     // :257
}
//
// Near-root mode, half_b >= 0 (the sphere's centre is not ahead of the origin): root1 = (-half_b - sqrt(disc)) * inv_a is
// <= 0 < 0.001 whatever disc is, so the lane is invalid in the reference too (:272-277) and the rest is skipped — exactly.
__device__ __forceinline__ void test_candidate(const ExactRay64& r, bool both_roots, double4 s, int index, double& best_t, int& best_i) {
    const double ocx = __dsub_rn(r.ox, s.x), ocy = __dsub_rn(r.oy, s.y), ocz = __dsub_rn(r.oz, s.z);   // :252
    const double hb = __fma_rn(ocz, r.dz, __fma_rn(ocy, r.dy, __dmul_rn(ocx, r.dx)));                  // :255
    if (!both_roots && !(hb < 0.0)) return;
    const double cc = __dsub_rn(__fma_rn(ocz, ocz, __fma_rn(ocy, ocy, __dmul_rn(ocx, ocx))), s.w);     // :256
    const double disc = __fma_rn(hb, hb, -__dmul_rn(r.a, cc));                                        // :257
    if (disc >= 0.0) {                                                                             // :259
        const double T_MIN_D = 0.001, INF_D = __longlong_as_double(0x7ff0000000000000LL);          // 0.001..f64::INFINITY (src/ray_tracing.rs:400)
        const double sq = __dsqrt_rn(disc);                                                        // :263
        double root1 = __dmul_rn(__dsub_rn(-hb, sq), r.inv_a);                                     // :270
        bool valid = (root1 >= T_MIN_D) && (root1 < INF_D);                                        // :272-277 (near root only, F4)
        if (!valid && both_roots) {                                                                // scalar Sphere::hit, :230-234
            root1 = __dmul_rn(__dadd_rn(-hb, sq), r.inv_a);                                        // :271
            valid = (root1 >= T_MIN_D) && (root1 < INF_D);
        }
        // :141 — `t <= best` applied in list order keeps, among the hits of smallest t, the LAST object (F5).  Stated as an
        // order-free rule (smaller t, or equal t and larger index) the candidates may be visited in any order.
        if (valid && (root1 < best_t || (root1 == best_t && index > best_i))) { best_t = root1; best_i = index; }
    }
}
__device__ __forceinline__ double4 load_exact(const double4* __restrict__ p) {                     // 32 bytes, one sector
    const double2 lo = __ldg(reinterpret_cast<const double2*>(p)), hi = __ldg(reinterpret_cast<const double2*>(p) + 1);
    return make_double4(lo.x, lo.y, hi.x, hi.y);
}
#define RTB_INF_D __longlong_as_double(0x7ff0000000000000LL)

// Where the sweep reads its warp-uniform sphere operand from.
//   SmemSpheres : the staged shared-memory array (LDS.128 -> 4 registers; a chain FFMA then reads 3 registers, which
//                 costs two register-file cycles unless the operand-reuse cache hits: ~12.5 cycles per warp-test).
//   ParamSpheres: the kernel's parameter space (constant bank 0; LDCU.64 x2 -> 4 UNIFORM registers; a chain FFMA then
//                 reads 2 registers + 1 uniform register: one cycle.  ~11 cycles per warp-test with two rays per lane;
//                 tools/filter_probe.cu, profiles/r1_filter_probe.txt).  Up to N spheres (32 KB of parameters).
struct SmemSpheres {
    const float4* p;
    __device__ __forceinline__ float4 at(int i) const { return p[i]; }
};

// Where the drain reads a lane's rays from: registers (resident kernel) or the lane's path state in shared memory.
struct RegisterRays {
    const ExactRay* ray;
    // r may be a run-time value (drain_pending): select, never index (an indexed register array would go to local memory)
    __device__ __forceinline__ ExactRay load(int r) const {
        ExactRay e = ray[0];
        if (r) { const ExactRay& o = ray[1]; e.ox = o.ox; e.oy = o.oy; e.oz = o.oz; e.dx = o.dx; e.dy = o.dy; e.dz = o.dz; e.inv_a = o.inv_a; }
        return e;
    }
    __device__ __forceinline__ f3 origin(int r) const { return mk3(ray[r].ox, ray[r].oy, ray[r].oz); }
    __device__ __forceinline__ bool both_roots() const { return ray[0].both_roots; }
};

// Drains one ray's deferred candidate list in sweep order.  list entry = {sign mask of up to 32 filter values (bit
// 31 = first sphere of the block, set = not a candidate), index of the block's first sphere in the staged array}.
// g_exact = the exact (f64) spheres of the staged range (global memory, L1-resident).
__device__ __forceinline__ void drain_candidates(const double4* __restrict__ g_exact, int index_base, const uint2* list, int list_stride, int& n_list,
                                                 const ExactRay& r32, double& best_t, int& best_i) {
    const ExactRay64 r = promote_ray(r32);
    int e = 0;
    uint32_t cand = 0;
    int cbase = 0;
    for (;;) {
        if (cand == 0) {
            if (e >= n_list) break;
            uint2 v = list[e * list_stride];
            e++;
            cand = ~v.x;
            cbase = (int)v.y;
        }
        int j = __clz(cand);
        cand &= ~(0x80000000u >> j);
        int local = cbase + j;
        test_candidate(r, r32.both_roots, load_exact(&g_exact[local]), index_base + local, best_t, best_i);
    }
    n_list = 0;
}

// The sweep of RAYS rays per lane over `count` spheres (global indices index_base .. index_base+count).
//
// Fast path per (sphere, ray), all lanes, no branch: the 8 FP32 operations of the conservative candidate filter
// (rt_filter.h) plus ONE funnel shift that appends the filter value's sign bit to a 32-sphere mask word; one sphere
// load serves all RAYS rays of the lane.  A set sign bit means "certainly disc < 0 in the reference's arithmetic".
// Slow path (deferred, lane-parallel): mask words with a clear bit are pushed to a per-ray list in shared memory and
// re-tested with the reference's exact arithmetic (discriminant, sqrt, root, range, `<=`) after the sweep, in sweep
// order — so the result is that of the reference's sequential `update` over the object list.
// list layout: entry e of ray r of this thread at list[(r * CAP + e) * list_stride].  CAP >= number of 32-blocks
// (CHECKED = false) removes the overflow test from the per-block path.
template <int RAYS, int CAP, bool CHECKED, class SphSrc, class RaySrc>
__device__ __forceinline__ void sweep_spheres(const SphSrc& sph, const RaySrc& rays, const double4* __restrict__ g_exact, int count,
                                              int index_base, uint2* list, int list_stride, const FilterRay (&f)[RAYS],
                                              double (&best_t)[RAYS], int (&best_i)[RAYS]) {
    int n_list[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) n_list[r] = 0;
    const int full = count & ~(SWEEP_BLOCK - 1);
    // The block loop holds no divergent inner loop (that would push its counters out of the uniform datapath): when some
    // lane's list fills up, the whole warp leaves the loop on a vote, the full lists are drained, and the loop resumes.
    int blk = 0;
    for (;;) {
        for (; blk < full; blk += SWEEP_BLOCK) {
            uint32_t mask[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = 0;
            constexpr int UNROLL = SWEEP_BLOCK / RAYS;  // keep the unrolled body at ~350 instructions (instruction cache)
#pragma unroll 1
            for (int sub = 0; sub < SWEEP_BLOCK; sub += UNROLL) {
#pragma unroll
                for (int j = 0; j < UNROLL; j++) {
                    float4 s = sph.at(blk + sub + j);                                             // warp-uniform = Simd::splat
#pragma unroll
                    for (int r = 0; r < RAYS; r++) {
                        float q = filter_value(f[r], s.x, s.y, s.z, s.w);                         // 8 FP32 instructions
                        mask[r] = __funnelshift_l(__float_as_uint(q), mask[r], 1);                // (mask << 1) | sign(q)
                    }
                }
            }
            bool list_full = false;
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                if (mask[r] != 0xffffffffu) {
                    list[(r * CAP + n_list[r]) * list_stride] = make_uint2(mask[r], (uint32_t)blk);
                    n_list[r]++;
                    list_full = list_full || n_list[r] == CAP;
                }
            }
            if (CHECKED && __any_sync(0xffffffffu, list_full)) { blk += SWEEP_BLOCK; break; }
        }
        if (CHECKED) {
#pragma unroll
            for (int r = 0; r < RAYS; r++)
                if (n_list[r] == CAP)
                    drain_candidates(g_exact, index_base, list + r * CAP * list_stride, list_stride, n_list[r], rays.load(r), best_t[r], best_i[r]);
        }
        if (blk >= full) break;
    }
    const int rem = count - full;                     // 0..31 spheres left: rolled loop, same arithmetic
    if (rem > 0) {
        uint32_t mask[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) mask[r] = 0;
#pragma unroll 2
        for (int j = 0; j < rem; j++) {
            float4 s = sph.at(full + j);
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                float q = filter_value(f[r], s.x, s.y, s.z, s.w);
                mask[r] = __funnelshift_l(__float_as_uint(q), mask[r], 1);
            }
        }
        const uint32_t fill = 0xffffffffu >> rem;     // the 32-rem unused low positions read as "not a candidate"
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            uint32_t m = (mask[r] << (32 - rem)) | fill;
            if (m != 0xffffffffu) {
                list[(r * CAP + n_list[r]) * list_stride] = make_uint2(m, (uint32_t)full);
                n_list[r]++;
            }
        }
    }
#pragma unroll
    for (int r = 0; r < RAYS; r++)
        if (n_list[r] > 0)
            drain_candidates(g_exact, index_base, list + r * CAP * list_stride, list_stride, n_list[r], rays.load(r), best_t[r], best_i[r]);
}

// The sweep of the parameter-space kernel (at most 32 blocks = 1024 spheres): same fast path, but a block loop with NO
// data-dependent exit, so that ptxas keeps its counters and the sphere loads in the uniform datapath (LDCU -> uniform
// registers).  Per ray a 32-bit `blocks` word gets one bit per 32-sphere block (bit b = block b has a candidate) and
// the masks of the first CAP such blocks are stored (4 bytes each, list[(r * CAP + e) * list_stride]).  The drain
// walks the set bits in block order: a stored mask is re-tested bit by bit, a block beyond the CAP-th (rare) is
// re-tested whole.  Exact tests therefore run in sweep order, as PackedHitRecords::update needs (`<=`: last wins).
template <int RAYS, int CAP, class SphSrc, class RaySrc>
__device__ __forceinline__ void sweep_spheres_bitmap(const SphSrc& sph, const RaySrc& rays, const double4* __restrict__ g_exact, int count,
                                                     uint32_t* list, int list_stride, const FilterRay (&f)[RAYS], double (&best_t)[RAYS],
                                                     int (&best_i)[RAYS]) {
    uint32_t blocks[RAYS];
    int n_list[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) { blocks[r] = 0u; n_list[r] = 0; }
    const int full = count & ~(SWEEP_BLOCK - 1);
    uint32_t bit = 1u;
    for (int blk = 0; blk < full; blk += SWEEP_BLOCK, bit <<= 1) {
        uint32_t mask[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) mask[r] = 0;
        constexpr int UNROLL = RAYS == 1 ? 32 : RAYS == 2 ? 16 : 8;   // divides SWEEP_BLOCK; body of ~350 instructions
#pragma unroll 1
        for (int sub = 0; sub < SWEEP_BLOCK; sub += UNROLL) {
#pragma unroll
            for (int j = 0; j < UNROLL; j++) {
                float4 s = sph.at(blk + sub + j);
#pragma unroll
                for (int r = 0; r < RAYS; r++) {
                    float q = filter_value(f[r], s.x, s.y, s.z, s.w);
                    mask[r] = __funnelshift_l(__float_as_uint(q), mask[r], 1);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            if (mask[r] != 0xffffffffu) {
                if (n_list[r] < CAP) list[(r * CAP + n_list[r]) * list_stride] = mask[r];
                n_list[r]++;
                blocks[r] |= bit;
            }
        }
    }
    const int rem = count - full;                     // 0..31 spheres left: rolled loop, same arithmetic
    if (rem > 0) {
        uint32_t mask[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) mask[r] = 0;
#pragma unroll 2
        for (int j = 0; j < rem; j++) {
            float4 s = sph.at(full + j);
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                float q = filter_value(f[r], s.x, s.y, s.z, s.w);
                mask[r] = __funnelshift_l(__float_as_uint(q), mask[r], 1);
            }
        }
        const uint32_t fill = 0xffffffffu >> rem;
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            uint32_t m = (mask[r] << (32 - rem)) | fill;
            if (m != 0xffffffffu) {
                if (n_list[r] < CAP) list[(r * CAP + n_list[r]) * list_stride] = m;
                n_list[r]++;
                blocks[r] |= bit;
            }
        }
    }
    // drain, lane-parallel, in sweep order
#pragma unroll
    for (int r = 0; r < RAYS; r++) {
        if (blocks[r] == 0u) continue;
        const ExactRay ray32 = rays.load(r);
        const ExactRay64 ray = promote_ray(ray32);
        const uint32_t fill = rem > 0 ? (0xffffffffu >> rem) : 0u;                                  // never-candidates of a partial last block
        uint32_t todo = blocks[r];
        int e = 0;
        while (todo) {
            const int b = __ffs(todo) - 1;
            todo &= todo - 1u;
            const int cbase = b * SWEEP_BLOCK;
            uint32_t cand = (e < CAP) ? ~list[(r * CAP + e) * list_stride] : (cbase == full ? ~fill : 0xffffffffu);
            e++;
            while (cand) {
                const int j = __clz(cand);
                cand &= ~(0x80000000u >> j);
                test_candidate(ray, ray32.both_roots, load_exact(&g_exact[cbase + j]), cbase + j, best_t[r], best_i[r]);
            }
        }
    }
}

// Stages 2 and 3 of the two-stage sweep for ONE ray and one chunk of up to 8 blocks: a lane-parallel walk over the
// stage-1 candidates in sweep order.  `todo` has one bit per block of the chunk that holds a candidate (bit b = block b);
// the block counter is derived from it, so it is a genuinely per-lane value.  (Written with a plain counter that only
// ever takes warp-uniform VALUES, at lane-dependent TIMES, ptxas 12.9 parked it in a uniform register which partial
// warps then advanced: lanes skipped blocks.)  The device arrays are padded to whole blocks with never-hit spheres, so
// a padding bit that slips through stage 1 (NaN ray) is harmless.
struct Best { double t; int i; };                          // the running closest hit: the f64 root and its object index
#ifndef RTB_REFINE_INLINE
#define RTB_REFINE_INLINE __forceinline__   // two inlined copies (one per ray of the lane) beat one out-of-line copy by 5 %
#endif

// Next stage-1 candidate of the walk.  maddr = shared-memory byte address of the ray's mask word of the
// chunk's block 0 (block b's word lies stride_bytes * b further); todo = bitmap of the blocks that hold candidates; the
// block index comes from that per-lane bitmap, highest block first (the visiting order is free: test_candidate's update
// rule does not depend on it).
__device__ __forceinline__ void walk_next(uint32_t maddr, uint32_t stride_bytes, uint32_t& todo, uint32_t& cand, int& last, int base31, int& idx) {
    // precondition: (cand | todo) != 0.  last = index of sphere 31 of the current block; base31 = that of the chunk's block 0.
    uint32_t hi, bit;
    // Refill without a branch: with ~1.1 candidates per block some lane of the warp needs the next block's mask word on
    // nearly every step, so the branch (ISETP + BSSY + BRA + BSYNC around 8 instructions) was taken anyway; predicated, the
    // refill costs its 8 instructions and no control flow.  bfind(0) = 0xffffffff only when no refill is needed (precondition).
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b32 hb, bb, w, a;\n\t"
        "setp.eq.u32 p, %1, 0;\n\t"
        "bfind.u32 hb, %0;\n\t"
        "bmsk.clamp.b32 bb, hb, 1;\n\t"
        "mad.lo.u32 a, hb, %4, %3;\n\t"
        "@p ld.shared.u32 w, [a];\n\t"
        "@p not.b32 %1, w;\n\t"
        "@p xor.b32 %0, %0, bb;\n\t"
        "@p shl.b32 a, hb, 5;\n\t"
        "@p add.s32 %2, a, %5;\n\t"
        "}"
        : "+r"(todo), "+r"(cand), "+r"(last)
        : "r"(maddr), "r"(stride_bytes), "r"(base31));
    asm("bfind.u32 %0, %1;" : "=r"(hi) : "r"(cand));       // sphere j of the block sits at bit 31 - j
    asm("bmsk.clamp.b32 %0, %1, 1;" : "=r"(bit) : "r"(hi));
    cand ^= bit;
    idx = last - (int)hi;
}

// Stage-2 survivors are not tested on the spot: a lane parks them in ONE register and the warp runs stage 3 once per
// chunk for the survivors of all its rays together (drain_pending).  On the spot, stage 3 ran with ~6 of 32 lanes active,
// 1.5 times per ray and chunk; deferred, the same lane-tests take ~1.7 passes per chunk for both rays.  The order is free
// (test_candidate's rule).  pend = a stack of 9-bit entries (ray << 8 | sphere index within the chunk; 10 bits in the 16-block
// chunks of k_trace_paths_l, Pend<CB>) above a sentinel 1: empty = 1, three entries fill it; a fourth survivor of one lane in one chunk (rare) is tested on the spot — inline in the
// parameter-space kernels (with a CALL inside the chunk loop ptxas moves the loop's counters and the stage-1 sphere operands
// out of the uniform datapath, LDC into vector registers instead of LDCU, and the sweep drops to 74 % issue-slot use: C2
// 54.9 ms), out of line in the kernels that read stage 1 from shared memory (no uniform operands to lose, and the walk
// loop stays small: C4 460 vs 472 ms inline, 785 spheres 12.04 vs 12.37 ms).
constexpr uint32_t PEND_EMPTY = 1u;
// CB = blocks per chunk (8 or 16): an entry is the sphere index within the chunk (8 or 9 bits) + the ray bit
template <int CB> struct Pend { static constexpr int index_bits = CB == 16 ? 9 : 8, entry_bits = index_bits + 1; static constexpr uint32_t full = 1u << (3 * entry_bits); };
// Chunk size of k_trace_paths_l: 16 blocks (512 spheres).  A walk runs as many steps as the fullest lane of the warp holds
// candidates; over a chunk twice as long the lanes' counts even out and half as many walks are set up (10 001 spheres: 20 chunks
// instead of 40): C4 424.1 -> 420.5 ms, 785 / 1 297 / 2 501 spheres 11.18 / 16.98 / 29.45 -> 10.77 / 16.26 / 28.26.  Its 16 mask
// words per ray (82 KB at 640 threads) fit beside the stage-1 data; the parameter-space kernels (four CTAs' path state per SM) and
// the streamed kernel (per-warp double buffers) have room for 8.
#ifndef RTB_L_CHUNK
#define RTB_L_CHUNK 16
#endif
constexpr int L_CHUNK_BLOCKS = RTB_L_CHUNK;

__device__ __noinline__ Best stage3_now(ExactRay ray32, double4 sphere, int idx, Best best) {
    test_candidate(promote_ray(ray32), ray32.both_roots, sphere, idx, best.t, best.i);
    return best;
}

// filter_value_h (rt_filter.h) as the walks evaluate it: the h and g chains as ONE chain of packed FMAs (fma.rn.f32x2 — an
// IEEE fma per element, the same operations in the same order), the sphere coordinate as the scalar-broadcast operand.  With
// the sphere in vector registers (from LDG) every other scalar chain FFMA read three fresh registers and lost a cycle to the
// two-bank register file: T 424.6 -> 419.3 ms, C2 43.42 -> 42.88, C5 115.1 -> 113.8.  fs = {cx, cy, cz, k}.
__device__ __forceinline__ float filter_value_packed(const FilterRay& f, float4 fs, float& h) {
    float2 hg = __ffma2_rn(make_float2(fs.z, fs.z), make_float2(f.hz, f.gz), make_float2(f.nod, f.oo));
    hg = __ffma2_rn(make_float2(fs.y, fs.y), make_float2(f.hy, f.gy), hg);
    hg = __ffma2_rn(make_float2(fs.x, fs.x), make_float2(f.hx, f.gx), hg);
    h = hg.x;
    return __fmaf_rn(h, h, -__fadd_rn(hg.y, fs.w));
}
// One element of SmemSlab::quad (the packed stage 1 of the shared-memory kernels): nrp = -R'', sign bit SET <=> candidate.
__device__ __forceinline__ float slab_value_packed(const SlabRay& s, float ca, float cb, float nrp) {
    const float2 p = __ffma2_rn(make_float2(ca, ca), make_float2(s.ua, s.ua), __ffma2_rn(make_float2(cb, cb), make_float2(s.ub, s.ub), make_float2(s.nou, s.nou)));
    return __ffma2_rn(p, p, make_float2(nrp, nrp)).x;
}

// Stage 2 of the two-stage sweep for ray r and one chunk: the lane-parallel walk over its stage-1 candidates.
template <bool CALL_OVERFLOW, int CB, class RaySrc>
__device__ RTB_REFINE_INLINE void refine_chunk(const uint32_t* masks_r, int stride, uint32_t todo, int first_sphere,
                                          const float4* __restrict__ g_filter, const double4* __restrict__ g_exact, float h_bound, FilterRay f,
                                          const RaySrc& rays, int r, uint32_t& pend, Best& best) {
    if (todo == 0u) return;
    // Near-root mode: a sphere whose centre lies behind the origin along the ray (half_b > 0) cannot give a valid root
    // (test_candidate).  The filter's own h = (c - o).D is that projection in FP32, off by at most 8u(|o| + |c| + r)
    // (rt_filter.h), so h < -16u(|o|_1 + max(|c| + r)) proves half_b > 0 and the f64 test is skipped: for a bounce ray
    // in the middle of the scene that is half of the stage-2 survivors.
    float h_min = -RTB_INF;
    if (!rays.both_roots()) {
        const f3 o = rays.origin(r);
        h_min = -__fmul_rn(9.5367431640625e-07f, __fadd_rn(__fadd_rn(fabsf(o.x), fabsf(o.y)), __fadd_rn(fabsf(o.z), h_bound)));
    }
    uint32_t cand = 0;
    int last = 0, idx = 0;                                 // last = index of the current block's sphere 31
    // the mask words' shared-memory address, held in ONE register for the whole walk (left to itself ptxas rebuilds it from
    // %tid and the shared window base on every block: five instructions per refill)
    uint32_t maddr = (uint32_t)__cvta_generic_to_shared(masks_r);
    asm volatile("" : "+r"(maddr));
    while ((cand | todo) != 0u) {                          // one exit, at the top: no break out of the refill branch
        walk_next(maddr, (uint32_t)stride * 4u, todo, cand, last, first_sphere + (SWEEP_BLOCK - 1), idx);
        const float4 fs = __ldg(&g_filter[idx]);
        float h;
        const float q = filter_value_packed(f, fs, h);                                             // stage 2 (FP32, conservative)
        if (!(__float_as_uint(q) >> 31) && !(h < h_min)) {
            if (pend < Pend<CB>::full) pend = ((pend << Pend<CB>::entry_bits) + (uint32_t)(r << Pend<CB>::index_bits)) | ((uint32_t)idx & ((1u << Pend<CB>::index_bits) - 1u));   // park it for drain_pending (chunks start at multiples of their size)
            else if (CALL_OVERFLOW) best = stage3_now(rays.load(r), load_exact(&g_exact[idx]), idx, best);   // the lane's fourth: now
            else test_candidate(promote_ray(rays.load(r)), rays.both_roots(), load_exact(&g_exact[idx]), idx, best.t, best.i);
        }
    }
}

// Stage 3 (f64, exact) for the parked survivors of all RAYS rays of the lane: one pass per entry of the fullest lane.  The
// ray is fetched and promoted HERE, so that nothing of it is live across the walks (the kernels run at the 64-register limit).
template <int RAYS, int CB, class RaySrc>
__device__ __forceinline__ void drain_pending(uint32_t pend, int first_sphere, const double4* __restrict__ g_exact, const RaySrc& rays,
                                              double (&best_t)[RAYS], int (&best_i)[RAYS]) {
    static_assert(RAYS <= 2, "a pending entry has one ray bit");
    while (pend != PEND_EMPTY) {
        const uint32_t e = pend & ((1u << Pend<CB>::entry_bits) - 1u);
        pend >>= Pend<CB>::entry_bits;
        const int r = (int)(e >> Pend<CB>::index_bits), idx = first_sphere + (int)(e & ((1u << Pend<CB>::index_bits) - 1u));
        const ExactRay ray32 = rays.load(r);
        double bt = best_t[0];
        int bi = best_i[0];
        if (RAYS > 1 && r) { bt = best_t[RAYS - 1]; bi = best_i[RAYS - 1]; }
        test_candidate(promote_ray(ray32), ray32.both_roots, load_exact(&g_exact[idx]), idx, bt, bi);
        if (RAYS > 1 && r) { best_t[RAYS - 1] = bt; best_i[RAYS - 1] = bi; }
        else { best_t[0] = bt; best_i[0] = bi; }
    }
}

// Stage-1 operand sources: evaluate the slab test of one 32-sphere block for the RAYS rays of a lane.
// ParamSlab: {c_a, c_b, R'} triples in kernel parameter space -> LDCU.64 x3 per sphere pair, uniform-register operands.
template <class PV_T>
struct ParamSlab {
    static constexpr bool overflow_by_call = false;   // refine_chunk: keep the chunk loop free of calls (uniform datapath)
    static constexpr int max_tail_groups = 3;         // sweep_two_stage: the last block takes the short path up to 24 existing spheres
    const PV_T& pv;
    template <int RAYS>
    __device__ __forceinline__ void block(int blk, const SlabRay (&sl)[RAYS], uint32_t (&mask)[RAYS]) const {
        const int pair0 = blk * (SWEEP_BLOCK / 2);
#pragma unroll
        for (int j = 0; j < SWEEP_BLOCK / 2; j++) {
            const float2 A = pv.s1[3 * (pair0 + j)], B = pv.s1[3 * (pair0 + j) + 1], C = pv.s1[3 * (pair0 + j) + 2];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = __funnelshift_l(__float_as_uint(slab_value(sl[r], A.x, A.y, B.x)), mask[r], 1);
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = __funnelshift_l(__float_as_uint(slab_value(sl[r], B.y, C.x, C.y)), mask[r], 1);
        }
    }
    // The scene's LAST block when fewer than 25 of its 32 spheres exist (the arrays are padded to whole blocks with
    // never-candidates): only the first `groups` (1..3) groups of 8 spheres are evaluated — a rolled loop over 8-sphere bodies,
    // still uniform control flow — and the rest of the mask word is filled with "not a candidate".  C2/T/C5 have 486/485
    // spheres = 15 blocks + 6 (5) spheres: the 16th block costs 79 instead of 319 instructions, 2.9 % of a loop trip.
    template <int RAYS>
    __device__ __forceinline__ void partial(int blk, int groups, const SlabRay (&sl)[RAYS], uint32_t (&mask)[RAYS]) const {
        const int pair0 = blk * (SWEEP_BLOCK / 2);
#pragma unroll 1
        for (int g = 0; g < groups; g++) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float2 A = pv.s1[3 * (pair0 + g * 4 + j)], B = pv.s1[3 * (pair0 + g * 4 + j) + 1], C = pv.s1[3 * (pair0 + g * 4 + j) + 2];
#pragma unroll
                for (int r = 0; r < RAYS; r++) mask[r] = __funnelshift_l(__float_as_uint(slab_value(sl[r], A.x, A.y, B.x)), mask[r], 1);
#pragma unroll
                for (int r = 0; r < RAYS; r++) mask[r] = __funnelshift_l(__float_as_uint(slab_value(sl[r], B.y, C.x, C.y)), mask[r], 1);
            }
        }
        const int sh = SWEEP_BLOCK - 8 * groups;             // 8, 16 or 24 spheres not evaluated: never candidates
#pragma unroll
        for (int r = 0; r < RAYS; r++) mask[r] = (mask[r] << sh) | ((1u << sh) - 1u);
    }
};

// SmemSlab: the stage-1 data staged in shared memory (scenes too large for parameter space): three LDS.128 broadcasts
// per four spheres, each serving the RAYS rays of the lane.
// Layout per four spheres: {c_a x4}, {c_b x4}, {-R'' x4} (R'' = R' stepped one ulp up, negated, by the host), evaluated with
// the packed FP32 FMA of sm_100 (fma.rn.f32x2: two spheres of one ray per instruction).  With the sphere in VECTOR registers
// (from LDS.128) a scalar chain FFMA reads three distinct registers and loses a cycle to the two-bank register file on
// every other instruction (0.36 dispatch stalls per issue on C4); the packed form reads three 64-bit pairs in the two
// cycles it occupies anyway.  Per element the arithmetic is the scalar one: p = fma(c_a, U_a, fma(c_b, U_b, -o.U)), then
// q' = fma(p, p, -R'') = -(R'' - p^2): sign bit SET <=> candidate, and p^2 <= R' implies p^2 < R'' (exact difference
// negative, so the rounded one is negative and non-zero): every candidate of the scalar test is a candidate here.
struct SmemSlab {
    static constexpr bool overflow_by_call = true;
    static constexpr int max_tail_groups = 2;
    const float4* s;
    template <int RAYS>
    __device__ __forceinline__ void quad(const float4 A, const float4 B, const float4 C, const SlabRay (&sl)[RAYS], uint32_t (&mask)[RAYS]) const {
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            const float2 ua = make_float2(sl[r].ua, sl[r].ua), ub = make_float2(sl[r].ub, sl[r].ub), nou = make_float2(sl[r].nou, sl[r].nou);
            const float2 p01 = __ffma2_rn(make_float2(A.x, A.y), ua, __ffma2_rn(make_float2(B.x, B.y), ub, nou));
            const float2 p23 = __ffma2_rn(make_float2(A.z, A.w), ua, __ffma2_rn(make_float2(B.z, B.w), ub, nou));
            const float2 q01 = __ffma2_rn(p01, p01, make_float2(C.x, C.y)), q23 = __ffma2_rn(p23, p23, make_float2(C.z, C.w));
            mask[r] = __funnelshift_l(__float_as_uint(q01.x), mask[r], 1);
            mask[r] = __funnelshift_l(__float_as_uint(q01.y), mask[r], 1);
            mask[r] = __funnelshift_l(__float_as_uint(q23.x), mask[r], 1);
            mask[r] = __funnelshift_l(__float_as_uint(q23.y), mask[r], 1);
        }
    }
    template <int RAYS>
    __device__ __forceinline__ void block(int blk, const SlabRay (&sl)[RAYS], uint32_t (&mask)[RAYS]) const {
        const float4* q = s + blk * (SWEEP_BLOCK / 4 * 3);
#pragma unroll
        for (int j = 0; j < SWEEP_BLOCK / 4; j++) quad<RAYS>(q[3 * j], q[3 * j + 1], q[3 * j + 2], sl, mask);
#pragma unroll
        for (int r = 0; r < RAYS; r++) mask[r] = ~mask[r];     // the callers' polarity: bit set = not a candidate
    }
    // the scene's last, mostly empty block (see ParamSlab::partial): `groups` (1..3) groups of 8 spheres
    template <int RAYS>
    __device__ __forceinline__ void partial(int blk, int groups, const SlabRay (&sl)[RAYS], uint32_t (&mask)[RAYS]) const {
        const float4* q = s + blk * (SWEEP_BLOCK / 4 * 3);
#pragma unroll 1
        for (int j = 0; j < 2 * groups; j++) quad<RAYS>(q[3 * j], q[3 * j + 1], q[3 * j + 2], sl, mask);
        const int sh = SWEEP_BLOCK - 8 * groups;
#pragma unroll
        for (int r = 0; r < RAYS; r++) mask[r] = (~mask[r] << sh) | ((1u << sh) - 1u);
    }
};

template <int RAYS, int CB, class S1, class RaySrc>
__device__ __forceinline__ void sweep_two_stage(const S1& s1, const RaySrc& rays, const float4* __restrict__ g_filter,
                                                const double4* __restrict__ g_exact, float h_bound, int count, uint32_t* masks, int stride,
                                                const SlabRay (&sl)[RAYS], const FilterRay (&f)[RAYS], double (&best_t)[RAYS], int (&best_i)[RAYS]) {
    const int n_blocks = (count + SWEEP_BLOCK - 1) / SWEEP_BLOCK;     // the stage-1 array is padded to whole blocks
    // groups of 8 spheres that exist in the scene's last block; 1..3: that block takes the short path (S1::partial)
    const int tail_groups = ((count & (SWEEP_BLOCK - 1)) + 7) >> 3;
    const bool short_tail = tail_groups >= 1 && tail_groups <= S1::max_tail_groups;
    for (int cb = 0; cb < n_blocks; cb += CB) {
        const int nb = n_blocks - cb < CB ? n_blocks - cb : CB;
        const bool tail_here = short_tail && cb + nb == n_blocks;
        const int nb_full = tail_here ? nb - 1 : nb;
        uint32_t todo[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) todo[r] = 0u;
        // ---- stage 1 ----
        uint32_t bit = 1u;
#pragma unroll 1
        for (int b = 0; b < nb_full; b++, bit <<= 1) {
            uint32_t mask[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = 0;
            s1.template block<RAYS>(cb + b, sl, mask);
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                masks[(r * CB + b) * stride] = mask[r];
                todo[r] |= mask[r] != 0xffffffffu ? bit : 0u;
            }
        }
        if (tail_here) {
            uint32_t mask[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = 0;
            s1.template partial<RAYS>(cb + nb_full, tail_groups, sl, mask);
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                masks[(r * CB + nb_full) * stride] = mask[r];
                todo[r] |= mask[r] != 0xffffffffu ? bit : 0u;
            }
        }
        // ---- stages 2 and 3 ----
        uint32_t pend = PEND_EMPTY;
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            Best bst{best_t[r], best_i[r]};
            refine_chunk<S1::overflow_by_call, CB>(masks + r * CB * stride, stride, todo[r], cb * SWEEP_BLOCK, g_filter, g_exact, h_bound, f[r], rays, r, pend, bst);
            best_t[r] = bst.t;
            best_i[r] = bst.i;
        }
        drain_pending<RAYS, CB>(pend, cb * SWEEP_BLOCK, g_exact, rays, best_t, best_i);
    }
}

// The two-stage sweep with the stage-1 triples STREAMED from global memory (L2-resident) through a per-warp double
// buffer: while the warp sweeps chunk c (8 blocks = 256 spheres = 3 KB of triples), `cp.async` brings chunk c + 1 into
// the other half of its 6 KB buffer.  No CTA barrier and no CTA-wide staging, so the scene size is unbounded and the
// warps of a CTA stay independent (the persistent loop's trips are not in lockstep).  A chunk sweep is ~2 250 issue
// slots of this warp, i.e. >= 10^4 cycles with five warps per sub-partition: the L2 latency is hidden many times over,
// and the fill traffic is 3 KB per chunk and warp (~5.5 B/clk/SM).
__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gmem_src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

constexpr int STREAM_CHUNK_F4 = 8 * SWEEP_BLOCK * 3 / 4;     // float4 per chunk of 8 blocks: 192 (3 KB)

template <int RAYS, class RaySrc>
__device__ __forceinline__ void sweep_two_stage_stream(float4* wbuf, const float4* __restrict__ g_s1, unsigned lane, const RaySrc& rays,
                                                       const float4* __restrict__ g_filter, const double4* __restrict__ g_exact, float h_bound, int n_padded,
                                                       uint32_t* masks, int stride, const SlabRay (&sl)[RAYS], const FilterRay (&f)[RAYS],
                                                       double (&best_t)[RAYS], int (&best_i)[RAYS]) {
    const int n_blocks = n_padded / SWEEP_BLOCK;              // the device arrays are padded to whole blocks
    const int n_chunks = (n_blocks + 7) / 8;
    auto issue = [&](int c) {                                 // chunk c -> half (c & 1) of the warp's buffer, one commit group
        const int nb = n_blocks - c * 8 < 8 ? n_blocks - c * 8 : 8;
        const float4* src = g_s1 + (size_t)c * STREAM_CHUNK_F4;
        float4* dst = wbuf + (c & 1) * STREAM_CHUNK_F4;
        for (int i = (int)lane; i < nb * (SWEEP_BLOCK * 3 / 4); i += 32) cp_async_16(dst + i, src + i);
        cp_async_commit();
    };
    issue(0);
    for (int c = 0; c < n_chunks; c++) {
        if (c + 1 < n_chunks) issue(c + 1);
        else cp_async_commit();                               // an empty group keeps "all but the newest" = chunk c
        cp_async_wait<1>();
        __syncwarp();                                         // every lane's part of chunk c has landed
        const int nb = n_blocks - c * 8 < 8 ? n_blocks - c * 8 : 8;
        const SmemSlab s1{wbuf + (c & 1) * STREAM_CHUNK_F4};
        uint32_t todo[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) todo[r] = 0u;
        uint32_t bit = 1u;
#pragma unroll 1
        for (int b = 0; b < nb; b++, bit <<= 1) {             // ---- stage 1 ----
            uint32_t mask[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = 0;
            s1.template block<RAYS>(b, sl, mask);
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                masks[(r * 8 + b) * stride] = mask[r];
                todo[r] |= mask[r] != 0xffffffffu ? bit : 0u;
            }
        }
        __syncwarp();                                         // all lanes are done reading this half before chunk c + 2 refills it
        uint32_t pend = PEND_EMPTY;
#pragma unroll
        for (int r = 0; r < RAYS; r++) {                      // ---- stages 2 and 3 ----
            Best bst{best_t[r], best_i[r]};
            refine_chunk<true, 8>(masks + r * 8 * stride, stride, todo[r], c * 8 * SWEEP_BLOCK, g_filter, g_exact, h_bound, f[r], rays, r, pend, bst);
            best_t[r] = bst.t;
            best_i[r] = bst.i;
        }
        drain_pending<RAYS, 8>(pend, c * 8 * SWEEP_BLOCK, g_exact, rays, best_t, best_i);
    }
    cp_async_wait<0>();
}

// ---- K1 ----------------------------------------------------------------------------------------------------------
// One thread per ray; spheres streamed through shared memory in tiles of `tile` (multiple of SWEEP_BLOCK).
// dynamic smem: tile float4 + LIST_CAP * blockDim uint2.
__global__ void __launch_bounds__(256) k_intersect_batch(const double4* __restrict__ g_sph, const float4* __restrict__ g_flt, int n_spheres, int tile,
                                                         const float* __restrict__ ox, const float* __restrict__ oy,
                                                         const float* __restrict__ oz, const float* __restrict__ dx,
                                                         const float* __restrict__ dy, const float* __restrict__ dz,
                                                         const uint8_t* __restrict__ enabled, long long n_rays,
                                                         int* __restrict__ hit_index, float* __restrict__ hit_t, int both_roots) {
    extern __shared__ float4 smem4[];
    float4* s_sph = smem4;
    uint2* list = reinterpret_cast<uint2*>(smem4 + tile) + threadIdx.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    bool live = i < n_rays;
    f3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f);
    if (live) { o = mk3(ox[i], oy[i], oz[i]); d = mk3(dx[i], dy[i], dz[i]); }
    ExactRay ray[1] = {make_exact_ray(o, d, both_roots != 0)};
    FilterRay flt[1] = {make_filter(o, d)};
    double best_t[1] = {RTB_INF_D};                            // PackedHitRecords::default, src/objects.rs:121-134
    int best_i[1] = {-1};
    for (int begin = 0; begin < n_spheres; begin += tile) {
        int count = min(tile, n_spheres - begin);
        __syncthreads();
        stage_spheres(s_sph, g_flt, begin, count);
        __syncthreads();
        sweep_spheres<1, LIST_CAP, true>(SmemSpheres{s_sph}, RegisterRays{ray}, g_sph + begin, count, begin, list, blockDim.x, flt, best_t, best_i);
    }
    if (live) {
        bool en = enabled ? enabled[i] != 0 : true;            // rays.enabled(), src/objects.rs:259,277
        hit_index[i] = en ? best_i[0] : -1;
        hit_t[i] = (en && best_i[0] >= 0) ? __double2float_rn(best_t[0]) : RTB_INF;   // the f64 root, rounded once
    }
}

// ---- K1d: the reference's routine at the reference's precision ---------------------------------------------------------
// Sphere::hit_packed + PackedHitRecords::update (src/objects.rs:249-290, :140-155) for one ray per thread in f64 — the
// reference's own scalar type (src/geometry.rs:9-13) — over f64 spheres {cx, cy, cz, r*r} in list order, EVERY pair, no
// filter.  Same operation order and FMA sites as the reference (mul_add at src/geometry.rs:435,467 and src/objects.rs:257),
// IEEE sqrt and division: the result equals the f64 routine bit for bit, for f64-valued rays.  The production sweep
// (K1, every K2 variant) runs the same test only on the candidates its FP32 filters let through; this kernel is the
// filter-free yardstick the tests hold it against (same index, t equal after one rounding to FP32), and the call for a
// caller whose rays are f64.  Spheres go through shared memory in tiles of 256 (8 KB).
__global__ void __launch_bounds__(256) k_intersect_batch_f64(const double4* __restrict__ g_sph, int n_spheres, const double* __restrict__ ox,
                                                             const double* __restrict__ oy, const double* __restrict__ oz,
                                                             const double* __restrict__ dx, const double* __restrict__ dy,
                                                             const double* __restrict__ dz, const uint8_t* __restrict__ enabled, long long n_rays,
                                                             int* __restrict__ hit_index, double* __restrict__ hit_t, int both_roots) {
    __shared__ double4 s_sph[256];
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < n_rays;
    double rox = 0.0, roy = 0.0, roz = 0.0, rdx = 0.0, rdy = 0.0, rdz = 1.0;
    if (live) { rox = ox[i]; roy = oy[i]; roz = oz[i]; rdx = dx[i]; rdy = dy[i]; rdz = dz[i]; }
    const double T_MIN_D = 0.001, INF_D = __longlong_as_double(0x7ff0000000000000LL);
    const double a = __fma_rn(rdz, rdz, __fma_rn(rdy, rdy, __dmul_rn(rdx, rdx)));                   // src/objects.rs:253
    const double inv_a = __ddiv_rn(1.0, a);                                                         // :254
    double best_t = INF_D;                                                                          // PackedHitRecords::default, :121-134
    int best_i = -1;
    for (int begin = 0; begin < n_spheres; begin += 256) {
        const int count = min(256, n_spheres - begin);
        __syncthreads();
        if ((int)threadIdx.x < count) s_sph[threadIdx.x] = g_sph[begin + threadIdx.x];
        __syncthreads();
        for (int j = 0; j < count; j++) {
            const double4 sp = s_sph[j];
            const double ocx = __dsub_rn(rox, sp.x), ocy = __dsub_rn(roy, sp.y), ocz = __dsub_rn(roz, sp.z);          // :252
            const double hb = __fma_rn(ocz, rdz, __fma_rn(ocy, rdy, __dmul_rn(ocx, rdx)));                            // :255
            const double cc = __dsub_rn(__fma_rn(ocz, ocz, __fma_rn(ocy, ocy, __dmul_rn(ocx, ocx))), sp.w);   // :256 (sp.w = radius.powi(2), host f64)
            const double disc = __fma_rn(hb, hb, -__dmul_rn(a, cc));                                                  // :257
            if (disc >= 0.0) {                                                                                        // :259
                const double sq = __dsqrt_rn(disc);                                                                   // :263
                double root1 = __dmul_rn(__dsub_rn(-hb, sq), inv_a);                                                  // :270
                bool valid = (root1 >= T_MIN_D) && (root1 < INF_D);                                                   // :272-277
                if (!valid && both_roots) {                                                                           // scalar rule, :230-234
                    root1 = __dmul_rn(__dadd_rn(-hb, sq), inv_a);                                                     // :271
                    valid = (root1 >= T_MIN_D) && (root1 < INF_D);
                }
                if (valid && root1 <= best_t) { best_t = root1; best_i = begin + j; }                                 // :141
            }
        }
    }
    if (live) {
        const bool en = enabled ? enabled[i] != 0 : true;                                           // rays.enabled(), :259,277
        hit_index[i] = en ? best_i : -1;
        hit_t[i] = (en && best_i >= 0) ? best_t : INF_D;
    }
}

// ---- K2 ----------------------------------------------------------------------------------------------------------
struct TraceParams {
    DeviceCamera cam;
    const float4* spheres;        // FP32 {cx,cy,cz,r*r}: shading (finalize_hit)
    const float4* filter;         // candidate filter: {cx,cy,cz,k} (rt_filter.h): the one-stage sweeps, stage 2 of the two-stage sweep
    const double4* exact;         // f64 {cx,cy,cz,r*r}: the exact candidate test (test_candidate)
    float h_bound;                // max over the scene of |c| + r, rounded up (refine_chunk's behind-the-origin rejection)
    const float4* mat;            // per sphere {albedo rgb, param}
    const int* mat_kind;          // per sphere kind | hollow<<8
    int n_spheres;
    int max_bounces;
    uint32_t seed_lo, seed_hi;
    int sample_begin, sample_end; // global sample indices traced by this launch
    int row_begin, row_end;       // image rows traced by this launch: row_begin, row_begin + row_stride, ... < row_end
    int row_stride;               // 1 = a contiguous band; R = every R-th scanline (interleaved row shards)
    int claim_chunk;              // work items a warp takes from the queue per atomic (1..64; slot_step)
    int both_roots;               // hit mode: 0 = near root only (the reference's packed path), 1 = scalar Sphere::hit rule
    int splits_log2;              // K = 1 << splits_log2: the sample range is cut into K contiguous sub-ranges, one work item each
    uint32_t width_magic;         // floor(2^32 / W) + 1 (0 when W == 1): row = pixel / W by multiply-high
    float4* out;                  // K == 1: accum image (W*H float4); else partial planes [K][rows*W]
    unsigned long long* work_counter;
    unsigned long long* progress; // mapped pinned host memory (or null): the queue head, published every 4096 items
    unsigned long long* counters; // [0] paths, [1] segments
};

// TraceParams + the filter spheres themselves in the kernel's parameter space (constant bank 0), for scenes of at most
// N spheres: the sweep then takes its sphere operand from uniform registers (ParamSpheres).
template <int N>
struct TraceParamsV {
    static constexpr bool two_stage = false;
    static constexpr int capacity = N;
    TraceParams b;
    float4 flt[N];
};

// TraceParams + the STAGE-1 (slab test, rt_filter.h) spheres in parameter space: three floats {c_a, c_b, R'} per sphere
// with the scene's `up` axis dropped, two spheres per three float2 (LDCU.64 x3 per pair), padded to a multiple of 32
// spheres with never-candidates (R' = -inf).  Up to N = 512 spheres (the triples must stay resident in the constant cache, see launch_trace).
template <int N>
struct TraceParamsS {
    static constexpr bool two_stage = true;
    static constexpr bool s1_smem = false;
    static constexpr int capacity = N;
    TraceParams b;
    int up_axis;
    float rp_min;                 // the scene's smallest (floored) R', make_slab_ray's scale
    int _pad[2];
    float2 s1[N * 3 / 2];
};

// TraceParams + a pointer to the stage-1 triples in global memory: the kernel stages them in SHARED memory (scenes of
// more than 512 spheres, up to what fits beside the path state: ~10 600 spheres).
struct TraceParamsL {
    static constexpr bool two_stage = true;
    static constexpr bool s1_smem = true;
    static constexpr int capacity = 1 << 30;
    TraceParams b;
    int up_axis;
    float rp_min;
    int n_padded;                 // spheres rounded up to whole 32-blocks
    int _pad;
    const float4* s1;             // n_padded * 3 floats, 16-byte aligned
};

template <int N>
struct ParamSpheres {
    const TraceParamsV<N>& pv;
    __device__ __forceinline__ float4 at(int i) const { return pv.flt[i]; }
};

constexpr uint32_t NO_ITEM = 0xffffffffu;

// One path slot of a lane.  A slot owns one work item (a pixel and a contiguous range of its samples) at a time and
// runs the samples one after another; the per-pixel sum `acc` is a private sum in sample order, so the image is
// deterministic and needs no atomics.  item = pixel * K + split (NO_ITEM: the slot is free).
struct PathSlot {
    f3 o, d;                      // current ray (direction NOT normalised after scatter, src/ray.rs:13-18)
    f3 thr, acc;                  // path throughput (`color`, src/ray_tracing.rs:319) and the pixel's running sum
    uint32_t item;
    int s;                        // next sample (global index)
    int bounce;
    bool has_path;
};

__device__ __forceinline__ void slot_reset(PathSlot& q) {
    q.o = mk3(0.f, 0.f, 0.f); q.d = mk3(0.f, 0.f, 1.f);
    q.thr = mk3(1.f, 1.f, 1.f); q.acc = mk3(0.f, 0.f, 0.f);
    q.item = NO_ITEM; q.s = 0; q.bounce = 0; q.has_path = false;
}

// Launch-wide constants derived from TraceParams once per thread.
struct TraceConsts {
    int W;
    uint32_t n_pix, n_items, n_samples, first_pixel, split_mask;
};

template <bool STRIDED>
__device__ __forceinline__ TraceConsts trace_consts(const TraceParams& P) {
    TraceConsts c;
    c.W = P.cam.width;
    if constexpr (STRIDED) c.n_pix = (uint32_t)((P.row_end - P.row_begin + P.row_stride - 1) / P.row_stride) * (uint32_t)c.W;   // traced rows x W
    else c.n_pix = (uint32_t)(P.row_end - P.row_begin) * (uint32_t)c.W;
    c.n_items = c.n_pix << P.splits_log2;                     // host guarantees < 2^32
    c.n_samples = (uint32_t)(P.sample_end - P.sample_begin);
    c.first_pixel = (uint32_t)P.row_begin * (uint32_t)c.W;
    c.split_mask = (1u << P.splits_log2) - 1u;
    return c;
}

// Sample range [begin, end) of split `split` of an item.
__device__ __forceinline__ int split_sample(const TraceParams& P, const TraceConsts& C, uint32_t split) {
    return P.sample_begin + (int)(((unsigned long long)split * C.n_samples) >> P.splits_log2);
}

// Local pixel of a work item.  The queue runs through the image BOTTOM-UP (item 0 = the last pixel): the launch ends
// when its slowest path ends (up to max_bounces loop trips after the queue ran dry), and for an upright camera the top
// rows are sky — one-segment paths — so the expensive pixels go first and the drain is made of the cheapest work.
#ifndef RTB_QUEUE_TOP_DOWN
__device__ __forceinline__ uint32_t item_pixel(const TraceParams& P, const TraceConsts& C, uint32_t item) {
    return C.n_pix - 1u - (item >> P.splits_log2);
}
#else
__device__ __forceinline__ uint32_t item_pixel(const TraceParams& P, const TraceConsts&, uint32_t item) { return item >> P.splits_log2; }
#endif

// Image pixel index (row * W + col) of the launch's local pixel `pix` (traced rows are numbered consecutively).
// STRIDED is a template parameter of the trace kernels: the contiguous-band instantiations (every single-GPU render)
// carry none of this code (as a run-time branch it cost 1.3-2 % of C2: 44.75 -> 45.3-45.7 ms).
template <bool STRIDED>
__device__ __forceinline__ uint32_t image_pixel(const TraceParams& P, const TraceConsts& C, uint32_t pix) {
    uint32_t gp = C.first_pixel + pix;
    if constexpr (STRIDED) {
        uint32_t lrow = P.width_magic ? __umulhi(pix, P.width_magic) : pix;   // pix / W
        if (lrow * (uint32_t)C.W > pix) lrow--;
        gp += lrow * (uint32_t)(P.row_stride - 1) * (uint32_t)C.W;
    }
    return gp;
}

// One loop trip of one slot, steps 1-3 (warp-collective: every lane of the warp calls it together):
//   1. after a sweep, terminate the path if it escaped (sky) or reached the depth limit (black);
//   2. write a finished item out and claim a new one (one atomic per warp);
//   3. ONE Philox block: the next camera sample (Camera::get_ray) or this bounce's scatter
//      (finalize + Material::get_hit_result).
#ifdef RTB_TIMELINE
// Debug builds only (tools/gpu_timeline.sh): when did the warps of a launch start, see the queue run dry, and exit?
// g_timeline: [0] min start, [1] max start, [2] min dry, [3] max dry, [4] min exit, [5] max exit (globaltimer ns),
// [8 + b]: warps whose exit falls into the b-th 50 us bin BEFORE the last exit is not known on the device, so the bins
// count exits by absolute time >> 16 (65.5 us) modulo 1024 from the launch's min start (host subtracts).
__device__ unsigned long long g_timeline[8 + 4096];
__device__ __forceinline__ unsigned long long tl_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ void tl_mark(int slot, unsigned lane) {
    if (lane == 0) { const unsigned long long t = tl_now(); atomicMin(&g_timeline[slot], t); atomicMax(&g_timeline[slot + 1], t); }
}
__device__ __forceinline__ void tl_exit(unsigned lane) {
    if (lane == 0) { const unsigned long long t = tl_now(); atomicMin(&g_timeline[4], t); atomicMax(&g_timeline[5], t);
                     atomicAdd(&g_timeline[8 + ((t >> 15) & 4095)], 1ull); }
}
#define RTB_TL_START(lane) tl_mark(0, lane)
#define RTB_TL_DRY(lane) tl_mark(2, lane)
#define RTB_TL_EXIT(lane) tl_exit(lane)
#else
#define RTB_TL_START(lane)
#define RTB_TL_DRY(lane)
#define RTB_TL_EXIT(lane)
#endif

struct WarpChunk { uint32_t next = 0, end = 0; };   // the warp's private slice of the work queue (warp-uniform values)

template <bool STRIDED>
__device__ __forceinline__ void slot_step(const TraceParams& P, const TraceConsts& C, PathSlot& q, bool swept, float best_t, int best_i,
                                          bool& exhausted, WarpChunk& chunk, unsigned& n_segments, unsigned lane) {
    // ---- 1. terminate (src/ray_tracing.rs:353-357, :365-370) ----
    if (swept && q.has_path) {
        n_segments++;
        if (best_i < 0) {                              // hit the sky
            f3 sk = sky(q.d);
            q.acc = add3(q.acc, mk3(__fmul_rn(q.thr.x, sk.x), __fmul_rn(q.thr.y, sk.y), __fmul_rn(q.thr.z, sk.z)));
            q.has_path = false;
            q.s++;
        } else if (q.bounce + 1 >= P.max_bounces) {    // still enabled after the last bounce => black (:370)
            q.has_path = false;
            q.s++;
        }
    }
    // ---- 2. finished items are written out; new items are claimed with one atomic per warp ----
    if (q.item != NO_ITEM && !q.has_path) {
        const uint32_t split = q.item & C.split_mask, pix = item_pixel(P, C, q.item);
        const int s_end = split_sample(P, C, split + 1);
        if (q.s >= s_end) {
            const int n_s = s_end - split_sample(P, C, split);
            const uint32_t out_off = (P.splits_log2 == 0) ? image_pixel<STRIDED>(P, C, pix) : split * C.n_pix + pix;
            P.out[out_off] = make_float4(q.acc.x, q.acc.y, q.acc.z, (float)n_s);
            q.item = NO_ITEM;
        }
    }
    const unsigned need_mask = __ballot_sync(0xffffffffu, q.item == NO_ITEM && !exhausted);
    if (need_mask) {
        // The warp takes P.claim_chunk consecutive items from the global queue at a time and hands them to its lanes as
        // they fall free (chunk.next/end are warp-uniform): one atomic per chunk, and the lanes of a warp keep working
        // on neighbouring pixels, whose coherent rays keep the sweep's refine walks and the shading branches in step.
        uint32_t avail = chunk.end - chunk.next;
        if (avail == 0u) {
            unsigned long long base64 = 0;
            if (lane == 0) {
                base64 = atomicAdd(P.work_counter, (unsigned long long)P.claim_chunk);
                // progress for the host (rtb200_set_progress): every 4096 items the new queue head goes to mapped host memory
                if (P.progress && ((base64 ^ (base64 + P.claim_chunk)) >> 12))
                    *reinterpret_cast<volatile unsigned long long*>(P.progress) = base64 + P.claim_chunk;
            }
            base64 = __shfl_sync(0xffffffffu, base64, 0);
            // n_items < 2^32 - 65536: a saturated head stays beyond the end
            const uint32_t base = base64 > 0xffffffffull ? 0xffffffffu : (uint32_t)base64;
            if (base < C.n_items) {
                chunk.next = base;
                chunk.end = C.n_items - base < (uint32_t)P.claim_chunk ? C.n_items : base + (uint32_t)P.claim_chunk;
            } else {
                exhausted = true;                      // the queue is empty: warp-uniform
                RTB_TL_DRY(lane);
            }
            avail = chunk.end - chunk.next;
        }
        if (q.item == NO_ITEM && !exhausted) {
            const uint32_t rank = __popc(need_mask & ((1u << lane) - 1u));
            if (rank < avail) {                        // item = pixel * K + split (lanes beyond `avail` claim next trip)
                const uint32_t item = chunk.next + rank;
                q.item = item;
                q.s = split_sample(P, C, item & C.split_mask);
                q.acc = mk3(0.f, 0.f, 0.f);
                if (P.max_bounces <= 0)                // `for _ in 0..0`: every ray stays enabled => black (:370)
                    q.s = split_sample(P, C, (item & C.split_mask) + 1);
            }
        }
        const uint32_t n_need = (uint32_t)__popc(need_mask);
        chunk.next += n_need < avail ? n_need : avail;
    }
    // ---- 3. one Philox block: the next camera sample, or this bounce's scatter ----
    bool regen = false;
    uint32_t pixel_id = 0;
    if (q.item != NO_ITEM) {
        pixel_id = image_pixel<STRIDED>(P, C, item_pixel(P, C, q.item));
        regen = !q.has_path && q.s < split_sample(P, C, (q.item & C.split_mask) + 1);
    }
    if (regen || q.has_path) {
        uint4 rnd = philox4x32_10(pixel_id, (uint32_t)q.s, regen ? RTB_DIM_CAMERA : 1u + (uint32_t)q.bounce, 0u, P.seed_lo, P.seed_hi);
        if (regen) {                                   // Camera::get_ray, src/ray_tracing.rs:77-89
            uint32_t rowu = P.width_magic ? __umulhi(pixel_id, P.width_magic) : pixel_id;   // pixel / W
            if (rowu * (uint32_t)C.W > pixel_id) rowu--;                                    // estimate is q or q+1
            int row = (int)rowu, col = (int)(pixel_id - rowu * (uint32_t)C.W);
            camera_get_ray(P.cam, col, row, rnd, q.o, q.d);
            q.thr = mk3(1.f, 1.f, 1.f);
            q.bounce = 0;
            q.has_path = true;
        } else {                                       // finalize + Material::get_hit_result (src/ray_tracing.rs:335-346)
            float4 sp = __ldg(&P.spheres[best_i]);
            HitRec rec = finalize_hit(q.o, q.d, best_t, mk3(sp.x, sp.y, sp.z));
            f3 att, nd;
            scatter(__ldg(&P.mat_kind[best_i]), __ldg(&P.mat[best_i]), q.d, rec, rnd, att, nd);
            q.thr = mk3(__fmul_rn(q.thr.x, att.x), __fmul_rn(q.thr.y, att.y), __fmul_rn(q.thr.z, att.z));
            q.o = rec.p;
            q.d = nd;
            q.bounce++;
        }
    }
}

// Adds a warp's per-lane segment counts to the launch counters (one atomic per warp).  The number of PATHS needs no
// counting: every sample of every traced pixel starts exactly one path (src/renderer.rs:148-160), so the launch's first
// thread adds n_pix x n_samples (a per-lane path counter cost a register the kernels do not have to spare).
__device__ __forceinline__ void flush_counters(const TraceParams& P, const TraceConsts& C, unsigned n_segments, unsigned lane) {
    unsigned long long seg64 = n_segments;
#pragma unroll
    for (int sh = 16; sh > 0; sh >>= 1) seg64 += __shfl_xor_sync(0xffffffffu, seg64, sh);
    if (lane == 0) atomicAdd(&P.counters[1], seg64);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&P.counters[0], (unsigned long long)C.n_pix * C.n_samples);
}

// K2, resident variant: path state in registers, filter spheres staged in shared memory.  Any scene that fits the
// staging (up to ~11 000 spheres with one 1024-thread CTA per SM).
// Persistent kernel: grid = SMs x resident CTAs.  Every loop trip runs slot_step for every slot of the lane, then makes
// ONE closest-hit sweep for all slots of the warp.
// dynamic smem: n_spheres float4 + RAYS * CAP * THREADS uint2.  CAP = LIST_CAP_SMALL needs n_spheres <= 32 * CAP.
template <int THREADS, int MIN_BLOCKS, int RAYS, int CAP, bool STRIDED>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS) k_trace_paths(const TraceParams P) {
    extern __shared__ float4 smem4[];
    float4* s_sph = smem4;
    uint2* list = reinterpret_cast<uint2*>(smem4 + ((P.n_spheres + 1) & ~1)) + threadIdx.x;
    stage_spheres(s_sph, P.filter, 0, P.n_spheres);
    __syncthreads();

    const TraceConsts C = trace_consts<STRIDED>(P);
    const unsigned lane = threadIdx.x & 31u;
    PathSlot slot[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) slot_reset(slot[r]);
    bool exhausted = false;
    WarpChunk chunk;
    unsigned n_segments = 0;                       // per-lane count (a lane traces far fewer than 2^32 segments)
    float best_t[RAYS];
    int best_i[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) { best_t[r] = RTB_INF; best_i[r] = -1; }
    bool swept = false;

    for (;;) {
        bool any_path = false, idle = true;
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            slot_step<STRIDED>(P, C, slot[r], swept, best_t[r], best_i[r], exhausted, chunk, n_segments, lane);
            any_path = any_path || slot[r].has_path;
            idle = idle && slot[r].item == NO_ITEM;
        }
        if (!__any_sync(0xffffffffu, any_path)) {
            if (__all_sync(0xffffffffu, idle && exhausted)) break;
            swept = false;
            continue;                                          // items with an empty sample range: write out, claim again
        }
        // the closest-hit sweep (all lanes; a slot without a path sweeps its stale ray and the result is ignored)
        ExactRay ray[RAYS];
        FilterRay flt[RAYS];
        double bt[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            ray[r] = make_exact_ray(slot[r].o, slot[r].d, P.both_roots != 0);
            flt[r] = make_filter(slot[r].o, slot[r].d);
            bt[r] = RTB_INF_D;                                 // PackedHitRecords::default, src/objects.rs:121-134
            best_i[r] = -1;
        }
        sweep_spheres<RAYS, CAP, (CAP < LIST_CAP_SMALL)>(SmemSpheres{s_sph}, RegisterRays{ray}, P.exact, P.n_spheres, 0, list, THREADS, flt,
                                                         bt, best_i);
#pragma unroll
        for (int r = 0; r < RAYS; r++) best_t[r] = __double2float_rn(bt[r]);   // the f64 root, rounded once for the FP32 path state
        swept = true;
    }
    flush_counters(P, C, n_segments, lane);
}

// K2, large scenes: two-stage sweep with the stage-1 triples staged in shared memory and the path state of the RAYS
// slots in REGISTERS (one CTA per SM has registers to spare: 65536 / THREADS per thread), so that shared memory holds
// only the triples and the stage-1 masks and THREADS can be larger than in the shared-state variant: more warps to hide
// the refine walks, whose sphere records come from L2 here (the L1 is all but given to shared memory).
// dynamic smem: L_CHUNK_BLOCKS * RAYS * THREADS mask words + n_padded * 3 floats.
template <int THREADS, int RAYS, bool STRIDED>
__global__ void __launch_bounds__(THREADS, 1) k_trace_paths_l(const __grid_constant__ TraceParamsL PV) {
    const TraceParams& P = PV.b;
    extern __shared__ float4 smem4[];
    uint32_t* masks = reinterpret_cast<uint32_t*>(smem4) + threadIdx.x;
    float4* s_slab = smem4 + L_CHUNK_BLOCKS * RAYS * THREADS / 4;
    {
        const int n4 = PV.n_padded * 3 / 4;
        for (int i = threadIdx.x; i < n4; i += THREADS) s_slab[i] = __ldg(&PV.s1[i]);
        __syncthreads();
    }
    const TraceConsts C = trace_consts<STRIDED>(P);
    const unsigned lane = threadIdx.x & 31u;
    PathSlot slot[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) slot_reset(slot[r]);
    bool exhausted = false;
    WarpChunk chunk;
    unsigned n_segments = 0;
    float best_t[RAYS];
    int best_i[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) { best_t[r] = RTB_INF; best_i[r] = -1; }
    bool swept = false;
    for (;;) {
        bool any_path = false, idle = true;
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            slot_step<STRIDED>(P, C, slot[r], swept, best_t[r], best_i[r], exhausted, chunk, n_segments, lane);
            any_path = any_path || slot[r].has_path;
            idle = idle && slot[r].item == NO_ITEM;
        }
        if (!__any_sync(0xffffffffu, any_path)) {
            if (__all_sync(0xffffffffu, idle && exhausted)) break;
            swept = false;
            continue;
        }
        ExactRay ray[RAYS];
        FilterRay flt[RAYS];
        SlabRay slab[RAYS];
        double bt[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            ray[r] = make_exact_ray(slot[r].o, slot[r].d, P.both_roots != 0);
            flt[r] = make_filter(slot[r].o, slot[r].d);
            slab[r] = make_slab_ray(slot[r].o.x, slot[r].o.y, slot[r].o.z, slot[r].d.x, slot[r].d.y, slot[r].d.z, plen2_3(slot[r].d), PV.up_axis,
                                    PV.rp_min);
            bt[r] = RTB_INF_D;
            best_i[r] = -1;
        }
        sweep_two_stage<RAYS, L_CHUNK_BLOCKS>(SmemSlab{s_slab}, RegisterRays{ray}, P.filter, P.exact, P.h_bound, P.n_spheres, masks, THREADS, slab, flt, bt, best_i);
#pragma unroll
        for (int r = 0; r < RAYS; r++) best_t[r] = __double2float_rn(bt[r]);
        swept = true;
    }
    flush_counters(P, C, n_segments, lane);
}

// K2, scenes of ANY size: k_trace_paths_l with the stage-1 triples streamed per warp from global memory
// (sweep_two_stage_stream) instead of staged once per CTA: nothing of the scene has to fit in shared memory.  Path state
// in registers, warps independent.  Measured against a CTA-lockstep variant that staged tiles of ~15 000 spheres per
// CTA behind two barriers per tile (40 001 spheres: 528 ms; this kernel 500 ms) and against k_trace_paths_l where that
// one fits (10 001 spheres: 508 vs 473 ms), which is why the resident kernel keeps the scenes it can hold.
// dynamic smem: 8 * RAYS * THREADS mask words + (THREADS / 32) * 2 * STREAM_CHUNK_F4 float4.
template <int THREADS, int RAYS, bool STRIDED>
__global__ void __launch_bounds__(THREADS, 1) k_trace_paths_g(const __grid_constant__ TraceParamsL PV) {
    const TraceParams& P = PV.b;
    extern __shared__ float4 smem4[];
    uint32_t* masks = reinterpret_cast<uint32_t*>(smem4) + threadIdx.x;
    float4* wbuf = smem4 + 8 * RAYS * THREADS / 4 + (threadIdx.x >> 5) * (2 * STREAM_CHUNK_F4);
    const TraceConsts C = trace_consts<STRIDED>(P);
    const unsigned lane = threadIdx.x & 31u;
    PathSlot slot[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) slot_reset(slot[r]);
    bool exhausted = false;
    WarpChunk chunk;
    unsigned n_segments = 0;
    float best_t[RAYS];
    int best_i[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) { best_t[r] = RTB_INF; best_i[r] = -1; }
    bool swept = false;
    for (;;) {
        bool any_path = false, idle = true;
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            slot_step<STRIDED>(P, C, slot[r], swept, best_t[r], best_i[r], exhausted, chunk, n_segments, lane);
            any_path = any_path || slot[r].has_path;
            idle = idle && slot[r].item == NO_ITEM;
        }
        if (!__any_sync(0xffffffffu, any_path)) {
            if (__all_sync(0xffffffffu, idle && exhausted)) break;
            swept = false;
            continue;
        }
        ExactRay ray[RAYS];
        FilterRay flt[RAYS];
        SlabRay slab[RAYS];
        double bt[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            ray[r] = make_exact_ray(slot[r].o, slot[r].d, P.both_roots != 0);
            flt[r] = make_filter(slot[r].o, slot[r].d);
            slab[r] = make_slab_ray(slot[r].o.x, slot[r].o.y, slot[r].o.z, slot[r].d.x, slot[r].d.y, slot[r].d.z, plen2_3(slot[r].d), PV.up_axis,
                                    PV.rp_min);
            bt[r] = RTB_INF_D;
            best_i[r] = -1;
        }
        sweep_two_stage_stream<RAYS>(wbuf, PV.s1, lane, RegisterRays{ray}, P.filter, P.exact, P.h_bound, PV.n_padded, masks, THREADS, slab, flt, bt, best_i);
#pragma unroll
        for (int r = 0; r < RAYS; r++) best_t[r] = __double2float_rn(bt[r]);
        swept = true;
    }
    flush_counters(P, C, n_segments, lane);
}

// K2, parameter-space variant (scenes of at most N spheres): the filter spheres are kernel parameters, so the sweep's
// sphere operand sits in uniform registers, and every lane runs RAYS path slots whose state lives in SHARED memory
// between trips (word w of slot r of thread t at st[(w * RAYS + r) * THREADS + t]: conflict-free).  The shading step
// is a rolled loop over the slots (one copy of the code), the sweep is unrolled over them (one sphere load serves RAYS
// filter evaluations), and only the RAYS x 8 filter constants are live in registers across the sweep.
// dynamic smem: (SLOT_WORDS + CAP) * RAYS * THREADS words.  N <= 1024 (one bit per 32-sphere block in a 32-bit word).
constexpr int SLOT_WORDS = 17;   // o 0-2, d 3-5, thr 6-8, acc 9-11, item 12, s 13, bounce|has_path<<31 14, best_t 15, best_i 16

template <int RAYS, int THREADS>
struct SmemRays {
    const uint32_t* st;
    bool both;
    __device__ __forceinline__ bool both_roots() const { return both; }
    __device__ __forceinline__ f3 origin(int r) const {
        return mk3(__uint_as_float(st[(0 * RAYS + r) * THREADS]), __uint_as_float(st[(1 * RAYS + r) * THREADS]), __uint_as_float(st[(2 * RAYS + r) * THREADS]));
    }
    __device__ __forceinline__ ExactRay load(int r) const {
        ExactRay e;
        e.both_roots = both;
        e.ox = __uint_as_float(st[(0 * RAYS + r) * THREADS]); e.oy = __uint_as_float(st[(1 * RAYS + r) * THREADS]);
        e.oz = __uint_as_float(st[(2 * RAYS + r) * THREADS]); e.dx = __uint_as_float(st[(3 * RAYS + r) * THREADS]);
        e.dy = __uint_as_float(st[(4 * RAYS + r) * THREADS]); e.dz = __uint_as_float(st[(5 * RAYS + r) * THREADS]);
        e.inv_a = __hiloint2double((int)st[(16 * RAYS + r) * THREADS], (int)st[(15 * RAYS + r) * THREADS]);   // parked by the kernel before the sweep
        return e;
    }
};

template <int THREADS, int MIN_BLOCKS, int RAYS, int CAP, class PV_T, bool STRIDED>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS) k_trace_paths_v(const __grid_constant__ PV_T PV) {
    const TraceParams& P = PV.b;
    extern __shared__ float4 smem4[];
    uint32_t* st = reinterpret_cast<uint32_t*>(smem4) + threadIdx.x;
    uint32_t* list = reinterpret_cast<uint32_t*>(smem4) + SLOT_WORDS * RAYS * THREADS + threadIdx.x;
    auto W = [&](int w, int r) -> uint32_t& { return st[(w * RAYS + r) * THREADS]; };
    auto F = [&](int w, int r) -> float { return __uint_as_float(st[(w * RAYS + r) * THREADS]); };
    [[maybe_unused]] const float4* s_slab = smem4 + (SLOT_WORDS + CAP) * RAYS * THREADS / 4;
    if constexpr (PV_T::two_stage) {
        if constexpr (PV_T::s1_smem) {                         // stage the stage-1 triples once per CTA
            float4* dst = smem4 + (SLOT_WORDS + CAP) * RAYS * THREADS / 4;
            const int n4 = PV.n_padded * 3 / 4;
            for (int i = threadIdx.x; i < n4; i += THREADS) dst[i] = __ldg(&PV.s1[i]);
            __syncthreads();
        }
    }

    const TraceConsts C = trace_consts<STRIDED>(P);
    const unsigned lane = threadIdx.x & 31u;
#pragma unroll 1
    for (int r = 0; r < RAYS; r++) {
#pragma unroll
        for (int w = 0; w < SLOT_WORDS; w++) W(w, r) = 0u;
        W(5, r) = __float_as_uint(1.0f);                       // stale ray of an idle slot: d = (0,0,1)
        W(12, r) = NO_ITEM;
    }
    bool exhausted = false;
    WarpChunk chunk;
    RTB_TL_START(lane);
    unsigned n_segments = 0;
    bool swept = false;

    for (;;) {
        bool any_path = false, idle = true;
#pragma unroll 1
        for (int r = 0; r < RAYS; r++) {
            PathSlot q;
            q.o = mk3(F(0, r), F(1, r), F(2, r)); q.d = mk3(F(3, r), F(4, r), F(5, r));
            q.thr = mk3(F(6, r), F(7, r), F(8, r)); q.acc = mk3(F(9, r), F(10, r), F(11, r));
            q.item = W(12, r); q.s = (int)W(13, r);
            const uint32_t bw = W(14, r);
            q.bounce = (int)(bw & 0x7fffffffu); q.has_path = (bw >> 31) != 0u;
            slot_step<STRIDED>(P, C, q, swept, F(15, r), (int)W(16, r), exhausted, chunk, n_segments, lane);
            W(0, r) = __float_as_uint(q.o.x); W(1, r) = __float_as_uint(q.o.y); W(2, r) = __float_as_uint(q.o.z);
            W(3, r) = __float_as_uint(q.d.x); W(4, r) = __float_as_uint(q.d.y); W(5, r) = __float_as_uint(q.d.z);
            W(6, r) = __float_as_uint(q.thr.x); W(7, r) = __float_as_uint(q.thr.y); W(8, r) = __float_as_uint(q.thr.z);
            W(9, r) = __float_as_uint(q.acc.x); W(10, r) = __float_as_uint(q.acc.y); W(11, r) = __float_as_uint(q.acc.z);
            W(12, r) = q.item; W(13, r) = (uint32_t)q.s;
            W(14, r) = (uint32_t)q.bounce | (q.has_path ? 0x80000000u : 0u);
            any_path = any_path || q.has_path;
            idle = idle && q.item == NO_ITEM;
        }
        if (!__any_sync(0xffffffffu, any_path)) {
            if (__all_sync(0xffffffffu, idle && exhausted)) break;
            swept = false;
            continue;
        }
        FilterRay flt[RAYS];
        double best_t[RAYS];
        int best_i[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            flt[r] = make_filter(mk3(F(0, r), F(1, r), F(2, r)), mk3(F(3, r), F(4, r), F(5, r)));
            // 1 / (d.d) in f64 for the exact test, once per ray and sweep: parked in the slot's two result words (15, 16), which
            // are free until the sweep's result overwrites them below
            const double inv_a = exact_inv_a(mk3(F(3, r), F(4, r), F(5, r)));
            W(15, r) = (uint32_t)__double2loint(inv_a); W(16, r) = (uint32_t)__double2hiint(inv_a);
            best_t[r] = RTB_INF_D;
            best_i[r] = -1;
        }
        if constexpr (PV_T::two_stage) {
            static_assert(CAP >= 8, "the two-stage sweep keeps 8 mask words per ray");
            SlabRay slab[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++)
                slab[r] = make_slab_ray(F(0, r), F(1, r), F(2, r), F(3, r), F(4, r), F(5, r), plen2_3(mk3(F(3, r), F(4, r), F(5, r))), PV.up_axis, PV.rp_min);
            if constexpr (PV_T::s1_smem)
                sweep_two_stage<RAYS, 8>(SmemSlab{s_slab}, SmemRays<RAYS, THREADS>{st, P.both_roots != 0}, P.filter, P.exact, P.h_bound, P.n_spheres, list, THREADS, slab, flt, best_t, best_i);
            else
                sweep_two_stage<RAYS, 8>(ParamSlab<PV_T>{PV}, SmemRays<RAYS, THREADS>{st, P.both_roots != 0}, P.filter, P.exact, P.h_bound, P.n_spheres, list, THREADS, slab,
                                      flt, best_t, best_i);
        } else {
            sweep_spheres_bitmap<RAYS, CAP>(ParamSpheres<PV_T::capacity>{PV}, SmemRays<RAYS, THREADS>{st, P.both_roots != 0}, P.exact, P.n_spheres, list, THREADS,
                                            flt, best_t, best_i);
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) { W(15, r) = __float_as_uint(__double2float_rn(best_t[r])); W(16, r) = (uint32_t)best_i[r]; }   // the f64 root, rounded once
        swept = true;
    }
    RTB_TL_EXIT(lane);
    flush_counters(P, C, n_segments, lane);
}
// ---- combine: accum[row range] = sum_k partial[k] in plane order (deterministic) -----------------------------------
// out = the image at the launch's first traced row; width / row_skip map a local pixel to its image pixel when the traced
// rows are interleaved (row_skip = (row_stride - 1) * width, 0 for a contiguous band).
__global__ void __launch_bounds__(256) k_combine(const float4* __restrict__ partial, int splits, long long n_pix, float4* __restrict__ out, int width,
                                                 long long row_skip) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    float4 acc = partial[i];
    for (int k = 1; k < splits; k++) {
        float4 v = partial[(long long)k * n_pix + i];
        acc.x = __fadd_rn(acc.x, v.x); acc.y = __fadd_rn(acc.y, v.y); acc.z = __fadd_rn(acc.z, v.z); acc.w = __fadd_rn(acc.w, v.w);
    }
    out[row_skip ? i + (i / width) * row_skip : i] = acc;
}

// ---- K3: resolve.  4 pixels per thread: four float4 loads -> three packed u32 stores (12 bytes of RGB8). -------------
__global__ void __launch_bounds__(256) k_resolve(const float4* __restrict__ accum, long long n_pix, float spp, uint8_t* __restrict__ rgb8) {
    long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // quad index
    long long p0 = q * 4;
    if (p0 >= n_pix) return;
    uint32_t b[12];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        float4 v = (p0 + k < n_pix) ? accum[p0 + k] : make_float4(0.f, 0.f, 0.f, 0.f);
        b[3 * k + 0] = quantise(__fdiv_rn(v.x, spp));               // pixel = sum / spp, src/renderer.rs:161
        b[3 * k + 1] = quantise(__fdiv_rn(v.y, spp));
        b[3 * k + 2] = quantise(__fdiv_rn(v.z, spp));
    }
    if (p0 + 4 <= n_pix && ((reinterpret_cast<uintptr_t>(rgb8) & 3) == 0)) {
        uint32_t* o32 = reinterpret_cast<uint32_t*>(rgb8 + p0 * 3);
        o32[0] = b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24);
        o32[1] = b[4] | (b[5] << 8) | (b[6] << 16) | (b[7] << 24);
        o32[2] = b[8] | (b[9] << 8) | (b[10] << 16) | (b[11] << 24);
    } else {
        for (int k = 0; k < 12 && p0 * 3 + k < n_pix * 3; k++) rgb8[p0 * 3 + k] = (uint8_t)b[k];
    }
}

// ---- K3p: fused peer reduce + resolve.  Each thread owns 4 consecutive pixels: it loads their float4 partial sums from
// every rank's accumulation image (its own HBM and, over NVLink, the peers'), adds them in rank order, and stores 12
// bytes of RGB8 (possibly into a peer's image) and, on request, the summed float4 (sum_out, null = not wanted).  All loads
// of a thread are issued before the first add.
struct PeerPtrs { const float4* p[16]; };

__global__ void __launch_bounds__(256) k_resolve_peers(const __grid_constant__ PeerPtrs peers, int n_peers, int row_groups, int width, long long pixel_begin,
                                                       long long pixel_end, float spp, uint8_t* __restrict__ rgb8, float4* __restrict__ sum_out) {
    // peers.p[g * n_peers + k]: partial image of sample group k of row group g; image row r belongs to row group
    // r % row_groups (interleaved scanlines).  row_groups == 1: every pixel sums peers.p[0 .. n_peers).
    long long p0 = pixel_begin + ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (p0 >= pixel_end) return;
    const int n = (int)((pixel_end - p0) < 4 ? (pixel_end - p0) : 4);
    int base[4];
#pragma unroll
    for (int k = 0; k < 4; k++) base[k] = row_groups > 1 ? (int)(((p0 + k) / width) % row_groups) * n_peers : 0;
    float4 acc[4];
#pragma unroll
    for (int k = 0; k < 4; k++) acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int g = 0; g < n_peers; g += 4) {                   // four peers' loads in flight at a time
        float4 v[4][4];
#pragma unroll
        for (int gg = 0; gg < 4; gg++)
#pragma unroll
            for (int k = 0; k < 4; k++)
                v[gg][k] = (g + gg < n_peers && k < n) ? peers.p[base[k] + g + gg][p0 + k] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int gg = 0; gg < 4; gg++)
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (g + gg < n_peers) {
                    acc[k].x = __fadd_rn(acc[k].x, v[gg][k].x); acc[k].y = __fadd_rn(acc[k].y, v[gg][k].y);
                    acc[k].z = __fadd_rn(acc[k].z, v[gg][k].z); acc[k].w = __fadd_rn(acc[k].w, v[gg][k].w);
                }
            }
    }
    if (sum_out) {                                           // the summed linear image (local or a peer's), for callers that want it
        for (int k = 0; k < n; k++) sum_out[p0 + k] = acc[k];
    }
    uint32_t b[12];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        b[3 * k + 0] = quantise(__fdiv_rn(acc[k].x, spp));          // pixel = sum / spp, src/renderer.rs:161
        b[3 * k + 1] = quantise(__fdiv_rn(acc[k].y, spp));
        b[3 * k + 2] = quantise(__fdiv_rn(acc[k].z, spp));
    }
    if (n == 4 && (((reinterpret_cast<uintptr_t>(rgb8) + (uintptr_t)(p0 * 3)) & 3) == 0)) {
        uint32_t* o32 = reinterpret_cast<uint32_t*>(rgb8 + p0 * 3);
        o32[0] = b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24);
        o32[1] = b[4] | (b[5] << 8) | (b[6] << 16) | (b[7] << 24);
        o32[2] = b[8] | (b[9] << 8) | (b[10] << 16) | (b[11] << 24);
    } else {
        for (int k = 0; k < 3 * n; k++) rgb8[p0 * 3 + k] = (uint8_t)b[k];
    }
}

// ---- peer barrier: a stream-ordered barrier between the GPUs of one box, through peer memory ------------------------------
// One process per GPU, each with a flag array of n_peers words that every peer has mapped (CUDA IPC).  Rank r's kernel
// publishes `epoch` into slot r of EVERY rank's array (peer stores over NVLink, after a system-scope fence so that everything
// the stream wrote before — the partial framebuffer — is visible first), then waits until all slots of its OWN array have
// reached `epoch`.  Kernels queued behind it on the stream therefore start only after every rank has finished what it queued
// before its barrier: what the exchange needs before the reduce (all traces done) and after it (all slices landed) — without
// a host round trip and without an NCCL call per step.  A rank that never arrives would hang the others, so the wait is
// bounded: after timeout_ns the kernel traps and the next CUDA call of that process fails loudly.
struct BarrierPtrs { unsigned int* p[16]; };

__global__ void __launch_bounds__(32) k_peer_barrier(const __grid_constant__ BarrierPtrs flags, int n_peers, int my_rank, unsigned int epoch,
                                                     unsigned long long timeout_ns) {
    const int g = (int)threadIdx.x;
    if (g >= n_peers) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flags.p[g] + my_rank), "r"(epoch) : "memory");
    unsigned long long t0;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    const unsigned int* mine = flags.p[my_rank] + g;
    for (;;) {
        unsigned int v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
        if ((int)(v - epoch) >= 0) break;                     // epochs only grow (wrap-safe comparison)
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        if (t - t0 > timeout_ns) __trap();
        __nanosleep(200);
    }
    __threadfence_system();
}

// ---- filter self-check: the conservative filters against the exact test, ON THE DEVICE --------------------------------
// The parity of every sweep rests on two FP32 filters never rejecting a pair the exact (f64) test accepts.  This kernel
// brute-forces that property with the arithmetic that actually runs: the device builds of make_filter / make_slab_ray
// (approximate rsqrtf included) / filter_value / slab_value and exact_discriminant, on adversarial pairs of the UPLOADED
// scene — rays aimed at a sphere's silhouette with relative perturbations 1e-9 .. 1e-1 to either side, from a camera
// point, from points on other spheres' surfaces (bounce rays) and from far out on the largest sphere, directions of
// length 0.1 .. 2 (the reference never normalises); each ray is tested against the aimed-at sphere and
// `per_ray - 1` random others, for the slab test along all three axes.
// out: [0] pairs, [1] exact-positive pairs (disc >= 0), [2] filter candidates, [3] filter FALSE NEGATIVES,
//      [4..6] slab candidates per dropped axis, [7] slab FALSE NEGATIVES, [8] min over exact-positive pairs of
//      filter value / margin, as float bits in the low word (atomicMin on the bit pattern of a non-negative float).
struct FilterCheckArgs {
    const float4* spheres; const float4* filter; const float* slab; const double4* exact;
    int n_spheres; float rp_min; int largest;      // index of the sphere with the largest radius ("ground")
    float cam[3];
    unsigned long long n_rays; int per_ray; uint32_t seed;
    unsigned long long* out;
};

__global__ void __launch_bounds__(256) k_filter_check(const FilterCheckArgs A) {
    unsigned long long cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float min_slack = 3.0e38f;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long ray_id = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; ray_id < A.n_rays; ray_id += stride) {
        uint32_t ctr = 0;
        auto rnd = [&]() { return philox4x32_10((uint32_t)ray_id, (uint32_t)(ray_id >> 32), ctr++, 0x51f7u, A.seed, 0x2545f491u); };
        auto u01d = [](uint32_t x) { return ((double)x + 0.5) * (1.0 / 4294967296.0); };
        auto unit_rand = [&](double v[3]) {
            const uint4 q = rnd();
            const double z = 1.0 - 2.0 * u01d(q.x), rr = sqrt(fmax(0.0, 1.0 - z * z)), ph = 6.283185307179586 * u01d(q.y);
            v[0] = rr * cos(ph); v[1] = rr * sin(ph); v[2] = z;
        };
        // ---- origin (f32-valued, as the path state is)
        const uint4 q0 = rnd();
        const double pick = u01d(q0.x);
        double o[3];
        if (pick < 0.25) { o[0] = A.cam[0]; o[1] = A.cam[1]; o[2] = A.cam[2]; }
        else {
            const int bi = pick < 0.45 ? A.largest : (int)(q0.y % (uint32_t)A.n_spheres);
            const float4 b = A.spheres[bi];
            const double br = sqrt(fmax(0.0, (double)b.w));
            double nn[3]; unit_rand(nn);
            if (bi == A.largest && pick < 0.45) {                      // near the "top" of the ground sphere, out to 0.6 radii
                const double rad = br * pow(10.0, -4.0 + 3.8 * u01d(q0.z)), ang = 6.283185307179586 * u01d(q0.w);
                nn[0] = rad * cos(ang) / br; nn[2] = rad * sin(ang) / br; nn[1] = sqrt(fmax(0.0, 1.0 - nn[0] * nn[0] - nn[2] * nn[2]));
            }
            o[0] = b.x + nn[0] * br; o[1] = b.y + nn[1] * br; o[2] = b.z + nn[2] * br;
        }
        const float of[3] = {(float)o[0], (float)o[1], (float)o[2]};
        // ---- the aimed-at sphere and a near-silhouette direction
        const uint4 q1 = rnd();
        const int ti = (int)(q1.x % (uint32_t)A.n_spheres);
        const double4 ts = A.exact[ti];
        const double tr = sqrt(fmax(0.0, ts.w));
        const double oc[3] = {ts.x - (double)of[0], ts.y - (double)of[1], ts.z - (double)of[2]};
        const double L = sqrt(oc[0] * oc[0] + oc[1] * oc[1] + oc[2] * oc[2]);
        double d[3];
        bool aimed = false;
        if (L > tr * (1.0 + 1e-9) && u01d(q1.y) < 0.9) {
            double ax[3]; unit_rand(ax);
            const double dp = (ax[0] * oc[0] + ax[1] * oc[1] + ax[2] * oc[2]) / (L * L);
            for (int k = 0; k < 3; k++) ax[k] -= dp * oc[k];
            const double al = sqrt(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]);
            if (al > 1e-9) {
                const double eps = pow(10.0, -9.0 + 8.0 * u01d(q1.z)) * ((q1.w & 1u) ? -1.0 : 1.0);
                const double theta = asin(tr / L) * (1.0 + eps);
                for (int k = 0; k < 3; k++) d[k] = cos(theta) * oc[k] / L + sin(theta) * ax[k] / al;
                aimed = true;
            }
        }
        if (!aimed) unit_rand(d);
        const double scale = 0.1 + 1.9 * u01d(q1.w);
        const f3 ro = mk3(of[0], of[1], of[2]), rd = mk3((float)(d[0] * scale), (float)(d[1] * scale), (float)(d[2] * scale));
        // ---- what the sweeps compute per ray
        const FilterRay fr = make_filter(ro, rd);
        SlabRay sr[3];
#pragma unroll
        for (int up = 0; up < 3; up++) sr[up] = make_slab_ray(ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, plen2_3(rd), up, A.rp_min);
        const ExactRay64 er = promote_ray(make_exact_ray(ro, rd, false));
        const double o2 = (double)of[0] * of[0] + (double)of[1] * of[1] + (double)of[2] * of[2];
        for (int rep = 0; rep < A.per_ray; rep++) {
            int si = ti;
            if (rep > 0) { const uint4 q = rnd(); si = (int)(q.x % (uint32_t)A.n_spheres); }
            const float4 fs = A.filter[si];
            const float4 sp = A.spheres[si];
            const float rp = A.slab[3 * si + 2];
            double hb;
            const bool ref = exact_discriminant(er, A.exact[si], hb) >= 0.0;
            const float qv = filter_value(fr, fs.x, fs.y, fs.z, fs.w);
            float hp;
            const float qp = filter_value_packed(fr, fs, hp);                     // what the walks run: must be the same bits
            const bool flag = !(__float_as_uint(qv) >> 31) && __float_as_uint(qp) == __float_as_uint(qv);
            cnt[0]++; cnt[1] += ref; cnt[2] += flag; cnt[3] += (ref && !flag);
            const float cc[3] = {sp.x, sp.y, sp.z};
#pragma unroll
            for (int up = 0; up < 3; up++) {
                const int ia = up == 2 ? 0 : up + 1, ib = ia == 2 ? 0 : ia + 1;
                const bool sflag = !(__float_as_uint(slab_value(sr[up], cc[ia], cc[ib], rp)) >> 31);
                // the packed form of the shared-memory kernels (R' one ulp up, negated): never fewer candidates than the scalar form
                const bool pflag = (__float_as_uint(slab_value_packed(sr[up], cc[ia], cc[ib], -nextafterf(rp, RTB_INF))) >> 31) != 0u;
                cnt[4 + up] += sflag;
                cnt[7] += (ref && !sflag) || (ref && !pflag) || (sflag && !pflag);
            }
            if (ref && flag) {
                const double b = sqrt((double)sp.x * sp.x + (double)sp.y * sp.y + (double)sp.z * sp.z) + sqrt(fmax(0.0, (double)sp.w));
                const double margin = (double)RTB_FILTER_MARGIN * (o2 + b * b);
                if (margin > 0.0) min_slack = fminf(min_slack, (float)((double)qv / margin));
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        unsigned long long v = cnt[k];
        for (int sh = 16; sh > 0; sh >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sh);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(&A.out[k], v);
    }
    for (int sh = 16; sh > 0; sh >>= 1) min_slack = fminf(min_slack, __shfl_xor_sync(0xffffffffu, min_slack, sh));
    if ((threadIdx.x & 31) == 0) atomicMin(&A.out[8], (unsigned long long)__float_as_uint(fmaxf(min_slack, 0.0f)));
}

// ---- K0: FP32 FMA peak probe ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024, 1) k_fp32_peak(float* out, int iters, float seed) {
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = seed + (float)i;
    float m = 0.999f + seed * 1e-9f, b = seed * 1e-3f;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) acc[i] = __fmaf_rn(acc[i], m, b);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i];
    if (s == 123.456f) out[0] = s;
}

}  // namespace rtb
