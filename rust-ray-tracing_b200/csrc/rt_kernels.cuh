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
constexpr float T_MIN = 0.001f;          // 0.001..f64::INFINITY (src/ray_tracing.rs:400)
#define RTB_INF __int_as_float(0x7f800000)

// ---- sphere staging ------------------------------------------------------------------------------------------
// Two device arrays per scene, both in list order: the EXACT spheres float4 {cx, cy, cz, r*r} (global memory, read
// only by the candidate re-test and by shading) and the FILTER spheres float4 {cx, cy, cz, k} (rt_filter.h), which
// are what the sweep stages in shared memory.  `count` spheres are copied as they are (no padding: the sweep handles
// the remainder of the last 32-block with a rolled loop).
__device__ __forceinline__ void stage_spheres(float4* s_sph, const float4* __restrict__ g_flt, int begin, int count) {
    for (int i = threadIdx.x; i < count; i += blockDim.x) s_sph[i] = __ldg(&g_flt[begin + i]);
}

// One ray of a lane, as the sweep sees it: origin, direction, and the ray's half of the candidate filter.
// a = d.d (src/objects.rs:253) and 1/a (:254) are recomputed where the exact test needs them (once per drain).
struct SweepRay { float ox, oy, oz, dx, dy, dz; FilterRay f; };

__device__ __forceinline__ SweepRay make_sweep_ray(f3 o, f3 d) {
    SweepRay r;
    r.ox = o.x; r.oy = o.y; r.oz = o.z; r.dx = d.x; r.dy = d.y; r.dz = d.z;
    r.f = make_filter_ray(o.x, o.y, o.z, d.x, d.y, d.z, plen2_3(d));
    return r;
}

// The discriminant of hit_packed in the reference's operation order (src/objects.rs:252-257):
// 3 FADD (oc), FMUL + 2 FFMA (half_b), FMUL + 2 FFMA + FADD (c), FMUL + FFMA (disc) = 12 FP32 instructions, 17 FLOP.
__device__ __forceinline__ float discriminant(const SweepRay& r, float a, float4 s, float& hb) {
    float ocx = __fsub_rn(r.ox, s.x), ocy = __fsub_rn(r.oy, s.y), ocz = __fsub_rn(r.oz, s.z);       // :252
    hb = __fmaf_rn(ocz, r.dz, __fmaf_rn(ocy, r.dy, __fmul_rn(ocx, r.dx)));                          // :255
    float cc = __fsub_rn(__fmaf_rn(ocz, ocz, __fmaf_rn(ocy, ocy, __fmul_rn(ocx, ocx))), s.w);       // :256
    return __fmaf_rn(hb, hb, -__fmul_rn(a, cc));                                                    // :257
}

// Exact per-candidate test: the body of hit_packed's `if discriminant_positive.any()` block for one lane
// (src/objects.rs:262-288) followed by PackedHitRecords::update (:140-147).
__device__ __forceinline__ void test_candidate(const SweepRay& r, float a, float inv_a, float4 s, int index, float& best_t, int& best_i) {
    float hb;
    float disc = discriminant(r, a, s, hb);
    if (disc >= 0.0f) {                                                                            // :259
        float sq = __fsqrt_rn(disc);                                                               // :263
        float root1 = __fmul_rn(__fsub_rn(-hb, sq), inv_a);                                        // :270
        bool valid = (root1 >= T_MIN) && (root1 < RTB_INF);                                        // :272-277 (near root only, F4)
        if (valid && root1 <= best_t) { best_t = root1; best_i = index; }                          // :141 (`<=`: last wins, F5)
    }
}

// Drains one ray's deferred candidate list in sweep order.  list entry = {sign mask of up to 32 filter values (bit
// 31 = first sphere of the block, set = not a candidate), index of the block's first sphere in the staged array}.
// g_exact = the exact spheres of the staged range (global memory, L1-resident).
__device__ __forceinline__ void drain_candidates(const float4* __restrict__ g_exact, int index_base, const uint2* list, int list_stride, int& n_list,
                                                 const SweepRay& r, float& best_t, int& best_i) {
    const float a = plen2_3(mk3(r.dx, r.dy, r.dz));                                                // :253
    const float inv_a = __fdiv_rn(1.0f, a);                                                        // :254
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
        test_candidate(r, a, inv_a, __ldg(&g_exact[local]), index_base + local, best_t, best_i);
    }
    n_list = 0;
}

// The sweep of RAYS rays per lane over `count` staged spheres (global indices index_base .. index_base+count).
//
// Fast path per (sphere, ray), all lanes, no branch: the 12 FP32 operations of the discriminant plus ONE funnel shift
// that appends the discriminant's sign bit to a 32-sphere mask word; one LDS.128 per sphere serves all RAYS rays of the
// lane.  A set sign bit means disc < 0 (or disc = -0.0, which needs half_b = +-0 and then root1 = +-0 < 0.001: not a hit
// either), so candidates are the clear bits.
// Slow path (deferred, lane-parallel): mask words with a clear bit are pushed to a per-ray list in shared memory and
// re-tested exactly (sqrt, root, range, `<=`) after the sweep, in sweep order — so the result is that of the
// reference's sequential `update` over the object list.
// list layout: entry e of ray r of this thread at list[(r * CAP + e) * list_stride].  CAP >= number of 32-blocks
// (CHECKED = false) removes the overflow test from the per-block path.
template <int RAYS, int CAP, bool CHECKED>
__device__ __forceinline__ void sweep_spheres(const float4* s_sph, const float4* __restrict__ g_exact, int count, int index_base,
                                              uint2* list, int list_stride,
                                              const SweepRay (&ray)[RAYS], float (&best_t)[RAYS], int (&best_i)[RAYS]) {
    int n_list[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) n_list[r] = 0;
    const int full = count & ~(SWEEP_BLOCK - 1);
    for (int blk = 0; blk < full; blk += SWEEP_BLOCK) {
        uint32_t mask[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) mask[r] = 0;
        constexpr int UNROLL = SWEEP_BLOCK / RAYS;      // keep the unrolled body at ~450 instructions (instruction cache)
#pragma unroll 1
        for (int sub = 0; sub < SWEEP_BLOCK; sub += UNROLL) {
#pragma unroll
            for (int j = 0; j < UNROLL; j++) {
                float4 s = s_sph[blk + sub + j];                                                  // LDS.128 broadcast = Simd::splat
#pragma unroll
                for (int r = 0; r < RAYS; r++) {
                    float q = filter_value(ray[r].f, s.x, s.y, s.z, s.w);                         // 8 FP32 instructions
                    mask[r] = __funnelshift_l(__float_as_uint(q), mask[r], 1);                    // (mask << 1) | sign(q)
                }
            }
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            if (mask[r] != 0xffffffffu) {
                list[(r * CAP + n_list[r]) * list_stride] = make_uint2(mask[r], (uint32_t)blk);
                n_list[r]++;
                if (CHECKED && n_list[r] == CAP)
                    drain_candidates(g_exact, index_base, list + r * CAP * list_stride, list_stride, n_list[r], ray[r], best_t[r], best_i[r]);
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
            float4 s = s_sph[full + j];
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                float q = filter_value(ray[r].f, s.x, s.y, s.z, s.w);
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
            drain_candidates(g_exact, index_base, list + r * CAP * list_stride, list_stride, n_list[r], ray[r], best_t[r], best_i[r]);
}

// ---- K1 ----------------------------------------------------------------------------------------------------------
// One thread per ray; spheres streamed through shared memory in tiles of `tile` (multiple of SWEEP_BLOCK).
// dynamic smem: tile float4 + LIST_CAP * blockDim uint2.
__global__ void __launch_bounds__(256) k_intersect_batch(const float4* __restrict__ g_sph, const float4* __restrict__ g_flt, int n_spheres, int tile,
                                                         const float* __restrict__ ox, const float* __restrict__ oy,
                                                         const float* __restrict__ oz, const float* __restrict__ dx,
                                                         const float* __restrict__ dy, const float* __restrict__ dz,
                                                         const uint8_t* __restrict__ enabled, long long n_rays,
                                                         int* __restrict__ hit_index, float* __restrict__ hit_t) {
    extern __shared__ float4 smem4[];
    float4* s_sph = smem4;
    uint2* list = reinterpret_cast<uint2*>(smem4 + tile) + threadIdx.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    bool live = i < n_rays;
    f3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f);
    if (live) { o = mk3(ox[i], oy[i], oz[i]); d = mk3(dx[i], dy[i], dz[i]); }
    SweepRay ray[1] = {make_sweep_ray(o, d)};
    float best_t[1] = {RTB_INF};                               // PackedHitRecords::default, src/objects.rs:121-134
    int best_i[1] = {-1};
    for (int begin = 0; begin < n_spheres; begin += tile) {
        int count = min(tile, n_spheres - begin);
        __syncthreads();
        stage_spheres(s_sph, g_flt, begin, count);
        __syncthreads();
        sweep_spheres<1, LIST_CAP, true>(s_sph, g_sph + begin, count, begin, list, blockDim.x, ray, best_t, best_i);
    }
    if (live) {
        bool en = enabled ? enabled[i] != 0 : true;            // rays.enabled(), src/objects.rs:259,277
        hit_index[i] = en ? best_i[0] : -1;
        hit_t[i] = (en && best_i[0] >= 0) ? best_t[0] : RTB_INF;
    }
}

// ---- K2 ----------------------------------------------------------------------------------------------------------
struct TraceParams {
    DeviceCamera cam;
    const float4* spheres;        // exact: {cx,cy,cz,r*r}
    const float4* filter;         // candidate filter: {cx,cy,cz,k} (rt_filter.h)
    const float4* mat;            // per sphere {albedo rgb, param}
    const int* mat_kind;          // per sphere kind | hollow<<8
    int n_spheres;
    int max_bounces;
    uint32_t seed_lo, seed_hi;
    int sample_begin, sample_end; // global sample indices traced by this launch
    int row_begin, row_end;
    int splits_log2;              // K = 1 << splits_log2: the sample range is cut into K contiguous sub-ranges, one work item each
    uint32_t width_magic;         // floor(2^32 / W) + 1 (0 when W == 1): row = pixel / W by multiply-high
    float4* out;                  // K == 1: accum image (W*H float4); else partial planes [K][rows*W]
    unsigned long long* work_counter;
    unsigned long long* counters; // [0] paths, [1] segments
};

// One path slot of a lane.  A slot owns one work item (a pixel and a contiguous range of its samples) at a time and
// runs the samples one after another; the per-pixel sum `acc` is a private sum in sample order, so the image is
// deterministic and needs no atomics.
struct PathSlot {
    f3 o, d;                      // current ray (direction NOT normalised after scatter, src/ray.rs:13-18)
    f3 thr, acc;                  // path throughput (`color`, src/ray_tracing.rs:319) and the pixel's running sum
    uint32_t pixel_id;            // row * W + col: Philox counter word 0
    uint32_t out_off;             // float4 index into P.out
    int s, s_end, n_s;            // next sample, end of the item's sample range, its length
    int bounce;
    bool has_item, has_path;
};

// Persistent kernel: grid = SMs x resident CTAs.  Every loop trip makes ONE closest-hit sweep for all slots of the
// warp, then per slot: terminate (sky / depth limit) or keep the hit; finished items are written out and new ones
// claimed with one atomic per warp; then ONE Philox block per slot feeds either the next camera ray or the scatter.
// dynamic smem: n_spheres float4 + RAYS * CAP * THREADS uint2.  CAP = LIST_CAP_SMALL needs n_spheres <= 32 * CAP.
template <int THREADS, int MIN_BLOCKS, int RAYS, int CAP>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS) k_trace_paths(const TraceParams P) {
    extern __shared__ float4 smem4[];
    float4* s_sph = smem4;
    uint2* list = reinterpret_cast<uint2*>(smem4 + ((P.n_spheres + 1) & ~1)) + threadIdx.x;
    stage_spheres(s_sph, P.filter, 0, P.n_spheres);
    __syncthreads();

    const int W = P.cam.width;
    const uint32_t n_pix = (uint32_t)(P.row_end - P.row_begin) * (uint32_t)W;
    const uint32_t n_items = n_pix << P.splits_log2;          // host guarantees < 2^32
    const uint32_t n_samples = (uint32_t)(P.sample_end - P.sample_begin);
    const uint32_t first_pixel = (uint32_t)P.row_begin * (uint32_t)W;
    const unsigned lane = threadIdx.x & 31u;

    PathSlot slot[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) {
        slot[r].o = mk3(0.f, 0.f, 0.f); slot[r].d = mk3(0.f, 0.f, 1.f);
        slot[r].thr = mk3(1.f, 1.f, 1.f); slot[r].acc = mk3(0.f, 0.f, 0.f);
        slot[r].pixel_id = 0; slot[r].out_off = 0; slot[r].s = 0; slot[r].s_end = 0; slot[r].n_s = 0; slot[r].bounce = 0;
        slot[r].has_item = false; slot[r].has_path = false;
    }
    bool exhausted = false;
    unsigned n_segments = 0, n_paths = 0;          // per-lane counts (a lane traces far fewer than 2^32 segments)
    float best_t[RAYS];
    int best_i[RAYS];
    bool swept = false;

    for (;;) {
        // ---- 1. after a sweep: terminate paths that escaped or reached the depth limit ----
        if (swept) {
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                PathSlot& q = slot[r];
                if (q.has_path) {
                    n_segments++;
                    if (best_i[r] < 0) {                       // hit the sky (src/ray_tracing.rs:353-357, :365-369)
                        f3 sk = sky(q.d);
                        q.acc = add3(q.acc, mk3(__fmul_rn(q.thr.x, sk.x), __fmul_rn(q.thr.y, sk.y), __fmul_rn(q.thr.z, sk.z)));
                        q.has_path = false;
                        q.s++;
                    } else if (q.bounce + 1 >= P.max_bounces) { // still enabled after the last bounce => black (:370)
                        q.has_path = false;
                        q.s++;
                    }
                }
            }
        }
        // ---- 2. finished items are written out; new items are claimed with one atomic per warp ----
        unsigned need_mask[RAYS];
        int n_need = 0;
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            PathSlot& q = slot[r];
            if (q.has_item && !q.has_path && q.s >= q.s_end) {
                P.out[q.out_off] = make_float4(q.acc.x, q.acc.y, q.acc.z, (float)q.n_s);
                q.has_item = false;
            }
            need_mask[r] = __ballot_sync(0xffffffffu, !q.has_item && !exhausted);
            n_need += __popc(need_mask[r]);
        }
        if (n_need) {
            unsigned long long base64 = 0;
            if (lane == 0) base64 = atomicAdd(P.work_counter, (unsigned long long)n_need);
            base64 = __shfl_sync(0xffffffffu, base64, 0);
            // items beyond n_items (< 2^32) saturate: the queue head only ever overshoots by the number of lanes
            uint32_t base = base64 > 0xffffffffull ? 0xffffffffu : (uint32_t)base64;
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                PathSlot& q = slot[r];
                if (!q.has_item && !exhausted) {
                    uint32_t rank = __popc(need_mask[r] & ((1u << lane) - 1u));
                    uint32_t item = base + rank;
                    if (item >= base && item < n_items) {      // item = pixel * K + split: a warp works on neighbouring pixels
                        uint32_t split = item & ((1u << P.splits_log2) - 1u), pix = item >> P.splits_log2;
                        q.pixel_id = first_pixel + pix;
                        q.out_off = (P.splits_log2 == 0) ? q.pixel_id : split * n_pix + pix;
                        q.s = P.sample_begin + (int)(((unsigned long long)split * n_samples) >> P.splits_log2);
                        q.s_end = P.sample_begin + (int)(((unsigned long long)(split + 1) * n_samples) >> P.splits_log2);
                        q.n_s = q.s_end - q.s;
                        q.acc = mk3(0.f, 0.f, 0.f);
                        q.has_item = true;
                        if (P.max_bounces <= 0) {              // `for _ in 0..0`: every ray stays enabled => black (:370)
                            n_paths += (unsigned)q.n_s;
                            q.s = q.s_end;
                        }
                    } else {
                        exhausted = true;
                    }
                }
                uint32_t adv = __popc(need_mask[r]);
                base = (base + adv < base) ? 0xffffffffu : base + adv;
            }
        }
        // ---- 3. one Philox block per slot: the next camera sample, or this bounce's scatter ----
        bool any_path = false;
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            PathSlot& q = slot[r];
            const bool regen = q.has_item && !q.has_path && q.s < q.s_end;
            if (regen || q.has_path) {
                uint4 rnd = philox4x32_10(q.pixel_id, (uint32_t)q.s, regen ? RTB_DIM_CAMERA : 1u + (uint32_t)q.bounce, 0u, P.seed_lo, P.seed_hi);
                if (regen) {                                   // Camera::get_ray, src/ray_tracing.rs:77-89
                    uint32_t rowu = P.width_magic ? __umulhi(q.pixel_id, P.width_magic) : q.pixel_id;   // pixel / W
                    if (rowu * (uint32_t)W > q.pixel_id) rowu--;                                       // estimate is q or q+1
                    int row = (int)rowu, col = (int)(q.pixel_id - rowu * (uint32_t)W);
                    camera_get_ray(P.cam, col, row, rnd, q.o, q.d);
                    q.thr = mk3(1.f, 1.f, 1.f);
                    q.bounce = 0;
                    q.has_path = true;
                    n_paths++;
                } else {                                       // finalize + Material::get_hit_result (src/ray_tracing.rs:335-346)
                    float4 sp = s_sph[best_i[r]];
                    HitRec rec = finalize_hit(q.o, q.d, best_t[r], mk3(sp.x, sp.y, sp.z));
                    f3 att, nd;
                    scatter(__ldg(&P.mat_kind[best_i[r]]), __ldg(&P.mat[best_i[r]]), q.d, rec, rnd, att, nd);
                    q.thr = mk3(__fmul_rn(q.thr.x, att.x), __fmul_rn(q.thr.y, att.y), __fmul_rn(q.thr.z, att.z));
                    q.o = rec.p;
                    q.d = nd;
                    q.bounce++;
                }
            }
            any_path = any_path || q.has_path;
        }
        if (!__any_sync(0xffffffffu, any_path)) {
            bool idle = exhausted;
#pragma unroll
            for (int r = 0; r < RAYS; r++) idle = idle && !slot[r].has_item;
            if (__all_sync(0xffffffffu, idle)) break;
            swept = false;
            continue;                                          // items with an empty sample range: write out, claim again
        }
        // ---- 4. the closest-hit sweep (all lanes; a slot without a path sweeps its stale ray and the result is ignored)
        SweepRay ray[RAYS];
#pragma unroll
        for (int r = 0; r < RAYS; r++) {
            ray[r] = make_sweep_ray(slot[r].o, slot[r].d);
            best_t[r] = RTB_INF;                               // PackedHitRecords::default, src/objects.rs:121-134
            best_i[r] = -1;
        }
        sweep_spheres<RAYS, CAP, (CAP < LIST_CAP_SMALL)>(s_sph, P.spheres, P.n_spheres, 0, list, THREADS, ray, best_t, best_i);
        swept = true;
    }

    // ---- counters: one atomic pair per warp ----
    unsigned long long seg64 = n_segments, path64 = n_paths;
#pragma unroll
    for (int sh = 16; sh > 0; sh >>= 1) {
        seg64 += __shfl_xor_sync(0xffffffffu, seg64, sh);
        path64 += __shfl_xor_sync(0xffffffffu, path64, sh);
    }
    if (lane == 0) {
        atomicAdd(&P.counters[0], path64);
        atomicAdd(&P.counters[1], seg64);
    }
}

// ---- combine: accum[row range] = sum_k partial[k] in plane order (deterministic) -----------------------------------
__global__ void __launch_bounds__(256) k_combine(const float4* __restrict__ partial, int splits, long long n_pix, float4* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    float4 acc = partial[i];
    for (int k = 1; k < splits; k++) {
        float4 v = partial[(long long)k * n_pix + i];
        acc.x = __fadd_rn(acc.x, v.x); acc.y = __fadd_rn(acc.y, v.y); acc.z = __fadd_rn(acc.z, v.z); acc.w = __fadd_rn(acc.w, v.w);
    }
    out[i] = acc;
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
