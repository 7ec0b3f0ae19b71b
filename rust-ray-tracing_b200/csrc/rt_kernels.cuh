// rt_kernels.cuh — the sm_100a kernels of the render hot path.
//
//   sweep_spheres     the closest-hit sweep of one ray against a shared-memory sphere array; the device mirror of
//                     `for object in &self.objects { object.hit_packed(..) }` (src/ray_tracing.rs:399-401) with
//                     Sphere::hit_packed (src/objects.rs:249-290) and PackedHitRecords::update (:140-155).
//                     A warp is the reference's PackedRays<N> with N = 32: one lane per ray, one sphere broadcast
//                     from shared memory per step (= Simd::splat, src/objects.rs:252-256).
//   k_intersect_batch K1: the sweep alone for a batch of rays (parity + throughput probe).
//   k_trace_paths     K2: persistent path-tracing kernel — Camera::get_ray + Scene::trace_vectorized* + materials
//                     + sky + per-pixel accumulation (src/renderer.rs:141-176, src/ray_tracing.rs:77-89, :312-628,
//                     src/materials.rs:53-147).  Lane = one pixel's sample stream; path regeneration replaces the
//                     reference's lane compaction (src/ray_tracing.rs:429-483, :561-607).
//   k_combine         sums the K sample-split partial planes in plane order.
//   k_resolve         K3: sum/spp + Color::to_u8_array + image assembly (src/renderer.rs:161-163, :376-384;
//                     src/color.rs:54-64).
//   k_fp32_peak       K0: dependent-free FFMA chains, the measured FP32 roofline for this path.
#pragma once
#include "rt_device.cuh"

namespace rtb {

constexpr int SWEEP_BLOCK = 32;          // spheres per candidate-mask word
constexpr int LIST_CAP = 6;              // deferred candidate-mask entries per lane before a forced flush
constexpr float T_MIN = 0.001f;          // 0.001..f64::INFINITY (src/ray_tracing.rs:400)

// ---- sphere staging ------------------------------------------------------------------------------------------
// Device sphere = float4 {cx, cy, cz, r*r}.  The array in shared memory is padded to a multiple of SWEEP_BLOCK with
// entries whose r*r = -inf: their discriminant is -inf (or NaN), never a hit.
__device__ __forceinline__ void stage_spheres(float4* s_sph, const float4* __restrict__ g_sph, int begin, int count, int padded) {
    for (int i = threadIdx.x; i < padded; i += blockDim.x) {
        float4 v = make_float4(0.0f, 0.0f, 0.0f, -__int_as_float(0x7f800000));
        if (i < count) v = __ldg(&g_sph[begin + i]);
        s_sph[i] = v;
    }
}

// Exact per-candidate test: the body of hit_packed's `if discriminant_positive.any()` block for one lane
// (src/objects.rs:262-288) followed by PackedHitRecords::update (:140-147).
__device__ __forceinline__ void test_candidate(float4 s, int index, float ox, float oy, float oz, float dx, float dy, float dz,
                                               float a, float inv_a, float& best_t, int& best_i) {
    float ocx = __fsub_rn(ox, s.x), ocy = __fsub_rn(oy, s.y), ocz = __fsub_rn(oz, s.z);            // :252
    float hb = __fmaf_rn(ocz, dz, __fmaf_rn(ocy, dy, __fmul_rn(ocx, dx)));                        // :255
    float cc = __fsub_rn(__fmaf_rn(ocz, ocz, __fmaf_rn(ocy, ocy, __fmul_rn(ocx, ocx))), s.w);     // :256
    float disc = __fmaf_rn(hb, hb, -__fmul_rn(a, cc));                                            // :257
    if (disc >= 0.0f) {                                                                            // :259
        float sq = __fsqrt_rn(disc);                                                               // :263
        float root1 = __fmul_rn(__fsub_rn(-hb, sq), inv_a);                                        // :270
        bool valid = (root1 >= T_MIN) && (root1 < __int_as_float(0x7f800000));                     // :272-277 (near root only, F4)
        if (valid && root1 <= best_t) { best_t = root1; best_i = index; }                          // :141 (`<=`: last wins, F5)
    }
}

// Drains this lane's deferred candidate list in sweep order.  list entry = {sign mask of 32 discriminants, index of
// the first sphere of that block within the staged array}.
__device__ __forceinline__ void drain_candidates(const float4* s_sph, int index_base, uint2* list, int list_stride, int& n_list,
                                                 float ox, float oy, float oz, float dx, float dy, float dz, float a, float inv_a,
                                                 float& best_t, int& best_i) {
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
        test_candidate(s_sph[local], index_base + local, ox, oy, oz, dx, dy, dz, a, inv_a, best_t, best_i);
    }
    n_list = 0;
}

// The sweep over `padded` staged spheres (global indices index_base .. index_base+padded).
//
// Fast path per sphere (all lanes, no branch): the 12 FP32 operations of the discriminant in the reference's order
// — 3 FADD (oc), FMUL+2 FFMA (half_b), FMUL+2 FFMA+FADD (c), FMUL+FFMA (disc) = 17 FLOP — plus one funnel shift that
// appends the discriminant's sign bit to a 32-sphere mask.  A set sign bit means disc < 0 (or disc = -0.0, which only
// arises when half_b = +-0 and then root1 = +-0 < 0.001: not a hit either), so candidates are the clear bits.
// Slow path (deferred, lane-parallel): masks with a clear bit are pushed to a per-lane list in shared memory and
// re-tested exactly (sqrt, root, range, `<=`) after the sweep, in sweep order.
__device__ __forceinline__ void sweep_spheres(const float4* s_sph, int padded, int index_base, uint2* list, int list_stride,
                                              float ox, float oy, float oz, float dx, float dy, float dz, float a, float inv_a,
                                              float& best_t, int& best_i) {
    int n_list = 0;
    for (int blk = 0; blk < padded; blk += SWEEP_BLOCK) {
        uint32_t mask = 0;
#pragma unroll
        for (int j = 0; j < SWEEP_BLOCK; j++) {
            float4 s = s_sph[blk + j];                                                            // LDS.128 broadcast = Simd::splat
            float ocx = __fsub_rn(ox, s.x), ocy = __fsub_rn(oy, s.y), ocz = __fsub_rn(oz, s.z);
            float hb = __fmaf_rn(ocz, dz, __fmaf_rn(ocy, dy, __fmul_rn(ocx, dx)));
            float cc = __fsub_rn(__fmaf_rn(ocz, ocz, __fmaf_rn(ocy, ocy, __fmul_rn(ocx, ocx))), s.w);
            float disc = __fmaf_rn(hb, hb, -__fmul_rn(a, cc));
            mask = __funnelshift_l(__float_as_uint(disc), mask, 1);                               // mask = (mask << 1) | sign(disc)
        }
        if (mask != 0xffffffffu) {
            list[n_list * list_stride] = make_uint2(mask, (uint32_t)blk);
            n_list++;
            if (n_list == LIST_CAP)
                drain_candidates(s_sph, index_base, list, list_stride, n_list, ox, oy, oz, dx, dy, dz, a, inv_a, best_t, best_i);
        }
    }
    if (n_list > 0)
        drain_candidates(s_sph, index_base, list, list_stride, n_list, ox, oy, oz, dx, dy, dz, a, inv_a, best_t, best_i);
}

// ---- K1 ----------------------------------------------------------------------------------------------------------
// One thread per ray; spheres streamed through shared memory in tiles of `tile` (multiple of SWEEP_BLOCK).
// dynamic smem: tile float4 + LIST_CAP * blockDim uint2.
__global__ void __launch_bounds__(256) k_intersect_batch(const float4* __restrict__ g_sph, int n_spheres, int tile,
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
    float rox = 0.f, roy = 0.f, roz = 0.f, rdx = 0.f, rdy = 0.f, rdz = 0.f;
    if (live) { rox = ox[i]; roy = oy[i]; roz = oz[i]; rdx = dx[i]; rdy = dy[i]; rdz = dz[i]; }
    float a = plen2_3(mk3(rdx, rdy, rdz));                     // src/objects.rs:253
    float inv_a = __fdiv_rn(1.0f, a);                          // :254
    float best_t = __int_as_float(0x7f800000);                 // PackedHitRecords::default, src/objects.rs:121-134
    int best_i = -1;
    for (int begin = 0; begin < n_spheres; begin += tile) {
        int count = min(tile, n_spheres - begin);
        int padded = (count + SWEEP_BLOCK - 1) / SWEEP_BLOCK * SWEEP_BLOCK;
        __syncthreads();
        stage_spheres(s_sph, g_sph, begin, count, padded);
        __syncthreads();
        sweep_spheres(s_sph, padded, begin, list, blockDim.x, rox, roy, roz, rdx, rdy, rdz, a, inv_a, best_t, best_i);
    }
    if (live) {
        bool en = enabled ? enabled[i] != 0 : true;            // rays.enabled(), src/objects.rs:259,277
        hit_index[i] = en ? best_i : -1;
        hit_t[i] = (en && best_i >= 0) ? best_t : __int_as_float(0x7f800000);
    }
}

// ---- K2 ----------------------------------------------------------------------------------------------------------
struct TraceParams {
    DeviceCamera cam;
    const float4* spheres;        // {cx,cy,cz,r*r}
    const float4* mat;            // per sphere {albedo rgb, param}
    const int* mat_kind;          // per sphere kind | hollow<<8
    int n_spheres;
    int padded;                   // n_spheres rounded up to SWEEP_BLOCK
    int max_bounces;
    uint32_t seed_lo, seed_hi;
    int sample_begin, sample_end; // global sample indices traced by this launch
    int row_begin, row_end;
    int splits;                   // K: sample range is cut into K contiguous sub-ranges, one work item each
    float4* out;                  // splits == 1: accum image (W*H float4); else partial planes [K][rows*W]
    unsigned long long* work_counter;
    unsigned long long* counters; // [0] paths, [1] segments
};

// Persistent kernel.  Work item = (pixel of the row range, sample split).  A lane owns one item at a time and runs its
// samples one after another (path regeneration); the per-pixel sum is a private register sum in sample order, so the
// image is deterministic and needs no atomics.  dynamic smem: padded float4 + LIST_CAP * blockDim uint2.
template <int THREADS, int MIN_BLOCKS>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS) k_trace_paths(const TraceParams P) {
    extern __shared__ float4 smem4[];
    float4* s_sph = smem4;
    uint2* list = reinterpret_cast<uint2*>(smem4 + P.padded) + threadIdx.x;
    stage_spheres(s_sph, P.spheres, 0, P.n_spheres, P.padded);
    __syncthreads();

    const int W = P.cam.width;
    const unsigned long long n_pix = (unsigned long long)(P.row_end - P.row_begin) * (unsigned long long)W;
    const unsigned long long n_items = n_pix * (unsigned long long)P.splits;
    const int n_samples = P.sample_end - P.sample_begin;
    const unsigned lane = threadIdx.x & 31u;

    // lane state
    bool has_item = false, has_path = false, exhausted = false;
    unsigned long long item = 0;
    int col = 0, row = 0, s = 0, s_first = 0, s_end = 0, bounce = 0;
    uint32_t pixel_id = 0;
    f3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f);
    f3 thr = mk3(1.f, 1.f, 1.f), acc = mk3(0.f, 0.f, 0.f);
    unsigned long long n_segments = 0, n_paths = 0;

    for (;;) {
        // ---- regeneration: finished items are written out, new items claimed with one atomic per warp ----
        if (has_item && !has_path && s >= s_end) {
            unsigned long long split = item / n_pix, pix = item % n_pix;
            size_t off = (P.splits == 1) ? ((size_t)P.row_begin * W + pix) : (size_t)(split * n_pix + pix);
            P.out[off] = make_float4(acc.x, acc.y, acc.z, (float)(s_end - s_first));
            has_item = false;
        }
        bool need = !has_item && !exhausted;
        unsigned need_mask = __ballot_sync(0xffffffffu, need);
        if (need_mask) {
            unsigned long long base = 0;
            int leader = __ffs(need_mask) - 1;
            if ((int)lane == leader) base = atomicAdd(P.work_counter, (unsigned long long)__popc(need_mask));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (need) {
                item = base + __popc(need_mask & ((1u << lane) - 1u));
                if (item < n_items) {
                    unsigned long long split = item / n_pix, pix = item % n_pix;
                    row = P.row_begin + (int)(pix / W);
                    col = (int)(pix % W);
                    pixel_id = (uint32_t)row * (uint32_t)W + (uint32_t)col;
                    s_first = P.sample_begin + (int)(split * n_samples / P.splits);
                    s = s_first;
                    s_end = P.sample_begin + (int)((split + 1) * n_samples / P.splits);
                    acc = mk3(0.f, 0.f, 0.f);
                    has_item = true;
                    if (P.max_bounces <= 0) {               // `for _ in 0..0`: every ray is still enabled => black (:370)
                        n_paths += (unsigned long long)(s_end - s);
                        s = s_end;
                    }
                } else {
                    exhausted = true;
                }
            }
        }
        if (has_item && !has_path && s < s_end) {
            uint4 rnd = philox4x32_10(pixel_id, (uint32_t)s, RTB_DIM_CAMERA, 0u, P.seed_lo, P.seed_hi);
            camera_get_ray(P.cam, col, row, rnd, o, d);
            thr = mk3(1.f, 1.f, 1.f);
            bounce = 0;
            has_path = true;
            n_paths++;
        }
        if (!__any_sync(0xffffffffu, has_path)) {
            if (__all_sync(0xffffffffu, exhausted && !has_item)) break;
            continue;   // items with an empty sample range
        }

        // ---- closest-hit sweep (all lanes; lanes without a path sweep a dummy ray and ignore the result) ----
        float a = plen2_3(d);                                   // src/objects.rs:253
        float inv_a = __fdiv_rn(1.0f, a);                       // :254
        float best_t = __int_as_float(0x7f800000);
        int best_i = -1;
        sweep_spheres(s_sph, P.padded, 0, list, THREADS, o.x, o.y, o.z, d.x, d.y, d.z, a, inv_a, best_t, best_i);

        // ---- shade ----
        if (has_path) {
            n_segments++;
            if (best_i < 0) {                                   // hit the sky (src/ray_tracing.rs:353-357, :365-369)
                f3 sk = sky(d);
                acc = add3(acc, mk3(__fmul_rn(thr.x, sk.x), __fmul_rn(thr.y, sk.y), __fmul_rn(thr.z, sk.z)));
                has_path = false;
                s++;
            } else {
                float4 sp = s_sph[best_i];
                HitRec rec = finalize_hit(o, d, best_t, mk3(sp.x, sp.y, sp.z));
                uint4 rnd = philox4x32_10(pixel_id, (uint32_t)s, 1u + (uint32_t)bounce, 0u, P.seed_lo, P.seed_hi);
                f3 att, nd;
                scatter(__ldg(&P.mat_kind[best_i]), __ldg(&P.mat[best_i]), d, rec, rnd, att, nd);
                thr = mk3(__fmul_rn(thr.x, att.x), __fmul_rn(thr.y, att.y), __fmul_rn(thr.z, att.z));
                o = rec.p;
                d = nd;
                bounce++;
                if (bounce >= P.max_bounces) {                  // still enabled at the depth limit => black (:370)
                    has_path = false;
                    s++;
                }
            }
        }
    }

    // ---- counters: one atomic pair per warp ----
#pragma unroll
    for (int sh = 16; sh > 0; sh >>= 1) {
        n_segments += __shfl_xor_sync(0xffffffffu, n_segments, sh);
        n_paths += __shfl_xor_sync(0xffffffffu, n_paths, sh);
    }
    if (lane == 0) {
        atomicAdd(&P.counters[0], n_paths);
        atomicAdd(&P.counters[1], n_segments);
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
