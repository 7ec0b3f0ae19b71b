// rtb200.cu — C ABI (include/rtb200.h) over the sm_100a kernels in rt_kernels.cuh.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -Xcompiler -fPIC -shared
// There is no CPU path in this file: every compute entry point launches a kernel or fails.
#include "../../include/rtb200.h"
#include "rt_kernels.cuh"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#ifndef RTB_LR_RAYS
#define RTB_LR_RAYS 2       // path slots per lane of the register-state kernels (build-time tuning aid)
#endif
#ifndef RTB_LR_THREADS
#define RTB_LR_THREADS 640
#endif

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

#define CU_TRY(expr)                                                                                           \
    do {                                                                                                       \
        cudaError_t _e = (expr);                                                                               \
        if (_e != cudaSuccess)                                                                                 \
            return fail(RTB200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));                   \
    } while (0)

constexpr int SMALL_THREADS = 256;    // scenes whose staging leaves room for several CTAs per SM
constexpr int LARGE_THREADS = 1024;   // one CTA per SM, up to ~11k spheres resident in shared memory

size_t trace_smem_bytes(int n_spheres, int threads, int rays, int cap) {
    return (size_t)((n_spheres + 1) & ~1) * sizeof(float4) + (size_t)rays * cap * threads * sizeof(uint2);
}

}  // namespace

struct rtb200_ctx {
    int device = 0;
    cudaDeviceProp prop{};
    // scene
    void* d_scene = nullptr; size_t scene_capacity = 0;   // arena holding the four arrays below
    float4* d_spheres = nullptr;
    float4* d_filter = nullptr;
    std::vector<float4> h_filter;                      // host copy of the filter spheres (kernel-parameter variant)
    std::vector<float> h_slab;                         // stage-1 spheres {c_a, c_b, R'} padded to whole 32-blocks (two-stage variant)
    int hit_mode = 0;                                  // RTB200_HIT_*
    rtb200_progress_fn progress_cb = nullptr;          // rtb200_set_progress
    void* progress_user = nullptr;
    int progress_interval_ms = 100;
    unsigned long long* h_progress = nullptr;          // mapped pinned host memory: the kernel publishes its queue head here
    unsigned long long* d_progress = nullptr;          // its device-side address
    unsigned long long items_total = 0;                // work items of the last trace launch
    int up_axis = 1;                                   // the axis the slab test drops
    float slab_rp_min = 1.0f;                          // smallest floored R' of the scene (rt_filter.h make_slab_ray)
    double slab_pass = 1.0;                            // estimated fraction of pairs that pass stage 1
    std::vector<unsigned char> h_params = std::vector<unsigned char>(sizeof(rtb::TraceParamsV<1024>) + 64);
    float4* d_mat = nullptr;
    int* d_mat_kind = nullptr;
    double4* d_exact = nullptr;                        // {cx, cy, cz, r*r} in f64, list order, padded: the exact candidate test of every sweep
    float4* d_slab = nullptr;                          // stage-1 triples, n_padded * 3 floats (k_filter_check; the parameter-space kernels take h_slab)
    float4* d_slabq = nullptr;                         // the same in SmemSlab's layout (shared-memory and streamed two-stage kernels)
    int n_padded = 32;
    float h_bound = 0.0f;                              // max |c| + r over the scene, rounded up (TraceParams::h_bound)
    int largest_sphere = 0;                            // index of the sphere with the largest radius (rtb200_filter_check)
    int n_spheres = 0;
    // counters: [0] paths [1] segments [2] work counter
    unsigned long long* d_counters = nullptr;
    // render-call scratch
    float4* d_accum = nullptr; size_t accum_pixels = 0;
    uint8_t* d_rgb8 = nullptr; size_t rgb8_bytes = 0;
    float4* d_partial = nullptr; size_t partial_elems = 0;
    void* d_rays = nullptr; size_t rays_bytes = 0;
    float* d_probe = nullptr;
    cudaEvent_t ev[4]{};
    // the trace kernel ALONE (without k_combine), per launch: a ring of event pairs recorded inside launch_trace
    static constexpr int TRACE_RING = 64;
    cudaEvent_t ev_trace[TRACE_RING][2]{};
    int64_t trace_launches = 0;                        // trace kernels launched so far (ring index = count % TRACE_RING)
    const char* last_kernel = "";                      // the trace-kernel variant of the last launch
    int last_splits = 1;                               // K of the last launch: sample sub-ranges per pixel (partial planes summed in plane order)
    int64_t launches = 0;
    size_t max_optin_smem = 0;
    bool attrs_set = false;
    bool intersect_attr_set = false;
    int default_rays = 0;                              // 0 = choose by scene size
    std::vector<void*> exchange_allocs;                // buffers handed out by rtb200_device_alloc (rtb200_device_free accepts only these)
    std::vector<void*> ipc_mappings;                   // peer buffers mapped by rtb200_ipc_open (rtb200_ipc_close accepts only these)
};

namespace {

int ensure(void** p, size_t* have, size_t want) {
    if (*have >= want && *p) return RTB200_OK;
    if (*p) cudaFree(*p);
    *p = nullptr; *have = 0;
    CU_TRY(cudaMalloc(p, want));
    *have = want;
    return RTB200_OK;
}

rtb::f3 to_f3(const double v[3]) { return rtb::f3{(float)v[0], (float)v[1], (float)v[2]}; }

rtb::DeviceCamera device_camera(const rtb200_camera& c) {
    rtb::DeviceCamera d;
    d.width = c.image_width; d.height = c.image_height;
    d.center = to_f3(c.center); d.ulc = to_f3(c.viewport_upper_left_corner);
    d.viewport_u = to_f3(c.viewport_u); d.viewport_v = to_f3(c.viewport_v);
    d.defocus_u = to_f3(c.defocus_u); d.defocus_v = to_f3(c.defocus_v);
    auto nz = [](rtb::f3 a) { return a.x != 0.f || a.y != 0.f || a.z != 0.f; };
    d.defocus_enabled = (nz(d.defocus_u) || nz(d.defocus_v)) ? 1 : 0;
    return d;
}

int validate_params(const rtb200_ctx* ctx, const rtb200_camera* cam, const rtb200_params* p) {
    if (!ctx || !cam || !p) return fail(RTB200_ERR_INVALID, "null argument");
    if (!ctx->d_spheres) return fail(RTB200_ERR_STATE, "no scene uploaded (call rtb200_upload_scene first)");
    if (cam->image_width <= 0 || cam->image_height <= 0) return fail(RTB200_ERR_INVALID, "image size must be positive");
    if ((int64_t)cam->image_width * cam->image_height > (int64_t)1 << 31) return fail(RTB200_ERR_INVALID, "image too large");
    if (p->max_bounces < 0) return fail(RTB200_ERR_INVALID, "max_bounces must be >= 0");
    if (p->samples_per_pixel <= 0) return fail(RTB200_ERR_INVALID, "samples_per_pixel must be positive");
    if (p->sample_begin < 0 || p->sample_end < p->sample_begin) return fail(RTB200_ERR_INVALID, "bad sample range");
    if (p->row_begin < 0 || p->row_end < p->row_begin || p->row_end > cam->image_height) return fail(RTB200_ERR_INVALID, "bad row range");
    if (p->row_stride < 0) return fail(RTB200_ERR_INVALID, "row_stride must be >= 0");
    return RTB200_OK;
}

// Chooses log2 of the number of sample splits K (a power of two): the persistent grid wants many more work items than
// path slots (tail balance) while an item keeps enough samples to amortise its claim and its partial-plane write.
// Measured on C2 with the chunked claims of slot_step (profiles/r1_splits_sweep.txt): 100 spp K = 2/4/8/16/32 take
// 49.7/45.0/44.7/44.8/46.0 ms, 1000 spp K = 4/8/16/32: 444/434/433/432 ms.  RTB200_SPLITS_LOG2 overrides (tuning aid).
int choose_splits_log2(const rtb200_ctx* ctx, long long n_pix, int n_samples) {
    if (const char* e = std::getenv("RTB200_SPLITS_LOG2")) { int v = std::atoi(e); if (v >= 0 && v <= 6) return v; }
    const long long slots = (long long)ctx->prop.multiProcessorCount * 2048;   // 1024 threads x 2 path slots per SM
    int lg = 0;
    // (a) fill the machine: at least two items per path slot, down to 4 samples per item (small images: C1)
    while (lg < 6 && (n_pix << lg) < 2 * slots && (n_samples >> (lg + 1)) >= 4) lg++;
    // (b) tail balance: ~64 items per slot.  What matters is the NUMBER of items, not their length: on C2's geometry
    //     (bottom-up queue, profiles/r1_splits_sweep.txt) 8 spp K = 1/2/4/8 take 5.96/4.96/4.45/4.25 ms, 16 spp K = 2/4/8/16
    //     9.75/8.92/7.71/7.77, 25 spp K = 4/8/16 12.84/12.34/11.97, 50 spp K = 8/16/32 23.19/23.39/23.47 — so an item may
    //     shrink to a single sample.  Longer sweeps (>= 1000 spheres) keep two samples per item: 1 297 spheres at 16 spp
    //     K = 4/8/16 18.50/18.20/18.55 ms, 10 001 spheres 118.6/118.5/121.1.
    const int n_sph = ctx->n_spheres > 0 ? ctx->n_spheres : 1;
    const int min_samples = n_sph >= 1000 ? 2 : 1;
    while (lg < 6 && (n_pix << lg) < 64 * slots && (n_samples >> (lg + 1)) >= min_samples) lg++;
    return lg;
}

// Scratch override of launch_trace (rtb200_trace_device_static): the counters / work-queue head and the partial planes live
// in caller-named memory of fixed capacity instead of the context's growable buffers, and no event pair is recorded.
struct TraceScratch {
    unsigned long long* counters;   // 3 words: paths, segments, work-queue head
    float4* partial;                // partial planes of the sample splits
    size_t partial_elems;           // capacity in float4
};

int launch_trace(rtb200_ctx* ctx, const rtb200_camera* cam, const rtb200_params* p, float4* accum, cudaStream_t st,
                 const TraceScratch* ext = nullptr) {
    const int W = cam->image_width;
    unsigned long long* const d_counters = ext ? ext->counters : ctx->d_counters;
    const int row_stride = p->row_stride > 1 ? p->row_stride : 1;
    const long long n_pix = (long long)((p->row_end - p->row_begin + row_stride - 1) / row_stride) * W;   // traced rows x W
    CU_TRY(cudaMemsetAsync(d_counters, 0, 3 * sizeof(unsigned long long), st));
    if (n_pix == 0) return RTB200_OK;
    const int n_samples = p->sample_end - p->sample_begin;
    rtb::TraceParams P;
    P.cam = device_camera(*cam);
    P.spheres = ctx->d_spheres; P.filter = ctx->d_filter; P.exact = ctx->d_exact; P.h_bound = ctx->h_bound; P.mat = ctx->d_mat; P.mat_kind = ctx->d_mat_kind;
    P.n_spheres = ctx->n_spheres;
    P.max_bounces = p->max_bounces;
    P.seed_lo = (uint32_t)(p->seed & 0xffffffffu); P.seed_hi = (uint32_t)(p->seed >> 32);
    P.sample_begin = p->sample_begin; P.sample_end = p->sample_end;
    P.row_begin = p->row_begin; P.row_end = p->row_end; P.row_stride = row_stride;
    P.both_roots = (ctx->hit_mode == RTB200_HIT_BOTH_ROOTS || (p->flags & RTB200_FLAG_BOTH_ROOTS)) ? 1 : 0;
    P.splits_log2 = choose_splits_log2(ctx, n_pix, n_samples);
    const int splits = 1 << P.splits_log2;
    P.width_magic = W > 1 ? (uint32_t)((1ull << 32) / (unsigned long long)W) + 1u : 0u;
    {   // claim chunk: 64 items per warp and atomic on large launches (coherent lanes, measured best: 16/32/64/128/256 ->
        // 45.7/44.8/44.9/45.5/47.8 ms on C2); smaller when the launch has fewer than ~4 chunks per warp (balance; C1:
        // 1/2/4/8/16/64 -> 0.61/0.49/0.43/0.39/0.35/0.41 ms)
        const long long warps = (long long)ctx->prop.multiProcessorCount * 32;
        long long c = ((long long)n_pix * splits) / (warps * 4);
        int chunk = 1;
        while (chunk < 64 && chunk * 2 <= c) chunk *= 2;
        if (const char* e = std::getenv("RTB200_CLAIM_CHUNK")) { int v = std::atoi(e); if (v >= 1 && v <= 4096) chunk = v; }
        P.claim_chunk = chunk;
    }
    P.work_counter = d_counters + 2;
    P.progress = (ctx->progress_cb && ctx->h_progress) ? ctx->d_progress : nullptr;
    if (ctx->h_progress) *ctx->h_progress = 0ull;
    P.counters = d_counters;
    float4* d_partial = nullptr;
    if (splits > 1 && ext) {
        const size_t want = (size_t)splits * (size_t)n_pix;
        if (want > ext->partial_elems)
            return fail(RTB200_ERR_INVALID, "rtb200_trace_device_static: this shard needs " + std::to_string(want * sizeof(float4)) +
                        " bytes of partial planes (" + std::to_string(splits) + " sample splits x " + std::to_string(n_pix) +
                        " pixels x 16 B), beyond the static scratch maximum of " + std::to_string(ext->partial_elems * sizeof(float4)) +
                        " bytes: render fewer rows per call (row_begin/row_end) or use rtb200_trace_device");
        P.out = d_partial = ext->partial;
    } else if (splits > 1) {
        size_t want = (size_t)splits * (size_t)n_pix;
        if (ctx->partial_elems < want) {
            if (ctx->d_partial) cudaFree(ctx->d_partial);
            ctx->d_partial = nullptr; ctx->partial_elems = 0;
            CU_TRY(cudaMalloc(&ctx->d_partial, want * sizeof(float4)));
            ctx->partial_elems = want;
        }
        P.out = d_partial = ctx->d_partial;
    } else {
        P.out = accum;
    }
    if ((unsigned long long)n_pix * (unsigned long long)splits >= (1ull << 32) - 65536)
        return fail(RTB200_ERR_INVALID, "too many work items for one launch");
    ctx->items_total = (unsigned long long)n_pix * (unsigned long long)splits;
    // Kernel variants.
    //   v  : k_trace_paths_v<threads, CTAs/SM, rays per lane, list capacity, N> — filter spheres in parameter space, path
    //        state in shared memory; scenes of at most N spheres (N = 512 or 1024: up to 16 KB of kernel parameters).
    //   m1 : resident kernel, 256 threads x 4 CTAs/SM, spheres staged in shared memory (up to ~2 000 spheres).
    //   l1 : resident kernel, 1024 threads x 1 CTA/SM (up to ~11 000 spheres).
    constexpr int CG = rtb::LIST_CAP, CV = rtb::LIST_CAP_V, VT = 256, VR = 2, LR = RTB_LR_RAYS;
    using PV512 = rtb::TraceParamsV<512>;
    using PV1024 = rtb::TraceParamsV<1024>;
    using PS512 = rtb::TraceParamsS<512>;
    // Every trace kernel exists in two instantiations: contiguous band (STRIDED = false: every single-GPU render) and
    // interleaved scanlines (row_stride > 1).  KSEL picks by this launch's stride.
    const bool strided = row_stride > 1;
#define KSEL(...) (strided ? rtb::__VA_ARGS__, true> : rtb::__VA_ARGS__, false>)
#define KBOTH(...) { CU_TRY(cudaFuncSetAttribute(rtb::__VA_ARGS__, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->max_optin_smem)); \
                     CU_TRY(cudaFuncSetAttribute(rtb::__VA_ARGS__, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->max_optin_smem)); }
    constexpr int LT = 512;                            // shared-memory two-stage variant: one 512-thread CTA per SM
    constexpr int LRT = RTB_LR_THREADS;                // register-state variant of the same
    auto k_v512 = KSEL(k_trace_paths_v<VT, 4, VR, CV, PV512);
    auto k_v1024 = KSEL(k_trace_paths_v<VT, 4, VR, CV, PV1024);
    auto k_s512 = KSEL(k_trace_paths_v<VT, 4, VR, CV, PS512);
    auto k_sl = KSEL(k_trace_paths_v<LT, 1, VR, CV, rtb::TraceParamsL);
    auto k_slr = KSEL(k_trace_paths_l<LRT, LR);
    auto k_g = KSEL(k_trace_paths_g<LRT, LR);          // any scene size: stage-1 triples streamed per warp from global memory
    auto k_m1 = KSEL(k_trace_paths<SMALL_THREADS, 4, 1, CG);
    auto k_l1 = KSEL(k_trace_paths<LARGE_THREADS, 1, 1, CG);
    if (!ctx->attrs_set) {
        KBOTH(k_trace_paths_v<VT, 4, VR, CV, PV512);
        KBOTH(k_trace_paths_v<VT, 4, VR, CV, PV1024);
        KBOTH(k_trace_paths_v<VT, 4, VR, CV, PS512);
        KBOTH(k_trace_paths_v<LT, 1, VR, CV, rtb::TraceParamsL);
        KBOTH(k_trace_paths_l<LRT, LR);
        KBOTH(k_trace_paths_g<LRT, LR);
        KBOTH(k_trace_paths<SMALL_THREADS, 4, 1, CG);
        KBOTH(k_trace_paths<LARGE_THREADS, 1, 1, CG);
        ctx->attrs_set = true;
    }
#undef KSEL
#undef KBOTH
    const int sms = ctx->prop.multiProcessorCount;
    const size_t per_sm = ctx->prop.sharedMemPerMultiprocessor;
    const size_t smem_v = (size_t)(rtb::SLOT_WORDS + CV) * VR * VT * sizeof(uint32_t);
    const size_t smem_m1 = trace_smem_bytes(P.n_spheres, SMALL_THREADS, 1, CG);
    const size_t smem_l1 = trace_smem_bytes(P.n_spheres, LARGE_THREADS, 1, CG);
    int want = (p->flags & RTB200_FLAG_RAYS2) ? 2 : (p->flags & RTB200_FLAG_RAYS1) ? 1 : ctx->default_rays;
    const bool fits_v = want != 1 && (smem_v + 1024) * 4 <= per_sm;
    // two-stage (slab test first) when the scene's layout lets stage 1 reject most pairs; RTB200_FLAG_ONE_STAGE forces
    // the single-stage parameter-space kernel, RTB200_FLAG_TWO_STAGE the two-stage one.
    // Stage-1 triples go through kernel parameter space (uniform-register operands) for scenes of up to PARAM_MAX = 512
    // spheres (6 KB: they must stay resident in the SM's constant cache; beyond ~10 KB the indexed LDCU sweep thrashes
    // it: 1 297 spheres 24.3 vs 18.9 ms, 2 501: 56.8 vs 33.0).  Larger scenes stage them in shared memory and run the
    // packed-FMA stage 1 (SmemSlab), which since round 2 is as fast per pair as the uniform-operand form: 485 / 577 / 677
    // spheres at 64 spp, parameter space vs shared memory: 26.46 / 32.34 / 37.52 vs 26.56 / 32.08 / 36.98 ms; C2 (486
    // spheres, 100 spp) 43.55 vs 44.11, T 425.2 vs 430.5 in favour of parameter space (eight warps per sub-partition hide
    // the refine walks better than five).  RTB200_PARAM_MAX overrides (tuning aid).
    constexpr int PARAM_MAX = 512;
    int param_max = PARAM_MAX;
    if (const char* e = std::getenv("RTB200_PARAM_MAX")) { int v = std::atoi(e); if (v >= 0 && v <= PARAM_MAX) param_max = v; }
    const bool two_stage = fits_v && P.n_spheres <= param_max && !(p->flags & RTB200_FLAG_ONE_STAGE) &&
                           ((p->flags & RTB200_FLAG_TWO_STAGE) || ctx->slab_pass <= 0.12);
    const bool use_v = fits_v && !two_stage && P.n_spheres <= 1024;
    const size_t smem_sl = (size_t)(rtb::SLOT_WORDS + CV) * VR * LT * sizeof(uint32_t) + (size_t)ctx->n_padded * 3 * sizeof(float);
    const bool prefer_two = !(p->flags & RTB200_FLAG_ONE_STAGE) && ((p->flags & RTB200_FLAG_TWO_STAGE) || ctx->slab_pass <= 0.12);
    const bool two_stage_smem = want != 1 && P.n_spheres > param_max && smem_sl <= ctx->max_optin_smem && prefer_two;
    // Streamed two-stage kernel: scenes whose stage-1 triples do not fit in shared memory at once (beyond ~10 600
    // spheres), whatever their layout (both filters are conservative, so it is correct for any scene; it is also the
    // last resort of the chain below), and RTB200_FLAG_STREAMED.
    const bool streamed = (p->flags & RTB200_FLAG_STREAMED) || (want != 1 && P.n_spheres > param_max && smem_sl > ctx->max_optin_smem && prefer_two);
    auto launch_streamed = [&]() {
        rtb::TraceParamsL pl;
        std::memset(&pl, 0, sizeof pl);
        pl.b = P; pl.up_axis = ctx->up_axis; pl.rp_min = ctx->slab_rp_min; pl.n_padded = ctx->n_padded; pl.s1 = ctx->d_slabq;
        const size_t smem_g = (size_t)8 * LR * LRT * sizeof(uint32_t) + (size_t)(LRT / 32) * 2 * rtb::STREAM_CHUNK_F4 * sizeof(float4);
        k_g<<<sms, LRT, smem_g, st>>>(pl);
    };
    auto fill_s = [&](auto* ps) {
        ps->b = P;
        ps->up_axis = ctx->up_axis;
        ps->rp_min = ctx->slab_rp_min;
        std::memcpy(ps->s1, ctx->h_slab.data(), ctx->h_slab.size() * sizeof(float));
    };
    cudaEvent_t* ring = ctx->ev_trace[ctx->trace_launches % rtb200_ctx::TRACE_RING];
    if (!ext) {
        if (!ring[0]) { CU_TRY(cudaEventCreate(&ring[0])); CU_TRY(cudaEventCreate(&ring[1])); }
        CU_TRY(cudaEventRecord(ring[0], st));
    }
    ctx->last_splits = splits;
    if (streamed) {
        ctx->last_kernel = "k_trace_paths_g<640,2> (two-stage, stage-1 triples streamed per warp from global memory)";
        launch_streamed();
    } else if (two_stage) {
        ctx->last_kernel = "k_trace_paths_v<256,4,2,8,TraceParamsS<512>> (two-stage, stage-1 triples in kernel parameter space)";
        auto* ps = reinterpret_cast<PS512*>(ctx->h_params.data());
        fill_s(ps);
        k_s512<<<sms * 4, VT, smem_v, st>>>(*ps);
    } else if (two_stage_smem) {
        rtb::TraceParamsL pl;
        std::memset(&pl, 0, sizeof pl);
        pl.b = P; pl.up_axis = ctx->up_axis; pl.rp_min = ctx->slab_rp_min; pl.n_padded = ctx->n_padded; pl.s1 = ctx->d_slabq;
        const size_t smem_slr = (size_t)rtb::L_CHUNK_BLOCKS * LR * LRT * sizeof(uint32_t) + (size_t)ctx->n_padded * 3 * sizeof(float);
        if (smem_slr <= ctx->max_optin_smem && !(p->flags & RTB200_FLAG_SHARED_STATE)) {
            ctx->last_kernel = "k_trace_paths_l<640,2> (two-stage, stage-1 triples staged in shared memory, path state in registers)";
            k_slr<<<sms, LRT, smem_slr, st>>>(pl);
        } else {
            ctx->last_kernel = "k_trace_paths_v<512,1,2,8,TraceParamsL> (two-stage, stage-1 triples and path state in shared memory)";
            k_sl<<<sms, LT, smem_sl, st>>>(pl);
        }
    } else if (use_v && P.n_spheres <= 512) {
        ctx->last_kernel = "k_trace_paths_v<256,4,2,8,TraceParamsV<512>> (one-stage filter sweep from kernel parameter space)";
        auto* pv = reinterpret_cast<PV512*>(ctx->h_params.data());
        pv->b = P;
        std::memcpy(pv->flt, ctx->h_filter.data(), (size_t)P.n_spheres * sizeof(float4));
        k_v512<<<sms * 4, VT, smem_v, st>>>(*pv);
    } else if (use_v) {
        ctx->last_kernel = "k_trace_paths_v<256,4,2,8,TraceParamsV<1024>> (one-stage filter sweep from kernel parameter space)";
        auto* pv = reinterpret_cast<PV1024*>(ctx->h_params.data());
        pv->b = P;
        std::memcpy(pv->flt, ctx->h_filter.data(), (size_t)P.n_spheres * sizeof(float4));
        k_v1024<<<sms * 4, VT, smem_v, st>>>(*pv);
    } else if ((smem_m1 + 1024) * 4 <= per_sm) {
        ctx->last_kernel = "k_trace_paths<256,4,1,6> (one-stage filter sweep, spheres staged in shared memory)";
        k_m1<<<sms * 4, SMALL_THREADS, smem_m1, st>>>(P);
    } else if (smem_l1 <= ctx->max_optin_smem) {
        ctx->last_kernel = "k_trace_paths<1024,1,1,6> (one-stage filter sweep, spheres staged in shared memory)";
        k_l1<<<sms, LARGE_THREADS, smem_l1, st>>>(P);
    } else {
        ctx->last_kernel = "k_trace_paths_g<640,2> (two-stage, stage-1 triples streamed per warp from global memory)";
        launch_streamed();                                 // nothing else holds the scene
    }
    CU_TRY(cudaGetLastError());
    if (!ext) {
        CU_TRY(cudaEventRecord(ring[1], st));
        ctx->trace_launches++;
    }
    ctx->launches++;
    if (splits > 1) {
        int blocks = (int)((n_pix + 255) / 256);
        rtb::k_combine<<<blocks, 256, 0, st>>>(d_partial, splits, n_pix, accum + (size_t)p->row_begin * W, W, (long long)(row_stride - 1) * W);
        CU_TRY(cudaGetLastError());
        ctx->launches++;
    }
    return RTB200_OK;
}

int launch_resolve(rtb200_ctx* ctx, const float4* accum, long long n_pix, int spp, uint8_t* rgb8, cudaStream_t st) {
    if (n_pix <= 0) return RTB200_OK;
    long long quads = (n_pix + 3) / 4;
    int blocks = (int)((quads + 255) / 256);
    rtb::k_resolve<<<blocks, 256, 0, st>>>(accum, n_pix, (float)spp, rgb8);
    CU_TRY(cudaGetLastError());
    ctx->launches++;
    return RTB200_OK;
}

int launch_intersect(rtb200_ctx* ctx, const float* ox, const float* oy, const float* oz, const float* dx, const float* dy,
                     const float* dz, const uint8_t* enabled, long long n, int* idx, float* t, cudaStream_t st) {
    if (n <= 0) return RTB200_OK;
    const int threads = 256;
    int padded_all = (ctx->n_spheres + rtb::SWEEP_BLOCK - 1) / rtb::SWEEP_BLOCK * rtb::SWEEP_BLOCK;
    int tile = padded_all < 2048 ? padded_all : 2048;    // multiple of SWEEP_BLOCK: only the last tile has a remainder
    if (tile < rtb::SWEEP_BLOCK) tile = rtb::SWEEP_BLOCK;
    size_t smem = (size_t)tile * sizeof(float4) + (size_t)rtb::LIST_CAP * threads * sizeof(uint2);
    if (!ctx->intersect_attr_set) {                    // function attributes are per device: one flag per context
        CU_TRY(cudaFuncSetAttribute(rtb::k_intersect_batch, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctx->max_optin_smem));
        ctx->intersect_attr_set = true;
    }
    long long blocks = (n + threads - 1) / threads;
    if (blocks > 0x7fffffffLL) return fail(RTB200_ERR_INVALID, "too many rays in one batch");
    rtb::k_intersect_batch<<<(unsigned)blocks, threads, smem, st>>>(ctx->d_exact, ctx->d_filter, ctx->n_spheres, tile, ox, oy, oz, dx, dy, dz, enabled, n, idx, t,
                                                                  ctx->hit_mode == RTB200_HIT_BOTH_ROOTS ? 1 : 0);
    CU_TRY(cudaGetLastError());
    ctx->launches++;
    return RTB200_OK;
}

}  // namespace

extern "C" {

const char* rtb200_last_error(void) { return g_last_error.c_str(); }

int rtb200_create(int32_t device, rtb200_ctx** out) {
    if (!out) return fail(RTB200_ERR_INVALID, "out is null");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(RTB200_ERR_CUDA, std::string("no CUDA device available: ") + cudaGetErrorString(e) + " (rtb200 has no CPU fallback)");
    if (device < 0 || device >= count) return fail(RTB200_ERR_INVALID, "device index out of range");
    CU_TRY(cudaSetDevice(device));
    rtb200_ctx* c = new rtb200_ctx();
    c->device = device;
    e = cudaGetDeviceProperties(&c->prop, device);
    if (e != cudaSuccess) { delete c; return fail(RTB200_ERR_CUDA, cudaGetErrorString(e)); }
    if (c->prop.major < 10) {
        std::string msg = "device compute capability " + std::to_string(c->prop.major) + "." + std::to_string(c->prop.minor) + " < 10.0: this library carries sm_100a code only";
        delete c;
        return fail(RTB200_ERR_CUDA, msg);
    }
    c->max_optin_smem = c->prop.sharedMemPerBlockOptin;
    e = cudaMalloc(&c->d_counters, 3 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(c->d_counters, 0, 3 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc(&c->d_probe, sizeof(float));
    for (int i = 0; i < 4 && e == cudaSuccess; i++) e = cudaEventCreate(&c->ev[i]);
    if (e != cudaSuccess) { rtb200_destroy(c); return fail(RTB200_ERR_CUDA, cudaGetErrorString(e)); }
    *out = c;
    return RTB200_OK;
}

void rtb200_destroy(rtb200_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    for (void* m : c->ipc_mappings) cudaIpcCloseMemHandle(m);     // what the caller left open
    for (void* b : c->exchange_allocs) cudaFree(b);
    cudaFree(c->d_scene); cudaFree(c->d_counters);
    cudaFree(c->d_accum); cudaFree(c->d_rgb8); cudaFree(c->d_partial); cudaFree(c->d_rays); cudaFree(c->d_probe);
    for (int i = 0; i < 4; i++) if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    for (auto& pr : c->ev_trace) for (auto& e : pr) if (e) cudaEventDestroy(e);
    if (c->h_progress) cudaFreeHost(c->h_progress);
    delete c;
}

int rtb200_host_alloc(size_t bytes, void** out) {
    if (!out) return fail(RTB200_ERR_INVALID, "out is null");
    *out = nullptr;
    CU_TRY(cudaHostAlloc(out, bytes > 0 ? bytes : 1, cudaHostAllocDefault));
    return RTB200_OK;
}

void rtb200_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

int rtb200_device_info(rtb200_ctx* ctx, int32_t* sm_count, int32_t* sm_clock_khz, int32_t* cc_major, int32_t* cc_minor) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "ctx is null");
    if (sm_count) *sm_count = ctx->prop.multiProcessorCount;
    if (sm_clock_khz) *sm_clock_khz = ctx->prop.clockRate;
    if (cc_major) *cc_major = ctx->prop.major;
    if (cc_minor) *cc_minor = ctx->prop.minor;
    return RTB200_OK;
}

// Camera::new, src/ray_tracing.rs:27-62 — f64 on the host, like the reference.
int rtb200_camera_init(rtb200_camera* out, int32_t image_width, int32_t image_height, double focal_length, double view_angle_deg,
                       const double center[3], const double look_at[3], const double up[3], double defocus_angle_deg) {
    if (!out || !center || !look_at || !up) return fail(RTB200_ERR_INVALID, "null argument");
    if (image_width <= 0 || image_height <= 0) return fail(RTB200_ERR_INVALID, "image size must be positive");
    struct D3 { double x, y, z; };
    auto sub = [](D3 a, D3 b) { return D3{a.x - b.x, a.y - b.y, a.z - b.z}; };
    auto scale = [](D3 a, double s) { return D3{a.x * s, a.y * s, a.z * s}; };
    auto divs = [](D3 a, double s) { return D3{a.x / s, a.y / s, a.z / s}; };
    auto unit = [&](D3 a) { return divs(a, std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z)); };
    auto cross = [](D3 a, D3 b) { return D3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; };
    const double deg = 3.14159265358979323846 / 180.0;
    D3 c{center[0], center[1], center[2]}, la{look_at[0], look_at[1], look_at[2]}, upv{up[0], up[1], up[2]};
    double aspect = (double)image_width / (double)image_height;
    double vh = std::tan((view_angle_deg * deg) / 2.0) * focal_length * 2.0;
    double vw = vh * aspect;
    D3 dir = unit(sub(la, c));
    D3 w{-dir.x, -dir.y, -dir.z};
    D3 u = unit(cross(upv, w));
    D3 v = cross(w, u);
    D3 vu = scale(u, vw);
    D3 vv = scale(D3{-v.x, -v.y, -v.z}, vh);
    D3 ulc = sub(sub(sub(c, scale(w, focal_length)), divs(vu, 2.0)), divs(vv, 2.0));
    double dr = focal_length * std::tan((defocus_angle_deg / 2.0) * deg);
    D3 du = scale(u, dr), dv = scale(v, dr);
    out->image_width = image_width; out->image_height = image_height;
    auto put = [](double* dst, D3 a) { dst[0] = a.x; dst[1] = a.y; dst[2] = a.z; };
    put(out->center, c); put(out->viewport_upper_left_corner, ulc); put(out->viewport_u, vu); put(out->viewport_v, vv);
    put(out->defocus_u, du); put(out->defocus_v, dv);
    return RTB200_OK;
}

// The flattened scene as the device wants it, built ONCE on the host (rtb200_multi_upload_scene sends the same arena to every
// device): arena = [exact f64 | spheres | filter | materials | stage-1 triples | kinds], each padded to whole 32-sphere blocks.
struct HostScene {
    int n = 0;
    size_t cnt = 32;
    std::vector<unsigned char> arena;
    size_t off_sph = 0, off_flt = 0, off_mat = 0, off_slab = 0, off_kind = 0;
    std::vector<float4> filter;                        // n entries (kernel-parameter variant of the one-stage sweep)
    std::vector<float> slab;                           // cnt * 3 (kernel-parameter variant of the two-stage sweep)
    size_t off_slabq = 0;                              // the same data for the shared-memory kernels (SmemSlab's layout)
    int up_axis = 1, largest = 0;
    float rp_min = 1.0f, h_bound = 0.0f;
    double slab_pass = 1.0;
};

static int build_host_scene(const rtb200_scene* scene, HostScene& hs) {
    if (!scene) return fail(RTB200_ERR_INVALID, "null argument");
    if (scene->n_spheres < 0 || scene->n_materials < 0) return fail(RTB200_ERR_INVALID, "negative count");
    if (scene->n_spheres > 0 && (!scene->spheres || !scene->materials)) return fail(RTB200_ERR_INVALID, "null scene arrays");
    const int n = scene->n_spheres;
    // device arrays are padded to whole 32-sphere blocks with spheres that can never be hit: f64 exact {0,0,0,r^2 = NaN}
    // (the discriminant is NaN, never >= 0, for every ray), filter {0,0,0,k = +inf} (never a candidate)
    const size_t cnt = (size_t)((n + 31) / 32 * 32 > 0 ? (n + 31) / 32 * 32 : 32);
    std::vector<float4> sph(cnt, make_float4(0.f, 0.f, 0.f, -1.0f)), flt(cnt, make_float4(0.f, 0.f, 0.f, INFINITY)),
        mat(cnt, make_float4(0.f, 0.f, 0.f, 0.f));
    std::vector<double4> ex(cnt, make_double4(0.0, 0.0, 0.0, std::nan("")));
    std::vector<int> kind(cnt, 0);
    for (int i = 0; i < n; i++) {
        const rtb200_sphere& s = scene->spheres[i];
        if (s.material < 0 || s.material >= scene->n_materials)
            return fail(RTB200_ERR_INVALID, "sphere " + std::to_string(i) + " names material " + std::to_string(s.material) + " which does not exist");
        const rtb200_material& m = scene->materials[s.material];
        float r = (float)s.radius;
        sph[i] = make_float4((float)s.center[0], (float)s.center[1], (float)s.center[2], r * r);
        ex[i] = make_double4(s.center[0], s.center[1], s.center[2], s.radius * s.radius);           // radius.powi(2) in f64, src/objects.rs:256
        // The f32 path squares coordinates: keep them far from overflow (the reference is f64 and has no such limit).
        if (!(std::fabs(sph[i].x) <= 1e15f && std::fabs(sph[i].y) <= 1e15f && std::fabs(sph[i].z) <= 1e15f && std::fabs(r) <= 1e15f))
            return fail(RTB200_ERR_INVALID, "sphere " + std::to_string(i) + ": centre/radius not finite or beyond 1e15 (f32 device path)");
        flt[i] = make_float4(sph[i].x, sph[i].y, sph[i].z, rtb::filter_sphere_k(sph[i].x, sph[i].y, sph[i].z, sph[i].w));
        double param = m.param;
        if (m.kind == RTB200_METAL) param = (param < 1.0) ? param : 1.0;                           // Metal::new, src/materials.rs:79-88
        else if (m.kind != RTB200_LAMBERTIAN && m.kind != RTB200_DIELECTRIC)
            return fail(RTB200_ERR_INVALID, "Unknown material type " + std::to_string(m.kind));     // src/materials.rs:31
        mat[i] = make_float4((float)m.albedo[0], (float)m.albedo[1], (float)m.albedo[2], (float)param);
        kind[i] = m.kind | ((m.kind == RTB200_DIELECTRIC && m.hollow) ? RTB_HOLLOW_BIT : 0);
    }
    {   // stage-1 data of the two-stage sweep: drop the axis along which the centres spread least (robust spread: the
        // 10%..90% range, so that one huge "ground" sphere does not decide), estimate how many pairs will pass
        int up = 1;
        double spread[3] = {0, 0, 0};
        if (n > 0) {
            std::vector<float> v((size_t)n);
            for (int ax = 0; ax < 3; ax++) {
                for (int i = 0; i < n; i++) v[i] = ax == 0 ? sph[i].x : ax == 1 ? sph[i].y : sph[i].z;
                const size_t lo = (size_t)((n - 1) * 0.1), hi = (size_t)((n - 1) * 0.9);
                std::nth_element(v.begin(), v.begin() + hi, v.end());
                const float vhi = v[hi];
                std::nth_element(v.begin(), v.begin() + lo, v.begin() + hi + 1);
                spread[ax] = (double)vhi - (double)v[lo];
            }
            up = spread[0] <= spread[1] && spread[0] <= spread[2] ? 0 : (spread[1] <= spread[2] ? 1 : 2);
        }
        const int ia = (up + 1) % 3, ib = (up + 2) % 3;
        hs.slab.assign(cnt * 3, 0.0f);
        const double extent = std::max(1e-30, 0.5 * (spread[ia] + spread[ib]));
        for (size_t i = 0; i < cnt; i++) {
            float* o3 = &hs.slab[i * 3];
            if ((int)i < n) {
                const float c[3] = {sph[i].x, sph[i].y, sph[i].z};
                o3[0] = c[ia]; o3[1] = c[ib];
                o3[2] = rtb::slab_sphere_r(sph[i].x, sph[i].y, sph[i].z, sph[i].w);
            } else {
                o3[2] = -INFINITY;                        // padding: never a candidate
            }
        }
        hs.rp_min = rtb::slab_apply_floor(hs.slab.data() + 2, n, 3);
        // expected pass rate of stage 1: disc diameter over the scene's extent, inflated as for a ray that starts one
        // extent away from the origin (the 1 + M_r/R_min of make_slab_ray)
        const double inflate = 1.0 + (double)RTB_FILTER_MARGIN * extent * extent / (double)hs.rp_min;
        double pass = 0.0;
        for (int i = 0; i < n; i++) pass += std::min(1.0, 2.0 * std::sqrt((double)hs.slab[(size_t)i * 3 + 2] * inflate) / extent);
        hs.up_axis = up;
        hs.slab_pass = n > 0 ? pass / n : 1.0;
    }
    hs.n = n; hs.cnt = cnt;
    hs.off_sph = cnt * sizeof(double4); hs.off_flt = hs.off_sph + cnt * sizeof(float4); hs.off_mat = hs.off_flt + cnt * sizeof(float4);
    hs.off_slab = hs.off_mat + cnt * sizeof(float4); hs.off_kind = hs.off_slab + cnt * 3 * sizeof(float);
    hs.off_slabq = hs.off_kind + cnt * sizeof(int);
    hs.arena.resize(hs.off_slabq + cnt * 3 * sizeof(float));   // cnt is a multiple of 32, so every array starts 32-byte aligned
    std::memcpy(hs.arena.data(), ex.data(), cnt * sizeof(double4));
    std::memcpy(hs.arena.data() + hs.off_sph, sph.data(), cnt * sizeof(float4));
    std::memcpy(hs.arena.data() + hs.off_flt, flt.data(), cnt * sizeof(float4));
    std::memcpy(hs.arena.data() + hs.off_mat, mat.data(), cnt * sizeof(float4));
    std::memcpy(hs.arena.data() + hs.off_slab, hs.slab.data(), cnt * 3 * sizeof(float));
    std::memcpy(hs.arena.data() + hs.off_kind, kind.data(), cnt * sizeof(int));
    {   // SmemSlab's layout (rt_kernels.cuh): per four spheres {c_a x4}, {c_b x4}, {-R'' x4}, R'' = R' one ulp up
        float* q = reinterpret_cast<float*>(hs.arena.data() + hs.off_slabq);
        for (size_t g = 0; g < cnt / 4; g++)
            for (int k = 0; k < 4; k++) {
                const float* t = &hs.slab[(g * 4 + k) * 3];
                q[g * 12 + k] = t[0]; q[g * 12 + 4 + k] = t[1];
                q[g * 12 + 8 + k] = -std::nextafterf(t[2], INFINITY);      // padding: R' = -inf -> -R'' = +max: never a candidate
            }
    }
    hs.filter.assign(flt.begin(), flt.begin() + n);
    hs.largest = 0;
    double hb = 0.0;
    for (int i = 0; i < n; i++) {
        hb = std::max(hb, std::sqrt((double)sph[i].x * sph[i].x + (double)sph[i].y * sph[i].y + (double)sph[i].z * sph[i].z) + std::fabs(scene->spheres[i].radius));
        if (sph[i].w > sph[hs.largest].w) hs.largest = i;
    }
    hs.h_bound = std::nextafterf((float)(hb * (1.0 + 1e-6)), INFINITY);
    return RTB200_OK;
}

// Sends the arena to ctx's device on `st` (one copy, reusing the device arena while it is large enough) and points the
// context at it.  The copy is asynchronous when the caller says so (the source is pageable memory that must stay alive until
// the stream has been synchronised: rtb200_multi_upload_scene overlaps the copies to all devices that way).
static int commit_host_scene(rtb200_ctx* ctx, const HostScene& hs, cudaStream_t st, bool sync) {
    CU_TRY(cudaSetDevice(ctx->device));
    const size_t bytes = hs.arena.size();
    if (ctx->scene_capacity < bytes) {
        cudaFree(ctx->d_scene);
        ctx->d_scene = nullptr; ctx->scene_capacity = 0;
        ctx->d_spheres = nullptr; ctx->d_filter = nullptr; ctx->d_mat = nullptr; ctx->d_mat_kind = nullptr; ctx->d_slab = nullptr; ctx->d_slabq = nullptr; ctx->d_exact = nullptr;
        CU_TRY(cudaMalloc(&ctx->d_scene, bytes));
        ctx->scene_capacity = bytes;
    }
    CU_TRY(cudaMemcpyAsync(ctx->d_scene, hs.arena.data(), bytes, cudaMemcpyHostToDevice, st));
    if (sync) CU_TRY(cudaStreamSynchronize(st));
    unsigned char* base = static_cast<unsigned char*>(ctx->d_scene);
    ctx->d_exact = reinterpret_cast<double4*>(base);
    ctx->d_spheres = reinterpret_cast<float4*>(base + hs.off_sph);
    ctx->d_filter = reinterpret_cast<float4*>(base + hs.off_flt);
    ctx->d_mat = reinterpret_cast<float4*>(base + hs.off_mat);
    ctx->d_slab = reinterpret_cast<float4*>(base + hs.off_slab);
    ctx->d_slabq = reinterpret_cast<float4*>(base + hs.off_slabq);
    ctx->d_mat_kind = reinterpret_cast<int*>(base + hs.off_kind);
    ctx->n_padded = (int)hs.cnt;
    ctx->n_spheres = hs.n;
    ctx->h_filter = hs.filter;
    ctx->h_slab = hs.slab;
    ctx->up_axis = hs.up_axis; ctx->slab_rp_min = hs.rp_min; ctx->slab_pass = hs.slab_pass;
    ctx->largest_sphere = hs.largest; ctx->h_bound = hs.h_bound;
    return RTB200_OK;
}

int rtb200_upload_scene(rtb200_ctx* ctx, const rtb200_scene* scene) {
    if (!ctx || !scene) return fail(RTB200_ERR_INVALID, "null argument");
    HostScene hs;
    int rc = build_host_scene(scene, hs);
    if (rc) return rc;
    return commit_host_scene(ctx, hs, 0, true);
}

int rtb200_trace_device(rtb200_ctx* ctx, const rtb200_camera* camera, const rtb200_params* params, void* accum_dev, void* stream) {
    int rc = validate_params(ctx, camera, params);
    if (rc) return rc;
    if (!accum_dev) return fail(RTB200_ERR_INVALID, "accum_dev is null");
    CU_TRY(cudaSetDevice(ctx->device));
    return launch_trace(ctx, camera, params, static_cast<float4*>(accum_dev), static_cast<cudaStream_t>(stream));
}

// ---- rtb200_trace_device_static: the trace step with its scratch in a file-scope device array (no allocation) -------------
// Layout of the array: a 256-byte header (3 counter words: paths, segments, work-queue head) followed by the partial planes.
// One array per GPU that loads this library, held until the process exits; every call on that GPU shares it.
#define RTB200_STATIC_HEADER_BYTES 256
static __device__ __align__(256) unsigned char g_static_scratch[RTB200_STATIC_HEADER_BYTES + RTB200_STATIC_SCRATCH_BYTES];

namespace {
// Device address of g_static_scratch per CUDA device.  In host code the NAME g_static_scratch is only a host-side stand-in
// for the symbol (not a device address): it is resolved here once per device, with that device current, and only the
// resolved pointer is handed to kernels and memsets.
constexpr int MAX_DEVICES_STATIC = 64;
unsigned char* g_static_scratch_dev[MAX_DEVICES_STATIC] = {};

int static_scratch_for(int device, unsigned char** out) {
    if (device < 0 || device >= MAX_DEVICES_STATIC) return fail(RTB200_ERR_INVALID, "device index beyond the static-scratch table");
    if (!g_static_scratch_dev[device]) {
        CU_TRY(cudaSetDevice(device));
        void* p = nullptr;
        CU_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return fail(RTB200_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_static_scratch_dev[device] = static_cast<unsigned char*>(p);
    }
    *out = g_static_scratch_dev[device];
    return RTB200_OK;
}
}  // namespace

int rtb200_trace_device_static(rtb200_ctx* ctx, const rtb200_camera* camera, const rtb200_params* params, void* accum_dev, void* stream,
                               uint64_t* scratch_bytes_used) {
    int rc = validate_params(ctx, camera, params);
    if (rc) return rc;
    if (!accum_dev) return fail(RTB200_ERR_INVALID, "accum_dev is null");
    CU_TRY(cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    rc = static_scratch_for(ctx->device, &base);
    if (rc) return rc;
    TraceScratch sc;
    sc.counters = reinterpret_cast<unsigned long long*>(base);
    sc.partial = reinterpret_cast<float4*>(base + RTB200_STATIC_HEADER_BYTES);
    sc.partial_elems = (size_t)RTB200_STATIC_SCRATCH_BYTES / sizeof(float4);
    rc = launch_trace(ctx, camera, params, static_cast<float4*>(accum_dev), static_cast<cudaStream_t>(stream), &sc);
    if (rc) return rc;
    if (scratch_bytes_used) {
        const int rs = params->row_stride > 1 ? params->row_stride : 1;
        const uint64_t n_pix = (uint64_t)((params->row_end - params->row_begin + rs - 1) / rs) * (uint64_t)camera->image_width;
        *scratch_bytes_used = RTB200_STATIC_HEADER_BYTES + (n_pix > 0 && ctx->last_splits > 1 ? n_pix * (uint64_t)ctx->last_splits * sizeof(float4) : 0);
    }
    return RTB200_OK;
}

int rtb200_resolve_device(rtb200_ctx* ctx, const void* accum_dev, int64_t n_pixels, int32_t samples_per_pixel, void* rgb8_dev, void* stream) {
    if (!ctx || !accum_dev || !rgb8_dev) return fail(RTB200_ERR_INVALID, "null argument");
    if (n_pixels < 0 || samples_per_pixel <= 0) return fail(RTB200_ERR_INVALID, "bad size");
    CU_TRY(cudaSetDevice(ctx->device));
    return launch_resolve(ctx, static_cast<const float4*>(accum_dev), n_pixels, samples_per_pixel, static_cast<uint8_t*>(rgb8_dev), static_cast<cudaStream_t>(stream));
}

int rtb200_intersect_batch_device(rtb200_ctx* ctx, const float* ox, const float* oy, const float* oz, const float* dx, const float* dy,
                                  const float* dz, const uint8_t* enabled, int64_t n, int32_t* hit_index, float* hit_t, void* stream) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "ctx is null");
    if (!ctx->d_spheres) return fail(RTB200_ERR_STATE, "no scene uploaded (call rtb200_upload_scene first)");
    if (n < 0) return fail(RTB200_ERR_INVALID, "negative ray count");
    if (n > 0 && (!ox || !oy || !oz || !dx || !dy || !dz || !hit_index || !hit_t)) return fail(RTB200_ERR_INVALID, "null ray buffer");
    CU_TRY(cudaSetDevice(ctx->device));
    return launch_intersect(ctx, ox, oy, oz, dx, dy, dz, enabled, n, hit_index, hit_t, static_cast<cudaStream_t>(stream));
}

int rtb200_intersect_batch(rtb200_ctx* ctx, const float* ox, const float* oy, const float* oz, const float* dx, const float* dy,
                           const float* dz, const uint8_t* enabled, int64_t n, int32_t* hit_index, float* hit_t) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "ctx is null");
    if (!ctx->d_spheres) return fail(RTB200_ERR_STATE, "no scene uploaded (call rtb200_upload_scene first)");
    if (n < 0) return fail(RTB200_ERR_INVALID, "negative ray count");
    if (n == 0) return RTB200_OK;
    if (!ox || !oy || !oz || !dx || !dy || !dz || !hit_index || !hit_t) return fail(RTB200_ERR_INVALID, "null ray buffer");
    CU_TRY(cudaSetDevice(ctx->device));
    // layout of the device scratch: 6 float planes, idx plane, t plane, enabled bytes
    size_t nn = (size_t)n;
    size_t want = nn * (6 * sizeof(float) + sizeof(int) + sizeof(float) + 1);
    int rc = ensure(&ctx->d_rays, &ctx->rays_bytes, want);
    if (rc) return rc;
    float* base = static_cast<float*>(ctx->d_rays);
    const float* src[6] = {ox, oy, oz, dx, dy, dz};
    for (int k = 0; k < 6; k++) CU_TRY(cudaMemcpyAsync(base + k * nn, src[k], nn * sizeof(float), cudaMemcpyHostToDevice, 0));
    int* d_idx = reinterpret_cast<int*>(base + 6 * nn);
    float* d_t = base + 7 * nn;
    uint8_t* d_en = nullptr;
    if (enabled) {
        d_en = reinterpret_cast<uint8_t*>(base + 8 * nn);
        CU_TRY(cudaMemcpyAsync(d_en, enabled, nn, cudaMemcpyHostToDevice, 0));
    }
    rc = launch_intersect(ctx, base, base + nn, base + 2 * nn, base + 3 * nn, base + 4 * nn, base + 5 * nn, d_en, n, d_idx, d_t, 0);
    if (rc) return rc;
    CU_TRY(cudaMemcpyAsync(hit_index, d_idx, nn * sizeof(int), cudaMemcpyDeviceToHost, 0));
    CU_TRY(cudaMemcpyAsync(hit_t, d_t, nn * sizeof(float), cudaMemcpyDeviceToHost, 0));
    CU_TRY(cudaStreamSynchronize(0));
    return RTB200_OK;
}

int rtb200_intersect_batch_f64(rtb200_ctx* ctx, const double* ox, const double* oy, const double* oz, const double* dx, const double* dy,
                               const double* dz, const uint8_t* enabled, int64_t n, int32_t* hit_index, double* hit_t) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "ctx is null");
    if (!ctx->d_spheres) return fail(RTB200_ERR_STATE, "no scene uploaded (call rtb200_upload_scene first)");
    if (n < 0) return fail(RTB200_ERR_INVALID, "negative ray count");
    if (n == 0) return RTB200_OK;
    if (!ox || !oy || !oz || !dx || !dy || !dz || !hit_index || !hit_t) return fail(RTB200_ERR_INVALID, "null ray buffer");
    CU_TRY(cudaSetDevice(ctx->device));
    // device scratch: 6 double planes, t plane (double), idx plane, enabled bytes
    const size_t nn = (size_t)n;
    int rc = ensure(&ctx->d_rays, &ctx->rays_bytes, nn * (7 * sizeof(double) + sizeof(int) + 1));
    if (rc) return rc;
    double* base = static_cast<double*>(ctx->d_rays);
    const double* src[6] = {ox, oy, oz, dx, dy, dz};
    for (int k = 0; k < 6; k++) CU_TRY(cudaMemcpyAsync(base + k * nn, src[k], nn * sizeof(double), cudaMemcpyHostToDevice, 0));
    double* d_t = base + 6 * nn;
    int* d_idx = reinterpret_cast<int*>(base + 7 * nn);
    uint8_t* d_en = nullptr;
    if (enabled) {
        d_en = reinterpret_cast<uint8_t*>(d_idx + nn);
        CU_TRY(cudaMemcpyAsync(d_en, enabled, nn, cudaMemcpyHostToDevice, 0));
    }
    const long long blocks = ((long long)n + 255) / 256;
    if (blocks > 0x7fffffffLL) return fail(RTB200_ERR_INVALID, "too many rays in one batch");
    rtb::k_intersect_batch_f64<<<(unsigned)blocks, 256, 0, 0>>>(ctx->d_exact, ctx->n_spheres, base, base + nn, base + 2 * nn, base + 3 * nn,
                                                               base + 4 * nn, base + 5 * nn, d_en, n, d_idx, d_t,
                                                               ctx->hit_mode == RTB200_HIT_BOTH_ROOTS ? 1 : 0);
    CU_TRY(cudaGetLastError());
    ctx->launches++;
    CU_TRY(cudaMemcpyAsync(hit_index, d_idx, nn * sizeof(int), cudaMemcpyDeviceToHost, 0));
    CU_TRY(cudaMemcpyAsync(hit_t, d_t, nn * sizeof(double), cudaMemcpyDeviceToHost, 0));
    CU_TRY(cudaStreamSynchronize(0));
    return RTB200_OK;
}

int rtb200_render(rtb200_ctx* ctx, const rtb200_camera* camera, const rtb200_params* params, uint8_t* rgb8_out, float* linear_rgba_out,
                  rtb200_stats* stats) {
    int rc = validate_params(ctx, camera, params);
    if (rc) return rc;
    auto t0 = std::chrono::steady_clock::now();
    CU_TRY(cudaSetDevice(ctx->device));
    const size_t n_pix = (size_t)camera->image_width * camera->image_height;
    {
        void* p = ctx->d_accum; size_t have = ctx->accum_pixels * sizeof(float4);
        rc = ensure(&p, &have, n_pix * sizeof(float4));
        ctx->d_accum = static_cast<float4*>(p); ctx->accum_pixels = have / sizeof(float4);
        if (rc) return rc;
        p = ctx->d_rgb8; have = ctx->rgb8_bytes;
        rc = ensure(&p, &have, n_pix * 3 + 16);
        ctx->d_rgb8 = static_cast<uint8_t*>(p); ctx->rgb8_bytes = have;
        if (rc) return rc;
    }
    cudaStream_t st = 0;
    const bool full_rows = params->row_begin == 0 && params->row_end == camera->image_height && params->row_stride <= 1;
    if (!full_rows) CU_TRY(cudaMemsetAsync(ctx->d_accum, 0, n_pix * sizeof(float4), st));
    CU_TRY(cudaEventRecord(ctx->ev[0], st));
    rc = launch_trace(ctx, camera, params, ctx->d_accum, st);
    if (rc) return rc;
    CU_TRY(cudaEventRecord(ctx->ev[1], st));
    if (ctx->progress_cb && ctx->h_progress) {
        // The reference redraws a progress grid while its workers run (src/renderer.rs:300-365).  The trace kernel
        // publishes its work-queue head in mapped host memory (an async device->host copy of the counter does not
        // complete while the persistent kernel runs, measured), so the poll is a plain host read.
        const unsigned long long total = ctx->items_total;
        unsigned long long seen = 0;
        auto last = std::chrono::steady_clock::now();
        for (;;) {
            cudaError_t q = cudaEventQuery(ctx->ev[1]);
            if (q == cudaSuccess) break;
            if (q != cudaErrorNotReady) { CU_TRY(q); }
            std::this_thread::sleep_for(std::chrono::microseconds(500));   // the end of the kernel is noticed within 0.5 ms
            auto now = std::chrono::steady_clock::now();
            if (std::chrono::duration<double, std::milli>(now - last).count() < (double)ctx->progress_interval_ms) continue;
            last = now;
            unsigned long long claimed = *reinterpret_cast<volatile unsigned long long*>(ctx->h_progress);   // items handed out so far
            if (claimed > total) claimed = total;                // the queue head runs past the end
            if (claimed < seen) claimed = seen;                  // warps publish out of order
            seen = claimed;
            ctx->progress_cb(ctx->progress_user, claimed, total, std::chrono::duration<double>(now - t0).count());
        }
        ctx->progress_cb(ctx->progress_user, total, total, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    }
    rc = launch_resolve(ctx, ctx->d_accum, (long long)n_pix, params->samples_per_pixel, ctx->d_rgb8, st);
    if (rc) return rc;
    CU_TRY(cudaEventRecord(ctx->ev[2], st));
    if (rgb8_out) CU_TRY(cudaMemcpyAsync(rgb8_out, ctx->d_rgb8, n_pix * 3, cudaMemcpyDeviceToHost, st));
    if (linear_rgba_out) CU_TRY(cudaMemcpyAsync(linear_rgba_out, ctx->d_accum, n_pix * sizeof(float4), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaEventRecord(ctx->ev[3], st));
    CU_TRY(cudaStreamSynchronize(st));
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        unsigned long long cnt[3] = {0, 0, 0};
        CU_TRY(cudaMemcpy(cnt, ctx->d_counters, sizeof(cnt), cudaMemcpyDeviceToHost));
        float ms = 0.f;
        CU_TRY(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1])); stats->trace_ms = ms;
        CU_TRY(cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2])); stats->resolve_ms = ms;
        CU_TRY(cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3])); stats->d2h_ms = ms;
        stats->paths = cnt[0]; stats->segments = cnt[1];
        stats->sphere_tests = cnt[1] * (unsigned long long)ctx->n_spheres;
        {   // RenderStat::new, src/renderer.rs:385: the pixels this call traced (every row_stride-th row of the range)
            const int rs = params->row_stride > 1 ? params->row_stride : 1;
            stats->pixels_rendered = (uint64_t)((params->row_end - params->row_begin + rs - 1) / rs) * camera->image_width;
        }
        stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        stats->pixels_per_second = stats->seconds > 0 ? (double)stats->pixels_rendered / stats->seconds : 0.0;
    }
    return RTB200_OK;
}

int rtb200_device_alloc(rtb200_ctx* ctx, size_t bytes, void** dev_out) {
    if (!ctx || !dev_out) return fail(RTB200_ERR_INVALID, "null argument");
    *dev_out = nullptr;
    CU_TRY(cudaSetDevice(ctx->device));
    CU_TRY(cudaMalloc(dev_out, bytes > 0 ? bytes : 1));
    CU_TRY(cudaMemset(*dev_out, 0, bytes > 0 ? bytes : 1));
    ctx->exchange_allocs.push_back(*dev_out);
    return RTB200_OK;
}

int rtb200_device_free(rtb200_ctx* ctx, void* dev) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "ctx is null");
    if (!dev) return RTB200_OK;
    auto it = std::find(ctx->exchange_allocs.begin(), ctx->exchange_allocs.end(), dev);
    if (it == ctx->exchange_allocs.end()) return fail(RTB200_ERR_INVALID, "rtb200_device_free: not a buffer of this context's rtb200_device_alloc");
    CU_TRY(cudaSetDevice(ctx->device));
    CU_TRY(cudaFree(dev));
    ctx->exchange_allocs.erase(it);
    return RTB200_OK;
}

int rtb200_ipc_export(rtb200_ctx* ctx, void* dev, uint8_t handle_out[RTB200_IPC_HANDLE_BYTES]) {
    if (!ctx || !dev || !handle_out) return fail(RTB200_ERR_INVALID, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == RTB200_IPC_HANDLE_BYTES, "IPC handle size");
    CU_TRY(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    CU_TRY(cudaIpcGetMemHandle(&h, dev));
    std::memcpy(handle_out, &h, sizeof h);
    return RTB200_OK;
}

int rtb200_ipc_open(rtb200_ctx* ctx, const uint8_t handle[RTB200_IPC_HANDLE_BYTES], void** dev_out) {
    if (!ctx || !handle || !dev_out) return fail(RTB200_ERR_INVALID, "null argument");
    *dev_out = nullptr;
    CU_TRY(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof h);
    CU_TRY(cudaIpcOpenMemHandle(dev_out, h, cudaIpcMemLazyEnablePeerAccess));
    ctx->ipc_mappings.push_back(*dev_out);
    return RTB200_OK;
}

int rtb200_ipc_close(rtb200_ctx* ctx, void* dev) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "ctx is null");
    auto it = std::find(ctx->ipc_mappings.begin(), ctx->ipc_mappings.end(), dev);
    if (it == ctx->ipc_mappings.end()) return fail(RTB200_ERR_INVALID, "rtb200_ipc_close: not a mapping of this context's rtb200_ipc_open");
    CU_TRY(cudaSetDevice(ctx->device));
    CU_TRY(cudaIpcCloseMemHandle(dev));
    ctx->ipc_mappings.erase(it);
    return RTB200_OK;
}

int rtb200_resolve_grid_sum_device(rtb200_ctx* ctx, const void* const* accum_ptrs, int32_t row_groups, int32_t sample_groups, int32_t image_width,
                                   int64_t pixel_begin, int64_t pixel_end, int32_t samples_per_pixel, void* rgb8_dev, void* sum_dev, void* stream) {
    if (!ctx || !accum_ptrs || !rgb8_dev) return fail(RTB200_ERR_INVALID, "null argument");
    if (row_groups < 1 || sample_groups < 1 || (int64_t)row_groups * sample_groups > RTB200_MAX_PEERS)
        return fail(RTB200_ERR_INVALID, "row_groups x sample_groups must be 1.." + std::to_string(RTB200_MAX_PEERS));
    if (image_width < 1) return fail(RTB200_ERR_INVALID, "image_width must be positive");
    if (pixel_begin < 0 || pixel_end < pixel_begin || samples_per_pixel <= 0) return fail(RTB200_ERR_INVALID, "bad pixel range or sample count");
    const int n_all = row_groups * sample_groups;
    rtb::PeerPtrs pp;
    for (int k = 0; k < RTB200_MAX_PEERS; k++) pp.p[k] = nullptr;
    for (int k = 0; k < n_all; k++) {
        if (!accum_ptrs[k]) return fail(RTB200_ERR_INVALID, "null peer pointer");
        pp.p[k] = static_cast<const float4*>(accum_ptrs[k]);
    }
    if (pixel_end == pixel_begin) return RTB200_OK;
    CU_TRY(cudaSetDevice(ctx->device));
    const long long quads = (pixel_end - pixel_begin + 3) / 4;
    const int blocks = (int)((quads + 255) / 256);
    rtb::k_resolve_peers<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(pp, sample_groups, row_groups, image_width, pixel_begin, pixel_end,
                                                                               (float)samples_per_pixel, static_cast<uint8_t*>(rgb8_dev),
                                                                               static_cast<float4*>(sum_dev));
    CU_TRY(cudaGetLastError());
    ctx->launches++;
    return RTB200_OK;
}

int rtb200_resolve_grid_device(rtb200_ctx* ctx, const void* const* accum_ptrs, int32_t row_groups, int32_t sample_groups, int32_t image_width,
                               int64_t pixel_begin, int64_t pixel_end, int32_t samples_per_pixel, void* rgb8_dev, void* stream) {
    return rtb200_resolve_grid_sum_device(ctx, accum_ptrs, row_groups, sample_groups, image_width, pixel_begin, pixel_end, samples_per_pixel, rgb8_dev,
                                          nullptr, stream);
}

int rtb200_resolve_peers_device(rtb200_ctx* ctx, const void* const* accum_ptrs, int32_t n_peers, int64_t pixel_begin, int64_t pixel_end,
                                int32_t samples_per_pixel, void* rgb8_dev, void* stream) {
    if (n_peers < 1 || n_peers > RTB200_MAX_PEERS) return fail(RTB200_ERR_INVALID, "n_peers must be 1.." + std::to_string(RTB200_MAX_PEERS));
    return rtb200_resolve_grid_device(ctx, accum_ptrs, 1, n_peers, 1, pixel_begin, pixel_end, samples_per_pixel, rgb8_dev, stream);
}

int rtb200_peer_barrier_device(rtb200_ctx* ctx, void* const* flag_ptrs, int32_t n_peers, int32_t my_rank, uint32_t epoch, double timeout_s, void* stream) {
    if (!ctx || !flag_ptrs) return fail(RTB200_ERR_INVALID, "null argument");
    if (n_peers < 1 || n_peers > RTB200_MAX_PEERS || my_rank < 0 || my_rank >= n_peers) return fail(RTB200_ERR_INVALID, "bad peer count or rank");
    rtb::BarrierPtrs bp;
    for (int k = 0; k < RTB200_MAX_PEERS; k++) bp.p[k] = nullptr;
    for (int k = 0; k < n_peers; k++) {
        if (!flag_ptrs[k]) return fail(RTB200_ERR_INVALID, "null flag pointer");
        bp.p[k] = static_cast<unsigned int*>(flag_ptrs[k]);
    }
    if (!(timeout_s > 0.0)) timeout_s = 20.0;
    CU_TRY(cudaSetDevice(ctx->device));
    rtb::k_peer_barrier<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(bp, n_peers, my_rank, epoch, (unsigned long long)(timeout_s * 1e9));
    CU_TRY(cudaGetLastError());
    ctx->launches++;
    return RTB200_OK;
}

int rtb200_read_counters(rtb200_ctx* ctx, uint64_t* paths, uint64_t* segments) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "ctx is null");
    CU_TRY(cudaSetDevice(ctx->device));
    unsigned long long cnt[2] = {0, 0};
    CU_TRY(cudaMemcpy(cnt, ctx->d_counters, sizeof(cnt), cudaMemcpyDeviceToHost));
    if (paths) *paths = cnt[0];
    if (segments) *segments = cnt[1];
    return RTB200_OK;
}

int64_t rtb200_launch_count(rtb200_ctx* ctx) { return ctx ? ctx->launches : 0; }

int rtb200_trace_kernel_stats(rtb200_ctx* ctx, int32_t n_last, double* avg_ms, int32_t* n_used, const char** kernel_name, int32_t* sample_splits) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "ctx is null");
    if (kernel_name) *kernel_name = ctx->last_kernel;
    if (sample_splits) *sample_splits = ctx->last_splits;
    int64_t n = ctx->trace_launches < rtb200_ctx::TRACE_RING ? ctx->trace_launches : rtb200_ctx::TRACE_RING;
    if (n_last > 0 && n_last < n) n = n_last;
    double sum = 0.0;
    CU_TRY(cudaSetDevice(ctx->device));
    for (int64_t k = 0; k < n; k++) {
        cudaEvent_t* pr = ctx->ev_trace[(ctx->trace_launches - 1 - k) % rtb200_ctx::TRACE_RING];
        CU_TRY(cudaEventSynchronize(pr[1]));
        float ms = 0.f;
        CU_TRY(cudaEventElapsedTime(&ms, pr[0], pr[1]));
        sum += ms;
    }
    if (avg_ms) *avg_ms = n > 0 ? sum / (double)n : 0.0;
    if (n_used) *n_used = (int32_t)n;
    return RTB200_OK;
}

int rtb200_set_progress(rtb200_ctx* ctx, rtb200_progress_fn fn, void* user, int32_t interval_ms) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "null context");
    if (interval_ms < 1) interval_ms = 1;
    if (fn) {   // created here, not inside rtb200_render: creating a stream or pinned memory waits for running kernels
        CU_TRY(cudaSetDevice(ctx->device));
        if (!ctx->h_progress) {
            CU_TRY(cudaHostAlloc(reinterpret_cast<void**>(&ctx->h_progress), sizeof(unsigned long long), cudaHostAllocMapped));
            *ctx->h_progress = 0ull;
            CU_TRY(cudaHostGetDevicePointer(reinterpret_cast<void**>(&ctx->d_progress), ctx->h_progress, 0));
        }
    }
    ctx->progress_cb = fn;
    ctx->progress_user = user;
    ctx->progress_interval_ms = interval_ms;
    return RTB200_OK;
}

int rtb200_set_hit_mode(rtb200_ctx* ctx, int32_t mode) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "null context");
    if (mode != RTB200_HIT_NEAR_ROOT && mode != RTB200_HIT_BOTH_ROOTS) return fail(RTB200_ERR_INVALID, "unknown hit mode " + std::to_string(mode));
    ctx->hit_mode = mode;
    return RTB200_OK;
}

int rtb200_filter_check(rtb200_ctx* ctx, uint64_t n_rays, int32_t per_ray, uint32_t seed, const double camera_point[3], uint64_t out[9],
                        double* min_slack) {
    if (!ctx || !out) return fail(RTB200_ERR_INVALID, "null argument");
    if (!ctx->d_spheres) return fail(RTB200_ERR_STATE, "no scene uploaded (call rtb200_upload_scene first)");
    if (ctx->n_spheres < 1) return fail(RTB200_ERR_INVALID, "the scene has no spheres");
    if (per_ray < 1 || per_ray > 64) return fail(RTB200_ERR_INVALID, "per_ray must be 1..64");
    CU_TRY(cudaSetDevice(ctx->device));
    unsigned long long* d_out = nullptr;
    CU_TRY(cudaMalloc(&d_out, 9 * sizeof(unsigned long long)));
    unsigned long long init[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0x7f7fffffull};
    cudaError_t e = cudaMemcpy(d_out, init, sizeof init, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        rtb::FilterCheckArgs A;
        A.spheres = ctx->d_spheres; A.filter = ctx->d_filter; A.slab = reinterpret_cast<const float*>(ctx->d_slab); A.exact = ctx->d_exact;
        A.n_spheres = ctx->n_spheres; A.rp_min = ctx->slab_rp_min; A.largest = ctx->largest_sphere;
        for (int k = 0; k < 3; k++) A.cam[k] = camera_point ? (float)camera_point[k] : (k == 0 ? 13.0f : k == 1 ? 2.0f : 3.0f);
        A.n_rays = n_rays; A.per_ray = per_ray; A.seed = seed; A.out = d_out;
        rtb::k_filter_check<<<ctx->prop.multiProcessorCount * 8, 256>>>(A);
        e = cudaGetLastError();
        ctx->launches++;
    }
    unsigned long long host[9] = {0};
    if (e == cudaSuccess) e = cudaMemcpy(host, d_out, sizeof host, cudaMemcpyDeviceToHost);
    cudaFree(d_out);
    if (e != cudaSuccess) return fail(RTB200_ERR_CUDA, std::string("rtb200_filter_check: ") + cudaGetErrorString(e));
    for (int k = 0; k < 9; k++) out[k] = host[k];
    if (min_slack) { uint32_t bits = (uint32_t)host[8]; float f; std::memcpy(&f, &bits, 4); *min_slack = (double)f; }
    return RTB200_OK;
}

int rtb200_fp32_peak(rtb200_ctx* ctx, int32_t iters, double* tflops, double* ms_out) {
    if (!ctx) return fail(RTB200_ERR_INVALID, "ctx is null");
    if (iters <= 0) iters = 2000;
    CU_TRY(cudaSetDevice(ctx->device));
    const int grid = ctx->prop.multiProcessorCount * 2;
    rtb::k_fp32_peak<<<grid, 1024>>>(ctx->d_probe, 64, 1.0f);   // warm-up
    CU_TRY(cudaGetLastError());
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        CU_TRY(cudaEventRecord(ctx->ev[0], 0));
        rtb::k_fp32_peak<<<grid, 1024>>>(ctx->d_probe, iters, 1.0f);
        CU_TRY(cudaEventRecord(ctx->ev[1], 0));
        CU_TRY(cudaEventSynchronize(ctx->ev[1]));
        float ms = 0.f;
        CU_TRY(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]));
        if (ms < best) best = ms;
    }
    ctx->launches += 4;
    double flop = 2.0 * 8.0 * 16.0 * (double)iters * 1024.0 * (double)grid;
    if (tflops) *tflops = flop / ((double)best * 1e-3) * 1e-12;
    if (ms_out) *ms_out = best;
    return RTB200_OK;
}

}  // extern "C"


// ---- one call, N GPUs (single process): rtb200_multi -------------------------------------------------------------------------
// Renderer::render is one synchronous call that owns its parallelism (src/renderer.rs:38-40, :272-296, :367-369).  Here one host
// thread queues every shard's trace kernel on its own device and stream; cross-device events order the fused reduce+resolve
// kernels (k_resolve_peers, one pixel slice per shard, peer loads over NVLink) behind ALL traces; the root's stream waits for all
// slices and copies the image out.  No host synchronisation between the first launch and the final copy.
struct rtb200_multi {
    int n = 0;
    std::vector<rtb200_ctx*> ctx;
    std::vector<int> device;
    std::vector<cudaStream_t> stream;
    std::vector<cudaEvent_t> ev_start, ev_traced, ev_res0, ev_resolved;
    std::vector<float4*> accum;                        // per shard: W*H float4 on its device
    std::vector<float4*> staging;                      // no-peer-access fallback: the shards' images copied to the root device
    size_t pixels = 0, staging_pixels = 0;
    uint8_t* root_rgb8 = nullptr; size_t rgb8_bytes = 0;
    float4* root_sum = nullptr; size_t sum_pixels = 0;
    bool peer_ok = true;
    rtb200_progress_fn progress_cb = nullptr; void* progress_user = nullptr; int progress_interval_ms = 100;
};

namespace {

void multi_noop_progress(void*, uint64_t, uint64_t, double) {}

// The multi-device calls switch the calling thread's current CUDA device as they go; the caller gets its own device back.
struct DeviceRestore {
    int dev = -1;
    DeviceRestore() { if (cudaGetDevice(&dev) != cudaSuccess) { dev = -1; cudaGetLastError(); } }
    ~DeviceRestore() { if (dev >= 0) cudaSetDevice(dev); }
};

int multi_ensure_buffers(rtb200_multi* m, size_t n_pix, bool want_sum) {
    if (m->pixels < n_pix) {
        for (int g = 0; g < m->n; g++) {
            CU_TRY(cudaSetDevice(m->device[g]));
            if (m->accum[g]) cudaFree(m->accum[g]);
            m->accum[g] = nullptr;
        }
        m->pixels = 0;
        for (int g = 0; g < m->n; g++) {
            CU_TRY(cudaSetDevice(m->device[g]));
            CU_TRY(cudaMalloc(&m->accum[g], n_pix * sizeof(float4)));
            CU_TRY(cudaMemset(m->accum[g], 0, n_pix * sizeof(float4)));
        }
        m->pixels = n_pix;
    }
    CU_TRY(cudaSetDevice(m->device[0]));
    if (m->rgb8_bytes < n_pix * 3 + 16) {
        if (m->root_rgb8) cudaFree(m->root_rgb8);
        m->root_rgb8 = nullptr; m->rgb8_bytes = 0;
        CU_TRY(cudaMalloc(&m->root_rgb8, n_pix * 3 + 16));
        m->rgb8_bytes = n_pix * 3 + 16;
    }
    if (want_sum && m->sum_pixels < n_pix) {
        if (m->root_sum) cudaFree(m->root_sum);
        m->root_sum = nullptr; m->sum_pixels = 0;
        CU_TRY(cudaMalloc(&m->root_sum, n_pix * sizeof(float4)));
        m->sum_pixels = n_pix;
    }
    if (!m->peer_ok && m->staging_pixels < n_pix) {
        for (int g = 1; g < m->n; g++) {
            if (m->staging[g]) cudaFree(m->staging[g]);
            m->staging[g] = nullptr;
            CU_TRY(cudaMalloc(&m->staging[g], n_pix * sizeof(float4)));
        }
        m->staging_pixels = n_pix;
    }
    return RTB200_OK;
}

}  // namespace

extern "C" {

int rtb200_multi_create(const int32_t* devices, int32_t n_devices, rtb200_multi** out) {
    if (!out) return fail(RTB200_ERR_INVALID, "out is null");
    DeviceRestore restore_device;
    *out = nullptr;
    if (n_devices < 1 || n_devices > RTB200_MAX_PEERS) return fail(RTB200_ERR_INVALID, "n_devices must be 1.." + std::to_string(RTB200_MAX_PEERS));
    rtb200_multi* m = new rtb200_multi();
    m->n = n_devices;
    m->ctx.assign(n_devices, nullptr); m->device.assign(n_devices, 0); m->stream.assign(n_devices, nullptr);
    m->ev_start.assign(n_devices, nullptr); m->ev_traced.assign(n_devices, nullptr); m->ev_res0.assign(n_devices, nullptr);
    m->ev_resolved.assign(n_devices, nullptr); m->accum.assign(n_devices, nullptr); m->staging.assign(n_devices, nullptr);
    for (int g = 0; g < n_devices; g++) {
        m->device[g] = devices ? devices[g] : g;
        int rc = rtb200_create(m->device[g], &m->ctx[g]);
        if (rc) { std::string keep = g_last_error; rtb200_multi_destroy(m); return fail(rc, "shard " + std::to_string(g) + ": " + keep); }
        cudaError_t e = cudaStreamCreateWithFlags(&m->stream[g], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreate(&m->ev_start[g]);
        if (e == cudaSuccess) e = cudaEventCreate(&m->ev_traced[g]);
        if (e == cudaSuccess) e = cudaEventCreate(&m->ev_res0[g]);
        if (e == cudaSuccess) e = cudaEventCreate(&m->ev_resolved[g]);
        if (e != cudaSuccess) { rtb200_multi_destroy(m); return fail(RTB200_ERR_CUDA, std::string("rtb200_multi_create: ") + cudaGetErrorString(e)); }
    }
    // direct peer access between every pair of distinct devices (NVLink / NVSwitch on a B200 box)
    for (int a = 0; a < n_devices && m->peer_ok; a++) {
        for (int b = 0; b < n_devices; b++) {
            if (m->device[a] == m->device[b]) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, m->device[a], m->device[b]) != cudaSuccess || !can) { m->peer_ok = false; break; }
            cudaSetDevice(m->device[a]);
            cudaError_t e = cudaDeviceEnablePeerAccess(m->device[b], 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
            if (e != cudaSuccess) { cudaGetLastError(); m->peer_ok = false; break; }
        }
    }
    *out = m;
    return RTB200_OK;
}

void rtb200_multi_destroy(rtb200_multi* m) {
    if (!m) return;
    DeviceRestore restore_device;
    for (int g = 0; g < m->n; g++) {
        if (!m->ctx[g]) continue;                      // a shard rtb200_multi_create never got to
        cudaSetDevice(m->device[g]);
        if (m->stream[g]) { cudaStreamSynchronize(m->stream[g]); cudaStreamDestroy(m->stream[g]); }
        if (m->ev_start[g]) cudaEventDestroy(m->ev_start[g]);
        if (m->ev_traced[g]) cudaEventDestroy(m->ev_traced[g]);
        if (m->ev_res0[g]) cudaEventDestroy(m->ev_res0[g]);
        if (m->ev_resolved[g]) cudaEventDestroy(m->ev_resolved[g]);
        cudaFree(m->accum[g]);
    }
    if (m->n > 0 && m->ctx[0]) {
        cudaSetDevice(m->device[0]);
        cudaFree(m->root_rgb8); cudaFree(m->root_sum);
        for (int g = 0; g < m->n; g++) cudaFree(m->staging[g]);
    }
    for (int g = 0; g < m->n; g++) rtb200_destroy(m->ctx[g]);
    cudaGetLastError();                                // leave no stale error behind for the next call's launch check
    delete m;
}

int rtb200_multi_info(rtb200_multi* m, int32_t* n_shards, int32_t* peer_access) {
    if (!m) return fail(RTB200_ERR_INVALID, "null argument");
    if (n_shards) *n_shards = m->n;
    if (peer_access) *peer_access = m->peer_ok ? 1 : 0;
    return RTB200_OK;
}

int rtb200_multi_upload_scene(rtb200_multi* m, const rtb200_scene* scene) {
    if (!m || !scene) return fail(RTB200_ERR_INVALID, "null argument");
    DeviceRestore restore_device;
    HostScene hs;                                      // flattened once; the same arena goes to every device, copies overlapped
    int rc = build_host_scene(scene, hs);
    if (rc) return rc;
    for (int g = 0; g < m->n; g++) {
        rc = commit_host_scene(m->ctx[g], hs, m->stream[g], false);
        if (rc) return rc;
    }
    for (int g = 0; g < m->n; g++) {
        CU_TRY(cudaSetDevice(m->device[g]));
        CU_TRY(cudaStreamSynchronize(m->stream[g]));
    }
    return RTB200_OK;
}

int rtb200_multi_set_hit_mode(rtb200_multi* m, int32_t mode) {
    if (!m) return fail(RTB200_ERR_INVALID, "null argument");
    for (int g = 0; g < m->n; g++) {
        int rc = rtb200_set_hit_mode(m->ctx[g], mode);
        if (rc) return rc;
    }
    return RTB200_OK;
}

int rtb200_multi_set_progress(rtb200_multi* m, rtb200_progress_fn fn, void* user, int32_t interval_ms) {
    if (!m) return fail(RTB200_ERR_INVALID, "null argument");
    DeviceRestore restore_device;
    if (interval_ms < 1) interval_ms = 1;
    for (int g = 0; g < m->n; g++) {   // the shards' kernels publish their queue heads; the polling is done by rtb200_multi_render
        int rc = rtb200_set_progress(m->ctx[g], fn ? multi_noop_progress : nullptr, nullptr, interval_ms);
        if (rc) return rc;
    }
    m->progress_cb = fn; m->progress_user = user; m->progress_interval_ms = interval_ms;
    return RTB200_OK;
}

int rtb200_multi_render(rtb200_multi* m, const rtb200_camera* camera, const rtb200_params* params, int32_t row_groups, uint8_t* rgb8_out,
                        float* linear_rgba_out, rtb200_stats* stats) {
    if (!m) return fail(RTB200_ERR_INVALID, "null argument");
    DeviceRestore restore_device;
    int rc = validate_params(m->ctx[0], camera, params);
    if (rc) return rc;
    const int N = m->n;
    if (row_groups < 1 || N % row_groups) return fail(RTB200_ERR_INVALID, "row_groups must divide the number of shards (" + std::to_string(N) + ")");
    const int R = row_groups, S = N / R;
    if (params->row_stride > 1) return fail(RTB200_ERR_INVALID, "rtb200_multi_render shards the rows itself: row_stride must be 0 or 1");
    if (R > 1 && params->row_begin != 0) return fail(RTB200_ERR_INVALID, "row_begin must be 0 when row_groups > 1");
    for (int g = 1; g < N; g++)
        if (!m->ctx[g]->d_spheres || m->ctx[g]->n_spheres != m->ctx[0]->n_spheres) return fail(RTB200_ERR_STATE, "scene not uploaded to every shard (rtb200_multi_upload_scene)");
    auto t0 = std::chrono::steady_clock::now();
    const int W = camera->image_width, H = camera->image_height;
    const size_t n_pix = (size_t)W * H;
    rc = multi_ensure_buffers(m, n_pix, linear_rgba_out != nullptr);
    if (rc) return rc;
    const bool full_rows = params->row_begin == 0 && params->row_end == H;
    const int n_samples = params->sample_end - params->sample_begin;
    // ---- every shard's trace, queued back to back by this one thread ----
    for (int g = 0; g < N; g++) {
        const int gr = g / S, gs = g % S;
        rtb200_params p = *params;
        p.sample_begin = params->sample_begin + (int)((long long)n_samples * gs / S);
        p.sample_end = params->sample_begin + (int)((long long)n_samples * (gs + 1) / S);
        p.row_begin = params->row_begin + gr;
        if (p.row_begin > p.row_end) p.row_begin = p.row_end;
        p.row_stride = R;
        CU_TRY(cudaSetDevice(m->device[g]));
        if (!full_rows) CU_TRY(cudaMemsetAsync(m->accum[g], 0, n_pix * sizeof(float4), m->stream[g]));
        CU_TRY(cudaEventRecord(m->ev_start[g], m->stream[g]));
        rc = launch_trace(m->ctx[g], camera, &p, m->accum[g], m->stream[g]);
        if (rc) return rc;
        CU_TRY(cudaEventRecord(m->ev_traced[g], m->stream[g]));
    }
    // ---- progress (optional): poll the shards' published queue heads until the last trace has finished ----
    if (m->progress_cb) {
        unsigned long long total = 0;
        for (int g = 0; g < N; g++) total += m->ctx[g]->items_total;
        unsigned long long seen = 0;
        auto last = std::chrono::steady_clock::now();
        for (;;) {
            bool done = true;
            for (int g = 0; g < N && done; g++) {
                cudaError_t q = cudaEventQuery(m->ev_traced[g]);
                if (q == cudaErrorNotReady) done = false;
                else if (q != cudaSuccess) { CU_TRY(q); }
            }
            if (done) break;
            std::this_thread::sleep_for(std::chrono::microseconds(500));
            auto now = std::chrono::steady_clock::now();
            if (std::chrono::duration<double, std::milli>(now - last).count() < (double)m->progress_interval_ms) continue;
            last = now;
            unsigned long long claimed = 0;
            for (int g = 0; g < N; g++) {
                unsigned long long c = m->ctx[g]->h_progress ? *reinterpret_cast<volatile unsigned long long*>(m->ctx[g]->h_progress) : 0ull;
                claimed += c < m->ctx[g]->items_total ? c : m->ctx[g]->items_total;
            }
            if (claimed < seen) claimed = seen;
            seen = claimed;
            m->progress_cb(m->progress_user, claimed, total, std::chrono::duration<double>(now - t0).count());
        }
        m->progress_cb(m->progress_user, total, total, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    }
    // ---- fused reduce + resolve: shard g owns the pixel slice [b_g, b_{g+1}) (4-pixel aligned), reads all shards' sums ----
    const int spp = params->samples_per_pixel;
    float4* sum_dst = linear_rgba_out ? m->root_sum : nullptr;
    if (m->peer_ok) {
        std::vector<const void*> ptrs((size_t)N);
        for (int k = 0; k < N; k++) ptrs[k] = m->accum[k];
        for (int g = 0; g < N; g++) {
            const long long b0 = g == 0 ? 0 : (long long)(n_pix * (size_t)g / N) & ~3LL;
            const long long b1 = g == N - 1 ? (long long)n_pix : (long long)(n_pix * (size_t)(g + 1) / N) & ~3LL;
            CU_TRY(cudaSetDevice(m->device[g]));
            for (int k = 0; k < N; k++) if (k != g) CU_TRY(cudaStreamWaitEvent(m->stream[g], m->ev_traced[k], 0));
            CU_TRY(cudaEventRecord(m->ev_res0[g], m->stream[g]));
            rc = rtb200_resolve_grid_sum_device(m->ctx[g], ptrs.data(), R, S, W, b0, b1, spp, m->root_rgb8, sum_dst, m->stream[g]);
            if (rc) return rc;
            CU_TRY(cudaEventRecord(m->ev_resolved[g], m->stream[g]));
        }
        CU_TRY(cudaSetDevice(m->device[0]));
        for (int k = 1; k < N; k++) CU_TRY(cudaStreamWaitEvent(m->stream[0], m->ev_resolved[k], 0));
    } else {
        // no peer access: bring the partial images to the root device, resolve everything there
        CU_TRY(cudaSetDevice(m->device[0]));
        std::vector<const void*> ptrs((size_t)N);
        ptrs[0] = m->accum[0];
        for (int k = 1; k < N; k++) {
            CU_TRY(cudaStreamWaitEvent(m->stream[0], m->ev_traced[k], 0));
            CU_TRY(cudaMemcpyPeerAsync(m->staging[k], m->device[0], m->accum[k], m->device[k], n_pix * sizeof(float4), m->stream[0]));
            ptrs[k] = m->staging[k];
        }
        CU_TRY(cudaEventRecord(m->ev_res0[0], m->stream[0]));
        rc = rtb200_resolve_grid_sum_device(m->ctx[0], ptrs.data(), R, S, W, 0, (long long)n_pix, spp, m->root_rgb8, sum_dst, m->stream[0]);
        if (rc) return rc;
        CU_TRY(cudaEventRecord(m->ev_resolved[0], m->stream[0]));
    }
    if (rgb8_out) CU_TRY(cudaMemcpyAsync(rgb8_out, m->root_rgb8, n_pix * 3, cudaMemcpyDeviceToHost, m->stream[0]));
    if (linear_rgba_out) CU_TRY(cudaMemcpyAsync(linear_rgba_out, m->root_sum, n_pix * sizeof(float4), cudaMemcpyDeviceToHost, m->stream[0]));
    CU_TRY(cudaStreamSynchronize(m->stream[0]));
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        for (int g = 0; g < N; g++) {
            CU_TRY(cudaSetDevice(m->device[g]));
            CU_TRY(cudaStreamSynchronize(m->stream[g]));
            unsigned long long cnt[2] = {0, 0};
            CU_TRY(cudaMemcpy(cnt, m->ctx[g]->d_counters, sizeof(cnt), cudaMemcpyDeviceToHost));
            stats->paths += cnt[0]; stats->segments += cnt[1];
            float ms = 0.f;
            CU_TRY(cudaEventElapsedTime(&ms, m->ev_start[g], m->ev_traced[g]));
            if (ms > stats->trace_ms) stats->trace_ms = ms;
            if (m->peer_ok || g == 0) {
                CU_TRY(cudaEventElapsedTime(&ms, m->ev_res0[g], m->ev_resolved[g]));
                if (ms > stats->resolve_ms) stats->resolve_ms = ms;
            }
        }
        stats->sphere_tests = stats->segments * (unsigned long long)m->ctx[0]->n_spheres;
        stats->pixels_rendered = (uint64_t)(params->row_end - params->row_begin) * W;              // RenderStat::new, src/renderer.rs:385
        stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        stats->pixels_per_second = stats->seconds > 0 ? (double)stats->pixels_rendered / stats->seconds : 0.0;
    }
    return RTB200_OK;
}

}  // extern "C"

#ifdef RTB_TIMELINE
// Debug builds only: reset / read the trace kernels' timeline (rt_kernels.cuh, g_timeline).
extern "C" int rtb200_debug_timeline(int reset, unsigned long long* out /* 8 + 4096 */) {
    static unsigned long long init[8 + 4096];
    if (reset) {
        for (int i = 0; i < 8 + 4096; i++) init[i] = 0;
        init[0] = init[2] = init[4] = ~0ull;
        return (int)cudaMemcpyToSymbol(rtb::g_timeline, init, sizeof init);
    }
    return (int)cudaMemcpyFromSymbol(out, rtb::g_timeline, sizeof init);
}
#endif
