/* rtb200.h — C ABI of the B200-native path-tracing backend (librtb200.so).
 *
 * Drop-in boundary for the render hot path of semicolonTransistor/rust-ray-tracing.  Each entry point cites the
 * reference interface it replaces (paths relative to the reference repository root).  Plain pointers and sizes
 * only; the library owns all device memory; nothing aborts across this boundary — every call returns
 * RTB200_OK (0) or a negative error code and rtb200_last_error() holds a message.
 *
 * There is no CPU fallback: every compute entry point needs a CUDA device and fails with
 * RTB200_ERR_CUDA when none is usable.
 */
#ifndef RTB200_H
#define RTB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTB200_OK 0
#define RTB200_ERR_INVALID (-1) /* bad argument (null pointer, non-positive size, unknown material, ...) */
#define RTB200_ERR_CUDA (-2)    /* CUDA runtime error, or no usable device */
#define RTB200_ERR_STATE (-3)   /* call order (e.g. render before upload_scene) */

/* Material kinds — mirror of the three `impl Material` types, src/materials.rs:43-147. */
#define RTB200_LAMBERTIAN 0 /* src/materials.rs:43-69  */
#define RTB200_METAL 1      /* src/materials.rs:73-107 */
#define RTB200_DIELECTRIC 2 /* src/materials.rs:111-155 */

/* Mirror of the reference's material structs (src/materials.rs:45, :75-76, :113-114).
 * `param` is Metal::fuzzy_factor (clamped to <= 1.0 on upload exactly as Metal::new does, :79-88)
 * or Dielectric::index_of_refraction.  `hollow` is Dielectric::hollow. */
typedef struct rtb200_material {
    int32_t kind;
    int32_t hollow;
    double albedo[3];
    double param;
} rtb200_material;

/* Mirror of Sphere { center, radius, material } (src/objects.rs:203-209).  `material` indexes
 * rtb200_scene::materials (the reference holds an Arc<dyn Material>). */
typedef struct rtb200_sphere {
    double center[3];
    double radius;
    int32_t material;
    int32_t _pad;
} rtb200_sphere;

/* Mirror of Scene { objects: Vec<Object> } (src/ray_tracing.rs:100-104).  Object order is significant:
 * the closest-hit tie-break is "last object wins" (src/objects.rs:141). */
typedef struct rtb200_scene {
    const rtb200_sphere* spheres;
    int32_t n_spheres;
    const rtb200_material* materials;
    int32_t n_materials;
} rtb200_scene;

/* Mirror of Camera (src/ray_tracing.rs:15-24): the eight derived quantities, in f64. */
typedef struct rtb200_camera {
    int32_t image_width;
    int32_t image_height;
    double center[3];
    double viewport_upper_left_corner[3];
    double viewport_u[3];
    double viewport_v[3];
    double defocus_u[3];
    double defocus_v[3];
} rtb200_camera;

/* Arguments of Renderer::render (src/renderer.rs:38-40) plus what the reference hard-codes or lacks:
 * the sample stream seed (the reference is OS-seeded, src/ray_tracing.rs:78), and the shard of the
 * job this call renders (the reference shards by 128x128 tile over threads, src/renderer.rs:254-266). */
typedef struct rtb200_params {
    int32_t max_bounces;       /* render(max_bounces, ...)        */
    int32_t samples_per_pixel; /* render(..., samples_per_pixel)  — divisor of the final mean */
    uint64_t seed;             /* Philox key                                                   */
    int32_t sample_begin;      /* this call traces samples [sample_begin, sample_end) of each pixel */
    int32_t sample_end;        /*   (0, samples_per_pixel for the whole job)                   */
    int32_t row_begin;         /* and image rows [row_begin, row_end)  (0, image_height)       */
    int32_t row_end;
    int32_t flags;             /* RTB200_FLAG_* */
    int32_t row_stride;        /* 0 or 1: every row of [row_begin, row_end); R > 1: rows row_begin, row_begin + R, ... —
                                  interleaved scanline shards, the balanced analogue of the reference's tile queue      */
} rtb200_params;

#define RTB200_FLAG_NONE 0
/* Kernel-variant overrides (the parity tests run every variant; 0 lets the library choose from the scene):        */
#define RTB200_FLAG_RAYS1 1 /* force the one-ray-per-lane resident trace kernel (one-stage sweep, spheres in shared memory) */
#define RTB200_FLAG_RAYS2 2 /* force the two-slots-per-lane trace kernels                                                  */
#define RTB200_FLAG_ONE_STAGE 4 /* all pairs through the 8-instruction filter (no slab stage)                               */
#define RTB200_FLAG_TWO_STAGE 8 /* force the slab stage (3-FFMA all-pairs test, then filter, then the exact arithmetic) even
                                   where the scene layout does not favour it; <= 512 spheres from parameter space, beyond
                                   that from shared memory                                                              */

#define RTB200_FLAG_SHARED_STATE 32 /* scenes beyond 512 spheres, two-stage sweep: keep the path state in shared memory (512
                                      threads per SM) instead of registers (640 threads per SM, the default)            */

#define RTB200_FLAG_STREAMED 64 /* force the streamed two-stage kernel: stage-1 data fetched per warp from global memory
                                  (cp.async double buffer) — the kernel every scene beyond ~10 600 spheres gets; there is
                                  no upper limit on the scene size, like the reference's Vec<Object>                    */

#define RTB200_FLAG_BOTH_ROOTS 16 /* this call only: RTB200_HIT_BOTH_ROOTS (see rtb200_set_hit_mode) */

/* Hit modes.  NEAR_ROOT is what the reference's live renderer does: Sphere::hit_packed derives root2's validity from
 * root1 (src/objects.rs:272-276), so only the near root ever counts and a ray that starts inside a sphere misses it
 *  — glass spheres do not refract a second time on the way out.  BOTH_ROOTS applies the rule of the
 * reference's scalar Sphere::hit (src/objects.rs:227-234) inside the same packed arithmetic: when the near root is
 * outside [0.001, inf) the far root is tried.  It is NOT what the reference's binary renders; it is the physically
 * right dielectric.  Parity for it is against the oracle run with the same option. */
#define RTB200_HIT_NEAR_ROOT 0
#define RTB200_HIT_BOTH_ROOTS 1

/* Mirror of RenderStat (src/renderer.rs:11-34) plus the counters the metric needs. */
typedef struct rtb200_stats {
    double seconds;             /* RenderStat::duration — wall time of the call                      */
    uint64_t pixels_rendered;   /* RenderStat::pixels_rendered                                       */
    double pixels_per_second;   /* RenderStat::pixels_per_second                                     */
    uint64_t paths;             /* (pixel, sample) paths traced                                      */
    uint64_t segments;          /* rays cast (one closest-hit sweep each)                            */
    uint64_t sphere_tests;      /* segments * n_spheres (brute force, src/ray_tracing.rs:399-401)    */
    double trace_ms;            /* device time of the trace kernel (CUDA events)                     */
    double resolve_ms;          /* device time of the resolve kernel                                 */
    double h2d_ms, d2h_ms;      /* host<->device copies inside the call                              */
} rtb200_stats;

typedef struct rtb200_ctx rtb200_ctx;
/* Progress callback of rtb200_set_progress / rtb200_multi_set_progress (see there). */
typedef void (*rtb200_progress_fn)(void* user, uint64_t items_claimed, uint64_t items_total, double elapsed_seconds);

/* ---- lifetime ------------------------------------------------------------------------------------------ */

/* TileRenderer::new (src/renderer.rs:232-240): makes a renderer.  Binds CUDA device `device`. */
int rtb200_create(int32_t device, rtb200_ctx** out);
void rtb200_destroy(rtb200_ctx* ctx);
const char* rtb200_last_error(void);
/* Library / device facts: SM count, max SM clock (kHz), compiled arch.  Any pointer may be null. */
int rtb200_device_info(rtb200_ctx* ctx, int32_t* sm_count, int32_t* sm_clock_khz, int32_t* cc_major, int32_t* cc_minor);

/* Page-locked host memory for the image buffers handed to rtb200_render (the RgbImage the reference returns is plain
 * heap memory, src/renderer.rs:376-384; with a pinned buffer the device->host copy of the result is one DMA).
 * Any host pointer is accepted by rtb200_render; pinned ones are faster. */
int rtb200_host_alloc(size_t bytes, void** out);
void rtb200_host_free(void* p);

/* Camera::new (src/ray_tracing.rs:27-62), evaluated in f64 on the host.  Pure host arithmetic (no device). */
int rtb200_camera_init(rtb200_camera* out, int32_t image_width, int32_t image_height, double focal_length,
                       double view_angle_deg, const double center[3], const double look_at[3], const double up[3],
                       double defocus_angle_deg);

/* Scene::from_list (src/ray_tracing.rs:222-224): flattens the scene to the device layout (per sphere: f64 {cx,cy,cz,r*r}
 * for the exact hit test, FP32 filter records for the all-pairs passes of the sweep, a material record) and uploads it. */
int rtb200_upload_scene(rtb200_ctx* ctx, const rtb200_scene* scene);

/* ---- the render path, host buffers (the reference-facing call) ------------------------------------------- */

/* Renderer::render (src/renderer.rs:38-40, impl :244-387): trace + resolve, host buffers in and out.
 * rgb8_out: image_width*image_height*3 bytes, row-major, row 0 = top (RgbImage layout, :376-384); may be null.
 * linear_rgba_out: image_width*image_height*4 floats, the un-quantised per-pixel SUM over the traced samples
 * (r,g,b, sample count); may be null. */
int rtb200_render(rtb200_ctx* ctx, const rtb200_camera* camera, const rtb200_params* params, uint8_t* rgb8_out,
                  float* linear_rgba_out, rtb200_stats* stats);

/* Scene sweep = Object::hit_packed over all objects + PackedHitRecords::update
 * (src/ray_tracing.rs:397-403, src/objects.rs:249-290, :140-155) for a batch of rays, host buffers.
 * rays: SoA float arrays of length n.  enabled: PackedRays::enabled lane mask (src/ray.rs:41), may be null (= all).
 * hit_index[i] = object index of the closest near-root hit or -1; hit_t[i] = its t, +inf on a miss.
 * Precision: the rays are FP32 (the path state of the render kernels); the hit test itself runs at the reference's
 * precision — f64 discriminant, root and closest-hit comparison on the f64 spheres of the scene description, in the
 * reference's operation order — so hit_index equals the f64 routine's for these rays and hit_t is its root rounded once
 * to FP32 (relative error <= 6e-8; BASELINE.json asks for 1e-5). */
int rtb200_intersect_batch(rtb200_ctx* ctx, const float* ox, const float* oy, const float* oz, const float* dx,
                           const float* dy, const float* dz, const uint8_t* enabled, int64_t n, int32_t* hit_index,
                           float* hit_t);

/* The same routine for f64 rays, with NO filter: f64 rays against the f64 spheres of the uploaded scene, every (ray, sphere)
 * pair in list order, in the reference's operation order with its mul_add sites (src/geometry.rs:435,467;
 * src/objects.rs:252-257) — the result equals Sphere::hit_packed + PackedHitRecords::update in f64 bit for bit.  The
 * production sweep (rtb200_intersect_batch, every render kernel) runs this very test on the pairs its FP32 filters let
 * through; this call is the filter-free yardstick for it and the call for a caller whose rays are f64. */
int rtb200_intersect_batch_f64(rtb200_ctx* ctx, const double* ox, const double* oy, const double* oz, const double* dx,
                               const double* dy, const double* dz, const uint8_t* enabled, int64_t n,
                               int32_t* hit_index, double* hit_t);

/* ---- the same kernels on device-resident buffers (for callers that own device memory and streams) -------- */

/* Trace kernel only: accumulates the per-pixel linear SUM of the shard named by `params` into accum_dev
 * (image_width*image_height float4 {r,g,b,n_samples}, device memory).  Rows outside [row_begin,row_end) are
 * left untouched.  `stream` is a cudaStream_t (null = default stream).  Asynchronous. */
int rtb200_trace_device(rtb200_ctx* ctx, const rtb200_camera* camera, const rtb200_params* params, void* accum_dev,
                        void* stream);

/* rtb200_trace_device WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same kernels, same arguments, same result (bit-identical accum_dev) as rtb200_trace_device, but the trace step's scratch —
 * the counters / work-queue head and the partial planes of the sample splits that k_combine sums into accum_dev — lives in
 * ONE file-scope `static __device__` array of RTB200_STATIC_SCRATCH_BYTES (+ a 256-byte header) that every GPU which loads
 * the library holds until the process exits.  The call makes no cudaMalloc / cudaFree / event creation: it clears the
 * counter words (stream-ordered memset), launches the trace kernel and, when the sample range was split, k_combine.
 *  - Limit: a shard whose partial planes (sample splits x traced pixels x 16 B) exceed RTB200_STATIC_SCRATCH_BYTES is
 *    rejected with RTB200_ERR_INVALID and a message naming both sizes; trace fewer rows per call (row_begin / row_end).
 *    The library picks the number of splits K <= 64 from the traced pixels and samples (more splits for fewer pixels), so a
 *    call that traces <= 32 768 pixels always fits; C1 in one call (400x225x16 spp, K = 16) needs 23 MB.
 *  - Every call on a GPU shares the array and it keeps the last call's contents; what must start at zero (the counters) is
 *    cleared by the call itself.  Calls that overlap on the GPU (different streams, or several contexts on one device)
 *    RACE on it: serialise them.  rtb200_read_counters / rtb200_trace_kernel_stats do not see these launches.
 *  - The scene (rtb200_upload_scene) and accum_dev are set up beforehand by the caller; the first call on a device resolves
 *    the array's device address (cudaGetSymbolAddress) and sets the kernels' shared-memory attribute.
 * *scratch_bytes_used (may be null) = bytes of the array this call used. */
#define RTB200_STATIC_SCRATCH_BYTES 33554432 /* 32 MiB */
int rtb200_trace_device_static(rtb200_ctx* ctx, const rtb200_camera* camera, const rtb200_params* params, void* accum_dev,
                               void* stream, uint64_t* scratch_bytes_used);

/* Resolve kernel: sum/spp + Color::to_u8_array (src/renderer.rs:161-163, src/color.rs:54-64).
 * accum_dev: n_pixels float4; rgb8_dev: n_pixels*3 bytes (device).  Asynchronous. */
int rtb200_resolve_device(rtb200_ctx* ctx, const void* accum_dev, int64_t n_pixels, int32_t samples_per_pixel,
                          void* rgb8_dev, void* stream);

/* intersect_batch on device pointers.  Asynchronous. */
int rtb200_intersect_batch_device(rtb200_ctx* ctx, const float* ox, const float* oy, const float* oz, const float* dx,
                                  const float* dy, const float* dz, const uint8_t* enabled, int64_t n,
                                  int32_t* hit_index, float* hit_t, void* stream);

/* ---- multi-GPU exchange over NVLink peer memory (one process per GPU) ---------------------------------------
 * The reference gathers tile results over a channel and assembles the image on the main thread
 * (src/renderer.rs:300-365, :376-384).  Across GPUs the equivalent is ONE kernel per rank that reads every rank's
 * partial accumulation image for its slice of pixels straight from peer memory, sums them in rank order, and writes
 * the quantised RGB8 slice — into the root rank's image, again through peer memory.  No staging copy, no separate
 * reduce pass. */

/* Plain cudaMalloc / cudaFree of an exchange buffer (IPC handles need whole allocations). */
int rtb200_device_alloc(rtb200_ctx* ctx, size_t bytes, void** dev_out);
int rtb200_device_free(rtb200_ctx* ctx, void* dev);

#define RTB200_IPC_HANDLE_BYTES 64
/* cudaIpcGetMemHandle / cudaIpcOpenMemHandle / cudaIpcCloseMemHandle for a buffer from rtb200_device_alloc. */
int rtb200_ipc_export(rtb200_ctx* ctx, void* dev, uint8_t handle_out[RTB200_IPC_HANDLE_BYTES]);
int rtb200_ipc_open(rtb200_ctx* ctx, const uint8_t handle[RTB200_IPC_HANDLE_BYTES], void** dev_out);
int rtb200_ipc_close(rtb200_ctx* ctx, void* dev);

#define RTB200_MAX_PEERS 16
/* Fused reduce + resolve for pixels [pixel_begin, pixel_end): rgb8_dev[3*p..] = quantise(sum_k accum_ptrs[k][p] / spp),
 * the sum taken in array order (deterministic).  accum_ptrs: n_peers device pointers (own buffer and peers' buffers
 * opened with rtb200_ipc_open), each image_width*image_height float4.  rgb8_dev: base of a WHOLE-image RGB8 buffer
 * (local or a peer's).  Asynchronous on `stream`. */
int rtb200_resolve_peers_device(rtb200_ctx* ctx, const void* const* accum_ptrs, int32_t n_peers, int64_t pixel_begin,
                                int64_t pixel_end, int32_t samples_per_pixel, void* rgb8_dev, void* stream);

/* The same for a job sharded over a GRID of row groups x sample groups (tile + sample sharding): image row r was traced by
 * row group r % row_groups (rtb200_params.row_stride = row_groups), each of whose sample_groups members holds a partial
 * sum of that row; accum_ptrs[g * sample_groups + k] is the image of member k of row group g.  row_groups == 1 is
 * rtb200_resolve_peers_device. */
int rtb200_resolve_grid_device(rtb200_ctx* ctx, const void* const* accum_ptrs, int32_t row_groups, int32_t sample_groups,
                               int32_t image_width, int64_t pixel_begin, int64_t pixel_end, int32_t samples_per_pixel,
                               void* rgb8_dev, void* stream);

/* ---- one call, N GPUs (single process) --------------------------------------------------------------------------------
 * Renderer::render is ONE synchronous call that owns all of its parallelism: it spawns its workers, feeds them tiles from
 * a queue and joins them before it returns (src/renderer.rs:38-40, :272-296, :367-369; sole caller src/main.rs:62-64).
 * rtb200_multi is that call over the GPUs of one box: one host thread queues the trace kernel of every shard on its own
 * device and stream, then every device runs the fused reduce+resolve kernel for its slice of the pixels, reading the
 * partial sums of all shards directly from peer memory over NVLink (cudaDeviceEnablePeerAccess; no NCCL, no staging copy)
 * and writing its RGB8 slice into the first device's image, which is copied to the caller's host buffer.  Without peer
 * access between some pair of devices the partial images are copied to the first device instead (cudaMemcpyPeerAsync) and
 * resolved there.
 *
 * Sharding = a grid of row_groups x sample_groups shards (row_groups * sample_groups = number of shards): shard g belongs
 * to row group g / sample_groups and sample group g % sample_groups; it traces the scanlines r with
 * r % row_groups == its row group (interleaved scanlines: the balanced static analogue of the reference's tile queue,
 * src/renderer.rs:254-266) and its contiguous share of the sample range.  row_groups = 1 is pure sample-range sharding
 * (the default: perfectly balanced), row_groups = number of shards is pure row sharding, anything between is tile + sample
 * (BASELINE configs[2]).  The Philox counter carries the image pixel and the global sample index, so the image does not
 * depend on the sharding beyond float summation order. */
typedef struct rtb200_multi rtb200_multi;

/* One shard per entry of `devices` (n_devices in 1..RTB200_MAX_PEERS; devices == NULL: devices 0 .. n_devices-1).  A device
 * may be named more than once (several shards share it: oversubscription, and how the multi-device path is exercised on a
 * one-GPU box).  The first entry is the root: it holds the final image. */
int rtb200_multi_create(const int32_t* devices, int32_t n_devices, rtb200_multi** out);
void rtb200_multi_destroy(rtb200_multi* m);
/* Number of shards; *peer_access (may be null) = 1 when every pair of distinct devices has direct peer access. */
int rtb200_multi_info(rtb200_multi* m, int32_t* n_shards, int32_t* peer_access);
/* rtb200_upload_scene / rtb200_set_hit_mode on every shard. */
int rtb200_multi_upload_scene(rtb200_multi* m, const rtb200_scene* scene);
int rtb200_multi_set_hit_mode(rtb200_multi* m, int32_t mode);
/* Progress over all shards (items claimed / total summed over the devices); see rtb200_set_progress. */
int rtb200_multi_set_progress(rtb200_multi* m, rtb200_progress_fn fn, void* user, int32_t interval_ms);

/* Renderer::render over all shards: the whole job named by `params` (its sample range and rows; row_stride must be 0 or 1
 * and row_begin 0 when row_groups > 1), host buffers out as for rtb200_render.  linear_rgba_out, when not null, receives
 * the per-pixel SUM over all shards (r, g, b, sample count) as the reduce kernel formed it.  stats: counters summed over
 * the shards, trace_ms / resolve_ms = the slowest shard's. */
int rtb200_multi_render(rtb200_multi* m, const rtb200_camera* camera, const rtb200_params* params, int32_t row_groups,
                        uint8_t* rgb8_out, float* linear_rgba_out, rtb200_stats* stats);

/* Stream-ordered barrier between the ranks of a one-process-per-GPU job, through peer memory (no NCCL call, no host round
 * trip): flag_ptrs[k] = rank k's flag array (RTB200_MAX_PEERS uint32 words from rtb200_device_alloc, zero-initialised; the
 * caller's own and the peers' opened with rtb200_ipc_open).  The kernel publishes `epoch` (the caller increases it by one per
 * barrier) into slot my_rank of every array and waits until every slot of its own array has reached it; work queued on
 * `stream` after this call starts only when every rank's stream has reached its barrier of the same epoch, and sees what the
 * peers wrote before theirs.  The wait is bounded by timeout_s (<= 0: 20 s): a missing rank makes the kernel trap — the
 * process fails loudly instead of hanging. */
int rtb200_peer_barrier_device(rtb200_ctx* ctx, void* const* flag_ptrs, int32_t n_peers, int32_t my_rank, uint32_t epoch,
                               double timeout_s, void* stream);

/* rtb200_resolve_grid_device that also stores the summed float4 image for its pixel range into sum_dev (a WHOLE-image float4
 * buffer, local or a peer's; null = not wanted): how a caller of the multi-GPU path gets the linear image before quantisation. */
int rtb200_resolve_grid_sum_device(rtb200_ctx* ctx, const void* const* accum_ptrs, int32_t row_groups, int32_t sample_groups,
                                   int32_t image_width, int64_t pixel_begin, int64_t pixel_end, int32_t samples_per_pixel,
                                   void* rgb8_dev, void* sum_dev, void* stream);

/* Device counters of the last trace on this context (valid after the stream has been synchronised). */
int rtb200_read_counters(rtb200_ctx* ctx, uint64_t* paths, uint64_t* segments);

/* Progress reporting (the reference redraws a per-tile progress grid on stderr while it renders,
 * src/renderer.rs:300-365; here it is a callback, off by default).  While rtb200_render waits for the trace kernel it
 * calls fn(user, items_claimed, items_total, elapsed_seconds) every interval_ms milliseconds from the calling thread,
 * and once more with items_claimed == items_total when the kernel has finished.  A work item is one pixel's share of
 * the sample range (what a path slot claims from the device work queue).  fn == NULL switches it off. */
int rtb200_set_progress(rtb200_ctx* ctx, rtb200_progress_fn fn, void* user, int32_t interval_ms);

/* Selects the hit mode (RTB200_HIT_*) of every later rtb200_render / rtb200_trace_device / rtb200_intersect_batch on
 * this context.  Default RTB200_HIT_NEAR_ROOT. */
int rtb200_set_hit_mode(rtb200_ctx* ctx, int32_t mode);

/* Number of kernel launches this context has issued since creation (trace, resolve, intersect, probe). */
int64_t rtb200_launch_count(rtb200_ctx* ctx);

/* The trace kernel ALONE (without the combine pass that follows it when the sample range is split): average device time of
 * the last n_last launches on this context (<= 0 or more than recorded: all of the last 64), from CUDA events recorded on
 * the launching stream around each launch; *kernel_name = the variant the last launch ran; *sample_splits = K of the last
 * launch: each pixel's sample range was traced as K contiguous sub-ranges whose FP32 partial sums were added in order
 * (what a checker needs to reproduce the summation bit for bit).  Any pointer may be null.  Waits for those launches. */
int rtb200_trace_kernel_stats(rtb200_ctx* ctx, int32_t n_last, double* avg_ms, int32_t* n_used, const char** kernel_name,
                              int32_t* sample_splits);

/* Self-check of the sweep's conservative FP32 filters, on the device, with the arithmetic the render kernels run.  The
 * reference tests every (ray, sphere) pair with Sphere::hit_packed's discriminant (src/objects.rs:252-259); the sweeps run
 * that test (in f64) only on pairs that pass two cheaper FP32 tests, which must therefore never reject a pair with
 * discriminant >= 0.  This call brute-forces n_rays * per_ray adversarial pairs of the uploaded scene (rays aimed at sphere
 * silhouettes with relative perturbations 1e-9 .. 1e-1, from camera_point (null: 13,2,3), from sphere surfaces and from
 * far out on the largest sphere) and counts: out[0] pairs, [1] pairs with discriminant >= 0, [2] candidates of the 8-FLOP
 * filter, [3] its FALSE NEGATIVES (must be 0), [4..6] candidates of the slab test per dropped axis, [7] its FALSE
 * NEGATIVES over the three axes (must be 0), [8] reserved.  *min_slack (may be null) = the smallest filter value over
 * discriminant >= 0 pairs in units of the proven margin (1 = no rounding error observed, 0 = margin used up). */
int rtb200_filter_check(rtb200_ctx* ctx, uint64_t n_rays, int32_t per_ray, uint32_t seed, const double camera_point[3],
                        uint64_t out[9], double* min_slack);

/* K0: measures sustained FP32 FMA throughput (dependent-free FFMA chains on every SM) — the roofline
 * denominator for this path (the driver's MEASURED_PEAKS.json holds only HBM and BF16 figures).
 * Returns TFLOP/s (FMA = 2 FLOP) in *tflops and the kernel's device time in *ms. */
int rtb200_fp32_peak(rtb200_ctx* ctx, int32_t iters, double* tflops, double* ms);

#ifdef __cplusplus
}
#endif
#endif /* RTB200_H */
