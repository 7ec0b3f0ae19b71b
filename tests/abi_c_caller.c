/* abi_c_caller.c — TEST PROGRAM.  A plain-C caller of include/rtb200.h that performs exactly the call sequence of the Rust
 * binding's `render` (integration/rust/src/cuda_renderer.rs): multi_create -> multi_upload_scene -> multi_render ->
 * multi_destroy, on the three-sphere scene of BASELINE configs[0], flattened the way the Rust code flattens Vec<Object> (one
 * material record per object, list order).  Also the single-context sequence create -> upload_scene -> render -> destroy.
 * Built with gcc against the header (no C++, no CUDA headers: the boundary really is plain C) and run under `-m gpu` by
 * tests/test_c_caller.py, which compares the images it writes with tests/golden/render_three_spheres.npz.
 *
 * usage: abi_c_caller <out_prefix> <n_shards> <row_groups>     writes <out_prefix>_single.rgb and <out_prefix>_multi.rgb */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rtb200.h"

#define CHECK(expr)                                                                  \
    do {                                                                             \
        int rc_ = (expr);                                                            \
        if (rc_ != RTB200_OK) {                                                      \
            fprintf(stderr, "%s -> %d: %s\n", #expr, rc_, rtb200_last_error());     \
            return 1;                                                                \
        }                                                                            \
    } while (0)

static int write_file(const char* prefix, const char* suffix, const uint8_t* data, size_t n) {
    char path[1024];
    snprintf(path, sizeof path, "%s_%s.rgb", prefix, suffix);
    FILE* f = fopen(path, "wb");
    if (!f) return 1;
    size_t w = fwrite(data, 1, n, f);
    fclose(f);
    return w == n ? 0 : 1;
}

int main(int argc, char** argv) {
    const char* prefix = argc > 1 ? argv[1] : "abi_c_caller";
    const int n_shards = argc > 2 ? atoi(argv[2]) : 2;
    const int row_groups = argc > 3 ? atoi(argv[3]) : 1;
    /* the scene of tests/golden/three_spheres.toml */
    const rtb200_material materials[4] = {
        {RTB200_LAMBERTIAN, 0, {0.5, 0.5, 0.5}, 0.0},
        {RTB200_LAMBERTIAN, 0, {0.4, 0.2, 0.1}, 0.0},
        {RTB200_METAL, 0, {0.7, 0.6, 0.5}, 0.0},
        {RTB200_DIELECTRIC, 0, {1.0, 1.0, 1.0}, 1.5},
    };
    const rtb200_sphere spheres[4] = {
        {{0.0, -1000.0, 0.0}, 1000.0, 0, 0},
        {{-4.0, 1.0, 0.0}, 1.0, 1, 0},
        {{4.0, 1.0, 0.0}, 1.0, 2, 0},
        {{0.0, 1.0, 0.0}, 1.0, 3, 0},
    };
    const rtb200_scene scene = {spheres, 4, materials, 4};
    /* Camera::new with main's constants (src/main.rs:51-58) at the golden fixture's size */
    const int W = 96, H = 54, spp = 8, depth = 50;
    const double center[3] = {16.0, 2.0, 18.5}, look_at[3] = {0.0, 0.0, 0.0}, up[3] = {0.0, 1.0, 0.0};
    rtb200_camera cam;
    CHECK(rtb200_camera_init(&cam, W, H, 10.0, 30.0, center, look_at, up, 0.0));
    rtb200_params p;
    memset(&p, 0, sizeof p);
    p.max_bounces = depth; p.samples_per_pixel = spp; p.seed = 1;
    p.sample_begin = 0; p.sample_end = spp; p.row_begin = 0; p.row_end = H; p.flags = RTB200_FLAG_NONE; p.row_stride = 1;
    const size_t n = (size_t)W * H * 3;
    uint8_t* rgb = (uint8_t*)malloc(n);
    rtb200_stats st;

    /* single context: TileRenderer::new -> render */
    rtb200_ctx* ctx = NULL;
    CHECK(rtb200_create(0, &ctx));
    CHECK(rtb200_upload_scene(ctx, &scene));
    CHECK(rtb200_render(ctx, &cam, &p, rgb, NULL, &st));
    rtb200_destroy(ctx);
    if (write_file(prefix, "single", rgb, n)) return 2;
    printf("single paths=%llu segments=%llu\n", (unsigned long long)st.paths, (unsigned long long)st.segments);

    /* the Rust render(): every shard behind one call (all on device 0 here, so it runs on a one-GPU box) */
    int32_t devices[RTB200_MAX_PEERS];
    for (int g = 0; g < n_shards && g < RTB200_MAX_PEERS; g++) devices[g] = 0;
    rtb200_multi* m = NULL;
    CHECK(rtb200_multi_create(devices, n_shards, &m));
    CHECK(rtb200_multi_upload_scene(m, &scene));
    memset(rgb, 0, n);
    CHECK(rtb200_multi_render(m, &cam, &p, row_groups, rgb, NULL, &st));
    rtb200_multi_destroy(m);
    if (write_file(prefix, "multi", rgb, n)) return 2;
    printf("multi paths=%llu segments=%llu\n", (unsigned long long)st.paths, (unsigned long long)st.segments);
    free(rgb);
    return 0;
}
