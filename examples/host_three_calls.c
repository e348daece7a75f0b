/* host_three_calls.c — a host that is NOT Python driving librt.so through the C ABI alone (include/rt_api.h).
 *
 * This is the shape of the reference's renderer (src/raytracer/raytracing_renderer.odin) reduced to its three calls on the
 * GPU scene: build (set_scene :98-106 -> gpu_scene_init), trace (render_scene :204-270 -> vkCmdTraceRaysKHR) and consume
 * (the RGBA32F image the blit reads, :271-339) — single-threaded, like the Odin host — with the GPUs listed in rt_config
 * (n_devices / device_ids): the multi-GPU fan-out happens inside the library.
 *
 *   gcc -std=c99 -O2 -Iinclude examples/host_three_calls.c -Lraytracer_b200/csrc -lrt -Wl,-rpath,$PWD/raytracer_b200/csrc -o host
 *   ./host scene.bin out.rgba32f [device ids ...]
 *
 * scene.bin (little endian, written by tests/test_gpu_multi.py): u32 n_meshes, n_objects, n_materials, width, height,
 * frames, max_depth; per mesh u32 n_vertices, n_indices, vertices (24 B each), indices; objects (rt_object, 72 B each);
 * materials (44 B each); camera UBO (64 f32); clear colour (3 f32).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rt_api.h"

#define CHECK(call)                                                                          \
    do {                                                                                     \
        int rc_ = (call);                                                                    \
        if (rc_ != RT_OK) {                                                                  \
            fprintf(stderr, "%s -> %d: %s\n", #call, rc_, rt_last_error(ctx));               \
            return 1;                                                                        \
        }                                                                                    \
    } while (0)

static int read_all(FILE* f, void* dst, size_t n) { return fread(dst, 1, n, f) == n; }

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s scene.bin out.rgba32f [device ids ...]\n", argv[0]); return 2; }
    FILE* f = fopen(argv[1], "rb");
    if (!f) { perror(argv[1]); return 2; }
    uint32_t hdr[7];
    if (!read_all(f, hdr, sizeof(hdr))) return 2;
    const uint32_t n_meshes = hdr[0], n_objects = hdr[1], n_materials = hdr[2], W = hdr[3], H = hdr[4], frames = hdr[5];
    rt_mesh* meshes = (rt_mesh*)calloc(n_meshes ? n_meshes : 1, sizeof(rt_mesh));
    for (uint32_t i = 0; i < n_meshes; i++) {
        uint32_t nv, ni;
        if (!read_all(f, &nv, 4) || !read_all(f, &ni, 4)) return 2;
        rt_vertex* v = (rt_vertex*)malloc((size_t)nv * sizeof(rt_vertex) + 1);
        uint32_t* idx = (uint32_t*)malloc((size_t)ni * 4 + 1);
        if (!read_all(f, v, (size_t)nv * sizeof(rt_vertex)) || !read_all(f, idx, (size_t)ni * 4)) return 2;
        meshes[i].vertices = v; meshes[i].n_vertices = nv; meshes[i].indices = idx; meshes[i].n_indices = ni;
    }
    rt_object* objects = (rt_object*)malloc((size_t)n_objects * sizeof(rt_object) + 1);
    rt_material* materials = (rt_material*)malloc((size_t)n_materials * sizeof(rt_material) + 1);
    float camera[64], clear[3];
    if (!read_all(f, objects, (size_t)n_objects * sizeof(rt_object)) || !read_all(f, materials, (size_t)n_materials * sizeof(rt_material)) ||
        !read_all(f, camera, sizeof(camera)) || !read_all(f, clear, sizeof(clear)))
        return 2;
    fclose(f);

    rt_config cfg;
    rt_default_config(&cfg);
    cfg.max_depth = (int32_t)hdr[6];
    cfg.frames_in_flight = frames;
    const int n_dev = argc - 3;
    if (n_dev >= 2) {                      /* the GPUs go into the configuration; nothing else in this file knows about them */
        cfg.n_devices = (uint32_t)n_dev;
        for (int k = 0; k < n_dev && k < 8; k++) cfg.device_ids[k] = atoi(argv[3 + k]);
        cfg.strip_balance = RT_BALANCE_COST;
    } else if (n_dev == 1) {
        cfg.device = atoi(argv[3]);
    }
    rt_ctx* ctx = NULL;
    if (rt_create(&cfg, &ctx) != RT_OK) { fprintf(stderr, "rt_create: %s\n", rt_last_error(NULL)); return 1; }
    CHECK(rt_build_scene(ctx, meshes, n_meshes, objects, n_objects, materials, n_materials));   /* build   */
    CHECK(rt_resize(ctx, W, H));
    for (uint32_t k = 0; k < frames; k++) CHECK(rt_trace_frame(ctx, camera, clear, k));          /* trace   */
    float* image = (float*)malloc((size_t)W * H * 16);
    CHECK(rt_readback(ctx, image, (size_t)W * H * 16));                                          /* consume */
    rt_counters cnt;
    CHECK(rt_get_counters(ctx, &cnt));
    rt_scene_info info;
    CHECK(rt_scene_info_get(ctx, &info));
    printf("%s: %u devices, %ux%u x %u frames, %u triangles, %llu rays, %llu launches\n", rt_version(), n_dev >= 2 ? (unsigned)n_dev : 1u, W, H,
           frames, info.n_triangles, (unsigned long long)(cnt.extend_rays + cnt.shadow_rays), (unsigned long long)cnt.launches);
    rt_destroy(ctx);
    f = fopen(argv[2], "wb");
    if (!f || fwrite(image, 16, (size_t)W * H, f) != (size_t)W * H) { perror(argv[2]); return 2; }
    fclose(f);
    return 0;
}
