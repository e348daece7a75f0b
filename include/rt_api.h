/* rt_api.h — C ABI of librt.so, the B200-native path-tracing core.
 *
 * The boundary is cut along the three calls the reference's renderer makes on its GPU scene
 * (reference paths are relative to ARTurleite6/RayTracer):
 *   build   gpu_scene_init / gpu_scene_bake            src/raytracer/gpu_scene.odin:45-118
 *   trace   raytracing_renderer_render_scene           src/raytracer/raytracing_renderer.odin:194-270
 *           -> command_buffer_trace_rays               src/raytracer/command_buffer.odin:261-274
 *   consume the RGBA32F output image (blit source)     src/raytracer/raytracing_renderer.odin:271-339
 * plus resize (:108-126), destroy (gpu_scene.odin:49-75) and the incremental update set
 * (gpu_scene.odin:430-601).  All record layouts are the reference's own, so an Odin host can
 * hand over raw_data(slice) pointers (INTEGRATION.md shows the `foreign import` block).
 *
 * Conventions: every entry point returns RT_OK (0) or a negative rt_status; the message of the
 * last failure is kept per context (rt_last_error).  Nothing aborts the host (the reference's
 * vk_check calls os.exit(1): src/raytracer/vulkan_context.odin:385-394).  The caller owns every
 * input array; the library copies during the call and keeps no host pointer.  One rt_ctx is
 * used from one host thread at a time; calls are ordered on the context's CUDA stream.  Every call makes the
 * context's device the calling thread's current CUDA device and leaves it so (multi-GPU contexts: the last device).
 * Different contexts may be driven from different threads; nothing is shared between them.
 * There is no CPU fallback: rt_create fails with RT_ERR_NO_DEVICE when no sm_100 GPU is usable.
 */
#ifndef RT_API_H
#define RT_API_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define RT_API __attribute__((visibility("default")))
#else
#define RT_API
#endif

typedef struct rt_ctx rt_ctx;

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID_ARGUMENT = -1,
    RT_ERR_NO_DEVICE = -2,
    RT_ERR_CUDA = -3,
    RT_ERR_OUT_OF_MEMORY = -4,
    RT_ERR_NO_SCENE = -5,
    RT_ERR_NO_TARGET = -6,
    RT_ERR_BVH_TOO_DEEP = -7,
    RT_ERR_INDEX_RANGE = -8
} rt_status;

/* flags mirror the compile-time switches of shaders/simple.rchit:9-11 */
enum {
    RT_FLAG_DIRECT_LIGHTING = 1u << 0, /* USE_DIRECT_LIGHTING 1 */
    RT_FLAG_MIS = 1u << 1,             /* USE_MIS 1 */
    RT_FLAG_COUNTERS = 1u << 2,        /* maintain device ray/node/triangle counters */
    /* Primary rays are normally traced as 128-ray packets with one shared traversal per 16x8 screen tile (same hit
     * records, fewer node/triangle fetches: `nodes`/`tris` then count one fetch per packet), and visibility rays pool
     * their triangle tests per warp (same result; a ray may test the rest of a node's triangles after its first
     * occluder).  With this flag every primary ray walks the BVH on its own and visibility rays stop at their first
     * occluder, which is the per-ray work definition of SURVEY.md §8(d): counters then equal the CPU oracle's. */
    RT_FLAG_PER_RAY_PRIMARY = 1u << 3,
    RT_FLAG_DEFAULT = (1u << 0) | (1u << 1) | (1u << 2)
};

/* frame pipelines */
enum {
    RT_PIPELINE_WAVEFRONT = 0, /* raygen -> extend -> shade -> shadow queues (default) */
    RT_PIPELINE_MEGAKERNEL = 1 /* one persistent kernel, one path per thread */
};

typedef struct rt_config {
    int32_t device;          /* CUDA ordinal; -1 = current device */
    int32_t max_depth;       /* shaders/simple.rgen:23  MAX_DEPTH = 3 */
    int32_t rr_start_depth;  /* shaders/simple.rgen:56  roulette is a no-op below depth 3 */
    float radiance_clamp;    /* shaders/simple.rgen:121 min(color, 5.0) */
    uint32_t flags;          /* RT_FLAG_* */
    uint32_t pipeline;       /* RT_PIPELINE_* */
    /* screen-tile sharding: this context renders the image rows y with
     * (y / strip_rows) % strip_count == strip_index; seeds and camera rays always use the full
     * launch size, so the union over ranks is bit-identical to strip_count == 1. */
    uint32_t strip_rows;   /* 0 -> 8 */
    uint32_t strip_index;  /* rank */
    uint32_t strip_count;  /* world size; 0 -> 1 */
    /* rt_trace_frames traces up to this many consecutive frames as one wavefront (deep, short bounces of
     * one frame overlap the long first bounces of the others); the image is bit-identical to
     * frame-by-frame dispatch.  0 -> 8; at most 1024, and capped so the path-state arrays stay below 56 GiB. */
    uint32_t frames_in_flight;
    /* Multi-GPU fan-out INSIDE the library: with n_devices >= 2 the context owns one sub-context per
     * listed CUDA ordinal, the scene and BVH are replicated on each (deterministic builder -> identical BVH), the image's
     * strips are dealt to the devices (device k = strip_index k of strip_count n_devices), every rt_* call fans out from
     * the ONE calling host thread, and rt_readback gathers all strips into the caller's image.  The host's three calls
     * stay three calls; `device`, `strip_index` and `strip_count` are ignored then.  n_devices 0 or 1 = `device` alone. */
    uint32_t n_devices;
    int32_t device_ids[8];
    /* RT_BALANCE_*: how strips are assigned to the devices / ranks of a strip partition (strip_count > 1) */
    uint32_t strip_balance;
    uint32_t reserved[4];
} rt_config;

enum {
    RT_BALANCE_ROUND_ROBIN = 0, /* strip s -> rank s % strip_count */
    /* Strips are dealt by measured cost: whenever an accumulation restarts (frame_number 0) one calibration pass counts
     * the rays every strip of the full frame costs, and the strips go to the ranks longest-processing-time first.  Every
     * rank runs the same deterministic pass and so derives the same table without communication; seeds use global pixel
     * indices, so the image stays bit-identical whatever the assignment. */
    RT_BALANCE_COST = 1
};

/* scene.odin:10-13 / ray_common.glsl:48-51 — 24 B */
typedef struct rt_vertex {
    float pos[3];
    float normal[3];
} rt_vertex;

/* scene.odin:56-60 */
typedef struct rt_mesh {
    const rt_vertex* vertices;
    uint32_t n_vertices;
    const uint32_t* indices; /* triangle list, 3 per triangle (acceleration_structure.odin:32-63) */
    uint32_t n_indices;
} rt_mesh;

/* scene.odin:41-54 — the object's model matrix is column-major (m[c*4+r]) */
typedef struct rt_object {
    uint32_t mesh_index;
    uint32_t material_index;
    float model[16];
} rt_object;

/* gpu_scene.odin:17-21 / ray_common.glsl:3-11 — 44 B scalar layout */
typedef struct rt_material {
    float albedo[3];
    float emission_color[3];
    float emission_power;
    float roughness;
    float metallic;
    float transmission;
    float ior;
} rt_material;

/* push constant of the trace (raytracing_renderer.odin:33-36 / simple.rgen:18-21) is passed as
 * (clear_color, frame_number); Camera_UBO (camera.odin:15-20) as 64 floats:
 * proj, view, inverse_view, inverse_proj, each column-major. */

typedef struct rt_counters {
    uint64_t extend_rays;   /* closest-hit rays issued (traceRayEXT, simple.rgen:92-104) */
    uint64_t shadow_rays;   /* visibility rays issued (isVisibleRQ, simple.rchit:350-385) */
    uint64_t nodes;         /* BVH8 nodes fetched (128 B each) */
    uint64_t tris;          /* triangles fetched (48 B each) */
    uint64_t hits;          /* closest-hit shader invocations */
    uint64_t nee;           /* light samples taken */
    uint64_t frames;        /* rt_trace_frame calls since the last reset */
    uint64_t shadow_nodes;  /* the part of `nodes` fetched by visibility rays */
    uint64_t shadow_tris;   /* the part of `tris` fetched by visibility rays */
    uint64_t launches;      /* kernels of this library launched since the last reset */
} rt_counters;

typedef struct rt_scene_info {
    uint32_t n_triangles;   /* instanced, world-space */
    uint32_t n_nodes;       /* BVH8 nodes */
    uint32_t bvh_depth;
    uint32_t n_lights;
    uint32_t n_objects;
    uint32_t n_materials;
    float build_ms;         /* device time of the last BVH build */
    uint32_t refits;        /* box refits (rt_update_object_transform) since that build */
} rt_scene_info;

/* 32-byte reservoir, shaders/restir/restir_structs.glsl:1-12 */
typedef struct rt_reservoir {
    float weight_sum;
    float target_pdf;
    uint32_t M;
    uint32_t light_index;   /* 0xFFFFFFFF = empty (RTXDI_EmptyDIReservoir, :14-23) */
    float uv[2];
    float distance;         /* carries the sampled light triangle index (as bits) in this build */
    float W;
} rt_reservoir;

enum {
    RT_RESTIR_TEMPORAL = 1u << 0,
    RT_RESTIR_SPATIAL = 1u << 1
};

typedef struct rt_restir_params {
    uint32_t flags;              /* RT_RESTIR_* */
    uint32_t initial_candidates; /* light candidates per pixel (default 8) */
    uint32_t spatial_samples;    /* neighbours per pixel (default 4) */
    float spatial_radius;        /* pixels (default 16) */
    uint32_t history_cap;        /* M clamp multiple for temporal reuse (default 20) */
    uint32_t reserved[3];
} rt_restir_params;

/* ---- lifetime ------------------------------------------------------------------------- */
RT_API int rt_create(const rt_config* cfg, rt_ctx** out);
RT_API void rt_destroy(rt_ctx* ctx);
RT_API void rt_default_config(rt_config* cfg);
RT_API const char* rt_last_error(const rt_ctx* ctx); /* ctx may be NULL: error of the last rt_create */
RT_API const char* rt_version(void);

/* ---- build: replaces gpu_scene_init (gpu_scene.odin:45-118) -------------------------------
 * Derives lights (objects whose material has emission_power > 0, :366-379), instance masks and
 * light_index exactly as the reference does, flattens instances to world space and builds the
 * BVH8 on the GPU. */
RT_API int rt_build_scene(rt_ctx* ctx, const rt_mesh* meshes, uint32_t n_meshes, const rt_object* objects,
                   uint32_t n_objects, const rt_material* materials, uint32_t n_materials);
RT_API int rt_scene_info_get(rt_ctx* ctx, rt_scene_info* out);

/* ---- incremental updates: gpu_scene.odin:430-601 (applied before the next trace) ----------- */
RT_API int rt_update_material(rt_ctx* ctx, uint32_t material_index, const rt_material* m);      /* :430-455 */
RT_API int rt_update_object_transform(rt_ctx* ctx, uint32_t object_index, const float model[16]); /* :457-515 */
RT_API int rt_update_object_material(rt_ctx* ctx, uint32_t object_index, uint32_t material_index); /* :517-560 */
RT_API int rt_update_object_mesh(rt_ctx* ctx, uint32_t object_index, uint32_t mesh_index);       /* :562-601 */

/* add / remove (gpu_scene.odin:430-455, dispatched at raytracing_renderer.odin:163-183).  rt_add_object instances a mesh the
 * scene already holds.  rt_remove_material follows scene_delete_material (scene.odin:104-116): the LAST material moves into the
 * freed slot, objects that used the removed one fall back to material 0.  rt_remove_object shifts later objects down by one. */
RT_API int rt_add_material(rt_ctx* ctx, const rt_material* m, uint32_t* index_out);
RT_API int rt_remove_material(rt_ctx* ctx, uint32_t material_index);
RT_API int rt_add_object(rt_ctx* ctx, const rt_object* o, uint32_t* index_out);
RT_API int rt_remove_object(rt_ctx* ctx, uint32_t object_index);

/* ---- target: raytracing_renderer_on_resize (raytracing_renderer.odin:108-126) -------------- */
RT_API int rt_resize(rt_ctx* ctx, uint32_t width, uint32_t height);

/* ---- trace: one sample per pixel, stream-ordered (like a recorded command buffer) ---------
 * frame_number == 0 overwrites the image, otherwise running mean (simple.rgen:127-136). */
RT_API int rt_trace_frame(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t frame_number);
/* n consecutive frames with an unchanged camera == n calls of rt_trace_frame
 * (application loop, application.odin:109-115); lets terminated paths regenerate across frames. */
RT_API int rt_trace_frames(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t first_frame,
                    uint32_t n_frames);
/* spp slices (the alternative to screen strips): traces the frames first_frame + k * frame_stride,
 * k < n_frames, and folds them into the image as samples first_sample + k (0 restarts the mean).  Rank r of N
 * calls it with first_frame = r, frame_stride = N; the ranks' means are then combined as sum(n_r / n * mean_r)
 * (raytracer_b200.multi_gpu.SppSliceReduce: one NCCL reduce).  Same estimator, float sums in another order. */
RT_API int rt_trace_frames_sliced(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t first_frame,
                                  uint32_t n_frames, uint32_t frame_stride, uint32_t first_sample);
RT_API int rt_synchronize(rt_ctx* ctx);

/* ---- consume ----------------------------------------------------------------------------
 * rt_readback: synchronises and copies this context's rows into a full-size W*H RGBA32F host
 * image (rows owned by other strips are left untouched).  A multi-GPU context gathers the strips of all its
 * devices: every device writes its rows straight into dst (side by side over their own PCIe links when dst is
 * pinned host memory; staged copies otherwise). */
RT_API int rt_readback(rt_ctx* ctx, float* dst_rgba32f, size_t bytes);
/* what the blit to B8G8R8A8_SRGB shows (raytracing_renderer.odin:299-316): clamp01 + sRGB OETF */
RT_API int rt_readback_srgb8(rt_ctx* ctx, uint8_t* dst_rgba8, size_t bytes);
/* device-resident compact strip buffer (local_rows * W * 4 floats) for NCCL gathers by the host */
RT_API int rt_device_image(rt_ctx* ctx, void** device_ptr, uint32_t* local_rows);
/* checkpoint/resume of a long accumulation: upload a full-size image as the running mean */
RT_API int rt_upload_accum(rt_ctx* ctx, const float* src_rgba32f, size_t bytes);
/* strip partition in force (strip_count > 1 or a multi-GPU context): owners[s] = rank / device slot that renders strip s
 * (rows s*strip_rows ...).  owners may be NULL to query n_strips.  With RT_BALANCE_COST the table changes when an
 * accumulation restarts after rt_resize or rt_rebalance; every rank of a partition derives the same table. */
RT_API int rt_get_strip_owners(rt_ctx* ctx, uint32_t* owners, uint32_t capacity, uint32_t* n_strips);
/* the calibration pass on its own: out4[4 y + k] = {traversal work of the primary rays, of the continuation rays, closest-hit
 * shader invocations, continuation rays} of image row y for the frame `frame_number` (capacity_rows >= height) */
RT_API int rt_measure_row_costs(rt_ctx* ctx, const float camera_ubo[64], uint32_t frame_number, uint32_t* out4, uint32_t capacity_rows);
/* a strip -> rank table decided by the host (measured step times, say); every rank keeps its round-robin row count; the
 * accumulation must restart with frame_number 0 afterwards, because image rows change hands */
RT_API int rt_set_strip_owners(rt_ctx* ctx, const uint32_t* owners, uint32_t n_strips);
/* RT_BALANCE_COST: measure the strips again at the next frame_number == 0 trace (the camera settled somewhere else) */
RT_API int rt_rebalance(rt_ctx* ctx);

/* ---- G-buffer pass: shaders/restir.rgen:43-92, restir.rchit:36-59, restir.rmiss:18-20 ------ */
RT_API int rt_trace_gbuffer(rt_ctx* ctx, const float camera_ubo[64]);
/* any pointer may be NULL.  albedo/output rgba8 (4 B), normal_rough/emission rgba16f (8 B),
 * position_metal rgba32f (16 B) per pixel, full-size W*H images. */
RT_API int rt_readback_gbuffer(rt_ctx* ctx, uint8_t* albedo, uint16_t* normal_rough, float* position_metal,
                        uint16_t* emission, uint8_t* output);

/* ---- rasterised G-buffer variant: shaders/restir/gbuffer.vert:22-37 + gbuffer.frag:26-39 (no host code of the reference draws
 * with them; the fixed-function state is stated in rt_raster.cu).  Five float4 attachments, full-size W*H images, any pointer
 * may be NULL: albedo|1, normalize(normal)|metallic, world position|roughness, emission_power*emission_color|isEmissive,
 * roughness|metallic|isEmissive|0.  Note gbuffer.vert mirrors position and normal in y before the model matrix. */
RT_API int rt_raster_gbuffer(rt_ctx* ctx, const float camera_ubo[64]);
RT_API int rt_readback_raster_gbuffer(rt_ctx* ctx, float* albedo, float* normal_metal, float* position_rough, float* emission, float* props);

/* ---- ReSTIR DI on the G-buffer (layout: shaders/restir/restir_structs.glsl; algorithm: the
 * reference has none — shaders/restir/restir_di.rgen:17-22 is a stub; parity unpinned) -------- */
RT_API void rt_default_restir_params(rt_restir_params* p);
RT_API int rt_restir_frame(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t frame_number,
                    const rt_restir_params* params);
RT_API int rt_readback_reservoirs(rt_ctx* ctx, rt_reservoir* dst, size_t bytes);

/* ---- instrumentation -------------------------------------------------------------------- */
RT_API int rt_get_counters(rt_ctx* ctx, rt_counters* out);
RT_API int rt_reset_counters(rt_ctx* ctx);
/* device time (ms) of the last rt_trace_frame(s)/rt_restir_frame/rt_trace_gbuffer call */
RT_API int rt_last_trace_ms(rt_ctx* ctx, float* ms);
/* a user event pair on the context's stream: record which = 0 (start) / 1 (stop); the elapsed call
 * synchronises on the stop event.  Lets a caller time any sequence of calls on the device. */
RT_API int rt_event_record(rt_ctx* ctx, int which);
RT_API int rt_event_elapsed_ms(rt_ctx* ctx, float* ms);
/* the same pair per device of a multi-GPU context (one entry otherwise); ms may be NULL to query n_devices */
RT_API int rt_event_elapsed_per_device(rt_ctx* ctx, float* ms, uint32_t capacity, uint32_t* n_devices);
/* per-kernel device times of the wavefront pipeline: when enabled, every launch is bracketed by CUDA
 * events on the launching stream; rt_get_kernel_profile sums them since the last enable/reset. */
/* RT_K_EXTEND: per-ray closest-hit traversal (all bounces but the first, and first-bounce tiles the packet kernel
 * hands back); RT_K_EXTEND_PACKET: first bounce traced as packets */
enum { RT_K_RAYGEN = 0, RT_K_EXTEND = 1, RT_K_SHADE = 2, RT_K_SHADOW = 3, RT_K_ACCUMULATE = 4, RT_K_EXTEND_PACKET = 5, RT_K_COUNT = 6 };
typedef struct rt_kernel_profile {
    float ms[RT_K_COUNT];
    uint32_t launches[RT_K_COUNT];
} rt_kernel_profile;
RT_API int rt_set_profiling(rt_ctx* ctx, int enable);
RT_API int rt_get_kernel_profile(rt_ctx* ctx, rt_kernel_profile* out);

/* ---- verification hooks (used by tests; same kernels as the frame path) --------------------
 * rays: n x 8 floats (ox,oy,oz,tmin, dx,dy,dz,tmax), host memory.
 * closest: out_tuv n x 3 floats, out_id n x u32 flat triangle id (0xFFFFFFFF = miss).
 * any: out_occluded n x u32 (1 = a triangle of an instance != ignore_inst[i] lies in range). */
RT_API int rt_trace_rays_closest(rt_ctx* ctx, const float* rays, uint32_t n, float* out_tuv, uint32_t* out_id);
RT_API int rt_trace_rays_any(rt_ctx* ctx, const float* rays, uint32_t n, const uint32_t* ignore_inst, uint32_t* out_occluded);
/* Closest hit without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, host pointers and bit-identical results as rt_trace_rays_closest, but the device staging of rays and
 * results is a file-scope static __device__ array of RT_STATIC_MAX_RAYS x 48 B + 256 B (3 MiB, held by every GPU that loads
 * the library until the process exits) instead of per-call cudaMalloc; the call makes no cudaMalloc / cudaFree and no
 * allocating library call.  n > RT_STATIC_MAX_RAYS -> RT_ERR_INVALID_ARGUMENT (split the batch).  The scene must already
 * be resident: pending edits (which need allocations to reach the device) -> RT_ERR_NO_SCENE; rt_scene_info_get or any
 * trace call makes it resident.  Every call on a device shares the one array, which keeps the last call's contents (the
 * work cursor is cleared per call): calls that overlap on one GPU race — one call at a time per device. */
#define RT_STATIC_MAX_RAYS 65536u
RT_API uint32_t rt_static_max_rays(void);
RT_API int rt_trace_rays_closest_static(rt_ctx* ctx, const float* rays, uint32_t n, float* out_tuv, uint32_t* out_id);
/* Diagnostic, host-only (needs no device and no context; NO COUNTERPART IN THE REFERENCE): 1 when an object made of this mesh, this
 * model matrix (column-major 4x4, as rt_object.model) and this material qualifies for the first-hit filter of the light-less path
 * tracer — metallic == 1 with albedo and roughness finite and at most 1e6 in magnitude; vertex normals finite, 0.5 - 2 long and, per triangle,
 * within a half space of their sum; mat3(model) finite, of ordinary size and conditioned below 1e4 — and 0 otherwise or for a
 * malformed mesh.  The image never depends on the answer: the filter only ends paths whose outcome is exactly what the hit shader
 * computes for them (simple.rchit:403-449 and :690-700, simple.rgen:111-118); objects that do not qualify go through the hit shader. */
RT_API int rt_filter_qualifies(const rt_mesh* mesh, const float model[16], const rt_material* material);
/* BVH download for the bit-exact comparison with the deterministic CPU twin.
 * nodes: n_nodes x 128 B, tris: n_triangles x 48 B, morton/sorted ids: n_triangles each; NULL skips. */
RT_API int rt_get_bvh(rt_ctx* ctx, void* nodes, size_t nodes_bytes, void* tris, size_t tris_bytes, uint64_t* morton,
               uint32_t* sorted_ids);

#ifdef __cplusplus
}
#endif
#endif /* RT_API_H */
