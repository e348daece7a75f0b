// rt_internal.h — host-side declarations shared by the translation units of librt.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "rt_types.cuh"

namespace rt {

#define RT_CUDA_TRY(expr)                                                                    \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                   \
            return (_e == cudaErrorMemoryAllocation) ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA;   \
        }                                                                                    \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t alloc(size_t count) {
        release();
        if (count == 0) count = 1;
        cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T));
        if (e == cudaSuccess) n = count; else p = nullptr;
        return e;
    }
    cudaError_t ensure(size_t count) { return (count <= n && p) ? cudaSuccess : alloc(count); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr; n = 0;
    }
    ~DevBuf() { release(); }
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
};

// Scratch that lives for one call: stream-ordered allocation from the CONTEXT's own memory pool (created in rt_create with
// an unlimited release threshold, destroyed in rt_destroy; the device's default pool, which the embedding host may be
// using, is left alone).  After the first build the ~25 temporaries of a rebuild cost microseconds instead of a
// cudaMalloc + a device-synchronising cudaFree each.  The pool in force is set per calling thread by PoolScope.
inline cudaMemPool_t& current_pool() {
    static thread_local cudaMemPool_t pool = nullptr;
    return pool;
}
struct PoolScope {
    cudaMemPool_t saved;
    explicit PoolScope(cudaMemPool_t p) : saved(current_pool()) { current_pool() = p; }
    ~PoolScope() { current_pool() = saved; }
};
template <typename T>
struct PoolBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaStream_t s = nullptr;
    cudaError_t alloc(size_t count, cudaStream_t stream) {
        release();
        s = stream;
        if (count == 0) count = 1;
        cudaMemPool_t pool = current_pool();
        cudaError_t e = pool ? cudaMallocFromPoolAsync((void**)&p, count * sizeof(T), pool, stream)
                             : cudaMallocAsync((void**)&p, count * sizeof(T), stream);
        if (e == cudaSuccess) n = count; else p = nullptr;
        return e;
    }
    void release() {
        if (p) cudaFreeAsync(p, s);
        p = nullptr; n = 0;
    }
    ~PoolBuf() { release(); }
    PoolBuf() = default;
    PoolBuf(const PoolBuf&) = delete;
    PoolBuf& operator=(const PoolBuf&) = delete;
};

// ---- BVH build (rt_build.cu) --------------------------------------------------------------
struct BuildInput {
    const DevObject* d_objects;  // device
    const uint32_t* d_tri_offsets;  // device, n_objects + 1 entries (prefix sums of triangle counts)
    const rt_vertex* d_vertices;
    const uint32_t* d_indices;
    uint32_t n_objects;
    uint32_t n_tris;  // instanced (flat) triangle count
};
struct BuildOutput {
    DevBuf<uint4> nodes;        // Node8 x n_nodes
    DevBuf<float4> tris;        // BvhTri x n_tris, in leaf order
    DevBuf<unsigned long long> morton;  // sorted keys
    DevBuf<uint32_t> sorted_ids;        // flat ids in key order
    uint32_t n_nodes = 0;
    uint32_t depth = 0;
    float build_ms = 0.0f;
    // kept for refits: the exact box of every wide node, the first node of every level (nodes are in level order, so a
    // bottom-up pass is "levels last to first"), and the surface-area sum of the tree as built
    DevBuf<float4> node_lo, node_hi;
    std::vector<uint32_t> level_first;  // level_first[d] .. level_first[d + 1] - 1 = nodes of depth d; size depth + 1
    double sa_build = 0.0;
    uint32_t refits = 0;                // refits since the last build
    float refit_ms = 0.0f;
};
// returns rt_status; err receives a message on failure
int build_bvh8(const BuildInput& in, BuildOutput& out, cudaStream_t stream, std::string& err);
// Refit after object transforms changed: the triangles of the objects flagged in d_dirty (one byte per object) are flattened
// again in place, then every node's boxes are recomputed bottom-up over the unchanged topology.  *sa_now = the surface-area
// sum afterwards (compare with out.sa_build to decide whether the tree has degraded enough to rebuild).
int refit_bvh8(const BuildInput& in, BuildOutput& out, const uint8_t* d_dirty, double* sa_now, cudaStream_t stream, std::string& err);

// ---- frame kernels (rt_frame.cu) -----------------------------------------------------------
struct WavefrontBuffers;  // opaque to the API layer
// CUDA-event brackets around every launch of a kernel class (rt_set_profiling)
struct KernelProfiler {
    struct Span { cudaEvent_t a, b; int k; };
    std::vector<Span> spans;
    size_t used = 0;
    bool enabled = false;
    void begin(int k, cudaStream_t s) {
        if (!enabled) return;
        if (used == spans.size()) {
            Span sp; sp.k = k;
            cudaEventCreate(&sp.a); cudaEventCreate(&sp.b);
            spans.push_back(sp);
        }
        spans[used].k = k;
        cudaEventRecord(spans[used].a, s);
    }
    void end(cudaStream_t s) {
        if (!enabled) return;
        cudaEventRecord(spans[used].b, s);
        used++;
    }
    void reset() { used = 0; }
    ~KernelProfiler() { for (auto& sp : spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); } }
};

// Persistent-grid sizes of the frame kernels, per context (= per device) and per COUNT variant of the templates: filled by
// the first frame a context launches (occupancy query on ITS device), never shared between contexts.
struct LaunchDims {
    int num_sms = 0;
    struct Set { bool ready = false; int ext = 0, shw = 0, shd = 0, shd0 = 0, shd0d = 0, pkt = 0, coop = 0, nee = 0; } set[2];
    bool srgb_table_ready = false;  // c_srgb_thresholds uploaded to this context's device
};

struct FrameResources {
    LaunchDims* dims = nullptr;
    float4* image = nullptr;      // local_rows x width RGBA32F running mean
    DevCounters* counters = nullptr;
    WavefrontBuffers* wf = nullptr;
    KernelProfiler* prof = nullptr;
    uint64_t* launches = nullptr;  // host-side launch counter
    cudaEvent_t wait_accumulate = nullptr;  // the accumulation of this batch waits for this event (the previous batch's, other stream)
    cudaEvent_t done_accumulate = nullptr;  // recorded after the accumulation of this batch
};
int wavefront_alloc(WavefrontBuffers** wf, uint32_t n_pixels, std::string& err);
// fills DevObject::spec_prob of every object with the hit shader's own specularProbability of its material (first-hit filter)
int launch_object_spec_prob(DevObject* objects, const rt_material* materials, uint32_t n, cudaStream_t stream);
void wavefront_free(WavefrontBuffers* wf);
int launch_frame_wavefront(const SceneView& sc, const FrameParams& fp, FrameResources& res, cudaStream_t stream,
                           std::string& err);
// closest hits of one un-jittered ray per pixel CORNER of the listed rows (restir.rgen:47-57): through the packet kernel into the
// wavefront state's hit array when that has a slot per ray, else one ray per thread into own_hits.  *hits = where they are,
// slot = row-list index * width + x: {u bits, v bits, instance (0xFFFFFFFF = miss), primitive}.  Same records either way.
int launch_corner_hits(const SceneView& sc, const FrameParams& fp, FrameResources& res, const uint32_t* d_rows, uint32_t n_rows,
                       uint4* own_hits, cudaStream_t stream, std::string& err, const uint4** hits);
uint32_t wavefront_slots(const WavefrontBuffers* w);
int launch_frame_megakernel(const SceneView& sc, const FrameParams& fp, FrameResources& res, cudaStream_t stream,
                            std::string& err);
// strip-cost calibration (RT_BALANCE_COST): one path per pixel of the FULL frame `fp` describes (strip_count must be 1 in
// it), no image; row_cost[4 y + k] += {primary traversal work, continuation traversal work, hits, continuation rays} of row y
int launch_row_cost(const SceneView& sc, const FrameParams& fp, uint32_t* d_row_cost, int num_sms, cudaStream_t stream, std::string& err);
int launch_trace_rays(const SceneView& sc, const float* d_rays, uint32_t n, bool any, const uint32_t* d_ignore,
                      float* d_tuv, uint32_t* d_id, DevCounters* counters, int num_sms, cudaStream_t stream, std::string& err);
struct GBufferImages {
    uchar4* albedo; ushort4* normal_rough; float4* position_metal; ushort4* emission; uchar4* output;
};
int launch_gbuffer(const SceneView& sc, const FrameParams& fp, const GBufferImages& g, DevCounters* counters, int num_sms,
                   cudaStream_t stream, std::string& err);
int launch_srgb8(const float4* image, uchar4* out, uint32_t n, LaunchDims& dims, cudaStream_t stream, std::string& err);

}  // namespace rt

// ---- the context behind the C ABI (rt_api.cu, rt_restir.cu) --------------------------------------
struct HostMesh {
    std::vector<rt_vertex> v;
    std::vector<uint32_t> idx;
    uint32_t vertex_base = 0, index_base = 0;
    bool normals_safe = false;  // every interpolated normal on the mesh is finite and non-zero (mesh_normals_safe, rt_api.cu)
};

#ifndef RT_N_STREAMS
#define RT_N_STREAMS 2
#endif
struct rt_ctx {
    rt_config cfg;
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string error;
    // multi-GPU fan-out (rt_config.n_devices >= 2): this context then owns no device state of its own; child k renders
    // strip_index k of strip_count n on device_ids[k], and every entry point loops over the children
    std::vector<rt_ctx*> children;
    rt::LaunchDims dims;
    cudaMemPool_t pool = nullptr;  // private stream-ordered pool for build scratch (PoolBuf)
    // strip -> rank assignment (RT_BALANCE_COST): global strip of every local band, host + device copies; empty = round robin
    std::vector<uint32_t> band_strip;
    std::vector<uint32_t> strip_owner;  // global strip -> rank, for every rank (the calibration is the same everywhere)
    rt::DevBuf<uint32_t> d_band_strip;
    bool balance_pending = false;  // calibrate at the next frame_number == 0 trace

    // host copies (the reference keeps the Scene on the host too and re-derives tables on edits)
    std::vector<HostMesh> meshes;
    std::vector<rt_object> objects;
    std::vector<rt_material> materials;
    std::vector<uint32_t> light_index;  // per object, 0 after a bake (gpu_scene.odin:333-336)
    bool has_scene = false;
    bool geometry_dirty = false;  // BVH must be rebuilt before the next trace
    std::vector<uint32_t> refit_objects;  // objects whose transform changed since the last flush (refit, not rebuild)
    bool tables_dirty = false;    // objects / lights / materials must be re-uploaded

    // device scene
    rt::DevBuf<rt_vertex> d_vertices;
    rt::DevBuf<uint32_t> d_indices;
    rt::DevBuf<rt::DevObject> d_objects;
    rt::DevBuf<uint32_t> d_tri_offsets;
    rt::DevBuf<rt::DevLight> d_lights;
    rt::DevBuf<rt_material> d_materials;
    rt::BuildOutput bvh;
    uint32_t n_tris = 0, n_lights = 0;
    bool any_dead_diffuse = false;  // upload_tables: some object qualifies for the first-hit filter (DevObject::dead_diffuse)

    // target
    uint32_t width = 0, height = 0, local_rows = 0;
    uint32_t lanes = 1;  // frames traced together by the wavefront pipeline (<= cfg.frames_in_flight)
    uint32_t lanes_for = 0;  // the frames in flight the path state was last sized for (1 after rt_resize, cfg.frames_in_flight after the first batch)
    rt::DevBuf<float4> image;
    rt::DevBuf<uchar4> srgb;
    rt::DevBuf<rt::DevCounters> counters;
    rt::WavefrontBuffers* wf = nullptr;
    // the other shares of the frames in flight: consecutive batches rotate over RT_N_STREAMS streams, each with its own path
    // state, so that the tail of one batch's kernels (few blocks left, SMs idle) is filled by the next batch's
    rt::WavefrontBuffers* wfx[RT_N_STREAMS - 1] = {};    // path state of streams 1 .. RT_N_STREAMS - 1 (stream 0 uses wf)
    cudaStream_t streamx[RT_N_STREAMS - 1] = {};
    cudaEvent_t ev_fork = nullptr, ev_join[RT_N_STREAMS - 1] = {}, ev_acc[RT_N_STREAMS] = {};
    // g-buffer images (allocated on first use)
    rt::DevBuf<uchar4> g_albedo, g_output;
    rt::DevBuf<ushort4> g_normal, g_emission;
    rt::DevBuf<float4> g_position;
    bool has_gbuffer = false;
    // rasterised G-buffer variant (rt_raster.cu): visibility keys, camera copy, five float4 attachments, full frame
    rt::DevBuf<unsigned long long> ras_vis;
    rt::DevBuf<float> ras_cam;
    rt::DevBuf<float4> ras_img[5];
    bool has_raster = false;
    // ReSTIR DI (rt_restir.cu): primary surfaces, this frame's reservoirs, the history
    rt::DevBuf<float4> rs_surf;
    rt::DevBuf<rt_reservoir> rs_cur, rs_hist[2];
    // world-space light triangles for light_point (rt_restir.cu: LightTable), rebuilt when the scene tables change
    rt::DevBuf<float4> rs_light_tris;
    rt::DevBuf<uint4> rs_light_aux;
    rt::DevBuf<uint32_t> rs_light_base;
    uint32_t rs_light_epoch = 0, tables_epoch = 0;
    std::vector<uint32_t> light_tris;  // triangle count of every light, in light order (host copy)
    rt::DevBuf<uint4> rs_hits;      // primary hits of the surface pass when the wavefront state cannot hold them
    rt::DevBuf<uint32_t> rs_rows;   // global rows passes 1 and 2 cover: own rows + halo (strip partitions only)
    uint32_t rs_n_rows = 0, rs_rows_halo = 0, rs_rows_epoch = 0;
    uint32_t band_epoch = 1;        // bumped whenever the strip table or the target changes
    int rs_flip = 0;
    bool rs_history = false;   // rs_hist[rs_flip] holds last frame's reservoirs for rs_cam
    float rs_cam[64] = {0};

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t user_ev[2] = {nullptr, nullptr};
    rt::KernelProfiler prof;
    uint64_t launches = 0;
    float last_ms = 0.0f;
    uint64_t frames_traced = 0;
    bool timing_pending = false;

    void set_error(const std::string& s) { error = s; }
};

// shared by the translation units behind the C ABI (rt_api.cu)
int rt_flush_scene(rt_ctx* ctx);
int rt_require_target(rt_ctx* ctx);
rt::SceneView rt_scene_view(const rt_ctx* ctx);
void rt_fill_frame_params(const rt_ctx* ctx, const float* cam, const float* clear, uint32_t frame, rt::FrameParams* fp);
void rt_begin_timing(rt_ctx* ctx);
void rt_end_timing(rt_ctx* ctx);

