// rt_api.cu — the C ABI of librt.so (include/rt_api.h): context, scene tables, frame loop plumbing.
// Host-side data preparation mirrors src/raytracer/gpu_scene.odin (:120-150 instances, :317-397
// objects and lights, :399-428 materials, :430-601 incremental updates).  All arithmetic that shapes
// pixels runs in the CUDA kernels; there is no CPU rendering path in this library.
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <string>
#include <thread>
#include <vector>

#include "rt_internal.h"

using namespace rt;

static thread_local std::string g_create_error;

// Every entry point makes the context's device current (and leaves it current: documented in rt_api.h).
#define CTX_CHECK(ctx)                          \
    if (!(ctx)) return RT_ERR_INVALID_ARGUMENT; \
    if ((ctx)->children.empty()) cudaSetDevice((ctx)->device);
#define set_error ctx->set_error
// Multi-GPU contexts (rt_config.n_devices >= 2) own one child per device; an entry point that simply applies to every
// device runs `call` (an expression in `c`) on each child from the calling thread — all of them are stream-ordered, so
// the devices work concurrently — and reports the first failure.
#define FANOUT(ctx, call)                                                   \
    if (!(ctx)->children.empty()) {                                         \
        for (rt_ctx* c : (ctx)->children) {                                 \
            const int _rc = (call);                                         \
            if (_rc != RT_OK) { (ctx)->error = c->error; return _rc; }      \
        }                                                                   \
        return RT_OK;                                                       \
    }

static bool create_side_streams(rt_ctx* ctx) {
    for (int k = 0; k < RT_N_STREAMS; k++)
        if (cudaEventCreateWithFlags(&ctx->ev_acc[k], cudaEventDisableTiming) != cudaSuccess) return false;
    for (int k = 0; k < RT_N_STREAMS - 1; k++)
        if (cudaStreamCreateWithFlags(&ctx->streamx[k], cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&ctx->ev_join[k], cudaEventDisableTiming) != cudaSuccess)
            return false;
    return true;
}

extern "C" {

const char* rt_version(void) { return "raytracer_b200 librt 0.1 (sm_100a)"; }

void rt_default_config(rt_config* cfg) {
    if (!cfg) return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->device = -1;
    cfg->max_depth = 3;        // shaders/simple.rgen:23
    cfg->rr_start_depth = 3;   // shaders/simple.rgen:56
    cfg->radiance_clamp = 5.0f;  // shaders/simple.rgen:121
    cfg->flags = RT_FLAG_DEFAULT;
    cfg->pipeline = RT_PIPELINE_WAVEFRONT;
    cfg->strip_rows = 8;
    cfg->strip_index = 0;
    cfg->strip_count = 1;
    cfg->frames_in_flight = 8;
    cfg->n_devices = 0;
    cfg->strip_balance = RT_BALANCE_ROUND_ROBIN;
}

const char* rt_last_error(const rt_ctx* ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

int rt_create(const rt_config* cfg_in, rt_ctx** out) {
    if (!out) return RT_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    rt_config cfg;
    if (cfg_in) cfg = *cfg_in; else rt_default_config(&cfg);
    if (cfg.strip_rows == 0) cfg.strip_rows = 8;
    if (cfg.strip_count == 0) cfg.strip_count = 1;
    if (cfg.frames_in_flight == 0) cfg.frames_in_flight = 8;
    if (cfg.frames_in_flight > 1024) cfg.frames_in_flight = 1024;
    if (cfg.max_depth < 1 || cfg.rr_start_depth < 1 || cfg.strip_index >= cfg.strip_count ||
        cfg.pipeline > RT_PIPELINE_MEGAKERNEL || !(cfg.radiance_clamp == cfg.radiance_clamp)) {
        g_create_error = "rt_create: invalid configuration";
        return RT_ERR_INVALID_ARGUMENT;
    }
    if ((cfg.flags & (RT_FLAG_DIRECT_LIGHTING | RT_FLAG_MIS)) != (RT_FLAG_DIRECT_LIGHTING | RT_FLAG_MIS)) {
        g_create_error = "rt_create: only USE_DIRECT_LIGHTING=1, USE_MIS=1 (the shipped shader) is built";
        return RT_ERR_INVALID_ARGUMENT;
    }
    if (cfg.n_devices > 8 || cfg.strip_balance > RT_BALANCE_COST) {
        g_create_error = "rt_create: n_devices > 8 or unknown strip_balance";
        return RT_ERR_INVALID_ARGUMENT;
    }
    if (cfg.n_devices >= 2) {
        // one sub-context per device: device k renders strip_index k of strip_count n_devices
        rt_ctx* parent = new rt_ctx();
        parent->cfg = cfg;
        parent->device = cfg.device_ids[0];
        for (uint32_t k = 0; k < cfg.n_devices; k++) {
            rt_config cc = cfg;
            cc.n_devices = 0;
            cc.device = cfg.device_ids[k];
            cc.strip_index = k;
            cc.strip_count = cfg.n_devices;
            rt_ctx* child = nullptr;
            const int rc = rt_create(&cc, &child);
            if (rc != RT_OK) {
                g_create_error = "rt_create: device " + std::to_string(cfg.device_ids[k]) + ": " + g_create_error;
                rt_destroy(parent);
                return rc;
            }
            parent->children.push_back(child);
        }
        *out = parent;
        return RT_OK;
    }
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
        cudaGetLastError();
        g_create_error = "rt_create: no CUDA device (this library has no CPU fallback)";
        return RT_ERR_NO_DEVICE;
    }
    int dev = cfg.device;
    if (dev < 0) { if (cudaGetDevice(&dev) != cudaSuccess) dev = 0; }
    if (dev >= n_dev) { g_create_error = "rt_create: device ordinal out of range"; return RT_ERR_NO_DEVICE; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major != 10) {
        g_create_error = "rt_create: device is not sm_100 (kernels are built for sm_100a only)";
        return RT_ERR_NO_DEVICE;
    }
    if (cudaSetDevice(dev) != cudaSuccess) { g_create_error = "rt_create: cudaSetDevice failed"; return RT_ERR_CUDA; }
    rt_ctx* ctx = new rt_ctx();
    ctx->cfg = cfg;
    ctx->device = dev;
    {   // a private pool keeps the freed build scratch (PoolBuf): rebuilds then reuse it without a driver round trip, and the
        // device's default pool — which the embedding host may use — keeps its own release policy
        cudaMemPoolProps props;
        memset(&props, 0, sizeof(props));
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        if (cudaMemPoolCreate(&ctx->pool, &props) == cudaSuccess) {
            uint64_t keep = ~0ull;
            cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep);
        } else {
            cudaGetLastError();
            ctx->pool = nullptr;  // PoolBuf then falls back to the default pool, untouched
        }
    }
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming) != cudaSuccess || !create_side_streams(ctx) ||
        cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess ||
        ctx->counters.alloc(1) != cudaSuccess) {
        g_create_error = std::string("rt_create: ") + cudaGetErrorString(cudaGetLastError());
        delete ctx;
        return RT_ERR_CUDA;
    }
    cudaMemsetAsync(ctx->counters.p, 0, sizeof(DevCounters), ctx->stream);
    *out = ctx;
    return RT_OK;
}

void rt_destroy(rt_ctx* ctx) {
    if (!ctx) return;
    if (!ctx->children.empty() || ctx->cfg.n_devices >= 2) {
        for (rt_ctx* c : ctx->children) rt_destroy(c);
        delete ctx;
        return;
    }
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (cudaStream_t st : ctx->streamx) if (st) cudaStreamSynchronize(st);
    if (ctx->wf) wavefront_free(ctx->wf);
    for (rt::WavefrontBuffers* w : ctx->wfx) if (w) wavefront_free(w);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    for (cudaEvent_t e : ctx->ev_join) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->ev_acc) if (e) cudaEventDestroy(e);
    for (cudaStream_t st : ctx->streamx) if (st) cudaStreamDestroy(st);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    for (int k = 0; k < 2; k++) if (ctx->user_ev[k]) cudaEventDestroy(ctx->user_ev[k]);
    ctx->d_band_strip.release();
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
    delete ctx;
}

}  // extern "C"

// ---- scene tables ------------------------------------------------------------------------------
// ---- first-hit filter: which objects qualify (DevObject::dead_diffuse) -----------------------------------------------------
// A hit that draws the Lambert lobe on a material with metallic == 1 multiplies the throughput by exactly zero
// (simple.rchit:416-422: albedo * (1 - metallic) * (1 - F) / pi), so the path ends without a colour term — PROVIDED every
// intermediate of the hit shader is finite: a NaN normal turns `payload.hit` false and the ray-gen shader adds the sky term
// instead (the object-without-normals rule).  The three checks below guarantee finiteness, conservatively.
// True when every vertex normal is finite with length in [0.5, 2] and, for every triangle, each of its three normals has a
// dot product >= 0.25 with their sum g: any barycentric mix n of them then has g . n >= 0.25, i.e. |n| >= 0.25 / |g| >= 0.04 —
// finite and safely non-zero (pairwise angles up to ~100 degrees pass, which coarse smooth-shaded meshes need).
static bool mesh_normals_safe(const rt_vertex* vtx, size_t nv, const uint32_t* idx, size_t ni) {
    for (size_t q = 0; q < nv; q++) {
        const rt_vertex& v = vtx[q];
        double l2 = 0.0;
        for (int k = 0; k < 3; k++) {
            if (!std::isfinite(v.normal[k])) return false;
            l2 += (double)v.normal[k] * v.normal[k];
        }
        if (!(l2 >= 0.25 && l2 <= 4.0)) return false;
    }
    for (size_t t = 0; t + 2 < ni; t += 3) {
        const float* n[3] = {vtx[idx[t]].normal, vtx[idx[t + 1]].normal, vtx[idx[t + 2]].normal};
        const double g[3] = {(double)n[0][0] + n[1][0] + n[2][0], (double)n[0][1] + n[1][1] + n[2][1], (double)n[0][2] + n[1][2] + n[2][2]};
        for (int i = 0; i < 3; i++)
            if (!(n[i][0] * g[0] + n[i][1] * g[1] + n[i][2] * g[2] >= 0.25)) return false;
    }
    return true;
}
// True when mat3(model) is finite, of ordinary size (Frobenius norm in [1e-6, 1e6]) and well conditioned (norm x norm of the
// inverse <= 1e4): the float normal matrix the device derives from it (cofactors / determinant) then maps a unit normal to a
// finite vector whose squared length neither overflows nor underflows.
static bool model_matrix_safe(const float* model16) {
    double a[3][3], fro2 = 0.0;
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++) {
            a[r][c] = model16[c * 4 + r];
            if (!std::isfinite(a[r][c])) return false;
            fro2 += a[r][c] * a[r][c];
        }
    const double c00 = a[1][1] * a[2][2] - a[1][2] * a[2][1], c01 = a[1][2] * a[2][0] - a[1][0] * a[2][2], c02 = a[1][0] * a[2][1] - a[1][1] * a[2][0];
    const double c10 = a[0][2] * a[2][1] - a[0][1] * a[2][2], c11 = a[0][0] * a[2][2] - a[0][2] * a[2][0], c12 = a[0][1] * a[2][0] - a[0][0] * a[2][1];
    const double c20 = a[0][1] * a[1][2] - a[0][2] * a[1][1], c21 = a[0][2] * a[1][0] - a[0][0] * a[1][2], c22 = a[0][0] * a[1][1] - a[0][1] * a[1][0];
    const double det = a[0][0] * c00 + a[0][1] * c01 + a[0][2] * c02;
    if (!std::isfinite(det) || det == 0.0) return false;
    const double cof2 = c00 * c00 + c01 * c01 + c02 * c02 + c10 * c10 + c11 * c11 + c12 * c12 + c20 * c20 + c21 * c21 + c22 * c22;
    const double fro = std::sqrt(fro2), inv_fro = std::sqrt(cof2) / std::fabs(det);
    return fro >= 1e-6 && fro <= 1e6 && std::isfinite(inv_fro) && fro * inv_fro <= 1e4;
}
// metallic == 1 makes albedo * (1 - metallic) an exact zero; albedo and roughness must be finite AND of ordinary size: a roughness
// beyond ~1e19 overflows D_GGX's alpha^2 into inf / inf = NaN, the pdf test of simple.rchit:693 then fails and the path gets a
// sky term instead of ending quietly (glTF keeps both in [0, 1]; the bound is generous).
static bool material_dead_diffuse(const rt_material& m) {
    for (int k = 0; k < 3; k++)
        if (!(std::fabs(m.albedo[k]) <= 1e6f)) return false;  // also rejects NaN
    return m.metallic == 1.0f && std::fabs(m.roughness) <= 1e6f;
}

static void make_dev_object(const rt_ctx* ctx, uint32_t i, uint32_t tri_offset, DevObject* o) {
    const rt_object& src = ctx->objects[i];
    const HostMesh& m = ctx->meshes[src.mesh_index];
    memset(o, 0, sizeof(*o));
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 3; r++) o->model[c * 3 + r] = src.model[c * 4 + r];
    o->material_index = src.material_index;
    o->light_index = ctx->light_index[i];
    o->vertex_base = m.vertex_base;
    o->index_base = m.index_base;
    o->tri_offset = tri_offset;
    o->n_tris = (uint32_t)(m.idx.size() / 3);
    o->mesh_index = src.mesh_index;
    o->dead_diffuse = (m.normals_safe && model_matrix_safe(src.model) && material_dead_diffuse(ctx->materials[src.material_index])) ? 1u : 0u;
}

__global__ void normal_matrix_kernel(DevObject* objects, uint32_t n) {
    // transpose(inverse(mat3(ObjectToWorld))) for simple.rchit:607, evaluated once per object with the
    // contract's arithmetic (cofactors / determinant)
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    DevObject* o = objects + i;
    float m16[16];
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 3; r++) m16[c * 4 + r] = o->model[c * 3 + r];
    cofactor_normal_matrix(m16, o->nrm);
}

static int upload_tables(rt_ctx* ctx) {
    const uint32_t no = (uint32_t)ctx->objects.size();
    std::vector<DevObject> dobj(no);
    std::vector<uint32_t> offs(no + 1);
    std::vector<DevLight> lights;
    uint32_t off = 0;
    ctx->any_dead_diffuse = false;
    for (uint32_t i = 0; i < no; i++) {
        offs[i] = off;
        make_dev_object(ctx, i, off, &dobj[i]);
        off += dobj[i].n_tris;
        if (dobj[i].dead_diffuse) ctx->any_dead_diffuse = true;
        // a light is any object whose material has emission_power > 0 (gpu_scene.odin:366-379)
        const rt_material& mat = ctx->materials[ctx->objects[i].material_index];
        if (mat.emission_power > 0.0f) {
            DevLight l;
            memcpy(l.transform, dobj[i].model, sizeof(l.transform));
            l.object_index = i;
            l.num_triangles = dobj[i].n_tris;
            l.instance_mask = 1u << (i & 7u);
            l.emission_power = mat.emission_power;
            lights.push_back(l);
        }
    }
    offs[no] = off;
    ctx->n_tris = off;
    ctx->n_lights = (uint32_t)lights.size();
    ctx->light_tris.clear();
    for (const DevLight& l : lights) ctx->light_tris.push_back(l.num_triangles);
    ctx->tables_epoch++;
    RT_CUDA_TRY(ctx->d_objects.ensure(no));
    RT_CUDA_TRY(ctx->d_tri_offsets.ensure(no + 1));
    RT_CUDA_TRY(ctx->d_lights.ensure(lights.size()));
    RT_CUDA_TRY(ctx->d_materials.ensure(ctx->materials.size()));
    // the copies below read pageable host vectors: synchronise before they go out of scope
    if (no) RT_CUDA_TRY(cudaMemcpyAsync(ctx->d_objects.p, dobj.data(), no * sizeof(DevObject), cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA_TRY(cudaMemcpyAsync(ctx->d_tri_offsets.p, offs.data(), (no + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (!lights.empty())
        RT_CUDA_TRY(cudaMemcpyAsync(ctx->d_lights.p, lights.data(), lights.size() * sizeof(DevLight), cudaMemcpyHostToDevice, ctx->stream));
    if (!ctx->materials.empty())
        RT_CUDA_TRY(cudaMemcpyAsync(ctx->d_materials.p, ctx->materials.data(), ctx->materials.size() * sizeof(rt_material),
                                    cudaMemcpyHostToDevice, ctx->stream));
    if (no) normal_matrix_kernel<<<(no + 127) / 128, 128, 0, ctx->stream>>>(ctx->d_objects.p, no);
    RT_CUDA_TRY(cudaGetLastError());
    if (no && launch_object_spec_prob(ctx->d_objects.p, ctx->d_materials.p, no, ctx->stream) != RT_OK) {
        set_error("upload_tables: object_spec_prob_kernel failed to launch");
        return RT_ERR_CUDA;
    }
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->tables_dirty = false;
    return RT_OK;
}

static int rebuild_bvh(rt_ctx* ctx) {
    BuildInput in;
    in.d_objects = ctx->d_objects.p;
    in.d_tri_offsets = ctx->d_tri_offsets.p;
    in.d_vertices = ctx->d_vertices.p;
    in.d_indices = ctx->d_indices.p;
    in.n_objects = (uint32_t)ctx->objects.size();
    in.n_tris = ctx->n_tris;
    std::string err;
    PoolScope pool_scope(ctx->pool);
    const int rc = build_bvh8(in, ctx->bvh, ctx->stream, err);
    if (rc != RT_OK) { set_error("rt_build_scene: " + err); return rc; }
    if (ctx->bvh.depth > RT_MAX_BVH_DEPTH) {
        set_error("rt_build_scene: BVH8 deeper than the traversal stack");
        return RT_ERR_BVH_TOO_DEEP;
    }
    ctx->geometry_dirty = false;
    return RT_OK;
}

// Object transforms changed and nothing else about the geometry: refit the boxes over the existing topology (the reference
// updates its TLAS in place: acceleration_structure.odin:91 `mode = update ? .UPDATE : .BUILD`, gpu_scene.odin:457-483).
// Closest hits do not depend on the tree, so the image is the one a rebuild gives; the tree is rebuilt when its
// surface-area sum has grown past RT_REFIT_REBUILD_RATIO times the one it was built with.
#ifndef RT_REFIT_REBUILD_RATIO
#define RT_REFIT_REBUILD_RATIO 1.5
#endif
static int refit_bvh(rt_ctx* ctx, bool* rebuild) {
    const uint32_t no = (uint32_t)ctx->objects.size();
    std::vector<uint8_t> dirty(no, 0);
    for (uint32_t i : ctx->refit_objects) if (i < no) dirty[i] = 1;
    PoolScope pool_scope(ctx->pool);
    PoolBuf<uint8_t> d_dirty;
    RT_CUDA_TRY(d_dirty.alloc(no, ctx->stream));
    RT_CUDA_TRY(cudaMemcpyAsync(d_dirty.p, dirty.data(), no, cudaMemcpyHostToDevice, ctx->stream));
    BuildInput in;
    in.d_objects = ctx->d_objects.p;
    in.d_tri_offsets = ctx->d_tri_offsets.p;
    in.d_vertices = ctx->d_vertices.p;
    in.d_indices = ctx->d_indices.p;
    in.n_objects = no;
    in.n_tris = ctx->n_tris;
    std::string err;
    double sa_now = 0.0;
    const int rc = refit_bvh8(in, ctx->bvh, d_dirty.p, &sa_now, ctx->stream, err);  // synchronises: `dirty` may go out of scope
    if (rc != RT_OK) { set_error("refit: " + err); return rc; }
    *rebuild = ctx->bvh.sa_build > 0.0 && sa_now > RT_REFIT_REBUILD_RATIO * ctx->bvh.sa_build;
    return RT_OK;
}

int rt_flush_scene(rt_ctx* ctx) {
    if (!ctx->has_scene) { set_error("no scene: call rt_build_scene first"); return RT_ERR_NO_SCENE; }
    if (ctx->tables_dirty || ctx->geometry_dirty) {
        bool geo = ctx->geometry_dirty;
        ctx->rs_history = false;  // reservoirs refer to the old light tables
        int rc = upload_tables(ctx);
        if (rc != RT_OK) return rc;
        if (!geo && !ctx->refit_objects.empty() && ctx->n_tris != 0) {
            rc = refit_bvh(ctx, &geo);
            if (rc != RT_OK) return rc;
        }
        ctx->refit_objects.clear();
        if (geo) { rc = rebuild_bvh(ctx); if (rc != RT_OK) return rc; }
    }
    return RT_OK;
}

SceneView rt_scene_view(const rt_ctx* ctx) {
    SceneView sc;
    sc.nodes = ctx->bvh.nodes.p;
    sc.tris = ctx->bvh.tris.p;
    sc.objects = ctx->d_objects.p;
    sc.lights = ctx->d_lights.p;
    sc.materials = ctx->d_materials.p;
    sc.vertices = ctx->d_vertices.p;
    sc.indices = ctx->d_indices.p;
    sc.n_lights = ctx->n_lights;
    sc.n_objects = (uint32_t)ctx->objects.size();
    sc.n_tris = ctx->n_tris;
    sc.n_nodes = ctx->bvh.n_nodes;
    sc.any_dead_diffuse = (ctx->any_dead_diffuse && ctx->n_lights == 0u) ? 1u : 0u;
    return sc;
}

extern "C" {

int rt_build_scene(rt_ctx* ctx, const rt_mesh* meshes, uint32_t n_meshes, const rt_object* objects, uint32_t n_objects,
                   const rt_material* materials, uint32_t n_materials) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_build_scene(c, meshes, n_meshes, objects, n_objects, materials, n_materials));
    if ((n_meshes && !meshes) || (n_objects && !objects) || (n_materials && !materials)) {
        set_error("rt_build_scene: null array");
        return RT_ERR_INVALID_ARGUMENT;
    }
    std::vector<HostMesh> hm(n_meshes);
    size_t nv = 0, ni = 0;
    for (uint32_t i = 0; i < n_meshes; i++) {
        const rt_mesh& m = meshes[i];
        if ((m.n_vertices && !m.vertices) || (m.n_indices && !m.indices) || (m.n_indices % 3u) != 0u) {
            set_error("rt_build_scene: malformed mesh");
            return RT_ERR_INVALID_ARGUMENT;
        }
        for (uint32_t k = 0; k < m.n_indices; k++)
            if (m.indices[k] >= m.n_vertices) { set_error("rt_build_scene: vertex index out of range"); return RT_ERR_INDEX_RANGE; }
        hm[i].v.assign(m.vertices, m.vertices + m.n_vertices);
        hm[i].idx.assign(m.indices, m.indices + m.n_indices);
        hm[i].vertex_base = (uint32_t)nv;
        hm[i].index_base = (uint32_t)ni;
        hm[i].normals_safe = mesh_normals_safe(hm[i].v.data(), hm[i].v.size(), hm[i].idx.data(), hm[i].idx.size());
        nv += m.n_vertices;
        ni += m.n_indices;
    }
    uint64_t total_tris = 0;
    for (uint32_t i = 0; i < n_objects; i++) {
        if (objects[i].mesh_index >= n_meshes) { set_error("rt_build_scene: Invalid mesh index"); return RT_ERR_INDEX_RANGE; }
        if (objects[i].material_index >= n_materials) { set_error("rt_build_scene: Invalid material index"); return RT_ERR_INDEX_RANGE; }
        total_tris += hm[objects[i].mesh_index].idx.size() / 3;
    }
    if (total_tris >= 0x7FFFFFFFull || nv >= 0xFFFFFFFFull || ni >= 0xFFFFFFFFull) {
        set_error("rt_build_scene: scene too large for 32-bit triangle ids");
        return RT_ERR_INVALID_ARGUMENT;
    }
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->meshes.swap(hm);
    ctx->objects.assign(objects, objects + n_objects);
    ctx->materials.assign(materials, materials + n_materials);
    ctx->light_index.assign(n_objects, 0u);
    RT_CUDA_TRY(ctx->d_vertices.ensure(nv));
    RT_CUDA_TRY(ctx->d_indices.ensure(ni));
    for (const HostMesh& m : ctx->meshes) {  // staging copies, like buffer_init_with_staging_buffer (gpu_scene.odin:80-112)
        if (!m.v.empty())
            RT_CUDA_TRY(cudaMemcpyAsync(ctx->d_vertices.p + m.vertex_base, m.v.data(), m.v.size() * sizeof(rt_vertex),
                                        cudaMemcpyHostToDevice, ctx->stream));
        if (!m.idx.empty())
            RT_CUDA_TRY(cudaMemcpyAsync(ctx->d_indices.p + m.index_base, m.idx.data(), m.idx.size() * 4, cudaMemcpyHostToDevice,
                                        ctx->stream));
    }
    ctx->has_scene = true;
    ctx->tables_dirty = true;
    ctx->geometry_dirty = true;
    return rt_flush_scene(ctx);
}

// host-only diagnostic: the three checks behind DevObject::dead_diffuse for one (mesh, matrix, material) combination
int rt_filter_qualifies(const rt_mesh* mesh, const float model[16], const rt_material* material) {
    if (!mesh || !model || !material) return 0;
    if ((mesh->n_vertices && !mesh->vertices) || (mesh->n_indices && !mesh->indices) || (mesh->n_indices % 3u) != 0u) return 0;
    for (uint32_t k = 0; k < mesh->n_indices; k++)
        if (mesh->indices[k] >= mesh->n_vertices) return 0;
    return (mesh_normals_safe(mesh->vertices, mesh->n_vertices, mesh->indices, mesh->n_indices) && model_matrix_safe(model) &&
            material_dead_diffuse(*material)) ? 1 : 0;
}

int rt_scene_info_get(rt_ctx* ctx, rt_scene_info* out) {
    CTX_CHECK(ctx);
    if (!out) return RT_ERR_INVALID_ARGUMENT;
    if (!ctx->children.empty()) {  // every device holds the same scene and (deterministic builder) the same BVH
        const int rc0 = rt_scene_info_get(ctx->children[0], out);
        if (rc0 != RT_OK) ctx->error = ctx->children[0]->error;
        return rc0;
    }
    const int rc = rt_flush_scene(ctx);
    if (rc != RT_OK) return rc;
    memset(out, 0, sizeof(*out));
    out->n_triangles = ctx->n_tris;
    out->n_nodes = ctx->bvh.n_nodes;
    out->bvh_depth = ctx->bvh.depth;
    out->n_lights = ctx->n_lights;
    out->n_objects = (uint32_t)ctx->objects.size();
    out->n_materials = (uint32_t)ctx->materials.size();
    out->build_ms = ctx->bvh.build_ms;
    out->refits = ctx->bvh.refits;
    return RT_OK;
}

// gpu_scene_update_material (gpu_scene.odin:562-601): overwrite the record, recompile the lights
int rt_update_material(rt_ctx* ctx, uint32_t material_index, const rt_material* m) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_update_material(c, material_index, m));
    if (!ctx->has_scene) { set_error("no scene"); return RT_ERR_NO_SCENE; }
    if (!m || material_index >= ctx->materials.size()) { set_error("rt_update_material: bad index"); return RT_ERR_INDEX_RANGE; }
    ctx->materials[material_index] = *m;
    ctx->tables_dirty = true;
    return RT_OK;
}

// gpu_scene_update_object_transform (gpu_scene.odin:457-483): instance transform + TLAS update
int rt_update_object_transform(rt_ctx* ctx, uint32_t object_index, const float model[16]) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_update_object_transform(c, object_index, model));
    if (!ctx->has_scene) { set_error("no scene"); return RT_ERR_NO_SCENE; }
    if (!model || object_index >= ctx->objects.size()) { set_error("rt_update_object_transform: bad index"); return RT_ERR_INDEX_RANGE; }
    memcpy(ctx->objects[object_index].model, model, 64);
    ctx->tables_dirty = true;
    ctx->refit_objects.push_back(object_index);  // boxes are refitted at the next flush; no rebuild unless the tree degrades
    return RT_OK;
}

// gpu_scene_update_object with changed_material (gpu_scene.odin:517-560): the one place light_index is set
int rt_update_object_material(rt_ctx* ctx, uint32_t object_index, uint32_t material_index) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_update_object_material(c, object_index, material_index));
    if (!ctx->has_scene) { set_error("no scene"); return RT_ERR_NO_SCENE; }
    if (object_index >= ctx->objects.size() || material_index >= ctx->materials.size()) {
        set_error("rt_update_object_material: bad index");
        return RT_ERR_INDEX_RANGE;
    }
    ctx->objects[object_index].material_index = material_index;
    uint32_t li = 0, k = 0;
    for (uint32_t i = 0; i < ctx->objects.size(); i++) {
        if (ctx->materials[ctx->objects[i].material_index].emission_power > 0.0f) {
            if (i == object_index) li = k;
            k++;
        }
    }
    ctx->light_index[object_index] = li;
    ctx->tables_dirty = true;
    return RT_OK;
}

// ---- add / remove: gpu_scene_add_object / remove_object / add_material / remove_material (gpu_scene.odin:430-455), which the
// renderer dispatches on Object_Added / Object_Removed / Material_Added / Material_Removed (raytracing_renderer.odin:163-183).
// The reference recompiles the object / light / material tables from its host scene; here the host scene lives behind the ABI
// (the copies rt_build_scene made), so the calls carry the change itself.  Meshes are not re-uploaded: an added object
// instances a mesh the scene already holds (scene.odin has no change event for meshes either).
int rt_add_material(rt_ctx* ctx, const rt_material* m, uint32_t* index_out) {
    CTX_CHECK(ctx);
    if (!ctx->children.empty()) {
        for (rt_ctx* c : ctx->children) {
            const int rc = rt_add_material(c, m, index_out);
            if (rc != RT_OK) { ctx->error = c->error; return rc; }
        }
        return RT_OK;
    }
    if (!ctx->has_scene) { set_error("no scene"); return RT_ERR_NO_SCENE; }
    if (!m) { set_error("rt_add_material: null material"); return RT_ERR_INVALID_ARGUMENT; }
    ctx->materials.push_back(*m);  // scene_add_material: append (scene.odin:95-98)
    if (index_out) *index_out = (uint32_t)ctx->materials.size() - 1u;
    ctx->tables_dirty = true;
    return RT_OK;
}

// scene_delete_material (scene.odin:104-116): unordered_remove — the LAST material moves into the freed slot — and every
// object that used the removed material falls back to material 0 (through scene_update_object_material, hence its
// light_index is recomputed as in rt_update_object_material).  Objects that used the moved material follow it to its new
// index (the reference leaves their index pointing past the end of the table: defined here).
int rt_remove_material(rt_ctx* ctx, uint32_t material_index) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_remove_material(c, material_index));
    if (!ctx->has_scene) { set_error("no scene"); return RT_ERR_NO_SCENE; }
    const uint32_t n = (uint32_t)ctx->materials.size();
    if (material_index >= n) { set_error("rt_remove_material: bad index"); return RT_ERR_INDEX_RANGE; }
    if (n == 1) { set_error("rt_remove_material: the scene's last material cannot go (objects fall back to material 0)"); return RT_ERR_INVALID_ARGUMENT; }
    const uint32_t last = n - 1u;
    ctx->materials[material_index] = ctx->materials[last];
    ctx->materials.pop_back();
    std::vector<uint32_t> fell_back;
    for (uint32_t i = 0; i < ctx->objects.size(); i++) {
        uint32_t& mi = ctx->objects[i].material_index;
        if (mi == material_index) { mi = 0u; fell_back.push_back(i); }
        else if (mi == last) mi = material_index;
    }
    ctx->tables_dirty = true;
    for (uint32_t i : fell_back) {
        const int rc = rt_update_object_material(ctx, i, 0u);
        if (rc != RT_OK) return rc;
    }
    return RT_OK;
}

// scene_add_object (scene.odin:165-194): append an instance of an existing mesh.  Its triangles enter the BVH (rebuild);
// light_index starts at 0 like every baked object (gpu_scene.odin:333-336).
int rt_add_object(rt_ctx* ctx, const rt_object* o, uint32_t* index_out) {
    CTX_CHECK(ctx);
    if (!ctx->children.empty()) {
        for (rt_ctx* c : ctx->children) {
            const int rc = rt_add_object(c, o, index_out);
            if (rc != RT_OK) { ctx->error = c->error; return rc; }
        }
        return RT_OK;
    }
    if (!ctx->has_scene) { set_error("no scene"); return RT_ERR_NO_SCENE; }
    if (!o) { set_error("rt_add_object: null object"); return RT_ERR_INVALID_ARGUMENT; }
    if (o->mesh_index >= ctx->meshes.size()) { set_error("rt_add_object: Invalid mesh index"); return RT_ERR_INDEX_RANGE; }          // scene.odin:176
    if (o->material_index >= ctx->materials.size()) { set_error("rt_add_object: Invalid material index"); return RT_ERR_INDEX_RANGE; }  // scene.odin:177
    uint64_t total = ctx->meshes[o->mesh_index].idx.size() / 3;
    for (const rt_object& q : ctx->objects) total += ctx->meshes[q.mesh_index].idx.size() / 3;
    if (total >= 0x7FFFFFFFull) { set_error("rt_add_object: scene too large for 32-bit triangle ids"); return RT_ERR_INVALID_ARGUMENT; }
    ctx->objects.push_back(*o);
    ctx->light_index.push_back(0u);
    if (index_out) *index_out = (uint32_t)ctx->objects.size() - 1u;
    ctx->tables_dirty = true;
    ctx->geometry_dirty = true;
    return RT_OK;
}

// Object_Removed: the objects behind it move down one index (instanceCustomIndex = object index, gpu_scene.odin:135, so the
// order of the remaining objects — and with it the order of the lights — is kept).
int rt_remove_object(rt_ctx* ctx, uint32_t object_index) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_remove_object(c, object_index));
    if (!ctx->has_scene) { set_error("no scene"); return RT_ERR_NO_SCENE; }
    if (object_index >= ctx->objects.size()) { set_error("rt_remove_object: bad index"); return RT_ERR_INDEX_RANGE; }
    ctx->objects.erase(ctx->objects.begin() + object_index);
    ctx->light_index.erase(ctx->light_index.begin() + object_index);
    ctx->refit_objects.clear();
    ctx->tables_dirty = true;
    ctx->geometry_dirty = true;
    return RT_OK;
}

// gpu_scene_update_object_mesh (gpu_scene.odin:485-515)
int rt_update_object_mesh(rt_ctx* ctx, uint32_t object_index, uint32_t mesh_index) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_update_object_mesh(c, object_index, mesh_index));
    if (!ctx->has_scene) { set_error("no scene"); return RT_ERR_NO_SCENE; }
    if (object_index >= ctx->objects.size() || mesh_index >= ctx->meshes.size()) {
        set_error("rt_update_object_mesh: bad index");
        return RT_ERR_INDEX_RANGE;
    }
    ctx->objects[object_index].mesh_index = mesh_index;
    ctx->tables_dirty = true;
    ctx->geometry_dirty = true;
    return RT_OK;
}

}  // extern "C"

// ---- target / frame loop -------------------------------------------------------------------------
static uint32_t count_local_rows(const rt_config& c, uint32_t height) {
    uint32_t n = 0;
    for (uint32_t y = 0; y < height; y++)
        if ((y / c.strip_rows) % c.strip_count == c.strip_index) n++;
    return n;
}

void rt_fill_frame_params(const rt_ctx* ctx, const float* cam, const float* clear, uint32_t frame, FrameParams* fp) {
    memcpy(fp->inv_view, cam + 32, 64);  // Camera_UBO: proj, view, inverse_view, inverse_proj (camera.odin:15-20)
    memcpy(fp->inv_proj, cam + 48, 64);
    if (clear) memcpy(fp->clear, clear, 12); else fp->clear[0] = fp->clear[1] = fp->clear[2] = 0.0f;
    fp->frame = frame;
    fp->n_lanes = 1;
    fp->frame_stride = 1;
    fp->sample = frame;
    fp->width = ctx->width;
    fp->height = ctx->height;
    fp->strip_rows = ctx->cfg.strip_rows;
    fp->strip_index = ctx->cfg.strip_index;
    fp->strip_count = ctx->cfg.strip_count;
    fp->band_strip = ctx->band_strip.empty() ? nullptr : ctx->d_band_strip.p;
    fp->local_rows = ctx->local_rows;
    fp->max_depth = ctx->cfg.max_depth;
    fp->rr_start = ctx->cfg.rr_start_depth;
    fp->clamp = ctx->cfg.radiance_clamp;
    fp->flags = ctx->cfg.flags;
}

int rt_require_target(rt_ctx* ctx) {
    if (ctx->width == 0 || ctx->height == 0) { set_error("no target: call rt_resize first"); return RT_ERR_NO_TARGET; }
    return RT_OK;
}

void rt_begin_timing(rt_ctx* ctx) { cudaEventRecord(ctx->ev0, ctx->stream); }
void rt_end_timing(rt_ctx* ctx) {
    cudaEventRecord(ctx->ev1, ctx->stream);
    ctx->timing_pending = true;
}

// ---- strips by cost (RT_BALANCE_COST) -------------------------------------------------------------------------------
// One calibration pass over the FULL frame (every rank runs the same deterministic pass and so derives the same table without
// talking to anyone), then longest-processing-time-first under the constraint that every rank keeps the NUMBER of strips the
// round-robin deal gives it — so the compact image, the path state and the frames in flight keep their sizes.
// cost of a strip = sum over its rows of W_PX * pixels + W_PRIMARY * c0 + W_SECONDARY * c1 + W_HIT * c2 (row_cost_kernel's
// counters); the weights are tenths of a picosecond of device time per unit, fitted by non-negative least squares to the 16
// per-rank step times of two 8-way partitions on a B200 (tools/fit_balance.py: rms error 1.0 %)
#ifndef RT_BALANCE_W_PX
#define RT_BALANCE_W_PX 267u
#define RT_BALANCE_W_PRIMARY 9u
#define RT_BALANCE_W_SECONDARY 12u
#define RT_BALANCE_W_HIT 1572u
#endif
static int measure_row_costs(rt_ctx* ctx, const float* camera_ubo, const float* clear_color, uint32_t frame, std::vector<uint32_t>& rows) {
    const uint32_t H = ctx->height;
    PoolScope pool_scope(ctx->pool);
    PoolBuf<uint32_t> d_cost;
    RT_CUDA_TRY(d_cost.alloc((size_t)H * 4, ctx->stream));
    RT_CUDA_TRY(cudaMemsetAsync(d_cost.p, 0, (size_t)H * 16, ctx->stream));
    FrameParams fp;
    rt_fill_frame_params(ctx, camera_ubo, clear_color, frame, &fp);
    fp.strip_count = 1; fp.strip_index = 0; fp.band_strip = nullptr; fp.local_rows = H;
    if (!ctx->dims.num_sms) cudaDeviceGetAttribute(&ctx->dims.num_sms, cudaDevAttrMultiProcessorCount, ctx->device);
    std::string err;
    const int rc = launch_row_cost(rt_scene_view(ctx), fp, d_cost.p, ctx->dims.num_sms > 0 ? ctx->dims.num_sms : 148, ctx->stream, err);
    if (rc != RT_OK) { set_error(err); return rc; }
    ctx->launches++;
    rows.resize((size_t)H * 4);
    RT_CUDA_TRY(cudaMemcpyAsync(rows.data(), d_cost.p, (size_t)H * 16, cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

// Installs a strip -> rank table: this context's bands (ascending strips), host + device copies.
static int install_owners(rt_ctx* ctx, const uint32_t* owners, uint32_t n_strips) {
    const uint32_t sr = ctx->cfg.strip_rows, H = ctx->height;
    std::vector<uint32_t> mine;
    uint32_t rows_mine = 0;
    for (uint32_t st = 0; st < n_strips; st++)
        if (owners[st] == ctx->cfg.strip_index) { mine.push_back(st); rows_mine += std::min(sr, H - st * sr); }
    const uint32_t n_full = H / sr;
    if (rows_mine != ctx->local_rows || (n_strips > n_full && !mine.empty() && owners[n_full] == ctx->cfg.strip_index && mine.back() != n_full)) {
        set_error("strip owners: every rank must keep the row count of the round-robin deal (the partial last strip stays where it is)");
        return RT_ERR_INVALID_ARGUMENT;
    }
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    for (cudaStream_t st : ctx->streamx) if (st) RT_CUDA_TRY(cudaStreamSynchronize(st));
    ctx->strip_owner.assign(owners, owners + n_strips);
    ctx->band_strip.swap(mine);
    RT_CUDA_TRY(ctx->d_band_strip.ensure(ctx->band_strip.size()));
    if (!ctx->band_strip.empty())
        RT_CUDA_TRY(cudaMemcpyAsync(ctx->d_band_strip.p, ctx->band_strip.data(), ctx->band_strip.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->rs_history = false;   // per-pixel histories belong to the rows that were here before
    ctx->has_gbuffer = false;
    ctx->band_epoch++;
    return RT_OK;
}

static int balance_strips(rt_ctx* ctx, const float* camera_ubo, const float* clear_color, uint32_t frame) {
    ctx->balance_pending = false;
    const uint32_t sr = ctx->cfg.strip_rows, N = ctx->cfg.strip_count, H = ctx->height, W = ctx->width;
    const uint32_t n_full = H / sr, n_strips = (H + sr - 1) / sr;
    if (N <= 1 || n_full == 0) return RT_OK;
    std::vector<uint32_t> rows;
    int rc = measure_row_costs(ctx, camera_ubo, clear_color, frame, rows);
    if (rc != RT_OK) return rc;
    std::vector<uint64_t> cost(n_full);
    for (uint32_t st = 0; st < n_full; st++) {
        uint64_t c = (uint64_t)RT_BALANCE_W_PX * sr * W;
        for (uint32_t r = 0; r < sr; r++) {
            const uint32_t* q = &rows[(size_t)(st * sr + r) * 4];
            c += (uint64_t)RT_BALANCE_W_PRIMARY * q[0] + (uint64_t)RT_BALANCE_W_SECONDARY * q[1] + (uint64_t)RT_BALANCE_W_HIT * q[2];
        }
        cost[st] = c;
    }
    std::vector<uint32_t> order(n_full), cap(N, 0u);
    for (uint32_t st = 0; st < n_full; st++) { order[st] = st; cap[st % N]++; }
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return cost[a] > cost[b]; });  // ties: lower strip first
    std::vector<uint64_t> load(N, 0ull);
    std::vector<uint32_t> owner(n_strips, 0u);
    for (uint32_t st : order) {
        uint32_t best = N;
        for (uint32_t r = 0; r < N; r++)
            if (cap[r] && (best == N || load[r] < load[best])) best = r;
        owner[st] = best;
        load[best] += cost[st];
        cap[best]--;
    }
    if (n_strips > n_full) owner[n_full] = n_full % N;  // the partial last strip stays its round-robin owner's last band
    return install_owners(ctx, owner.data(), n_strips);
}

#ifndef RT_TWO_STREAM_MIN_LANES
#define RT_TWO_STREAM_MIN_LANES 8
#endif
#ifndef RT_STATE_CAP_GIB
#define RT_STATE_CAP_GIB 56  /* of the B200's 180 GB: 128 frames in flight at 1080p; 41 -> 128 frames: 0.540 -> 0.525 ms/frame */
#endif
// (Re)allocates the wavefront path state for up to `frames` frames in flight (~200 B per slot; all lanes together stay under
// RT_STATE_CAP_GIB and 2^31 slots).  From RT_TWO_STREAM_MIN_LANES frames on, the state is split into equal shares for
// RT_N_STREAMS streams.
static int alloc_path_state(rt_ctx* ctx, uint32_t frames) {
    if (ctx->wf) { wavefront_free(ctx->wf); ctx->wf = nullptr; }
    for (rt::WavefrontBuffers*& w : ctx->wfx) if (w) { wavefront_free(w); w = nullptr; }
    ctx->lanes = 1;
    ctx->lanes_for = frames;
    if (ctx->cfg.pipeline != RT_PIPELINE_WAVEFRONT) return RT_OK;
    const size_t n = (size_t)ctx->local_rows * ctx->width;
    std::string err;
    uint64_t lanes = std::max<uint32_t>(frames, 1u);
    const uint64_t per_lane = std::max<uint64_t>(n, 1) * 200ull;
    lanes = std::min<uint64_t>(lanes, std::max<uint64_t>(1, ((uint64_t)RT_STATE_CAP_GIB << 30) / per_lane));
    lanes = std::min<uint64_t>(lanes, std::max<uint64_t>(1, 0x7FFFFFFFull / std::max<uint64_t>(n, 1)));
    const bool split = lanes >= RT_TWO_STREAM_MIN_LANES;
    if (split) lanes = (lanes + RT_N_STREAMS - 1) / RT_N_STREAMS;
    ctx->lanes = (uint32_t)lanes;
    int rc = wavefront_alloc(&ctx->wf, (uint32_t)(n * lanes), err);
    for (int k = 0; k < RT_N_STREAMS - 1 && rc == RT_OK && split; k++) rc = wavefront_alloc(&ctx->wfx[k], (uint32_t)(n * lanes), err);
    if (rc != RT_OK) { set_error(err); return rc; }
    return RT_OK;
}

extern "C" {

// raytracing_renderer_on_resize (raytracing_renderer.odin:108-126): new output image, accumulation restarts
int rt_resize(rt_ctx* ctx, uint32_t width, uint32_t height) {
    CTX_CHECK(ctx);
    if (width == 0 || height == 0 || (uint64_t)width * height > 0x7FFFFFFFull) {
        set_error("rt_resize: bad extent");
        return RT_ERR_INVALID_ARGUMENT;
    }
    if (!ctx->children.empty()) { ctx->width = width; ctx->height = height; }
    FANOUT(ctx, rt_resize(c, width, height));
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->width = width;
    ctx->height = height;
    ctx->local_rows = count_local_rows(ctx->cfg, height);
    // back to the round-robin assignment; with RT_BALANCE_COST the next frame-0 trace measures the strips of the new target
    ctx->band_strip.clear();
    ctx->strip_owner.clear();
    ctx->band_epoch++;
    ctx->balance_pending = ctx->cfg.strip_balance == RT_BALANCE_COST && ctx->cfg.strip_count > 1;
    const size_t n = (size_t)ctx->local_rows * width;
    RT_CUDA_TRY(ctx->image.alloc(n));
    RT_CUDA_TRY(cudaMemsetAsync(ctx->image.p, 0, std::max<size_t>(n, 1) * sizeof(float4), ctx->stream));
    ctx->srgb.release();
    ctx->has_gbuffer = false;
    ctx->has_raster = false;
    ctx->ras_vis.release();
    for (auto& b : ctx->ras_img) b.release();
    ctx->rs_history = false;
    ctx->rs_surf.release(); ctx->rs_cur.release(); ctx->rs_hist[0].release(); ctx->rs_hist[1].release();
    // path state for ONE frame in flight: what rt_trace_frame (the reference's application loop, application.odin:109-115) needs.
    // The first rt_trace_frames call with n > 1 grows it to cfg.frames_in_flight (alloc_path_state).
    return alloc_path_state(ctx, 1u);
}

// Frames first_frame + k * frame_stride (k < n_frames) folded into the image as samples first_sample + k.
static int trace_frames_impl(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t first_frame,
                             uint32_t n_frames, uint32_t frame_stride, uint32_t first_sample) {
    CTX_CHECK(ctx);
    if (ctx->children.size() > 1) {
        // A frame is ~60 launches per device and batch; one thread queueing them device after device makes the last device start
        // a few hundred microseconds after the first, which an end-to-end step (trace, then read back) pays in full.  One short-lived
        // launcher thread per device queues them side by side; the caller still sees one synchronous call.
        const size_t n = ctx->children.size();
        std::vector<int> rcs(n, RT_OK);
        std::vector<std::thread> launchers;
        launchers.reserve(n);
        for (size_t k = 0; k < n; k++)
            launchers.emplace_back([&, k] {
                rcs[k] = trace_frames_impl(ctx->children[k], camera_ubo, clear_color, first_frame, n_frames, frame_stride, first_sample);
            });
        for (std::thread& t : launchers) t.join();
        for (size_t k = 0; k < n; k++)
            if (rcs[k] != RT_OK) { ctx->error = ctx->children[k]->error; return rcs[k]; }
        return RT_OK;
    }
    FANOUT(ctx, trace_frames_impl(c, camera_ubo, clear_color, first_frame, n_frames, frame_stride, first_sample));
    if (!camera_ubo || !clear_color) { set_error("rt_trace_frame: null argument"); return RT_ERR_INVALID_ARGUMENT; }
    if (frame_stride == 0) { set_error("rt_trace_frames_sliced: frame_stride must be >= 1"); return RT_ERR_INVALID_ARGUMENT; }
    int rc = rt_flush_scene(ctx);
    if (rc != RT_OK) return rc;
    rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    // an accumulation restarts: the strips may change hands (the image is overwritten by frame 0 anyway)
    if (ctx->balance_pending && first_sample == 0u) {
        rc = balance_strips(ctx, camera_ubo, clear_color, first_frame);
        if (rc != RT_OK) return rc;
    }
    if (n_frames > 1u && ctx->lanes_for < ctx->cfg.frames_in_flight && ctx->cfg.pipeline == RT_PIPELINE_WAVEFRONT) {
        // first batched call: the path state grows from one frame in flight to the configured number
        RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        for (cudaStream_t st2 : ctx->streamx) if (st2) RT_CUDA_TRY(cudaStreamSynchronize(st2));
        rc = alloc_path_state(ctx, ctx->cfg.frames_in_flight);
        if (rc != RT_OK) return rc;
    }
    const SceneView sc = rt_scene_view(ctx);
    FrameResources res;
    res.dims = &ctx->dims;
    res.image = ctx->image.p;
    res.counters = ctx->counters.p;
    res.wf = ctx->wf;
    res.prof = &ctx->prof;
    res.launches = &ctx->launches;
    std::string err;
    rt_begin_timing(ctx);
    const bool mega = ctx->cfg.pipeline == RT_PIPELINE_MEGAKERNEL;
    // equal batches: 256 frames with room for 145 in flight run as 128 + 128, not 145 + 111
    const uint32_t n_batches = mega ? n_frames : (n_frames + ctx->lanes - 1u) / std::max(ctx->lanes, 1u);
    const uint32_t per_batch = mega ? 1u : (n_frames + std::max(n_batches, 1u) - 1u) / std::max(n_batches, 1u);
    // Several streams when there is more than one batch (and no per-kernel profiling, whose event pairs want the launches
    // serialised): batch b runs on stream b % RT_N_STREAMS with its own path state; only the accumulations are ordered
    // across streams.
    const bool multi = !mega && RT_N_STREAMS > 1 && ctx->wfx[0] && n_batches > 1 && !ctx->prof.enabled;
    const uint32_t n_str = multi ? std::min<uint32_t>(RT_N_STREAMS, n_batches) : 1u;
    auto cuda_fail = [&](cudaError_t e) {
        if (e == cudaSuccess) return false;
        set_error(std::string("rt_trace_frames: ") + cudaGetErrorString(e));
        return true;
    };
    if (multi) {
        if (cuda_fail(cudaEventRecord(ctx->ev_fork, ctx->stream))) return RT_ERR_CUDA;
        for (uint32_t k = 1; k < n_str; k++)
            if (cuda_fail(cudaStreamWaitEvent(ctx->streamx[k - 1], ctx->ev_fork, 0))) return RT_ERR_CUDA;
    }
    uint32_t b = 0;
    for (uint32_t f = 0; f < n_frames; b++) {
        FrameParams fp;
        rt_fill_frame_params(ctx, camera_ubo, clear_color, first_frame + f * frame_stride, &fp);
        fp.frame_stride = frame_stride;
        fp.sample = first_sample + f;
        fp.n_lanes = mega ? 1u : std::min(per_batch, n_frames - f);
        const uint32_t s = b % n_str;
        res.wf = s ? ctx->wfx[s - 1] : ctx->wf;
        res.wait_accumulate = (multi && b > 0) ? ctx->ev_acc[(b - 1u) % n_str] : nullptr;
        res.done_accumulate = multi ? ctx->ev_acc[s] : nullptr;
        cudaStream_t st = s ? ctx->streamx[s - 1] : ctx->stream;
        rc = mega ? launch_frame_megakernel(sc, fp, res, st, err) : launch_frame_wavefront(sc, fp, res, st, err);
        if (rc != RT_OK) { set_error(err); return rc; }
        f += fp.n_lanes;
    }
    if (multi) {  // the context's stream continues once every side stream has finished its batches
        for (uint32_t k = 1; k < n_str; k++)
            if (cuda_fail(cudaEventRecord(ctx->ev_join[k - 1], ctx->streamx[k - 1])) ||
                cuda_fail(cudaStreamWaitEvent(ctx->stream, ctx->ev_join[k - 1], 0)))
                return RT_ERR_CUDA;
    }
    ctx->frames_traced += n_frames;
    rt_end_timing(ctx);
    return RT_OK;
}

int rt_trace_frames(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t first_frame,
                    uint32_t n_frames) {
    return trace_frames_impl(ctx, camera_ubo, clear_color, first_frame, n_frames, 1u, first_frame);
}

int rt_trace_frames_sliced(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t first_frame,
                           uint32_t n_frames, uint32_t frame_stride, uint32_t first_sample) {
    return trace_frames_impl(ctx, camera_ubo, clear_color, first_frame, n_frames, frame_stride, first_sample);
}

int rt_trace_frame(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t frame_number) {
    return rt_trace_frames(ctx, camera_ubo, clear_color, frame_number, 1);
}

int rt_synchronize(rt_ctx* ctx) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_synchronize(c));
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

int rt_last_trace_ms(rt_ctx* ctx, float* ms) {
    CTX_CHECK(ctx);
    if (!ms) return RT_ERR_INVALID_ARGUMENT;
    if (!ctx->children.empty()) {  // the devices work side by side: the call took as long as the slowest
        float worst = 0.0f;
        for (rt_ctx* c : ctx->children) {
            float m = 0.0f;
            const int rc = rt_last_trace_ms(c, &m);
            if (rc != RT_OK) { ctx->error = c->error; return rc; }
            worst = std::max(worst, m);
        }
        *ms = worst;
        return RT_OK;
    }
    if (ctx->timing_pending) {
        RT_CUDA_TRY(cudaEventSynchronize(ctx->ev1));
        RT_CUDA_TRY(cudaEventElapsedTime(&ctx->last_ms, ctx->ev0, ctx->ev1));
        ctx->timing_pending = false;
    }
    *ms = ctx->last_ms;
    return RT_OK;
}

}  // extern "C"

// ---- rows between a context's compact strip image and a full-size host image --------------------------------------
// global strip of local band b
static inline uint32_t band_to_strip(const rt_ctx* ctx, uint32_t band) {
    return ctx->band_strip.empty() ? band * ctx->cfg.strip_count + ctx->cfg.strip_index : ctx->band_strip[band];
}

// One thread per 16 bytes: the rows of this context's compact image written straight into a full-size image in PINNED host
// memory (mapped into the device's address space).  One launch per device instead of one copy per band, and with several
// devices every PCIe link carries its own rows at the same time.
__global__ void scatter_rows_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, uint32_t row_words, uint32_t local_rows,
                                    uint32_t strip_rows, uint32_t strip_index, uint32_t strip_count, const uint32_t* band_strip) {
    const size_t n = (size_t)row_words * local_rows;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t lrow = (uint32_t)(i / row_words), w = (uint32_t)(i - (size_t)lrow * row_words);
        const uint32_t band = lrow / strip_rows, r = lrow - band * strip_rows;
        const uint32_t strip = band_strip ? band_strip[band] : band * strip_count + strip_index;
        dst[(size_t)(strip * strip_rows + r) * row_words + w] = src[i];
    }
}

// Queues the copies on the context's stream; the caller synchronises.  px_bytes is 4, 8, 16 or 32.
static int scatter_rows_async(rt_ctx* ctx, const void* d_src, void* dst, size_t bytes, size_t px_bytes) {
    const size_t need = (size_t)ctx->width * ctx->height * px_bytes;
    if (!dst || bytes < need) { set_error("readback: destination too small"); return RT_ERR_INVALID_ARGUMENT; }
    const size_t row = (size_t)ctx->width * px_bytes;
    if (ctx->local_rows == 0) return RT_OK;
    if (ctx->cfg.strip_count == 1) {
        RT_CUDA_TRY(cudaMemcpyAsync(dst, d_src, need, cudaMemcpyDeviceToHost, ctx->stream));
        return RT_OK;
    }
    cudaPointerAttributes attr;
    memset(&attr, 0, sizeof(attr));
    if (cudaPointerGetAttributes(&attr, dst) != cudaSuccess) { cudaGetLastError(); attr.type = cudaMemoryTypeUnregistered; }
    if (attr.type == cudaMemoryTypeHost && attr.devicePointer && (row % 16u) == 0u && ((uintptr_t)attr.devicePointer % 16u) == 0u) {
        const uint32_t row_words = (uint32_t)(row / 16u);
        const size_t n = (size_t)row_words * ctx->local_rows;
        const uint32_t grid = (uint32_t)std::min<size_t>((n + 255) / 256, 4096);
        scatter_rows_kernel<<<grid, 256, 0, ctx->stream>>>(reinterpret_cast<const uint4*>(d_src), reinterpret_cast<uint4*>(attr.devicePointer),
                                                           row_words, ctx->local_rows, ctx->cfg.strip_rows, ctx->cfg.strip_index,
                                                           ctx->cfg.strip_count, ctx->band_strip.empty() ? nullptr : ctx->d_band_strip.p);
        RT_CUDA_TRY(cudaGetLastError());
        ctx->launches++;
        return RT_OK;
    }
    const uint32_t sr = ctx->cfg.strip_rows;
    if (ctx->band_strip.empty()) {
        // round robin: the bands sit at a constant pitch in the full image — one 2-D copy for the full bands, one for the rest
        const uint32_t full = ctx->local_rows / sr, rest = ctx->local_rows - full * sr;
        const uint32_t y0 = ctx->cfg.strip_index * sr;
        if (full)
            RT_CUDA_TRY(cudaMemcpy2DAsync((char*)dst + (size_t)y0 * row, (size_t)ctx->cfg.strip_count * sr * row, d_src, (size_t)sr * row,
                                          (size_t)sr * row, full, cudaMemcpyDeviceToHost, ctx->stream));
        if (rest)
            RT_CUDA_TRY(cudaMemcpyAsync((char*)dst + (size_t)(band_to_strip(ctx, full) * sr) * row, (const char*)d_src + (size_t)full * sr * row,
                                        (size_t)rest * row, cudaMemcpyDeviceToHost, ctx->stream));
        return RT_OK;
    }
    for (uint32_t lrow = 0; lrow < ctx->local_rows; lrow += sr) {  // balanced assignment: one copy per band
        const uint32_t rows = std::min(sr, ctx->local_rows - lrow);
        const uint32_t y = band_to_strip(ctx, lrow / sr) * sr;
        RT_CUDA_TRY(cudaMemcpyAsync((char*)dst + (size_t)y * row, (const char*)d_src + (size_t)lrow * row, rows * row,
                                    cudaMemcpyDeviceToHost, ctx->stream));
    }
    return RT_OK;
}

static int scatter_rows_to_host(rt_ctx* ctx, const void* d_src, void* dst, size_t bytes, size_t px_bytes) {
    const int rc = scatter_rows_async(ctx, d_src, dst, bytes, px_bytes);
    if (rc != RT_OK) return rc;
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

// the devices of a multi-GPU context copy their rows side by side: queue everything, then wait for everyone
template <typename Queue>
static int gather_children(rt_ctx* ctx, Queue queue) {
    for (rt_ctx* c : ctx->children) {
        cudaSetDevice(c->device);
        const int rc = queue(c);
        if (rc != RT_OK) { ctx->error = c->error; return rc; }
    }
    for (rt_ctx* c : ctx->children) {
        cudaSetDevice(c->device);
        const cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) { ctx->error = std::string("readback: ") + cudaGetErrorString(e); return RT_ERR_CUDA; }
    }
    return RT_OK;
}

extern "C" {

int rt_readback(rt_ctx* ctx, float* dst, size_t bytes) {
    CTX_CHECK(ctx);
    if (!ctx->children.empty())
        return gather_children(ctx, [&](rt_ctx* c) {
            const int rc = rt_require_target(c);
            return rc != RT_OK ? rc : scatter_rows_async(c, c->image.p, dst, bytes, sizeof(float4));
        });
    const int rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    return scatter_rows_to_host(ctx, ctx->image.p, dst, bytes, sizeof(float4));
}

static int srgb8_async(rt_ctx* ctx, uint8_t* dst, size_t bytes) {
    int rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    const size_t n = (size_t)ctx->local_rows * ctx->width;
    RT_CUDA_TRY(ctx->srgb.ensure(n));
    std::string err;
    rc = launch_srgb8(ctx->image.p, ctx->srgb.p, (uint32_t)n, ctx->dims, ctx->stream, err);
    if (rc != RT_OK) { set_error(err); return rc; }
    return scatter_rows_async(ctx, ctx->srgb.p, dst, bytes, sizeof(uchar4));
}

int rt_readback_srgb8(rt_ctx* ctx, uint8_t* dst, size_t bytes) {
    CTX_CHECK(ctx);
    if (!ctx->children.empty()) return gather_children(ctx, [&](rt_ctx* c) { return srgb8_async(c, dst, bytes); });
    const int rc = srgb8_async(ctx, dst, bytes);
    if (rc != RT_OK) return rc;
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

int rt_device_image(rt_ctx* ctx, void** device_ptr, uint32_t* local_rows) {
    CTX_CHECK(ctx);
    if (!ctx->children.empty()) { set_error("rt_device_image: a multi-GPU context has one image per device; use rt_readback"); return RT_ERR_INVALID_ARGUMENT; }
    const int rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    if (device_ptr) *device_ptr = ctx->image.p;
    if (local_rows) *local_rows = ctx->local_rows;
    return RT_OK;
}

int rt_upload_accum(rt_ctx* ctx, const float* src, size_t bytes) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_upload_accum(c, src, bytes));
    const int rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    const size_t row = (size_t)ctx->width * sizeof(float4);
    if (!src || bytes < row * ctx->height) { set_error("rt_upload_accum: source too small"); return RT_ERR_INVALID_ARGUMENT; }
    const uint32_t sr = ctx->cfg.strip_rows;
    for (uint32_t lrow = 0; lrow < ctx->local_rows; lrow += sr) {
        const uint32_t rows = std::min(sr, ctx->local_rows - lrow);
        const uint32_t y = band_to_strip(ctx, lrow / sr) * sr;
        RT_CUDA_TRY(cudaMemcpyAsync((char*)ctx->image.p + (size_t)lrow * row, (const char*)src + (size_t)y * row, rows * row,
                                    cudaMemcpyHostToDevice, ctx->stream));
    }
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->balance_pending = false;  // the rows now hold a running mean: they stay where they are
    return RT_OK;
}

// The strip assignment in force: owners[s] = rank (strip_index, or device slot of a multi-GPU context) that renders strip s.
int rt_get_strip_owners(rt_ctx* ctx, uint32_t* owners, uint32_t capacity, uint32_t* n_strips) {
    CTX_CHECK(ctx);
    rt_ctx* c0 = ctx->children.empty() ? ctx : ctx->children[0];
    const int rc = rt_require_target(c0);
    if (rc != RT_OK) { ctx->error = c0->error; return rc; }
    const uint32_t sr = c0->cfg.strip_rows, n = (c0->height + sr - 1) / sr;
    if (n_strips) *n_strips = n;
    if (!owners) return RT_OK;
    if (capacity < n) { set_error("rt_get_strip_owners: capacity too small"); return RT_ERR_INVALID_ARGUMENT; }
    if (c0->strip_owner.empty()) for (uint32_t s2 = 0; s2 < n; s2++) owners[s2] = s2 % c0->cfg.strip_count;
    else memcpy(owners, c0->strip_owner.data(), n * sizeof(uint32_t));
    return RT_OK;
}

// The calibration pass on its own (tools/fit_balance.py): four counters per image row, see row_cost_kernel.
int rt_measure_row_costs(rt_ctx* ctx, const float camera_ubo[64], uint32_t frame_number, uint32_t* out4, uint32_t capacity_rows) {
    CTX_CHECK(ctx);
    if (!ctx->children.empty()) {
        const int rc0 = rt_measure_row_costs(ctx->children[0], camera_ubo, frame_number, out4, capacity_rows);
        if (rc0 != RT_OK) ctx->error = ctx->children[0]->error;
        return rc0;
    }
    if (!camera_ubo || !out4) return RT_ERR_INVALID_ARGUMENT;
    int rc = rt_flush_scene(ctx);
    if (rc != RT_OK) return rc;
    rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    if (capacity_rows < ctx->height) { set_error("rt_measure_row_costs: capacity too small"); return RT_ERR_INVALID_ARGUMENT; }
    const float clear[3] = {0.0f, 0.0f, 0.0f};
    std::vector<uint32_t> rows;
    rc = measure_row_costs(ctx, camera_ubo, clear, frame_number, rows);
    if (rc != RT_OK) return rc;
    memcpy(out4, rows.data(), rows.size() * sizeof(uint32_t));
    return RT_OK;
}

// A strip -> rank table decided by the HOST (e.g. from measured per-rank step times: raytracer_b200.multi_gpu.StripBalancer).
// Every rank must keep the number of rows the round-robin deal gives it.  The image rows change hands: the accumulation
// must restart (frame_number 0) after this call; a pending cost calibration is cancelled.
int rt_set_strip_owners(rt_ctx* ctx, const uint32_t* owners, uint32_t n_strips) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_set_strip_owners(c, owners, n_strips));
    const int rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    const uint32_t sr = ctx->cfg.strip_rows;
    if (!owners || n_strips != (ctx->height + sr - 1) / sr) { set_error("rt_set_strip_owners: one owner per strip"); return RT_ERR_INVALID_ARGUMENT; }
    for (uint32_t st = 0; st < n_strips; st++)
        if (owners[st] >= ctx->cfg.strip_count) { set_error("rt_set_strip_owners: owner out of range"); return RT_ERR_INDEX_RANGE; }
    ctx->balance_pending = false;
    return install_owners(ctx, owners, n_strips);
}

// Ask for a new cost calibration at the next accumulation restart (e.g. after the camera settled somewhere else).
int rt_rebalance(rt_ctx* ctx) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_rebalance(c));
    ctx->balance_pending = ctx->cfg.strip_balance == RT_BALANCE_COST && ctx->cfg.strip_count > 1 && ctx->width != 0;
    return RT_OK;
}

// ---- G-buffer --------------------------------------------------------------------------------------
int rt_trace_gbuffer(rt_ctx* ctx, const float camera_ubo[64]) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_trace_gbuffer(c, camera_ubo));
    if (!camera_ubo) return RT_ERR_INVALID_ARGUMENT;
    int rc = rt_flush_scene(ctx);
    if (rc != RT_OK) return rc;
    rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    const size_t n = (size_t)ctx->local_rows * ctx->width;
    RT_CUDA_TRY(ctx->g_albedo.ensure(n));
    RT_CUDA_TRY(ctx->g_output.ensure(n));
    RT_CUDA_TRY(ctx->g_normal.ensure(n));
    RT_CUDA_TRY(ctx->g_emission.ensure(n));
    RT_CUDA_TRY(ctx->g_position.ensure(n));
    FrameParams fp;
    rt_fill_frame_params(ctx, camera_ubo, nullptr, 0, &fp);
    GBufferImages g;
    g.albedo = ctx->g_albedo.p; g.output = ctx->g_output.p; g.normal_rough = ctx->g_normal.p;
    g.emission = ctx->g_emission.p; g.position_metal = ctx->g_position.p;
    std::string err;
    rt_begin_timing(ctx);
    if (!ctx->dims.num_sms) cudaDeviceGetAttribute(&ctx->dims.num_sms, cudaDevAttrMultiProcessorCount, ctx->device);
    rc = launch_gbuffer(rt_scene_view(ctx), fp, g, (ctx->cfg.flags & RT_FLAG_COUNTERS) ? ctx->counters.p : nullptr, ctx->dims.num_sms,
                        ctx->stream, err);
    rt_end_timing(ctx);
    if (rc != RT_OK) { set_error(err); return rc; }
    ctx->has_gbuffer = true;
    return RT_OK;
}

int rt_readback_gbuffer(rt_ctx* ctx, uint8_t* albedo, uint16_t* normal_rough, float* position_metal, uint16_t* emission,
                        uint8_t* output) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_readback_gbuffer(c, albedo, normal_rough, position_metal, emission, output));
    if (!ctx->has_gbuffer) { set_error("rt_readback_gbuffer: no G-buffer traced"); return RT_ERR_NO_TARGET; }
    const size_t px = (size_t)ctx->width * ctx->height;
    int rc = RT_OK;
    if (albedo && rc == RT_OK) rc = scatter_rows_to_host(ctx, ctx->g_albedo.p, albedo, px * 4, 4);
    if (normal_rough && rc == RT_OK) rc = scatter_rows_to_host(ctx, ctx->g_normal.p, normal_rough, px * 8, 8);
    if (position_metal && rc == RT_OK) rc = scatter_rows_to_host(ctx, ctx->g_position.p, position_metal, px * 16, 16);
    if (emission && rc == RT_OK) rc = scatter_rows_to_host(ctx, ctx->g_emission.p, emission, px * 8, 8);
    if (output && rc == RT_OK) rc = scatter_rows_to_host(ctx, ctx->g_output.p, output, px * 4, 4);
    return rc;
}

// ---- instrumentation -------------------------------------------------------------------------------
int rt_get_counters(rt_ctx* ctx, rt_counters* out) {
    CTX_CHECK(ctx);
    if (!out) return RT_ERR_INVALID_ARGUMENT;
    if (!ctx->children.empty()) {  // sums over the devices; `frames` is the number of calls, the same on each
        memset(out, 0, sizeof(*out));
        for (rt_ctx* c : ctx->children) {
            rt_counters k;
            const int rc = rt_get_counters(c, &k);
            if (rc != RT_OK) { ctx->error = c->error; return rc; }
            out->extend_rays += k.extend_rays; out->shadow_rays += k.shadow_rays; out->nodes += k.nodes; out->tris += k.tris;
            out->hits += k.hits; out->nee += k.nee; out->shadow_nodes += k.shadow_nodes; out->shadow_tris += k.shadow_tris;
            out->launches += k.launches; out->frames = k.frames;
        }
        return RT_OK;
    }
    DevCounters c;
    RT_CUDA_TRY(cudaMemcpyAsync(&c, ctx->counters.p, sizeof(c), cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    out->extend_rays = c.extend_rays; out->shadow_rays = c.shadow_rays; out->nodes = c.nodes; out->tris = c.tris;
    out->hits = c.hits; out->nee = c.nee; out->frames = ctx->frames_traced;
    out->shadow_nodes = c.shadow_nodes; out->shadow_tris = c.shadow_tris; out->launches = ctx->launches;
    return RT_OK;
}

int rt_reset_counters(rt_ctx* ctx) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_reset_counters(c));
    RT_CUDA_TRY(cudaMemsetAsync(ctx->counters.p, 0, sizeof(DevCounters), ctx->stream));
    ctx->frames_traced = 0;
    ctx->launches = 0;
    return RT_OK;
}

int rt_event_record(rt_ctx* ctx, int which) {
    CTX_CHECK(ctx);
    if (which < 0 || which > 1) return RT_ERR_INVALID_ARGUMENT;
    FANOUT(ctx, rt_event_record(c, which));
    if (!ctx->user_ev[which]) RT_CUDA_TRY(cudaEventCreate(&ctx->user_ev[which]));
    RT_CUDA_TRY(cudaEventRecord(ctx->user_ev[which], ctx->stream));
    return RT_OK;
}

int rt_event_elapsed_ms(rt_ctx* ctx, float* ms) {
    CTX_CHECK(ctx);
    if (ms && !ctx->children.empty()) {  // max over the devices, each timed on its own stream
        float worst = 0.0f;
        for (rt_ctx* c : ctx->children) {
            float m = 0.0f;
            const int rc = rt_event_elapsed_ms(c, &m);
            if (rc != RT_OK) { ctx->error = c->error; return rc; }
            worst = std::max(worst, m);
        }
        *ms = worst;
        return RT_OK;
    }
    if (!ms || !ctx->user_ev[0] || !ctx->user_ev[1]) { set_error("rt_event_elapsed_ms: record both events first"); return RT_ERR_INVALID_ARGUMENT; }
    RT_CUDA_TRY(cudaEventSynchronize(ctx->user_ev[1]));
    RT_CUDA_TRY(cudaEventElapsedTime(ms, ctx->user_ev[0], ctx->user_ev[1]));
    return RT_OK;
}

// the same event pair per device of a multi-GPU context (a single-device context has one entry): what a host-side strip
// balancer feeds on
int rt_event_elapsed_per_device(rt_ctx* ctx, float* ms, uint32_t capacity, uint32_t* n_devices) {
    CTX_CHECK(ctx);
    const uint32_t n = ctx->children.empty() ? 1u : (uint32_t)ctx->children.size();
    if (n_devices) *n_devices = n;
    if (!ms) return RT_OK;
    if (capacity < n) { set_error("rt_event_elapsed_per_device: capacity too small"); return RT_ERR_INVALID_ARGUMENT; }
    if (ctx->children.empty()) return rt_event_elapsed_ms(ctx, ms);
    for (uint32_t k = 0; k < n; k++) {
        const int rc = rt_event_elapsed_ms(ctx->children[k], ms + k);
        if (rc != RT_OK) { ctx->error = ctx->children[k]->error; return rc; }
    }
    return RT_OK;
}

int rt_set_profiling(rt_ctx* ctx, int enable) {
    CTX_CHECK(ctx);
    FANOUT(ctx, rt_set_profiling(c, enable));
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->prof.enabled = enable != 0;
    ctx->prof.reset();
    return RT_OK;
}

int rt_get_kernel_profile(rt_ctx* ctx, rt_kernel_profile* out) {
    CTX_CHECK(ctx);
    if (!out) return RT_ERR_INVALID_ARGUMENT;
    memset(out, 0, sizeof(*out));
    if (!ctx->children.empty()) {  // device time summed over the devices
        for (rt_ctx* c : ctx->children) {
            rt_kernel_profile k;
            const int rc = rt_get_kernel_profile(c, &k);
            if (rc != RT_OK) { ctx->error = c->error; return rc; }
            for (int i = 0; i < RT_K_COUNT; i++) { out->ms[i] += k.ms[i]; out->launches[i] += k.launches[i]; }
        }
        return RT_OK;
    }
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < ctx->prof.used; i++) {
        float ms = 0.0f;
        RT_CUDA_TRY(cudaEventElapsedTime(&ms, ctx->prof.spans[i].a, ctx->prof.spans[i].b));
        out->ms[ctx->prof.spans[i].k] += ms;
        out->launches[ctx->prof.spans[i].k]++;
    }
    return RT_OK;
}

// ---- verification hooks ----------------------------------------------------------------------------
static int trace_rays_common(rt_ctx* ctx, const float* rays, uint32_t n, bool any, const uint32_t* ignore, float* out_tuv,
                             uint32_t* out_id) {
    int rc = rt_flush_scene(ctx);
    if (rc != RT_OK) return rc;
    if (n == 0) return RT_OK;
    if (!ctx->dims.num_sms) cudaDeviceGetAttribute(&ctx->dims.num_sms, cudaDevAttrMultiProcessorCount, ctx->device);
    if (!rays || !out_id || (!any && !out_tuv)) { set_error("rt_trace_rays: null argument"); return RT_ERR_INVALID_ARGUMENT; }
    DevBuf<float> d_rays, d_tuv;
    DevBuf<uint32_t> d_id, d_ig;
    RT_CUDA_TRY(d_rays.alloc((size_t)n * 8));
    RT_CUDA_TRY(d_id.alloc(n));
    RT_CUDA_TRY(d_tuv.alloc((size_t)n * 3));
    RT_CUDA_TRY(cudaMemcpyAsync(d_rays.p, rays, (size_t)n * 32, cudaMemcpyHostToDevice, ctx->stream));
    if (any && ignore) {
        RT_CUDA_TRY(d_ig.alloc(n));
        RT_CUDA_TRY(cudaMemcpyAsync(d_ig.p, ignore, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    std::string err;
    rc = launch_trace_rays(rt_scene_view(ctx), d_rays.p, n, any, (any && ignore) ? d_ig.p : nullptr, d_tuv.p, d_id.p,
                           (ctx->cfg.flags & RT_FLAG_COUNTERS) ? ctx->counters.p : nullptr, ctx->dims.num_sms, ctx->stream, err);
    if (rc != RT_OK) { set_error(err); return rc; }
    RT_CUDA_TRY(cudaMemcpyAsync(out_id, d_id.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (!any) RT_CUDA_TRY(cudaMemcpyAsync(out_tuv, d_tuv.p, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

// the verification hooks of a multi-GPU context run on its first device (every device holds the same BVH)
#define FIRST_CHILD(ctx, call)                                  \
    if (!(ctx)->children.empty()) {                             \
        rt_ctx* c = (ctx)->children[0];                         \
        const int _rc = (call);                                 \
        if (_rc != RT_OK) (ctx)->error = c->error;              \
        return _rc;                                             \
    }

int rt_trace_rays_closest(rt_ctx* ctx, const float* rays, uint32_t n, float* out_tuv, uint32_t* out_id) {
    CTX_CHECK(ctx);
    FIRST_CHILD(ctx, rt_trace_rays_closest(c, rays, n, out_tuv, out_id));
    return trace_rays_common(ctx, rays, n, false, nullptr, out_tuv, out_id);
}

int rt_trace_rays_any(rt_ctx* ctx, const float* rays, uint32_t n, const uint32_t* ignore_inst, uint32_t* out_occluded) {
    CTX_CHECK(ctx);
    FIRST_CHILD(ctx, rt_trace_rays_any(c, rays, n, ignore_inst, out_occluded));
    return trace_rays_common(ctx, rays, n, true, ignore_inst, nullptr, out_occluded);
}

int rt_get_bvh(rt_ctx* ctx, void* nodes, size_t nodes_bytes, void* tris, size_t tris_bytes, uint64_t* morton,
               uint32_t* sorted_ids) {
    CTX_CHECK(ctx);
    FIRST_CHILD(ctx, rt_get_bvh(c, nodes, nodes_bytes, tris, tris_bytes, morton, sorted_ids));
    const int rc = rt_flush_scene(ctx);
    if (rc != RT_OK) return rc;
    const size_t nb = (size_t)ctx->bvh.n_nodes * sizeof(Node8), tb = (size_t)ctx->n_tris * 48;
    if ((nodes && nodes_bytes < nb) || (tris && tris_bytes < tb)) { set_error("rt_get_bvh: buffer too small"); return RT_ERR_INVALID_ARGUMENT; }
    if (nodes) RT_CUDA_TRY(cudaMemcpyAsync(nodes, ctx->bvh.nodes.p, nb, cudaMemcpyDeviceToHost, ctx->stream));
    if (tris && tb) RT_CUDA_TRY(cudaMemcpyAsync(tris, ctx->bvh.tris.p, tb, cudaMemcpyDeviceToHost, ctx->stream));
    if (morton && ctx->n_tris) RT_CUDA_TRY(cudaMemcpyAsync(morton, ctx->bvh.morton.p, (size_t)ctx->n_tris * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (sorted_ids && ctx->n_tris) RT_CUDA_TRY(cudaMemcpyAsync(sorted_ids, ctx->bvh.sorted_ids.p, (size_t)ctx->n_tris * 4, cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    if (nodes) {  // device nodes keep their 32-byte chunks rotated by the node index (rt_types.cuh); callers get struct Node8
        Node8* hn = static_cast<Node8*>(nodes);
        for (uint32_t k = 0; k < ctx->bvh.n_nodes; k++) {
            Node8 raw = hn[k];
            const uint4* src = reinterpret_cast<const uint4*>(&raw);
            uint4* dst = reinterpret_cast<uint4*>(&hn[k]);
            for (uint32_t c = 0; c < 4u; c++) {
                const uint32_t at = node_chunk(k, c);
                dst[2u * c] = src[at];
                dst[2u * c + 1u] = src[at + 1u];
            }
        }
    }
    return RT_OK;
}

}  // extern "C"
