// rt_raster.cu — the rasterised G-buffer variant: shaders/restir/gbuffer.vert:22-37 + gbuffer.frag:26-39 as CUDA.
//
// The reference ships a vertex / fragment pair that fills the same kind of G-buffer as restir.rgen by rasterisation (SURVEY §8f
// row 4).  No host code draws with it, so the fixed-function state around it (formats, clear values, depth test, culling) is not
// recorded anywhere; this file states one:
//   vertex stage    gbuffer.vert as written: position and normal are mirrored in y BEFORE the model matrix (flippedPos,
//                   flippedNormal), world position = model * pos, gl_Position = (projection * view) * world position (GLSL
//                   multiplies left to right: the matrix product first), world normal = normalize(mat3(normal_matrix) * normal)
//                   with normal_matrix = transpose(inverse(model)) (the push constant's name; the matrix the ray-traced path uses)
//   rasterisation   homogeneous edge functions (Olano & Greer 1997) over clip-space (x, y, w), sampled at pixel centres
//                   (px + 0.5, py + 0.5) of a W x H viewport; no culling (both windings); a sample is covered when the three
//                   edge functions share a sign (zero counts as either).  lambda_i = E_i / (E_0 + E_1 + E_2) are the
//                   perspective-correct barycentrics; samples with interpolated w <= 0 or depth z/w outside [0, 1] are dropped
//                   (Vulkan clip volume, no depth clamp).  Depth test LESS; equal depths go to the lowest flat triangle id, so
//                   the result does not depend on the order triangles are drawn in.
//   fragment stage  gbuffer.frag as written, five float4 attachments: albedo|1, normalize(normal)|metallic, position|roughness,
//                   emission_power * emission_color|isEmissive, roughness|metallic|isEmissive|0; pixels no triangle covers are 0.
// The CPU twin in oracle/oracle_render.cpp states the same arithmetic in the same order (bit-exact: tests/test_gpu_parity.py), and
// tools/spirv_golden.py runs gbuffer_vert.spv / gbuffer_frag.spv themselves for the two programmable stages (tests/test_oracle.py).
#include <string.h>

#include <algorithm>

#include "rt_internal.h"
#include "rt_shade.cuh"

namespace rt {

struct RasterVertex {
    vec3 wp, wn;        // vertex-stage outputs: world position, world normal (normalised)
    float cx, cy, cz, cw;  // gl_Position
};

// (projection * view), column-major, every element ((p0 v0 + p1 v1) + p2 v2) + p3 v3
__device__ __forceinline__ void proj_view(const float* P, const float* V, float* PV) {
#pragma unroll
    for (int c = 0; c < 4; c++)
#pragma unroll
        for (int r = 0; r < 4; r++)
            PV[c * 4 + r] = ((P[0 * 4 + r] * V[c * 4 + 0] + P[1 * 4 + r] * V[c * 4 + 1]) + P[2 * 4 + r] * V[c * 4 + 2]) + P[3 * 4 + r] * V[c * 4 + 3];
}

__device__ __forceinline__ RasterVertex raster_vertex(const DevObject* o, const rt_vertex* v, const float* PV) {
    RasterVertex out;
    // gbuffer.vert:24-25: mirrored in y
    const float px = v->pos[0], py = -v->pos[1], pz = v->pos[2];
    const float nx = v->normal[0], ny = -v->normal[1], nz = v->normal[2];
    const float* m = o->model;  // three rows of four columns: model[c * 3 + r]
    out.wp = mk3(((m[0] * px + m[3] * py) + m[6] * pz) + m[9], ((m[1] * px + m[4] * py) + m[7] * pz) + m[10],
                 ((m[2] * px + m[5] * py) + m[8] * pz) + m[11]);
    out.wn = normalize3(mat3_mul(o->nrm, mk3(nx, ny, nz)));
    const float x = out.wp.x, y = out.wp.y, z = out.wp.z;
    out.cx = ((PV[0] * x + PV[4] * y) + PV[8] * z) + PV[12];
    out.cy = ((PV[1] * x + PV[5] * y) + PV[9] * z) + PV[13];
    out.cz = ((PV[2] * x + PV[6] * y) + PV[10] * z) + PV[14];
    out.cw = ((PV[3] * x + PV[7] * y) + PV[11] * z) + PV[15];
    return out;
}

struct EdgeSetup { float a[3], b[3], c[3]; };

__device__ __forceinline__ EdgeSetup edge_setup(const RasterVertex* v) {
    EdgeSetup e;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const RasterVertex& p = v[(i + 1) % 3];
        const RasterVertex& q = v[(i + 2) % 3];
        e.a[i] = p.cy * q.cw - p.cw * q.cy;
        e.b[i] = p.cw * q.cx - p.cx * q.cw;
        e.c[i] = p.cx * q.cy - p.cy * q.cx;
    }
    return e;
}

// covered?  lambda: perspective-correct barycentrics; depth = z / w
__device__ __forceinline__ bool raster_sample(const EdgeSetup& e, const RasterVertex* v, uint32_t px, uint32_t py, uint32_t W, uint32_t H,
                                              float* lambda, float* depth) {
    const float X = (((float)px + 0.5f) / (float)W) * 2.0f - 1.0f;
    const float Y = (((float)py + 0.5f) / (float)H) * 2.0f - 1.0f;
    const float E0 = (e.a[0] * X + e.b[0] * Y) + e.c[0];
    const float E1 = (e.a[1] * X + e.b[1] * Y) + e.c[1];
    const float E2 = (e.a[2] * X + e.b[2] * Y) + e.c[2];
    if (!((E0 >= 0.0f && E1 >= 0.0f && E2 >= 0.0f) || (E0 <= 0.0f && E1 <= 0.0f && E2 <= 0.0f))) return false;
    const float S = (E0 + E1) + E2;
    if (!(S != 0.0f)) return false;  // also rejects NaN
    lambda[0] = E0 / S; lambda[1] = E1 / S; lambda[2] = E2 / S;
    const float w = (lambda[0] * v[0].cw + lambda[1] * v[1].cw) + lambda[2] * v[2].cw;
    if (!(w > 0.0f)) return false;
    const float z = (lambda[0] * v[0].cz + lambda[1] * v[1].cz) + lambda[2] * v[2].cz;
    const float d = z / w;
    if (!(d >= 0.0f && d <= 1.0f)) return false;
    *depth = d;
    return true;
}

__device__ __forceinline__ void flat_to_object(const SceneView& sc, const uint32_t* tri_offsets, uint32_t i, uint32_t* obj, uint32_t* prim) {
    uint32_t a = 0, b = sc.n_objects;  // last object with tri_offsets[obj] <= i
    while (b - a > 1) {
        const uint32_t m = (a + b) >> 1;
        if (tri_offsets[m] <= i) a = m; else b = m;
    }
    *obj = a;
    *prim = i - tri_offsets[a];
}

__device__ __forceinline__ void load_raster_triangle(const SceneView& sc, const uint32_t* tri_offsets, uint32_t flat, const float* PV,
                                                     RasterVertex* v, uint32_t* obj_out) {
    uint32_t obj, prim;
    flat_to_object(sc, tri_offsets, flat, &obj, &prim);
    const DevObject* o = sc.objects + obj;
    const uint32_t* ip = sc.indices + o->index_base + 3u * prim;
    const rt_vertex* vb = sc.vertices + o->vertex_base;
#pragma unroll
    for (int k = 0; k < 3; k++) v[k] = raster_vertex(o, vb + ip[k], PV);
    *obj_out = obj;
}

// One warp per triangle: the lanes walk the pixels of its screen bounding box.  vis[pixel] = min(depth bits << 32 | flat id).
__global__ void __launch_bounds__(128) raster_visibility_kernel(SceneView sc, const uint32_t* tri_offsets, FrameParams fp, const float* cam,
                                                                unsigned long long* vis) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
    float PV[16];
    proj_view(cam, cam + 16, PV);
    for (uint32_t flat = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; flat < sc.n_tris; flat += warps) {
        RasterVertex v[3];
        uint32_t obj;
        load_raster_triangle(sc, tri_offsets, flat, PV, v, &obj);
        // conservative screen box: the projected vertices when all are in front of the eye (one pixel of slack), else the
        // part of the triangle with w >= eps clipped edge by edge; a triangle entirely behind is skipped
        int x0 = 0, y0 = 0, x1 = (int)fp.width - 1, y1 = (int)fp.height - 1;
        {
            const float eps = 1e-6f;
            float lo_x = INFINITY, hi_x = -INFINITY, lo_y = INFINITY, hi_y = -INFINITY;
            int n_front = 0;
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const RasterVertex& p = v[k];
                const RasterVertex& q = v[(k + 1) % 3];
                if (p.cw >= eps) {
                    n_front++;
                    const float sx = (p.cx / p.cw * 0.5f + 0.5f) * (float)fp.width, sy = (p.cy / p.cw * 0.5f + 0.5f) * (float)fp.height;
                    lo_x = fminf(lo_x, sx); hi_x = fmaxf(hi_x, sx); lo_y = fminf(lo_y, sy); hi_y = fmaxf(hi_y, sy);
                }
                if ((p.cw >= eps) != (q.cw >= eps)) {  // the edge crosses w = eps
                    const float t = (eps - p.cw) / (q.cw - p.cw);
                    const float ex = p.cx + t * (q.cx - p.cx), ey = p.cy + t * (q.cy - p.cy);
                    const float sx = (ex / eps * 0.5f + 0.5f) * (float)fp.width, sy = (ey / eps * 0.5f + 0.5f) * (float)fp.height;
                    lo_x = fminf(lo_x, sx); hi_x = fmaxf(hi_x, sx); lo_y = fminf(lo_y, sy); hi_y = fmaxf(hi_y, sy);
                }
            }
            if (n_front == 0) continue;
            if (lo_x == lo_x && hi_x == hi_x && lo_y == lo_y && hi_y == hi_y) {  // NaN anywhere: keep the whole viewport
                x0 = max(x0, (int)fmaxf(fminf(floorf(lo_x) - 1.0f, 1e9f), -1e9f));
                y0 = max(y0, (int)fmaxf(fminf(floorf(lo_y) - 1.0f, 1e9f), -1e9f));
                x1 = min(x1, (int)fmaxf(fminf(ceilf(hi_x) + 1.0f, 1e9f), -1e9f));
                y1 = min(y1, (int)fmaxf(fminf(ceilf(hi_y) + 1.0f, 1e9f), -1e9f));
            }
            if (x0 > x1 || y0 > y1) continue;
        }
        const EdgeSetup e = edge_setup(v);
        const uint32_t bw = (uint32_t)(x1 - x0 + 1), n = bw * (uint32_t)(y1 - y0 + 1);
        for (uint32_t i = lane; i < n; i += 32u) {
            const uint32_t py = (uint32_t)y0 + i / bw, px = (uint32_t)x0 + i % bw;
            float lambda[3], depth;
            if (raster_sample(e, v, px, py, fp.width, fp.height, lambda, &depth))
                atomicMin(&vis[(size_t)py * fp.width + px], ((unsigned long long)__float_as_uint(depth) << 32) | flat);
        }
    }
}

struct RasterImages { float4 *albedo, *normal_metal, *position_rough, *emission, *props; };

// One thread per pixel: the winning triangle's vertex stage again, the interpolated varyings, gbuffer.frag:26-39
__global__ void __launch_bounds__(128) raster_resolve_kernel(SceneView sc, const uint32_t* tri_offsets, FrameParams fp, const float* cam,
                                                             const unsigned long long* vis, RasterImages out) {
    const uint32_t n_px = fp.width * fp.height;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_px) return;
    const unsigned long long key = vis[i];
    float4 o0 = make_float4(0.f, 0.f, 0.f, 0.f), o1 = o0, o2 = o0, o3 = o0, o4 = o0;
    if (key != ~0ull) {
        float PV[16];
        proj_view(cam, cam + 16, PV);
        RasterVertex v[3];
        uint32_t obj;
        load_raster_triangle(sc, tri_offsets, (uint32_t)(key & 0xFFFFFFFFull), PV, v, &obj);
        const EdgeSetup e = edge_setup(v);
        float l[3], depth;
        raster_sample(e, v, i % fp.width, i / fp.width, fp.width, fp.height, l, &depth);  // covered: the visibility pass said so
        const vec3 in_pos = add3(add3(scale3(v[0].wp, l[0]), scale3(v[1].wp, l[1])), scale3(v[2].wp, l[2]));
        const vec3 in_nrm = add3(add3(scale3(v[0].wn, l[0]), scale3(v[1].wn, l[1])), scale3(v[2].wn, l[2]));
        const Mat mat = load_material(sc.materials + sc.objects[obj].material_index);
        const vec3 n = normalize3(in_nrm);
        const float emissive = mat.emission_power > 0.0f ? 1.0f : 0.0f;
        o0 = make_float4(mat.albedo.x, mat.albedo.y, mat.albedo.z, 1.0f);
        o1 = make_float4(n.x, n.y, n.z, mat.metallic);
        o2 = make_float4(in_pos.x, in_pos.y, in_pos.z, mat.roughness);
        o3 = make_float4(mat.emission_power * mat.emission_color.x, mat.emission_power * mat.emission_color.y,
                         mat.emission_power * mat.emission_color.z, emissive);
        o4 = make_float4(mat.roughness, mat.metallic, emissive, 0.0f);
    }
    out.albedo[i] = o0; out.normal_metal[i] = o1; out.position_rough[i] = o2; out.emission[i] = o3; out.props[i] = o4;
}

}  // namespace rt

#define set_error ctx->set_error

extern "C" {

// The whole frame on the context's device (a multi-GPU context uses its first device: this is a sub-millisecond primary-visibility
// pass, not worth partitioning).  Results stay on the device until rt_readback_raster_gbuffer.
int rt_raster_gbuffer(rt_ctx* ctx, const float camera_ubo[64]) {
    if (!ctx) return RT_ERR_INVALID_ARGUMENT;
    if (!ctx->children.empty()) {
        const int rc0 = rt_raster_gbuffer(ctx->children[0], camera_ubo);
        if (rc0 != RT_OK) ctx->error = ctx->children[0]->error;
        return rc0;
    }
    cudaSetDevice(ctx->device);
    if (!camera_ubo) { set_error("rt_raster_gbuffer: null camera"); return RT_ERR_INVALID_ARGUMENT; }
    int rc = rt_flush_scene(ctx);
    if (rc != RT_OK) return rc;
    rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    const size_t n = (size_t)ctx->width * ctx->height;
    RT_CUDA_TRY(ctx->ras_vis.ensure(n));
    RT_CUDA_TRY(ctx->ras_cam.ensure(64));
    for (auto& b : ctx->ras_img) RT_CUDA_TRY(b.ensure(n));
    rt::FrameParams fp;
    rt_fill_frame_params(ctx, camera_ubo, nullptr, 0, &fp);
    const rt::SceneView sc = rt_scene_view(ctx);
    rt_begin_timing(ctx);
    RT_CUDA_TRY(cudaMemcpyAsync(ctx->ras_cam.p, camera_ubo, 64 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA_TRY(cudaMemsetAsync(ctx->ras_vis.p, 0xFF, n * sizeof(unsigned long long), ctx->stream));
    if (sc.n_tris) {
        const uint32_t warps_per_block = 4, blocks = (uint32_t)std::min<size_t>(((size_t)sc.n_tris + warps_per_block - 1) / warps_per_block, 65535u * 16u);
        rt::raster_visibility_kernel<<<blocks, 128, 0, ctx->stream>>>(sc, ctx->d_tri_offsets.p, fp, ctx->ras_cam.p, ctx->ras_vis.p);
    }
    rt::RasterImages out{ctx->ras_img[0].p, ctx->ras_img[1].p, ctx->ras_img[2].p, ctx->ras_img[3].p, ctx->ras_img[4].p};
    rt::raster_resolve_kernel<<<(uint32_t)((n + 127) / 128), 128, 0, ctx->stream>>>(sc, ctx->d_tri_offsets.p, fp, ctx->ras_cam.p, ctx->ras_vis.p, out);
    rt_end_timing(ctx);
    RT_CUDA_TRY(cudaGetLastError());
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));  // camera_ubo is the caller's memory
    ctx->launches += 2;
    ctx->has_raster = true;
    return RT_OK;
}

int rt_readback_raster_gbuffer(rt_ctx* ctx, float* albedo, float* normal_metal, float* position_rough, float* emission, float* props) {
    if (!ctx) return RT_ERR_INVALID_ARGUMENT;
    if (!ctx->children.empty()) {
        const int rc0 = rt_readback_raster_gbuffer(ctx->children[0], albedo, normal_metal, position_rough, emission, props);
        if (rc0 != RT_OK) ctx->error = ctx->children[0]->error;
        return rc0;
    }
    cudaSetDevice(ctx->device);
    if (!ctx->has_raster) { set_error("rt_readback_raster_gbuffer: nothing rasterised"); return RT_ERR_NO_TARGET; }
    const size_t bytes = (size_t)ctx->width * ctx->height * sizeof(float4);
    float* dst[5] = {albedo, normal_metal, position_rough, emission, props};
    for (int k = 0; k < 5; k++)
        if (dst[k]) RT_CUDA_TRY(cudaMemcpyAsync(dst[k], ctx->ras_img[k].p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

}  // extern "C"
