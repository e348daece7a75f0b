// rt_restir.cu — ReSTIR DI on the primary-visibility pass of shaders/restir.rgen/rchit/rmiss.
//
// What the reference holds: the G-buffer pass (restir.rgen:43-92: one un-jittered ray through the pixel CORNER),
// the reservoir record RTXDI_DIReservoir {weightSum, targetPdf, M, lightIndex, uv, distance, W}
// (shaders/restir/restir_structs.glsl:1-12) with its empty constructor (:14-23), and a stub ray-gen that does not
// compile (shaders/restir/restir_di.rgen:17-22).  There is no resampling algorithm to be faithful to, so this
// file follows Bitterli et al., "Spatiotemporal reservoir resampling" (2020), built from the reference's own
// pieces: lights are picked with the power / distance^2 weights of simple.rchit:507-520, a triangle and a point
// on it uniformly (:239-322), the BSDF is evaluateFullBRDF (:118-166), visibility is the any-hit query of
// :350-385.  PARITY UNPINNED: the check is statistical, against the oracle's brute-force direct
// lighting (tests/test_gpu_parity.py::test_restir_*).
//
// Estimator.  Target function p^(y) = luminance(f(x, wo, wi) Le(y) cos_x cos_y / d^2), UNSHADOWED; the one
// visibility ray per pixel is traced for the finally selected sample only, and reservoirs are combined with the
// 1/Z normalisation (Z = sum of M over the sources whose own target function is positive at the selected sample),
// so candidates-only, temporal and spatial reuse are all unbiased; reuse only lowers the variance.
//
// Passes per frame:  surface (primary hit -> position, normal, material)  ->  candidates + temporal reuse  ->
// spatial reuse + shading + accumulation (simple.rgen:121-136 running mean, so the image is comparable).
#include <string.h>

#include <algorithm>

#include "rt_internal.h"
#include "rt_shade.cuh"
#include "rt_trav.cuh"

namespace rt {

struct RSurf {     // 48 B per pixel, three float4
    vec3 pos;      // world position of the primary hit
    uint32_t inst; // object index, 0xFFFFFFFF = miss
    vec3 N;        // shading normal, facing the viewer
    uint32_t mat;  // material index
    vec3 wo;       // unit vector towards the camera
    float depth;   // distance from the camera
};

__device__ __forceinline__ RSurf load_surf(const float4* s, uint32_t px) {
    const float4 a = s[3 * (size_t)px], b = s[3 * (size_t)px + 1], c = s[3 * (size_t)px + 2];
    RSurf r;
    r.pos = mk3(a.x, a.y, a.z); r.inst = __float_as_uint(a.w);
    r.N = mk3(b.x, b.y, b.z); r.mat = __float_as_uint(b.w);
    r.wo = mk3(c.x, c.y, c.z); r.depth = c.w;
    return r;
}

__device__ __forceinline__ float luminance(vec3 c) { return dot3(c, mk3(0.2126f, 0.7152f, 0.0722f)); }

// ---- pass 1: primary visibility (restir.rgen:43-59 + restir.rchit:36-59 / restir.rmiss:18-20) ---------------
// Passes 1 and 2 run over a LIST of global image rows (rows == nullptr: all of them): the rows this context owns plus the
// halo of ceil(spatial_radius) rows around them, so that pass 3 finds every neighbour it may pick — computed here, not fetched
// from the rank that owns it ("recompute G-buffer + candidates for the halo rows locally (no comms)").  Every value
// of a pixel depends on its GLOBAL coordinates only, so two ranks that both process a row get the same bits, and the image
// does not depend on how many ranks share the frame.  All per-pixel arrays of this file are indexed by global pixel.
// the closest hits come from rt_frame.cu (launch_corner_hits: the contract's traversal); this kernel only shades them
__global__ void __launch_bounds__(128) restir_surface_kernel(SceneView sc, FrameParams fp, const uint32_t* rows, uint32_t n_rows,
                                                             const uint4* hits, float4* surf, DevCounters* counters) {
    const uint32_t n_px = n_rows * fp.width;
    uint32_t c_nodes = 0, c_tris = 0, c_rays = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_px; i += gridDim.x * blockDim.x) {
        const uint32_t r = i / fp.width, x = i - r * fp.width;
        const uint32_t y = rows ? __ldg(rows + r) : r;
        const uint32_t px = y * fp.width + x;
        vec3 o, d;
        camera_ray(fp, (float)x, (float)y, false, &o, &d);  // pixel corner, direction not re-normalised
        Hit h;
        {
            const uint4 hr = hits[i];
            h.t = 0.0f; h.u = __uint_as_float(hr.x); h.v = __uint_as_float(hr.y); h.inst = hr.z; h.prim = hr.w;
            h.id = hr.z == 0xFFFFFFFFu ? 0xFFFFFFFFu : 0u;
        }
        float4 a = make_float4(0.f, 0.f, 0.f, __uint_as_float(0xFFFFFFFFu)), b = make_float4(0.f, 0.f, 0.f, 0.f),
               c = make_float4(0.f, 0.f, 0.f, 0.f);
        if (h.id != 0xFFFFFFFFu) {
            const DevObject* obj = sc.objects + h.inst;
            const uint32_t mi = __ldg(&obj->material_index);
            const uint32_t* ip = sc.indices + __ldg(&obj->index_base) + 3u * h.prim;
            const rt_vertex* vb = sc.vertices + __ldg(&obj->vertex_base);
            const rt_vertex *v0 = vb + __ldg(ip), *v1 = vb + __ldg(ip + 1), *v2 = vb + __ldg(ip + 2);
            const float bx = (1.0f - h.u) - h.v, by = h.u, bz = h.v;
            float model[12], n9[9];
#pragma unroll
            for (int k = 0; k < 12; k++) model[k] = __ldg(&obj->model[k]);
#pragma unroll
            for (int k = 0; k < 9; k++) n9[k] = __ldg(&obj->nrm[k]);
            const vec3 pos = add3(add3(scale3(vertex_pos(v0), bx), scale3(vertex_pos(v1), by)), scale3(vertex_pos(v2), bz));
            const vec3 wp = xf_point(model, pos);
            const vec3 nrm = normalize3(add3(add3(scale3(vertex_nrm(v0), bx), scale3(vertex_nrm(v1), by)), scale3(vertex_nrm(v2), bz)));
            vec3 N = normalize3(mat3_mul(n9, nrm));
            const vec3 wo = neg3(normalize3(d));
            if (!(dot3(N, wo) > 0.0f)) N = neg3(N);
            a = make_float4(wp.x, wp.y, wp.z, __uint_as_float(h.inst));
            b = make_float4(N.x, N.y, N.z, __uint_as_float(mi));
            c = make_float4(wo.x, wo.y, wo.z, length3(sub3(wp, o)));
        }
        surf[3 * (size_t)px] = a; surf[3 * (size_t)px + 1] = b; surf[3 * (size_t)px + 2] = c;
    }
    if (counters) {
        unsigned long long* dst[3] = {&counters->extend_rays, &counters->nodes, &counters->tris};
        uint32_t v[3] = {c_rays, c_nodes, c_tris};
#pragma unroll
        for (int k = 0; k < 3; k++) {
            uint32_t s = v[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
            if ((threadIdx.x & 31u) == 0u && s) atomicAdd(dst[k], (unsigned long long)s);
        }
    }
}

// ---- light samples --------------------------------------------------------------------------------------
struct LightPoint {
    vec3 y, n;   // world position and unit geometric normal (unflipped)
    vec3 Le;     // emission_color * emission_power of the light's material
    float area;  // world-space area of the triangle
    uint32_t object;
};

// World-space light triangles, flattened once per scene flush (light_table_kernel): a reservoir names a light point as
// (light, triangle, barycentrics), and rebuilding it from the scene tables is a chain of five dependent loads (light -> object ->
// indices -> vertices -> material) that the resolve kernel — which waits on memory, ncu r2f — walks up to five times per pixel.
// Here it is two: aux[light] (32 B: first triangle, triangle count, object, Le) -> the triangle's record (64 B, two LDG.256:
// three world-space vertices, unit normal, area).
struct LightTable {
    const float4* tris;  // 4 x float4 per light triangle: w0.xyz area | w1.xyz n.x | w2.xyz n.y | n.z - - -
    const uint4* aux;    // 2 x uint4 per light: tri_base n_tris object - | Le.xyz -
};

__global__ void light_table_kernel(SceneView sc, const uint32_t* tri_base, float4* tris, uint4* aux) {
    const uint32_t li = blockIdx.x;
    const DevLight* light = sc.lights + li;
    const uint32_t lobj = light->object_index, ntri = light->num_triangles;
    const DevObject* lo = sc.objects + lobj;
    if (threadIdx.x == 0) {
        const rt_material* lm = sc.materials + lo->material_index;
        aux[2 * li] = make_uint4(tri_base[li], ntri, lobj, 0u);
        aux[2 * li + 1] = make_uint4(__float_as_uint(lm->emission_color[0] * lm->emission_power), __float_as_uint(lm->emission_color[1] * lm->emission_power),
                                     __float_as_uint(lm->emission_color[2] * lm->emission_power), 0u);
    }
    float tr[12];
    for (int k = 0; k < 12; k++) tr[k] = light->transform[k];
    for (uint32_t t = threadIdx.x; t < ntri; t += blockDim.x) {
        const uint32_t* ip = sc.indices + lo->index_base + 3u * t;
        const rt_vertex* vb = sc.vertices + lo->vertex_base;
        const vec3 w0 = xf_point(tr, vertex_pos(vb + ip[0])), w1 = xf_point(tr, vertex_pos(vb + ip[1])), w2 = xf_point(tr, vertex_pos(vb + ip[2]));
        const vec3 cr = cross3(sub3(w1, w0), sub3(w2, w0));
        const vec3 n = normalize3(cr);
        float4* dst = tris + 4 * (size_t)(tri_base[li] + t);
        dst[0] = make_float4(w0.x, w0.y, w0.z, 0.5f * length3(cr));
        dst[1] = make_float4(w1.x, w1.y, w1.z, n.x);
        dst[2] = make_float4(w2.x, w2.y, w2.z, n.y);
        dst[3] = make_float4(n.z, 0.0f, 0.0f, 0.0f);
    }
}

// (light, triangle, barycentrics) -> point; the reservoir stores exactly this triple
__device__ __forceinline__ LightPoint light_point(const LightTable& lt, uint32_t light_idx, uint32_t tri, float bu, float bv) {
    const uint4 a0 = __ldg(lt.aux + 2 * light_idx), a1 = __ldg(lt.aux + 2 * light_idx + 1);
    const uint4* tp = reinterpret_cast<const uint4*>(lt.tris + 4 * (size_t)(a0.x + tri));
    uint4 q0, q1, q2, q3;
    ldg256(tp, q0, q1);
    ldg256(tp + 2, q2, q3);
    const vec3 w0 = mk3(__uint_as_float(q0.x), __uint_as_float(q0.y), __uint_as_float(q0.z));
    const vec3 w1 = mk3(__uint_as_float(q1.x), __uint_as_float(q1.y), __uint_as_float(q1.z));
    const vec3 w2 = mk3(__uint_as_float(q2.x), __uint_as_float(q2.y), __uint_as_float(q2.z));
    const float bw = (1.0f - bu) - bv;
    LightPoint lp;
    lp.y = add3(add3(scale3(w0, bu), scale3(w1, bv)), scale3(w2, bw));
    lp.area = __uint_as_float(q0.w);
    lp.n = mk3(__uint_as_float(q1.w), __uint_as_float(q2.w), __uint_as_float(q3.x));
    lp.Le = mk3(__uint_as_float(a1.x), __uint_as_float(a1.y), __uint_as_float(a1.z));
    lp.object = a0.z;
    return lp;
}
__device__ __forceinline__ uint32_t light_num_triangles(const LightTable& lt, uint32_t light_idx) { return __ldg(lt.aux + 2 * light_idx).y; }

// unshadowed contribution f Le cos_x cos_y / d^2 of a light point at a surface; zero when either side faces away
__device__ __forceinline__ vec3 unshadowed(const SceneView& sc, const RSurf& s, const Basis& basis, const vec3 wo_l, const Mat& mat,
                                          const LightPoint& lp) {
    const vec3 to = sub3(lp.y, s.pos);
    const float d2 = dot3(to, to);
    if (!(d2 > 1e-12f) || !(lp.area > 0.0f)) return splat3(0.0f);
    const float d = sqrtf(d2);
    const vec3 wi_w = div3(to, d);
    const float cos_y = fabsf(dot3(lp.n, wi_w));  // lights emit from both faces (sampleLight flips the normal)
    const vec3 wi = world_to_local(wi_w, basis);
    if (!(wi.z > 1e-4f) || !(cos_y > 1e-6f)) return splat3(0.0f);
    const vec3 f = eval_full_brdf(wo_l, wi, normalize3(add3(wo_l, wi)), mat);
    return scale3(mul3(f, lp.Le), (wi.z * cos_y) / d2);
}

__device__ __forceinline__ rt_reservoir empty_reservoir() {  // RTXDI_EmptyDIReservoir, restir_structs.glsl:14-23
    rt_reservoir r;
    r.weight_sum = 0.0f; r.target_pdf = 0.0f; r.M = 0u; r.light_index = 0xFFFFFFFFu;
    r.uv[0] = 0.0f; r.uv[1] = 0.0f; r.distance = 0.0f; r.W = 0.0f;
    return r;
}

struct SurfShade {  // what every target-function evaluation at a surface needs
    RSurf s;
    Mat mat;
    Basis basis;
    vec3 wo_l;
};
__device__ __forceinline__ SurfShade prepare_surface(const SceneView& sc, const float4* surf, uint32_t px) {
    SurfShade ss;
    ss.s = load_surf(surf, px);
    if (ss.s.inst != 0xFFFFFFFFu) {
        ss.mat = load_material(sc.materials + ss.s.mat);
        ss.basis = create_basis(ss.s.N);
        ss.wo_l = world_to_local(ss.s.wo, ss.basis);
    }
    return ss;
}
__device__ __forceinline__ bool reservoir_sample_valid(const SceneView& sc, const LightTable& lt, const rt_reservoir& r) {
    if (r.light_index >= sc.n_lights) return false;
    return __float_as_uint(r.distance) < light_num_triangles(lt, r.light_index);
}
// target function of a light point that has already been reconstructed (the same sample judged at several surfaces)
__device__ __forceinline__ float target_of_point(const SceneView& sc, const SurfShade& ss, const LightPoint& lp) {
    if (lp.object == ss.s.inst) return 0.0f;  // a light does not light itself (simple.rchit:548-553 skips it)
    const float p = luminance(unshadowed(sc, ss.s, ss.basis, ss.wo_l, ss.mat, lp));
    return (p > 0.0f && p == p && !is_inf1(p)) ? p : 0.0f;
}
// Is the target function of `lp` positive at surface `s`?  Same conditions as unshadowed() / target_of_point() — not the light
// itself, a proper distance and area, the point above the surface's horizon (wi.z > 1e-4: the basis' z axis is N) and the
// surface on an emitting side of the light — without evaluating the BSDF: f > 0 whenever both cosines are (the GGX term never
// vanishes for alpha >= 0.02 and N faces the viewer by construction).  The 1/Z normalisation only needs the sign.
__device__ __forceinline__ bool target_positive(const RSurf& s, const LightPoint& lp) {
    if (lp.object == s.inst) return false;
    const vec3 to = sub3(lp.y, s.pos);
    const float d2 = dot3(to, to);
    if (!(d2 > 1e-12f) || !(lp.area > 0.0f)) return false;
    const vec3 wi_w = div3(to, sqrtf(d2));
    return dot3(s.N, wi_w) > 1e-4f && fabsf(dot3(lp.n, wi_w)) > 1e-6f;
}
__device__ __forceinline__ float target_of(const SceneView& sc, const LightTable& lt, const SurfShade& ss, const rt_reservoir& r) {
    if (!reservoir_sample_valid(sc, lt, r)) return 0.0f;
    return target_of_point(sc, ss, light_point(lt, r.light_index, __float_as_uint(r.distance), r.uv[0], r.uv[1]));
}

// ---- pass 2: initial candidates (RIS) + temporal reuse ---------------------------------------------------
__global__ void __launch_bounds__(128) restir_candidates_kernel(SceneView sc, LightTable lt, FrameParams fp, rt_restir_params prm,
                                                                const uint32_t* rows, uint32_t n_rows, const float4* surf,
                                                                const rt_reservoir* prev, rt_reservoir* out, DevCounters* counters) {
    const uint32_t n_px = n_rows * fp.width;
    uint32_t c_nee = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_px; i += gridDim.x * blockDim.x) {
        const uint32_t r = i / fp.width, x = i - r * fp.width;
        const uint32_t y = rows ? __ldg(rows + r) : r;
        const uint32_t px = y * fp.width + x;
        const SurfShade ss = prepare_surface(sc, surf, px);
        rt_reservoir res = empty_reservoir();
        if (ss.s.inst == 0xFFFFFFFFu) { out[px] = res; continue; }
        uint32_t seed = tea16((y * fp.width + x) ^ 0x52455354u, fp.frame);  // 'REST': a stream of its own
        const uint32_t L = sc.n_lights < RT_MAXLIGHTS ? sc.n_lights : RT_MAXLIGHTS;
        float total = 0.0f;
        for (uint32_t i = 0; i < L; i++)
            if (ss.s.inst != __ldg(&sc.lights[i].object_index)) total += light_weight(sc, ss.s.pos, i);
        float wsum = 0.0f, sel_p = 0.0f;
        uint32_t M = 0;
        if (total > 0.0f) {
            for (uint32_t c = 0; c < prm.initial_candidates; c++) {
                M++;
                // light: power / distance^2 roulette of simple.rchit:543-583
                float r1 = rnd01(seed) * total;
                uint32_t sel = 0xFFFFFFFFu;
                float wsel = 0.0f;
                for (uint32_t i = 0; i < L; i++) {
                    const float w = (ss.s.inst == __ldg(&sc.lights[i].object_index)) ? 0.0f : light_weight(sc, ss.s.pos, i);
                    r1 -= w;
                    if (r1 <= 0.0f && w > 0.0f) { sel = i; wsel = w; break; }
                }
                const float ut = rnd01(seed), q1 = rnd01(seed), q2 = rnd01(seed), uu = rnd01(seed);
                if (sel == 0xFFFFFFFFu) continue;
                const uint32_t ntri = light_num_triangles(lt, sel);
                if (ntri == 0u) continue;  // a light whose mesh is empty has no point to sample
                uint32_t tri = (uint32_t)(ut * (float)ntri);
                if (tri > ntri - 1u) tri = ntri - 1u;
                const float sq = sqrtf(q1);
                const float bu = 1.0f - sq, bv = sq * (1.0f - q2);  // uniform point, simple.rchit:262-270
                const LightPoint lp = light_point(lt, sel, tri, bu, bv);
                c_nee++;
                if (!(lp.area > 0.0f)) continue;
                const float src_pdf = ((wsel / total) / (float)ntri) / lp.area;  // area measure
                float p = luminance(unshadowed(sc, ss.s, ss.basis, ss.wo_l, ss.mat, lp));
                if (!(p > 0.0f) || !(src_pdf > 0.0f) || is_inf1(p)) continue;
                const float w = p / src_pdf;
                wsum += w;
                if (uu * wsum < w) {
                    res.light_index = sel; res.distance = __uint_as_float(tri); res.uv[0] = bu; res.uv[1] = bv;
                    sel_p = p;
                }
            }
        }
        float Z = (float)M;
        // temporal reuse: the reservoir this pixel kept last frame (static camera and scene: same pixel)
        if (prev && (prm.flags & RT_RESTIR_TEMPORAL)) {
            const rt_reservoir rp = prev[px];
            if (rp.M > 0u) {
                const uint32_t cap = prm.history_cap * (M ? M : 1u);
                const uint32_t Mp = rp.M < cap ? rp.M : cap;
                const float pp = target_of(sc, lt, ss, rp);
                const float w = pp * rp.W * (float)Mp;
                const float uu = rnd01(seed);
                if (w > 0.0f && !is_inf1(w)) {
                    wsum += w;
                    if (uu * wsum < w) {
                        res.light_index = rp.light_index; res.distance = rp.distance; res.uv[0] = rp.uv[0]; res.uv[1] = rp.uv[1];
                        sel_p = pp;
                    }
                }
                M += Mp;
                Z += (float)Mp;  // same surface, same target function: the source counts whenever the winner does
            }
        }
        res.M = M;
        res.weight_sum = wsum;
        res.target_pdf = sel_p;
        res.W = (sel_p > 0.0f && Z > 0.0f) ? wsum / (Z * sel_p) : 0.0f;
        if (!(res.W == res.W) || is_inf1(res.W)) res.W = 0.0f;
        out[px] = res;
    }
    if (counters) {
        uint32_t s = c_nee;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
        if ((threadIdx.x & 31u) == 0u && s) atomicAdd(&counters->nee, (unsigned long long)s);
    }
}

// ---- pass 3: spatial reuse + shading + accumulation -----------------------------------------------------------
#define RT_RESTIR_MAX_NEIGHBOURS 8

#ifndef RT_RESOLVE_MIN_BLOCKS
#define RT_RESOLVE_MIN_BLOCKS 4  /* 128 registers, no spills: 2.015 -> 1.886 ms/frame (3 blocks x 152 registers before) */
#endif
__global__ void __launch_bounds__(128, RT_RESOLVE_MIN_BLOCKS) restir_resolve_kernel(SceneView sc, LightTable lt, FrameParams fp, rt_restir_params prm, const float4* surf,
                                                             const rt_reservoir* in, rt_reservoir* out, float4* image,
                                                             DevCounters* counters) {
    const uint32_t n_px = fp.local_rows * fp.width;
    uint32_t c_nodes = 0, c_tris = 0, c_rays = 0;
    const vec3 clearc = mk3(fp.clear[0], fp.clear[1], fp.clear[2]);
    // this context's own rows only; lpx indexes its compact image, px (global) everything else
    for (uint32_t lpx = blockIdx.x * blockDim.x + threadIdx.x; lpx < n_px; lpx += gridDim.x * blockDim.x) {
        const uint32_t lrow = lpx / fp.width, x = lpx - lrow * fp.width;
        const uint32_t y = strip_global_row(fp, lrow);
        const uint32_t px = y * fp.width + x;
        const SurfShade ss = prepare_surface(sc, surf, px);
        vec3 color = splat3(0.0f);
        rt_reservoir res = in[px];
        if (ss.s.inst == 0xFFFFFFFFu) {
            color = clearc;  // restir.rmiss: nothing hit -> background
            res = empty_reservoir();
        } else {
            uint32_t seed = tea16((y * fp.width + x) ^ 0x53504154u, fp.frame);  // 'SPAT'
            LightPoint lpw;  // the surviving sample's light point, once it has been reconstructed
            bool have_lpw = false;
            if ((prm.flags & RT_RESTIR_SPATIAL) && prm.spatial_samples > 0u) {
                // the pixel's own reservoir enters as one sample of weight p^ W M
                float sel_p = res.target_pdf;
                float wsum = sel_p * res.W * (float)res.M;
                uint32_t M = res.M;
                uint32_t nb[RT_RESTIR_MAX_NEIGHBOURS];
                uint32_t n_nb = 0;
                const uint32_t k_max = prm.spatial_samples < RT_RESTIR_MAX_NEIGHBOURS ? prm.spatial_samples : RT_RESTIR_MAX_NEIGHBOURS;
                for (uint32_t k = 0; k < k_max; k++) {
                    const float ur = rnd01(seed), ua = rnd01(seed), us = rnd01(seed);
                    float sn, cs;
                    sincos_rt(RT_TWO_PI * ua, &sn, &cs);
                    const float rad = prm.spatial_radius * sqrtf(ur);
                    const int qx = (int)x + (int)rintf(rad * cs), qy = (int)y + (int)rintf(rad * sn);
                    // |qy - y| <= ceil(spatial_radius): inside the halo passes 1 and 2 covered, whoever owns the row
                    if (qx < 0 || qx >= (int)fp.width || qy < 0 || qy >= (int)fp.height) continue;
                    const uint32_t qpx = (uint32_t)qy * fp.width + (uint32_t)qx;
                    if (qpx == px) continue;
                    const RSurf qs = load_surf(surf, qpx);
                    // reuse only across similar surfaces: normals within ~25 degrees, depths within 10 %
                    if (qs.inst == 0xFFFFFFFFu || dot3(qs.N, ss.s.N) < 0.9f || fabsf(qs.depth - ss.s.depth) > 0.1f * ss.s.depth) continue;
                    const rt_reservoir rq = in[qpx];
                    if (rq.M == 0u) continue;
                    nb[n_nb++] = qpx;
                    M += rq.M;
                    const float pq = target_of(sc, lt, ss, rq);  // the neighbour's sample, judged at THIS surface
                    const float w = pq * rq.W * (float)rq.M;
                    if (w > 0.0f && !is_inf1(w)) {
                        wsum += w;
                        if (us * wsum < w) {
                            res.light_index = rq.light_index; res.distance = rq.distance; res.uv[0] = rq.uv[0]; res.uv[1] = rq.uv[1];
                            sel_p = pq;
                        }
                    }
                }
                // 1/Z: count the sources that could have produced the winner (their own target function is > 0 there)
                float Z = (sel_p > 0.0f) ? (float)res.M : 0.0f;
                if (sel_p > 0.0f && reservoir_sample_valid(sc, lt, res)) {  // one light point for all sources (an invalid sample has target 0 everywhere)
                    lpw = light_point(lt, res.light_index, __float_as_uint(res.distance), res.uv[0], res.uv[1]);
                    have_lpw = true;
                    for (uint32_t k = 0; k < n_nb; k++)
                        if (target_positive(load_surf(surf, nb[k]), lpw)) Z += (float)in[nb[k]].M;
                }
                res.M = M;
                res.weight_sum = wsum;
                res.target_pdf = sel_p;
                res.W = (sel_p > 0.0f && Z > 0.0f) ? wsum / (Z * sel_p) : 0.0f;
                if (!(res.W == res.W) || is_inf1(res.W)) res.W = 0.0f;
            }
            // shading: f Le G V W for the surviving sample, one visibility ray (isVisibleRQ, simple.rchit:350-385)
            if (res.W > 0.0f && reservoir_sample_valid(sc, lt, res)) {
                const LightPoint lp = have_lpw ? lpw : light_point(lt, res.light_index, __float_as_uint(res.distance), res.uv[0], res.uv[1]);
                const vec3 c = scale3(unshadowed(sc, ss.s, ss.basis, ss.wo_l, ss.mat, lp), res.W);
                const vec3 tl = normalize3(sub3(lp.y, ss.s.pos));
                const vec3 from = add3(ss.s.pos, scale3(scale3(ss.s.N, 0.001f), sign1(dot3(ss.s.N, tl))));
                vec3 sdir = sub3(lp.y, from);
                const float sdist = length3(sdir);
                if (sdist > 0.0f && lp.object != ss.s.inst) {
                    sdir = div3(sdir, sdist);
                    c_rays++;
                    const Hit sh = traverse<true, true>(sc, from, sdir, 0.001f, sdist * 0.999f, lp.object, c_nodes, c_tris);
                    if (sh.id == 0xFFFFFFFFu) color = add3(color, c);
                }
            }
            if (ss.mat.emission_power > 0.0f) color = add3(color, scale3(ss.mat.emission_color, ss.mat.emission_power));
        }
        out[px] = res;
        image[lpx] = accumulate_pixel(fp.frame ? image[lpx] : make_float4(0.f, 0.f, 0.f, 0.f), color, fp.clamp, fp.frame);
    }
    if (counters) {
        unsigned long long* dst[5] = {&counters->shadow_rays, &counters->nodes, &counters->tris, &counters->shadow_nodes, &counters->shadow_tris};
        uint32_t v[5] = {c_rays, c_nodes, c_tris, c_nodes, c_tris};
#pragma unroll
        for (int k = 0; k < 5; k++) {
            uint32_t s = v[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
            if ((threadIdx.x & 31u) == 0u && s) atomicAdd(dst[k], (unsigned long long)s);
        }
    }
}

}  // namespace rt

#define set_error ctx->set_error

extern "C" {

void rt_default_restir_params(rt_restir_params* p) {
    if (!p) return;
    p->flags = RT_RESTIR_TEMPORAL | RT_RESTIR_SPATIAL;
    p->initial_candidates = 8;
    p->spatial_samples = 4;
    p->spatial_radius = 16.0f;
    p->history_cap = 20;
    p->reserved[0] = p->reserved[1] = p->reserved[2] = 0;
}

// One ReSTIR DI frame: direct lighting of the primary hits, accumulated into the same RGBA32F image as
// rt_trace_frame (frame_number == 0 overwrites).  The temporal history is this context's previous rt_restir_frame
// with the same camera; any other camera, a resize or a scene edit restarts it (the frame number does not).
int rt_restir_frame(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t frame_number,
                    const rt_restir_params* params) {
    if (!ctx) return RT_ERR_INVALID_ARGUMENT;
    for (rt_ctx* c : ctx->children) {  // multi-GPU context: every device shades its own strips
        const int rc = rt_restir_frame(c, camera_ubo, clear_color, frame_number, params);
        if (rc != RT_OK) { ctx->error = c->error; return rc; }
    }
    if (!ctx->children.empty()) return RT_OK;
    cudaSetDevice(ctx->device);
    if (!camera_ubo || !clear_color) { set_error("rt_restir_frame: null argument"); return RT_ERR_INVALID_ARGUMENT; }
    rt_restir_params prm;
    if (params) prm = *params; else rt_default_restir_params(&prm);
    if (prm.initial_candidates == 0 || prm.initial_candidates > 1024 || !(prm.spatial_radius >= 0.0f) || prm.history_cap == 0) {
        set_error("rt_restir_frame: bad parameters");
        return RT_ERR_INVALID_ARGUMENT;
    }
    int rc = rt_flush_scene(ctx);
    if (rc != RT_OK) return rc;
    rc = rt_require_target(ctx);
    if (rc != RT_OK) return rc;
    if (ctx->local_rows == 0) return RT_OK;
    const size_t n = (size_t)ctx->height * ctx->width;  // every per-pixel array is indexed by GLOBAL pixel (halo rows included)
    RT_CUDA_TRY(ctx->rs_surf.ensure(3 * n));
    RT_CUDA_TRY(ctx->rs_cur.ensure(n));
    RT_CUDA_TRY(ctx->rs_hist[0].ensure(n));
    RT_CUDA_TRY(ctx->rs_hist[1].ensure(n));
    // the rows of passes 1 and 2: own rows + halo.  A strip partition with few ranks and 8-row strips makes that the whole
    // frame (radius 16 >= the gap between a rank's strips up to N = 5); beyond, it is 5 x the own rows at most.
    const uint32_t halo = (prm.flags & RT_RESTIR_SPATIAL) && prm.spatial_samples ? (uint32_t)ceilf(prm.spatial_radius) : 0u;
    uint32_t n_rows = ctx->height;
    const uint32_t* d_rows = nullptr;
    if (ctx->cfg.strip_count > 1) {
        if (ctx->rs_rows_halo != halo || ctx->rs_rows_epoch != ctx->band_epoch || ctx->rs_rows.n == 0) {
            std::vector<uint8_t> mark(ctx->height, 0);
            const uint32_t sr = ctx->cfg.strip_rows;
            for (uint32_t lrow = 0; lrow < ctx->local_rows; lrow++) {
                const uint32_t band = lrow / sr;
                const uint32_t strip = ctx->band_strip.empty() ? band * ctx->cfg.strip_count + ctx->cfg.strip_index : ctx->band_strip[band];
                const uint32_t y = strip * sr + lrow % sr;
                const uint32_t y0 = y > halo ? y - halo : 0u, y1 = std::min(ctx->height - 1u, y + halo);
                for (uint32_t q = y0; q <= y1; q++) mark[q] = 1;
            }
            std::vector<uint32_t> rows;
            for (uint32_t y = 0; y < ctx->height; y++) if (mark[y]) rows.push_back(y);
            RT_CUDA_TRY(ctx->rs_rows.ensure(rows.size()));
            RT_CUDA_TRY(cudaMemcpyAsync(ctx->rs_rows.p, rows.data(), rows.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
            RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
            if (halo > ctx->rs_rows_halo || ctx->rs_rows_epoch != ctx->band_epoch) ctx->rs_history = false;  // rows without a history joined
            ctx->rs_n_rows = (uint32_t)rows.size();
            ctx->rs_rows_halo = halo;
            ctx->rs_rows_epoch = ctx->band_epoch;
        }
        n_rows = ctx->rs_n_rows;
        d_rows = ctx->rs_rows.p;
    }
    // the history follows the camera, not the accumulation counter: frame 0 overwrites the image but may still reuse
    if (ctx->rs_history && memcmp(ctx->rs_cam, camera_ubo, sizeof(ctx->rs_cam)) != 0) ctx->rs_history = false;
    rt::FrameParams fp;
    rt_fill_frame_params(ctx, camera_ubo, clear_color, frame_number, &fp);
    const rt::SceneView sc = rt_scene_view(ctx);
    rt::DevCounters* cnt = (ctx->cfg.flags & RT_FLAG_COUNTERS) ? ctx->counters.p : nullptr;
    if (ctx->rs_light_epoch != ctx->tables_epoch) {  // the light tables changed (build, material / object / transform edit): flatten again
        const uint32_t L = ctx->n_lights;
        std::vector<uint32_t> base(L + 1, 0u);
        for (uint32_t i = 0; i < L; i++) base[i + 1] = base[i] + ctx->light_tris[i];
        RT_CUDA_TRY(ctx->rs_light_base.ensure(L + 1));
        RT_CUDA_TRY(ctx->rs_light_tris.ensure((size_t)base[L] * 4));
        RT_CUDA_TRY(ctx->rs_light_aux.ensure((size_t)L * 2));
        RT_CUDA_TRY(cudaMemcpyAsync(ctx->rs_light_base.p, base.data(), (L + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
        if (L) rt::light_table_kernel<<<L, 128, 0, ctx->stream>>>(sc, ctx->rs_light_base.p, ctx->rs_light_tris.p, ctx->rs_light_aux.p);
        RT_CUDA_TRY(cudaGetLastError());
        RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));  // `base` is on this stack frame
        ctx->rs_light_epoch = ctx->tables_epoch;
        ctx->launches++;
    }
    const rt::LightTable lt{ctx->rs_light_tris.p, ctx->rs_light_aux.p};
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
    const uint32_t grid = (uint32_t)std::min<size_t>(((size_t)n_rows * ctx->width + 127) / 128, (size_t)sms * 32);
    const uint32_t grid3 = (uint32_t)std::min<size_t>(((size_t)ctx->local_rows * ctx->width + 127) / 128, (size_t)sms * 32);
    // History = the reservoirs after candidates + temporal reuse.  The spatially combined reservoirs are used for
    // shading only: feeding them back lets one lucky sample spread to its neighbours and return through the
    // temporal pass every frame (measured: fireflies persisting for the whole history length, mean +37 % after
    // 64 frames on conference.glb), which the 1/Z normalisation does not damp.
    const rt_reservoir* prev = ctx->rs_history ? ctx->rs_hist[ctx->rs_flip].p : nullptr;
    rt_reservoir* temporal = ctx->rs_hist[ctx->rs_flip ^ 1].p;
    rt_begin_timing(ctx);
    // primary visibility: rt_frame.cu finds the hits (packets when the wavefront state has a slot per covered pixel — always with
    // one context per frame — else one ray per thread); both under the arithmetic contract, so every context gets the same records
    const uint4* hits = nullptr;
    {
        const bool packets = ctx->wf && rt::wavefront_slots(ctx->wf) >= (size_t)n_rows * ctx->width && !(ctx->cfg.flags & RT_FLAG_PER_RAY_PRIMARY);
        if (!packets) RT_CUDA_TRY(ctx->rs_hits.ensure((size_t)n_rows * ctx->width));
        rt::FrameResources fr;
        fr.dims = &ctx->dims;
        fr.counters = ctx->counters.p;
        fr.wf = ctx->wf;
        fr.launches = &ctx->launches;
        std::string err;
        rc = rt::launch_corner_hits(sc, fp, fr, d_rows, n_rows, packets ? nullptr : ctx->rs_hits.p, ctx->stream, err, &hits);
        if (rc != RT_OK) { set_error(err); return rc; }
    }
    rt::restir_surface_kernel<<<grid, 128, 0, ctx->stream>>>(sc, fp, d_rows, n_rows, hits, ctx->rs_surf.p, cnt);
    rt::restir_candidates_kernel<<<grid, 128, 0, ctx->stream>>>(sc, lt, fp, prm, d_rows, n_rows, ctx->rs_surf.p, prev, temporal, cnt);
    rt::restir_resolve_kernel<<<grid3, 128, 0, ctx->stream>>>(sc, lt, fp, prm, ctx->rs_surf.p, temporal, ctx->rs_cur.p, ctx->image.p, cnt);
    rt_end_timing(ctx);
    RT_CUDA_TRY(cudaGetLastError());
    ctx->launches += 3;
    ctx->frames_traced += 1;
    ctx->rs_flip ^= 1;
    ctx->rs_history = true;
    memcpy(ctx->rs_cam, camera_ubo, sizeof(ctx->rs_cam));
    return RT_OK;
}

// the reservoirs the last rt_restir_frame shaded with (after spatial reuse; full-size W*H array; rows of other strips are left untouched)
int rt_readback_reservoirs(rt_ctx* ctx, rt_reservoir* dst, size_t bytes) {
    if (!ctx) return RT_ERR_INVALID_ARGUMENT;
    for (rt_ctx* c : ctx->children) {
        const int rc = rt_readback_reservoirs(c, dst, bytes);
        if (rc != RT_OK) { ctx->error = c->error; return rc; }
    }
    if (!ctx->children.empty()) return RT_OK;
    cudaSetDevice(ctx->device);
    if (!ctx->rs_history) { set_error("rt_readback_reservoirs: no ReSTIR frame traced"); return RT_ERR_NO_TARGET; }
    const size_t row = (size_t)ctx->width * sizeof(rt_reservoir);
    if (!dst || bytes < row * ctx->height) { set_error("rt_readback_reservoirs: destination too small"); return RT_ERR_INVALID_ARGUMENT; }
    const rt_reservoir* src = ctx->rs_cur.p;  // indexed by global pixel: this context's own rows are the ones it resolved
    const uint32_t sr = ctx->cfg.strip_rows;
    for (uint32_t lrow = 0; lrow < ctx->local_rows; lrow += sr) {
        const uint32_t rows = std::min(sr, ctx->local_rows - lrow);
        const uint32_t band = lrow / sr;
        const uint32_t y = (ctx->band_strip.empty() ? band * ctx->cfg.strip_count + ctx->cfg.strip_index : ctx->band_strip[band]) * sr;
        RT_CUDA_TRY(cudaMemcpyAsync((char*)dst + (size_t)y * row, (const char*)src + (size_t)y * row, rows * row,
                                    cudaMemcpyDeviceToHost, ctx->stream));
    }
    RT_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

}  // extern "C"
