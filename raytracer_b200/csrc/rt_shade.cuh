// rt_shade.cuh — device restatement of the reference's hit shading and ray generation:
//   shaders/simple.rchit  (BSDF :69-217, light sampling :239-322, visibility setup :350-385,
//                          sampleBRDF :403-449, NEE + MIS :472-583, main :590-706)
//   shaders/simple.rgen   (camera ray :41-53, jitter :25-38, roulette :55-68)
//   shaders/math.glsl     (createBasis / worldToLocal / localToWorld :9-29)
// Arithmetic follows rt_math.cuh's contract (no contraction, fixed association order).
//
// The reference resolves visibility inline (rayQueryEXT) in the middle of the hit shader.  No random
// number and no branch after that point depends on the answer, so the wavefront pipeline splits the
// hit shader at the query: shade_hit() returns the NEE contribution "as if visible" together with the
// shadow ray, and the caller adds it (before the emissive / sky terms, to keep the float summation
// order of the reference) once the any-hit traversal has answered.
#pragma once
#include "rt_types.cuh"

namespace rt {

#define RT_MAXLIGHTS 256u       /* simple.rchit:13 */
#define RT_EPS_PDF 1e-6f        /* simple.rchit:63 */
#define RT_EPS_COS 1e-4f        /* simple.rchit:64 */
#define RT_EPS_VOH 1e-4f        /* simple.rchit:65 */
#define RT_MIN_ROUGHNESS 0.02f  /* simple.rchit:66 */

struct Mat {  // registers copy of rt_material
    vec3 albedo, emission_color;
    float emission_power, roughness, metallic;
};

__device__ __forceinline__ Mat load_material(const rt_material* m) {
    Mat r;
    r.albedo = mk3(__ldg(&m->albedo[0]), __ldg(&m->albedo[1]), __ldg(&m->albedo[2]));
    r.emission_color = mk3(__ldg(&m->emission_color[0]), __ldg(&m->emission_color[1]), __ldg(&m->emission_color[2]));
    r.emission_power = __ldg(&m->emission_power);
    r.roughness = __ldg(&m->roughness);
    r.metallic = __ldg(&m->metallic);
    return r;
}

// 3x4 column-major transform stored as m[c*3+r]
__device__ __forceinline__ vec3 xf_point(const float* m, vec3 p) {
    return mk3(((m[0] * p.x + m[3] * p.y) + m[6] * p.z) + m[9], ((m[1] * p.x + m[4] * p.y) + m[7] * p.z) + m[10],
               ((m[2] * p.x + m[5] * p.y) + m[8] * p.z) + m[11]);
}

struct Basis { vec3 b0, b1, b2; };
// math.glsl:9-15
__device__ __forceinline__ Basis create_basis(vec3 n) {
    Basis b;
    b.b2 = normalize3(n);
    vec3 a = (fabsf(b.b2.x) > 0.9f) ? mk3(0.0f, 1.0f, 0.0f) : mk3(1.0f, 0.0f, 0.0f);
    b.b1 = normalize3(cross3(b.b2, a));
    b.b0 = cross3(b.b2, b.b1);
    return b;
}
__device__ __forceinline__ vec3 world_to_local(vec3 v, const Basis& b) { return mk3(dot3(v, b.b0), dot3(v, b.b1), dot3(v, b.b2)); }
__device__ __forceinline__ vec3 local_to_world(vec3 v, const Basis& b) {
    return add3(add3(scale3(b.b0, v.x), scale3(b.b1, v.y)), scale3(b.b2, v.z));
}

__device__ __forceinline__ vec3 f0_of(const Mat& m) { return mix3(splat3(0.04f), m.albedo, m.metallic); }
// simple.rchit:69-75
__device__ __forceinline__ float specular_probability(const Mat& m) {
    float base = max3c(f0_of(m));
    float ri = smoothstep1(0.0f, 1.0f, m.roughness * 0.7f);
    return mix1(base, base * 0.5f, ri);
}
// simple.rchit:77-83
__device__ __forceinline__ float d_ggx(float NoH, float roughness) {
    float a = glsl_max(roughness, RT_MIN_ROUGHNESS);
    float a2 = a * a;
    float nh = glsl_clamp(NoH, 0.0f, 1.0f);
    float denom = (nh * nh) * (a2 - 1.0f) + 1.0f;
    return a2 / ((RT_PI * denom) * denom);
}
// simple.rchit:85-93
__device__ __forceinline__ float g_smith(float NoV, float NoL, float roughness) {
    float a = glsl_max(roughness, RT_MIN_ROUGHNESS);
    float k = a * 0.5f;
    float nv = glsl_clamp(NoV, RT_EPS_COS, 1.0f);
    float nl = glsl_clamp(NoL, RT_EPS_COS, 1.0f);
    float g1v = nv / (nv * (1.0f - k) + k);
    float g1l = nl / (nl * (1.0f - k) + k);
    return g1v * g1l;
}
// simple.rchit:96-98
__device__ __forceinline__ vec3 f_schlick(vec3 F0, float VoH) {
    return add3(F0, scale3(sub3(splat3(1.0f), F0), pow5f(1.0f - VoH)));
}
// simple.rchit:100-110
// (s, c) = sincos_rt(2 pi rx): both lobes turn the same random number into the same azimuth, so the caller evaluates it
// once ahead of the lobe branch instead of once per divergent side
__device__ __forceinline__ vec3 cosine_direction(float s, float c, float ry) {
    float ct = sqrtf(ry);
    float st = sqrtf(1.0f - ry);
    return mk3(c * st, s * st, ct);
}
// simple.rchit:118-166
// h = normalize(wo + wi), which the caller needs for the pdf as well
__device__ __forceinline__ vec3 eval_full_brdf(vec3 wo, vec3 wi, vec3 h, const Mat& m) {
    float NoL = wi.z, NoV = wo.z;
    if (NoL <= 0.0f || NoV <= 0.0f) return splat3(0.0f);
    float NoH = h.z;
    float VoH = dot3(wo, h);
    vec3 F = f_schlick(f0_of(m), VoH);
    vec3 da = scale3(m.albedo, 1.0f - m.metallic);
    vec3 diffuse = div3z(mul3(da, sub3(splat3(1.0f), F)), RT_PI);
    float D = d_ggx(NoH, m.roughness);
    float G = g_smith(NoV, NoL, m.roughness);
    vec3 specular = div3(scale3(F, D * G), (4.0f * NoV) * NoL);
    return add3(diffuse, specular);
}
// simple.rchit:195-200
__device__ __forceinline__ float microfacet_pdf(vec3 wo, vec3 h, float roughness) {
    float nh = glsl_max(h.z, RT_EPS_COS);
    float voh = glsl_max(dot3(wo, h), RT_EPS_VOH);
    float D = d_ggx(nh, roughness);
    return glsl_max((D * nh) / (4.0f * voh), RT_EPS_PDF);
}
// simple.rchit:202-217
__device__ __forceinline__ vec3 sample_ggx(float s, float c, float ry, float roughness) {
    float a = roughness * roughness;
    float ct = sqrtf((1.0f - ry) / (1.0f + (a * a - 1.0f) * ry));
    float st = sqrtf(1.0f - ct * ct);
    return mk3(st * c, st * s, ct);
}
// simple.rchit:234-237
__device__ __forceinline__ float mis_weight_power(float p1, float p2) {
    if (p1 <= 0.0f || p2 <= 0.0f) return 0.0f;
    return (p1 * p1) / (p1 * p1 + p2 * p2);
}

__device__ __forceinline__ vec3 vertex_pos(const rt_vertex* v) { return mk3(__ldg(&v->pos[0]), __ldg(&v->pos[1]), __ldg(&v->pos[2])); }
// position and normal of a 24-byte vertex as three 64-bit loads
__device__ __forceinline__ void load_vertex(const rt_vertex* v, vec3* pos, vec3* nrm) {
    const float2* q = reinterpret_cast<const float2*>(v);
    const float2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    *pos = mk3(a.x, a.y, b.x);
    *nrm = mk3(b.y, c.x, c.y);
}
__device__ __forceinline__ vec3 vertex_nrm(const rt_vertex* v) {
    return mk3(__ldg(&v->normal[0]), __ldg(&v->normal[1]), __ldg(&v->normal[2]));
}

// simple.rchit:507-520
__device__ __forceinline__ float light_weight(const SceneView& sc, vec3 x, uint32_t i) {
    const DevLight* l = sc.lights + i;
    vec3 centre = mk3(__ldg(&l->transform[9]), __ldg(&l->transform[10]), __ldg(&l->transform[11]));
    vec3 to = sub3(centre, x);
    float d2 = dot3(to, to);
    float f = 1.0f / glsl_max(d2, 0.001f);
    return __ldg(&l->emission_power) * f;
}

struct BsdfSample {
    vec3 direction, value;
    float pdf;
    bool is_specular;
};

// simple.rchit:403-449.  The reference evaluates each lobe inside its own branch; a warp holds both kinds of lane, so the
// parts the lobes share are hoisted out of the branch here — the cosine direction (diffuse lanes and the specular lanes
// that fall back to it), the half vector of the chosen direction (the Lambert lobe's and the pdf's are the same value),
// and the Schlick term (same F0, each lane's own cosine).  Every lane still performs exactly the reference's operations.
__device__ __forceinline__ BsdfSample sample_brdf(vec3 wo, const Mat& m, float rx, float ry, uint32_t& seed) {
    BsdfSample r;
    float sw = specular_probability(m);
    float sn, cs;
    sincos_rt(RT_TWO_PI * rx, &sn, &cs);
    r.direction = cosine_direction(sn, cs, ry);
    r.is_specular = false;
    vec3 hs = mk3(0.0f, 0.0f, 0.0f);
    if (rnd01(seed) < sw) {
        hs = sample_ggx(sn, cs, ry, m.roughness);
        vec3 wi = reflect3(neg3(wo), hs);
        if (!(wi.z <= 0.0f)) {
            r.is_specular = true;
            r.direction = wi;
        }
    }
    vec3 h = normalize3(add3(wo, r.direction));
    float woh = dot3(wo, h);
    vec3 F = f_schlick(f0_of(m), r.is_specular ? dot3(wo, hs) : glsl_clamp(woh, 0.0f, 1.0f));
    if (r.is_specular) {  // microfacetF, simple.rchit:168-193
        float NoL = r.direction.z, NoV = wo.z;
        if (NoL <= 0.0f || NoV <= 0.0f) {
            r.value = splat3(0.0f);
        } else {
            float D = d_ggx(hs.z, m.roughness);
            float G = g_smith(NoV, NoL, m.roughness);
            r.value = div3(scale3(F, D * G), (4.0f * NoV) * NoL);
        }
    } else {  // the Lambert lobe with the Fresnel complement, simple.rchit:416-422 and :434-440
        vec3 da = scale3(m.albedo, 1.0f - m.metallic);
        r.value = div3z(mul3(da, sub3(splat3(1.0f), F)), RT_PI);
    }
    float nh = glsl_max(h.z, RT_EPS_COS);  // microfacet_pdf(wo, h, roughness)
    float voh = glsl_max(woh, RT_EPS_VOH);
    float sp = glsl_max((d_ggx(nh, m.roughness) * nh) / (4.0f * voh), RT_EPS_PDF);
    float dp = glsl_max(r.direction.z, 0.0f) / RT_PI;
    r.pdf = sw * sp + (1.0f - sw) * dp;
    r.pdf = glsl_max(r.pdf, RT_EPS_PDF);
    return r;
}

// RayPayload of shaders/ray_common.glsl:13-26 (the fields that live across bounces)
struct PathState {
    vec3 color, throughput;
    vec3 prev_hit_pos;
    float prev_brdf_pdf, p_sample_light;
    uint32_t seed;      // payload.seed
    uint32_t rr_seed;   // the ray-gen shader's own copy of the seed (simple.rgen:71, used by :65)
    bool first_bounce, is_specular, did_direct;
};

struct ShadowRequest {
    bool pending;          // a visibility ray must be traced
    vec3 origin, dir;
    float tmax;            // tmin is 1e-3 (simple.rchit:368)
    uint32_t light_object; // hits on this instance do not occlude (simple.rchit:378-382)
    vec3 contrib;          // throughput * direct / p_sample_light, to add when visible
};

// A light sample that has been drawn but not evaluated yet (wavefront pipeline, see nee_select / nee_evaluate): everything
// after the draws of sampleLight is a pure function of these values and of the hit, so it can run in a later kernel.
struct NeeRequest {
    bool pending;
    uint32_t light;         // index of the selected light
    float sel_pdf;          // its selection probability
    float r_tri, q1, q2;    // the three draws of sampleLight (triangle, point)
    vec3 x, N, wo, throughput;  // hit point, shading normal, local view direction, throughput at the time of the sample
    uint32_t material;      // material index of the hit
};

struct ShadeResult {
    bool alive;           // payload.hit after the hit shader
    vec3 next_origin, next_dir;
    bool emissive_add;    // an emissive term follows the NEE term
    vec3 emissive;
};

// NEE: simple.rchit:543-583 (light selection) + :472-505 (evaluateLightMIS) + :239-322 (sampleLight)
// + the ray of :350-371, in two halves.  nee_select takes every random draw the reference takes (selection, triangle,
// point: no draw is skipped that the reference takes) and nee_evaluate turns them into the unshadowed contribution and the
// visibility ray; nothing in the second half feeds back into the path, so the wavefront pipeline runs it one thread per
// sample in its own kernel (wf_nee_kernel) instead of inside the divergent hit shader.
__device__ __forceinline__ bool nee_select(const SceneView& sc, vec3 x, uint32_t current_object, uint32_t& seed, uint32_t* light,
                                           float* sel_pdf, float* r_tri, float* q1, float* q2) {
    const uint32_t L = sc.n_lights;
    if (L == 0u) return false;
    const uint32_t maxL = L < RT_MAXLIGHTS ? L : RT_MAXLIGHTS;
    float total = 0.0f;
    for (uint32_t i = 0; i < maxL; i++) {
        if (current_object == __ldg(&sc.lights[i].object_index)) continue;
        total += light_weight(sc, x, i);
    }
    if (total <= 0.0f) return false;
    float r1 = rnd01(seed) * total;
    uint32_t sel = 0xFFFFFFFFu;
    float wsel = 0.0f;
    for (uint32_t i = 0; i < maxL; i++) {
        // the skipped light's weight is DEFINED as 0 (uninitialised in simple.rchit:555-574)
        float w = (current_object == __ldg(&sc.lights[i].object_index)) ? 0.0f : light_weight(sc, x, i);
        r1 -= w;
        if (r1 <= 0.0f) { sel = i; wsel = w; break; }
    }
    if (sel == 0xFFFFFFFFu) return false;
    *light = sel;
    *sel_pdf = wsel / total;
    *r_tri = rnd01(seed);
    *q1 = rnd01(seed);
    *q2 = rnd01(seed);
    return true;
}

__device__ __forceinline__ bool nee_evaluate(const SceneView& sc, vec3 x, vec3 N, const Basis& basis, vec3 wo, const Mat& mat,
                                             uint32_t sel, float sel_pdf, float r_tri, float q1, float q2, vec3* direct,
                                             ShadowRequest* sr) {
    // ---- sampleLight ----
    const DevLight* light = sc.lights + sel;
    const uint32_t lobj = __ldg(&light->object_index);
    const DevObject* lo = sc.objects + lobj;
    const uint32_t ntri = __ldg(&light->num_triangles);
    if (ntri == 0u) return false;  // DEFINED: a light whose mesh is empty gives an invalid sample (the shader would index triangle 0xFFFFFFFF)
    uint32_t tri = (uint32_t)(r_tri * (float)ntri);
    if (tri > ntri - 1u) tri = ntri - 1u;
    const uint32_t* ip = sc.indices + __ldg(&lo->index_base) + 3u * tri;
    const rt_vertex* vb = sc.vertices + __ldg(&lo->vertex_base);
    vec3 p0 = vertex_pos(vb + __ldg(ip)), p1 = vertex_pos(vb + __ldg(ip + 1)), p2 = vertex_pos(vb + __ldg(ip + 2));
    float sq = sqrtf(q1);
    float bu = 1.0f - sq, bv = sq * (1.0f - q2), bw = sq * q2;
    vec3 local_pos = add3(add3(scale3(p0, bu), scale3(p1, bv)), scale3(p2, bw));
    float tr[12];
#pragma unroll
    for (int k = 0; k < 12; k++) tr[k] = __ldg(&light->transform[k]);
    vec3 w0 = xf_point(tr, p0), w1 = xf_point(tr, p1), w2 = xf_point(tr, p2);
    vec3 wp = xf_point(tr, local_pos);
    vec3 cr = cross3(sub3(w1, w0), sub3(w2, w0));
    vec3 ln = normalize3(cr);
    // normalize(x - wp) of simple.rchit:283 is -normalize(wp - x) bit for bit (negation commutes with every rounding)
    vec3 tl = normalize3(sub3(wp, x));
    float cosL = dot3(ln, neg3(tl));
    if (cosL < 0.0f) { cosL = fabsf(cosL); ln = neg3(ln); }
    if (!(cosL > 0.0f)) return false;
    vec3 to_light = sub3(wp, x);
    float dist = glsl_max(length3(to_light), 0.01f);
    vec3 ldir = div3(to_light, dist);
    float area = 0.5f * length3(cr);
    float cos_t = glsl_max(0.0f, dot3(neg3(ldir), ln));
    if (cos_t <= 1e-6f) return false;
    float area_pdf = 1.0f / area;
    float tri_pdf = 1.0f / (float)ntri;
    float lpdf_geo = ((tri_pdf * area_pdf) * (dist * dist)) / cos_t;
    const rt_material* lm = sc.materials + __ldg(&lo->material_index);
    vec3 Le = scale3(mk3(__ldg(&lm->emission_color[0]), __ldg(&lm->emission_color[1]), __ldg(&lm->emission_color[2])),
                     __ldg(&lm->emission_power));
    // ---- evaluateLightMIS ----
    vec3 wi = world_to_local(ldir, basis);
    if (!(wi.z > 1e-4f)) return false;
    // isVisibleRQ ray (simple.rchit:350-371)
    vec3 from = add3(x, scale3(scale3(N, 0.001f), sign1(dot3(N, tl))));
    vec3 sdir = sub3(wp, from);
    float sdist = length3(sdir);
    if (sdist <= 0.0f) return false;
    sdir = div3(sdir, sdist);
    vec3 h = normalize3(add3(wo, wi));
    vec3 brdf = eval_full_brdf(wo, wi, h, mat);
    float light_pdf = lpdf_geo * sel_pdf;
    float sp = specular_probability(mat);
    float spec_pdf = microfacet_pdf(wo, h, mat.roughness);
    float diff_pdf = wi.z / RT_PI;
    float brdf_pdf = sp * spec_pdf + (1.0f - sp) * diff_pdf;
    float w = mis_weight_power(light_pdf, brdf_pdf);
    *direct = div3z(scale3(scale3(mul3(brdf, Le), wi.z), w), glsl_max(light_pdf, 1e-6f));
    sr->origin = from;
    sr->dir = sdir;
    sr->tmax = sdist * 0.999f;
    sr->light_object = lobj;
    return true;
}

__device__ __forceinline__ bool nee_sample(const SceneView& sc, vec3 x, vec3 N, const Basis& basis, vec3 wo, const Mat& mat,
                                           uint32_t current_object, uint32_t& seed, vec3* direct, ShadowRequest* sr,
                                           bool* sampled) {
    uint32_t sel;
    float sel_pdf, r_tri, q1, q2;
    *sampled = nee_select(sc, x, current_object, seed, &sel, &sel_pdf, &r_tri, &q1, &q2);
    if (!*sampled) return false;
    return nee_evaluate(sc, x, N, basis, wo, mat, sel, sel_pdf, r_tri, q1, q2, direct, sr);
}

// simple.rchit:536-541 with :522-534 (no light is skipped here)
__device__ __forceinline__ float light_selection_pdf(const SceneView& sc, vec3 x, uint32_t light_idx) {
    float w = light_weight(sc, x, light_idx);
    const uint32_t L = sc.n_lights;
    const uint32_t maxL = L < RT_MAXLIGHTS ? L : RT_MAXLIGHTS;
    float total = 0.0f;
    for (uint32_t i = 0; i < maxL; i++) total += light_weight(sc, x, i);
    if (total <= 0.0f) return 0.0f;
    return w / total;
}

// simple.rchit:590-706.  `flags`: RT_FLAG_DIRECT_LIGHTING / RT_FLAG_MIS (both set = the shipped shader).
// LIGHTS = false: the variant for scenes without lights (sc.n_lights == 0, hence no emissive material either: a light IS
// an object whose material has emission_power > 0, gpu_scene.odin:366-379).  The NEE and emissive-MIS code drops out at
// compile time — the same draws are taken and the same values produced, in a kernel half the size.
// DEFER = true (wavefront pipeline): the light sample is drawn here and handed back in `nq` for wf_nee_kernel to evaluate.
template <bool LIGHTS = true, bool DEFER = false>
__device__ __forceinline__ ShadeResult shade_hit(const SceneView& sc, PathState& ps, const Hit& hit, vec3 ray_dir,
                                                 ShadowRequest& sr, uint32_t& n_nee, NeeRequest* nq = nullptr) {
    ShadeResult out;
    sr.pending = false;
    if (DEFER) nq->pending = false;
    out.emissive_add = false;
    const DevObject* obj = sc.objects + hit.inst;
    // The object record and the vertices with wide loads: after the first bounce every lane sits in another object and
    // triangle, and such a load costs the L1 pipe the same whatever its width (the bounce-1 launch ran at 80 % of that pipe
    // with 56 scalar loads per hit: ncu r1q).  DevObject: words 0..11 model, 12..20 nrm, 21 material, 22 light, 23 vertex base.
    // The variant with the light-sampling code has no registers to spare for 24 words in flight (measured: 0.64 -> 0.74 ms
    // per frame on config 4b) and keeps the scalar loads.
    float model[12], n9[9];
    uint32_t material_index, light_index, vertex_base;
    if (!LIGHTS) {
        uint4 oa, ob, oc, od, oe, of;
        ldg256(reinterpret_cast<const uint4*>(obj), oa, ob);
        ldg256(reinterpret_cast<const uint4*>(obj) + 2, oc, od);
        ldg256(reinterpret_cast<const uint4*>(obj) + 4, oe, of);
        const uint32_t w[21] = {oa.x, oa.y, oa.z, oa.w, ob.x, ob.y, ob.z, ob.w, oc.x, oc.y, oc.z, oc.w, od.x, od.y, od.z, od.w, oe.x, oe.y, oe.z, oe.w, of.x};
#pragma unroll
        for (int k = 0; k < 12; k++) model[k] = __uint_as_float(w[k]);
#pragma unroll
        for (int k = 0; k < 9; k++) n9[k] = __uint_as_float(w[12 + k]);
        material_index = of.y; light_index = of.z; vertex_base = of.w;
    } else {
#pragma unroll
        for (int k = 0; k < 12; k++) model[k] = __ldg(&obj->model[k]);
#pragma unroll
        for (int k = 0; k < 9; k++) n9[k] = __ldg(&obj->nrm[k]);
        material_index = __ldg(&obj->material_index); light_index = __ldg(&obj->light_index); vertex_base = __ldg(&obj->vertex_base);
    }
    const Mat mat = load_material(sc.materials + material_index);
    const uint32_t* ip = sc.indices + __ldg(&obj->index_base) + 3u * hit.prim;
    const rt_vertex* vb = sc.vertices + vertex_base;
    vec3 p0, p1, p2, m0, m1, m2;
    if (!LIGHTS) {
        load_vertex(vb + __ldg(ip), &p0, &m0);
        load_vertex(vb + __ldg(ip + 1), &p1, &m1);
        load_vertex(vb + __ldg(ip + 2), &p2, &m2);
    } else {
        const rt_vertex *v0 = vb + __ldg(ip), *v1 = vb + __ldg(ip + 1), *v2 = vb + __ldg(ip + 2);
        p0 = vertex_pos(v0); p1 = vertex_pos(v1); p2 = vertex_pos(v2);
        m0 = vertex_nrm(v0); m1 = vertex_nrm(v1); m2 = vertex_nrm(v2);
    }
    const float bx = (1.0f - hit.u) - hit.v, by = hit.u, bz = hit.v;
    vec3 pos = add3(add3(scale3(p0, bx), scale3(p1, by)), scale3(p2, bz));
    vec3 world_pos = xf_point(model, pos);
    vec3 nrm = normalize3(add3(add3(scale3(m0, bx), scale3(m1, by)), scale3(m2, bz)));
    vec3 N = normalize3(mat3_mul(n9, nrm));
    if (!(dot3(N, neg3(ray_dir)) > 0.0f)) N = neg3(N);
    const bool is_emissive = mat.emission_power > 0.0f;

    bool did_direct = false;
    const float psl = glsl_clamp(mat.roughness, 0.1f, 0.9f);
    const float p = rnd01(ps.seed);
    const Basis basis = create_basis(N);
    const vec3 wo = world_to_local(neg3(ray_dir), basis);
    if (p < psl) {
        if (LIGHTS && DEFER) {
            if (nee_select(sc, world_pos, hit.inst, ps.seed, &nq->light, &nq->sel_pdf, &nq->r_tri, &nq->q1, &nq->q2)) {
                nq->pending = true;
                nq->x = world_pos; nq->N = N; nq->wo = wo; nq->throughput = ps.throughput;
                nq->material = material_index;
                n_nee++;
            }
        } else if (LIGHTS) {
            vec3 direct;
            bool sampled;
            if (nee_sample(sc, world_pos, N, basis, wo, mat, hit.inst, ps.seed, &direct, &sr, &sampled)) {
                sr.pending = true;
                sr.contrib = div3z(mul3(ps.throughput, direct), psl);
            }
            if (sampled) n_nee++;
        }
        did_direct = true;
    }

    float rx = rnd01(ps.seed);
    float ry = rnd01(ps.seed);
    BsdfSample bs = sample_brdf(wo, mat, rx, ry, ps.seed);

    if (LIGHTS && is_emissive) {
        if (ps.first_bounce || ps.is_specular) {
            out.emissive_add = true;
            out.emissive = scale3(mul3(ps.throughput, mat.emission_color), mat.emission_power);
        } else if (!ps.did_direct) {
            const uint32_t li = light_index;
            const DevLight* light = sc.lights + li;
            float tr[12];
#pragma unroll
            for (int k = 0; k < 12; k++) tr[k] = __ldg(&light->transform[k]);
            float dist = length3(sub3(world_pos, ps.prev_hit_pos));
            float dist2 = dist * dist;
            float cos_l = glsl_max(0.0f, dot3(N, neg3(ray_dir)));
            vec3 w0 = xf_point(tr, p0), w1 = xf_point(tr, p1), w2 = xf_point(tr, p2);
            float area = 0.5f * length3(cross3(sub3(w1, w0), sub3(w2, w0)));
            float area_pdf = 1.0f / area;
            float tri_pdf = 1.0f / (float)__ldg(&light->num_triangles);
            float pdf_geo = ((tri_pdf * area_pdf) * dist2) / cos_l;
            float sel = light_selection_pdf(sc, world_pos, li);
            float light_pdf = sel * pdf_geo;
            float w = mis_weight_power(ps.prev_brdf_pdf, light_pdf);
            out.emissive_add = true;
            out.emissive = div3(scale3(scale3(mul3(ps.throughput, mat.emission_color), mat.emission_power), w),
                                1.0f - ps.p_sample_light);
        }
    }
    ps.did_direct = did_direct;
    ps.p_sample_light = psl;

    if (bs.pdf > 0.0f && bs.direction.z > 0.0f) {
        ps.throughput = mul3(ps.throughput, div3z(scale3(bs.value, bs.direction.z), bs.pdf));
        out.next_dir = local_to_world(bs.direction, basis);
        out.next_origin = world_pos;
        out.alive = true;
        ps.is_specular = bs.is_specular;
        ps.prev_brdf_pdf = bs.pdf;
        ps.prev_hit_pos = world_pos;
    } else {
        out.alive = false;
    }
    ps.first_bounce = false;
    return out;
}

// simple.rgen:41-53; px,py = pixel + jitter.  restir.rgen:47-57 uses the same matrices without the
// final normalisation (renormalise = false).
__device__ __forceinline__ void camera_ray(const FrameParams& fp, float px, float py, bool renormalise, vec3* o, vec3* d) {
    const float* iv = fp.inv_view;
    const float* ip = fp.inv_proj;
    float ux = px / (float)fp.width, uy = py / (float)fp.height;
    float dx = ux * 2.0f - 1.0f, dy = uy * 2.0f - 1.0f;
    *o = mk3(iv[12], iv[13], iv[14]);
    vec3 target = mk3(((ip[0] * dx + ip[4] * dy) + ip[8]) + ip[12], ((ip[1] * dx + ip[5] * dy) + ip[9]) + ip[13],
                      ((ip[2] * dx + ip[6] * dy) + ip[10]) + ip[14]);
    vec3 dl = mat_dir(iv, normalize3(target));
    *d = renormalise ? normalize3(dl) : dl;
}

// simple.rgen:70-85: seed, jitter, primary ray, payload initialisation
__device__ __forceinline__ void raygen_ray(const FrameParams& fp, uint32_t frame, uint32_t x, uint32_t y, uint32_t* seed_out, vec3* o, vec3* d) {
    uint32_t seed = tea16(y * fp.width + x, frame);
    float jx = 0.5f, jy = 0.5f;
    if (frame != 0u) {
        float r1 = rnd01(seed);
        float r2 = rnd01(seed);
        jx = 0.5f + (r1 - 0.5f) * 0.4f;
        jy = 0.5f + (r2 - 0.5f) * 0.4f;
    }
    camera_ray(fp, (float)x + jx, (float)y + jy, true, o, d);
    *seed_out = seed;
}

__device__ __forceinline__ void raygen_init(const FrameParams& fp, uint32_t x, uint32_t y, PathState& ps, vec3* o, vec3* d) {
    uint32_t seed;
    raygen_ray(fp, fp.frame, x, y, &seed, o, d);
    ps.seed = seed;
    ps.rr_seed = seed;
    ps.color = splat3(0.0f);
    ps.throughput = splat3(1.0f);
    ps.first_bounce = true;
    ps.is_specular = false;
    ps.prev_brdf_pdf = 1.0f;
    ps.did_direct = false;
    ps.p_sample_light = 0.0f;
    ps.prev_hit_pos = splat3(0.0f);
}

// simple.rgen:55-68; true = the path continues
__device__ __forceinline__ bool russian_roulette(PathState& ps, int depth, int rr_start) {
    if (depth < rr_start) return true;
    float lum = dot3(ps.throughput, mk3(0.2126f, 0.7152f, 0.0722f));
    float p = glsl_min(glsl_max(lum, 0.05f), 0.95f);
    if (rnd01(ps.rr_seed) > p) return false;
    ps.throughput = div3(ps.throughput, p);
    return true;
}

// simple.rgen:115-118
__device__ __forceinline__ bool throughput_dead(vec3 T) {
    return any_nan3(T) || any_inf3(T) || (T.x < 0.001f && T.y < 0.001f && T.z < 0.001f);
}

// simple.rgen:121-136: clamp, scrub, progressive mean
__device__ __forceinline__ float4 accumulate_pixel(float4 old, vec3 color, float clampv, uint32_t frame) {
    vec3 fc = mk3(glsl_min(color.x, clampv), glsl_min(color.y, clampv), glsl_min(color.z, clampv));
    if (any_nan3(fc) || any_inf3(fc)) fc = splat3(0.0f);
    if (frame == 0u) return make_float4(fc.x, fc.y, fc.z, 1.0f);
    float a = 1.0f / (float)(frame + 1u);
    return make_float4(mix1(old.x, fc.x, a), mix1(old.y, fc.y, a), mix1(old.z, fc.z, a), 1.0f);
}

}  // namespace rt
