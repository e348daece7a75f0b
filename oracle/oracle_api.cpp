// ORACLE — test infrastructure only (see oracle_math.h header).  Parity status: pinned to the shipped SPIR-V (tests/golden/spirv_frames.npz); driver-defined parts unpinned.
// C entry points for ctypes (tests/, bench.py cpu_baseline / --impl reference, smoke()).
#include <cstdio>
#include <cstdlib>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "oracle_scene.h"

namespace orc {
struct Cfg { int max_depth; float clamp; int use_bvh; };
void render_rows(const Scene& s, const Cfg& cfg, const float* cam, const float* clear, uint32_t frame, uint32_t W,
                 uint32_t H, uint32_t y0, uint32_t y1, float* image_rgba, Counters* total);
void direct_reference_rows(const Scene& s, int use_bvh, const float* cam, const float* clear, uint32_t W, uint32_t H,
                           uint32_t y0, uint32_t y1, int spt, float* image_rgba);
void gbuffer_rows(const Scene& s, int use_bvh, const float* cam, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1,
                  uint8_t* albedo_rgba8, uint16_t* normal_rough_f16, float* position_metal_f32, uint16_t* emission_f16,
                  uint8_t* output_rgba8);
void raster_gbuffer(const Scene& s, const float* cam, uint32_t W, uint32_t H, float* albedo, float* normal_metal, float* position_rough,
                    float* emission, float* props);
void unit_raster_vertex(const float* model16, const float* pos3, const float* normal3, const float* cam, float out[10]);
void unit_sample_brdf(const Material& m, V3 wo, float rx, float ry, uint32_t* seed, float out[8]);
void unit_eval_brdf(const Material& m, V3 wo, V3 wi, float out[4]);
float unit_specular_probability(const Material& m);
}  // namespace orc

using namespace orc;

extern "C" {

void* orc_scene_create(uint32_t n_meshes, const void* const* vtx, const uint32_t* n_vtx, const uint32_t* const* idx,
                       const uint32_t* n_idx, uint32_t n_objects, const uint32_t* obj_mesh, const uint32_t* obj_mat,
                       const float* obj_model16, uint32_t n_materials, const void* materials44, int* err) {
    Scene* s = new Scene();
    s->meshes.resize(n_meshes);
    for (uint32_t i = 0; i < n_meshes; i++) {
        s->meshes[i].v.assign((const Vertex*)vtx[i], (const Vertex*)vtx[i] + n_vtx[i]);
        s->meshes[i].idx.assign(idx[i], idx[i] + n_idx[i]);
    }
    s->materials.assign((const Material*)materials44, (const Material*)materials44 + n_materials);
    s->objects.resize(n_objects);
    for (uint32_t i = 0; i < n_objects; i++) {
        s->objects[i].mesh_index = obj_mesh[i];
        s->objects[i].material_index = obj_mat[i];
        memcpy(s->objects[i].model, obj_model16 + 16 * (size_t)i, 64);
    }
    int rc = scene_finalize(*s);
    if (err) *err = rc;
    if (rc != 0) { delete s; return nullptr; }
    return s;
}
void orc_scene_destroy(void* h) { delete (Scene*)h; }
int orc_build_bvh(void* h) { build_bvh8(*(Scene*)h); return 0; }
uint32_t orc_num_nodes(void* h) { return (uint32_t)((Scene*)h)->nodes.size(); }
uint32_t orc_num_tris(void* h) { return (uint32_t)((Scene*)h)->tris.size(); }
uint32_t orc_num_lights(void* h) { return (uint32_t)((Scene*)h)->lights.size(); }
uint32_t orc_bvh_depth(void* h) { return ((Scene*)h)->bvh_depth; }
void orc_get_nodes(void* h, void* dst) { Scene* s = (Scene*)h; memcpy(dst, s->nodes.data(), s->nodes.size() * sizeof(Node8)); }
void orc_get_btris(void* h, void* dst) { Scene* s = (Scene*)h; memcpy(dst, s->btris.data(), s->btris.size() * sizeof(BvhTri)); }
void orc_get_morton(void* h, uint64_t* keys, uint32_t* ids) {
    Scene* s = (Scene*)h;
    memcpy(keys, s->morton.data(), s->morton.size() * 8);
    memcpy(ids, s->sorted_ids.data(), s->sorted_ids.size() * 4);
}
void orc_get_lights(void* h, void* dst76) { Scene* s = (Scene*)h; memcpy(dst76, s->lights.data(), s->lights.size() * sizeof(Light)); }
// gpu_scene_update_object (gpu_scene.odin:517-536) is the only place light_index becomes non-zero
void orc_set_light_index(void* h, uint32_t object, uint32_t light_index) { ((Scene*)h)->objects[object].light_index = light_index; }
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

int orc_render(void* h, const float* cam64, const float* clear3, uint32_t frame, uint32_t W, uint32_t H, uint32_t y0,
               uint32_t y1, int max_depth, float clamp, int use_bvh, float* image_rgba, uint64_t* counters6) {
    Scene* s = (Scene*)h;
    if (use_bvh && !s->has_bvh) build_bvh8(*s);
    Cfg cfg{max_depth, clamp, use_bvh};
    Counters c{0, 0, 0, 0, 0, 0};
    render_rows(*s, cfg, cam64, clear3, frame, W, H, y0, y1, image_rgba, counters6 ? &c : nullptr);
    if (counters6) {
        counters6[0] += c.extend_rays; counters6[1] += c.shadow_rays; counters6[2] += c.nodes;
        counters6[3] += c.tris; counters6[4] += c.hits; counters6[5] += c.nee;
    }
    return 0;
}

int orc_gbuffer(void* h, const float* cam64, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1, int use_bvh, uint8_t* albedo,
                uint16_t* normal_rough, float* position_metal, uint16_t* emission, uint8_t* output) {
    Scene* s = (Scene*)h;
    if (use_bvh && !s->has_bvh) build_bvh8(*s);
    gbuffer_rows(*s, use_bvh, cam64, W, H, y0, y1, albedo, normal_rough, position_metal, emission, output);
    return 0;
}

int orc_direct_reference(void* h, const float* cam64, const float* clear3, uint32_t W, uint32_t H, int use_bvh, int spt, float* image_rgba) {
    Scene* s = (Scene*)h;
    if (use_bvh && !s->has_bvh) build_bvh8(*s);
    direct_reference_rows(*s, use_bvh, cam64, clear3, W, H, 0, H, spt, image_rgba);
    return 0;
}

int orc_raster_gbuffer(void* h, const float* cam64, uint32_t W, uint32_t H, float* albedo, float* normal_metal, float* position_rough,
                       float* emission, float* props) {
    raster_gbuffer(*(Scene*)h, cam64, W, H, albedo, normal_metal, position_rough, emission, props);
    return 0;
}
void orc_unit_raster_vertex(const float* model16, const float* pos3, const float* normal3, const float* cam64, float* out10) {
    unit_raster_vertex(model16, pos3, normal3, cam64, out10);
}

// the same on the rows y0 .. y1-1 only (full-size image; the other rows are left alone): full-resolution checks of a band
int orc_direct_reference_rows(void* h, const float* cam64, const float* clear3, uint32_t W, uint32_t H, uint32_t y0, uint32_t y1,
                              int use_bvh, int spt, float* image_rgba) {
    Scene* s = (Scene*)h;
    if (use_bvh && !s->has_bvh) build_bvh8(*s);
    direct_reference_rows(*s, use_bvh, cam64, clear3, W, H, y0, y1 < H ? y1 : H, spt, image_rgba);
    return 0;
}

// rays: n x 8 floats (ox,oy,oz,tmin, dx,dy,dz,tmax). mode: 0 brute closest, 1 bvh closest, 2 brute any, 3 bvh any.
// out_tuv: n x 3 (closest modes), out_id: flat id or 0xFFFFFFFF (closest); 1/0 occluded flag (any modes).
int orc_trace_rays(void* h, const float* rays, uint32_t n, int mode, const uint32_t* ignore_inst, float* out_tuv,
                   uint32_t* out_id, uint64_t* counters6) {
    Scene* s = (Scene*)h;
    if ((mode & 1) && !s->has_bvh) build_bvh8(*s);
    Counters tot{0, 0, 0, 0, 0, 0};
#pragma omp parallel
    {
        Counters c{0, 0, 0, 0, 0, 0};
#pragma omp for schedule(dynamic, 64)
        for (int i = 0; i < (int)n; i++) {
            const float* r = rays + 8 * (size_t)i;
            V3 o = v3(r[0], r[1], r[2]), d = v3(r[4], r[5], r[6]);
            if (mode < 2) {
                Hit hit = mode == 1 ? closest_bvh(*s, o, d, r[3], r[7], &c) : closest_brute(*s, o, d, r[3], r[7]);
                out_id[i] = hit.id;
                if (out_tuv) {
                    bool ok = hit.id != 0xFFFFFFFFu;
                    out_tuv[3 * i] = ok ? hit.t : 0.0f; out_tuv[3 * i + 1] = ok ? hit.u : 0.0f; out_tuv[3 * i + 2] = ok ? hit.v : 0.0f;
                }
            } else {
                uint32_t ig = ignore_inst ? ignore_inst[i] : 0xFFFFFFFFu;
                bool occ = mode == 3 ? occluded_bvh(*s, o, d, r[3], r[7], ig, &c) : occluded_brute(*s, o, d, r[3], r[7], ig);
                out_id[i] = occ ? 1u : 0u;
            }
        }
#pragma omp critical
        { tot.nodes += c.nodes; tot.tris += c.tris; }
    }
    if (counters6) { counters6[2] += tot.nodes; counters6[3] += tot.tris; }
    return 0;
}

uint32_t orc_tea(uint32_t a, uint32_t b) { return tea(a, b); }
void orc_rnd_seq(uint32_t seed, uint32_t n, float* out, uint32_t* seed_out) {
    for (uint32_t i = 0; i < n; i++) out[i] = rnd(seed);
    if (seed_out) *seed_out = seed;
}
void orc_sincos(const float* a, uint32_t n, float* s, float* c) {
    for (uint32_t i = 0; i < n; i++) sincos_contract(a[i], &s[i], &c[i]);
}
void orc_unit_sample_brdf(const void* material44, const float* wo3, float rx, float ry, uint32_t* seed, float* out8) {
    unit_sample_brdf(*(const Material*)material44, v3(wo3[0], wo3[1], wo3[2]), rx, ry, seed, out8);
}
float orc_unit_specular_probability(const void* material44) { return unit_specular_probability(*(const Material*)material44); }
void orc_unit_eval_brdf(const void* material44, const float* wo3, const float* wi3, float* out4) {
    unit_eval_brdf(*(const Material*)material44, v3(wo3[0], wo3[1], wo3[2]), v3(wi3[0], wi3[1], wi3[2]), out4);
}

}  // extern "C"
