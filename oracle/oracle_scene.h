// ORACLE — test infrastructure only (see oracle_math.h header).  Parity status: pinned to the shipped SPIR-V (tests/golden/spirv_frames.npz); driver-defined parts unpinned.
//
// Scene tables as the reference's shaders see them, derived from the host arrays exactly like
// src/raytracer/gpu_scene.odin does, plus the flattened world-space triangle list that stands in
// for the driver-owned BLAS/TLAS (acceleration_structure.odin:32-143; opaque to the app).
#pragma once
#include <cstdint>
#include <vector>

#include "oracle_math.h"

namespace orc {

struct Vertex {  // scene.odin:10-13, ray_common.glsl:48-51 (24 B)
    float pos[3];
    float normal[3];
};
struct Material {  // gpu_scene.odin:17-21, ray_common.glsl:3-11 (44 B)
    float albedo[3];
    float emission_color[3];
    float emission_power, roughness, metallic, transmission, ior;
};
struct Mesh {
    std::vector<Vertex> v;
    std::vector<uint32_t> idx;
};
struct Object {  // Object_GPU_Data (gpu_scene.odin:23-29) + the instance transform
    uint32_t mesh_index, material_index, light_index;
    float model[16];  // column-major ObjectToWorld
    float nrm[9];     // transpose(inverse(mat3(model))) columns
    uint32_t tri_offset;  // first flat triangle id
};
struct Light {  // Light_GPU_Data (gpu_scene.odin:31-36)
    float transform[16];
    uint32_t object_index, num_triangles, instance_mask;
};
struct FlatTri {
    V3 p0, p1, p2;  // world space: xform_point(model, v.pos)
    uint32_t inst, prim;
};

struct Hit {
    float t, u, v;   // u,v = barycentric weights of vertex 1 and 2 (hitAttributeEXT attribs)
    uint32_t id;     // flat triangle id, 0xFFFFFFFF = miss
};

// 128-byte compressed wide node (own layout): grid coordinates of the child boxes as binary16
#define ORC_QMAX 2046
#define ORC_QSTEPS 2044.0f
struct Node8 {
    float p[3];
    uint8_t e[3];
    uint8_t imask;
    uint32_t child_base;
    uint32_t tri_base;
    uint8_t meta[8];
    uint16_t q[3][2][8];  // [axis][0 = lo, 1 = hi][slot]
};
static_assert(sizeof(Node8) == 128, "Node8 must be 128 bytes");

struct BvhTri {  // 48 B: three float4
    float p0[3]; uint32_t id;
    float p1[3]; uint32_t inst;
    float p2[3]; uint32_t prim;
};
static_assert(sizeof(BvhTri) == 48, "BvhTri must be 48 bytes");

struct Counters {
    uint64_t extend_rays, shadow_rays, nodes, tris, hits, nee;
};

struct Scene {
    std::vector<Mesh> meshes;
    std::vector<Object> objects;
    std::vector<Material> materials;
    std::vector<Light> lights;
    std::vector<FlatTri> tris;
    // BVH8
    std::vector<Node8> nodes;
    std::vector<BvhTri> btris;
    std::vector<uint64_t> morton;      // sorted keys (debug/inspection)
    std::vector<uint32_t> sorted_ids;  // flat ids in sorted order
    uint32_t bvh_depth = 0;
    bool has_bvh = false;
};

// build steps
int scene_finalize(Scene& s);   // objects' derived data, lights, flat triangles
void build_bvh8(Scene& s);

// traversal
Hit closest_brute(const Scene& s, V3 o, V3 d, float tmin, float tmax);
bool occluded_brute(const Scene& s, V3 o, V3 d, float tmin, float tmax, uint32_t ignore_inst);
Hit closest_bvh(const Scene& s, V3 o, V3 d, float tmin, float tmax, Counters* c);
bool occluded_bvh(const Scene& s, V3 o, V3 d, float tmin, float tmax, uint32_t ignore_inst, Counters* c);

}  // namespace orc
