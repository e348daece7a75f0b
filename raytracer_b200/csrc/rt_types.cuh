// rt_types.cuh — device-side record layouts of librt.so.
//
// Input layouts are the reference's (include/rt_api.h); the derived tables below are what the
// shading kernels read instead of the Vulkan descriptors of the reference:
//   ObjectData  (shaders/ray_common.glsl:34-40, gpu_scene.odin:23-29)  -> DevObject
//   LightData   (shaders/ray_common.glsl:42-46, gpu_scene.odin:31-36)  -> DevLight
//   TLAS/BLAS   (acceleration_structure.odin:65-199, driver-opaque)    -> Node8 + BvhTri
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "../../include/rt_api.h"
#include "rt_math.cuh"

// traversal stack entries per ray; a BVH8 deeper than this is rejected at build time
#define RT_STACK_SIZE 32
#define RT_MAX_BVH_DEPTH RT_STACK_SIZE

namespace rt {

// 128-byte compressed 8-wide node = one cache line = eight 16-byte words:
//   word 0: p.x p.y p.z | e.x e.y e.z imask
//   word 1: child_base tri_base meta[0..3] meta[4..7]
//   word 2: lo.x[8]   word 3: hi.x[8]   word 4: lo.y[8]   word 5: hi.y[8]   word 6: lo.z[8]   word 7: hi.z[8]
// Child box on axis a = p[a] + 2^(e[a]-127) * [lo, hi], the grid coordinates lo/hi in 0..RT_QMAX stored as
// binary16 (exact for integers <= 2048).  binary16 rather than bytes because the B200 converts half->float on
// the FMA pipe (HADD2.F32, free .H0/.H1 operand select) while byte->float (I2F.U8) runs on the 16-lane XU
// pipe, which the profile of the byte format showed to be the busiest pipe of the traversal (profiles/).
// The near / far plane of an axis is picked by ADDRESS (word 2 + 2a + sign), not by a select per node.
// meta: 0 = empty slot; internal child = 0x20 | (24 + slot) (one bit, position = slot); leaf = unary
// triangle count << 5 | offset from tri_base.
#define RT_QMAX 2046
#define RT_QSTEPS 2044.0f
struct __align__(16) Node8 {
    float p[3];
    uint8_t e[3];
    uint8_t imask;
    uint32_t child_base;
    uint32_t tri_base;
    uint8_t meta[8];
    uint16_t q[3][2][8];  // [axis][0 = lo, 1 = hi][slot], binary16 bits
};
static_assert(sizeof(Node8) == 128, "Node8 is 128 bytes");
#define RT_NODE_WORDS 8u
// In memory the four 32-byte chunks of a node (header, x planes, y planes, z planes) are rotated by the node's
// index: chunk c of node n sits at chunk position (c + n) & 3.  Incoherent rays fetch chunk c of 32 DIFFERENT
// nodes with one LDG.256; unrotated, all 32 accesses fall on the same eight L1 banks and are served one thread
// per cycle, rotated they spread over the four bank groups (tools/mb/gather2.cu: 128.8 -> 70.1 cycles per warp
// and node, L1-resident).  Struct Node8 above is the UNROTATED form (builder registers, rt_get_bvh, oracle).
#ifndef RT_NODE_ROTATE
#define RT_NODE_ROTATE 1
#endif
__host__ __device__ __forceinline__ uint32_t node_chunk(uint32_t node, uint32_t c) {  // in uint4 units
#if RT_NODE_ROTATE
    return ((c + node) & 3u) * 2u;
#else
    (void)node;
    return c * 2u;
#endif
}

// 48-byte triangle: three float4 (world-space vertices + ids), fetched as LDG.256 + LDG.128 (ldg_tri below).
struct __align__(16) BvhTri {
    float p0[3]; uint32_t id;    // flat triangle id = object.tri_offset + primitive
    float p1[3]; uint32_t inst;  // object index (instanceCustomIndex, gpu_scene.odin:135)
    float p2[3]; uint32_t prim;  // gl_PrimitiveID
};
static_assert(sizeof(BvhTri) == 48, "BvhTri is 48 bytes");

struct __align__(16) DevObject {  // 128 B
    float model[12];      // columns 0..3 of ObjectToWorld, 3 rows each (model[c*3+r])
    float nrm[9];         // transpose(inverse(mat3(model))) columns
    uint32_t material_index;
    uint32_t light_index; // 0 after a bake (gpu_scene.odin:333-336 writes it into a copy)
    uint32_t vertex_base; // into the concatenated vertex array
    uint32_t index_base;  // into the concatenated index array
    uint32_t tri_offset;  // first flat triangle id
    uint32_t n_tris;
    uint32_t mesh_index;
    // First-hit filter of the hit shader's first launch (rt_frame.cu, wf_shade_kernel DENSE, "dead diffuse lobe"): 1 when a hit on this object that draws the Lambert
    // lobe ends its path without a colour term whatever the geometry — metallic == 1 makes the lobe's value exactly zero, and the
    // object's normals and matrix are checked (host side, rt_api.cu) to keep every intermediate finite.  spec_prob is
    // specularProbability (simple.rchit:69-75) of the object's material, computed on the device by the shader's own function.
    uint32_t dead_diffuse;
    float spec_prob;
    uint32_t pad[2];
};
static_assert(sizeof(DevObject) == 128, "DevObject is 128 bytes");
static_assert(offsetof(DevObject, nrm) == 48 && offsetof(DevObject, material_index) == 84 && offsetof(DevObject, light_index) == 88 &&
                  offsetof(DevObject, vertex_base) == 92 && offsetof(DevObject, index_base) == 96,
              "shade_hit reads DevObject as six 16-byte words");
static_assert(offsetof(DevObject, dead_diffuse) == 112 && offsetof(DevObject, spec_prob) == 116, "the filter reads (dead_diffuse, spec_prob) as one 64-bit word");

struct __align__(16) DevLight {  // 64 B
    float transform[12];  // as DevObject.model
    uint32_t object_index;
    uint32_t num_triangles;
    uint32_t instance_mask;  // 1 << (object & 7), gpu_scene.odin:129-132 (no ray uses it)
    float emission_power;    // materials[objects[object_index].material].emission_power
};
static_assert(sizeof(DevLight) == 64, "DevLight is 64 bytes");

struct Hit {
    float t, u, v;
    uint32_t id;    // flat triangle id, 0xFFFFFFFF = miss
    uint32_t inst, prim;
};

struct DevCounters {  // mirrors rt_counters
    unsigned long long extend_rays, shadow_rays, nodes, tris, hits, nee, frames, shadow_nodes, shadow_tris, launches;
};

// Everything a traversal / shading kernel needs, passed by value.
struct SceneView {
    const uint4* nodes;        // Node8 as 8 x uint4
    const float4* tris;        // BvhTri as 3 x float4
    const DevObject* objects;
    const DevLight* lights;
    const rt_material* materials;
    const rt_vertex* vertices;
    const uint32_t* indices;
    uint32_t n_lights;
    uint32_t n_objects;
    uint32_t n_tris;
    uint32_t n_nodes;
    uint32_t any_dead_diffuse;  // some object has DevObject::dead_diffuse set (and the scene no light): the first-hit filter is worth running
};

struct FrameParams {
    float inv_view[16];
    float inv_proj[16];
    float clear[3];
    uint32_t frame;                                 // first frame of the launch
    uint32_t n_lanes;                               // frames traced together (wavefront pipeline): frame + l * frame_stride
    uint32_t frame_stride;                          // 1, or the world size when ranks take spp slices
    uint32_t sample;                                // running-mean index of lane 0 (simple.rgen:127-136); == frame unless sliced
    uint32_t width, height;
    uint32_t strip_rows, strip_index, strip_count;  // rows owned: (y / strip_rows) % strip_count == strip_index ...
    const uint32_t* band_strip;                     // ... unless this table (device; local band -> global strip) says otherwise (RT_BALANCE_COST)
    uint32_t local_rows;                            // rows in this context's compact image
    int32_t max_depth;
    int32_t rr_start;
    float clamp;
    uint32_t flags;
};

// local row <-> global row of the strip partition
RT_HD uint32_t strip_global_row(uint32_t local_row, uint32_t strip_rows, uint32_t strip_index, uint32_t strip_count) {
    uint32_t band = local_row / strip_rows, r = local_row % strip_rows;
    return (band * strip_count + strip_index) * strip_rows + r;
}
#if defined(__CUDACC__)
__device__ __forceinline__ uint32_t strip_global_row(const FrameParams& fp, uint32_t local_row) {
    if (fp.strip_count == 1u) return local_row;
    const uint32_t band = local_row / fp.strip_rows, r = local_row - band * fp.strip_rows;
    const uint32_t strip = fp.band_strip ? __ldg(fp.band_strip + band) : band * fp.strip_count + fp.strip_index;
    return strip * fp.strip_rows + r;
}
#endif

#if defined(__CUDACC__)
// 256-bit read-only load (sm_100 LDG.256); p must be 32-byte aligned.  A load whose threads sit in different lines costs
// the SM's L1 pipe one cycle per thread whatever its width (profiles/r1q_gather_microbench.md), so wide loads are how
// a divergent gather gets cheaper.
__device__ __forceinline__ void ldg256(const uint4* p, uint4& a, uint4& b) {
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
                 : "l"(p));
}
__device__ __forceinline__ void ldg256f(const float4* p, float4& a, float4& b) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
}
// One 48-byte triangle as a 256-bit and a 128-bit load (which of the two comes first depends on the parity of the
// index: only even triangles start on a 32-byte boundary) instead of three 128-bit ones: two L1 requests per thread
// instead of three when the lanes of a warp fetch different triangles (tools/mb/gather2.cu: 62.7 -> 33.2 cycles per
// warp, L1-resident; 90.1 -> 63.7 from L2).  `tris` must be 32-byte aligned (every device allocation is).
#ifndef RT_TRI_LDG256
#define RT_TRI_LDG256 1
#endif
#ifndef RT_TRI_NOALLOC
#define RT_TRI_NOALLOC 0 /* measured: -1 % on bounces >= 1: triangles do get L1 hits */
#endif
__device__ __forceinline__ void ldg256f_na(const float4* p, float4& a, float4& b) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
}
__device__ __forceinline__ float4 ldg128f_na(const float4* p) {
    float4 a;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "l"(p));
    return a;
}
__device__ __forceinline__ void ldg_tri(const float4* tris, uint32_t ti, float4& q0, float4& q1, float4& q2) {
    const float4* tp = tris + (size_t)ti * 3u;
#if RT_TRI_LDG256 && RT_TRI_NOALLOC
    if (ti & 1u) { q0 = ldg128f_na(tp); ldg256f_na(tp + 1, q1, q2); }
    else { ldg256f_na(tp, q0, q1); q2 = ldg128f_na(tp + 2); }
#elif RT_TRI_LDG256
    if (ti & 1u) { q0 = __ldg(tp); ldg256f(tp + 1, q1, q2); }
    else { ldg256f(tp, q0, q1); q2 = __ldg(tp + 2); }
#else
    q0 = __ldg(tp); q1 = __ldg(tp + 1); q2 = __ldg(tp + 2);
#endif
}
#endif

}  // namespace rt
