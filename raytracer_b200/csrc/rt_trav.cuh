// rt_trav.cuh — BVH8 traversal: the CUDA replacement of traceRayEXT (shaders/simple.rgen:92-104,
// closest opaque hit) and of the inline rayQueryEXT of isVisibleRQ (shaders/simple.rchit:350-385,
// "any hit whose instance is not the light").  B200 has no RT cores; this is the hot loop.
//
// Contract: hits are accepted for tmin < t < tmax (strictly), ties on t go to the
// lowest flat triangle id, so the result is independent of traversal order.  The ray/triangle test is
// the watertight test of Woop, Benthin & Wald (2013) with the evaluation order fixed below.
#pragma once
#include "rt_types.cuh"

namespace rt {

struct TravRay {
    float ox, oy, oz;
    float idx, idy, idz;   // 1/d (clamped for |d| < 1e-20)
    float Sx, Sy, Sz;      // watertight shear
    uint32_t kx, ky, kz;   // watertight axis permutation
    uint32_t rbits;        // sign bits of d: traversal octant
};

__device__ __forceinline__ float comp3(float x, float y, float z, uint32_t k) { return k == 0 ? x : (k == 1 ? y : z); }

__device__ __forceinline__ TravRay trav_setup(vec3 o, vec3 d) {
    TravRay r;
    r.ox = o.x; r.oy = o.y; r.oz = o.z;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    uint32_t kz = (ax > ay) ? ((ax > az) ? 0u : 2u) : ((ay > az) ? 1u : 2u);
    uint32_t kx = (kz == 2u) ? 0u : kz + 1u;
    uint32_t ky = (kx == 2u) ? 0u : kx + 1u;
    float dz = comp3(d.x, d.y, d.z, kz);
    if (dz < 0.0f) { uint32_t t = kx; kx = ky; ky = t; }
    r.kx = kx; r.ky = ky; r.kz = kz;
    r.Sx = comp3(d.x, d.y, d.z, kx) / dz;
    r.Sy = comp3(d.x, d.y, d.z, ky) / dz;
    r.Sz = 1.0f / dz;
    r.idx = (fabsf(d.x) >= 1e-20f) ? 1.0f / d.x : copysignf(1e20f, d.x);
    r.idy = (fabsf(d.y) >= 1e-20f) ? 1.0f / d.y : copysignf(1e20f, d.y);
    r.idz = (fabsf(d.z) >= 1e-20f) ? 1.0f / d.z : copysignf(1e20f, d.z);
    r.rbits = (__float_as_uint(d.x) >> 31) | ((__float_as_uint(d.y) >> 31) << 1) | ((__float_as_uint(d.z) >> 31) << 2);
    return r;
}

// Watertight intersection; returns false when the ray misses the triangle's plane projection.
__device__ __forceinline__ bool tri_isect(const TravRay& r, float4 q0, float4 q1, float4 q2, float* t, float* u,
                                          float* v) {
    float ax = q0.x - r.ox, ay = q0.y - r.oy, az = q0.z - r.oz;
    float bx = q1.x - r.ox, by = q1.y - r.oy, bz = q1.z - r.oz;
    float cx = q2.x - r.ox, cy = q2.y - r.oy, cz = q2.z - r.oz;
    float Akz = comp3(ax, ay, az, r.kz), Bkz = comp3(bx, by, bz, r.kz), Ckz = comp3(cx, cy, cz, r.kz);
    float Ax = fmaf(-r.Sx, Akz, comp3(ax, ay, az, r.kx)), Ay = fmaf(-r.Sy, Akz, comp3(ax, ay, az, r.ky));
    float Bx = fmaf(-r.Sx, Bkz, comp3(bx, by, bz, r.kx)), By = fmaf(-r.Sy, Bkz, comp3(bx, by, bz, r.ky));
    float Cx = fmaf(-r.Sx, Ckz, comp3(cx, cy, cz, r.kx)), Cy = fmaf(-r.Sy, Ckz, comp3(cx, cy, cz, r.ky));
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {  // edge case: redo the edge functions in binary64
        U = (float)__dsub_rn(__dmul_rn((double)Cx, (double)By), __dmul_rn((double)Cy, (double)Bx));
        V = (float)__dsub_rn(__dmul_rn((double)Ax, (double)Cy), __dmul_rn((double)Ay, (double)Cx));
        W = (float)__dsub_rn(__dmul_rn((double)Bx, (double)Ay), __dmul_rn((double)By, (double)Ax));
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    float det = (U + V) + W;
    if (det == 0.0f) return false;
    float Az = r.Sz * Akz, Bz = r.Sz * Bkz, Cz = r.Sz * Ckz;
    float T = (U * Az + V * Bz) + W * Cz;
    float inv = 1.0f / det;
    *t = T * inv;
    *u = V * inv;
    *v = W * inv;
    return true;
}

__device__ __forceinline__ float byte_f(uint32_t w, int k) { return (float)((w >> (8 * k)) & 0xFFu); }

// Slab test of the 8 quantised child boxes.  Result: bits [31:24] = internal children that are hit, at
// position 24 + (slot ^ octant) (so the highest bit is the nearest octant); bits [23:0] = triangles.
__device__ __forceinline__ uint32_t node_hits(const uint4 w0, const uint4 w1, const uint4 w2, const uint4 w3,
                                              const uint4 w4, const TravRay& r, float tmin, float tmax) {
    const uint32_t ew = w0.w;  // e.x e.y e.z imask
    const float scx = __uint_as_float((ew & 0xFFu) << 23) * r.idx;
    const float scy = __uint_as_float(((ew >> 8) & 0xFFu) << 23) * r.idy;
    const float scz = __uint_as_float(((ew >> 16) & 0xFFu) << 23) * r.idz;
    const uint32_t imask = ew >> 24;
    const float orx = (__uint_as_float(w0.x) - r.ox) * r.idx;
    const float ory = (__uint_as_float(w0.y) - r.oy) * r.idy;
    const float orz = (__uint_as_float(w0.z) - r.oz) * r.idz;
    // near/far byte planes per axis, chosen by the ray's sign
    const bool nx = r.rbits & 1u, ny = r.rbits & 2u, nz = r.rbits & 4u;
    // qlo.x = w2.xy, qlo.y = w2.zw, qlo.z = w3.xy, qhi.x = w3.zw, qhi.y = w4.xy, qhi.z = w4.zw
    const uint32_t nX0 = nx ? w3.z : w2.x, nX1 = nx ? w3.w : w2.y, fX0 = nx ? w2.x : w3.z, fX1 = nx ? w2.y : w3.w;
    const uint32_t nY0 = ny ? w4.x : w2.z, nY1 = ny ? w4.y : w2.w, fY0 = ny ? w2.z : w4.x, fY1 = ny ? w2.w : w4.y;
    const uint32_t nZ0 = nz ? w4.z : w3.x, nZ1 = nz ? w4.w : w3.y, fZ0 = nz ? w3.x : w4.z, fZ1 = nz ? w3.y : w4.w;
    uint32_t mask = 0;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const uint32_t metaw = half ? w1.w : w1.z;
        const uint32_t nX = half ? nX1 : nX0, fX = half ? fX1 : fX0;
        const uint32_t nY = half ? nY1 : nY0, fY = half ? fY1 : fY0;
        const uint32_t nZ = half ? nZ1 : nZ0, fZ = half ? fZ1 : fZ0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t meta = (metaw >> (8 * k)) & 0xFFu;
            const float tnx = fmaf(byte_f(nX, k), scx, orx), tfx = fmaf(byte_f(fX, k), scx, orx);
            const float tny = fmaf(byte_f(nY, k), scy, ory), tfy = fmaf(byte_f(fY, k), scy, ory);
            const float tnz = fmaf(byte_f(nZ, k), scz, orz), tfz = fmaf(byte_f(fZ, k), scz, orz);
            const float tnear = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
            const float tfar = fminf(fminf(tfx, tfy), fminf(tfz, tmax));
            if (meta != 0u && tnear <= tfar) {
                const uint32_t sl = (uint32_t)(half * 4 + k);
                const uint32_t bits = meta >> 5;
                const uint32_t idx = ((imask >> sl) & 1u) ? (24u + (sl ^ r.rbits)) : (meta & 31u);
                mask |= bits << idx;
            }
        }
    }
    return mask;
}

// One ray per lane, while-while with a (base, mask) group stack.  ANY: stop at the first accepted hit
// whose instance differs from ignore_inst.  n_nodes / n_tris are incremented when COUNT.
template <bool ANY, bool COUNT>
__device__ __forceinline__ Hit traverse(const SceneView& sc, vec3 o, vec3 d, float tmin, float tmax,
                                        uint32_t ignore_inst, uint32_t& n_nodes, uint32_t& n_tris) {
    const TravRay r = trav_setup(o, d);
    Hit best;
    best.t = tmax; best.u = 0.0f; best.v = 0.0f; best.id = 0xFFFFFFFFu; best.inst = 0u; best.prim = 0u;
    uint2 stack[RT_STACK_SIZE];
    int sp = 0;
    uint2 ng = make_uint2(0u, 0x80000000u);  // root group: bit 31, relative index 0
    uint2 tg = make_uint2(0u, 0u);
    for (;;) {
        if (ng.y & 0xFF000000u) {
            const int bit = 31 - __clz(ng.y);
            ng.y &= ~(1u << bit);
            const uint32_t slot = ((uint32_t)bit - 24u) ^ r.rbits;
            const uint32_t rel = __popc(ng.y & 0xFFu & ((1u << slot) - 1u));
            const uint32_t node = ng.x + rel;
            if (ng.y & 0xFF000000u) {
                stack[sp++] = ng;
            }
            const uint4* np = sc.nodes + (size_t)node * 5u;
            const uint4 w0 = __ldg(np), w1 = __ldg(np + 1), w2 = __ldg(np + 2), w3 = __ldg(np + 3), w4 = __ldg(np + 4);
            if (COUNT) n_nodes++;
            const uint32_t hm = node_hits(w0, w1, w2, w3, w4, r, tmin, best.t);
            ng = make_uint2(w1.x, (hm & 0xFF000000u) | (w0.w >> 24));
            tg = make_uint2(w1.y, hm & 0x00FFFFFFu);
        } else {
            tg = make_uint2(0u, 0u);
            ng = make_uint2(0u, 0u);
        }
        while (tg.y) {
            const int bit = __ffs(tg.y) - 1;
            tg.y &= tg.y - 1u;
            const float4* tp = sc.tris + (size_t)(tg.x + (uint32_t)bit) * 3u;
            const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
            if (COUNT) n_tris++;
            const uint32_t inst = __float_as_uint(q1.w);
            if (ANY && inst == ignore_inst) continue;
            float tt, uu, vv;
            if (!tri_isect(r, q0, q1, q2, &tt, &uu, &vv)) continue;
            if (!(tt > tmin)) continue;
            const uint32_t id = __float_as_uint(q0.w);
            if (ANY) {
                if (tt < tmax) {
                    best.t = tt; best.u = uu; best.v = vv; best.id = id; best.inst = inst; best.prim = __float_as_uint(q2.w);
                    return best;
                }
            } else {
                const bool take = (best.id == 0xFFFFFFFFu) ? (tt < best.t) : ((tt < best.t) || (tt == best.t && id < best.id));
                if (take) {
                    best.t = tt; best.u = uu; best.v = vv; best.id = id; best.inst = inst; best.prim = __float_as_uint(q2.w);
                }
            }
        }
        if (!(ng.y & 0xFF000000u)) {
            if (sp == 0) break;
            ng = stack[--sp];
        }
    }
    return best;
}

}  // namespace rt
