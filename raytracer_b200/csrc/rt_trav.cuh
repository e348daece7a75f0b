// rt_trav.cuh — BVH8 traversal: the CUDA replacement of traceRayEXT (shaders/simple.rgen:92-104,
// closest opaque hit) and of the inline rayQueryEXT of isVisibleRQ (shaders/simple.rchit:350-385,
// "any hit whose instance is not the light").  B200 has no RT cores; this is the hot loop.
//
// Contract: hits are accepted for tmin < t < tmax (strictly), ties on t go to the
// lowest flat triangle id, so the result is independent of traversal order.  The ray/triangle test is
// the watertight test of Woop, Benthin & Wald (2013) with the evaluation order fixed below.
#pragma once
#include <cuda_fp16.h>

#include "rt_types.cuh"

#ifndef RT_USE_FFMA2
#define RT_USE_FFMA2 1
#endif

namespace rt {

// The per-ray part of the watertight test on its own, so that the packet kernel can keep several rays per lane
// without the slab-test members of TravRay.
struct Shear { float2 px, py, pz, cxy; };

struct TravRay {
    float ox, oy, oz;
    float idx, idy, idz;   // 1/d (clamped for |d| < 1e-20)
    // Watertight shear as an affine map: row pairs (mx_i, my_i) for i = x, y, z with
    // mx = e_kx - Sx e_kz, my = e_ky - Sy e_kz, and cxy = (mx . o, my . o).
    // The entries that are exactly 0 or 1 make the fmaf chains below reproduce "pick a component" without
    // any per-triangle selects (the selects cost 54 ALU-pipe instructions per triangle before).
    Shear sh;
    uint32_t rbits;        // bits 0..2: sign bits of d (traversal octant); bits 3..4: dominant axis kz
};

// Woop et al. permute the axes so that kz is the dominant direction and kx, ky follow cyclically, and swap
// kx, ky when d[kz] < 0 to keep the winding for back-face culling.  Nothing is culled on this path and the
// swap only negates U, V, W, det and T together, so it is left out here and in the oracle.
__device__ __forceinline__ TravRay trav_setup(vec3 o, vec3 d) {
    TravRay r;
    r.ox = o.x; r.oy = o.y; r.oz = o.z;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    const uint32_t kz = (ax > ay) ? ((ax > az) ? 0u : 2u) : ((ay > az) ? 1u : 2u);
    // rotate (x, y, z) -> (kx, ky, kz) components
    const float dkx = kz == 0u ? d.y : (kz == 1u ? d.z : d.x);
    const float dky = kz == 0u ? d.z : (kz == 1u ? d.x : d.y);
    const float dz = kz == 0u ? d.x : (kz == 1u ? d.y : d.z);
    const float nSx = -(dkx / dz), nSy = -(dky / dz);
    // axis i is kx when i == (kz + 1) % 3, ky when i == (kz + 2) % 3
    r.sh.px = kz == 0u ? make_float2(nSx, nSy) : (kz == 1u ? make_float2(0.0f, 1.0f) : make_float2(1.0f, 0.0f));
    r.sh.py = kz == 0u ? make_float2(1.0f, 0.0f) : (kz == 1u ? make_float2(nSx, nSy) : make_float2(0.0f, 1.0f));
    r.sh.pz = kz == 0u ? make_float2(0.0f, 1.0f) : (kz == 1u ? make_float2(1.0f, 0.0f) : make_float2(nSx, nSy));
    r.sh.cxy.x = fmaf(o.z, r.sh.pz.x, fmaf(o.y, r.sh.py.x, o.x * r.sh.px.x));
    r.sh.cxy.y = fmaf(o.z, r.sh.pz.y, fmaf(o.y, r.sh.py.y, o.x * r.sh.px.y));
    r.idx = (fabsf(d.x) >= 1e-20f) ? 1.0f / d.x : copysignf(1e20f, d.x);
    r.idy = (fabsf(d.y) >= 1e-20f) ? 1.0f / d.y : copysignf(1e20f, d.y);
    r.idz = (fabsf(d.z) >= 1e-20f) ? 1.0f / d.z : copysignf(1e20f, d.z);
    r.rbits = (__float_as_uint(d.x) >> 31) | ((__float_as_uint(d.y) >> 31) << 1) | ((__float_as_uint(d.z) >> 31) << 2) | (kz << 3);
    return r;
}

// sheared (x, y) of a vertex: three packed FFMA2 (sm_100 fma.rn.f32x2), each half rounded exactly like fmaf
__device__ __forceinline__ float2 shear_xy(const Shear& r, float4 q) {
    float2 a = __ffma2_rn(make_float2(q.x, q.x), r.px, make_float2(-r.cxy.x, -r.cxy.y));
    a = __ffma2_rn(make_float2(q.y, q.y), r.py, a);
    return __ffma2_rn(make_float2(q.z, q.z), r.pz, a);
}

// Edge functions of the sheared triangle; false when the ray misses.  det = (U + V) + W is non-zero on success.
__device__ __forceinline__ bool tri_edges(const Shear& r, float4 q0, float4 q1, float4 q2, float* pU, float* pV, float* pW,
                                          float* pdet) {
    const float2 A = shear_xy(r, q0), B = shear_xy(r, q1), C = shear_xy(r, q2);
    float U = C.x * B.y - C.y * B.x;
    float V = A.x * C.y - A.y * C.x;
    float W = B.x * A.y - B.y * A.x;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {  // edge case: redo the edge functions in binary64
        U = (float)__dsub_rn(__dmul_rn((double)C.x, (double)B.y), __dmul_rn((double)C.y, (double)B.x));
        V = (float)__dsub_rn(__dmul_rn((double)A.x, (double)C.y), __dmul_rn((double)A.y, (double)C.x));
        W = (float)__dsub_rn(__dmul_rn((double)B.x, (double)A.y), __dmul_rn((double)B.y, (double)A.x));
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    const float det = (U + V) + W;
    if (det == 0.0f) return false;
    *pU = U; *pV = V; *pW = W; *pdet = det;
    return true;
}

// distance: the hit point's coordinate along the dominant axis kz (barycentric mix of the three vertices),
// then t = (p_kz - o_kz) / d_kz.  kzbits: bit 0 = (kz == 1), bit 1 = (kz == 2).
__device__ __forceinline__ void tri_finish(float U, float V, float W, float det, float4 q0, float4 q1, float4 q2, uint32_t kzbits,
                                           float okz, float ikz, float* t, float* u, float* v) {
    const bool k0 = kzbits == 0u, k1 = (kzbits & 1u) != 0u;
    const float Az = k0 ? q0.x : (k1 ? q0.y : q0.z), Bz = k0 ? q1.x : (k1 ? q1.y : q1.z), Cz = k0 ? q2.x : (k1 ? q2.y : q2.z);
    const float T = fmaf(W, Cz, fmaf(V, Bz, U * Az));
    const float inv = 1.0f / det;
    *t = (T * inv - okz) * ikz;
    *u = V * inv;
    *v = W * inv;
}

// Watertight intersection; returns false when the ray misses the triangle's plane projection.
__device__ __forceinline__ bool tri_isect(const TravRay& r, float4 q0, float4 q1, float4 q2, float* t, float* u,
                                          float* v) {
    float U, V, W, det;
    if (!tri_edges(r.sh, q0, q1, q2, &U, &V, &W, &det)) return false;
    const uint32_t kzbits = (r.rbits >> 3) & 3u;
    const bool k0 = kzbits == 0u, k1 = (kzbits & 1u) != 0u;
    tri_finish(U, V, W, det, q0, q1, q2, kzbits, k0 ? r.ox : (k1 ? r.oy : r.oz), k0 ? r.idx : (k1 ? r.idy : r.idz), t, u, v);
    return true;
}

// binary16 -> binary32 of the low / high half of a word: HADD2.F32 with an operand selector (FMA pipe)
__device__ __forceinline__ float half_lo(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w & 0xFFFFu))); }
__device__ __forceinline__ float half_hi(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w >> 16))); }

// Slab test of the 8 quantised child boxes.  nx/fx .. nz/fz: the near and far plane words of each axis (8 grid
// coordinates each), already chosen by the ray's sign.  Result: bits [31:24] = internal children that are
// hit, at position 24 + (slot ^ octant) (so the highest bit is the nearest octant); bits [23:0] = triangles.
__device__ __forceinline__ uint32_t node_hits(const uint4 w0, const uint4 w1, const uint4 nx, const uint4 fx, const uint4 ny,
                                              const uint4 fy, const uint4 nz, const uint4 fz, const TravRay& r, float tmin,
                                              float tmax) {
    const uint32_t ew = w0.w;  // e.x e.y e.z imask
    const float scx = __uint_as_float((ew & 0xFFu) << 23) * r.idx;
    const float scy = __uint_as_float(((ew >> 8) & 0xFFu) << 23) * r.idy;
    const float scz = __uint_as_float(((ew >> 16) & 0xFFu) << 23) * r.idz;
    const float orx = (__uint_as_float(w0.x) - r.ox) * r.idx;
    const float ory = (__uint_as_float(w0.y) - r.oy) * r.idy;
    const float orz = (__uint_as_float(w0.z) - r.oz) * r.idz;
    const float2 scx2 = make_float2(scx, scx), scy2 = make_float2(scy, scy), scz2 = make_float2(scz, scz);
    const float2 orx2 = make_float2(orx, orx), ory2 = make_float2(ory, ory), orz2 = make_float2(orz, orz);
    // meta bytes, four children at a time: internal children (24..31 in the low five bits) get their slot
    // XORed with the ray octant; empty children (meta 0) carry no bits and drop out without a test
    const uint32_t oct4 = (r.rbits & 7u) * 0x01010101u;
    uint32_t mask = 0;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const uint32_t meta4 = half ? w1.w : w1.z;
        const uint32_t inner4 = ((meta4 & (meta4 << 1)) & 0x10101010u) >> 4;
        const uint32_t idx4 = (meta4 ^ (oct4 & (inner4 * 0xFFu))) & 0x1F1F1F1Fu;
        const uint32_t bits4 = (meta4 >> 5) & 0x07070707u;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            // children 2j, 2j+1 of an axis share word j of its plane vector
            const int j = half * 2 + (k >> 1);
            const uint32_t wnx = j == 0 ? nx.x : (j == 1 ? nx.y : (j == 2 ? nx.z : nx.w));
            const uint32_t wfx = j == 0 ? fx.x : (j == 1 ? fx.y : (j == 2 ? fx.z : fx.w));
            const uint32_t wny = j == 0 ? ny.x : (j == 1 ? ny.y : (j == 2 ? ny.z : ny.w));
            const uint32_t wfy = j == 0 ? fy.x : (j == 1 ? fy.y : (j == 2 ? fy.z : fy.w));
            const uint32_t wnz = j == 0 ? nz.x : (j == 1 ? nz.y : (j == 2 ? nz.z : nz.w));
            const uint32_t wfz = j == 0 ? fz.x : (j == 1 ? fz.y : (j == 2 ? fz.z : fz.w));
            const bool hi = (k & 1) != 0;
            // near and far plane of an axis share scale and offset: one packed FFMA2 (sm_100 f32x2) per axis,
            // each half rounded exactly like fmaf
            const float2 tx = __ffma2_rn(make_float2(hi ? half_hi(wnx) : half_lo(wnx), hi ? half_hi(wfx) : half_lo(wfx)), scx2, orx2);
            const float2 ty = __ffma2_rn(make_float2(hi ? half_hi(wny) : half_lo(wny), hi ? half_hi(wfy) : half_lo(wfy)), scy2, ory2);
            const float2 tz = __ffma2_rn(make_float2(hi ? half_hi(wnz) : half_lo(wnz), hi ? half_hi(wfz) : half_lo(wfz)), scz2, orz2);
            const float tnear = fmaxf(fmaxf(tx.x, ty.x), fmaxf(tz.x, tmin));
            const float tfar = fminf(fminf(tx.y, ty.y), fminf(tz.y, tmax));
            const uint32_t hit = ((bits4 >> (8 * k)) & 0xFFu) << ((idx4 >> (8 * k)) & 0xFFu);
            mask |= (tnear <= tfar) ? hit : 0u;
        }
    }
    return mask;
}

#ifndef RT_NODE_LDG256
#define RT_NODE_LDG256 1
#endif
__device__ __forceinline__ uint4 sel4(bool c, uint4 a, uint4 b) {
    return make_uint4(c ? a.x : b.x, c ? a.y : b.y, c ? a.z : b.z, c ? a.w : b.w);
}

// One node = one 128-byte line: header words w0, w1 and, per axis, the near and far plane vectors for this ray's
// direction signs (near plane of axis a: word 2 + 2a + sign(d_a), far plane: the other one).  The four chunks are
// stored rotated by the node index (rt_types.cuh: node_chunk).
__device__ __forceinline__ void load_node(const uint4* nodes, uint32_t node, uint32_t sx, uint32_t sy, uint32_t sz, uint4& w0, uint4& w1,
                                          uint4& nx, uint4& fx, uint4& ny, uint4& fy, uint4& nz, uint4& fz) {
    const uint4* np = nodes + (size_t)node * RT_NODE_WORDS;
#if RT_NODE_LDG256
    // four 256-bit loads per node (sm_100 LDG.256) instead of eight 128-bit ones: fewer L1 requests when the lanes
    // of a warp sit in different nodes; the near / far planes are then picked by register selects
    uint4 lx, hx, ly, hy, lz, hz;
    ldg256(np + node_chunk(node, 0u), w0, w1);
    ldg256(np + node_chunk(node, 1u), lx, hx);
    ldg256(np + node_chunk(node, 2u), ly, hy);
    ldg256(np + node_chunk(node, 3u), lz, hz);
    nx = sel4(sx != 0u, hx, lx); fx = sel4(sx != 0u, lx, hx);
    ny = sel4(sy != 0u, hy, ly); fy = sel4(sy != 0u, ly, hy);
    nz = sel4(sz != 0u, hz, lz); fz = sel4(sz != 0u, lz, hz);
#else
    const uint4 *c0 = np + node_chunk(node, 0u), *cx = np + node_chunk(node, 1u), *cy = np + node_chunk(node, 2u),
                *cz = np + node_chunk(node, 3u);
    w0 = __ldg(c0); w1 = __ldg(c0 + 1);
    nx = __ldg(cx + sx); fx = __ldg(cx + 1 - sx);
    ny = __ldg(cy + sy); fy = __ldg(cy + 1 - sy);
    nz = __ldg(cz + sz); fz = __ldg(cz + 1 - sz);
#endif
}

// One ray per lane with a (base, mask) group stack.  The traversal is a resumable state machine so that the
// persistent kernels can swap a finished ray for a new one without waiting for the rest of the warp; one
// step = one node (its eight child boxes), all triangles that node references directly, and a stack pop.
// The visiting order of a ray depends on the ray alone, so node/triangle counters are reproducible.
// ANY: stop at the first accepted hit whose instance differs from ignore_inst.
struct TravState {
    TravRay r;
    Hit best;
    uint2 ng;       // current node group: child_base, hit bits [31:24] | imask [7:0]
    float tmin;
    uint32_t ignore_inst;
    int sp;
};
// the group stack lives beside the state (a dynamically indexed member would drag the whole struct into local memory)
struct TravStack { uint2 e[RT_STACK_SIZE]; };
// The cooperative kernels keep the newest stack entry in registers (`top`, valid while top.y != 0: a pushed group
// always has hit bits left): an entry that is popped right after it was pushed never touches local memory.  ncu on
// the Sponza proxy: the stack was 20 % of the kernel's L1 sectors and 13 % of its L1 misses (node traffic evicts it).
#ifndef RT_STACK_TOP
#define RT_STACK_TOP 0 /* measured: 2073.7 -> 2049.2 Mrays/s on bounces >= 1 (the extra selects cost more issue slots than the L1 relief returns) */
#endif
__device__ __forceinline__ void stk_push(TravStack& stk, int& sp, uint2& top, uint2 x) {
#if RT_STACK_TOP
    if (top.y) stk.e[sp++] = top;
    top = x;
#else
    (void)top;
    stk.e[sp++] = x;
#endif
}
// false: the stack is empty
__device__ __forceinline__ bool stk_pop(TravStack& stk, int& sp, uint2& top, uint2& out) {
#if RT_STACK_TOP
    if (top.y) { out = top; top.y = 0u; return true; }
#else
    (void)top;
#endif
    if (sp == 0) return false;
    out = stk.e[--sp];
    return true;
}

__device__ __forceinline__ void trav_init(TravState& st, vec3 o, vec3 d, float tmin, float tmax, uint32_t ignore_inst) {
    st.r = trav_setup(o, d);
    st.best.t = tmax; st.best.u = 0.0f; st.best.v = 0.0f; st.best.id = 0xFFFFFFFFu; st.best.inst = 0u; st.best.prim = 0u;
    st.ng = make_uint2(0u, 0x80000000u);  // root group: bit 31, relative index 0
    st.tmin = tmin;
    st.ignore_inst = ignore_inst;
    st.sp = 0;
}

// returns true when the traversal is complete (st.best holds the answer)
template <bool ANY, bool COUNT>
__device__ __forceinline__ bool trav_step(const SceneView& sc, TravState& st, TravStack& stk, uint32_t& n_nodes,
                                          uint32_t& n_tris) {
    const TravRay& r = st.r;
    uint2 ng = st.ng;
    const int bit = 31 - __clz(ng.y);
    ng.y &= ~(1u << bit);
    const uint32_t slot = ((uint32_t)bit - 24u) ^ (r.rbits & 7u);
    const uint32_t rel = __popc(ng.y & 0xFFu & ((1u << slot) - 1u));
    const uint32_t node = ng.x + rel;
    if (ng.y & 0xFF000000u) stk.e[st.sp++] = ng;
    // near plane of axis a: word 2 + 2a + sign(d_a), far plane: the other one
    const uint32_t sx = r.rbits & 1u, sy = (r.rbits >> 1) & 1u, sz = (r.rbits >> 2) & 1u;
    uint4 w0, w1, nx, fx, ny, fy, nz, fz;
    load_node(sc.nodes, node, sx, sy, sz, w0, w1, nx, fx, ny, fy, nz, fz);
    if (COUNT) n_nodes++;
    const uint32_t hm = node_hits(w0, w1, nx, fx, ny, fy, nz, fz, r, st.tmin, st.best.t);
    ng = make_uint2(w1.x, (hm & 0xFF000000u) | (w0.w >> 24));
    uint32_t tmask = hm & 0x00FFFFFFu;
    while (tmask) {
        const int tb = __ffs(tmask) - 1;
        tmask &= tmask - 1u;
        float4 q0, q1, q2;
        ldg_tri(sc.tris, w1.y + (uint32_t)tb, q0, q1, q2);
        if (COUNT) n_tris++;
        const uint32_t inst = __float_as_uint(q1.w);
        if (ANY && inst == st.ignore_inst) continue;
        float tt, uu, vv;
        if (!tri_isect(r, q0, q1, q2, &tt, &uu, &vv)) continue;
        if (!(tt > st.tmin)) continue;
        const uint32_t id = __float_as_uint(q0.w);
        if (ANY) {
            if (tt < st.best.t) {  // best.t stays tmax until the first accepted hit
                st.best.t = tt; st.best.u = uu; st.best.v = vv; st.best.id = id; st.best.inst = inst; st.best.prim = __float_as_uint(q2.w);
                return true;
            }
        } else {
            const bool take = (st.best.id == 0xFFFFFFFFu) ? (tt < st.best.t) : ((tt < st.best.t) || (tt == st.best.t && id < st.best.id));
            if (take) {
                st.best.t = tt; st.best.u = uu; st.best.v = vv; st.best.id = id; st.best.inst = inst; st.best.prim = __float_as_uint(q2.w);
            }
        }
    }
    if (!(ng.y & 0xFF000000u)) {
        if (st.sp == 0) return true;
        ng = stk.e[--st.sp];
    }
    st.ng = ng;
    return false;
}

// ---- warp-cooperative triangle tests (closest hit) --------------------------------------------------
// With incoherent rays only a few lanes of a warp reach triangles in the same node step, and each runs its own
// loop over them: ncu showed half of the kernel's warp instructions issued in that loop with 2 of 32 lanes
// active, every iteration exposing one more load latency.  Here the (ray, triangle) pairs of all lanes are
// pooled in shared memory after each node step and dealt out evenly: lane w tests pair w for whichever lane
// owns it.  The owner's closest hit lives in shared memory as one 64-bit key (t bits << 32 | flat id): an
// unsigned atomicMin on it IS the (t, id)-lexicographic rule of the per-lane loop (t > 0, so the bit patterns
// order like the floats), hence the same hit record bit for bit.
struct CoopWarp {
    unsigned long long key[32];  // t bits << 32 | flat triangle id of the closest accepted hit; (tmax bits << 32) = none yet
    float4 ra[32];               // shear px.x px.y py.x py.y
    float4 rb[32];               // shear pz.x pz.y cxy.x cxy.y
    float4 rc[32];               // o_kz, 1/d_kz, kz bits, tmax (any-hit)
    float2 uv[32];               // barycentrics of the hit in `key`
    uint32_t tri[32];            // its index in the BVH triangle array, 0xFFFFFFFF = miss; any-hit: 0 = occluded
    uint32_t ignore[32];         // any-hit: triangles of this instance do not occlude (simple.rchit:378-382)
    uint4 ent[32];               // one per lane with triangles this step: triangle base, bits, first pool slot, lane
};

// slim per-lane state of the cooperative kernel: everything the node test needs, nothing of the triangle test
struct NodeRay {
    float ox, oy, oz, idx, idy, idz;
    uint32_t oct;
};

__device__ __forceinline__ void coop_init_ray(CoopWarp& cw, uint32_t lane, const TravRay& r, float tmax,
                                              uint32_t ignore_inst = 0xFFFFFFFFu) {
    cw.key[lane] = (unsigned long long)__float_as_uint(tmax) << 32;
    cw.ra[lane] = make_float4(r.sh.px.x, r.sh.px.y, r.sh.py.x, r.sh.py.y);
    cw.rb[lane] = make_float4(r.sh.pz.x, r.sh.pz.y, r.sh.cxy.x, r.sh.cxy.y);
    const uint32_t kzbits = (r.rbits >> 3) & 3u;
    const bool k0 = kzbits == 0u, k1 = (kzbits & 1u) != 0u;
    cw.rc[lane] = make_float4(k0 ? r.ox : (k1 ? r.oy : r.oz), k0 ? r.idx : (k1 ? r.idy : r.idz), __uint_as_float(kzbits), tmax);
    cw.tri[lane] = 0xFFFFFFFFu;
    cw.ignore[lane] = ignore_inst;
}

// one (ray of lane `owner`, triangle ti) pair; returns the key offered (0 = nothing offered).
// ANY (visibility rays): a hit in (tmin, tmax) on another instance than `ignore` marks the owner occluded — a flag,
// so the order of the tests cannot matter; nothing is offered.
template <bool ANY>
__device__ __forceinline__ unsigned long long coop_test(const SceneView& sc, CoopWarp& cw, uint32_t owner, uint32_t ti, float tmin,
                                                        float* uu, float* vv) {
    float4 q0, q1, q2;
    ldg_tri(sc.tris, ti, q0, q1, q2);
    if (ANY && __float_as_uint(q1.w) == cw.ignore[owner]) return 0ull;
    const float4 a = cw.ra[owner], b = cw.rb[owner];
    Shear sh;
    sh.px = make_float2(a.x, a.y); sh.py = make_float2(a.z, a.w); sh.pz = make_float2(b.x, b.y); sh.cxy = make_float2(b.z, b.w);
    float U, V, W, det;
    if (!tri_edges(sh, q0, q1, q2, &U, &V, &W, &det)) return 0ull;
    const float4 c = cw.rc[owner];  // only the few pairs that pass the edge test need the distance terms
    float tt;
    tri_finish(U, V, W, det, q0, q1, q2, __float_as_uint(c.z), c.x, c.y, &tt, uu, vv);
    if (!(tt > tmin)) return 0ull;
    if (ANY) {
        if (tt < c.w) cw.tri[owner] = 0u;
        return 0ull;
    }
    const unsigned long long key = ((unsigned long long)__float_as_uint(tt) << 32) | __float_as_uint(q0.w);
    const unsigned long long old = atomicMin(&cw.key[owner], key);
    return key < old ? key : 0ull;
}

// All 32 lanes call this after a node step; tmask / tbase: the lane's triangle bits and triangle base (0 when the
// lane has no ray).  On return every owner's closest hit (key, tri, uv) includes the pooled triangles.
//
// Dealing the pairs out: the lanes that reached triangles each leave ONE entry (triangle base, bits, first pool slot,
// lane) in shared memory, in lane order, hence sorted by first slot; the lane that tests pool slot w finds its entry
// by counting the entries that start at or before w (a REDUX.OR of the start slots + POPC) and clears w - start low
// bits of the entry's mask.  (Before: every owner wrote one item per pair in a loop that ran max-pairs-per-lane times
// with 2.7 lanes active, 9 % of the kernel's issue slots; the counts were scanned with five dependent SHFL.UP.)
template <bool ANY, bool COUNT>
__device__ __forceinline__ void coop_triangles(const SceneView& sc, CoopWarp& cw, uint32_t lane, uint32_t tmask, uint32_t tbase,
                                               float tmin, uint32_t& n_tris) {
    const uint32_t cnt = __popc(tmask);  // 0..24
    // inclusive scan of the counts by bit planes: five independent ballots instead of a five-step shuffle chain
    const uint32_t le = 0xFFFFFFFFu >> (31u - lane);
    uint32_t pre = 0u, total = 0u, owners = 0u;
#pragma unroll
    for (int b = 0; b < 5; b++) {
        const uint32_t plane = __ballot_sync(0xFFFFFFFFu, ((cnt >> b) & 1u) != 0u);
        pre += (uint32_t)__popc(plane & le) << b;
        total += (uint32_t)__popc(plane) << b;
        owners |= plane;
    }
    if (total == 0u) return;
    const uint32_t start = pre - cnt;
    if (cnt) cw.ent[__popc(owners & (le >> 1))] = make_uint4(tbase, tmask, start, lane);
    __syncwarp();
    for (uint32_t base = 0; base < total; base += 32u) {
        const uint32_t rel = start - base;  // wraps for entries that start before this round
        const uint32_t heads = __reduce_or_sync(0xFFFFFFFFu, (cnt != 0u && rel < 32u) ? (1u << rel) : 0u);
        const uint32_t before = base ? (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, cnt != 0u && start < base)) : 0u;
        const uint32_t w = base + lane;
        unsigned long long key = 0ull;
        uint32_t owner = 0, ti = 0;
        float uu = 0.0f, vv = 0.0f;
        if (w < total) {
            const uint4 e = cw.ent[before + (uint32_t)__popc(heads & le) - 1u];
            uint32_t m = e.y;
            for (uint32_t k = w - e.z; k; k--) m &= m - 1u;
            ti = e.x + (uint32_t)__ffs(m) - 1u;
            owner = e.w;
            if (COUNT) n_tris++;
            key = coop_test<ANY>(sc, cw, owner, ti, tmin, &uu, &vv);
        }
        __syncwarp();
        if (!ANY) {
            // of the lanes that lowered an owner's key this round, the one whose key stayed publishes the rest
            if (key != 0ull && cw.key[owner] == key) { cw.tri[owner] = ti; cw.uv[owner] = make_float2(uu, vv); }
            __syncwarp();
        }
    }
}

// node part of a traversal step for the cooperative kernel: pop the next child, slab-test its eight boxes.
// ng / sp / stk as in TravState; returns the triangle bits, *tbase the node's triangle base.
template <bool COUNT>
__device__ __forceinline__ uint32_t coop_node(const SceneView& sc, const NodeRay& nr, uint2& ng_io, int& sp, TravStack& stk, uint2& top,
                                              float tmin, float tmax, uint32_t* tbase, uint32_t& n_nodes) {
    TravRay r;  // node_hits reads only the members set here
    r.ox = nr.ox; r.oy = nr.oy; r.oz = nr.oz; r.idx = nr.idx; r.idy = nr.idy; r.idz = nr.idz; r.rbits = nr.oct;
    uint2 ng = ng_io;
    const int bit = 31 - __clz(ng.y);
    ng.y &= ~(1u << bit);
    const uint32_t slot = ((uint32_t)bit - 24u) ^ nr.oct;
    const uint32_t node = ng.x + __popc(ng.y & 0xFFu & ((1u << slot) - 1u));
    if (ng.y & 0xFF000000u) stk_push(stk, sp, top, ng);
    const uint32_t sx = nr.oct & 1u, sy = (nr.oct >> 1) & 1u, sz = (nr.oct >> 2) & 1u;
    uint4 w0, w1, nx, fx, ny, fy, nz, fz;
    load_node(sc.nodes, node, sx, sy, sz, w0, w1, nx, fx, ny, fy, nz, fz);
    if (COUNT) n_nodes++;
    const uint32_t hm = node_hits(w0, w1, nx, fx, ny, fy, nz, fz, r, tmin, tmax);
    ng_io = make_uint2(w1.x, (hm & 0xFF000000u) | (w0.w >> 24));
    *tbase = w1.y;
    return hm & 0x00FFFFFFu;
}

template <bool ANY, bool COUNT>
__device__ __forceinline__ Hit traverse(const SceneView& sc, vec3 o, vec3 d, float tmin, float tmax,
                                        uint32_t ignore_inst, uint32_t& n_nodes, uint32_t& n_tris) {
    TravState st;
    TravStack stk;
    trav_init(st, o, d, tmin, tmax, ignore_inst);
    while (!trav_step<ANY, COUNT>(sc, st, stk, n_nodes, n_tris)) {}
    return st.best;
}

}  // namespace rt
