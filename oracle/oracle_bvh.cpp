// ORACLE — test infrastructure only (see oracle_math.h header).  Parity status: pinned to the shipped SPIR-V (tests/golden/spirv_frames.npz); driver-defined parts unpinned.
//
// Stand-in for what the reference delegates to the Vulkan driver: acceleration-structure build
// (src/raytracer/acceleration_structure.odin:65-199, gpu_scene.odin:120-150) and the traversal
// behind traceRayEXT (shaders/simple.rgen:92-104) / rayQueryEXT (shaders/simple.rchit:362-371).
// The driver's algorithm is opaque, so only the API contract is restated: closest opaque hit in
// (tmin, tmax), no culling, all instances visible to mask 0xFF.  Ties on t are broken toward the
// lowest flat triangle id so brute force, this BVH8 and the GPU BVH8 agree bit for bit.
//
// The BVH8 builder is the deterministic CPU twin of the GPU builder: 63-bit Morton
// keys -> stable sort -> Karras radix tree -> surface-area guided collapse to 8-wide nodes ->
// octant slot assignment -> 8-bit quantised child boxes with one quantum of padding.
#include <algorithm>
#include <cstdio>

#include "oracle_scene.h"

namespace orc {

// --------------------------------------------------------------------------------------------
// scene finalisation: gpu_scene.odin:317-397 (objects, lights), flattening of instances
// --------------------------------------------------------------------------------------------
int scene_finalize(Scene& s) {
    s.lights.clear();
    s.tris.clear();
    uint32_t off = 0;
    for (size_t i = 0; i < s.objects.size(); i++) {
        Object& o = s.objects[i];
        if (o.mesh_index >= s.meshes.size()) return -2;
        if (o.material_index >= s.materials.size()) return -3;
        normal_matrix(o.model, o.nrm);
        o.tri_offset = off;
        // light_index stays 0 for every object after a bake: the reference writes it into a copy
        // (gpu_scene.odin:333-336)
        o.light_index = 0;
        const Mesh& m = s.meshes[o.mesh_index];
        uint32_t nt = (uint32_t)(m.idx.size() / 3);
        for (uint32_t p = 0; p < nt; p++) {
            uint32_t i0 = m.idx[3 * p], i1 = m.idx[3 * p + 1], i2 = m.idx[3 * p + 2];
            if (i0 >= m.v.size() || i1 >= m.v.size() || i2 >= m.v.size()) return -4;
            FlatTri t;
            t.p0 = xform_point(o.model, v3(m.v[i0].pos[0], m.v[i0].pos[1], m.v[i0].pos[2]));
            t.p1 = xform_point(o.model, v3(m.v[i1].pos[0], m.v[i1].pos[1], m.v[i1].pos[2]));
            t.p2 = xform_point(o.model, v3(m.v[i2].pos[0], m.v[i2].pos[1], m.v[i2].pos[2]));
            t.inst = (uint32_t)i;
            t.prim = p;
            s.tris.push_back(t);
        }
        off += nt;
        // a light is any object whose material has emission_power > 0 (gpu_scene.odin:366-379)
        if (s.materials[o.material_index].emission_power > 0.0f) {
            Light l;
            memcpy(l.transform, o.model, sizeof(l.transform));
            l.object_index = (uint32_t)i;
            l.num_triangles = nt;
            l.instance_mask = 1u << ((uint32_t)i & 7u);
            s.lights.push_back(l);
        }
    }
    s.has_bvh = false;
    return 0;
}

// --------------------------------------------------------------------------------------------
// watertight ray/triangle test (Woop, Benthin, Wald 2013), evaluation order per DESIGN.md §3
// --------------------------------------------------------------------------------------------
// The shear of Woop et al. is applied as an affine map so that no component has to be picked per triangle:
//   x' = mx . p - cx,  y' = my . p - cy   with   mx = e_kx - Sx e_kz,  my = e_ky - Sy e_kz,  c* = m* . o.
// Every dot product is the chain fmaf(p.z, m.z, fmaf(p.y, m.y, fmaf(p.x, m.x, -c))) (c itself:
// fmaf(o.z, m.z, fmaf(o.y, m.y, o.x * m.x))); entries that are exactly 0 or 1 contribute exactly.
struct RaySetup {
    float mx[3], my[3];
    float cx, cy;
    int kz;
    float o_kz, inv_d_kz;  // origin and 1/d along the dominant axis (1/d exactly as the slab test's idir)
};
static inline float comp(V3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
static inline float chain(const float m[3], V3 p, float c) { return fmaf(p.z, m[2], fmaf(p.y, m[1], fmaf(p.x, m[0], -c))); }
static inline float chain0(const float m[3], V3 p) { return fmaf(p.z, m[2], fmaf(p.y, m[1], p.x * m[0])); }

static inline RaySetup ray_setup(V3 o, V3 d) {
    RaySetup r;
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    // Woop et al. swap kx and ky here when d[kz] < 0 to keep the winding for back-face culling.  Nothing is
    // culled on this path (simple.rgen:92-104 traces with gl_RayFlagsOpaqueEXT only), and the swap merely
    // negates U, V, W, det and T together, so it is left out (it could only flip the sign of a zero u or v).
    float dz = comp(d, kz);
    for (int i = 0; i < 3; i++) r.mx[i] = r.my[i] = 0.0f;
    r.mx[kx] = 1.0f; r.mx[kz] = -(comp(d, kx) / dz);
    r.my[ky] = 1.0f; r.my[kz] = -(comp(d, ky) / dz);
    r.kz = kz;
    r.o_kz = comp(o, kz);
    r.inv_d_kz = (fabsf(dz) >= 1e-20f) ? 1.0f / dz : copysignf(1e20f, dz);
    r.cx = chain0(r.mx, o);
    r.cy = chain0(r.my, o);
    return r;
}

static inline bool tri_isect(const RaySetup& r, V3 p0, V3 p1, V3 p2, float* t, float* u, float* v) {
    float Ax = chain(r.mx, p0, r.cx), Ay = chain(r.my, p0, r.cy);
    float Bx = chain(r.mx, p1, r.cx), By = chain(r.my, p1, r.cy);
    float Cx = chain(r.mx, p2, r.cx), Cy = chain(r.my, p2, r.cy);
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    float det = (U + V) + W;
    if (det == 0.0f) return false;
    // distance from the hit point's coordinate along the dominant axis: t = (p_kz - o_kz) / d_kz
    float Az = comp(p0, r.kz), Bz = comp(p1, r.kz), Cz = comp(p2, r.kz);
    float T = fmaf(W, Cz, fmaf(V, Bz, U * Az));
    float inv = 1.0f / det;
    *t = (T * inv - r.o_kz) * r.inv_d_kz;
    *u = V * inv;
    *v = W * inv;
    return true;
}

static inline bool closer(float t, uint32_t id, const Hit& best) {
    return (t < best.t) || (t == best.t && id < best.id);
}

Hit closest_brute(const Scene& s, V3 o, V3 d, float tmin, float tmax) {
    RaySetup r = ray_setup(o, d);
    Hit best{tmax, 0.0f, 0.0f, 0xFFFFFFFFu};
    for (size_t i = 0; i < s.tris.size(); i++) {
        float t, u, v;
        const FlatTri& tr = s.tris[i];
        if (!tri_isect(r, tr.p0, tr.p1, tr.p2, &t, &u, &v)) continue;
        if (!(t > tmin)) continue;
        if (best.id == 0xFFFFFFFFu ? (t < best.t) : closer(t, (uint32_t)i, best)) best = Hit{t, u, v, (uint32_t)i};
    }
    return best;
}

bool occluded_brute(const Scene& s, V3 o, V3 d, float tmin, float tmax, uint32_t ignore_inst) {
    RaySetup r = ray_setup(o, d);
    for (size_t i = 0; i < s.tris.size(); i++) {
        const FlatTri& tr = s.tris[i];
        if (tr.inst == ignore_inst) continue;
        float t, u, v;
        if (!tri_isect(r, tr.p0, tr.p1, tr.p2, &t, &u, &v)) continue;
        if (t > tmin && t < tmax) return true;
    }
    return false;
}

// --------------------------------------------------------------------------------------------
// BVH8 build
// --------------------------------------------------------------------------------------------
static inline uint64_t spread21(uint32_t v) {
    uint64_t x = v & 0x1FFFFFu;
    x = (x | (x << 32)) & 0x1F00000000FFFFull;
    x = (x | (x << 16)) & 0x1F0000FF0000FFull;
    x = (x | (x << 8)) & 0x100F00F00F00F00Full;
    x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}

struct Box {
    V3 lo, hi;
};
static inline Box box_union(const Box& a, const Box& b) {
    return Box{v3(fminf(a.lo.x, b.lo.x), fminf(a.lo.y, b.lo.y), fminf(a.lo.z, b.lo.z)),
               v3(fmaxf(a.hi.x, b.hi.x), fmaxf(a.hi.y, b.hi.y), fmaxf(a.hi.z, b.hi.z))};
}
static inline float half_area(const Box& b) {
    V3 e = b.hi - b.lo;
    return (e.x * e.y + e.y * e.z) + e.z * e.x;
}

struct BinTree {
    uint32_t n = 0;  // leaves
    std::vector<uint32_t> left, right, first, count;  // internal nodes 0..n-2
    std::vector<Box> box;                             // all 2n-1 nodes
    bool is_leaf(uint32_t node) const { return n == 1 || node >= n - 1; }
    uint32_t cnt(uint32_t node) const { return is_leaf(node) ? 1u : count[node]; }
    uint32_t fst(uint32_t node) const { return is_leaf(node) ? (n == 1 ? 0u : node - (n - 1)) : first[node]; }
};

static inline int delta(const std::vector<uint64_t>& k, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    if (k[i] == k[j]) return 64 + __builtin_clz((uint32_t)(i ^ j));
    return __builtin_clzll(k[i] ^ k[j]);
}

static Box bin_box_rec(BinTree& bt, uint32_t node) {
    if (bt.is_leaf(node)) return bt.box[node];
    Box b = box_union(bin_box_rec(bt, bt.left[node]), bin_box_rec(bt, bt.right[node]));
    bt.box[node] = b;
    return b;
}

// binary16 <-> small non-negative integers (exact up to 2048): written out so the oracle needs no half type
static inline uint16_t half_of_int(int q) {
    if (q <= 0) return 0;
    int e = 31 - __builtin_clz((uint32_t)q);              // q = 1.m x 2^e, e <= 11
    uint32_t man = ((uint32_t)q << (10 - e)) & 0x3FFu;    // q <= 2048 loses no bits
    return (uint16_t)(((uint32_t)(e + 15) << 10) | man);
}
static inline float int_of_half(uint16_t h) {
    if (h == 0) return 0.0f;
    int e = (int)(h >> 10) - 15;
    uint32_t man = (h & 0x3FFu) | 0x400u;
    return (float)man * u2f((uint32_t)(e - 10 + 127) << 23);
}

static void quantize_axis(float lo, float hi, int* e_out, float* p_out) {
    float ext = hi - lo;
    float f = ext / ORC_QSTEPS;
    uint32_t bits = f2u(f);
    int e = (int)((bits >> 23) & 0xFFu) - 127 + ((bits & 0x7FFFFFu) ? 1 : 0);
    float mag = gmax(fabsf(lo), fabsf(hi));
    int mb = (int)((f2u(mag) >> 23) & 0xFFu) - 127;
    if (e < mb - 20) e = mb - 20;
    if (e < -100) e = -100;
    if (e > 100) e = 100;
    float scale = u2f((uint32_t)(e + 127) << 23);
    *e_out = e;
    *p_out = lo - scale;
}

void build_bvh8(Scene& s) {
    const uint32_t n = (uint32_t)s.tris.size();
    s.nodes.clear(); s.btris.clear(); s.morton.clear(); s.sorted_ids.clear();
    s.bvh_depth = 1;
    s.has_bvh = true;
    if (n == 0) {
        Node8 nd; memset(&nd, 0, sizeof(nd));
        for (int a = 0; a < 3; a++) { nd.e[a] = 127; for (int c = 0; c < 8; c++) { nd.q[a][0][c] = half_of_int(ORC_QMAX); nd.q[a][1][c] = half_of_int(0); } }
        s.nodes.push_back(nd);
        return;
    }
    // 1. triangle boxes, scene box
    std::vector<Box> tb(n);
    Box sb{v3s(INFINITY), v3s(-INFINITY)};
    for (uint32_t i = 0; i < n; i++) {
        const FlatTri& t = s.tris[i];
        tb[i].lo = v3(fminf(fminf(t.p0.x, t.p1.x), t.p2.x), fminf(fminf(t.p0.y, t.p1.y), t.p2.y), fminf(fminf(t.p0.z, t.p1.z), t.p2.z));
        tb[i].hi = v3(fmaxf(fmaxf(t.p0.x, t.p1.x), t.p2.x), fmaxf(fmaxf(t.p0.y, t.p1.y), t.p2.y), fmaxf(fmaxf(t.p0.z, t.p1.z), t.p2.z));
        sb = box_union(sb, tb[i]);
    }
    // 2. Morton keys, 21 bits per axis, from the box centre
    V3 ext = sb.hi - sb.lo;
    V3 sc = v3(ext.x > 0.0f ? 2097152.0f / ext.x : 0.0f, ext.y > 0.0f ? 2097152.0f / ext.y : 0.0f,
               ext.z > 0.0f ? 2097152.0f / ext.z : 0.0f);
    std::vector<std::pair<uint64_t, uint32_t>> keyed(n);
    for (uint32_t i = 0; i < n; i++) {
        V3 c = (tb[i].lo + tb[i].hi) * 0.5f;
        V3 q = (c - sb.lo) * sc;
        uint32_t qx = (uint32_t)gmin(gmax(q.x, 0.0f), 2097151.0f);
        uint32_t qy = (uint32_t)gmin(gmax(q.y, 0.0f), 2097151.0f);
        uint32_t qz = (uint32_t)gmin(gmax(q.z, 0.0f), 2097151.0f);
        keyed[i] = {(spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz), i};
    }
    // 3. sort by (key, flat id)
    std::sort(keyed.begin(), keyed.end());
    s.morton.resize(n); s.sorted_ids.resize(n);
    for (uint32_t i = 0; i < n; i++) { s.morton[i] = keyed[i].first; s.sorted_ids[i] = keyed[i].second; }

    // 4. Karras radix tree; leaves are nodes n-1 .. 2n-2
    BinTree bt;
    bt.n = n;
    bt.box.resize(n == 1 ? 1 : 2 * n - 1);
    if (n > 1 && getenv("ORC_SAH_EXPERIMENT")) {
        // EXPERIMENT (not the twin of anything on the GPU): top-down binned SAH over the same leaves, to measure what a
        // SAH-quality binary tree would buy the collapsed BVH8 before building it on the device
        std::vector<uint32_t> perm(s.sorted_ids);
        bt.left.assign(n - 1, 0); bt.right.assign(n - 1, 0); bt.first.assign(n - 1, 0); bt.count.assign(n - 1, 0);
        uint32_t next_internal = 0;
        struct Job { uint32_t lo, hi, node; };
        std::vector<Job> stack;
        auto alloc = [&](uint32_t lo, uint32_t hi) -> uint32_t {  // node id of the range
            if (hi - lo == 1) return n - 1 + lo;
            const uint32_t id = next_internal++;
            stack.push_back({lo, hi, id});
            return id;
        };
        alloc(0, n);
        const int NB = 32;
        while (!stack.empty()) {
            Job j = stack.back(); stack.pop_back();
            Box cb{v3s(INFINITY), v3s(-INFINITY)};
            for (uint32_t k = j.lo; k < j.hi; k++) { V3 c = (tb[perm[k]].lo + tb[perm[k]].hi) * 0.5f; cb.lo = v3(fminf(cb.lo.x, c.x), fminf(cb.lo.y, c.y), fminf(cb.lo.z, c.z)); cb.hi = v3(fmaxf(cb.hi.x, c.x), fmaxf(cb.hi.y, c.y), fmaxf(cb.hi.z, c.z)); }
            float best = INFINITY; int best_axis = -1, best_bin = -1;
            for (int ax = 0; ax < 3; ax++) {
                const float lo = comp(cb.lo, ax), hi = comp(cb.hi, ax);
                if (!(hi > lo)) continue;
                Box bb[NB]; uint32_t bc[NB];
                for (int b = 0; b < NB; b++) { bb[b] = Box{v3s(INFINITY), v3s(-INFINITY)}; bc[b] = 0; }
                const float sc2 = NB / (hi - lo);
                for (uint32_t k = j.lo; k < j.hi; k++) {
                    const Box& t = tb[perm[k]];
                    int b = (int)((comp((t.lo + t.hi) * 0.5f, ax) - lo) * sc2); if (b >= NB) b = NB - 1; if (b < 0) b = 0;
                    bb[b] = box_union(bb[b], t); bc[b]++;
                }
                Box acc{v3s(INFINITY), v3s(-INFINITY)}; float la[NB]; uint32_t lc[NB]; uint32_t c = 0;
                for (int b = 0; b < NB - 1; b++) { acc = box_union(acc, bb[b]); c += bc[b]; la[b] = c ? half_area(acc) : 0.0f; lc[b] = c; }
                acc = Box{v3s(INFINITY), v3s(-INFINITY)}; c = 0;
                for (int b = NB - 1; b > 0; b--) {
                    acc = box_union(acc, bb[b]); c += bc[b];
                    if (c == 0 || lc[b - 1] == 0) continue;
                    const float cost = la[b - 1] * lc[b - 1] + half_area(acc) * c;
                    if (cost < best) { best = cost; best_axis = ax; best_bin = b; }
                }
            }
            uint32_t mid;
            if (best_axis < 0) mid = (j.lo + j.hi) / 2;
            else {
                const float lo = comp(cb.lo, best_axis), sc2 = NB / (comp(cb.hi, best_axis) - lo);
                auto it = std::stable_partition(perm.begin() + j.lo, perm.begin() + j.hi, [&](uint32_t id) {
                    int b = (int)((comp((tb[id].lo + tb[id].hi) * 0.5f, best_axis) - lo) * sc2); if (b >= NB) b = NB - 1; if (b < 0) b = 0;
                    return b < best_bin; });
                mid = (uint32_t)(it - perm.begin());
                if (mid == j.lo || mid == j.hi) mid = (j.lo + j.hi) / 2;
            }
            bt.first[j.node] = j.lo; bt.count[j.node] = j.hi - j.lo;
            const uint32_t l = alloc(j.lo, mid), r = alloc(mid, j.hi);
            bt.left[j.node] = l; bt.right[j.node] = r;
        }
        s.sorted_ids = perm;
        for (uint32_t k = 0; k < n; k++) bt.box[n - 1 + k] = tb[s.sorted_ids[k]];
        std::vector<uint32_t> order; std::vector<uint32_t> st{0};
        while (!st.empty()) { uint32_t nd = st.back(); st.pop_back(); order.push_back(nd);
            if (!bt.is_leaf(bt.left[nd])) st.push_back(bt.left[nd]);
            if (!bt.is_leaf(bt.right[nd])) st.push_back(bt.right[nd]); }
        for (size_t q = order.size(); q-- > 0;) bt.box[order[q]] = box_union(bt.box[bt.left[order[q]]], bt.box[bt.right[order[q]]]);
    } else if (n == 1) {
        bt.box[0] = tb[s.sorted_ids[0]];
    } else {
        bt.left.resize(n - 1); bt.right.resize(n - 1); bt.first.resize(n - 1); bt.count.resize(n - 1);
        for (uint32_t k = 0; k < n; k++) bt.box[n - 1 + k] = tb[s.sorted_ids[k]];
        const std::vector<uint64_t>& K = s.morton;
        int N = (int)n;
        for (int i = 0; i < N - 1; i++) {
            int d = (delta(K, N, i, i + 1) - delta(K, N, i, i - 1)) >= 0 ? 1 : -1;
            int dmin = delta(K, N, i, i - d);
            int lmax = 2;
            while (delta(K, N, i, i + lmax * d) > dmin) lmax *= 2;
            int l = 0;
            for (int t = lmax / 2; t >= 1; t /= 2)
                if (delta(K, N, i, i + (l + t) * d) > dmin) l += t;
            int j = i + l * d;
            int dnode = delta(K, N, i, j);
            int sp = 0, t = l;
            do {
                t = (t + 1) >> 1;
                if (delta(K, N, i, i + (sp + t) * d) > dnode) sp += t;
            } while (t > 1);
            int gamma = i + sp * d + std::min(d, 0);
            int lo = std::min(i, j), hi = std::max(i, j);
            bt.left[i] = (lo == gamma) ? (uint32_t)(N - 1 + gamma) : (uint32_t)gamma;
            bt.right[i] = (hi == gamma + 1) ? (uint32_t)(N - 1 + gamma + 1) : (uint32_t)(gamma + 1);
            bt.first[i] = (uint32_t)lo;
            bt.count[i] = (uint32_t)(hi - lo + 1);
        }
        // 5. boxes bottom-up (exact min/max, order independent); iterative to survive deep trees
        {
            std::vector<uint32_t> order; order.reserve(n - 1);
            std::vector<uint32_t> st{0};
            while (!st.empty()) {
                uint32_t nd = st.back(); st.pop_back();
                order.push_back(nd);
                if (!bt.is_leaf(bt.left[nd])) st.push_back(bt.left[nd]);
                if (!bt.is_leaf(bt.right[nd])) st.push_back(bt.right[nd]);
            }
            for (size_t q = order.size(); q-- > 0;) {
                uint32_t nd = order[q];
                bt.box[nd] = box_union(bt.box[bt.left[nd]], bt.box[bt.right[nd]]);
            }
        }
    }

    // 6. collapse, level by level (BFS numbering makes node and triangle order deterministic)
    struct Wide { uint32_t bin; };
    std::vector<uint32_t> level{0u};  // binary node of each wide node in this level (root = 0)
    if (n == 1) level[0] = 0;
    uint32_t nodes_total = 1, depth = 0;
    while (!level.empty()) {
        depth++;
        std::vector<uint32_t> next;
        uint32_t level_first = (uint32_t)s.nodes.size();
        (void)level_first;
        for (size_t w = 0; w < level.size(); w++) {
            uint32_t b = level[w];
            uint32_t ch[8]; int m = 1; ch[0] = b;
            for (int phase = 0; phase < 2; phase++) {
                while (m < 8) {
                    int best = -1; float best_a = -1.0f;
                    for (int c = 0; c < m; c++) {
                        if (bt.is_leaf(ch[c])) continue;
                        uint32_t cn = bt.cnt(ch[c]);
                        if (phase == 0 ? (cn <= 3) : (cn < 2)) continue;
                        float a = half_area(bt.box[ch[c]]);
                        if (best < 0 || a > best_a) { best = c; best_a = a; }
                    }
                    if (best < 0) break;
                    uint32_t nd = ch[best];
                    for (int c = m; c > best + 1; c--) ch[c] = ch[c - 1];
                    ch[best] = bt.left[nd]; ch[best + 1] = bt.right[nd];
                    m++;
                }
            }
            // a lone leaf child can only happen for n == 1
            Box nb = bt.box[b];
            V3 centre = (nb.lo + nb.hi) * 0.5f;
            // 7. greedy octant slot assignment: slot bit k set <=> child sits on the negative side of axis k
            int slot_of[8]; bool slot_used[8] = {false}; bool child_done[8] = {false};
            float cost[8][8];
            for (int c = 0; c < m; c++) {
                const Box& cb = bt.box[ch[c]];
                V3 dc = ((cb.lo + cb.hi) * 0.5f) - centre;
                for (int sl = 0; sl < 8; sl++) {
                    float sx = (sl & 1) ? -dc.x : dc.x, sy = (sl & 2) ? -dc.y : dc.y, sz = (sl & 4) ? -dc.z : dc.z;
                    cost[c][sl] = (sx + sy) + sz;
                }
            }
            for (int it = 0; it < m; it++) {
                int bc = -1, bs = -1; float bv = 0.0f;
                for (int c = 0; c < m; c++) {
                    if (child_done[c]) continue;
                    for (int sl = 0; sl < 8; sl++) {
                        if (slot_used[sl]) continue;
                        if (bc < 0 || cost[c][sl] > bv) { bc = c; bs = sl; bv = cost[c][sl]; }
                    }
                }
                slot_of[bc] = bs; child_done[bc] = true; slot_used[bs] = true;
            }
            int child_in_slot[8];
            for (int sl = 0; sl < 8; sl++) child_in_slot[sl] = -1;
            for (int c = 0; c < m; c++) child_in_slot[slot_of[c]] = c;

            // 8. emit node
            Node8 nd; memset(&nd, 0, sizeof(nd));
            int e[3];
            quantize_axis(nb.lo.x, nb.hi.x, &e[0], &nd.p[0]);
            quantize_axis(nb.lo.y, nb.hi.y, &e[1], &nd.p[1]);
            quantize_axis(nb.lo.z, nb.hi.z, &e[2], &nd.p[2]);
            float inv_scale[3];
            for (int a = 0; a < 3; a++) { nd.e[a] = (uint8_t)(e[a] + 127); inv_scale[a] = u2f((uint32_t)(127 - e[a]) << 23); }
            nd.child_base = nodes_total;
            nd.tri_base = (uint32_t)s.btris.size();
            uint32_t n_int = 0, tri_off = 0;
            for (int sl = 0; sl < 8; sl++) {
                int c = child_in_slot[sl];
                if (c < 0) {
                    nd.meta[sl] = 0;
                    for (int a = 0; a < 3; a++) { nd.q[a][0][sl] = half_of_int(ORC_QMAX); nd.q[a][1][sl] = half_of_int(0); }
                    continue;
                }
                const Box& cb = bt.box[ch[c]];
                const float clo[3] = {cb.lo.x, cb.lo.y, cb.lo.z}, chi[3] = {cb.hi.x, cb.hi.y, cb.hi.z};
                for (int a = 0; a < 3; a++) {
                    int ql = (int)floorf((clo[a] - nd.p[a]) * inv_scale[a]) - 1;
                    int qh = (int)ceilf((chi[a] - nd.p[a]) * inv_scale[a]) + 1;
                    nd.q[a][0][sl] = half_of_int(std::min(std::max(ql, 0), ORC_QMAX));
                    nd.q[a][1][sl] = half_of_int(std::min(std::max(qh, 0), ORC_QMAX));
                }
                uint32_t cn = bt.cnt(ch[c]);
                if (!bt.is_leaf(ch[c]) && cn > 3) {
                    nd.imask |= (uint8_t)(1u << sl);
                    nd.meta[sl] = (uint8_t)(0x20u | (24u + (uint32_t)sl));
                    next.push_back(ch[c]);
                    n_int++;
                } else {
                    uint32_t unary = (cn == 1) ? 1u : (cn == 2 ? 3u : 7u);
                    nd.meta[sl] = (uint8_t)((unary << 5) | tri_off);
                    uint32_t f0 = bt.fst(ch[c]);
                    for (uint32_t q = 0; q < cn; q++) {
                        uint32_t id = s.sorted_ids[f0 + q];
                        const FlatTri& ft = s.tris[id];
                        BvhTri t;
                        t.p0[0] = ft.p0.x; t.p0[1] = ft.p0.y; t.p0[2] = ft.p0.z; t.id = id;
                        t.p1[0] = ft.p1.x; t.p1[1] = ft.p1.y; t.p1[2] = ft.p1.z; t.inst = ft.inst;
                        t.p2[0] = ft.p2.x; t.p2[1] = ft.p2.y; t.p2[2] = ft.p2.z; t.prim = ft.prim;
                        s.btris.push_back(t);
                    }
                    tri_off += cn;
                }
            }
            nodes_total += n_int;
            s.nodes.push_back(nd);
        }
        level.swap(next);
    }
    s.bvh_depth = depth;
}

// --------------------------------------------------------------------------------------------
// BVH8 traversal (scalar).  Group stack: (base, mask) with internal-child hit bits in [31:24]
// ordered by slot ^ ray_octant, triangle hit bits in [23:0].
// --------------------------------------------------------------------------------------------
struct TravRay {
    RaySetup rs;
    float idir[3];
    float o[3];
    uint32_t rbits;
};
static inline TravRay trav_setup(V3 o, V3 d) {
    TravRay r;
    r.rs = ray_setup(o, d);
    const float dd[3] = {d.x, d.y, d.z};
    r.o[0] = o.x; r.o[1] = o.y; r.o[2] = o.z;
    r.rbits = 0;
    for (int a = 0; a < 3; a++) {
        r.idir[a] = (fabsf(dd[a]) >= 1e-20f) ? 1.0f / dd[a] : copysignf(1e20f, dd[a]);
        if (std::signbit(dd[a])) r.rbits |= 1u << a;
    }
    return r;
}

// returns hit mask: bits [31:24] internal children at priority positions, [23:0] triangles
static inline uint32_t node_hits(const Node8& nd, const TravRay& r, float tmin, float tmax) {
    float sc[3], orr[3];
    for (int a = 0; a < 3; a++) {
        sc[a] = u2f((uint32_t)nd.e[a] << 23) * r.idir[a];
        orr[a] = (nd.p[a] - r.o[a]) * r.idir[a];
    }
    uint32_t mask = 0;
    for (int sl = 0; sl < 8; sl++) {
        uint32_t meta = nd.meta[sl];
        if (!meta) continue;
        float tn[3], tf[3];
        for (int a = 0; a < 3; a++) {
            bool negd = (r.rbits >> a) & 1u;
            float qn = int_of_half(nd.q[a][negd ? 1 : 0][sl]);
            float qf = int_of_half(nd.q[a][negd ? 0 : 1][sl]);
            tn[a] = fmaf(qn, sc[a], orr[a]);
            tf[a] = fmaf(qf, sc[a], orr[a]);
        }
        float tnear = fmaxf(fmaxf(tn[0], tn[1]), fmaxf(tn[2], tmin));
        float tfar = fminf(fminf(tf[0], tf[1]), fminf(tf[2], tmax));
        if (tnear <= tfar) {
            uint32_t bits = meta >> 5, idx = meta & 31u;
            if (nd.imask & (1u << sl)) idx = 24u + ((uint32_t)sl ^ r.rbits);
            mask |= bits << idx;
        }
    }
    return mask;
}

template <bool ANY>
static inline Hit traverse(const Scene& s, V3 o, V3 d, float tmin, float tmax, uint32_t ignore_inst, Counters* cnt) {
    TravRay r = trav_setup(o, d);
    Hit best{tmax, 0.0f, 0.0f, 0xFFFFFFFFu};
    struct G { uint32_t base, mask; };
    G stack[64]; int sp = 0;
    G ng{0u, 0x80000000u};  // root: bit 31, imask 0 -> node 0
    G tg{0u, 0u};
    uint64_t n_nodes = 0, n_tris = 0;
    for (;;) {
        if (ng.mask & 0xFF000000u) {
            int bit = 31 - __builtin_clz(ng.mask);
            ng.mask &= ~(1u << bit);
            uint32_t slot = ((uint32_t)bit - 24u) ^ r.rbits;
            uint32_t rel = (uint32_t)__builtin_popcount(ng.mask & 0xFFu & ((1u << slot) - 1u));
            uint32_t node = ng.base + rel;
            if (ng.mask & 0xFF000000u) stack[sp++] = ng;
            const Node8& nd = s.nodes[node];
            n_nodes++;
            uint32_t hm = node_hits(nd, r, tmin, best.t);
            ng = G{nd.child_base, (hm & 0xFF000000u) | nd.imask};
            tg = G{nd.tri_base, hm & 0x00FFFFFFu};
        } else {
            tg = G{0u, 0u};
            ng = G{0u, 0u};
        }
        while (tg.mask) {
            int bit = __builtin_ctz(tg.mask);
            tg.mask &= tg.mask - 1u;
            const BvhTri& t = s.btris[tg.base + (uint32_t)bit];
            n_tris++;
            if (ANY && t.inst == ignore_inst) continue;
            float tt, uu, vv;
            if (!tri_isect(r.rs, v3(t.p0[0], t.p0[1], t.p0[2]), v3(t.p1[0], t.p1[1], t.p1[2]), v3(t.p2[0], t.p2[1], t.p2[2]), &tt, &uu, &vv)) continue;
            if (!(tt > tmin)) continue;
            if (ANY) {
                if (tt < tmax) { best = Hit{tt, uu, vv, t.id}; goto done; }
            } else {
                if (best.id == 0xFFFFFFFFu ? (tt < best.t) : closer(tt, t.id, best)) best = Hit{tt, uu, vv, t.id};
            }
        }
        if (!(ng.mask & 0xFF000000u)) {
            if (sp == 0) break;
            ng = stack[--sp];
        }
    }
done:
    if (cnt) { cnt->nodes += n_nodes; cnt->tris += n_tris; }
    return best;
}

Hit closest_bvh(const Scene& s, V3 o, V3 d, float tmin, float tmax, Counters* c) {
    return traverse<false>(s, o, d, tmin, tmax, 0xFFFFFFFFu, c);
}
bool occluded_bvh(const Scene& s, V3 o, V3 d, float tmin, float tmax, uint32_t ignore_inst, Counters* c) {
    return traverse<true>(s, o, d, tmin, tmax, ignore_inst, c).id != 0xFFFFFFFFu;
}

}  // namespace orc
