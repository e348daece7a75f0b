// rt_build.cu — GPU BVH builder: the CUDA replacement of the driver-side BLAS/TLAS build the
// reference requests in src/raytracer/acceleration_structure.odin:65-199 and gpu_scene.odin:120-150
// (vkCmdBuildAccelerationStructuresKHR; the driver's algorithm is opaque, only the API contract —
// every instanced triangle is in the structure — carries over).
//
// Pipeline (all on the device, deterministic: no result depends on thread scheduling):
//   1 flatten instances to world space, per-triangle boxes, scene box (exact min/max)
//   2 63-bit Morton keys of the box centres
//   3 stable LSD radix sort (8 passes x 8 bits) of (key, flat id)
//   4 Karras (2012) radix tree, ties broken by index
//   5 boxes bottom-up (atomic arrival counters; min/max are order independent)
//   6 level-by-level collapse to 8-wide nodes guided by surface area, octant slot assignment,
//     8-bit quantised child boxes (one quantum of padding), triangles emitted in leaf order
// The oracle holds a CPU twin of the same specification for the bit-exact node/triangle comparison.
#include <stdio.h>

#include "rt_internal.h"

namespace rt {

// ------------------------------------------------------------------------------------------------
// exclusive scan of uint32 (in place), 1024 threads x 4 items per block, recursive on block sums
// ------------------------------------------------------------------------------------------------
#define SCAN_THREADS 1024
#define SCAN_ITEMS 4
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

__global__ void __launch_bounds__(SCAN_THREADS) scan_block_kernel(uint32_t* data, uint32_t n, uint32_t* block_sums) {
    __shared__ uint32_t warp_sums[32];
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t local = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = (base + k < n) ? data[base + k] : 0u;
        local += v[k];
    }
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= (uint32_t)o) incl += t;
    }
    if (lane == 31u) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = warp_sums[lane];
        uint32_t wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xFFFFFFFFu, wi, o);
            if (lane >= (uint32_t)o) wi += t;
        }
        warp_sums[lane] = wi - w;  // exclusive
        if (lane == 31u && block_sums) block_sums[blockIdx.x] = wi;
    }
    __syncthreads();
    uint32_t run = warp_sums[warp] + (incl - local);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) data[base + k] = run;
        run += v[k];
    }
}

__global__ void scan_add_kernel(uint32_t* data, uint32_t n, const uint32_t* block_offsets) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) data[i] += block_offsets[i / SCAN_TILE];
}

struct Scanner {
    PoolBuf<uint32_t> lvl1, lvl2;
    cudaError_t reserve(size_t n, cudaStream_t s) {
        size_t b1 = (n + SCAN_TILE - 1) / SCAN_TILE;
        size_t b2 = (b1 + SCAN_TILE - 1) / SCAN_TILE;
        cudaError_t e = lvl1.alloc(b1 + 1, s);
        if (e != cudaSuccess) return e;
        return lvl2.alloc(b2 + 1, s);
    }
    // exclusive scan in place; n <= SCAN_TILE^3
    void run(uint32_t* data, uint32_t n, cudaStream_t s) {
        if (n == 0) return;
        const uint32_t b1 = (n + SCAN_TILE - 1) / SCAN_TILE;
        scan_block_kernel<<<b1, SCAN_THREADS, 0, s>>>(data, n, lvl1.p);
        if (b1 > 1) {
            const uint32_t b2 = (b1 + SCAN_TILE - 1) / SCAN_TILE;
            scan_block_kernel<<<b2, SCAN_THREADS, 0, s>>>(lvl1.p, b1, lvl2.p);
            if (b2 > 1) {
                scan_block_kernel<<<1, SCAN_THREADS, 0, s>>>(lvl2.p, b2, nullptr);
                scan_add_kernel<<<(b1 + 255) / 256, 256, 0, s>>>(lvl1.p, b1, lvl2.p);
            }
            scan_add_kernel<<<(n + 255) / 256, 256, 0, s>>>(data, n, lvl1.p);
        }
    }
};

// ------------------------------------------------------------------------------------------------
// stable LSD radix sort of (u64 key, u32 value): 8 bits per pass
// ------------------------------------------------------------------------------------------------
#define SORT_THREADS 256
#define SORT_ITEMS 8
#define SORT_TILE (SORT_THREADS * SORT_ITEMS)

__global__ void __launch_bounds__(SORT_THREADS) sort_hist_kernel(const unsigned long long* keys, uint32_t n, int shift,
                                                                uint32_t* hist, uint32_t n_tiles) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * SORT_TILE;
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        const uint32_t i = base + k * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(uint32_t)(keys[i] >> shift) & 0xFFu], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];  // digit-major
}

// Each warp owns SORT_ITEMS consecutive rows of 32 keys; ranks inside the tile keep input order.
__global__ void __launch_bounds__(SORT_THREADS) sort_scatter_kernel(const unsigned long long* keys_in,
                                                                   const uint32_t* vals_in, unsigned long long* keys_out,
                                                                   uint32_t* vals_out, uint32_t n, int shift,
                                                                   const uint32_t* offsets, uint32_t n_tiles) {
    __shared__ uint32_t cnt[SORT_THREADS / 32][256];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (uint32_t i = threadIdx.x; i < (SORT_THREADS / 32) * 256; i += SORT_THREADS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t wbase = blockIdx.x * SORT_TILE + warp * (32 * SORT_ITEMS);
    unsigned long long key[SORT_ITEMS];
    uint32_t rank[SORT_ITEMS];
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        const uint32_t i = wbase + k * 32 + lane;
        const bool valid = i < n;
        key[k] = valid ? keys_in[i] : 0ull;
        const uint32_t d = valid ? ((uint32_t)(key[k] >> shift) & 0xFFu) : 0xFFFFFFFFu;
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);
        const uint32_t leader = __ffs(peers) - 1;
        uint32_t b = 0;
        if (valid && lane == leader) {
            b = cnt[warp][d];
            cnt[warp][d] = b + __popc(peers);
        }
        b = __shfl_sync(0xFFFFFFFFu, b, leader);
        rank[k] = b + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();
    {   // per digit: exclusive prefix over the warps, plus the tile's global offset
        const uint32_t d = threadIdx.x;
        uint32_t run = offsets[d * n_tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < SORT_THREADS / 32; w++) {
            const uint32_t c = cnt[w][d];
            cnt[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        const uint32_t i = wbase + k * 32 + lane;
        if (i < n) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & 0xFFu;
            const uint32_t dst = cnt[warp][d] + rank[k];
            keys_out[dst] = key[k];
            vals_out[dst] = vals_in[i];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// build kernels
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t enc_f(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float dec_f(uint32_t e) {
    return __uint_as_float((e & 0x80000000u) ? (e & 0x7FFFFFFFu) : ~e);
}

struct BuildScratch {
    uint32_t scene_box[6];   // encoded lo.xyz, hi.xyz
    uint32_t level_int;      // internal children of the level just collapsed
    uint32_t level_tris;     // triangles emitted by that level
    double sa_sum;           // half surface areas of all wide nodes (refit quality measure only: summation order is free)
};

__global__ void init_scratch_kernel(BuildScratch* s) {
    s->scene_box[0] = s->scene_box[1] = s->scene_box[2] = 0xFFFFFFFFu;
    s->scene_box[3] = s->scene_box[4] = s->scene_box[5] = 0u;
    s->level_int = 0; s->level_tris = 0;
    s->sa_sum = 0.0;
}

// world-space vertex of an instance: the one evaluation order the flatten pass, the refit pass and the CPU twin share
__device__ __forceinline__ vec3 instance_vertex(const DevObject* o, const rt_vertex* v) {
    const vec3 q = mk3(v->pos[0], v->pos[1], v->pos[2]);
    const float* m = o->model;
    return mk3(((m[0] * q.x + m[3] * q.y) + m[6] * q.z) + m[9], ((m[1] * q.x + m[4] * q.y) + m[7] * q.z) + m[10],
               ((m[2] * q.x + m[5] * q.y) + m[8] * q.z) + m[11]);
}

// 1. flat triangle i -> (object, primitive) by binary search over the object offsets
__global__ void flatten_kernel(BuildInput in, float4* flat, float4* box_lo, float4* box_hi, BuildScratch* scratch) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (i < in.n_tris) {
        uint32_t a = 0, b = in.n_objects;  // last object with tri_offsets[obj] <= i
        while (b - a > 1) {
            const uint32_t m = (a + b) >> 1;
            if (in.d_tri_offsets[m] <= i) a = m; else b = m;
        }
        const DevObject* o = in.d_objects + a;
        const uint32_t prim = i - in.d_tri_offsets[a];
        const uint32_t* ip = in.d_indices + o->index_base + 3u * prim;
        const rt_vertex* vb = in.d_vertices + o->vertex_base;
        vec3 p[3];
#pragma unroll
        for (int k = 0; k < 3; k++) p[k] = instance_vertex(o, vb + ip[k]);
        flat[3 * (size_t)i + 0] = make_float4(p[0].x, p[0].y, p[0].z, __uint_as_float(i));
        flat[3 * (size_t)i + 1] = make_float4(p[1].x, p[1].y, p[1].z, __uint_as_float(a));
        flat[3 * (size_t)i + 2] = make_float4(p[2].x, p[2].y, p[2].z, __uint_as_float(prim));
        lo[0] = fminf(fminf(p[0].x, p[1].x), p[2].x); lo[1] = fminf(fminf(p[0].y, p[1].y), p[2].y);
        lo[2] = fminf(fminf(p[0].z, p[1].z), p[2].z);
        hi[0] = fmaxf(fmaxf(p[0].x, p[1].x), p[2].x); hi[1] = fmaxf(fmaxf(p[0].y, p[1].y), p[2].y);
        hi[2] = fmaxf(fmaxf(p[0].z, p[1].z), p[2].z);
        box_lo[i] = make_float4(lo[0], lo[1], lo[2], 0.0f);
        box_hi[i] = make_float4(hi[0], hi[1], hi[2], 0.0f);
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
        float l = lo[a], h = hi[a];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            l = fminf(l, __shfl_xor_sync(0xFFFFFFFFu, l, o));
            h = fmaxf(h, __shfl_xor_sync(0xFFFFFFFFu, h, o));
        }
        if ((threadIdx.x & 31u) == 0u) {
            if (l != INFINITY) atomicMin(&scratch->scene_box[a], enc_f(l));
            if (h != -INFINITY) atomicMax(&scratch->scene_box[3 + a], enc_f(h));
        }
    }
}

__device__ __forceinline__ unsigned long long spread21(uint32_t v) {
    unsigned long long x = v & 0x1FFFFFu;
    x = (x | (x << 32)) & 0x1F00000000FFFFull;
    x = (x | (x << 16)) & 0x1F0000FF0000FFull;
    x = (x | (x << 8)) & 0x100F00F00F00F00Full;
    x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}

// 2. Morton keys: 21 bits per axis of the box centre, x in the highest position
__global__ void morton_kernel(const float4* box_lo, const float4* box_hi, uint32_t n, const BuildScratch* scratch,
                              unsigned long long* keys, uint32_t* ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const vec3 slo = mk3(dec_f(scratch->scene_box[0]), dec_f(scratch->scene_box[1]), dec_f(scratch->scene_box[2]));
    const vec3 shi = mk3(dec_f(scratch->scene_box[3]), dec_f(scratch->scene_box[4]), dec_f(scratch->scene_box[5]));
    const vec3 ext = sub3(shi, slo);
    const vec3 sc = mk3(ext.x > 0.0f ? 2097152.0f / ext.x : 0.0f, ext.y > 0.0f ? 2097152.0f / ext.y : 0.0f,
                        ext.z > 0.0f ? 2097152.0f / ext.z : 0.0f);
    const float4 l = box_lo[i], h = box_hi[i];
    const vec3 c = scale3(add3(mk3(l.x, l.y, l.z), mk3(h.x, h.y, h.z)), 0.5f);
    const vec3 q = mul3(sub3(c, slo), sc);
    const uint32_t qx = (uint32_t)glsl_min(glsl_max(q.x, 0.0f), 2097151.0f);
    const uint32_t qy = (uint32_t)glsl_min(glsl_max(q.y, 0.0f), 2097151.0f);
    const uint32_t qz = (uint32_t)glsl_min(glsl_max(q.z, 0.0f), 2097151.0f);
    keys[i] = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
    ids[i] = i;
}

// 4. Karras radix tree.  Internal nodes 0..n-2, leaf k is node n-1+k.
__device__ __forceinline__ int delta_fn(const unsigned long long* k, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = k[i], b = k[j];
    if (a == b) return 64 + __clz((uint32_t)(i ^ j));
    return __clzll((long long)(a ^ b));
}

__global__ void karras_kernel(const unsigned long long* keys, int n, uint32_t* left, uint32_t* right, uint32_t* first,
                              uint32_t* count, uint32_t* parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta_fn(keys, n, i, i + 1) - delta_fn(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta_fn(keys, n, i, i - d);
    int lmax = 2;
    while (delta_fn(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta_fn(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta_fn(keys, n, i, j);
    int sp = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta_fn(keys, n, i, i + (sp + t) * d) > dnode) sp += t;
    } while (t > 1);
    const int gamma = i + sp * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const uint32_t lc = (lo == gamma) ? (uint32_t)(n - 1 + gamma) : (uint32_t)gamma;
    const uint32_t rc = (hi == gamma + 1) ? (uint32_t)(n - 1 + gamma + 1) : (uint32_t)(gamma + 1);
    left[i] = lc; right[i] = rc;
    first[i] = (uint32_t)lo; count[i] = (uint32_t)(hi - lo + 1);
    parent[lc] = (uint32_t)i; parent[rc] = (uint32_t)i;
    if (i == 0) parent[0] = 0xFFFFFFFFu;
}

// 5. leaf boxes, then climb: the second thread to arrive at a node merges its children
__global__ void boxes_kernel(const float4* tri_lo, const float4* tri_hi, const uint32_t* sorted_ids, uint32_t n,
                             const uint32_t* left, const uint32_t* right, const uint32_t* parent, uint32_t* arrive,
                             float4* nlo, float4* nhi) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t id = sorted_ids[k];
    if (n == 1) { nlo[0] = tri_lo[id]; nhi[0] = tri_hi[id]; return; }
    uint32_t node = n - 1 + k;
    nlo[node] = tri_lo[id]; nhi[node] = tri_hi[id];
    __threadfence();
    uint32_t p = parent[node];
    while (p != 0xFFFFFFFFu) {
        if (atomicAdd(&arrive[p], 1u) == 0u) return;
        __threadfence();
        const uint32_t a = left[p], b = right[p];
        const volatile float4* vlo = nlo; const volatile float4* vhi = nhi;
        float4 al, ah, bl, bh;
        al.x = vlo[a].x; al.y = vlo[a].y; al.z = vlo[a].z; ah.x = vhi[a].x; ah.y = vhi[a].y; ah.z = vhi[a].z;
        bl.x = vlo[b].x; bl.y = vlo[b].y; bl.z = vlo[b].z; bh.x = vhi[b].x; bh.y = vhi[b].y; bh.z = vhi[b].z;
        nlo[p] = make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), 0.0f);
        nhi[p] = make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0.0f);
        __threadfence();
        p = parent[p];
    }
}

struct TreeView {
    const uint32_t *left, *right, *first, *count;
    const float4 *lo, *hi;
    uint32_t n;
    __device__ __forceinline__ bool is_leaf(uint32_t node) const { return n == 1u || node >= n - 1u; }
    __device__ __forceinline__ uint32_t cnt(uint32_t node) const { return is_leaf(node) ? 1u : count[node]; }
    __device__ __forceinline__ uint32_t fst(uint32_t node) const { return is_leaf(node) ? (n == 1u ? 0u : node - (n - 1u)) : first[node]; }
    __device__ __forceinline__ float half_area(uint32_t node) const {
        const float4 l = lo[node], h = hi[node];
        const float ex = h.x - l.x, ey = h.y - l.y, ez = h.z - l.z;
        return (ex * ey + ey * ez) + ez * ex;
    }
};

// 6a. one thread per wide node of the level: choose up to 8 children, assign octant slots, count
__global__ void collapse_kernel(TreeView bt, const uint32_t* level, uint32_t n_level, uint32_t* child_slots,
                                uint32_t* n_int, uint32_t* n_tri) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_level) return;
    const uint32_t b = level[w];
    uint32_t ch[8];
    int m = 1;
    ch[0] = b;
    for (int phase = 0; phase < 2; phase++) {
        while (m < 8) {
            int best = -1;
            float best_a = -1.0f;
            for (int c = 0; c < m; c++) {
                if (bt.is_leaf(ch[c])) continue;
                const uint32_t cn = bt.cnt(ch[c]);
                if (phase == 0 ? (cn <= 3u) : (cn < 2u)) continue;
                const float a = bt.half_area(ch[c]);
                if (best < 0 || a > best_a) { best = c; best_a = a; }
            }
            if (best < 0) break;
            const uint32_t nd = ch[best];
            for (int c = m; c > best + 1; c--) ch[c] = ch[c - 1];
            ch[best] = bt.left[nd];
            ch[best + 1] = bt.right[nd];
            m++;
        }
    }
    const float4 nlo = bt.lo[b], nhi = bt.hi[b];
    const vec3 centre = scale3(add3(mk3(nlo.x, nlo.y, nlo.z), mk3(nhi.x, nhi.y, nhi.z)), 0.5f);
    vec3 dc[8];
    for (int c = 0; c < m; c++) {
        const float4 cl = bt.lo[ch[c]], chh = bt.hi[ch[c]];
        dc[c] = sub3(scale3(add3(mk3(cl.x, cl.y, cl.z), mk3(chh.x, chh.y, chh.z)), 0.5f), centre);
    }
    // greedy assignment: repeatedly take the (child, free slot) pair with the largest signed projection
    uint32_t slot_used = 0, child_done = 0;
    int child_in_slot[8];
    for (int sl = 0; sl < 8; sl++) child_in_slot[sl] = -1;
    for (int it = 0; it < m; it++) {
        int bc = -1, bs = -1;
        float bv = 0.0f;
        for (int c = 0; c < m; c++) {
            if (child_done & (1u << c)) continue;
            for (int sl = 0; sl < 8; sl++) {
                if (slot_used & (1u << sl)) continue;
                const float sx = (sl & 1) ? -dc[c].x : dc[c].x, sy = (sl & 2) ? -dc[c].y : dc[c].y,
                            sz = (sl & 4) ? -dc[c].z : dc[c].z;
                const float cost = (sx + sy) + sz;
                if (bc < 0 || cost > bv) { bc = c; bs = sl; bv = cost; }
            }
        }
        child_in_slot[bs] = bc;
        child_done |= 1u << bc;
        slot_used |= 1u << bs;
    }
    uint32_t ni = 0, nt = 0;
    for (int sl = 0; sl < 8; sl++) {
        const int c = child_in_slot[sl];
        uint32_t v = 0xFFFFFFFFu;
        if (c >= 0) {
            v = ch[c];
            const uint32_t cn = bt.cnt(v);
            if (!bt.is_leaf(v) && cn > 3u) ni++; else nt += cn;
        }
        child_slots[(size_t)w * 8 + sl] = v;
    }
    n_int[w] = ni;
    n_tri[w] = nt;
}

// binary16 bits of an integer 0..2048 (exact)
__device__ __forceinline__ uint16_t half_of_int(int q) { return half_bits_rne((float)q); }

__device__ __forceinline__ void quantize_axis(float lo, float hi, int* e_out, float* p_out) {
    const float ext = hi - lo;
    const float f = ext / RT_QSTEPS;
    const uint32_t bits = __float_as_uint(f);
    int e = (int)((bits >> 23) & 0xFFu) - 127 + ((bits & 0x7FFFFFu) ? 1 : 0);
    const float mag = glsl_max(fabsf(lo), fabsf(hi));
    const int mb = (int)((__float_as_uint(mag) >> 23) & 0xFFu) - 127;
    if (e < mb - 20) e = mb - 20;
    if (e < -100) e = -100;
    if (e > 100) e = 100;
    const float scale = __uint_as_float((uint32_t)(e + 127) << 23);
    *e_out = e;
    *p_out = lo - scale;
}

// A node in memory: its four 32-byte chunks rotated by the node index (rt_types.cuh: node_chunk).
__device__ __forceinline__ void store_node(Node8* nodes, uint32_t idx, const Node8& nd) {
    const uint4* src = reinterpret_cast<const uint4*>(&nd);
    uint4* dst = reinterpret_cast<uint4*>(nodes + idx);
#pragma unroll
    for (uint32_t c = 0; c < 4u; c++) {
        const uint32_t at = node_chunk(idx, c);
        dst[at] = src[2u * c];
        dst[at + 1u] = src[2u * c + 1u];
    }
}
__device__ __forceinline__ void fetch_node(const Node8* nodes, uint32_t idx, Node8& nd) {
    const uint4* src = reinterpret_cast<const uint4*>(nodes + idx);
    uint4* dst = reinterpret_cast<uint4*>(&nd);
#pragma unroll
    for (uint32_t c = 0; c < 4u; c++) {
        const uint32_t at = node_chunk(idx, c);
        dst[2u * c] = src[at];
        dst[2u * c + 1u] = src[at + 1u];
    }
}

// 6b. emit the level's nodes, its leaf triangles and the next level's node list.
// n_int / n_tri hold exclusive prefix sums over the level.
__global__ void emit_kernel(TreeView bt, const uint32_t* level, uint32_t n_level, const uint32_t* child_slots,
                            const uint32_t* int_prefix, const uint32_t* tri_prefix, uint32_t level_first,
                            uint32_t nodes_total, uint32_t tris_total, const uint32_t* sorted_ids, const float4* flat,
                            Node8* nodes, float4* btris, uint32_t* next_level, BuildScratch* scratch, float4* node_lo,
                            float4* node_hi) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_level) return;
    const uint32_t b = level[w];
    const float4 nlo = bt.lo[b], nhi = bt.hi[b];
    node_lo[level_first + w] = nlo;
    node_hi[level_first + w] = nhi;
    atomicAdd(&scratch->sa_sum, (double)bt.half_area(b));
    Node8 nd;
    int e[3];
    quantize_axis(nlo.x, nhi.x, &e[0], &nd.p[0]);
    quantize_axis(nlo.y, nhi.y, &e[1], &nd.p[1]);
    quantize_axis(nlo.z, nhi.z, &e[2], &nd.p[2]);
    float inv_scale[3];
    for (int a = 0; a < 3; a++) {
        nd.e[a] = (uint8_t)(e[a] + 127);
        inv_scale[a] = __uint_as_float((uint32_t)(127 - e[a]) << 23);
    }
    nd.imask = 0;
    nd.child_base = nodes_total + int_prefix[w];
    nd.tri_base = tris_total + tri_prefix[w];
    uint32_t n_int = 0, tri_off = 0;
    for (int sl = 0; sl < 8; sl++) {
        const uint32_t c = child_slots[(size_t)w * 8 + sl];
        if (c == 0xFFFFFFFFu) {
            nd.meta[sl] = 0;
            for (int a = 0; a < 3; a++) { nd.q[a][0][sl] = half_of_int(RT_QMAX); nd.q[a][1][sl] = half_of_int(0); }
            continue;
        }
        const float4 cl = bt.lo[c], chh = bt.hi[c];
        const float clo[3] = {cl.x, cl.y, cl.z}, chi[3] = {chh.x, chh.y, chh.z};
        for (int a = 0; a < 3; a++) {
            const int ql = (int)floorf((clo[a] - nd.p[a]) * inv_scale[a]) - 1;
            const int qh = (int)ceilf((chi[a] - nd.p[a]) * inv_scale[a]) + 1;
            nd.q[a][0][sl] = half_of_int(min(max(ql, 0), RT_QMAX));
            nd.q[a][1][sl] = half_of_int(min(max(qh, 0), RT_QMAX));
        }
        const uint32_t cn = bt.cnt(c);
        if (!bt.is_leaf(c) && cn > 3u) {
            nd.imask |= (uint8_t)(1u << sl);
            nd.meta[sl] = (uint8_t)(0x20u | (24u + (uint32_t)sl));
            next_level[int_prefix[w] + n_int] = c;
            n_int++;
        } else {
            const uint32_t unary = (cn == 1u) ? 1u : (cn == 2u ? 3u : 7u);
            nd.meta[sl] = (uint8_t)((unary << 5) | tri_off);
            const uint32_t f0 = bt.fst(c);
            for (uint32_t q = 0; q < cn; q++) {
                const uint32_t id = sorted_ids[f0 + q];
                const size_t dst = 3 * (size_t)(nd.tri_base + tri_off + q);
                btris[dst + 0] = flat[3 * (size_t)id + 0];
                btris[dst + 1] = flat[3 * (size_t)id + 1];
                btris[dst + 2] = flat[3 * (size_t)id + 2];
            }
            tri_off += cn;
        }
    }
    store_node(nodes, level_first + w, nd);
    if (w == n_level - 1) {  // level totals for the host loop
        scratch->level_int = int_prefix[w] + n_int;
        scratch->level_tris = tri_prefix[w] + tri_off;
    }
}

__global__ void empty_root_kernel(Node8* nodes) {
    Node8 nd;
    memset(&nd, 0, sizeof(nd));
    for (int a = 0; a < 3; a++) {
        nd.e[a] = 127;
        for (int c = 0; c < 8; c++) { nd.q[a][0][c] = half_of_int(RT_QMAX); nd.q[a][1][c] = half_of_int(0); }
    }
    nodes[0] = nd;
}

// ------------------------------------------------------------------------------------------------
// host orchestration
// ------------------------------------------------------------------------------------------------
#define BUILD_TRY(expr)                                                              \
    do {                                                                             \
        cudaError_t _e = (expr);                                                     \
        if (_e != cudaSuccess) {                                                     \
            err = std::string(#expr) + ": " + cudaGetErrorString(_e);                \
            return (_e == cudaErrorMemoryAllocation) ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA; \
        }                                                                            \
    } while (0)

int build_bvh8(const BuildInput& in, BuildOutput& out, cudaStream_t stream, std::string& err) {
    const uint32_t n = in.n_tris;
    cudaEvent_t ev0, ev1;
    BUILD_TRY(cudaEventCreate(&ev0));
    BUILD_TRY(cudaEventCreate(&ev1));
    BUILD_TRY(cudaEventRecord(ev0, stream));
    const uint32_t max_nodes = n / 3 + 2;  // every wide node below the root covers > 3 triangles: W(n) <= (n-1)/3
    BUILD_TRY(out.nodes.ensure((size_t)max_nodes * RT_NODE_WORDS));
    BUILD_TRY(out.tris.ensure((size_t)n * 3));
    BUILD_TRY(out.morton.ensure(n));
    BUILD_TRY(out.sorted_ids.ensure(n));
    BUILD_TRY(out.node_lo.ensure(max_nodes));
    BUILD_TRY(out.node_hi.ensure(max_nodes));
    out.n_nodes = 1;
    out.depth = 1;
    out.level_first.assign(1, 0u);
    out.sa_build = 0.0;
    out.refits = 0;
    if (n == 0) {
        empty_root_kernel<<<1, 1, 0, stream>>>(reinterpret_cast<Node8*>(out.nodes.p));
        BUILD_TRY(cudaGetLastError());
        BUILD_TRY(cudaStreamSynchronize(stream));
        cudaEventDestroy(ev0); cudaEventDestroy(ev1);
        return RT_OK;
    }
    PoolBuf<BuildScratch> scratch;
    PoolBuf<float4> flat, tlo, thi, nlo, nhi;
    PoolBuf<unsigned long long> keys_b;
    PoolBuf<uint32_t> ids_b, hist, left, right, first, count, parent, arrive;
    PoolBuf<uint32_t> level_a, level_b, child_slots, n_int, n_tri;
    Scanner scanner;
    const uint32_t n_tiles = (n + SORT_TILE - 1) / SORT_TILE;
    BUILD_TRY(scratch.alloc(1, stream));
    BUILD_TRY(flat.alloc((size_t)n * 3, stream));
    BUILD_TRY(tlo.alloc(n, stream));
    BUILD_TRY(thi.alloc(n, stream));
    BUILD_TRY(keys_b.alloc(n, stream));
    BUILD_TRY(ids_b.alloc(n, stream));
    BUILD_TRY(hist.alloc((size_t)n_tiles * 256, stream));
    BUILD_TRY(scanner.reserve(std::max<size_t>((size_t)n_tiles * 256, n), stream));
    const uint32_t n_bin = (n == 1) ? 1 : 2 * n - 1;
    BUILD_TRY(nlo.alloc(n_bin, stream));
    BUILD_TRY(nhi.alloc(n_bin, stream));
    BUILD_TRY(left.alloc(n, stream));
    BUILD_TRY(right.alloc(n, stream));
    BUILD_TRY(first.alloc(n, stream));
    BUILD_TRY(count.alloc(n, stream));
    BUILD_TRY(parent.alloc(n_bin, stream));
    BUILD_TRY(arrive.alloc(n, stream));

    const uint32_t T = 256;
    init_scratch_kernel<<<1, 1, 0, stream>>>(scratch.p);
    flatten_kernel<<<(n + T - 1) / T, T, 0, stream>>>(in, flat.p, tlo.p, thi.p, scratch.p);
    morton_kernel<<<(n + T - 1) / T, T, 0, stream>>>(tlo.p, thi.p, n, scratch.p, out.morton.p, out.sorted_ids.p);
    BUILD_TRY(cudaGetLastError());
    // 8 passes: the result of an even number of ping-pongs lands back in out.*
    {
        unsigned long long* kin = out.morton.p; unsigned long long* kout = keys_b.p;
        uint32_t* vin = out.sorted_ids.p; uint32_t* vout = ids_b.p;
        for (int pass = 0; pass < 8; pass++) {
            const int shift = pass * 8;
            sort_hist_kernel<<<n_tiles, SORT_THREADS, 0, stream>>>(kin, n, shift, hist.p, n_tiles);
            scanner.run(hist.p, n_tiles * 256, stream);
            sort_scatter_kernel<<<n_tiles, SORT_THREADS, 0, stream>>>(kin, vin, kout, vout, n, shift, hist.p, n_tiles);
            std::swap(kin, kout);
            std::swap(vin, vout);
        }
        BUILD_TRY(cudaGetLastError());
    }
    if (n > 1) {
        karras_kernel<<<(n - 1 + T - 1) / T, T, 0, stream>>>(out.morton.p, (int)n, left.p, right.p, first.p, count.p, parent.p);
        BUILD_TRY(cudaMemsetAsync(arrive.p, 0, (size_t)n * 4, stream));
    }
    boxes_kernel<<<(n + T - 1) / T, T, 0, stream>>>(tlo.p, thi.p, out.sorted_ids.p, n, left.p, right.p, parent.p, arrive.p,
                                                    nlo.p, nhi.p);
    BUILD_TRY(cudaGetLastError());

    TreeView bt;
    bt.left = left.p; bt.right = right.p; bt.first = first.p; bt.count = count.p;
    bt.lo = nlo.p; bt.hi = nhi.p; bt.n = n;
    BUILD_TRY(level_a.alloc(max_nodes, stream));
    BUILD_TRY(level_b.alloc(max_nodes, stream));
    BUILD_TRY(child_slots.alloc((size_t)max_nodes * 8, stream));
    BUILD_TRY(n_int.alloc(max_nodes, stream));
    BUILD_TRY(n_tri.alloc(max_nodes, stream));
    const uint32_t zero = 0;
    BUILD_TRY(cudaMemcpyAsync(level_a.p, &zero, 4, cudaMemcpyHostToDevice, stream));
    uint32_t n_level = 1, level_first = 0, nodes_total = 1, tris_total = 0, depth = 0;
    uint32_t* cur = level_a.p;
    uint32_t* nxt = level_b.p;
    while (n_level > 0) {
        depth++;
        const uint32_t g = (n_level + 127) / 128;
        collapse_kernel<<<g, 128, 0, stream>>>(bt, cur, n_level, child_slots.p, n_int.p, n_tri.p);
        scanner.run(n_int.p, n_level, stream);
        scanner.run(n_tri.p, n_level, stream);
        emit_kernel<<<g, 128, 0, stream>>>(bt, cur, n_level, child_slots.p, n_int.p, n_tri.p, level_first, nodes_total,
                                           tris_total, out.sorted_ids.p, flat.p, reinterpret_cast<Node8*>(out.nodes.p),
                                           out.tris.p, nxt, scratch.p, out.node_lo.p, out.node_hi.p);
        BUILD_TRY(cudaGetLastError());
        BuildScratch hs;
        BUILD_TRY(cudaMemcpyAsync(&hs, scratch.p, sizeof(hs), cudaMemcpyDeviceToHost, stream));
        BUILD_TRY(cudaStreamSynchronize(stream));
        level_first += n_level;
        out.level_first.push_back(level_first);
        out.sa_build = hs.sa_sum;
        n_level = hs.level_int;
        nodes_total += hs.level_int;
        tris_total += hs.level_tris;
        if (nodes_total > max_nodes) { err = "BVH8 node bound exceeded"; return RT_ERR_CUDA; }
        std::swap(cur, nxt);
    }
    if (tris_total != n) { err = "BVH8 build lost triangles"; return RT_ERR_CUDA; }
    out.n_nodes = nodes_total;
    out.depth = depth;
    BUILD_TRY(cudaEventRecord(ev1, stream));
    BUILD_TRY(cudaEventSynchronize(ev1));
    cudaEventElapsedTime(&out.build_ms, ev0, ev1);
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    return RT_OK;
}

// ------------------------------------------------------------------------------------------------
// refit: object transforms changed, the topology stays (the TLAS `mode = .UPDATE` of the reference:
// src/raytracer/acceleration_structure.odin:91, gpu_scene.odin:457-483)
// ------------------------------------------------------------------------------------------------
__global__ void refit_tris_kernel(BuildInput in, const uint8_t* dirty, float4* btris) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= in.n_tris) return;
    const uint32_t inst = __float_as_uint(btris[3 * (size_t)i + 1].w);
    if (!dirty[inst]) return;
    const uint32_t prim = __float_as_uint(btris[3 * (size_t)i + 2].w);
    const DevObject* o = in.d_objects + inst;
    const uint32_t* ip = in.d_indices + o->index_base + 3u * prim;
    const rt_vertex* vb = in.d_vertices + o->vertex_base;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const vec3 p = instance_vertex(o, vb + ip[k]);
        float4 q = btris[3 * (size_t)i + k];
        q.x = p.x; q.y = p.y; q.z = p.z;
        btris[3 * (size_t)i + k] = q;
    }
}

// one thread per wide node of a level (deepest level first): exact child boxes from the triangles (leaves) or from the
// children's own boxes (already refitted), the node's box as their union, then the same quantisation as emit_kernel
__global__ void refit_level_kernel(Node8* nodes, const float4* btris, float4* node_lo, float4* node_hi, uint32_t first, uint32_t n_level,
                                   double* sa_sum) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_level) return;
    Node8 nd;
    fetch_node(nodes, first + w, nd);
    float clo[8][3], chi[8][3];
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    uint32_t n_int = 0;
    for (int sl = 0; sl < 8; sl++) {
        const uint32_t meta = nd.meta[sl];
        if (meta == 0u) continue;
        float l[3] = {INFINITY, INFINITY, INFINITY}, h[3] = {-INFINITY, -INFINITY, -INFINITY};
        if ((meta & 0x38u) == 0x38u && (meta >> 5) == 1u) {   // internal child: 0x20 | (24 + slot)
            const uint32_t c = nd.child_base + n_int++;
            const float4 a = node_lo[c], b = node_hi[c];
            l[0] = a.x; l[1] = a.y; l[2] = a.z; h[0] = b.x; h[1] = b.y; h[2] = b.z;
        } else {                                               // leaf: unary count << 5 | offset from tri_base
            const uint32_t un = meta >> 5, cnt = un == 1u ? 1u : (un == 3u ? 2u : 3u), off = meta & 31u;
            for (uint32_t q = 0; q < cnt; q++)
                for (int k = 0; k < 3; k++) {
                    const float4 p = btris[3 * (size_t)(nd.tri_base + off + q) + k];
                    l[0] = fminf(l[0], p.x); l[1] = fminf(l[1], p.y); l[2] = fminf(l[2], p.z);
                    h[0] = fmaxf(h[0], p.x); h[1] = fmaxf(h[1], p.y); h[2] = fmaxf(h[2], p.z);
                }
        }
        for (int a = 0; a < 3; a++) { clo[sl][a] = l[a]; chi[sl][a] = h[a]; lo[a] = fminf(lo[a], l[a]); hi[a] = fmaxf(hi[a], h[a]); }
    }
    if (lo[0] == INFINITY) return;  // a node without children (the root of an empty scene)
    node_lo[first + w] = make_float4(lo[0], lo[1], lo[2], 0.0f);
    node_hi[first + w] = make_float4(hi[0], hi[1], hi[2], 0.0f);
    {
        const float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        atomicAdd(sa_sum, (double)((ex * ey + ey * ez) + ez * ex));
    }
    int e[3];
    float inv_scale[3];
    for (int a = 0; a < 3; a++) {
        quantize_axis(lo[a], hi[a], &e[a], &nd.p[a]);
        nd.e[a] = (uint8_t)(e[a] + 127);
        inv_scale[a] = __uint_as_float((uint32_t)(127 - e[a]) << 23);
    }
    for (int sl = 0; sl < 8; sl++) {
        if (nd.meta[sl] == 0u) continue;  // empty slots keep their inverted box
        for (int a = 0; a < 3; a++) {
            const int ql = (int)floorf((clo[sl][a] - nd.p[a]) * inv_scale[a]) - 1;
            const int qh = (int)ceilf((chi[sl][a] - nd.p[a]) * inv_scale[a]) + 1;
            nd.q[a][0][sl] = half_of_int(min(max(ql, 0), RT_QMAX));
            nd.q[a][1][sl] = half_of_int(min(max(qh, 0), RT_QMAX));
        }
    }
    store_node(nodes, first + w, nd);
}

int refit_bvh8(const BuildInput& in, BuildOutput& out, const uint8_t* d_dirty, double* sa_now, cudaStream_t stream, std::string& err) {
    const uint32_t n = in.n_tris;
    if (n == 0 || out.level_first.size() < 2) { *sa_now = out.sa_build; return RT_OK; }
    cudaEvent_t ev0, ev1;
    BUILD_TRY(cudaEventCreate(&ev0));
    BUILD_TRY(cudaEventCreate(&ev1));
    BUILD_TRY(cudaEventRecord(ev0, stream));
    PoolBuf<double> sa;
    BUILD_TRY(sa.alloc(1, stream));
    BUILD_TRY(cudaMemsetAsync(sa.p, 0, sizeof(double), stream));
    refit_tris_kernel<<<(n + 255) / 256, 256, 0, stream>>>(in, d_dirty, out.tris.p);
    for (size_t d = out.level_first.size() - 1; d-- > 0;) {
        const uint32_t first = out.level_first[d], n_level = out.level_first[d + 1] - first;
        if (n_level == 0) continue;
        refit_level_kernel<<<(n_level + 127) / 128, 128, 0, stream>>>(reinterpret_cast<Node8*>(out.nodes.p), out.tris.p, out.node_lo.p,
                                                                      out.node_hi.p, first, n_level, sa.p);
    }
    BUILD_TRY(cudaGetLastError());
    BUILD_TRY(cudaMemcpyAsync(sa_now, sa.p, sizeof(double), cudaMemcpyDeviceToHost, stream));
    BUILD_TRY(cudaEventRecord(ev1, stream));
    BUILD_TRY(cudaEventSynchronize(ev1));
    cudaEventElapsedTime(&out.refit_ms, ev0, ev1);
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    out.refits++;
    return RT_OK;
}

}  // namespace rt
