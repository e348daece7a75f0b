// rt_frame.cu — the frame: CUDA replacement of vkCmdTraceRaysKHR(W, H, 1) as recorded by
// raytracing_renderer_render_scene (src/raytracer/raytracing_renderer.odin:204-270) and of the
// shaders it drives (shaders/simple.rgen / simple.rchit / simple.rmiss, inline isVisibleRQ).
//
// Two pipelines produce bit-identical images:
//   wavefront  raygen -> [extend -> shade -> shadow] x max_depth -> accumulate.  Path state lives in
//              SoA arrays indexed by pixel slot; queues hold slots, filled with warp-aggregated
//              atomics (ballot compaction); traversal kernels are persistent and fetch work per warp.
//   megakernel one path per thread, shadow rays inline (kept as the cross-check and for tiny frames).
#include "rt_internal.h"
#include "rt_shade.cuh"
#include "rt_trav.cuh"

namespace rt {

#define FRAME_TRY(expr)                                                              \
    do {                                                                             \
        cudaError_t _e = (expr);                                                     \
        if (_e != cudaSuccess) {                                                     \
            err = std::string(#expr) + ": " + cudaGetErrorString(_e);                \
            return (_e == cudaErrorMemoryAllocation) ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA; \
        }                                                                            \
    } while (0)

// SM count of the CURRENT device (the context's: every entry point sets it); cached per context in LaunchDims
static int query_num_sms() {
    int dev = 0, n = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n > 0 ? n : 148;
}

// ---- counters: one atomic per warp -------------------------------------------------------------
__device__ __forceinline__ void warp_add(unsigned long long* dst, uint32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    if ((threadIdx.x & 31u) == 0u && v) atomicAdd(dst, (unsigned long long)v);
}

// warp-aggregated queue append: one atomicAdd per warp, slots keep lane order
__device__ __forceinline__ uint32_t queue_push(uint32_t* counter, bool pred) {
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, pred);
    if (ballot == 0u) return 0u;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t leader = __ffs(ballot) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(ballot));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    return base + __popc(ballot & ((1u << lane) - 1u));
}

// ================================================================================================
// verification hook: arbitrary rays through the same traversal code (rt_trace_rays_closest / _any)
// ================================================================================================
template <bool ANY>
__global__ void __launch_bounds__(128) trace_rays_kernel(SceneView sc, const float4* rays, uint32_t n,
                                                         const uint32_t* ignore, float* tuv, uint32_t* ids,
                                                         DevCounters* counters) {
    uint32_t nn = 0, nt = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = rays[2 * (size_t)i], b = rays[2 * (size_t)i + 1];
        const Hit h = traverse<ANY, true>(sc, mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), a.w, b.w,
                                          ANY ? (ignore ? ignore[i] : 0xFFFFFFFFu) : 0xFFFFFFFFu, nn, nt);
        if (ANY) {
            ids[i] = (h.id != 0xFFFFFFFFu) ? 1u : 0u;
        } else {
            ids[i] = h.id;
            const bool ok = h.id != 0xFFFFFFFFu;
            tuv[3 * (size_t)i + 0] = ok ? h.t : 0.0f;
            tuv[3 * (size_t)i + 1] = ok ? h.u : 0.0f;
            tuv[3 * (size_t)i + 2] = ok ? h.v : 0.0f;
        }
    }
    if (counters) {
        atomicAdd(&counters->nodes, (unsigned long long)nn);
        atomicAdd(&counters->tris, (unsigned long long)nt);
    }
}

int launch_trace_rays(const SceneView& sc, const float* d_rays, uint32_t n, bool any, const uint32_t* d_ignore,
                      float* d_tuv, uint32_t* d_id, DevCounters* counters, int num_sms, cudaStream_t stream, std::string& err) {
    if (n == 0) return RT_OK;
    const uint32_t grid = std::min<uint32_t>((n + 127) / 128, (uint32_t)num_sms * 16);
    if (any)
        trace_rays_kernel<true><<<grid, 128, 0, stream>>>(sc, reinterpret_cast<const float4*>(d_rays), n, d_ignore, d_tuv,
                                                          d_id, counters);
    else
        trace_rays_kernel<false><<<grid, 128, 0, stream>>>(sc, reinterpret_cast<const float4*>(d_rays), n, d_ignore, d_tuv,
                                                           d_id, counters);
    FRAME_TRY(cudaGetLastError());
    return RT_OK;
}

// ================================================================================================
// megakernel: simple.rgen:70-137 as one thread per pixel
// ================================================================================================
template <bool COUNT>
__global__ void __launch_bounds__(128) megakernel_frame(SceneView sc, FrameParams fp, float4* image, DevCounters* counters) {
    const uint32_t n_slots = fp.local_rows * fp.width;
    uint32_t c_ext = 0, c_sh = 0, c_nodes = 0, c_tris = 0, c_hits = 0, c_nee = 0, c_snodes = 0, c_stris = 0;
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n_slots; slot += gridDim.x * blockDim.x) {
        const uint32_t lrow = slot / fp.width, x = slot - lrow * fp.width;
        const uint32_t y = strip_global_row(fp, lrow);
        PathState ps;
        vec3 o, d;
        raygen_init(fp, x, y, ps, &o, &d);
        const vec3 clearc = mk3(fp.clear[0], fp.clear[1], fp.clear[2]);
        for (int depth = 0; depth < fp.max_depth; depth++) {
            if (!russian_roulette(ps, depth, fp.rr_start)) break;
            c_ext++;
            const Hit h = traverse<false, COUNT>(sc, o, d, 0.001f, 10000.0f, 0xFFFFFFFFu, c_nodes, c_tris);
            bool alive = false;
            if (h.id != 0xFFFFFFFFu) {
                c_hits++;
                ShadowRequest sr;
                const ShadeResult res = shade_hit(sc, ps, h, d, sr, c_nee);
                if (sr.pending) {
                    c_sh++;
                    const Hit sh = traverse<true, COUNT>(sc, sr.origin, sr.dir, 0.001f, sr.tmax, sr.light_object, c_snodes, c_stris);
                    if (sh.id == 0xFFFFFFFFu) ps.color = add3(ps.color, sr.contrib);
                }
                if (res.emissive_add) ps.color = add3(ps.color, res.emissive);
                alive = res.alive;
                o = res.next_origin;
                d = res.next_dir;
            }
            if (!alive) {
                ps.color = add3(ps.color, mul3(ps.throughput, clearc));
                break;
            }
            if (throughput_dead(ps.throughput)) break;
        }
        image[slot] = accumulate_pixel(fp.sample ? image[slot] : make_float4(0.f, 0.f, 0.f, 0.f), ps.color, fp.clamp, fp.sample);
    }
    if (COUNT) {
        warp_add(&counters->extend_rays, c_ext); warp_add(&counters->shadow_rays, c_sh);
        warp_add(&counters->nodes, c_nodes + c_snodes); warp_add(&counters->tris, c_tris + c_stris);
        warp_add(&counters->shadow_nodes, c_snodes); warp_add(&counters->shadow_tris, c_stris);
        warp_add(&counters->hits, c_hits); warp_add(&counters->nee, c_nee);
    }
}

int launch_frame_megakernel(const SceneView& sc, const FrameParams& fp, FrameResources& res, cudaStream_t stream,
                            std::string& err) {
    const uint32_t n_slots = fp.local_rows * fp.width;
    if (n_slots == 0) return RT_OK;
    if (res.dims && !res.dims->num_sms) res.dims->num_sms = query_num_sms();
    const uint32_t grid = std::min<uint32_t>((n_slots + 127) / 128, (uint32_t)(res.dims ? res.dims->num_sms : query_num_sms()) * 64);
    if (fp.flags & RT_FLAG_COUNTERS)
        megakernel_frame<true><<<grid, 128, 0, stream>>>(sc, fp, res.image, res.counters);
    else
        megakernel_frame<false><<<grid, 128, 0, stream>>>(sc, fp, res.image, res.counters);
    if (res.launches) (*res.launches)++;
    FRAME_TRY(cudaGetLastError());
    return RT_OK;
}

// Strip-cost calibration (RT_BALANCE_COST): the paths of the full frame once more, nothing stored but what each image row
// costs: row_cost[4 y + 0] = nodes + triangles a per-ray walk of its PRIMARY rays fetches (the proxy for the packet kernel's
// work), [4 y + 1] = nodes + triangles of its continuation rays (wf_extend_coop_kernel), [4 y + 2] = closest-hit shader
// invocations, [4 y + 3] = continuation rays.  rt_api.cu (balance_strips) weighs them with the measured cost of each
// (tools/fit_balance.py).  Light samples are drawn (the random stream must stay the path's) but not evaluated;
// visibility rays are not traced.
__global__ void __launch_bounds__(128) row_cost_kernel(SceneView sc, FrameParams fp, uint32_t* row_cost) {
    const uint32_t n_px = fp.height * fp.width;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n_px; base += gridDim.x * blockDim.x) {
        const uint32_t px = base + threadIdx.x;
        uint32_t c[4] = {0u, 0u, 0u, 0u};
        uint32_t y = 0;
        if (px < n_px) {
            y = px / fp.width;
            const uint32_t x = px - y * fp.width;
            PathState ps;
            vec3 o, d;
            raygen_init(fp, x, y, ps, &o, &d);
            uint32_t nee = 0;
            for (int depth = 0; depth < fp.max_depth; depth++) {
                if (!russian_roulette(ps, depth, fp.rr_start)) break;
                uint32_t nn = 0, nt = 0;
                const Hit h = traverse<false, true>(sc, o, d, 0.001f, 10000.0f, 0xFFFFFFFFu, nn, nt);
                c[depth == 0 ? 0 : 1] += nn + nt;
                if (depth > 0) c[3]++;
                if (h.id == 0xFFFFFFFFu) break;
                c[2]++;
                ShadowRequest sr;
                NeeRequest nq;
                const ShadeResult res = shade_hit<true, true>(sc, ps, h, d, sr, nee, &nq);
                if (!res.alive || throughput_dead(ps.throughput)) break;
                o = res.next_origin;
                d = res.next_dir;
            }
        }
        // a warp's pixels lie in at most two rows when the image is at least a warp wide: one atomic per row, warp and counter
        const uint32_t y0 = __shfl_sync(0xFFFFFFFFu, y, 0);
        const bool in0 = px < n_px && y == y0, in1 = px < n_px && y != y0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            uint32_t a = in0 ? c[k] : 0u, b = in1 ? c[k] : 0u;
#pragma unroll
            for (int o2 = 16; o2 > 0; o2 >>= 1) { a += __shfl_xor_sync(0xFFFFFFFFu, a, o2); b += __shfl_xor_sync(0xFFFFFFFFu, b, o2); }
            if (fp.width >= 32u) {
                if ((threadIdx.x & 31u) == 0u) {
                    if (a) atomicAdd(&row_cost[4u * y0 + k], a);
                    if (b && y0 + 1u < fp.height) atomicAdd(&row_cost[4u * (y0 + 1u) + k], b);
                }
            } else if (px < n_px && c[k]) {
                atomicAdd(&row_cost[4u * y + k], c[k]);
            }
        }
    }
}

int launch_row_cost(const SceneView& sc, const FrameParams& fp, uint32_t* d_row_cost, int num_sms, cudaStream_t stream, std::string& err) {
    const uint32_t n_px = fp.height * fp.width;
    if (n_px == 0) return RT_OK;
    const uint32_t grid = std::min<uint32_t>((n_px + 127) / 128, (uint32_t)num_sms * 64);
    row_cost_kernel<<<grid, 128, 0, stream>>>(sc, fp, d_row_cost);
    FRAME_TRY(cudaGetLastError());
    return RT_OK;
}

// ================================================================================================
// wavefront pipeline
// ================================================================================================
// Path state, SoA over pixel slots.  flag bits in ray_o.w (scenes with lights):
enum { PF_FIRST = 1u, PF_SPECULAR = 2u, PF_DID_DIRECT = 4u };

struct WavefrontBuffers {
    uint32_t n_slots = 0;
    DevBuf<float4> ray_o;   // origin.xyz = previousHitPos of the bounce that reads it, flags
    DevBuf<float4> ray_d;   // direction.xyz, -
    DevBuf<float4> thr;     // throughput.xyz, previousBrdfPdf
    DevBuf<float4> col;     // color.xyz, p_sample_light
    DevBuf<uint2> seeds;    // payload.seed, ray-gen seed
    DevBuf<uint4> hit;      // u bits, v bits, inst (0xFFFFFFFF = miss), prim
    DevBuf<uint32_t> queue[2];
    DevBuf<float4> sh_o;    // shadow origin.xyz, tmax
    DevBuf<float4> sh_d;    // shadow dir.xyz, light object bits
    DevBuf<float4> sh_c;    // contribution.xyz, slot bits
    DevBuf<float4> sh_e;    // emissive term.xyz, flags (1 = emissive add, 2 = sky add)
    DevBuf<float4> sh_n;    // light sample waiting for wf_nee_kernel: its three draws, selection pdf
    DevBuf<uint32_t> ctrl;  // [0..1] queue counts, [2] shadow count, [3..5] work cursors (extend, shade, shadow)
    // how many paths the last batch on this state still had at bounce RT_TAIL_DEPTH: written by the GPU into mapped host memory,
    // read by the host WITHOUT synchronisation when it queues the next batch (a hint for the choice of kernels, never for results)
    uint32_t* tail_hint_h = nullptr;
    uint32_t* tail_hint_d = nullptr;
    ~WavefrontBuffers() { if (tail_hint_h) cudaFreeHost(tail_hint_h); }
};

int wavefront_alloc(WavefrontBuffers** out, uint32_t n_slots, std::string& err) {
    WavefrontBuffers* w = new WavefrontBuffers();
    w->n_slots = n_slots;
    cudaError_t e = cudaSuccess;
    auto A = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
    A(w->ray_o.alloc(n_slots)); A(w->ray_d.alloc(n_slots)); A(w->thr.alloc(n_slots)); A(w->col.alloc(n_slots));
    A(w->seeds.alloc(n_slots)); A(w->hit.alloc(n_slots));
    A(w->queue[0].alloc(n_slots)); A(w->queue[1].alloc(n_slots));
    A(w->sh_o.alloc(n_slots)); A(w->sh_d.alloc(n_slots)); A(w->sh_c.alloc(n_slots)); A(w->sh_e.alloc(n_slots)); A(w->sh_n.alloc(n_slots));
    A(w->ctrl.alloc(8));
    A(cudaHostAlloc(reinterpret_cast<void**>(&w->tail_hint_h), sizeof(uint32_t), cudaHostAllocMapped));
    if (e == cudaSuccess) {
        *w->tail_hint_h = 0xFFFFFFFFu;  // unknown
        A(cudaHostGetDevicePointer(reinterpret_cast<void**>(&w->tail_hint_d), w->tail_hint_h, 0));
    }
    if (e != cudaSuccess) {
        err = std::string("wavefront_alloc: ") + cudaGetErrorString(e);
        delete w;
        return e == cudaErrorMemoryAllocation ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA;
    }
    *out = w;
    return RT_OK;
}
void wavefront_free(WavefrontBuffers* w) { delete w; }

struct WfView {
    float4 *ray_o, *ray_d, *thr, *col;
    uint2* seeds;
    uint4* hit;
    float4 *sh_o, *sh_d, *sh_c, *sh_e, *sh_n;
    uint32_t* ctrl;
};

// ctrl layout
#define CT_Q0 0
#define CT_Q1 1
#define CT_SH 2
#define CT_CUR_EXT 3
#define CT_CUR_SHD 4
#define CT_CUR_SHW 5
#define CT_FB 6       /* rays the packet kernel handed back to the per-ray kernel */
#define CT_CUR_PKT 7  /* packet cursor */

__global__ void wf_reset_kernel(uint32_t* ctrl, uint32_t q_in, uint32_t q_in_count, uint32_t* report = nullptr) {
    // before a bounce: the out queue, the shadow queue and the cursors are empty
    const uint32_t t = threadIdx.x;
    if (t == 0) {
        ctrl[q_in ^ 1u] = 0; ctrl[CT_SH] = 0; ctrl[CT_CUR_EXT] = 0; ctrl[CT_CUR_SHD] = 0; ctrl[CT_CUR_SHW] = 0;
        ctrl[CT_FB] = 0; ctrl[CT_CUR_PKT] = 0;
        if (q_in_count != 0xFFFFFFFFu) ctrl[q_in] = q_in_count;
        if (report) *report = ctrl[q_in];  // the tail hint (mapped host memory)
    }
}

// DevObject::spec_prob = specularProbability of the object's material, by the function the hit shader itself calls: the first-hit
// filter compares the lobe draw with these bits, so they must be the ones sample_brdf computes (same code, -fmad=false).
__global__ void object_spec_prob_kernel(DevObject* objects, const rt_material* materials, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    objects[i].spec_prob = specular_probability(load_material(materials + objects[i].material_index));
}
int launch_object_spec_prob(DevObject* objects, const rt_material* materials, uint32_t n, cudaStream_t stream) {
    object_spec_prob_kernel<<<(n + 127u) / 128u, 128, 0, stream>>>(objects, materials, n);
    return cudaGetLastError() == cudaSuccess ? RT_OK : RT_ERR_CUDA;
}

// raygen: simple.rgen:70-85.  All slots enter bounce 0 (unless roulette is enabled from depth 0).
// A launch covers fp.n_lanes consecutive frames ("lanes"): slot = lane * pixels + pixel, frame = fp.frame + lane * fp.frame_stride.
__device__ __forceinline__ void raygen_slot(const FrameParams& fp, const WfView& wf, uint32_t x, uint32_t lrow, uint32_t lane_f) {
    const uint32_t slot = (lane_f * fp.local_rows + lrow) * fp.width + x;
    const uint32_t y = strip_global_row(fp, lrow);
    uint32_t seed;
    vec3 o, d;
    raygen_ray(fp, fp.frame + lane_f * fp.frame_stride, x, y, &seed, &o, &d);
    // throughput, colour and previous-hit state start from constants: the first shade launch (FIRST = true)
    // supplies them itself instead of reading them back (saves 48 B written here and 48 B read there per ray)
    wf.ray_o[slot] = make_float4(o.x, o.y, o.z, 0.0f);
    wf.ray_d[slot] = make_float4(d.x, d.y, d.z, __uint_as_float(seed));  // w: the seed again, for the light-less shade variant
    wf.seeds[slot] = make_uint2(seed, seed);
}

__global__ void __launch_bounds__(256) wf_raygen_kernel(FrameParams fp, WfView wf) {
    // grid = (blocks across a row, local rows, lanes): the kernel is issue-bound (16 TEA rounds, IEEE divisions), and four
    // integer divisions by run-time values to take a linear index apart were a sixth of its instructions
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= fp.width) return;
    if (gridDim.y == fp.local_rows && gridDim.z == fp.n_lanes) {
        raygen_slot(fp, wf, x, blockIdx.y, blockIdx.z);
    } else {  // grid dimensions y and z stop at 65535: taller images / more lanes loop (the loop costs the usual case 5 %)
        for (uint32_t lane_f = blockIdx.z; lane_f < fp.n_lanes; lane_f += gridDim.z)
            for (uint32_t lrow = blockIdx.y; lrow < fp.local_rows; lrow += gridDim.y) raygen_slot(fp, wf, x, lrow, lane_f);
    }
}

// ---- persistent ray fetch -------------------------------------------------------------------------
// A warp claims RT_FETCH_CHUNK queue entries at a time (one atomic) and hands them to its lanes as they
// run dry: a lane whose ray has finished gets the next entry as soon as RT_REFILL_IDLE lanes are idle,
// instead of waiting for the slowest ray of its batch (Aila & Laine 2009, "dynamic fetch").
#define RT_FETCH_CHUNK 128u
#ifndef RT_REFILL_IDLE
#define RT_REFILL_IDLE 8u
#endif
#ifndef RT_TRAV_MIN_BLOCKS
#define RT_TRAV_MIN_BLOCKS 7  /* 72 registers: the occupancy sweep in profiles/ favours 7 blocks of 128 */
#endif

struct WarpFetch {
    uint32_t next, end;  // the warp's private range of queue indices (warp-uniform)
    bool drained;        // the global cursor has passed the end of the queue
};

// Gives every lane with want == true a queue index below n, or 0xFFFFFFFF once the queue is exhausted.
__device__ __forceinline__ uint32_t warp_fetch(WarpFetch& wfetch, uint32_t* cursor, uint32_t n, bool want) {
    const uint32_t lane = threadIdx.x & 31u;
    // short queues are spread over all warps, long ones are claimed in larger pieces (fewer atomics)
    const uint32_t chunk = (n > gridDim.x * blockDim.x * 4u) ? RT_FETCH_CHUNK : 32u;
    uint32_t need = __ballot_sync(0xFFFFFFFFu, want);
    uint32_t idx = 0xFFFFFFFFu;
    while (need) {
        if (wfetch.next == wfetch.end) {
            if (wfetch.drained) break;
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(cursor, chunk);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            if (base >= n) { wfetch.drained = true; break; }
            wfetch.next = base;
            wfetch.end = min(base + chunk, n);
        }
        const uint32_t avail = wfetch.end - wfetch.next;
        const uint32_t rank = __popc(need & ((1u << lane) - 1u));
        const bool mine = ((need >> lane) & 1u) && rank < avail && idx == 0xFFFFFFFFu;
        if (mine) idx = wfetch.next + rank;
        const uint32_t served = __ballot_sync(0xFFFFFFFFu, mine);
        wfetch.next += __popc(served);
        need &= ~served;
    }
    return idx;
}

// First bounce: every slot is live and there is no queue, so the i-th ray is chosen here.  Consecutive rays
// are the pixels of a 4-wide x 8-tall screen tile rather than of a 32 x 1 row segment: the rays of a warp
// then share more of their path through the BVH (measured: -4 % extend time).  Rows that do not fill a band
// of 8 keep the linear order.
#define RT_TILE_W 4u
#define RT_TILE_H (32u / RT_TILE_W)
__device__ __forceinline__ uint32_t primary_slot(uint32_t i, uint32_t width, uint32_t n_px) {
    const uint32_t lane_f = i / n_px, r = i - lane_f * n_px;
    const uint32_t band_px = width * RT_TILE_H;
    const uint32_t band = r / band_px, b = r - band * band_px;
    const uint32_t wt = (width / RT_TILE_W) * 32u;  // pixels of the band covered by whole tiles
    if ((band + 1u) * band_px > n_px || b >= wt) return i;
    const uint32_t tile = b >> 5, t = b & 31u;
    const uint32_t x = tile * RT_TILE_W + (t % RT_TILE_W), y = band * RT_TILE_H + (t / RT_TILE_W);
    return lane_f * n_px + y * width + x;
}

// extend: closest hit for every queued path (traceRayEXT of simple.rgen:92-104).
template <bool COUNT>
__global__ void __launch_bounds__(128, RT_TRAV_MIN_BLOCKS) wf_extend_kernel(SceneView sc, WfView wf, const uint32_t* queue, uint32_t q_idx,
                                                        uint32_t width, uint32_t n_px, DevCounters* counters) {
    const uint32_t n = wf.ctrl[q_idx];
    uint32_t c_nodes = 0, c_tris = 0, c_rays = 0;
    WarpFetch wfetch;
    wfetch.next = wfetch.end = 0u;
    wfetch.drained = false;
    TravState st;
    TravStack stk;
    bool has_ray = false;
    uint32_t slot = 0;
    for (;;) {
        const uint32_t i = warp_fetch(wfetch, &wf.ctrl[CT_CUR_EXT], n, !has_ray);
        if (i != 0xFFFFFFFFu) {
            slot = queue ? queue[i] : primary_slot(i, width, n_px);
            const float4 o = wf.ray_o[slot], d = wf.ray_d[slot];
            trav_init(st, mk3(o.x, o.y, o.z), mk3(d.x, d.y, d.z), 0.001f, 10000.0f, 0xFFFFFFFFu);
            has_ray = true;
            c_rays++;
        }
        if (__ballot_sync(0xFFFFFFFFu, has_ray) == 0u) break;
        // while the queue still has entries, return to the fetch once enough lanes are idle
        const uint32_t min_active = wfetch.drained ? 1u : (32u - RT_REFILL_IDLE + 1u);
        while (has_ray) {
            if (trav_step<false, COUNT>(sc, st, stk, c_nodes, c_tris)) {
                const Hit& h = st.best;
                wf.hit[slot] = make_uint4(__float_as_uint(h.u), __float_as_uint(h.v), h.id == 0xFFFFFFFFu ? 0xFFFFFFFFu : h.inst,
                                          h.prim);
                has_ray = false;
            }
            if (__popc(__activemask()) < min_active) break;
        }
    }
    if (COUNT) {
        warp_add(&counters->extend_rays, c_rays);
        warp_add(&counters->nodes, c_nodes);
        warp_add(&counters->tris, c_tris);
    }
}

// extend, cooperative triangle tests (rt_trav.cuh: coop_*): the kernel for incoherent rays (bounces >= 1 and what the
// packet kernel hands back).  The loop is warp-uniform: lanes without a ray skip the node step but take their share
// of the pooled triangle tests.
#ifndef RT_COOP_MIN_BLOCKS
#define RT_COOP_MIN_BLOCKS 8
#endif
template <bool COUNT>
__global__ void __launch_bounds__(128, RT_COOP_MIN_BLOCKS) wf_extend_coop_kernel(SceneView sc, WfView wf, const uint32_t* queue,
                                                                                 uint32_t q_idx, uint32_t width, uint32_t n_px,
                                                                                 DevCounters* counters) {
    __shared__ CoopWarp s_cw[4];
    const uint32_t n = wf.ctrl[q_idx];
    const uint32_t lane = threadIdx.x & 31u;
    CoopWarp& cw = s_cw[threadIdx.x >> 5];
    uint32_t c_nodes = 0, c_tris = 0, c_rays = 0;
    WarpFetch wfetch;
    wfetch.next = wfetch.end = 0u;
    wfetch.drained = false;
    NodeRay nr;
    nr.ox = nr.oy = nr.oz = nr.idx = nr.idy = nr.idz = 0.0f; nr.oct = 0u;
    TravStack stk;
    uint2 ng = make_uint2(0u, 0u), top = make_uint2(0u, 0u);
    int sp = 0;
    float best_t = 0.0f;
    bool has_ray = false;
    uint32_t slot = 0;
    for (;;) {
        const uint32_t i = warp_fetch(wfetch, &wf.ctrl[CT_CUR_EXT], n, !has_ray);
        if (i != 0xFFFFFFFFu) {
            slot = queue ? queue[i] : primary_slot(i, width, n_px);
            const float4 o = wf.ray_o[slot], d = wf.ray_d[slot];
            const TravRay r = trav_setup(mk3(o.x, o.y, o.z), mk3(d.x, d.y, d.z));
            coop_init_ray(cw, lane, r, 10000.0f);
            nr.ox = r.ox; nr.oy = r.oy; nr.oz = r.oz; nr.idx = r.idx; nr.idy = r.idy; nr.idz = r.idz; nr.oct = r.rbits & 7u;
            ng = make_uint2(0u, 0x80000000u);
            sp = 0;
            top.y = 0u;
            best_t = 10000.0f;
            has_ray = true;
            c_rays++;
        }
        __syncwarp();
        if (__ballot_sync(0xFFFFFFFFu, has_ray) == 0u) break;
        // while the queue still has entries, return to the fetch once enough lanes are idle
        const uint32_t max_idle = wfetch.drained ? 31u : (RT_REFILL_IDLE - 1u);
        for (;;) {
            uint32_t tmask = 0u, tbase = 0u;
            if (has_ray) tmask = coop_node<COUNT>(sc, nr, ng, sp, stk, top, 0.001f, best_t, &tbase, c_nodes);
            coop_triangles<false, COUNT>(sc, cw, lane, tmask, tbase, 0.001f, c_tris);
            if (has_ray) {
                best_t = __uint_as_float((uint32_t)(cw.key[lane] >> 32));
                if (!(ng.y & 0xFF000000u) && !stk_pop(stk, sp, top, ng)) {
                    {  // traversal complete
                        const uint32_t ti = cw.tri[lane];
                        uint32_t inst = 0xFFFFFFFFu, prim = 0u;
                        float2 uv = make_float2(0.0f, 0.0f);
                        if (ti != 0xFFFFFFFFu) {
                            const float* tq = reinterpret_cast<const float*>(sc.tris + (size_t)ti * 3u);
                            inst = __float_as_uint(__ldg(tq + 7));
                            prim = __float_as_uint(__ldg(tq + 11));
                            uv = cw.uv[lane];
                        }
                        wf.hit[slot] = make_uint4(__float_as_uint(uv.x), __float_as_uint(uv.y), inst, prim);
                        has_ray = false;
                    }
                }
            }
            const uint32_t live = __ballot_sync(0xFFFFFFFFu, has_ray);
            if (32u - (uint32_t)__popc(live) > max_idle) break;
        }
    }
    if (COUNT) {
        warp_add(&counters->extend_rays, c_rays);
        warp_add(&counters->nodes, c_nodes);
        warp_add(&counters->tris, c_tris);
    }
}

// ---- primary rays as packets ------------------------------------------------------------------------
// The primary rays of a screen tile leave the same point and differ by a fraction of a degree, so a warp walks
// the BVH ONCE for RT_PKT_RPL x 32 of them (RT_PKT_RPL horizontally adjacent 4 x 8 tiles of one frame, see
// primary_slot) with a shared group stack instead of once per ray: lane (l & 7) tests child box (l & 7)
// against the interval hull of the packet's reciprocal directions (Wald et al.'s interval-arithmetic slab
// test: conservative for every ray of the packet; the hull is widened by RT_PKT_MARGIN to cover the rounding of
// this test), the hit bits are OR-reduced over the warp (REDUX), and every triangle the hull reaches is
// tested by all 32 lanes, each against its own RT_PKT_RPL rays with the same watertight edge test and the same
// (t, id) acceptance rule as the per-ray kernel.  Closest hits do not depend on which extra triangles were
// tested, so the hit records are bit-identical to per-ray traversal
// (tests/test_gpu_parity.py::test_packet_primary_matches_per_ray_1080p).  On the Sponza proxy a 16 x 8 packet
// visits 21.8 nodes and 17.9 triangles where ONE ray visits 17.4 and 6.2 (tools/packet_proto.py).
// Packets whose rays do not share origin bits and direction signs (tiles across an axis plane) are handed to
// the per-ray kernel through a queue.
#ifndef RT_PKT_RPL
#define RT_PKT_RPL 4
#endif
#ifndef RT_PKT_MIN_BLOCKS
/* measured (ms/frame of this kernel, RPL 4): 3 blocks / 128 registers 0.171, 5 / 96 0.152, 6 / 80 0.145, 8 / 64 (spills) 0.140 */
#define RT_PKT_MIN_BLOCKS 8
#endif
#define RT_PKT_MARGIN 4e-6f
#ifndef RT_PKT_GROUP_SLOT
#define RT_PKT_GROUP_SLOT 1
#endif

// min / max over the warp of floats that all carry the sign `neg` (unsigned order of the bit patterns)
__device__ __forceinline__ float warp_fmin_signed(float v, bool neg) {
    const uint32_t b = __float_as_uint(v);
    return __uint_as_float(neg ? __reduce_max_sync(0xFFFFFFFFu, b) : __reduce_min_sync(0xFFFFFFFFu, b));
}
__device__ __forceinline__ float warp_fmax_signed(float v, bool neg) {
    const uint32_t b = __float_as_uint(v);
    return __uint_as_float(neg ? __reduce_min_sync(0xFFFFFFFFu, b) : __reduce_max_sync(0xFFFFFFFFu, b));
}

#ifndef RT_FUSED_RAYGEN
#define RT_FUSED_RAYGEN 0 /* measured r1t: packet kernel 8.21 -> 11.97 ms per 64-frame batch, ray-gen launch saved 1.69: 0.506 -> 0.544 ms/frame; off */
#endif
template <bool COUNT>
__global__ void __launch_bounds__(128, RT_PKT_MIN_BLOCKS) wf_extend_packet_kernel(SceneView sc, FrameParams fp, WfView wf, uint32_t* fb_queue,
                                                                                  uint32_t width, uint32_t n_px,
                                                                                  DevCounters* counters) {
    constexpr int R = RT_PKT_RPL;
    const uint32_t n = wf.ctrl[CT_Q0];
    const uint32_t n_groups = (n + 32u * R - 1u) / (32u * R);
    const uint32_t lane = threadIdx.x & 31u, c = lane & 7u;
    uint32_t c_nodes = 0, c_tris = 0, c_rays = 0;
    uint2 stk[RT_STACK_SIZE];
#if RT_PKT_GROUP_SLOT
    const bool grp_fast = n % (32u * R) == 0u && n_px % (32u * R) == 0u && (width * RT_TILE_H) % (32u * R) == 0u && width % RT_TILE_W == 0u;
#endif
    for (;;) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(&wf.ctrl[CT_CUR_PKT], 1u);
        g = __shfl_sync(0xFFFFFFFFu, g, 0);
        if (g >= n_groups) break;
        // ---- the packet's rays: ray k of this lane is entry g * 32 R + 32 k + lane of the tiled pixel order
        uint32_t slot[R];
        Shear sh[R];
        float ikz[R], bt[R], bu[R], bv[R];
        vec3 idl[R];  // 1/d of the lane's rays, until the hulls are built
        uint32_t bid[R], btri[R];
        uint32_t kzs = 0u, valid = 0u;
        float ox, oy, oz;
        float ilx, ihx, ily, ihy, ilz, ihz;
        uint32_t oct;
        bool ok = true;
        {
            uint32_t sb0 = 0, ox0 = 0, oy0 = 0, oz0 = 0;
#if RT_PKT_GROUP_SLOT
            // primary_slot() for the whole group at once: when frames and 8-row bands are whole numbers of groups, the group's
            // 32 R entries lie in one band of one frame, so the two run-time divisions are taken once per group instead of once per ray
            // (they were 2 % of the kernel's instructions).  Same slots as primary_slot(): tile (b0 >> 5) + k, pixel `lane` of the tile.
            uint32_t gs_base = 0u, gs_step = 0u;
            if (grp_fast) {
                const uint32_t i0 = g * (32u * R);
                const uint32_t lane_f = i0 / n_px, r0 = i0 - lane_f * n_px;
                const uint32_t band_px = width * RT_TILE_H;
                const uint32_t band = r0 / band_px, b0 = r0 - band * band_px;
                const bool tiled = (band + 1u) * band_px <= n_px;  // rows that do not fill a band keep the linear order
                gs_base = lane_f * n_px + (tiled ? (band * RT_TILE_H + lane / RT_TILE_W) * width + (b0 >> 5) * RT_TILE_W + (lane % RT_TILE_W)
                                                 : r0 + lane);
                gs_step = tiled ? RT_TILE_W : 32u;
            }
#endif
#pragma unroll
            for (int k = 0; k < R; k++) {
                const uint32_t i = (g * R + (uint32_t)k) * 32u + lane;
                if (i < n) valid |= 1u << k;
#if RT_PKT_GROUP_SLOT
                slot[k] = grp_fast ? gs_base + (uint32_t)k * gs_step : primary_slot(i < n ? i : n - 1u, width, n_px);
#else
                slot[k] = primary_slot(i < n ? i : n - 1u, width, n_px);
#endif
#if RT_FUSED_RAYGEN
                // the ray-gen shader runs here (simple.rgen:70-85): no ray-gen launch, no primary ray read back.  The hit
                // shader wants the direction and the seeds, the per-ray kernel (packets handed back) the origin as well.
                float4 o4, d4;
                {
                    const uint32_t lane_f = slot[k] / n_px, px = slot[k] - lane_f * n_px;
                    const uint32_t lrow = px / width, x = px - lrow * width;
                    const uint32_t y = strip_global_row(fp, lrow);
                    uint32_t seed;
                    vec3 o, d;
                    raygen_ray(fp, fp.frame + lane_f * fp.frame_stride, x, y, &seed, &o, &d);
                    o4 = make_float4(o.x, o.y, o.z, 0.0f);
                    d4 = make_float4(d.x, d.y, d.z, __uint_as_float(seed));
                    if (i < n) {
                        wf.ray_d[slot[k]] = d4;
                        wf.seeds[slot[k]] = make_uint2(seed, seed);
                    }
                }
#else
                const float4 o4 = wf.ray_o[slot[k]], d4 = wf.ray_d[slot[k]];
#endif
                const uint32_t sb = (__float_as_uint(d4.x) >> 31) | ((__float_as_uint(d4.y) >> 31) << 1) | ((__float_as_uint(d4.z) >> 31) << 2);
                if (k == 0) {  // every lane executes every shuffle: no short-circuit around a *_sync call
                    sb0 = __shfl_sync(0xFFFFFFFFu, sb, 0);
                    ox0 = __shfl_sync(0xFFFFFFFFu, __float_as_uint(o4.x), 0);
                    oy0 = __shfl_sync(0xFFFFFFFFu, __float_as_uint(o4.y), 0);
                    oz0 = __shfl_sync(0xFFFFFFFFu, __float_as_uint(o4.z), 0);
                }
                ok = ok && sb == sb0 && __float_as_uint(o4.x) == ox0 && __float_as_uint(o4.y) == oy0 && __float_as_uint(o4.z) == oz0 &&
                     fabsf(d4.x) >= 1e-20f && fabsf(d4.y) >= 1e-20f && fabsf(d4.z) >= 1e-20f;  // the last three also reject NaN
                const TravRay r = trav_setup(mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z));
                sh[k] = r.sh;
                const uint32_t kz = (r.rbits >> 3) & 3u;
                kzs |= kz << (2 * k);
                ikz[k] = kz == 0u ? r.idx : (kz == 1u ? r.idy : r.idz);
                idl[k] = mk3(r.idx, r.idy, r.idz);
                bt[k] = 10000.0f; bu[k] = 0.0f; bv[k] = 0.0f; bid[k] = 0xFFFFFFFFu; btri[k] = 0u;
            }
            ox = __uint_as_float(ox0); oy = __uint_as_float(oy0); oz = __uint_as_float(oz0);
            oct = sb0;
        }
        if (!__all_sync(0xFFFFFFFFu, ok)) {
#pragma unroll
            for (int k = 0; k < R; k++) {
                const bool v = (valid >> k) & 1u;
                const uint32_t bal = __ballot_sync(0xFFFFFFFFu, v);
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(&wf.ctrl[CT_FB], (uint32_t)__popc(bal));
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                if (v) fb_queue[base + __popc(bal & ((1u << lane) - 1u))] = slot[k];
            }
#if RT_FUSED_RAYGEN
            // origins are not compared lane by lane here (all equal inv_view's translation): the lane's own value is the ray's
            if (valid) {
                const float4 o4 = make_float4(fp.inv_view[12], fp.inv_view[13], fp.inv_view[14], 0.0f);
#pragma unroll
                for (int k = 0; k < R; k++)
                    if ((valid >> k) & 1u) wf.ray_o[slot[k]] = o4;
            }
#endif
            continue;
        }
        if (COUNT) c_rays += __popc(valid);
        const bool s0 = (oct & 1u) != 0u, s1 = (oct & 2u) != 0u, s2 = (oct & 4u) != 0u;
        // Hulls of the reciprocal directions, one per sub-packet (sub-packet k = ray k of every lane = one 4 x 8 tile),
        // widened so that the products below err on the safe side.  Lane l tests child (l & 7) against the hull of
        // sub-packet (l >> 3): a child is entered when any sub-packet reaches it, and a triangle is tested only by the rays
        // of the sub-packets that reach its leaf box (on the Sponza proxy 2.2 of 4: tools/packet_proto.py).
        const uint32_t sub = lane >> 3;
        {
            float lx = 0.f, hx = 0.f, ly = 0.f, hy = 0.f, lz = 0.f, hz = 0.f;
#pragma unroll
            for (int k = 0; k < R; k++) {
                const float a = warp_fmin_signed(idl[k].x, s0), b = warp_fmax_signed(idl[k].x, s0);
                const float c2 = warp_fmin_signed(idl[k].y, s1), d2 = warp_fmax_signed(idl[k].y, s1);
                const float e2 = warp_fmin_signed(idl[k].z, s2), f2 = warp_fmax_signed(idl[k].z, s2);
                if (R == 4 ? (sub == (uint32_t)k) : (k == 0)) { lx = a; hx = b; ly = c2; hy = d2; lz = e2; hz = f2; }
                else if (R != 4) { lx = fminf(lx, a); hx = fmaxf(hx, b); ly = fminf(ly, c2); hy = fmaxf(hy, d2); lz = fminf(lz, e2); hz = fmaxf(hz, f2); }
            }
            ilx = lx; ihx = hx; ily = ly; ihy = hy; ilz = lz; ihz = hz;
        }
        ilx -= fabsf(ilx) * RT_PKT_MARGIN; ihx += fabsf(ihx) * RT_PKT_MARGIN;
        ily -= fabsf(ily) * RT_PKT_MARGIN; ihy += fabsf(ihy) * RT_PKT_MARGIN;
        ilz -= fabsf(ilz) * RT_PKT_MARGIN; ihz += fabsf(ihz) * RT_PKT_MARGIN;
        float tmaxp = 10000.0f;  // max over the lane's sub-packet of the rays' current closest distances
        uint2 ng = make_uint2(0u, 0x80000000u);
        int sp = 0;
        for (;;) {
            // pop the nearest child of the current group (warp-uniform)
            const int bit = 31 - __clz(ng.y);
            ng.y &= ~(1u << bit);
            const uint32_t sl = ((uint32_t)bit - 24u) ^ oct;
            const uint32_t node = ng.x + __popc(ng.y & 0xFFu & ((1u << sl) - 1u));
            if (ng.y & 0xFF000000u) stk[sp++] = ng;
            const uint4* np = sc.nodes + (size_t)node * RT_NODE_WORDS;
            const uint4* hp = np + node_chunk(node, 0u);  // the node's chunks are stored rotated (rt_types.cuh)
            const uint4 w0 = __ldg(hp);
            const uint2 w1 = __ldg(reinterpret_cast<const uint2*>(hp + 1));  // child_base, tri_base
            const uint32_t meta = __ldg(reinterpret_cast<const unsigned char*>(hp) + 24 + c);
            const unsigned short* qx = reinterpret_cast<const unsigned short*>(np + node_chunk(node, 1u)) + c;  // q[axis][lo/hi][slot]
            const unsigned short* qy = reinterpret_cast<const unsigned short*>(np + node_chunk(node, 2u)) + c;
            const unsigned short* qz = reinterpret_cast<const unsigned short*>(np + node_chunk(node, 3u)) + c;
            const float qxl = __half2float(__ushort_as_half(__ldg(qx))), qxh = __half2float(__ushort_as_half(__ldg(qx + 8)));
            const float qyl = __half2float(__ushort_as_half(__ldg(qy))), qyh = __half2float(__ushort_as_half(__ldg(qy + 8)));
            const float qzl = __half2float(__ushort_as_half(__ldg(qz))), qzh = __half2float(__ushort_as_half(__ldg(qz + 8)));
            if (COUNT && lane == 0) c_nodes++;
            const uint32_t ew = w0.w;
            const float scx = __uint_as_float((ew & 0xFFu) << 23), scy = __uint_as_float(((ew >> 8) & 0xFFu) << 23),
                        scz = __uint_as_float(((ew >> 16) & 0xFFu) << 23);
            // child box relative to the common origin
            const float rx = __uint_as_float(w0.x) - ox, ry = __uint_as_float(w0.y) - oy, rz = __uint_as_float(w0.z) - oz;
            const float lx = fmaf(qxl, scx, rx), hx = fmaf(qxh, scx, rx);
            const float ly = fmaf(qyl, scy, ry), hy = fmaf(qyh, scy, ry);
            const float lz = fmaf(qzl, scz, rz), hz = fmaf(qzh, scz, rz);
            // entry / exit of the hull per axis: extremes of the four plane x reciprocal-direction products
            const float tn = fmaxf(fmaxf(fminf(fminf(lx * ilx, lx * ihx), fminf(hx * ilx, hx * ihx)),
                                         fminf(fminf(ly * ily, ly * ihy), fminf(hy * ily, hy * ihy))),
                                   fminf(fminf(lz * ilz, lz * ihz), fminf(hz * ilz, hz * ihz)));
            const float tf = fminf(fminf(fmaxf(fmaxf(lx * ilx, lx * ihx), fmaxf(hx * ilx, hx * ihx)),
                                         fmaxf(fmaxf(ly * ily, ly * ihy), fmaxf(hy * ily, hy * ihy))),
                                   fmaxf(fmaxf(lz * ilz, lz * ihz), fmaxf(hz * ilz, hz * ihz)));
            const uint32_t idx = (meta & 31u) ^ (((meta & 0x18u) == 0x18u) ? oct : 0u);
            const bool hit = fmaxf(tn, 0.001f) <= fminf(tf, tmaxp);
            const uint32_t mybits = hit ? ((meta >> 5) << idx) : 0u;
            const uint32_t hm = __reduce_or_sync(0xFFFFFFFFu, mybits);
            ng = make_uint2(w1.x, (hm & 0xFF000000u) | (w0.w >> 24));
            uint32_t tmask = hm & 0x00FFFFFFu;
            uint32_t sm[R];  // triangle bits reached by each sub-packet
            if (R == 4 && tmask) {
#pragma unroll
                for (int k = 0; k < R; k++) sm[k] = __reduce_or_sync(0xFFFFFFFFu, sub == (uint32_t)k ? mybits : 0u);
            } else {
#pragma unroll
                for (int k = 0; k < R; k++) sm[k] = tmask;
            }
            if (tmask) {
                do {  // warp-uniform
                    const int tb = __ffs(tmask) - 1;
                    tmask &= tmask - 1u;
                    const uint32_t ti = w1.y + (uint32_t)tb;
                    const float4* tp = sc.tris + (size_t)ti * 3u;
                    const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
                    if (COUNT && lane == 0) c_tris++;
                    const uint32_t id = __float_as_uint(q0.w);
#pragma unroll
                    for (int k = 0; k < R; k++) {
                        if (!((sm[k] >> tb) & 1u)) continue;  // warp-uniform: this sub-packet's hull misses the leaf box
                        float U, V, W, det;
                        if (tri_edges(sh[k], q0, q1, q2, &U, &V, &W, &det)) {
                            const uint32_t kz = (kzs >> (2 * k)) & 3u;
                            float tt, uu, vv;
                            tri_finish(U, V, W, det, q0, q1, q2, kz, kz == 0u ? ox : (kz == 1u ? oy : oz), ikz[k], &tt, &uu, &vv);
                            const bool take = (tt > 0.001f) && ((bid[k] == 0xFFFFFFFFu) ? (tt < bt[k]) : ((tt < bt[k]) || (tt == bt[k] && id < bid[k])));
                            if (take) { bt[k] = tt; bu[k] = uu; bv[k] = vv; bid[k] = id; btri[k] = ti; }
                        }
                    }
                } while (tmask);
                if (R == 4) {  // farthest current hit of the lane's sub-packet (distances are positive)
#pragma unroll
                    for (int k = 0; k < R; k++) {
                        const float m = __uint_as_float(__reduce_max_sync(0xFFFFFFFFu, __float_as_uint(bt[k])));
                        if (sub == (uint32_t)k) tmaxp = m;
                    }
                } else {
                    float m = bt[0];
#pragma unroll
                    for (int k = 1; k < R; k++) m = fmaxf(m, bt[k]);
                    tmaxp = __uint_as_float(__reduce_max_sync(0xFFFFFFFFu, __float_as_uint(m)));
                }
            }
            if (!(ng.y & 0xFF000000u)) {
                if (sp == 0) break;
                ng = stk[--sp];
            }
        }
#pragma unroll
        for (int k = 0; k < R; k++) {
            if ((valid >> k) & 1u) {
                uint32_t inst = 0xFFFFFFFFu, prim = 0u;
                if (bid[k] != 0xFFFFFFFFu) {
                    const float* tq = reinterpret_cast<const float*>(sc.tris + (size_t)btri[k] * 3u);
                    inst = __float_as_uint(__ldg(tq + 7));
                    prim = __float_as_uint(__ldg(tq + 11));
                }
                wf.hit[slot[k]] = make_uint4(__float_as_uint(bu[k]), __float_as_uint(bv[k]), inst, prim);
            }
        }
    }
    if (COUNT) {
        warp_add(&counters->extend_rays, c_rays);
        warp_add(&counters->nodes, c_nodes);
        warp_add(&counters->tris, c_tris);
    }
}

// The shade kernel wants all of a path's state in flight at once; ptxas sinks plain loads below the miss test, so the
// lines behind it are requested with a prefetch it leaves in place.
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// shade: simple.rmiss / simple.rchit for every queued path + the ray-gen loop tail (simple.rgen:106-118).
// Colour terms are added in the reference's order: NEE, emissive, sky.  When the NEE term waits for
// its visibility ray, the later terms travel with the shadow-queue entry and wf_shadow_kernel adds them.
#ifndef RT_SHADE_MIN_BLOCKS
#define RT_SHADE_MIN_BLOCKS 8  /* light-less variant.  6 blocks (80 registers) beat 4 (128): 0.219 -> 0.195 ms/frame; since the path-state diet (r1t) it fits 56 registers
                                  without spills: 7 / 8 / 9 / 10 blocks per SM -> frame 0.4762 / 0.4717 / 0.4692 / 0.4715 ms (r1u A/B); config 5 (4K, depth 3): 0.752 / 0.744 / 0.850 / - -> 8 */
#endif
#ifndef RT_SHADE_LIT_MIN_BLOCKS
#define RT_SHADE_LIT_MIN_BLOCKS 7  /* the variant with the NEE code: 72 registers + 176 B of spills still beat 6 blocks (1.279 -> 1.262 ms/frame, config 4b) */
#endif
#ifndef RT_PACK_SEEDS
#define RT_PACK_SEEDS 1
#endif
// DENSE = the first bounce of a light-less scene with the FIRST-HIT FILTER: a warp scans 128 consecutive slots at a time and ends,
// without running the hit shader, the paths whose outcome it can tell from the hit record and the path's seed alone:
// (a) a miss: the sky term of simple.rgen:106-109 on the initial payload (colour 0, throughput 1);
// (b) a hit on an object with DevObject::dead_diffuse whose lobe draw picks the Lambert lobe with a direction above the surface
//     (cosine direction: z = sqrt(ry) > 0 iff ry > 0): sample_brdf returns value = albedo * (1 - metallic) * (1 - F) / pi = 0 exactly
//     with a finite positive pdf, shade_hit leaves payload.hit true and the throughput 0, the ray-gen loop stops at
//     simple.rgen:115-118 and the pixel keeps colour 0.  The draws are the hit shader's own, in its order: light-sample decision
//     (simple.rchit:622), rx, ry (:638), lobe (:408).  What makes (b) exact is checked per object on the host (rt_api.cu).
// The scan writes every slot's initial colour (coalesced); the survivors (43 % on the Sponza proxy, where the reference's loader
// leaves every material metallic) wait in a per-warp ring until 32 of them fill a warp, so the hit shader runs with all lanes and
// in pixel order.  History (profiles/r2q_filter.md): a global survivor queue filled at the end of the packet kernel interleaved
// the tiles of all packets in flight — hit shader 11 % faster for 57 % fewer entries, bounce-1 kernels 32 % slower; marking the
// hit records there instead cost the packet kernel 8 % (scattered colour writes, a dependent table load per ray).
#ifndef RT_SHADE_DENSE_MIN_BLOCKS
#define RT_SHADE_DENSE_MIN_BLOCKS RT_SHADE_MIN_BLOCKS /* 8 / 9 / 10 blocks per SM (64 / 56 / 48 registers): frame 0.4207 / 0.4205 / 0.4207 ms (r2v A/B, twice each) */
#endif
template <bool COUNT, bool FIRST, bool LIGHTS, bool DENSE = false>
__global__ void __launch_bounds__(128, LIGHTS ? RT_SHADE_LIT_MIN_BLOCKS : (DENSE ? RT_SHADE_DENSE_MIN_BLOCKS : RT_SHADE_MIN_BLOCKS)) wf_shade_kernel(SceneView sc, FrameParams fp, WfView wf, const uint32_t* q_in,
                                                       uint32_t q_idx, uint32_t* q_out, int depth, DevCounters* counters) {
    constexpr bool PACKED = !LIGHTS && RT_PACK_SEEDS;
    const uint32_t n = wf.ctrl[q_idx];
    uint32_t c_hits = 0, c_nee = 0;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n_round = (n + 31u) & ~31u;  // whole warps iterate together (ballots below)
    const vec3 clearc = mk3(fp.clear[0], fp.clear[1], fp.clear[2]);
    // one entry per lane; called by all 32 lanes of a warp together (the queue appends are warp-wide)
    auto shade_entry = [&](const bool valid, const uint32_t slot_in) {
        bool push_next = false, push_shadow = false, sky = false;
        uint32_t slot = 0;
        ShadowRequest sr;
        sr.pending = false;
        NeeRequest nq;
        nq.pending = false;
        ShadeResult res;
        res.emissive_add = false;
        res.alive = false;
        PathState ps;
        if (valid) {
            slot = slot_in;
            // the state behind the miss test is requested together with the hit record (one memory latency, not two)
            prefetch_l1(&wf.ray_d[slot]);
            if (!PACKED) prefetch_l1(&wf.seeds[slot]);
            if (!FIRST && LIGHTS) prefetch_l1(&wf.ray_o[slot]);
            const uint4 hrec = wf.hit[slot];
            // first bounce: the payload initialisation of simple.rgen:76-85
            const float4 tv = FIRST ? make_float4(1.0f, 1.0f, 1.0f, 1.0f) : wf.thr[slot];
            // without lights a path's colour is zero until the sky term that ends it: nothing to read, and only that term
            // (or the first bounce, which initialises the slot) to write
            // With lights the colour changes here only through an emissive or sky term that is not deferred to the shadow
            // pass: it is read then (below), not up front, and p_sample_light rides in direction.w instead of colour.w.
            const float4 cv = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            ps.throughput = mk3(tv.x, tv.y, tv.z);
            ps.prev_brdf_pdf = tv.w;
            ps.color = mk3(cv.x, cv.y, cv.z);
            ps.p_sample_light = cv.w;
            if (hrec.z == 0xFFFFFFFFu) {  // simple.rmiss:26-29 -> simple.rgen:106-109
                sky = true;
            } else {
                // previous hit point and bounce flags feed the emissive rules of simple.rchit:641-686 only: a scene without
                // lights has no emissive surface, so its variant neither reads nor writes that record (32 B per path and bounce)
                // With lights: the previous hit point IS the origin of the ray that got here (shade_hit sets both to the hit's
                // world position), so the record is the ray-origin array itself, bounce flags in its w.
                const float4 pv = (FIRST || !LIGHTS) ? make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(PF_FIRST)) : wf.ray_o[slot];
                const float4 dv = wf.ray_d[slot];
                // light-less variant: throughput.w (previousBrdfPdf) is dead, so the two seeds ride in direction.w / throughput.w
                // and the seeds array is not touched (first bounce: ray-gen left the one initial seed in direction.w)
                const uint2 sd = PACKED ? make_uint2(__float_as_uint(dv.w), FIRST ? __float_as_uint(dv.w) : __float_as_uint(tv.w)) : wf.seeds[slot];
                if (LIGHTS && !FIRST) ps.p_sample_light = dv.w;
                const uint32_t fl = __float_as_uint(pv.w);
                ps.prev_hit_pos = mk3(pv.x, pv.y, pv.z);
                ps.first_bounce = (fl & PF_FIRST) != 0u;
                ps.is_specular = (fl & PF_SPECULAR) != 0u;
                ps.did_direct = (fl & PF_DID_DIRECT) != 0u;
                ps.seed = sd.x;
                ps.rr_seed = sd.y;
                Hit h;
                h.t = 0.0f; h.u = __uint_as_float(hrec.x); h.v = __uint_as_float(hrec.y); h.id = 0u; h.inst = hrec.z; h.prim = hrec.w;
                c_hits++;
                res = shade_hit<LIGHTS, true>(sc, ps, h, mk3(dv.x, dv.y, dv.z), sr, c_nee, &nq);
                if (!res.alive) {
                    sky = true;  // payload.hit = false -> the ray-gen shader adds the sky term
                } else {
                    // simple.rgen:111-118, then the head of the next iteration (:87, :55-68)
                    bool go = !throughput_dead(ps.throughput) && (depth + 1 < fp.max_depth);
                    if (go) go = russian_roulette(ps, depth + 1, fp.rr_start);
                    push_next = go;
                }
            }
            push_shadow = LIGHTS && nq.pending;
            const bool col_term = !push_shadow && (res.emissive_add || sky);
            if (LIGHTS && !FIRST && col_term) {
                const float4 c = wf.col[slot];
                ps.color = mk3(c.x, c.y, c.z);
            }
            if (!push_shadow) {
                if (res.emissive_add) ps.color = add3(ps.color, res.emissive);
                if (sky) ps.color = add3(ps.color, mul3(ps.throughput, clearc));
            }
            // DENSE: the scan has initialised the slot's colour (zero), only a sky term changes it
            if ((FIRST && !DENSE) || (LIGHTS ? col_term : sky)) wf.col[slot] = make_float4(ps.color.x, ps.color.y, ps.color.z, 0.0f);
            if (push_next) {
                const uint32_t fl = LIGHTS ? ((ps.is_specular ? PF_SPECULAR : 0u) | (ps.did_direct ? PF_DID_DIRECT : 0u)) : 0u;
                wf.ray_o[slot] = make_float4(res.next_origin.x, res.next_origin.y, res.next_origin.z, __uint_as_float(fl));
                wf.ray_d[slot] = make_float4(res.next_dir.x, res.next_dir.y, res.next_dir.z, PACKED ? __uint_as_float(ps.seed) : (LIGHTS ? ps.p_sample_light : 0.0f));
                wf.thr[slot] = make_float4(ps.throughput.x, ps.throughput.y, ps.throughput.z, PACKED ? __uint_as_float(ps.rr_seed) : ps.prev_brdf_pdf);
                if (!PACKED) wf.seeds[slot] = make_uint2(ps.seed, ps.rr_seed);
            }
        }
        const uint32_t qi = queue_push(&wf.ctrl[q_idx ^ 1u], push_next);
        if (push_next) q_out[qi] = slot;
        const uint32_t si = LIGHTS ? queue_push(&wf.ctrl[CT_SH], push_shadow) : 0u;
        if (LIGHTS && push_shadow) {
            // a sky term is only pending when the hit shader ended the path, in which case the throughput
            // in the state array is still the one the term needs (on the first bounce it has not been stored yet)
            if (FIRST && sky) wf.thr[slot] = make_float4(ps.throughput.x, ps.throughput.y, ps.throughput.z, ps.prev_brdf_pdf);
            // the light sample as drawn (NeeRequest); wf_nee_kernel replaces it by the visibility ray and the contribution.
            // It reads the hit point from the ray-origin array, which continuing paths have just written.
            if (!push_next) wf.ray_o[slot] = make_float4(nq.x.x, nq.x.y, nq.x.z, 0.0f);
            wf.sh_o[si] = make_float4(nq.N.x, nq.N.y, nq.N.z, __uint_as_float(nq.material));
            wf.sh_d[si] = make_float4(nq.wo.x, nq.wo.y, nq.wo.z, __uint_as_float(nq.light));
            wf.sh_c[si] = make_float4(nq.throughput.x, nq.throughput.y, nq.throughput.z, __uint_as_float(slot));
            wf.sh_n[si] = make_float4(nq.r_tri, nq.q1, nq.q2, nq.sel_pdf);
            const uint32_t fl = (res.emissive_add ? 1u : 0u) | (sky ? 2u : 0u);
            wf.sh_e[si] = make_float4(res.emissive_add ? res.emissive.x : 0.0f, res.emissive_add ? res.emissive.y : 0.0f,
                                      res.emissive_add ? res.emissive.z : 0.0f, __uint_as_float(fl));
        }
    };
    if (!DENSE) {
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
            const bool valid = i < n;
            shade_entry(valid, valid ? (q_in ? q_in[i] : i) : 0u);
        }
    } else {
        // survivors wait in a per-warp ring until 32 of them fill a warp: every round of the hit shader runs with all lanes
        // (taking each 128-slot chunk's survivors on their own left the chunk's last round half empty: 21.6 of 32 threads per
        // instruction, ncu r2p)
        __shared__ uint32_t s_ring[4][160];
        uint32_t* ring = s_ring[threadIdx.x >> 5];
        const uint32_t lane = threadIdx.x & 31u;
        const uint32_t lt = (1u << lane) - 1u;
        uint32_t head = 0u, count = 0u;  // warp-uniform
        uint32_t c0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 128u;
        for (;;) {
            while (count < 32u && c0 < n) {  // warp-uniform: take in the survivors of the warp's next chunk
                // all of the chunk's loads first (eight per lane in flight), then the decisions
                uint32_t insts[4], seeds[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint32_t i = c0 + 32u * (uint32_t)j + lane;
                    insts[j] = i < n ? wf.hit[i].z : 0xFFFFFFFFu;
                    seeds[j] = i < n ? __float_as_uint(wf.ray_d[i].w) : 0u;  // the path's seed, as ray-gen left it (PACKED)
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint32_t i = c0 + 32u * (uint32_t)j + lane;
                    bool live = false;
                    if (i < n) {
                        const uint32_t inst = insts[j];
                        float4 c4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                        if (inst == 0xFFFFFFFFu) {
                            const vec3 c = add3(mk3(0.0f, 0.0f, 0.0f), mul3(mk3(1.0f, 1.0f, 1.0f), clearc));
                            c4 = make_float4(c.x, c.y, c.z, 0.0f);
                        } else {
                            live = true;
                            const uint2 dd = __ldg(reinterpret_cast<const uint2*>(&sc.objects[inst].dead_diffuse));
                            if (dd.x != 0u) {
                                uint32_t s = seeds[j];
                                rnd01(s);
                                rnd01(s);
                                const float ry = rnd01(s);
                                const float lobe = rnd01(s);
                                if (!(lobe < __uint_as_float(dd.y)) && ry > 0.0f) {
                                    live = false;
                                    if (COUNT) c_hits++;
                                }
                            }
                        }
                        wf.col[i] = c4;
                    }
                    const uint32_t b = __ballot_sync(0xFFFFFFFFu, live);
                    if (live) ring[(head + count + (uint32_t)__popc(b & lt)) % 160u] = i;
                    count += (uint32_t)__popc(b);
                }
                c0 += (stride >> 5) * 128u;
            }
            __syncwarp();
            if (count == 0u) break;
            const uint32_t take = count < 32u ? count : 32u;  // fewer than 32 only once the warp's chunks are used up
            const bool valid = lane < take;
            const uint32_t slot_in = valid ? ring[(head + lane) % 160u] : 0u;
            __syncwarp();
            head = (head + take) % 160u;
            count -= take;
            shade_entry(valid, slot_in);
        }
    }
    if (COUNT) {
        warp_add(&counters->hits, c_hits);
        warp_add(&counters->nee, c_nee);
    }
}

// tail: the few paths that are still alive after RT_TAIL_DEPTH bounces are each run to their end by one thread, in ONE launch,
// instead of three to five launches per remaining bounce.  With a handful of rays a launch lasts as long as its slowest ray
// (60 - 100 us for a cooperative extend launch on the Sponza proxy, whatever the queue holds), so bounces 3..7 cost a
// 64-frame batch 0.43 ms in 25 launches — 1.4 % of the batch on one GPU and 5 % of an eighth of it on eight.  The loop is the
// megakernel's (simple.rgen:87-118 with the hit shader and the visibility query inline), entered with the state the shade
// kernel of bounce RT_TAIL_DEPTH - 1 left behind: the same operations on the same values in the same order, hence the same
// image bit for bit.  The host uses it only when the previous batch on this state reported a small queue (WavefrontBuffers::
// tail_hint_h); with many live paths the wavefront kernels are faster.
#ifndef RT_TAIL_DEPTH
#define RT_TAIL_DEPTH 3
#endif
template <bool COUNT, bool LIGHTS>
__global__ void __launch_bounds__(128) wf_tail_kernel(SceneView sc, FrameParams fp, WfView wf, const uint32_t* q_in, uint32_t q_idx, int depth0,
                                                      uint32_t* report, DevCounters* counters) {
    constexpr bool PACKED = !LIGHTS && RT_PACK_SEEDS;
    const uint32_t n = wf.ctrl[q_idx];
    if (report && blockIdx.x == 0 && threadIdx.x == 0) *report = n;
    uint32_t c_ext = 0, c_sh = 0, c_nodes = 0, c_tris = 0, c_hits = 0, c_nee = 0, c_snodes = 0, c_stris = 0;
    const vec3 clearc = mk3(fp.clear[0], fp.clear[1], fp.clear[2]);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t slot = q_in[i];
        // the path's state as wf_shade_kernel<., false, LIGHTS> reads it
        const float4 tv = wf.thr[slot], dv = wf.ray_d[slot], ov = wf.ray_o[slot];
        PathState ps;
        ps.throughput = mk3(tv.x, tv.y, tv.z);
        ps.prev_brdf_pdf = tv.w;
        ps.color = mk3(0.0f, 0.0f, 0.0f);
        ps.p_sample_light = LIGHTS ? dv.w : 0.0f;
        if (LIGHTS) {
            const float4 c = wf.col[slot];
            ps.color = mk3(c.x, c.y, c.z);
        }
        const uint32_t fl = LIGHTS ? __float_as_uint(ov.w) : PF_FIRST;
        ps.prev_hit_pos = LIGHTS ? mk3(ov.x, ov.y, ov.z) : mk3(0.0f, 0.0f, 0.0f);
        ps.first_bounce = (fl & PF_FIRST) != 0u;
        ps.is_specular = (fl & PF_SPECULAR) != 0u;
        ps.did_direct = (fl & PF_DID_DIRECT) != 0u;
        const uint2 sd = PACKED ? make_uint2(__float_as_uint(dv.w), __float_as_uint(tv.w)) : wf.seeds[slot];
        ps.seed = sd.x;
        ps.rr_seed = sd.y;
        vec3 o = mk3(ov.x, ov.y, ov.z), d = mk3(dv.x, dv.y, dv.z);
        bool sky = false;
        for (int depth = depth0;;) {
            c_ext++;
            const Hit h = traverse<false, COUNT>(sc, o, d, 0.001f, 10000.0f, 0xFFFFFFFFu, c_nodes, c_tris);
            bool alive = false;
            if (h.id != 0xFFFFFFFFu) {
                c_hits++;
                ShadowRequest sr;
                const ShadeResult res = shade_hit<LIGHTS, false>(sc, ps, h, d, sr, c_nee);
                if (LIGHTS && sr.pending) {
                    c_sh++;
                    const Hit sh = traverse<true, COUNT>(sc, sr.origin, sr.dir, 0.001f, sr.tmax, sr.light_object, c_snodes, c_stris);
                    if (sh.id == 0xFFFFFFFFu) ps.color = add3(ps.color, sr.contrib);
                }
                if (res.emissive_add) ps.color = add3(ps.color, res.emissive);
                alive = res.alive;
                o = res.next_origin;
                d = res.next_dir;
            }
            if (!alive) {  // simple.rmiss / payload.hit = false: the sky term ends the path
                ps.color = add3(ps.color, mul3(ps.throughput, clearc));
                sky = true;
                break;
            }
            if (throughput_dead(ps.throughput)) break;
            if (++depth >= fp.max_depth) break;
            if (!russian_roulette(ps, depth, fp.rr_start)) break;
        }
        // without lights a path's colour is zero (as the first shade launch stored it) until its sky term
        if (LIGHTS || sky) wf.col[slot] = make_float4(ps.color.x, ps.color.y, ps.color.z, 0.0f);
    }
    if (COUNT) {
        warp_add(&counters->extend_rays, c_ext); warp_add(&counters->shadow_rays, c_sh);
        warp_add(&counters->nodes, c_nodes + c_snodes); warp_add(&counters->tris, c_tris + c_stris);
        warp_add(&counters->shadow_nodes, c_snodes); warp_add(&counters->shadow_tris, c_stris);
        warp_add(&counters->hits, c_hits); warp_add(&counters->nee, c_nee);
    }
}

// nee: the second half of the light sample (sampleLight geometry, evaluateLightMIS, the isVisibleRQ ray) for every entry
// of the shadow queue, one thread per sample.  Inside the hit shader this code ran with 6 of 32 lanes active (ncu r1p,
// config 4b) and made the kernel 3 700 instructions long; here every lane has a sample.  An entry whose sample
// contributes nothing (light facing away, below the horizon ...) gets tmax < 0: no ray, only the later colour terms.
template <bool COUNT>
__global__ void __launch_bounds__(128) wf_nee_kernel(SceneView sc, WfView wf) {
    const uint32_t n = wf.ctrl[CT_SH];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = wf.sh_o[i], b = wf.sh_d[i], c = wf.sh_c[i], r = wf.sh_n[i];
        const uint32_t slot = __float_as_uint(c.w);
        const float4 xo = wf.ray_o[slot];
        const Mat mat = load_material(sc.materials + __float_as_uint(a.w));
        const vec3 N = mk3(a.x, a.y, a.z);
        const Basis basis = create_basis(N);
        const float psl = glsl_clamp(mat.roughness, 0.1f, 0.9f);
        vec3 direct;
        ShadowRequest sr;
        if (nee_evaluate(sc, mk3(xo.x, xo.y, xo.z), N, basis, mk3(b.x, b.y, b.z), mat, __float_as_uint(b.w), r.w, r.x, r.y, r.z, &direct,
                         &sr)) {
            const vec3 contrib = div3z(mul3(mk3(c.x, c.y, c.z), direct), psl);
            wf.sh_o[i] = make_float4(sr.origin.x, sr.origin.y, sr.origin.z, sr.tmax);
            wf.sh_d[i] = make_float4(sr.dir.x, sr.dir.y, sr.dir.z, __uint_as_float(sr.light_object));
            wf.sh_c[i] = make_float4(contrib.x, contrib.y, contrib.z, c.w);
        } else {
            wf.sh_o[i] = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
        }
    }
}

// the colour terms that follow the NEE term of a shadow-queue entry, in the reference's order (emissive, sky)
__device__ __forceinline__ void shadow_entry_finish(const FrameParams& fp, const WfView& wf, uint32_t entry, bool visible) {
    const float4 c = wf.sh_c[entry], e = wf.sh_e[entry];
    const uint32_t slot = __float_as_uint(c.w), fl = __float_as_uint(e.w);
    float4 cv = wf.col[slot];
    vec3 col = mk3(cv.x, cv.y, cv.z);
    if (visible) col = add3(col, mk3(c.x, c.y, c.z));
    if (fl & 1u) col = add3(col, mk3(e.x, e.y, e.z));
    if (fl & 2u) {
        const float4 tv = wf.thr[slot];
        col = add3(col, mul3(mk3(tv.x, tv.y, tv.z), mk3(fp.clear[0], fp.clear[1], fp.clear[2])));
    }
    wf.col[slot] = make_float4(col.x, col.y, col.z, cv.w);
}

// shadow: isVisibleRQ (simple.rchit:350-385) for every queued NEE sample, then the ordered colour adds
template <bool COUNT>
__global__ void __launch_bounds__(128, RT_TRAV_MIN_BLOCKS) wf_shadow_kernel(SceneView sc, FrameParams fp, WfView wf, DevCounters* counters) {
    const uint32_t n = wf.ctrl[CT_SH];
    uint32_t c_nodes = 0, c_tris = 0, c_rays = 0;
    WarpFetch wfetch;
    wfetch.next = wfetch.end = 0u;
    wfetch.drained = false;
    TravState st;
    TravStack stk;
    bool has_ray = false;
    uint32_t entry = 0;
    for (;;) {
        const uint32_t i = warp_fetch(wfetch, &wf.ctrl[CT_CUR_SHW], n, !has_ray);
        if (i != 0xFFFFFFFFu) {
            entry = i;
            const float4 o = wf.sh_o[i], d = wf.sh_d[i];
            if (o.w < 0.0f) {  // the sample contributes nothing: no ray
                shadow_entry_finish(fp, wf, i, false);
            } else {
                trav_init(st, mk3(o.x, o.y, o.z), mk3(d.x, d.y, d.z), 0.001f, o.w, __float_as_uint(d.w));
                has_ray = true;
                c_rays++;
            }
        }
        if (__ballot_sync(0xFFFFFFFFu, has_ray) == 0u) {
            if (__ballot_sync(0xFFFFFFFFu, i != 0xFFFFFFFFu) == 0u) break;  // nothing fetched: the queue is empty
            continue;                                                         // only ray-less entries this time
        }
        const uint32_t min_active = wfetch.drained ? 1u : (32u - RT_REFILL_IDLE + 1u);
        while (has_ray) {
            if (trav_step<true, COUNT>(sc, st, stk, c_nodes, c_tris)) {
                shadow_entry_finish(fp, wf, entry, st.best.id == 0xFFFFFFFFu);
                has_ray = false;
            }
            if (__popc(__activemask()) < min_active) break;
        }
    }
    if (COUNT) {
        warp_add(&counters->shadow_rays, c_rays);
        warp_add(&counters->nodes, c_nodes);
        warp_add(&counters->tris, c_tris);
        warp_add(&counters->shadow_nodes, c_nodes);
        warp_add(&counters->shadow_tris, c_tris);
    }
}

// shadow, cooperative triangle tests: isVisibleRQ for every queued NEE sample with the (ray, triangle) pairs of a warp
// pooled as in wf_extend_coop_kernel; "occluded" is a flag, so pooling cannot change it.
template <bool COUNT>
__global__ void __launch_bounds__(128, RT_COOP_MIN_BLOCKS) wf_shadow_coop_kernel(SceneView sc, FrameParams fp, WfView wf,
                                                                                 DevCounters* counters) {
    __shared__ CoopWarp s_cw[4];
    const uint32_t n = wf.ctrl[CT_SH];
    const uint32_t lane = threadIdx.x & 31u;
    CoopWarp& cw = s_cw[threadIdx.x >> 5];
    uint32_t c_nodes = 0, c_tris = 0, c_rays = 0;
    WarpFetch wfetch;
    wfetch.next = wfetch.end = 0u;
    wfetch.drained = false;
    NodeRay nr;
    nr.ox = nr.oy = nr.oz = nr.idx = nr.idy = nr.idz = 0.0f; nr.oct = 0u;
    TravStack stk;
    uint2 ng = make_uint2(0u, 0u), top = make_uint2(0u, 0u);
    int sp = 0;
    float tmax = 0.0f;
    bool has_ray = false;
    uint32_t entry = 0;
    for (;;) {
        const uint32_t i = warp_fetch(wfetch, &wf.ctrl[CT_CUR_SHW], n, !has_ray);
        if (i != 0xFFFFFFFFu) {
            entry = i;
            const float4 o = wf.sh_o[i], d = wf.sh_d[i];
            if (o.w < 0.0f) {  // the sample contributes nothing: no ray
                shadow_entry_finish(fp, wf, i, false);
            } else {
                const TravRay r = trav_setup(mk3(o.x, o.y, o.z), mk3(d.x, d.y, d.z));
                tmax = o.w;
                coop_init_ray(cw, lane, r, tmax, __float_as_uint(d.w));
                nr.ox = r.ox; nr.oy = r.oy; nr.oz = r.oz; nr.idx = r.idx; nr.idy = r.idy; nr.idz = r.idz; nr.oct = r.rbits & 7u;
                ng = make_uint2(0u, 0x80000000u);
                sp = 0;
                top.y = 0u;
                has_ray = true;
                c_rays++;
            }
        }
        __syncwarp();
        if (__ballot_sync(0xFFFFFFFFu, has_ray) == 0u) {
            if (__ballot_sync(0xFFFFFFFFu, i != 0xFFFFFFFFu) == 0u) break;  // nothing fetched: the queue is empty
            continue;                                                         // only ray-less entries this time
        }
        const uint32_t max_idle = wfetch.drained ? 31u : (RT_REFILL_IDLE - 1u);
        for (;;) {
            uint32_t tmask = 0u, tbase = 0u;
            if (has_ray) tmask = coop_node<COUNT>(sc, nr, ng, sp, stk, top, 0.001f, tmax, &tbase, c_nodes);
            coop_triangles<true, COUNT>(sc, cw, lane, tmask, tbase, 0.001f, c_tris);
            if (has_ray) {
                const bool occluded = cw.tri[lane] == 0u;
                bool done = occluded;
                if (!done && !(ng.y & 0xFF000000u)) done = !stk_pop(stk, sp, top, ng);
                if (done) {
                    shadow_entry_finish(fp, wf, entry, !occluded);
                    has_ray = false;
                }
            }
            const uint32_t live = __ballot_sync(0xFFFFFFFFu, has_ray);
            if (32u - (uint32_t)__popc(live) > max_idle) break;
        }
    }
    if (COUNT) {
        warp_add(&counters->shadow_rays, c_rays);
        warp_add(&counters->nodes, c_nodes);
        warp_add(&counters->tris, c_tris);
        warp_add(&counters->shadow_nodes, c_nodes);
        warp_add(&counters->shadow_tris, c_tris);
    }
}

// accumulate: simple.rgen:121-136
// the lanes of a pixel are folded into the running mean in frame order, so the float result is the one of
// n_lanes successive single-frame dispatches
__global__ void __launch_bounds__(256) wf_accumulate_kernel(FrameParams fp, WfView wf, float4* image) {
    const uint32_t n_px = fp.local_rows * fp.width;
    const uint32_t px = blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= n_px) return;
    float4 acc = fp.sample ? image[px] : make_float4(0.f, 0.f, 0.f, 0.f);
    for (uint32_t l = 0; l < fp.n_lanes; l++) {
        const float4 c = wf.col[(size_t)l * n_px + px];
        acc = accumulate_pixel(acc, mk3(c.x, c.y, c.z), fp.clamp, fp.sample + l);
    }
    image[px] = acc;
}

template <typename K>
static int persistent_grid(K kernel, int threads, int num_sms) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess || per_sm <= 0) per_sm = 4;
    return per_sm * num_sms;
}

template <bool COUNT>
static int frame_wavefront_t(const SceneView& sc, const FrameParams& fp, FrameResources& res, cudaStream_t stream,
                             std::string& err) {
    WavefrontBuffers* w = res.wf;
    const uint32_t n_px = fp.local_rows * fp.width;
    const uint32_t n_slots = n_px * fp.n_lanes;
    if (n_slots == 0) return RT_OK;
    if (!w || w->n_slots < n_slots) { err = "wavefront buffers not allocated"; return RT_ERR_NO_TARGET; }
    WfView v;
    v.ray_o = w->ray_o.p; v.ray_d = w->ray_d.p; v.thr = w->thr.p; v.col = w->col.p;
    v.seeds = w->seeds.p; v.hit = w->hit.p; v.sh_o = w->sh_o.p; v.sh_d = w->sh_d.p; v.sh_c = w->sh_c.p; v.sh_e = w->sh_e.p; v.sh_n = w->sh_n.p;
    v.ctrl = w->ctrl.p;
#ifndef RT_EXTEND_COOP
#define RT_EXTEND_COOP 1
#endif
    // persistent grids: per context (res.dims), i.e. per device — nothing process-wide
    LaunchDims local_dims;
    LaunchDims& dims = res.dims ? *res.dims : local_dims;
    if (!dims.num_sms) dims.num_sms = query_num_sms();
    LaunchDims::Set& ds = dims.set[COUNT ? 1 : 0];
    if (!ds.ready) {
        const int sms = dims.num_sms;
        ds.coop = persistent_grid(wf_extend_coop_kernel<COUNT>, 128, sms);
        ds.pkt = persistent_grid(wf_extend_packet_kernel<COUNT>, 128, sms);
        ds.ext = persistent_grid(wf_extend_kernel<COUNT>, 128, sms);
        ds.shw = persistent_grid(wf_shadow_kernel<COUNT>, 128, sms);
        ds.shd = persistent_grid(wf_shade_kernel<COUNT, false, true>, 128, sms);
        ds.shd0 = persistent_grid(wf_shade_kernel<COUNT, false, false>, 128, sms);
        ds.shd0d = persistent_grid(wf_shade_kernel<COUNT, true, false, true>, 128, sms);
        ds.nee = persistent_grid(wf_nee_kernel<COUNT>, 128, sms);
        ds.ready = true;
    }
    const int g_ext = ds.ext, g_shw = ds.shw, g_shd = ds.shd, g_shd0 = ds.shd0, g_pkt = ds.pkt, g_coop = ds.coop, g_nee = ds.nee;
    const uint32_t gb_px = (n_px + 255) / 256;
    KernelProfiler none;
    KernelProfiler& pf = res.prof ? *res.prof : none;
    uint64_t n_launch = 0;
    // with packets on the first bounce the ray-gen shader runs inside the packet kernel
    if (!RT_FUSED_RAYGEN || (fp.flags & RT_FLAG_PER_RAY_PRIMARY)) {
        pf.begin(RT_K_RAYGEN, stream);
        wf_raygen_kernel<<<dim3((fp.width + 255u) / 256u, std::min(fp.local_rows, 65535u), std::min(fp.n_lanes, 65535u)), 256, 0, stream>>>(fp, v);
        pf.end(stream);
        n_launch++;
    }
    uint32_t q_in = 0;
    // no more blocks than the widest possible queue needs
    const int need = (int)((n_slots + 127) / 128);
    // the tail kernel takes over at bounce RT_TAIL_DEPTH when at least two bounces remain and the last batch on this state had
    // few paths left there (the hint is read unsynchronised: it steers the choice of kernels, not the result)
    const uint32_t tail_hint = *reinterpret_cast<volatile uint32_t*>(w->tail_hint_h);
    const char* tail_env = getenv("RT_TAIL");  // tests and A/B runs: "0" = never, "1" = whenever the depths allow it
    const bool tail_small = tail_env ? tail_env[0] == '1' : tail_hint <= (uint32_t)dims.num_sms * 1024u;
    const bool tail_ok = RT_TAIL_DEPTH > 0 && !(fp.flags & RT_FLAG_PER_RAY_PRIMARY) && fp.max_depth - RT_TAIL_DEPTH >= 2 && tail_small;
    // first-hit filter (the scan of wf_shade_kernel<., true, false, true>): light-less scenes with qualifying objects.  RT_FILTER=0
    // turns it off (tests, A/B runs); the image does not depend on it.
    const char* filter_env = getenv("RT_FILTER");
    const bool filter = sc.any_dead_diffuse != 0u && sc.n_lights == 0u && RT_PACK_SEEDS && !RT_FUSED_RAYGEN && !(filter_env && filter_env[0] == '0');
    for (int depth = 0; depth < fp.max_depth; depth++) {
        if (depth == RT_TAIL_DEPTH && tail_ok) {
            const int g = std::min(dims.num_sms * 8, need);
            pf.begin(RT_K_EXTEND, stream);
            if (sc.n_lights != 0u)
                wf_tail_kernel<COUNT, true><<<g, 128, 0, stream>>>(sc, fp, v, w->queue[q_in].p, q_in, depth, w->tail_hint_d, res.counters);
            else
                wf_tail_kernel<COUNT, false><<<g, 128, 0, stream>>>(sc, fp, v, w->queue[q_in].p, q_in, depth, w->tail_hint_d, res.counters);
            pf.end(stream);
            n_launch++;
            break;
        }
        wf_reset_kernel<<<1, 32, 0, stream>>>(v.ctrl, q_in, depth == 0 ? n_slots : 0xFFFFFFFFu,
                                             depth == RT_TAIL_DEPTH ? w->tail_hint_d : nullptr);
        const uint32_t* qi = depth == 0 ? nullptr : w->queue[q_in].p;
        if (depth == 0 && !(fp.flags & RT_FLAG_PER_RAY_PRIMARY)) {
            // packets first; what they hand back (tiles across an axis plane) goes through the per-ray kernel
            pf.begin(RT_K_EXTEND_PACKET, stream);
            wf_extend_packet_kernel<COUNT><<<std::min(g_pkt, need), 128, 0, stream>>>(sc, fp, v, w->queue[0].p, fp.width, n_px,
                                                                                     res.counters);
            pf.end(stream);
            pf.begin(RT_K_EXTEND, stream);
            if (RT_EXTEND_COOP)
                wf_extend_coop_kernel<COUNT><<<std::min(g_coop, need), 128, 0, stream>>>(sc, v, w->queue[0].p, CT_FB, fp.width, n_px,
                                                                                        res.counters);
            else
                wf_extend_kernel<COUNT><<<std::min(g_ext, need), 128, 0, stream>>>(sc, v, w->queue[0].p, CT_FB, fp.width, n_px,
                                                                                   res.counters);
            n_launch++;
        } else {
            pf.begin(RT_K_EXTEND, stream);
            if (RT_EXTEND_COOP)
                wf_extend_coop_kernel<COUNT><<<std::min(g_coop, need), 128, 0, stream>>>(sc, v, qi, q_in, fp.width, n_px, res.counters);
            else
                wf_extend_kernel<COUNT><<<std::min(g_ext, need), 128, 0, stream>>>(sc, v, qi, q_in, fp.width, n_px, res.counters);
        }
        pf.end(stream);
        pf.begin(RT_K_SHADE, stream);
        {
            const bool lit = sc.n_lights != 0u;
            const int g = std::min(lit ? g_shd : ((depth == 0 && filter) ? ds.shd0d : g_shd0), need);
            uint32_t* qo = w->queue[q_in ^ 1u].p;
            if (depth == 0 && lit) wf_shade_kernel<COUNT, true, true><<<g, 128, 0, stream>>>(sc, fp, v, qi, q_in, qo, depth, res.counters);
            else if (depth == 0 && filter) wf_shade_kernel<COUNT, true, false, true><<<g, 128, 0, stream>>>(sc, fp, v, qi, q_in, qo, depth, res.counters);
            else if (depth == 0) wf_shade_kernel<COUNT, true, false><<<g, 128, 0, stream>>>(sc, fp, v, qi, q_in, qo, depth, res.counters);
            else if (lit) wf_shade_kernel<COUNT, false, true><<<g, 128, 0, stream>>>(sc, fp, v, qi, q_in, qo, depth, res.counters);
            else wf_shade_kernel<COUNT, false, false><<<g, 128, 0, stream>>>(sc, fp, v, qi, q_in, qo, depth, res.counters);
        }
        pf.end(stream);
        if (sc.n_lights != 0u) {  // no light, no visibility ray
            pf.begin(RT_K_SHADE, stream);  // the deferred half of the hit shader's light sample
            wf_nee_kernel<COUNT><<<std::min(g_nee, need), 128, 0, stream>>>(sc, v);
            pf.end(stream);
            n_launch++;
            pf.begin(RT_K_SHADOW, stream);
#ifndef RT_SHADOW_COOP
#define RT_SHADOW_COOP 1
#endif
            if (RT_SHADOW_COOP && !(fp.flags & RT_FLAG_PER_RAY_PRIMARY))
                wf_shadow_coop_kernel<COUNT><<<std::min(g_coop, need), 128, 0, stream>>>(sc, fp, v, res.counters);
            else
                wf_shadow_kernel<COUNT><<<std::min(g_shw, need), 128, 0, stream>>>(sc, fp, v, res.counters);
            pf.end(stream);
            n_launch++;
        }
        n_launch += 3;
        q_in ^= 1u;
    }
    // batches fold into the image in frame order, whichever stream they ran on
    if (res.wait_accumulate) FRAME_TRY(cudaStreamWaitEvent(stream, res.wait_accumulate, 0));
    pf.begin(RT_K_ACCUMULATE, stream);
    wf_accumulate_kernel<<<gb_px, 256, 0, stream>>>(fp, v, res.image);
    pf.end(stream);
    if (res.done_accumulate) FRAME_TRY(cudaEventRecord(res.done_accumulate, stream));
    n_launch++;
    if (res.launches) *res.launches += n_launch;
    FRAME_TRY(cudaGetLastError());
    return RT_OK;
}

// ---- primary visibility through pixel CORNERS (restir.rgen:47-57) with the packet kernel ---------------------------------------
// The ReSTIR surface pass shoots one un-jittered ray per pixel: as coherent as the path tracer's first bounce, so it goes through
// the same 128-ray packets (hand-backs through the cooperative kernel) instead of one BVH walk per ray.  Slot i = (index in the row
// list) * width + x; hit records land in the wavefront state's hit array.
__global__ void __launch_bounds__(256) corner_raygen_kernel(FrameParams fp, const uint32_t* rows, uint32_t n_rows, WfView wf) {
    const uint32_t n = n_rows * fp.width;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t r = i / fp.width, x = i - r * fp.width;
        const uint32_t y = rows ? __ldg(rows + r) : r;
        vec3 o, d;
        camera_ray(fp, (float)x, (float)y, false, &o, &d);  // pixel corner, direction not re-normalised
        wf.ray_o[i] = make_float4(o.x, o.y, o.z, 0.0f);
        wf.ray_d[i] = make_float4(d.x, d.y, d.z, 0.0f);
    }
}

// the same rays one per thread (a strip context whose halo covers more rows than its wavefront state has slots for, and
// RT_FLAG_PER_RAY_PRIMARY): the same traversal code under the same arithmetic contract, hence the same hit records bit for bit
template <bool COUNT>
__global__ void __launch_bounds__(128) corner_hits_per_ray_kernel(SceneView sc, FrameParams fp, const uint32_t* rows, uint32_t n_rows, uint4* hits,
                                                                  DevCounters* counters) {
    const uint32_t n = n_rows * fp.width;
    uint32_t c_nodes = 0, c_tris = 0, c_rays = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t r = i / fp.width, x = i - r * fp.width;
        const uint32_t y = rows ? __ldg(rows + r) : r;
        vec3 o, d;
        camera_ray(fp, (float)x, (float)y, false, &o, &d);
        c_rays++;
        const Hit h = traverse<false, COUNT>(sc, o, d, 0.001f, 10000.0f, 0xFFFFFFFFu, c_nodes, c_tris);
        hits[i] = make_uint4(__float_as_uint(h.u), __float_as_uint(h.v), h.id == 0xFFFFFFFFu ? 0xFFFFFFFFu : h.inst, h.prim);
    }
    if (COUNT) {
        warp_add(&counters->extend_rays, c_rays);
        warp_add(&counters->nodes, c_nodes);
        warp_add(&counters->tris, c_tris);
    }
}

template <bool COUNT>
static int corner_hits_t(const SceneView& sc, const FrameParams& fp, FrameResources& res, const uint32_t* d_rows, uint32_t n_rows,
                         uint4* own_hits, cudaStream_t stream, std::string& err, const uint4** hits) {
    WavefrontBuffers* w = res.wf;
    const uint32_t n = n_rows * fp.width;
    if (n == 0) { *hits = own_hits; return RT_OK; }
    if (!w || w->n_slots < n || (fp.flags & RT_FLAG_PER_RAY_PRIMARY)) {
        if (!own_hits) { err = "corner pass: no hit buffer"; return RT_ERR_NO_TARGET; }
        int sms = res.dims && res.dims->num_sms ? res.dims->num_sms : query_num_sms();
        corner_hits_per_ray_kernel<COUNT><<<std::min<uint32_t>((n + 127) / 128, (uint32_t)sms * 32), 128, 0, stream>>>(sc, fp, d_rows, n_rows, own_hits,
                                                                                                                  res.counters);
        FRAME_TRY(cudaGetLastError());
        if (res.launches) *res.launches += 1;
        *hits = own_hits;
        return RT_OK;
    }
    WfView v;
    v.ray_o = w->ray_o.p; v.ray_d = w->ray_d.p; v.thr = w->thr.p; v.col = w->col.p;
    v.seeds = w->seeds.p; v.hit = w->hit.p; v.sh_o = w->sh_o.p; v.sh_d = w->sh_d.p; v.sh_c = w->sh_c.p; v.sh_e = w->sh_e.p; v.sh_n = w->sh_n.p;
    v.ctrl = w->ctrl.p;
    LaunchDims local_dims;
    LaunchDims& dims = res.dims ? *res.dims : local_dims;
    if (!dims.num_sms) dims.num_sms = query_num_sms();
    const int g_pkt = persistent_grid(wf_extend_packet_kernel<COUNT>, 128, dims.num_sms), g_coop = persistent_grid(wf_extend_coop_kernel<COUNT>, 128, dims.num_sms);
    const int need = (int)((n + 127) / 128);
    corner_raygen_kernel<<<std::min<uint32_t>((n + 255) / 256, (uint32_t)dims.num_sms * 16), 256, 0, stream>>>(fp, d_rows, n_rows, v);
    wf_reset_kernel<<<1, 32, 0, stream>>>(v.ctrl, 0u, n);
    wf_extend_packet_kernel<COUNT><<<std::min(g_pkt, need), 128, 0, stream>>>(sc, fp, v, w->queue[0].p, fp.width, n, res.counters);
    wf_extend_coop_kernel<COUNT><<<std::min(g_coop, need), 128, 0, stream>>>(sc, v, w->queue[0].p, CT_FB, fp.width, n, res.counters);
    FRAME_TRY(cudaGetLastError());
    if (res.launches) *res.launches += 4;
    *hits = w->hit.p;
    return RT_OK;
}

int launch_corner_hits(const SceneView& sc, const FrameParams& fp, FrameResources& res, const uint32_t* d_rows, uint32_t n_rows,
                       uint4* own_hits, cudaStream_t stream, std::string& err, const uint4** hits) {
    return (fp.flags & RT_FLAG_COUNTERS) ? corner_hits_t<true>(sc, fp, res, d_rows, n_rows, own_hits, stream, err, hits)
                                         : corner_hits_t<false>(sc, fp, res, d_rows, n_rows, own_hits, stream, err, hits);
}
uint32_t wavefront_slots(const WavefrontBuffers* w) { return w ? w->n_slots : 0u; }

int launch_frame_wavefront(const SceneView& sc, const FrameParams& fp, FrameResources& res, cudaStream_t stream,
                           std::string& err) {
    return (fp.flags & RT_FLAG_COUNTERS) ? frame_wavefront_t<true>(sc, fp, res, stream, err)
                                         : frame_wavefront_t<false>(sc, fp, res, stream, err);
}

// ================================================================================================
// G-buffer pass: shaders/restir.rgen:43-92 + restir.rchit:36-59 + restir.rmiss:18-20
// ================================================================================================
__global__ void __launch_bounds__(128) gbuffer_kernel(SceneView sc, FrameParams fp, GBufferImages g, DevCounters* counters) {
    const uint32_t n_slots = fp.local_rows * fp.width;
    uint32_t c_nodes = 0, c_tris = 0, c_rays = 0;
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < n_slots; slot += gridDim.x * blockDim.x) {
        const uint32_t lrow = slot / fp.width, x = slot - lrow * fp.width;
        const uint32_t y = strip_global_row(fp, lrow);
        vec3 o, d;
        camera_ray(fp, (float)x, (float)y, false, &o, &d);  // pixel corner, direction not re-normalised
        c_rays++;
        const Hit h = traverse<false, true>(sc, o, d, 0.001f, 10000.0f, 0xFFFFFFFFu, c_nodes, c_tris);
        if (h.id != 0xFFFFFFFFu) {
            const DevObject* obj = sc.objects + h.inst;
            const Mat mat = load_material(sc.materials + __ldg(&obj->material_index));
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
            const vec3 N = normalize3(mat3_mul(n9, nrm));
            const uchar4 alb = make_uchar4(unorm8_rne(mat.albedo.x), unorm8_rne(mat.albedo.y), unorm8_rne(mat.albedo.z), 255);
            g.albedo[slot] = alb;
            g.output[slot] = alb;
            g.normal_rough[slot] = make_ushort4(half_bits_rne(N.x), half_bits_rne(N.y), half_bits_rne(N.z), half_bits_rne(mat.roughness));
            g.position_metal[slot] = make_float4(wp.x, wp.y, wp.z, mat.metallic);
            g.emission[slot] = make_ushort4(half_bits_rne(mat.emission_color.x), half_bits_rne(mat.emission_color.y),
                                            half_bits_rne(mat.emission_color.z), half_bits_rne(mat.emission_power));
        } else {
            g.albedo[slot] = make_uchar4(0, 0, 0, 255);
            g.output[slot] = make_uchar4(0, 0, 0, 255);
            g.normal_rough[slot] = make_ushort4(0, 0, 0, 0x3C00);
            g.position_metal[slot] = make_float4(0.0f, 0.0f, 0.0f, 1.0f);
            g.emission[slot] = make_ushort4(0, 0, 0, 0x3C00);
        }
    }
    if (counters) {
        warp_add(&counters->extend_rays, c_rays);
        warp_add(&counters->nodes, c_nodes);
        warp_add(&counters->tris, c_tris);
    }
}

int launch_gbuffer(const SceneView& sc, const FrameParams& fp, const GBufferImages& g, DevCounters* counters, int num_sms,
                   cudaStream_t stream, std::string& err) {
    const uint32_t n_slots = fp.local_rows * fp.width;
    if (n_slots == 0) return RT_OK;
    const uint32_t grid = std::min<uint32_t>((n_slots + 127) / 128, (uint32_t)num_sms * 32);
    gbuffer_kernel<<<grid, 128, 0, stream>>>(sc, fp, g, counters);
    FRAME_TRY(cudaGetLastError());
    return RT_OK;
}

// ================================================================================================
// what the blit to B8G8R8A8_SRGB shows (raytracing_renderer.odin:299-316): clamp to [0,1], sRGB OETF,
// 8-bit.  The OETF is evaluated as a count of thresholds (the linear values at which the 8-bit code
// steps), so the result does not depend on a device powf.
// ================================================================================================
__constant__ float c_srgb_thresholds[256];  // [k] = smallest linear value that encodes to >= k (k >= 1)

__device__ __forceinline__ uint8_t srgb8(float v) {
    if (!(v > 0.0f)) return 0;
    int lo = 0, hi = 255;  // largest k with thresholds[k] <= v
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (c_srgb_thresholds[mid] <= v) lo = mid; else hi = mid - 1;
    }
    return (uint8_t)lo;
}

__global__ void srgb8_kernel(const float4* image, uchar4* out, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 c = image[i];
    out[i] = make_uchar4(srgb8(c.x), srgb8(c.y), srgb8(c.z), 255);
}

int launch_srgb8(const float4* image, uchar4* out, uint32_t n, LaunchDims& dims, cudaStream_t stream, std::string& err) {
    if (!dims.srgb_table_ready) {  // __constant__ memory is per device: once per context
        float t[256];
        t[0] = 0.0f;
        for (int k = 1; k < 256; k++) {
            const double s = ((double)k - 0.5) / 255.0;  // encoded value at the rounding boundary
            const double lin = (s <= 0.04045) ? s / 12.92 : pow((s + 0.055) / 1.055, 2.4);
            t[k] = (float)lin;
        }
        FRAME_TRY(cudaMemcpyToSymbolAsync(c_srgb_thresholds, t, sizeof(t), 0, cudaMemcpyHostToDevice, stream));
        FRAME_TRY(cudaStreamSynchronize(stream));  // `t` is on this stack frame
        dims.srgb_table_ready = true;
    }
    if (n == 0) return RT_OK;
    srgb8_kernel<<<(n + 255) / 256, 256, 0, stream>>>(image, out, n);
    FRAME_TRY(cudaGetLastError());
    return RT_OK;
}

}  // namespace rt
