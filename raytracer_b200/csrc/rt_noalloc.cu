// rt_noalloc.cu — rt_trace_rays_closest_static: the closest-hit verification hook without any allocation.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  rt_trace_rays_closest (rt_api.cu)
// stages the caller's rays and results in three cudaMalloc'ed device buffers per call; here that scratch is ONE file-scope
// __device__ array of fixed size, held by every GPU that loads librt.so until the process exits:
//     256 B control block (work cursor) + RT_STATIC_MAX_RAYS x (32 B ray + 12 B t,u,v + 4 B id) = 3 MiB + 256 B.
// The array is shared by every call and every context on a device and keeps the last call's contents: the cursor is
// cleared at the start of each call, rays / results are overwritten up to n.  Two calls that overlap on one GPU (two host
// threads, two contexts on one device) race on it — callers serialise them; one call at a time per device is the contract.
//
// Same traversal code (rt_trav.cuh: traverse<false, .>) and the same arithmetic flags as the existing hook, so the
// results are bit-identical to rt_trace_rays_closest.
#include <mutex>

#include "rt_internal.h"
#include "rt_trav.cuh"

using namespace rt;

#define RT_STATIC_CTRL_BYTES 256u
#define RT_STATIC_RAY_BYTES 32u
#define RT_STATIC_BYTES (RT_STATIC_CTRL_BYTES + (size_t)RT_STATIC_MAX_RAYS * (RT_STATIC_RAY_BYTES + 12u + 4u))

// In host code this name is a host-side stand-in, NOT a device address: the address a kernel may use comes from
// cudaGetSymbolAddress (static_scratch_ptr below), once per device, with that device current.
static __device__ __align__(256) unsigned char g_static_scratch[RT_STATIC_BYTES];

namespace {

struct StaticView {
    uint32_t* cursor;     // next unclaimed ray; must be 0 when the kernel starts
    const float4* rays;   // n x 2 float4
    float* tuv;           // n x 3
    uint32_t* ids;        // n
};

__host__ __device__ inline StaticView static_view(unsigned char* base) {
    StaticView v;
    v.cursor = reinterpret_cast<uint32_t*>(base);
    unsigned char* p = base + RT_STATIC_CTRL_BYTES;
    v.rays = reinterpret_cast<const float4*>(p);
    p += (size_t)RT_STATIC_MAX_RAYS * RT_STATIC_RAY_BYTES;
    v.tuv = reinterpret_cast<float*>(p);
    p += (size_t)RT_STATIC_MAX_RAYS * 12u;
    v.ids = reinterpret_cast<uint32_t*>(p);
    return v;
}

// Variant of trace_rays_kernel<false> (rt_frame.cu): rays and results live in the static scratch, and a warp claims 32
// rays at a time from the cursor in its control block instead of walking a fixed grid stride (rays differ widely in cost).
__global__ void __launch_bounds__(128) trace_rays_static_kernel(SceneView sc, unsigned char* scratch, uint32_t n) {
    const StaticView v = static_view(scratch);
    const uint32_t lane = threadIdx.x & 31u;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(v.cursor, 32u);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= n) break;
        const uint32_t i = base + lane;
        if (i < n) {
            const float4 a = v.rays[2 * (size_t)i], b = v.rays[2 * (size_t)i + 1];
            uint32_t nn = 0, nt = 0;
            const Hit h = traverse<false, false>(sc, mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), a.w, b.w, 0xFFFFFFFFu, nn, nt);
            const bool ok = h.id != 0xFFFFFFFFu;
            v.ids[i] = h.id;
            v.tuv[3 * (size_t)i + 0] = ok ? h.t : 0.0f;
            v.tuv[3 * (size_t)i + 1] = ok ? h.u : 0.0f;
            v.tuv[3 * (size_t)i + 2] = ok ? h.v : 0.0f;
        }
        __syncwarp();
    }
}

// device address of g_static_scratch on `device` (which must be current); resolved once per device
constexpr int kMaxDevices = 64;
std::mutex g_ptr_mutex;
unsigned char* g_ptr[kMaxDevices] = {};

cudaError_t static_scratch_ptr(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_ptr_mutex);
    if (!g_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_ptr[device];
    return cudaSuccess;
}

}  // namespace

#define NA_TRY(expr)                                                                   \
    do {                                                                               \
        const cudaError_t _e = (expr);                                                 \
        if (_e != cudaSuccess) {                                                       \
            ctx->error = std::string(#expr) + ": " + cudaGetErrorString(_e);           \
            return RT_ERR_CUDA;                                                        \
        }                                                                              \
    } while (0)

extern "C" {

uint32_t rt_static_max_rays(void) { return RT_STATIC_MAX_RAYS; }

int rt_trace_rays_closest_static(rt_ctx* ctx, const float* rays, uint32_t n, float* out_tuv, uint32_t* out_id) {
    if (!ctx) return RT_ERR_INVALID_ARGUMENT;
    if (!ctx->children.empty()) {  // multi-GPU context: the hook runs on its first device (every device holds the same BVH)
        rt_ctx* c = ctx->children[0];
        const int rc = rt_trace_rays_closest_static(c, rays, n, out_tuv, out_id);
        if (rc != RT_OK) ctx->error = c->error;
        return rc;
    }
    if (n > RT_STATIC_MAX_RAYS) {
        ctx->error = "rt_trace_rays_closest_static: n = " + std::to_string(n) + " exceeds the static scratch (" +
                     std::to_string(RT_STATIC_MAX_RAYS) + " rays per call): split the batch or use rt_trace_rays_closest";
        return RT_ERR_INVALID_ARGUMENT;
    }
    if (n == 0) return RT_OK;
    if (!rays || !out_tuv || !out_id) { ctx->error = "rt_trace_rays_closest_static: null argument"; return RT_ERR_INVALID_ARGUMENT; }
    if (!ctx->has_scene) { ctx->error = "no scene: call rt_build_scene first"; return RT_ERR_NO_SCENE; }
    // bringing edits to the device (table upload, BVH refit / rebuild) allocates: this entry point never does it
    if (ctx->tables_dirty || ctx->geometry_dirty || !ctx->refit_objects.empty()) {
        ctx->error = "rt_trace_rays_closest_static: the scene has edits that are not on the device yet; make it resident first "
                     "(rt_scene_info_get, or any trace call) — this entry point does not allocate";
        return RT_ERR_NO_SCENE;
    }
    NA_TRY(cudaSetDevice(ctx->device));
    unsigned char* scratch = nullptr;
    NA_TRY(static_scratch_ptr(ctx->device, &scratch));
    const StaticView v = static_view(scratch);
    int sms = ctx->dims.num_sms;
    if (sms <= 0) {
        NA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
        if (sms <= 0) sms = 148;
    }
    NA_TRY(cudaMemsetAsync(v.cursor, 0, RT_STATIC_CTRL_BYTES, ctx->stream));  // the cursor keeps the last call's value
    NA_TRY(cudaMemcpyAsync(const_cast<float4*>(v.rays), rays, (size_t)n * RT_STATIC_RAY_BYTES, cudaMemcpyHostToDevice, ctx->stream));
    const uint32_t grid = std::min<uint32_t>((n + 127u) / 128u, (uint32_t)sms * 8u);
    trace_rays_static_kernel<<<grid, 128, 0, ctx->stream>>>(rt_scene_view(ctx), scratch, n);
    NA_TRY(cudaGetLastError());
    NA_TRY(cudaMemcpyAsync(out_id, v.ids, (size_t)n * 4u, cudaMemcpyDeviceToHost, ctx->stream));
    NA_TRY(cudaMemcpyAsync(out_tuv, v.tuv, (size_t)n * 12u, cudaMemcpyDeviceToHost, ctx->stream));
    NA_TRY(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

}  // extern "C"
