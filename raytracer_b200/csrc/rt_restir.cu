// rt_restir.cu — ReSTIR DI on the G-buffer.  Reservoir layout: shaders/restir/restir_structs.glsl:1-23.
// The reference holds no algorithm (shaders/restir/restir_di.rgen:17-22 is a non-compiling stub), so this
// path is parity-unpinned; see DESIGN.md §7.
#include "rt_internal.h"

extern "C" {

void rt_default_restir_params(rt_restir_params* p) {
    if (!p) return;
    p->flags = RT_RESTIR_TEMPORAL | RT_RESTIR_SPATIAL;
    p->initial_candidates = 8;
    p->spatial_samples = 4;
    p->spatial_radius = 16.0f;
    p->history_cap = 20;
    p->reserved[0] = p->reserved[1] = p->reserved[2] = 0;
}

int rt_restir_frame(rt_ctx* ctx, const float camera_ubo[64], const float clear_color[3], uint32_t frame_number,
                    const rt_restir_params* params) {
    if (!ctx) return RT_ERR_INVALID_ARGUMENT;
    (void)camera_ubo; (void)clear_color; (void)frame_number; (void)params;
    ctx->set_error("rt_restir_frame: ReSTIR kernels are not part of this build yet");
    return RT_ERR_INVALID_ARGUMENT;
}

int rt_readback_reservoirs(rt_ctx* ctx, rt_reservoir* dst, size_t bytes) {
    if (!ctx) return RT_ERR_INVALID_ARGUMENT;
    (void)dst; (void)bytes;
    ctx->set_error("rt_readback_reservoirs: ReSTIR kernels are not part of this build yet");
    return RT_ERR_NO_TARGET;
}

}  // extern "C"
