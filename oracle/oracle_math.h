// ORACLE — test infrastructure only.  Never linked, imported or executed by the product path
// (raytracer_b200/csrc).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load it.
//
// Parity status: PINNED TO THE SHIPPED SHADER BINARIES for the shader code; driver-defined parts unpinned.
// The reference (ARTurleite6/RayTracer) has no CPU implementation, no tests and no golden vectors for this path,
// and its Vulkan pipeline cannot be run here (no graphics stack).  What CAN be run is the SPIR-V the reference ships and
// loads (shaders/rgen.spv, rchit.spv, rmiss.spv, restir_*.spv): tools/spirv_interp.py executes those binaries instruction
// by instruction, tools/spirv_golden.py commits the frames, G-buffers and function vectors they compute
// (tests/golden/spirv_frames.npz), and tests/test_oracle.py requires this oracle to reproduce them bit for bit — which pins
// evaluation order, association, branch structure and the order of the random draws.  Not pinned by anything the reference
// holds: the precision of the GLSL built-ins Vulkan leaves to the driver (normalize, dot, sin, cos, pow ...), the BVH
// and its traversal (inside the driver), and which hit a terminate-on-first-hit ray query commits.  For those this file
// states plain IEEE-754 binary32 arithmetic with a fixed evaluation order ("the arithmetic contract"), so that
// a second, independently written implementation (the CUDA kernels) can be compared bit for bit.
//
// Build: g++ -O2 -ffp-contract=off -mfma (no -ffast-math).  fmaf() appears only where the
// contract says "fused".
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

struct V3 {
    float x, y, z;
};

static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 v3s(float s) { return V3{s, s, s}; }
static inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
static inline V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
static inline V3 neg(V3 a) { return V3{-a.x, -a.y, -a.z}; }

// GLSL min/max/clamp (GLSL 4.60 §8.3): min(x,y) = y < x ? y : x ; max(x,y) = x < y ? y : x
static inline float gmin(float x, float y) { return (y < x) ? y : x; }
static inline float gmax(float x, float y) { return (x < y) ? y : x; }
static inline float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }

// dot = (x*x' + y*y') + z*z', no fusion
static inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
static inline float length(V3 a) { return sqrtf(dot(a, a)); }
// normalize = v * (1 / length(v)); a zero vector gives NaN, as IEEE says
static inline V3 normalize(V3 a) {
    float inv = 1.0f / length(a);
    return a * inv;
}
// mix(a,b,t) = a*(1-t) + b*t
static inline float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
static inline V3 mix3(V3 a, V3 b, float t) { return a * (1.0f - t) + b * t; }
static inline float smoothstep(float e0, float e1, float x) {
    float t = gclamp((x - e0) / (e1 - e0), 0.0f, 1.0f);
    return (t * t) * (3.0f - 2.0f * t);
}
// reflect(I,N) = I - N * (2*dot(N,I))
static inline V3 reflect(V3 I, V3 N) {
    float k = 2.0f * dot(N, I);
    return I - N * k;
}
static inline float pow5(float x) {
    float x2 = x * x;
    float x4 = x2 * x2;
    return x4 * x;
}
static inline float signf(float x) { return (float)((x > 0.0f) - (x < 0.0f)); }

static const float kPi = 3.14159265359f;      // math.glsl:1 (M_PI) rounded to binary32
static const float kTwoPi = 6.28318530718f;   // 2.0 * M_PI folded, rounded to binary32

// sin/cos contract: Cody-Waite reduction by pi/2 in three fused steps, Cephes single-precision
// minimax polynomials on [-pi/4, pi/4] in fused Horner form.  Argument is phi >= 0 (2*pi*xi).
static inline void sincos_contract(float a, float* s_out, float* c_out) {
    int j = (int)(a * 0.636619772f + 0.5f);
    float kf = (float)j;
    float r = fmaf(kf, -1.5703125f, a);
    r = fmaf(kf, -4.837512969970703125e-4f, r);
    r = fmaf(kf, -7.54978995489188e-8f, r);
    float r2 = r * r;
    float ps = fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    ps = fmaf(r2, ps, -1.6666654611e-1f);
    float sn = fmaf(r * r2, ps, r);
    float pc = fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc = fmaf(r2, pc, 4.166664568298827e-2f);
    float cs = fmaf(r2 * r2, pc, fmaf(-0.5f, r2, 1.0f));
    switch (j & 3) {
        case 0: *s_out = sn; *c_out = cs; break;
        case 1: *s_out = cs; *c_out = -sn; break;
        case 2: *s_out = -sn; *c_out = -cs; break;
        default: *s_out = -cs; *c_out = sn; break;
    }
}

// ---- RNG: shaders/random.glsl:23-53 (TEA-16 seed, Numerical-Recipes LCG stream) ----
static inline uint32_t tea(uint32_t val0, uint32_t val1) {
    uint32_t v0 = val0, v1 = val1, s0 = 0;
    for (int n = 0; n < 16; n++) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}
static inline uint32_t lcg(uint32_t& prev) {
    prev = 1664525u * prev + 1013904223u;
    return prev & 0x00FFFFFFu;
}
static inline float rnd(uint32_t& prev) { return (float)lcg(prev) / 16777216.0f; }

// ---- column-major 4x4 helpers (m[c*4+r]) ----
// point: ((c0*x + c1*y) + c2*z) + c3 ; direction: (c0*x + c1*y) + c2*z
static inline V3 xform_point(const float* m, V3 p) {
    return V3{((m[0] * p.x + m[4] * p.y) + m[8] * p.z) + m[12],
              ((m[1] * p.x + m[5] * p.y) + m[9] * p.z) + m[13],
              ((m[2] * p.x + m[6] * p.y) + m[10] * p.z) + m[14]};
}
static inline V3 xform_dir(const float* m, V3 d) {
    return V3{(m[0] * d.x + m[4] * d.y) + m[8] * d.z, (m[1] * d.x + m[5] * d.y) + m[9] * d.z,
              (m[2] * d.x + m[6] * d.y) + m[10] * d.z};
}
// transpose(inverse(mat3(M))) as three columns: (a1 x a2, a2 x a0, a0 x a1) / det
static inline void normal_matrix(const float* m, float out9[9]) {
    V3 a0 = v3(m[0], m[1], m[2]), a1 = v3(m[4], m[5], m[6]), a2 = v3(m[8], m[9], m[10]);
    V3 c0 = cross(a1, a2), c1 = cross(a2, a0), c2 = cross(a0, a1);
    float inv = 1.0f / dot(a0, c0);
    c0 = c0 * inv; c1 = c1 * inv; c2 = c2 * inv;
    out9[0] = c0.x; out9[1] = c0.y; out9[2] = c0.z;
    out9[3] = c1.x; out9[4] = c1.y; out9[5] = c1.z;
    out9[6] = c2.x; out9[7] = c2.y; out9[8] = c2.z;
}
static inline V3 mul_nm(const float* n9, V3 v) {
    return V3{(n9[0] * v.x + n9[3] * v.y) + n9[6] * v.z, (n9[1] * v.x + n9[4] * v.y) + n9[7] * v.z,
              (n9[2] * v.x + n9[5] * v.y) + n9[8] * v.z};
}

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

}  // namespace orc
