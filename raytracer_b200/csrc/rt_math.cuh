// rt_math.cuh — the arithmetic contract of the device path.
//
// The reference's shaders rely on GLSL built-ins whose precision is driver-defined
// (shaders/simple.rchit: pow, sin, cos, sqrt, normalize, mix, smoothstep, reflect ...).  This
// header fixes one IEEE-754 binary32 evaluation order for each of them so that the CUDA kernels
// and the independently written CPU oracle can be compared bit for bit.  Compile with
// `-fmad=false` (nvcc) / `-ffp-contract=off` (g++): a fused multiply-add appears only where
// fmaf() is written.  No --use_fast_math: '/' and sqrtf are correctly rounded.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD inline
#endif

namespace rt {

struct vec3 {
    float x, y, z;
};

RT_HD vec3 mk3(float x, float y, float z) { vec3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_HD vec3 splat3(float s) { return mk3(s, s, s); }
RT_HD vec3 add3(vec3 a, vec3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_HD vec3 sub3(vec3 a, vec3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_HD vec3 mul3(vec3 a, vec3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
RT_HD vec3 scale3(vec3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
RT_HD vec3 div3(vec3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
RT_HD vec3 neg3(vec3 a) { return mk3(-a.x, -a.y, -a.z); }

RT_HD uint32_t float_bits(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
RT_HD float bits_float(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

// GLSL 4.60 §8.3: min(x,y) = y < x ? y : x;  max(x,y) = x < y ? y : x;  clamp = min(max(x,lo),hi)
RT_HD float glsl_min(float x, float y) { return (y < x) ? y : x; }
RT_HD float glsl_max(float x, float y) { return (x < y) ? y : x; }
RT_HD float glsl_clamp(float x, float lo, float hi) { return glsl_min(glsl_max(x, lo), hi); }

// a / b, same bits as the plain quotient.  A zero numerator sends the compiler's division sequence down its out-of-line
// slow path (~35 instructions), and a metal's Lambert lobe is exactly zero, so hit shading took that path nine times per
// hit on Sponza's all-metal materials.  0 / b is a zero signed sign(a) ^ sign(b) for every finite non-zero b: that case
// is answered directly and the division runs on a harmless numerator.
RT_HD float div_zn(float a, float b) {
    const bool z = (a == 0.0f) && (fabsf(b) > 0.0f) && (fabsf(b) < INFINITY);
    const float q = (z ? 1.0f : a) / b;
    return z ? bits_float((float_bits(a) ^ float_bits(b)) & 0x80000000u) : q;
}
RT_HD vec3 div3z(vec3 a, float s) { return mk3(div_zn(a.x, s), div_zn(a.y, s), div_zn(a.z, s)); }

// dot = (x x' + y y') + z z'
RT_HD float dot3(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
RT_HD vec3 cross3(vec3 a, vec3 b) {
    return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
RT_HD float length3(vec3 a) { return sqrtf(dot3(a, a)); }
// normalize(v) = v * (1 / |v|); the zero vector yields NaN
RT_HD vec3 normalize3(vec3 a) { return scale3(a, 1.0f / length3(a)); }
// mix(a,b,t) = a (1-t) + b t
RT_HD float mix1(float a, float b, float t) { return a * (1.0f - t) + b * t; }
RT_HD vec3 mix3(vec3 a, vec3 b, float t) {
    float s = 1.0f - t;
    return mk3(a.x * s + b.x * t, a.y * s + b.y * t, a.z * s + b.z * t);
}
RT_HD float smoothstep1(float e0, float e1, float x) {
    float t = glsl_clamp((x - e0) / (e1 - e0), 0.0f, 1.0f);
    return (t * t) * (3.0f - 2.0f * t);
}
// reflect(I,N) = I - N (2 dot(N,I))
RT_HD vec3 reflect3(vec3 I, vec3 N) { return sub3(I, scale3(N, 2.0f * dot3(N, I))); }
// pow(x, 5.0) as ((x^2)^2) x — also defines the negative-base case of simple.rchit:97
RT_HD float pow5f(float x) {
    float x2 = x * x;
    return (x2 * x2) * x;
}
RT_HD float sign1(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }
RT_HD float max3c(vec3 v) { return glsl_max(v.x, glsl_max(v.y, v.z)); }  // math.glsl:39-41

#define RT_PI 3.14159265359f      /* math.glsl:1 */
#define RT_TWO_PI 6.28318530718f  /* 2.0 * M_PI */

// sin & cos of phi >= 0: three-step Cody-Waite reduction by pi/2 (fused), then the Cephes
// single-precision minimax polynomials on [-pi/4, pi/4] in fused Horner form.
RT_HD void sincos_rt(float a, float* s_out, float* c_out) {
    int q = (int)(a * 0.636619772f + 0.5f);
    float k = (float)q;
    float r = fmaf(k, -1.5703125f, a);
    r = fmaf(k, -4.837512969970703125e-4f, r);
    r = fmaf(k, -7.54978995489188e-8f, r);
    float z = r * r;
    float sp = fmaf(z, -1.9515295891e-4f, 8.3321608736e-3f);
    sp = fmaf(z, sp, -1.6666654611e-1f);
    float sv = fmaf(r * z, sp, r);
    float cp = fmaf(z, 2.443315711809948e-5f, -1.388731625493765e-3f);
    cp = fmaf(z, cp, 4.166664568298827e-2f);
    float cv = fmaf(z * z, cp, fmaf(-0.5f, z, 1.0f));
    int m = q & 3;
    float ss = (m & 1) ? cv : sv;
    float cc = (m & 1) ? sv : cv;
    *s_out = (m & 2) ? -ss : ss;
    *c_out = (m == 1 || m == 2) ? -cc : cc;
}

// ---- shaders/random.glsl:23-53 ----------------------------------------------------------------
RT_HD uint32_t tea16(uint32_t a, uint32_t b) {
    uint32_t v0 = a, v1 = b, sum = 0u;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}
RT_HD float rnd01(uint32_t& state) {
    state = 1664525u * state + 1013904223u;
    return (float)(state & 0x00FFFFFFu) / 16777216.0f;
}

// ---- column-major mat4 (m[c*4+r]) -------------------------------------------------------------
RT_HD vec3 mat_point(const float* m, vec3 p) {
    return mk3(((m[0] * p.x + m[4] * p.y) + m[8] * p.z) + m[12], ((m[1] * p.x + m[5] * p.y) + m[9] * p.z) + m[13],
               ((m[2] * p.x + m[6] * p.y) + m[10] * p.z) + m[14]);
}
RT_HD vec3 mat_dir(const float* m, vec3 d) {
    return mk3((m[0] * d.x + m[4] * d.y) + m[8] * d.z, (m[1] * d.x + m[5] * d.y) + m[9] * d.z,
               (m[2] * d.x + m[6] * d.y) + m[10] * d.z);
}
// transpose(inverse(mat3(M))): cofactor columns / determinant
RT_HD void cofactor_normal_matrix(const float* m, float* n9) {
    vec3 a0 = mk3(m[0], m[1], m[2]), a1 = mk3(m[4], m[5], m[6]), a2 = mk3(m[8], m[9], m[10]);
    vec3 c0 = cross3(a1, a2), c1 = cross3(a2, a0), c2 = cross3(a0, a1);
    float inv = 1.0f / dot3(a0, c0);
    c0 = scale3(c0, inv); c1 = scale3(c1, inv); c2 = scale3(c2, inv);
    n9[0] = c0.x; n9[1] = c0.y; n9[2] = c0.z;
    n9[3] = c1.x; n9[4] = c1.y; n9[5] = c1.z;
    n9[6] = c2.x; n9[7] = c2.y; n9[8] = c2.z;
}
RT_HD vec3 mat3_mul(const float* n9, vec3 v) {
    return mk3((n9[0] * v.x + n9[3] * v.y) + n9[6] * v.z, (n9[1] * v.x + n9[4] * v.y) + n9[7] * v.z,
               (n9[2] * v.x + n9[5] * v.y) + n9[8] * v.z);
}

RT_HD bool any_nan3(vec3 v) { return (v.x != v.x) || (v.y != v.y) || (v.z != v.z); }
RT_HD bool is_inf1(float f) { return (float_bits(f) & 0x7FFFFFFFu) == 0x7F800000u; }
RT_HD bool any_inf3(vec3 v) { return is_inf1(v.x) || is_inf1(v.y) || is_inf1(v.z); }

// float -> half (round to nearest even), float -> unorm8: imageStore conversions of restir.rgen
RT_HD uint16_t half_bits_rne(float f) {
    uint32_t x = float_bits(f);
    uint32_t sgn = (x >> 16) & 0x8000u;
    uint32_t ax = x & 0x7FFFFFFFu;
    if (ax > 0x7F800000u) return (uint16_t)0x7E00u;  // NaN: canonical quiet NaN (sign and payload are not portable)
    if (ax == 0x7F800000u) return (uint16_t)(sgn | 0x7C00u);
    if (ax >= 0x477FF000u) return (uint16_t)(sgn | 0x7C00u);
    if (ax < 0x33000001u) return (uint16_t)sgn;
    int e = (int)(ax >> 23) - 127;
    uint32_t man = (ax & 0x7FFFFFu) | 0x800000u;
    int shift = 13;
    uint32_t base = 0u;
    if (e < -14) shift = 13 + (-14 - e);
    else { base = (uint32_t)(e + 15) << 10; man &= 0x7FFFFFu; }
    uint32_t q = man >> shift;
    uint32_t rem = man & ((1u << shift) - 1u);
    uint32_t halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (q & 1u))) q++;
    return (uint16_t)(sgn | (base + q));
}
RT_HD uint8_t unorm8_rne(float f) {
    float c = glsl_clamp(f, 0.0f, 1.0f);
    if (c != c) c = 0.0f;
    return (uint8_t)(int)rintf(c * 255.0f);
}

}  // namespace rt
