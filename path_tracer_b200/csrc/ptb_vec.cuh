// Device vector kit of the B200 path-tracing integrator.
//
// Numerical contract: every operator is one IEEE binary32
// operation with the association order written here; the translation units are
// compiled with --fmad=false so nothing is contracted behind our back, and the
// places that WANT a fused multiply-add (wide-BVH slab tests) say so with
// fmaf().  GLSL built-ins follow the GLSL 4.50 §8 formulas the reference's
// shaders are specified against (src/core/common.glsl.inc uses them throughout).
//
// The functions are __host__ __device__ so that tests/host_emu can compile the
// per-thread device logic with g++ for debugging in GPU-less containers.  That
// build is test infrastructure; libptb.so contains no CPU execution path.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define PTB_HD __host__ __device__ __forceinline__
#define PTB_D __device__ __forceinline__
// Large helpers with many call sites.  Fully inlined, k_scatter is ~300 KB of SASS and stalls on
// instruction fetch (profiles/r01_v3b_ncu_summary.txt, no_instruction); making them real functions
// (-DPTB_NOINLINE_HELPERS) was measured and is worse (ABI spills: scatter 0.30 -> 0.43 ms,
// profiles/r01_v3_ab_builds.txt), so the code size is cut by scene-feature specialisation instead.
#if defined(PTB_NOINLINE_HELPERS)
#define PTB_HDN static __host__ __device__ __noinline__
#else
#define PTB_HDN PTB_HD
#endif
#else
#define PTB_HD inline
#define PTB_D inline
#define PTB_HDN inline
#endif

#if defined(__CUDA_ARCH__)
#define PTB_LDG(p) __ldg(p)
#else
#define PTB_LDG(p) (*(p))
#endif

namespace ptb {

typedef uint32_t u32;

struct f2 { float x, y; };
struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };

PTB_HD f2 mk2(float x, float y) { f2 r; r.x = x; r.y = y; return r; }
PTB_HD f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
PTB_HD f3 mk3(float s) { return mk3(s, s, s); }
PTB_HD f4 mk4(float x, float y, float z, float w) { f4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
PTB_HD f4 mk4(float s) { return mk4(s, s, s, s); }
PTB_HD f4 mk4(f3 v, float w) { return mk4(v.x, v.y, v.z, w); }
PTB_HD f3 xyz(f4 v) { return mk3(v.x, v.y, v.z); }

// ---- f2 ----
PTB_HD f2 operator+(f2 a, f2 b) { return mk2(a.x + b.x, a.y + b.y); }
PTB_HD f2 operator*(f2 a, f2 b) { return mk2(a.x * b.x, a.y * b.y); }
PTB_HD f2 operator*(f2 a, float s) { return mk2(a.x * s, a.y * s); }
PTB_HD f2 operator/(f2 a, f2 b) { return mk2(a.x / b.x, a.y / b.y); }
PTB_HD float dot(f2 a, f2 b) { return a.x * b.x + a.y * b.y; }

// ---- f3 ----
PTB_HD f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
PTB_HD f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
PTB_HD f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
PTB_HD f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
PTB_HD f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
PTB_HD f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
PTB_HD f3 operator/(f3 a, f3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
// vector / scalar = vector * (1 / scalar), correctly rounded reciprocal: the same pinned reading as oracle/glsl_shim.h
PTB_HD f3 operator/(f3 a, float s) { const float r = 1.0f / s; return mk3(a.x * r, a.y * r, a.z * r); }
PTB_HD float dot(f3 a, f3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
PTB_HD f3 cross(f3 a, f3 b) { return mk3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
PTB_HD f3 normalize(f3 a) { return a / sqrtf(dot(a, a)); }
PTB_HD f3 abs3(f3 a) { return mk3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
PTB_HD f3 min3(f3 a, f3 b) { return mk3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
PTB_HD f3 max3(f3 a, f3 b) { return mk3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }

// ---- f4 ----
PTB_HD f4 operator+(f4 a, f4 b) { return mk4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
PTB_HD f4 operator-(f4 a, f4 b) { return mk4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
PTB_HD f4 operator-(f4 a) { return mk4(-a.x, -a.y, -a.z, -a.w); }
PTB_HD f4 operator*(f4 a, f4 b) { return mk4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
PTB_HD f4 operator*(f4 a, float s) { return mk4(a.x * s, a.y * s, a.z * s, a.w * s); }
PTB_HD f4 operator*(float s, f4 a) { return mk4(s * a.x, s * a.y, s * a.z, s * a.w); }
PTB_HD f4 operator/(f4 a, f4 b) { return mk4(a.x / b.x, a.y / b.y, a.z / b.z, a.w / b.w); }
PTB_HD f4 operator/(f4 a, float s) { const float r = 1.0f / s; return mk4(a.x * r, a.y * r, a.z * r, a.w * r); }
PTB_HD f4 operator+(float s, f4 a) { return mk4(s + a.x, s + a.y, s + a.z, s + a.w); }
PTB_HD f4 operator+(f4 a, float s) { return mk4(a.x + s, a.y + s, a.z + s, a.w + s); }
PTB_HD f4 operator-(float s, f4 a) { return mk4(s - a.x, s - a.y, s - a.z, s - a.w); }
PTB_HD f4 operator-(f4 a, float s) { return mk4(a.x - s, a.y - s, a.z - s, a.w - s); }
PTB_HD float dot(f4 a, f4 b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }
PTB_HD f4 abs4(f4 a) { return mk4(fabsf(a.x), fabsf(a.y), fabsf(a.z), fabsf(a.w)); }
PTB_HD f4 sqrt4(f4 a) { return mk4(sqrtf(a.x), sqrtf(a.y), sqrtf(a.z), sqrtf(a.w)); }
PTB_HD f4 exp4(f4 a) { return mk4(expf(a.x), expf(a.y), expf(a.z), expf(a.w)); }
PTB_HD f4 log4(f4 a) { return mk4(logf(a.x), logf(a.y), logf(a.z), logf(a.w)); }
PTB_HD f4 max4s(f4 a, float s) { return mk4(fmaxf(a.x, s), fmaxf(a.y, s), fmaxf(a.z, s), fmaxf(a.w, s)); }
PTB_HD float hmax4(f4 v) { return fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)); }   // common.glsl.inc:107-110

// ---- scalar GLSL built-ins ----
PTB_HD float gsign(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
PTB_HD f4 gsign(f4 a) { return mk4(gsign(a.x), gsign(a.y), gsign(a.z), gsign(a.w)); }
PTB_HD float gfract(float x) { return x - floorf(x); }
PTB_HD float gmix(float x, float y, float a) { return x * (1.0f - a) + y * a; }
PTB_HD float gclamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

PTB_HD float u2f(u32 u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
PTB_HD u32 f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    u32 u; memcpy(&u, &f, 4); return u;
#endif
}

// Column-major 4x4 (std430 mat4) addressed as 16 floats, m[4*col + row].
// GLSL `M * vec4(v, w)` = sum of columns weighted left to right.
PTB_HD f3 mat_mul_point(const float* m, f3 p) {       // (M * vec4(p,1)).xyz
    return mk3(((m[0] * p.x + m[4] * p.y) + m[8] * p.z) + m[12] * 1.0f,
               ((m[1] * p.x + m[5] * p.y) + m[9] * p.z) + m[13] * 1.0f,
               ((m[2] * p.x + m[6] * p.y) + m[10] * p.z) + m[14] * 1.0f);
}
PTB_HD f3 mat_mul_vector(const float* m, f3 v) {      // (M * vec4(v,0)).xyz
    return mk3(((m[0] * v.x + m[4] * v.y) + m[8] * v.z) + m[12] * 0.0f,
               ((m[1] * v.x + m[5] * v.y) + m[9] * v.z) + m[13] * 0.0f,
               ((m[2] * v.x + m[6] * v.y) + m[10] * v.z) + m[14] * 0.0f);
}
PTB_HD f3 vector_mul_mat(f3 v, const float* m) {      // (vec4(v,0) * M).xyz: component j = dot(v4, column j)
    return mk3(((v.x * m[0] + v.y * m[1]) + v.z * m[2]) + 0.0f * m[3],
               ((v.x * m[4] + v.y * m[5]) + v.z * m[6]) + 0.0f * m[7],
               ((v.x * m[8] + v.y * m[9]) + v.z * m[10]) + 0.0f * m[11]);
}

// ---- pack / unpack (GLSL 4.50 §8.4) ----
PTB_HD u32 pack_snorm2x16(f2 v) {
    int x = (int)roundf(gclamp(v.x, -1.0f, 1.0f) * 32767.0f);
    int y = (int)roundf(gclamp(v.y, -1.0f, 1.0f) * 32767.0f);
    return ((u32)x & 0xFFFFu) | (((u32)y & 0xFFFFu) << 16);
}
// x / 32767.0f, correctly rounded, for x the value of an int16: q = RN(x * RN(1/32767)) and one correction step with
// the exact residual (two fused multiply-adds) instead of the general IEEE division sequence.  Equal to the division
// for all 65536 inputs (tests/test_host_layer.py checks every one through the host build of this header).
PTB_HD float snorm16_to_float(float x) {
    const float r = 0x1.0002p-15f;
    const float q = x * r;
    return fmaf(fmaf(-q, 32767.0f, x), r, q);
}
PTB_HD f2 unpack_snorm2x16(u32 p) {
    float x = (float)(int16_t)(p & 0xFFFFu);
    float y = (float)(int16_t)(p >> 16);
    return mk2(gclamp(snorm16_to_float(x), -1.0f, 1.0f), gclamp(snorm16_to_float(y), -1.0f, 1.0f));
}

// binary16 -> binary32, exact (unpackHalf2x16).
PTB_HD float half_bits_to_float(u32 h) {
    u32 s = (h & 0x8000u) << 16;
    u32 e = (h >> 10) & 0x1Fu;
    u32 m = h & 0x3FFu;
    if (e == 0) {
        if (m == 0) return u2f(s);
        float f = (float)m * 5.9604644775390625e-8f;    // m * 2^-24
        return u2f(s | f2u(f));
    }
    if (e == 31) return u2f(s | 0x7F800000u | (m << 13));
    return u2f(s | ((e + 112u) << 23) | (m << 13));
}
PTB_HD f2 unpack_half2x16(u32 p) { return mk2(half_bits_to_float(p & 0xFFFFu), half_bits_to_float(p >> 16)); }

// Octahedral unit vectors, snorm16x2 (common.glsl.inc:127-151).
PTB_HD f2 sign_not_zero(f2 v) { return mk2(v.x >= 0.0f ? 1.0f : -1.0f, v.y >= 0.0f ? 1.0f : -1.0f); }
PTB_HD u32 pack_unit_vector(f3 v) {
    float s = 1.0f / ((fabsf(v.x) + fabsf(v.y)) + fabsf(v.z));
    f2 p = mk2(v.x * s, v.y * s);
    if (v.z <= 0.0f) p = mk2(1.0f - fabsf(p.y), 1.0f - fabsf(p.x)) * sign_not_zero(p);
    return pack_snorm2x16(p);
}
PTB_HD f3 unpack_unit_vector(u32 packed) {
    f2 p = unpack_snorm2x16(packed);
    float z = (1.0f - fabsf(p.x)) - fabsf(p.y);
    if (z < 0.0f) p = mk2(1.0f - fabsf(p.y), 1.0f - fabsf(p.x)) * sign_not_zero(p);
    return normalize(mk3(p.x, p.y, z));
}

// common.glsl.inc:93-100
PTB_HD f3 safe_normalize(f3 v) {
    float l2 = dot(v, v);
    if (l2 < 1e-12f) return mk3(0, 0, 1);
    return v / sqrtf(l2);
}

}  // namespace ptb
