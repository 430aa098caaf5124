// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product; nothing under
// path_tracer_b200/ may include, link or call this.  See oracle/README.md.
//
// Minimal GLSL-4.50-flavoured vector shim for restating the reference's
// shaders as plain host C++.  Every arithmetic operator here is a single
// IEEE-754 binary32 operation with a fixed association order; the oracle is
// compiled with -ffp-contract=off so no multiply-add is fused.  The CUDA
// product pins the same association order (see DESIGN.md "numerical contract").
//
// Built-ins follow the GLSL 4.50 spec formulas (section 8.x):
//   mix(x,y,a) = x*(1-a) + y*a          fract(x) = x - floor(x)
//   sign(x)    = 1, 0, -1               clamp(x,lo,hi) = min(max(x,lo),hi)
//   normalize(v) = v / sqrt(dot(v,v))   (pinned; GLSL allows inversesqrt)
//   vector / scalar = vector * (1 / scalar) with a correctly rounded reciprocal (pinned; one division instead of
//   three or four — GLSL only asks 2.5 ULP of a division, so both readings are admissible; the CUDA kernels do the same)
//   min/max: pinned to C fminf/fmaxf (NaN-dropping), reference quirk Q19.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

typedef uint32_t uint;

struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };

inline vec2 V2(float x, float y) { return vec2{x, y}; }
inline vec3 V3(float x, float y, float z) { return vec3{x, y, z}; }
inline vec3 V3(float s) { return vec3{s, s, s}; }
inline vec4 V4(float x, float y, float z, float w) { return vec4{x, y, z, w}; }
inline vec4 V4(float s) { return vec4{s, s, s, s}; }
inline vec4 V4(vec3 v, float w) { return vec4{v.x, v.y, v.z, w}; }
inline vec3 xyz(vec4 v) { return vec3{v.x, v.y, v.z}; }

// ---- vec2 ----
inline vec2 operator+(vec2 a, vec2 b) { return {a.x + b.x, a.y + b.y}; }
inline vec2 operator-(vec2 a, vec2 b) { return {a.x - b.x, a.y - b.y}; }
inline vec2 operator*(vec2 a, vec2 b) { return {a.x * b.x, a.y * b.y}; }
inline vec2 operator*(vec2 a, float s) { return {a.x * s, a.y * s}; }
inline vec2 operator*(float s, vec2 a) { return {s * a.x, s * a.y}; }
inline vec2 operator/(vec2 a, vec2 b) { return {a.x / b.x, a.y / b.y}; }
inline vec2 operator/(float s, vec2 a) { return {s / a.x, s / a.y}; }
inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }

// ---- vec3 ----
inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec3 operator/(vec3 a, float s) { float r = 1.0f / s; return {a.x * r, a.y * r, a.z * r}; }   // see header: vector / scalar
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// GLSL cross(): (x1*y2 - y1*x2, ...) per spec.
inline vec3 cross(vec3 a, vec3 b) {
    return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
inline float length(vec3 a) { return sqrtf(dot(a, a)); }
inline vec3 normalize(vec3 a) { return a / sqrtf(dot(a, a)); }
inline vec3 vabs(vec3 a) { return {fabsf(a.x), fabsf(a.y), fabsf(a.z)}; }
inline vec3 vmin(vec3 a, vec3 b) { return {fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)}; }
inline vec3 vmax(vec3 a, vec3 b) { return {fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)}; }

// ---- vec4 ----
inline vec4 operator+(vec4 a, vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline vec4 operator-(vec4 a, vec4 b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
inline vec4 operator-(vec4 a) { return {-a.x, -a.y, -a.z, -a.w}; }
inline vec4 operator*(vec4 a, vec4 b) { return {a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }
inline vec4 operator*(vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline vec4 operator*(float s, vec4 a) { return {s * a.x, s * a.y, s * a.z, s * a.w}; }
inline vec4 operator/(vec4 a, vec4 b) { return {a.x / b.x, a.y / b.y, a.z / b.z, a.w / b.w}; }
inline vec4 operator/(vec4 a, float s) { float r = 1.0f / s; return {a.x * r, a.y * r, a.z * r, a.w * r}; }
inline vec4 operator/(float s, vec4 a) { return {s / a.x, s / a.y, s / a.z, s / a.w}; }
inline vec4 operator+(vec4 a, float s) { return {a.x + s, a.y + s, a.z + s, a.w + s}; }
inline vec4 operator+(float s, vec4 a) { return {s + a.x, s + a.y, s + a.z, s + a.w}; }
inline vec4 operator-(float s, vec4 a) { return {s - a.x, s - a.y, s - a.z, s - a.w}; }
inline vec4 operator-(vec4 a, float s) { return {a.x - s, a.y - s, a.z - s, a.w - s}; }
inline vec4& operator*=(vec4& a, vec4 b) { a = a * b; return a; }
inline vec4& operator*=(vec4& a, float s) { a = a * s; return a; }
inline vec4& operator/=(vec4& a, float s) { a = a / s; return a; }
inline float dot(vec4 a, vec4 b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }
inline vec4 vabs(vec4 a) { return {fabsf(a.x), fabsf(a.y), fabsf(a.z), fabsf(a.w)}; }
inline vec4 vsqrt(vec4 a) { return {sqrtf(a.x), sqrtf(a.y), sqrtf(a.z), sqrtf(a.w)}; }
inline vec4 vexp(vec4 a) { return {expf(a.x), expf(a.y), expf(a.z), expf(a.w)}; }
inline vec4 vlog(vec4 a) { return {logf(a.x), logf(a.y), logf(a.z), logf(a.w)}; }
inline vec4 vmax(vec4 a, float s) { return {fmaxf(a.x, s), fmaxf(a.y, s), fmaxf(a.z, s), fmaxf(a.w, s)}; }

// ---- scalar built-ins ----
inline float gl_sign(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
inline vec4 gl_sign(vec4 a) { return {gl_sign(a.x), gl_sign(a.y), gl_sign(a.z), gl_sign(a.w)}; }
inline float gl_fract(float x) { return x - floorf(x); }
inline float gl_mix(float x, float y, float a) { return x * (1.0f - a) + y * a; }
inline float gl_clamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
inline int gl_clampi(int x, int lo, int hi) { return x < lo ? lo : (x > hi ? hi : x); }

inline float uintBitsToFloat(uint u) { float f; memcpy(&f, &u, 4); return f; }
inline uint floatBitsToUint(float f) { uint u; memcpy(&u, &f, 4); return u; }

// Column-major 4x4 matrix, c[j] = column j (std430 mat4).
struct mat4 { vec4 c[4]; };

// GLSL  M * v  : linear combination of columns, summed left to right.
inline vec4 mul(const mat4& m, vec4 v) {
    return ((m.c[0] * v.x + m.c[1] * v.y) + m.c[2] * v.z) + m.c[3] * v.w;
}
// GLSL  v * M  : component j = dot(v, column j).
inline vec4 mul(vec4 v, const mat4& m) {
    return {dot(v, m.c[0]), dot(v, m.c[1]), dot(v, m.c[2]), dot(v, m.c[3])};
}

// ---- pack/unpack built-ins (GLSL 4.50 section 8.4) ----
// packSnorm2x16: round(clamp(c,-1,+1) * 32767.0).  Rounding of exact .5 ties is
// implementation-defined in GLSL; pinned to round-half-away (glm::round, roundf).
inline uint packSnorm2x16(vec2 v) {
    int x = (int)roundf(gl_clamp(v.x, -1.0f, 1.0f) * 32767.0f);
    int y = (int)roundf(gl_clamp(v.y, -1.0f, 1.0f) * 32767.0f);
    return ((uint)(uint16_t)(int16_t)x) | (((uint)(uint16_t)(int16_t)y) << 16);
}
inline vec2 unpackSnorm2x16(uint p) {
    float x = (float)(int16_t)(p & 0xFFFFu);
    float y = (float)(int16_t)(p >> 16);
    return {gl_clamp(x / 32767.0f, -1.0f, 1.0f), gl_clamp(y / 32767.0f, -1.0f, 1.0f)};
}

// IEEE binary16 -> binary32 (exact).
inline float half_to_float(uint16_t h) {
    uint s = (uint)(h >> 15) << 31;
    uint e = (h >> 10) & 0x1Fu;
    uint m = h & 0x3FFu;
    uint bits;
    if (e == 0) {
        if (m == 0) bits = s;
        else {
            // subnormal: value = m * 2^-24
            float f = (float)m * 5.9604644775390625e-8f;
            bits = s | floatBitsToUint(f);
        }
    } else if (e == 31) {
        bits = s | 0x7F800000u | (m << 13);
    } else {
        bits = s | ((e + 112u) << 23) | (m << 13);
    }
    return uintBitsToFloat(bits);
}
// binary32 -> binary16, round-to-nearest-even.
inline uint16_t float_to_half(float f) {
    uint x = floatBitsToUint(f);
    uint s = (x >> 16) & 0x8000u;
    uint ax = x & 0x7FFFFFFFu;
    if (ax >= 0x7F800000u) return (uint16_t)(s | 0x7C00u | ((ax > 0x7F800000u) ? 0x200u : 0u));
    if (ax >= 0x477FF000u) return (uint16_t)(s | 0x7C00u);          // overflow -> inf (after rounding)
    if (ax < 0x33000001u) return (uint16_t)s;                       // underflow -> 0  (<= 2^-25)
    int e = (int)(ax >> 23) - 127;
    uint m = (ax & 0x7FFFFFu) | 0x800000u;
    int shift;
    uint he;
    if (e < -14) { shift = 13 + (-14 - e); he = 0; }
    else { shift = 13; he = (uint)(e + 15); }
    uint hm = m >> shift;
    uint rem = m & ((1u << shift) - 1u);
    uint halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (hm & 1u))) hm++;
    uint h;
    if (he == 0) h = hm;                    // subnormal (hm may carry into exponent bit: correct)
    else h = ((he << 10) + (hm - 0x400u)) ; // normal: remove implicit bit; carry bumps exponent
    return (uint16_t)(s | h);
}
inline vec2 unpackHalf2x16(uint p) {
    return {half_to_float((uint16_t)(p & 0xFFFFu)), half_to_float((uint16_t)(p >> 16))};
}
inline uint packHalf2x16(vec2 v) {
    return (uint)float_to_half(v.x) | ((uint)float_to_half(v.y) << 16);
}

}  // namespace orc
