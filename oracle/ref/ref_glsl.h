// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// GLSL 4.50 vocabulary for compiling the REFERENCE'S OWN shader sources as C++
// (oracle/ref/transpile.py rewrites the few things that are not also C++ and writes
// the result under oracle/_ref/gen/; nothing of the reference enters the repository).
// This file is the "driver" side of that build: vector and matrix types with GLSL
// constructors and the swizzles the shaders use, the built-in functions they call,
// samplers and images, and the compute-shader built-in variables.
//
// Every operator is one IEEE binary32 operation in the association order that
// oracle/glsl_shim.h pins for the restatement (same latitude choices: vector / scalar as
// multiplication by the correctly rounded reciprocal, min/max = fminf/fmaxf, round-half-
// away in packSnorm2x16, fp32 bilinear weights), so the two can be compared bit for bit.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#undef INFINITY                 // the shaders define their own (common.glsl.inc:4)

namespace glsl {

typedef uint32_t uint;

struct ivec2;
struct vec2 {
    union { float x; float r; };
    union { float y; float g; };
    vec2() : x(0), y(0) {}
    inline vec2(const ivec2& v);                 // GLSL converts ivec2 -> vec2 implicitly
    explicit vec2(float s) : x(s), y(s) {}
    vec2(float x_, float y_) : x(x_), y(y_) {}
    vec2 xy() const { return *this; }
    vec2 yx() const { return vec2(y, x); }
};
struct vec3 {
    union { float x; float r; };
    union { float y; float g; };
    union { float z; float b; };
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    vec3(vec2 v, float z_) : x(v.x), y(v.y), z(z_) {}
    vec3 xyz() const { return *this; }
    vec3 rgb() const { return *this; }
    vec2 xy() const { return vec2(x, y); }
    vec2 yx() const { return vec2(y, x); }
    vec2 yz() const { return vec2(y, z); }
    vec2 xz() const { return vec2(x, z); }
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};
struct vec4 {
    union { float x; float r; };
    union { float y; float g; };
    union { float z; float b; };
    union { float w; float a; };
    vec4() : x(0), y(0), z(0), w(0) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
    vec4(vec3 v, float w_) : x(v.x), y(v.y), z(v.z), w(w_) {}
    vec4(vec2 a, vec2 b) : x(a.x), y(a.y), z(b.x), w(b.y) {}
    vec3 xyz() const { return vec3(x, y, z); }
    vec3 rgb() const { return vec3(x, y, z); }
    vec2 xy() const { return vec2(x, y); }
    vec2 yx() const { return vec2(y, x); }
    vec2 yz() const { return vec2(y, z); }
    vec2 xz() const { return vec2(x, z); }
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};
struct uvec2 { uint x, y; uvec2() : x(0), y(0) {} uvec2(uint a, uint b) : x(a), y(b) {} inline uvec2(const ivec2& v); };
struct uvec3 {
    uint x, y, z;
    uvec3() : x(0), y(0), z(0) {}
    uvec3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
    uvec2 xy() const { return uvec2(x, y); }
};
struct ivec2 {
    int x, y;
    ivec2() : x(0), y(0) {}
    ivec2(int a, int b) : x(a), y(b) {}
    explicit ivec2(uvec2 v) : x((int)v.x), y((int)v.y) {}
};
inline uvec2::uvec2(const ivec2& v) : x((uint)v.x), y((uint)v.y) {}     // GLSL converts ivec2 -> uvec2 implicitly
inline vec2::vec2(const ivec2& v) : x((float)v.x), y((float)v.y) {}

// ---- vec2 ----
inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(vec2 a, vec2 b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator-(vec2 a) { return vec2(-a.x, -a.y); }
inline vec2 operator*(vec2 a, vec2 b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator*(vec2 a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator*(float s, vec2 a) { return vec2(s * a.x, s * a.y); }
inline vec2 operator/(vec2 a, vec2 b) { return vec2(a.x / b.x, a.y / b.y); }
inline vec2 operator/(vec2 a, float s) { float r = 1.0f / s; return vec2(a.x * r, a.y * r); }
inline vec2 operator/(float s, vec2 a) { return vec2(s / a.x, s / a.y); }
inline vec2 operator+(vec2 a, float s) { return vec2(a.x + s, a.y + s); }
inline vec2 operator-(vec2 a, float s) { return vec2(a.x - s, a.y - s); }
inline vec2 operator+(float s, vec2 a) { return vec2(s + a.x, s + a.y); }
inline vec2 operator-(float s, vec2 a) { return vec2(s - a.x, s - a.y); }
inline vec2& operator+=(vec2& a, vec2 b) { a = a + b; return a; }
inline vec2& operator-=(vec2& a, vec2 b) { a = a - b; return a; }
inline vec2& operator*=(vec2& a, float s) { a = a * s; return a; }
inline vec2& operator*=(vec2& a, vec2 b) { a = a * b; return a; }
inline vec2& operator/=(vec2& a, float s) { a = a / s; return a; }
inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
inline float length(vec2 a) { return sqrtf(dot(a, a)); }

// ---- vec3 ----
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator/(vec3 a, float s) { float r = 1.0f / s; return vec3(a.x * r, a.y * r, a.z * r); }
inline vec3 operator/(float s, vec3 a) { return vec3(s / a.x, s / a.y, s / a.z); }
inline vec3 operator+(vec3 a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3 operator-(vec3 a, float s) { return vec3(a.x - s, a.y - s, a.z - s); }
inline vec3 operator+(float s, vec3 a) { return vec3(s + a.x, s + a.y, s + a.z); }
inline vec3 operator-(float s, vec3 a) { return vec3(s - a.x, s - a.y, s - a.z); }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
inline vec3& operator-=(vec3& a, vec3 b) { a = a - b; return a; }
inline vec3& operator*=(vec3& a, float s) { a = a * s; return a; }
inline vec3& operator*=(vec3& a, vec3 b) { a = a * b; return a; }
inline vec3& operator/=(vec3& a, float s) { a = a / s; return a; }
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) { return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
inline float length(vec3 a) { return sqrtf(dot(a, a)); }
inline vec3 normalize(vec3 a) { return a / sqrtf(dot(a, a)); }

// ---- vec4 ----
inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(vec4 a, vec4 b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline vec4 operator-(vec4 a) { return vec4(-a.x, -a.y, -a.z, -a.w); }
inline vec4 operator*(vec4 a, vec4 b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator*(float s, vec4 a) { return vec4(s * a.x, s * a.y, s * a.z, s * a.w); }
inline vec4 operator/(vec4 a, vec4 b) { return vec4(a.x / b.x, a.y / b.y, a.z / b.z, a.w / b.w); }
inline vec4 operator/(vec4 a, float s) { float r = 1.0f / s; return vec4(a.x * r, a.y * r, a.z * r, a.w * r); }
inline vec4 operator/(float s, vec4 a) { return vec4(s / a.x, s / a.y, s / a.z, s / a.w); }
inline vec4 operator+(vec4 a, float s) { return vec4(a.x + s, a.y + s, a.z + s, a.w + s); }
inline vec4 operator+(float s, vec4 a) { return vec4(s + a.x, s + a.y, s + a.z, s + a.w); }
inline vec4 operator-(vec4 a, float s) { return vec4(a.x - s, a.y - s, a.z - s, a.w - s); }
inline vec4 operator-(float s, vec4 a) { return vec4(s - a.x, s - a.y, s - a.z, s - a.w); }
inline vec4& operator+=(vec4& a, vec4 b) { a = a + b; return a; }
inline vec4& operator-=(vec4& a, vec4 b) { a = a - b; return a; }
inline vec4& operator*=(vec4& a, vec4 b) { a = a * b; return a; }
inline vec4& operator*=(vec4& a, float s) { a = a * s; return a; }
inline vec4& operator/=(vec4& a, float s) { a = a / s; return a; }
inline vec4& operator/=(vec4& a, vec4 b) { a = a / b; return a; }
inline float dot(vec4 a, vec4 b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }
inline float length(vec4 a) { return sqrtf(dot(a, a)); }

// ---- matrices: column-major, c[j] = column j (std430 mat4, 16-byte aligned) ----
struct alignas(16) mat4 {
    vec4 c[4];
    vec4& operator[](int j) { return c[j]; }
    const vec4& operator[](int j) const { return c[j]; }
};
inline vec4 operator*(const mat4& m, vec4 v) { return ((m.c[0] * v.x + m.c[1] * v.y) + m.c[2] * v.z) + m.c[3] * v.w; }
inline vec4 operator*(vec4 v, const mat4& m) { return vec4(dot(v, m.c[0]), dot(v, m.c[1]), dot(v, m.c[2]), dot(v, m.c[3])); }
struct mat3 {
    vec3 c[3];
    mat3() {}
    mat3(vec3 a, vec3 b, vec3 d) { c[0] = a; c[1] = b; c[2] = d; }
    mat3(float a0, float a1, float a2, float b0, float b1, float b2, float d0, float d1, float d2) { c[0] = vec3(a0, a1, a2); c[1] = vec3(b0, b1, b2); c[2] = vec3(d0, d1, d2); }
    vec3& operator[](int j) { return c[j]; }
    const vec3& operator[](int j) const { return c[j]; }
};
inline vec3 operator*(const mat3& m, vec3 v) { return (m.c[0] * v.x + m.c[1] * v.y) + m.c[2] * v.z; }
inline vec3 operator*(vec3 v, const mat3& m) { return vec3(dot(v, m.c[0]), dot(v, m.c[1]), dot(v, m.c[2])); }

struct mat4x3 {                                  // 4 columns of 3 rows
    vec3 c[4];
    mat4x3() {}
    mat4x3(vec3 a, vec3 b, vec3 d, vec3 e) { c[0] = a; c[1] = b; c[2] = d; c[3] = e; }
    vec3& operator[](int j) { return c[j]; }
    const vec3& operator[](int j) const { return c[j]; }
};
inline vec3 operator*(const mat4x3& m, vec4 v) { return ((m.c[0] * v.x + m.c[1] * v.y) + m.c[2] * v.z) + m.c[3] * v.w; }

// ---- built-in functions (GLSL 4.50 section 8) ----
// Scalars first; C's names for the same functions are taken over where the semantics agree.
inline float gl_sign1(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
inline float sign(float x) { return gl_sign1(x); }
inline float fract(float x) { return x - floorf(x); }
inline float mix(float x, float y, float a) { return x * (1.0f - a) + y * a; }
inline float clamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
inline int clamp(int x, int lo, int hi) { return x < lo ? lo : (x > hi ? hi : x); }
inline uint clamp(uint x, uint lo, uint hi) { return x < lo ? lo : (x > hi ? hi : x); }
inline float min(float a, float b) { return fminf(a, b); }
inline float max(float a, float b) { return fmaxf(a, b); }
inline int min(int a, int b) { return a < b ? a : b; }
inline int max(int a, int b) { return a > b ? a : b; }
inline uint min(uint a, uint b) { return a < b ? a : b; }
inline uint max(uint a, uint b) { return a > b ? a : b; }
inline float abs(float x) { return fabsf(x); }
inline float sqrt(float x) { return sqrtf(x); }
inline float inversesqrt(float x) { return 1.0f / sqrtf(x); }
inline float exp(float x) { return expf(x); }
inline float log(float x) { return logf(x); }
inline float pow(float x, float y) { return powf(x, y); }
inline float sin(float x) { return sinf(x); }
inline float cos(float x) { return cosf(x); }
inline float tan(float x) { return tanf(x); }
inline float asin(float x) { return asinf(x); }
inline float acos(float x) { return acosf(x); }
inline float atan(float y, float x) { return atan2f(y, x); }
inline float atan(float x) { return atanf(x); }
inline float floor(float x) { return floorf(x); }
// guards against double overloads being picked for mixed int/float literals
inline float max(float a, int b) { return fmaxf(a, (float)b); }
inline float max(int a, float b) { return fmaxf((float)a, b); }
inline float min(float a, int b) { return fminf(a, (float)b); }
inline float min(int a, float b) { return fminf((float)a, b); }
inline float max(float a, double b) { return fmaxf(a, (float)b); }
inline float max(double a, float b) { return fmaxf((float)a, b); }
inline float min(float a, double b) { return fminf(a, (float)b); }
inline float min(double a, float b) { return fminf((float)a, b); }
inline float pow(float x, double y) { return powf(x, (float)y); }
inline float pow(float x, int y) { return powf(x, (float)y); }
inline float mix(float x, float y, double a) { return mix(x, y, (float)a); }
inline float clamp(float x, double lo, double hi) { return clamp(x, (float)lo, (float)hi); }
inline float clamp(float x, int lo, int hi) { return clamp(x, (float)lo, (float)hi); }

#define REF_COMPONENTWISE1(F) \
    inline vec2 F(vec2 a) { return vec2(F(a.x), F(a.y)); } \
    inline vec3 F(vec3 a) { return vec3(F(a.x), F(a.y), F(a.z)); } \
    inline vec4 F(vec4 a) { return vec4(F(a.x), F(a.y), F(a.z), F(a.w)); }
REF_COMPONENTWISE1(abs) REF_COMPONENTWISE1(sqrt) REF_COMPONENTWISE1(exp) REF_COMPONENTWISE1(log) REF_COMPONENTWISE1(sin)
REF_COMPONENTWISE1(cos) REF_COMPONENTWISE1(fract) REF_COMPONENTWISE1(floor) REF_COMPONENTWISE1(sign)
#define REF_COMPONENTWISE2(F) \
    inline vec2 F(vec2 a, vec2 b) { return vec2(F(a.x, b.x), F(a.y, b.y)); } \
    inline vec3 F(vec3 a, vec3 b) { return vec3(F(a.x, b.x), F(a.y, b.y), F(a.z, b.z)); } \
    inline vec4 F(vec4 a, vec4 b) { return vec4(F(a.x, b.x), F(a.y, b.y), F(a.z, b.z), F(a.w, b.w)); } \
    inline vec2 F(vec2 a, float s) { return vec2(F(a.x, s), F(a.y, s)); } \
    inline vec3 F(vec3 a, float s) { return vec3(F(a.x, s), F(a.y, s), F(a.z, s)); } \
    inline vec4 F(vec4 a, float s) { return vec4(F(a.x, s), F(a.y, s), F(a.z, s), F(a.w, s)); }
REF_COMPONENTWISE2(min) REF_COMPONENTWISE2(max) REF_COMPONENTWISE2(pow)
inline vec2 mix(vec2 a, vec2 b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
inline vec4 mix(vec4 a, vec4 b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 mix(vec3 a, vec3 b, vec3 t) { return vec3(mix(a.x, b.x, t.x), mix(a.y, b.y, t.y), mix(a.z, b.z, t.z)); }
inline vec4 mix(vec4 a, vec4 b, vec4 t) { return vec4(mix(a.x, b.x, t.x), mix(a.y, b.y, t.y), mix(a.z, b.z, t.z), mix(a.w, b.w, t.w)); }
inline vec2 clamp(vec2 a, float lo, float hi) { return vec2(clamp(a.x, lo, hi), clamp(a.y, lo, hi)); }
inline vec3 clamp(vec3 a, float lo, float hi) { return vec3(clamp(a.x, lo, hi), clamp(a.y, lo, hi), clamp(a.z, lo, hi)); }
inline vec4 clamp(vec4 a, float lo, float hi) { return vec4(clamp(a.x, lo, hi), clamp(a.y, lo, hi), clamp(a.z, lo, hi), clamp(a.w, lo, hi)); }
inline vec3 clamp(vec3 a, vec3 lo, vec3 hi) { return vec3(clamp(a.x, lo.x, hi.x), clamp(a.y, lo.y, hi.y), clamp(a.z, lo.z, hi.z)); }

inline float uintBitsToFloat(uint u) { float f; memcpy(&f, &u, 4); return f; }
inline uint floatBitsToUint(float f) { uint u; memcpy(&u, &f, 4); return u; }

// packSnorm2x16: round(clamp(c,-1,+1) * 32767.0), ties away from zero (glm::round);  unpack: clamp(v / 32767, -1, 1).
inline uint packSnorm2x16(vec2 v) {
    int x = (int)roundf(clamp(v.x, -1.0f, 1.0f) * 32767.0f);
    int y = (int)roundf(clamp(v.y, -1.0f, 1.0f) * 32767.0f);
    return ((uint)(uint16_t)(int16_t)x) | (((uint)(uint16_t)(int16_t)y) << 16);
}
inline vec2 unpackSnorm2x16(uint p) {
    float x = (float)(int16_t)(p & 0xFFFFu), y = (float)(int16_t)(p >> 16);
    return vec2(clamp(x / 32767.0f, -1.0f, 1.0f), clamp(y / 32767.0f, -1.0f, 1.0f));
}
inline float ref_half_to_float(uint16_t h) {          // IEEE binary16 -> binary32, exact
    uint s = (uint)(h >> 15) << 31, e = (h >> 10) & 0x1Fu, m = h & 0x3FFu, bits;
    if (e == 0) bits = m == 0 ? s : (s | floatBitsToUint((float)m * 5.9604644775390625e-8f));
    else if (e == 31) bits = s | 0x7F800000u | (m << 13);
    else bits = s | ((e + 112u) << 23) | (m << 13);
    return uintBitsToFloat(bits);
}
inline vec2 unpackHalf2x16(uint p) { return vec2(ref_half_to_float((uint16_t)(p & 0xFFFFu)), ref_half_to_float((uint16_t)(p >> 16))); }

// ---- samplers and images ----
// Vulkan texel addressing for a 2D array view: REPEAT, LOD 0, unnormalised u = s * size; NEAREST takes floor(u);
// LINEAR takes texels floor(u - 0.5) and + 1 with weights frac(u - 0.5), kept in fp32.
struct sampler2DArray {
    const float* const* layers = nullptr;      // RGBA32F
    int width = 0, height = 0;
    bool nearest = false;
    static int wrap(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }
    vec4 texel(uint layer, int x, int y) const {
        const float* L = layers[layer];
        size_t o = ((size_t)wrap(y, height) * width + (size_t)wrap(x, width)) * 4;
        return vec4(L[o], L[o + 1], L[o + 2], L[o + 3]);
    }
};
inline vec4 textureLod(const sampler2DArray& S, vec3 uvw, float) {
    const uint layer = (uint)uvw.z;
    float u = uvw.x * (float)S.width, v = uvw.y * (float)S.height;
    if (S.nearest) return S.texel(layer, (int)floorf(u), (int)floorf(v));
    float ub = u - 0.5f, vb = v - 0.5f;
    float fu = floorf(ub), fv = floorf(vb);
    float a = ub - fu, b = vb - fv;
    int i0 = (int)fu, j0 = (int)fv;
    vec4 t00 = S.texel(layer, i0, j0), t10 = S.texel(layer, i0 + 1, j0), t01 = S.texel(layer, i0, j0 + 1), t11 = S.texel(layer, i0 + 1, j0 + 1);
    vec4 r0 = t00 * (1.0f - a) + t10 * a;
    vec4 r1 = t01 * (1.0f - a) + t11 * a;
    return r0 * (1.0f - b) + r1 * b;
}
struct image2D {
    float* rgba = nullptr;
    int width = 0, height = 0;
};
inline ivec2 imageSize(const image2D& I) { return ivec2(I.width, I.height); }
inline vec4 imageLoad(const image2D& I, ivec2 p) { const float* t = I.rgba + ((size_t)p.y * I.width + p.x) * 4; return vec4(t[0], t[1], t[2], t[3]); }
inline void imageStore(image2D& I, ivec2 p, vec4 v) { float* t = I.rgba + ((size_t)p.y * I.width + p.x) * 4; t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w; }

// ---- compute-shader built-in variables (set by the dispatch loops in ref_api.cpp) ----
extern uvec3 gl_GlobalInvocationID, gl_WorkGroupID, gl_LocalInvocationID;

}  // namespace glsl
