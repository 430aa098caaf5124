// Host-side small vector/matrix kit for the scene layer (stands in for the
// reference's un-vendored glm 1.0.1).  Column-major mat4,
// fp32, fixed left-to-right association so results are reproducible.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <limits>

namespace ptb { namespace host {

struct vec2 { float x = 0, y = 0; };
struct vec3 {
    float x = 0, y = 0, z = 0;
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};
struct vec4 {
    float x = 0, y = 0, z = 0, w = 0;
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 min3(vec3 a, vec3 b) { return {std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)}; }
inline vec3 max3(vec3 a, vec3 b) { return {std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)}; }
inline float dot3(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length3(vec3 a) { return std::sqrt(dot3(a, a)); }
inline vec3 cross3(vec3 a, vec3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline vec3 normalize3(vec3 a) { float l = length3(a); return l > 0 ? a / l : vec3{0, 0, 1}; }

inline vec4 operator+(vec4 a, vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline vec4 operator*(vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }

struct mat4 {
    vec4 c[4];   // columns
    static mat4 identity() { mat4 m; m.c[0] = {1, 0, 0, 0}; m.c[1] = {0, 1, 0, 0}; m.c[2] = {0, 0, 1, 0}; m.c[3] = {0, 0, 0, 1}; return m; }
};

inline vec4 mul(const mat4& m, vec4 v) { return ((m.c[0] * v.x + m.c[1] * v.y) + m.c[2] * v.z) + m.c[3] * v.w; }
inline mat4 mul(const mat4& a, const mat4& b) { mat4 r; for (int j = 0; j < 4; j++) r.c[j] = mul(a, b.c[j]); return r; }

// Rz(rz) * Ry(ry) * Rx(rx)  — role of glm::eulerAngleZYX(Rotation.z, Rotation.y, Rotation.x)
// in /root/reference/src/core/common.hpp:62-82.
inline mat4 euler_zyx(float rz, float ry, float rx) {
    float c1 = std::cos(rz), s1 = std::sin(rz);
    float c2 = std::cos(ry), s2 = std::sin(ry);
    float c3 = std::cos(rx), s3 = std::sin(rx);
    mat4 m = mat4::identity();
    m.c[0] = {c1 * c2, c2 * s1, -s2, 0};
    m.c[1] = {c1 * s2 * s3 - c3 * s1, c1 * c3 + s1 * s2 * s3, c2 * s3, 0};
    m.c[2] = {s1 * s3 + c1 * c3 * s2, c3 * s1 * s2 - c1 * s3, c2 * c3, 0};
    return m;
}

// T(position) * R(rotation) * S(scale)   (common.hpp:72-82)
inline mat4 make_transform(vec3 position, vec3 rotation, vec3 scale) {
    mat4 r = euler_zyx(rotation.z, rotation.y, rotation.x);
    mat4 m;
    m.c[0] = r.c[0] * scale.x;
    m.c[1] = r.c[1] * scale.y;
    m.c[2] = r.c[2] * scale.z;
    m.c[3] = {position.x, position.y, position.z, 1.0f};
    return m;
}

// General 4x4 inverse by cofactors (role of glm::inverse at scene.hpp:405).
inline mat4 inverse(const mat4& m) {
    const float a00 = m.c[0].x, a01 = m.c[0].y, a02 = m.c[0].z, a03 = m.c[0].w;
    const float a10 = m.c[1].x, a11 = m.c[1].y, a12 = m.c[1].z, a13 = m.c[1].w;
    const float a20 = m.c[2].x, a21 = m.c[2].y, a22 = m.c[2].z, a23 = m.c[2].w;
    const float a30 = m.c[3].x, a31 = m.c[3].y, a32 = m.c[3].z, a33 = m.c[3].w;
    const float b00 = a00 * a11 - a01 * a10, b01 = a00 * a12 - a02 * a10, b02 = a00 * a13 - a03 * a10;
    const float b03 = a01 * a12 - a02 * a11, b04 = a01 * a13 - a03 * a11, b05 = a02 * a13 - a03 * a12;
    const float b06 = a20 * a31 - a21 * a30, b07 = a20 * a32 - a22 * a30, b08 = a20 * a33 - a23 * a30;
    const float b09 = a21 * a32 - a22 * a31, b10 = a21 * a33 - a23 * a31, b11 = a22 * a33 - a23 * a32;
    const float det = ((b00 * b11 - b01 * b10) + (b02 * b09 + b03 * b08)) - (b04 * b07 - b05 * b06);
    const float id = 1.0f / det;
    mat4 r;
    r.c[0] = {(a11 * b11 - a12 * b10 + a13 * b09) * id, (a02 * b10 - a01 * b11 - a03 * b09) * id,
              (a31 * b05 - a32 * b04 + a33 * b03) * id, (a22 * b04 - a21 * b05 - a23 * b03) * id};
    r.c[1] = {(a12 * b08 - a10 * b11 - a13 * b07) * id, (a00 * b11 - a02 * b08 + a03 * b07) * id,
              (a32 * b02 - a30 * b05 - a33 * b01) * id, (a20 * b05 - a22 * b02 + a23 * b01) * id};
    r.c[2] = {(a10 * b10 - a11 * b08 + a13 * b06) * id, (a01 * b08 - a00 * b10 - a03 * b06) * id,
              (a30 * b04 - a31 * b02 + a33 * b00) * id, (a21 * b02 - a20 * b04 - a23 * b00) * id};
    r.c[3] = {(a11 * b07 - a10 * b09 - a12 * b06) * id, (a00 * b09 - a01 * b07 + a02 * b06) * id,
              (a31 * b01 - a30 * b03 - a32 * b00) * id, (a20 * b03 - a21 * b01 + a22 * b00) * id};
    return r;
}

// ---- pack helpers (common.hpp:90-114, scene.cpp:1297) ----
inline uint32_t pack_snorm2x16(float x, float y) {
    auto q = [](float v) { v = std::min(std::max(v, -1.0f), 1.0f); return (int)std::round(v * 32767.0f); };
    return ((uint32_t)(uint16_t)(int16_t)q(x)) | (((uint32_t)(uint16_t)(int16_t)q(y)) << 16);
}
// Octahedral unit-vector encode (common.hpp:100-105).
inline uint32_t pack_unit_vector(vec3 v) {
    float s = 1.0f / ((std::fabs(v.x) + std::fabs(v.y)) + std::fabs(v.z));
    float px = v.x * s, py = v.y * s;
    if (v.z <= 0.0f) {
        float qx = (1.0f - std::fabs(py)) * (px >= 0.0f ? 1.0f : -1.0f);
        float qy = (1.0f - std::fabs(px)) * (py >= 0.0f ? 1.0f : -1.0f);
        px = qx; py = qy;
    }
    return pack_snorm2x16(px, py);
}
// binary32 -> binary16, round to nearest even (glm::packHalf2x16 role).
inline uint16_t float_to_half(float f) {
    uint32_t x; std::memcpy(&x, &f, 4);
    uint32_t s = (x >> 16) & 0x8000u, ax = x & 0x7FFFFFFFu;
    if (ax >= 0x7F800000u) return (uint16_t)(s | 0x7C00u | (ax > 0x7F800000u ? 0x200u : 0u));
    if (ax >= 0x477FF000u) return (uint16_t)(s | 0x7C00u);
    if (ax < 0x33000001u) return (uint16_t)s;
    int e = (int)(ax >> 23) - 127;
    uint32_t m = (ax & 0x7FFFFFu) | 0x800000u;
    int shift = e < -14 ? 13 + (-14 - e) : 13;
    uint32_t he = e < -14 ? 0u : (uint32_t)(e + 15);
    uint32_t hm = m >> shift, rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (hm & 1u))) hm++;
    uint32_t h = he == 0 ? hm : ((he << 10) + (hm - 0x400u));
    return (uint16_t)(s | h);
}
inline uint32_t pack_half2x16(float x, float y) { return (uint32_t)float_to_half(x) | ((uint32_t)float_to_half(y) << 16); }

}}  // namespace ptb::host
