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

// glm 1.0.1's association: a matrix-vector product sums its four column terms in pairs, a matrix-matrix product sums
// them left to right (both pinned against the reference's own PackSceneData, tests/test_reference_host.py).
inline vec4 mul(const mat4& m, vec4 v) { return (m.c[0] * v.x + m.c[1] * v.y) + (m.c[2] * v.z + m.c[3] * v.w); }
inline mat4 mul(const mat4& a, const mat4& b) {
    mat4 r;
    for (int j = 0; j < 4; j++) r.c[j] = ((a.c[0] * b.c[j].x + a.c[1] * b.c[j].y) + a.c[2] * b.c[j].z) + a.c[3] * b.c[j].w;
    return r;
}

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

// T(position) * R(rotation) * S(scale)   (common.hpp:72-82), as the reference computes it: glm::translate and glm::scale of the
// identity, then two full matrix products.  The values are those of "columns of R times the scale, position in column 3"; what the
// products add is the SIGN OF ZEROS: a -0 of R (sin of a zero angle, negated) or of a negative scale meets the +0 terms of the sums
// and comes out +0.  Found by tools/fuzz_pack.py against the reference's own PackSceneData (byte for byte, zeros included).
inline mat4 make_transform(vec3 position, vec3 rotation, vec3 scale) {
    const mat4 I = mat4::identity();
    mat4 t = I;                                              // glm::translate: m[0] v[0] + m[1] v[1] + m[2] v[2] + m[3]
    t.c[3] = ((I.c[0] * position.x + I.c[1] * position.y) + I.c[2] * position.z) + I.c[3];
    mat4 sc;                                                 // glm::scale: columns times the factors
    sc.c[0] = I.c[0] * scale.x; sc.c[1] = I.c[1] * scale.y; sc.c[2] = I.c[2] * scale.z; sc.c[3] = I.c[3];
    return mul(mul(t, euler_zyx(rotation.z, rotation.y, rotation.x)), sc);
}

// glm::inverse(mat4) as the reference calls it (scene.hpp:403-406 PackTransform), in glm 1.0.1's operation order:
// eighteen 2x2 minors of rows 1..3, four cofactor columns as sums of three products, signs by multiplication,
// the determinant from column 0 with pairwise association, and one reciprocal.  Pinned bit for bit against the
// reference's own PackTransform compiled in oracle/_ref/libref_host.so (tests/test_reference_host.py).
inline mat4 inverse(const mat4& m) {
    auto e = [&](int col, int row) { return m.c[col][row]; };
    auto minor2 = [&](int c0, int c1, int r0, int r1) { return e(c0, r0) * e(c1, r1) - e(c1, r0) * e(c0, r1); };
    // minors named by (columns, rows): the pairs glm calls Coef00..Coef23
    const float k00 = minor2(2, 3, 2, 3), k02 = minor2(1, 3, 2, 3), k03 = minor2(1, 2, 2, 3);
    const float k04 = minor2(2, 3, 1, 3), k06 = minor2(1, 3, 1, 3), k07 = minor2(1, 2, 1, 3);
    const float k08 = minor2(2, 3, 1, 2), k10 = minor2(1, 3, 1, 2), k11 = minor2(1, 2, 1, 2);
    const float k12 = minor2(2, 3, 0, 3), k14 = minor2(1, 3, 0, 3), k15 = minor2(1, 2, 0, 3);
    const float k16 = minor2(2, 3, 0, 2), k18 = minor2(1, 3, 0, 2), k19 = minor2(1, 2, 0, 2);
    const float k20 = minor2(2, 3, 0, 1), k22 = minor2(1, 3, 0, 1), k23 = minor2(1, 2, 0, 1);
    const float f0[4] = {k00, k00, k02, k03}, f1[4] = {k04, k04, k06, k07}, f2[4] = {k08, k08, k10, k11};
    const float f3[4] = {k12, k12, k14, k15}, f4[4] = {k16, k16, k18, k19}, f5[4] = {k20, k20, k22, k23};
    float v[4][4];                                       // v[r][i]: element of row r taken from column (i == 0 ? 1 : 0)
    for (int r = 0; r < 4; r++) { v[r][0] = e(1, r); v[r][1] = v[r][2] = v[r][3] = e(0, r); }
    mat4 adj;
    for (int i = 0; i < 4; i++) {
        const float sa = (i & 1) ? -1.0f : 1.0f, sb = -sa;
        adj.c[0][i] = (v[1][i] * f0[i] - v[2][i] * f1[i] + v[3][i] * f2[i]) * sa;
        adj.c[1][i] = (v[0][i] * f0[i] - v[2][i] * f3[i] + v[3][i] * f4[i]) * sb;
        adj.c[2][i] = (v[0][i] * f1[i] - v[1][i] * f3[i] + v[3][i] * f5[i]) * sa;
        adj.c[3][i] = (v[0][i] * f2[i] - v[1][i] * f4[i] + v[2][i] * f5[i]) * sb;
    }
    const float d0 = e(0, 0) * adj.c[0][0], d1 = e(0, 1) * adj.c[1][0], d2 = e(0, 2) * adj.c[2][0], d3 = e(0, 3) * adj.c[3][0];
    const float inv_det = 1.0f / ((d0 + d1) + (d2 + d3));
    for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) adj.c[c][r] = adj.c[c][r] * inv_det;
    return adj;
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
// glm::packHalf2x16 (scene.cpp:1297) converts with glm::detail::toFloat16 (glm 1.0.1, detail/type_half.inl): round to nearest with
// TIES AWAY FROM ZERO ("round 0.5 up" on the magnitude), values below 2^-25 flushed to zero, overflow to infinity - not IEEE
// round-to-nearest-even.  A tie needs the 13 dropped bits to be exactly 0x1000: one texture coordinate in 8192 for arbitrary floats,
// every one of the form (2k + 1) / 4096 in [0.5, 1).  Found by tools/fuzz_pack.py against the reference's own PackSceneData.
inline uint16_t float_to_half(float f) {
    uint32_t i; std::memcpy(&i, &f, 4);
    int s = (int)((i >> 16) & 0x00008000u);
    int e = (int)((i >> 23) & 0x000000ffu) - (127 - 15);
    int m = (int)(i & 0x007fffffu);
    if (e <= 0) {
        if (e < -10) return (uint16_t)s;
        m = (m | 0x00800000) >> (1 - e);
        if (m & 0x00001000) m += 0x00002000;
        return (uint16_t)(s | (m >> 13));
    }
    if (e == 0xff - (127 - 15)) {
        if (m == 0) return (uint16_t)(s | 0x7c00);
        m >>= 13;
        return (uint16_t)(s | 0x7c00 | m | (m == 0));
    }
    if (m & 0x00001000) { m += 0x00002000; if (m & 0x00800000) { m = 0; e += 1; } }
    if (e > 30) return (uint16_t)(s | 0x7c00);
    return (uint16_t)(s | (e << 10) | (m >> 13));
}
inline uint32_t pack_half2x16(float x, float y) { return (uint32_t)float_to_half(x) | ((uint32_t)float_to_half(y) << 16); }

}}  // namespace ptb::host
