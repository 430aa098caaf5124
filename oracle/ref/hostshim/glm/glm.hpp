// ORACLE - TEST INFRASTRUCTURE ONLY.
//
// Stand-in for glm 1.0.1 (the reference's un-vendored dependency, /root/reference/CMakeLists.txt:9
// `add_subdirectory (lib/glm-1.0.1)`), so that the reference's own host sources (src/scene/scene.cpp up to its
// Vulkan half, src/core/spectrum.cpp, the material packers) can be compiled here into oracle/_ref/libref_host.so.
// Only the subset those files use.  The arithmetic restates glm's published scalar code paths (GLM_FORCE_INTRINSICS
// is not defined by the reference, so glm's SIMD paths are off): operation order of mat4 * mat4, mat4 * vec4,
// inverse, determinant, eulerAngleZYX, translate, scale, dot, cross, normalize, mix, the packing functions.
// What this library pins is therefore the reference's host code; glm itself is restated from its documentation.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>

namespace glm {

typedef unsigned int uint;

template <typename T> struct tvec2;
template <typename T> struct tvec3;
template <typename T> struct tvec4;

template <typename T> struct tvec2 {
    T x, y;
    tvec2() : x(0), y(0) {}
    explicit tvec2(T s) : x(s), y(s) {}
    tvec2(T a, T b) : x(a), y(b) {}
    template <typename U> explicit tvec2(const tvec2<U>& v) : x((T)v.x), y((T)v.y) {}
    tvec2(const tvec3<T>& v);
    tvec2(const tvec4<T>& v);
    T& operator[](int i) { return (&x)[i]; }
    const T& operator[](int i) const { return (&x)[i]; }
    tvec2 yx() const { return tvec2(y, x); }
    tvec2 xy() const { return *this; }
    tvec2& operator+=(const tvec2& o) { x += o.x; y += o.y; return *this; }
    tvec2& operator-=(const tvec2& o) { x -= o.x; y -= o.y; return *this; }
    tvec2& operator*=(T s) { x *= s; y *= s; return *this; }
    tvec2& operator/=(T s) { x /= s; y /= s; return *this; }
};

template <typename T> struct swizzle_xy {          // V.xy of GLM_FORCE_SWIZZLE: reads the first two components in place
    T x, y;
    operator tvec2<T>() const { return tvec2<T>(x, y); }
    tvec2<T> operator*(T s) const { return tvec2<T>(x * s, y * s); }
    tvec2<T> operator()() const { return tvec2<T>(x, y); }
};

template <typename T> struct tvec3 {
    union {
        struct { T x, y, z; };
        struct { T r, g, b; };
        swizzle_xy<T> xy;
    };
    tvec3() : x(0), y(0), z(0) {}
    explicit tvec3(T s) : x(s), y(s), z(s) {}
    tvec3(T a, T b, T c) : x(a), y(b), z(c) {}
    template <typename A, typename B, typename C> tvec3(A a, B b, C c) : x((T)a), y((T)b), z((T)c) {}
    tvec3(const tvec2<T>& v, T c) : x(v.x), y(v.y), z(c) {}
    tvec3(const tvec3& o) : x(o.x), y(o.y), z(o.z) {}
    tvec3& operator=(const tvec3& o) { x = o.x; y = o.y; z = o.z; return *this; }
    template <typename U> tvec3(const tvec3<U>& v) : x((T)v.x), y((T)v.y), z((T)v.z) {}
    tvec3(const tvec4<T>& v);                   // glm's truncating conversions are implicit unless GLM_FORCE_EXPLICIT_CTOR
    tvec3 xyz() const { return *this; }
    tvec3 rgb() const { return *this; }
    T& operator[](int i) { return (&x)[i]; }
    const T& operator[](int i) const { return (&x)[i]; }
    tvec3& operator+=(const tvec3& o) { x += o.x; y += o.y; z += o.z; return *this; }
    tvec3& operator-=(const tvec3& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
    tvec3& operator*=(const tvec3& o) { x *= o.x; y *= o.y; z *= o.z; return *this; }
    tvec3& operator*=(T s) { x *= s; y *= s; z *= s; return *this; }
    tvec3& operator/=(T s) { x /= s; y /= s; z /= s; return *this; }
};

template <typename T> struct tvec4 {
    union {
        struct { T x, y, z, w; };
        struct { T r, g, b, a; };
    };
    tvec4() : x(0), y(0), z(0), w(0) {}
    explicit tvec4(T s) : x(s), y(s), z(s), w(s) {}
    tvec4(T a_, T b_, T c_, T d_) : x(a_), y(b_), z(c_), w(d_) {}
    template <typename A, typename B, typename C, typename D> tvec4(A a_, B b_, C c_, D d_) : x((T)a_), y((T)b_), z((T)c_), w((T)d_) {}
    tvec4(const tvec3<T>& v, T d_) : x(v.x), y(v.y), z(v.z), w(d_) {}
    tvec4(const tvec4& o) : x(o.x), y(o.y), z(o.z), w(o.w) {}
    tvec4& operator=(const tvec4& o) { x = o.x; y = o.y; z = o.z; w = o.w; return *this; }
    T& operator[](int i) { return (&x)[i]; }
    const T& operator[](int i) const { return (&x)[i]; }
    tvec3<T> xyz() const { return tvec3<T>(x, y, z); }
    tvec3<T> rgb() const { return tvec3<T>(x, y, z); }
    tvec4& operator+=(const tvec4& o) { x += o.x; y += o.y; z += o.z; w += o.w; return *this; }
    tvec4& operator*=(T s) { x *= s; y *= s; z *= s; w *= s; return *this; }
};
template <typename T> tvec3<T>::tvec3(const tvec4<T>& v) : x(v.x), y(v.y), z(v.z) {}
template <typename T> tvec2<T>::tvec2(const tvec3<T>& v) : x(v.x), y(v.y) {}
template <typename T> tvec2<T>::tvec2(const tvec4<T>& v) : x(v.x), y(v.y) {}

#define GLM_SHIM_OPS(V, N, ...)                                                                                          \
    template <typename T> V<T> operator+(const V<T>& a, const V<T>& b) { V<T> r; for (int i = 0; i < N; i++) r[i] = a[i] + b[i]; return r; } \
    template <typename T> V<T> operator-(const V<T>& a, const V<T>& b) { V<T> r; for (int i = 0; i < N; i++) r[i] = a[i] - b[i]; return r; } \
    template <typename T> V<T> operator*(const V<T>& a, const V<T>& b) { V<T> r; for (int i = 0; i < N; i++) r[i] = a[i] * b[i]; return r; } \
    template <typename T> V<T> operator/(const V<T>& a, const V<T>& b) { V<T> r; for (int i = 0; i < N; i++) r[i] = a[i] / b[i]; return r; } \
    template <typename T> V<T> operator-(const V<T>& a) { V<T> r; for (int i = 0; i < N; i++) r[i] = -a[i]; return r; }                      \
    template <typename T> V<T> operator*(const V<T>& a, T s) { V<T> r; for (int i = 0; i < N; i++) r[i] = a[i] * s; return r; }              \
    template <typename T> V<T> operator*(T s, const V<T>& a) { V<T> r; for (int i = 0; i < N; i++) r[i] = s * a[i]; return r; }              \
    template <typename T> V<T> operator/(const V<T>& a, T s) { V<T> r; for (int i = 0; i < N; i++) r[i] = a[i] / s; return r; }              \
    template <typename T> V<T> operator/(T s, const V<T>& a) { V<T> r; for (int i = 0; i < N; i++) r[i] = s / a[i]; return r; }              \
    template <typename T> V<T> operator+(const V<T>& a, T s) { V<T> r; for (int i = 0; i < N; i++) r[i] = a[i] + s; return r; }              \
    template <typename T> V<T> operator+(T s, const V<T>& a) { V<T> r; for (int i = 0; i < N; i++) r[i] = s + a[i]; return r; }              \
    template <typename T> V<T> operator-(const V<T>& a, T s) { V<T> r; for (int i = 0; i < N; i++) r[i] = a[i] - s; return r; }              \
    template <typename T> V<T> operator-(T s, const V<T>& a) { V<T> r; for (int i = 0; i < N; i++) r[i] = s - a[i]; return r; }              \
    template <typename T> bool operator==(const V<T>& a, const V<T>& b) { for (int i = 0; i < N; i++) if (!(a[i] == b[i])) return false; return true; } \
    template <typename T> bool operator!=(const V<T>& a, const V<T>& b) { return !(a == b); }                                                \
    template <typename T> V<T> abs(const V<T>& a) { V<T> r; for (int i = 0; i < N; i++) r[i] = std::fabs(a[i]); return r; }                  \
    template <typename T> V<T> min(const V<T>& a, const V<T>& b) { V<T> r; for (int i = 0; i < N; i++) r[i] = (b[i] < a[i]) ? b[i] : a[i]; return r; } \
    template <typename T> V<T> max(const V<T>& a, const V<T>& b) { V<T> r; for (int i = 0; i < N; i++) r[i] = (a[i] < b[i]) ? b[i] : a[i]; return r; } \
    template <typename T> V<T> clamp(const V<T>& v, const V<T>& lo, const V<T>& hi) { return min(max(v, lo), hi); }                          \
    template <typename T> V<T> clamp(const V<T>& v, T lo, T hi) { return min(max(v, V<T>(lo)), V<T>(hi)); }                                  \
    template <typename T> V<T> mix(const V<T>& x, const V<T>& y, T a) { return x * (T(1) - a) + y * a; }                                     \
    template <typename T> V<T> floor(const V<T>& a) { V<T> r; for (int i = 0; i < N; i++) r[i] = std::floor(a[i]); return r; }               \
    template <typename T> V<T> round(const V<T>& a) { V<T> r; for (int i = 0; i < N; i++) r[i] = std::round(a[i]); return r; }

GLM_SHIM_OPS(tvec2, 2)
GLM_SHIM_OPS(tvec3, 3)
GLM_SHIM_OPS(tvec4, 4)

typedef tvec2<float> vec2;  typedef tvec3<float> vec3;  typedef tvec4<float> vec4;
typedef tvec2<double> dvec2; typedef tvec3<double> dvec3; typedef tvec4<double> dvec4;
typedef tvec2<int> ivec2;   typedef tvec3<int> ivec3;   typedef tvec4<int> ivec4;
typedef tvec2<unsigned> uvec2; typedef tvec3<unsigned> uvec3; typedef tvec4<unsigned> uvec4;

// scalars
inline float abs(float x) { return std::fabs(x); }
inline double abs(double x) { return std::fabs(x); }
inline int abs(int x) { return x < 0 ? -x : x; }
template <typename T> T min(T a, T b) { return (b < a) ? b : a; }
template <typename T> T max(T a, T b) { return (a < b) ? b : a; }
template <typename T> T clamp(T x, T lo, T hi) { return min(max(x, lo), hi); }
template <typename T, typename U> T mix(T x, T y, U a) { return static_cast<T>(static_cast<U>(x) * (static_cast<U>(1) - a) + static_cast<U>(y) * a); }
inline float sqrt(float x) { return std::sqrt(x); }
inline double sqrt(double x) { return std::sqrt(x); }
inline float pow(float x, float y) { return std::pow(x, y); }
inline double pow(double x, double y) { return std::pow(x, y); }
inline float exp(float x) { return std::exp(x); }
inline double exp(double x) { return std::exp(x); }
inline float sin(float x) { return std::sin(x); }
inline float cos(float x) { return std::cos(x); }
inline float tan(float x) { return std::tan(x); }
inline double sin(double x) { return std::sin(x); }
inline double cos(double x) { return std::cos(x); }
inline float floor(float x) { return std::floor(x); }
inline float fract(float x) { return x - std::floor(x); }
inline float radians(float d) { return d * 0.01745329251994329576923690768489f; }
inline float degrees(float r) { return r * 57.295779513082320876798154814105f; }
inline uint floatBitsToUint(float f) { uint u; std::memcpy(&u, &f, 4); return u; }
inline int floatBitsToInt(float f) { int u; std::memcpy(&u, &f, 4); return u; }
inline float uintBitsToFloat(uint u) { float f; std::memcpy(&f, &u, 4); return f; }

// geometric
template <typename T> T dot(const tvec2<T>& a, const tvec2<T>& b) { tvec2<T> t(a * b); return t.x + t.y; }
template <typename T> T dot(const tvec3<T>& a, const tvec3<T>& b) { tvec3<T> t(a * b); return t.x + t.y + t.z; }
template <typename T> T dot(const tvec4<T>& a, const tvec4<T>& b) { tvec4<T> t(a * b); return (t.x + t.y) + (t.z + t.w); }
template <typename T> tvec3<T> cross(const tvec3<T>& x, const tvec3<T>& y) {
    return tvec3<T>(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
template <typename V> auto length(const V& v) -> decltype(dot(v, v)) { return std::sqrt(dot(v, v)); }
template <typename T> T inversesqrt(T x) { return T(1) / std::sqrt(x); }
template <typename V> V normalize(const V& v) { return v * inversesqrt(dot(v, v)); }

// matrices (column major)
template <typename T> struct tmat3 {
    tvec3<T> c[3];
    tmat3() : tmat3(T(1)) { c[0] = tvec3<T>(); c[1] = tvec3<T>(); c[2] = tvec3<T>(); }
    explicit tmat3(T s) { c[0] = tvec3<T>(s, 0, 0); c[1] = tvec3<T>(0, s, 0); c[2] = tvec3<T>(0, 0, s); }
    tmat3(T x0, T y0, T z0, T x1, T y1, T z1, T x2, T y2, T z2) { c[0] = tvec3<T>(x0, y0, z0); c[1] = tvec3<T>(x1, y1, z1); c[2] = tvec3<T>(x2, y2, z2); }
    tmat3(const tvec3<T>& a, const tvec3<T>& b, const tvec3<T>& d) { c[0] = a; c[1] = b; c[2] = d; }
    tvec3<T>& operator[](int i) { return c[i]; }
    const tvec3<T>& operator[](int i) const { return c[i]; }
};
template <typename T> tvec3<T> operator*(const tmat3<T>& m, const tvec3<T>& v) {
    return tvec3<T>(m[0][0] * v.x + m[1][0] * v.y + m[2][0] * v.z,
                    m[0][1] * v.x + m[1][1] * v.y + m[2][1] * v.z,
                    m[0][2] * v.x + m[1][2] * v.y + m[2][2] * v.z);
}
template <typename T> T determinant(const tmat3<T>& m) {
    return +m[0][0] * (m[1][1] * m[2][2] - m[2][1] * m[1][2])
           - m[1][0] * (m[0][1] * m[2][2] - m[2][1] * m[0][2])
           + m[2][0] * (m[0][1] * m[1][2] - m[1][1] * m[0][2]);
}
template <typename T> tmat3<T> inverse(const tmat3<T>& m) {
    T OneOverDeterminant = T(1) / (+m[0][0] * (m[1][1] * m[2][2] - m[2][1] * m[1][2])
                                   - m[1][0] * (m[0][1] * m[2][2] - m[2][1] * m[0][2])
                                   + m[2][0] * (m[0][1] * m[1][2] - m[1][1] * m[0][2]));
    tmat3<T> I(T(0));
    I[0][0] = +(m[1][1] * m[2][2] - m[2][1] * m[1][2]) * OneOverDeterminant;
    I[1][0] = -(m[1][0] * m[2][2] - m[2][0] * m[1][2]) * OneOverDeterminant;
    I[2][0] = +(m[1][0] * m[2][1] - m[2][0] * m[1][1]) * OneOverDeterminant;
    I[0][1] = -(m[0][1] * m[2][2] - m[2][1] * m[0][2]) * OneOverDeterminant;
    I[1][1] = +(m[0][0] * m[2][2] - m[2][0] * m[0][2]) * OneOverDeterminant;
    I[2][1] = -(m[0][0] * m[2][1] - m[2][0] * m[0][1]) * OneOverDeterminant;
    I[0][2] = +(m[0][1] * m[1][2] - m[1][1] * m[0][2]) * OneOverDeterminant;
    I[1][2] = -(m[0][0] * m[1][2] - m[1][0] * m[0][2]) * OneOverDeterminant;
    I[2][2] = +(m[0][0] * m[1][1] - m[1][0] * m[0][1]) * OneOverDeterminant;
    return I;
}

template <typename T> struct tmat4 {
    tvec4<T> c[4];
    tmat4() { for (int i = 0; i < 4; i++) c[i] = tvec4<T>(); }
    explicit tmat4(T s) { c[0] = tvec4<T>(s, 0, 0, 0); c[1] = tvec4<T>(0, s, 0, 0); c[2] = tvec4<T>(0, 0, s, 0); c[3] = tvec4<T>(0, 0, 0, s); }
    tmat4(const tvec4<T>& a, const tvec4<T>& b, const tvec4<T>& d, const tvec4<T>& e) { c[0] = a; c[1] = b; c[2] = d; c[3] = e; }
    tvec4<T>& operator[](int i) { return c[i]; }
    const tvec4<T>& operator[](int i) const { return c[i]; }
};
template <typename T> tvec4<T> operator*(const tmat4<T>& m, const tvec4<T>& v) {
    const tvec4<T> Mov0(v[0]), Mov1(v[1]);
    const tvec4<T> Mul0 = m[0] * Mov0, Mul1 = m[1] * Mov1;
    const tvec4<T> Add0 = Mul0 + Mul1;
    const tvec4<T> Mov2(v[2]), Mov3(v[3]);
    const tvec4<T> Mul2 = m[2] * Mov2, Mul3 = m[3] * Mov3;
    const tvec4<T> Add1 = Mul2 + Mul3;
    return Add0 + Add1;
}
template <typename T> tmat4<T> operator*(const tmat4<T>& m1, const tmat4<T>& m2) {
    tmat4<T> R;
    for (int i = 0; i < 4; i++) {
        tvec4<T> t = m1[0] * m2[i][0];
        t += m1[1] * m2[i][1];
        t += m1[2] * m2[i][2];
        t += m1[3] * m2[i][3];
        R[i] = t;
    }
    return R;
}
template <typename T> tmat4<T> operator*(const tmat4<T>& m, T s) { return tmat4<T>(m[0] * s, m[1] * s, m[2] * s, m[3] * s); }
template <typename T> tmat4<T> inverse(const tmat4<T>& m) {
    T Coef00 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
    T Coef02 = m[1][2] * m[3][3] - m[3][2] * m[1][3];
    T Coef03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    T Coef04 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
    T Coef06 = m[1][1] * m[3][3] - m[3][1] * m[1][3];
    T Coef07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    T Coef08 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
    T Coef10 = m[1][1] * m[3][2] - m[3][1] * m[1][2];
    T Coef11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    T Coef12 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
    T Coef14 = m[1][0] * m[3][3] - m[3][0] * m[1][3];
    T Coef15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    T Coef16 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
    T Coef18 = m[1][0] * m[3][2] - m[3][0] * m[1][2];
    T Coef19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    T Coef20 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
    T Coef22 = m[1][0] * m[3][1] - m[3][0] * m[1][1];
    T Coef23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];
    tvec4<T> Fac0(Coef00, Coef00, Coef02, Coef03), Fac1(Coef04, Coef04, Coef06, Coef07), Fac2(Coef08, Coef08, Coef10, Coef11);
    tvec4<T> Fac3(Coef12, Coef12, Coef14, Coef15), Fac4(Coef16, Coef16, Coef18, Coef19), Fac5(Coef20, Coef20, Coef22, Coef23);
    tvec4<T> Vec0(m[1][0], m[0][0], m[0][0], m[0][0]), Vec1(m[1][1], m[0][1], m[0][1], m[0][1]);
    tvec4<T> Vec2(m[1][2], m[0][2], m[0][2], m[0][2]), Vec3(m[1][3], m[0][3], m[0][3], m[0][3]);
    tvec4<T> Inv0(Vec1 * Fac0 - Vec2 * Fac1 + Vec3 * Fac2);
    tvec4<T> Inv1(Vec0 * Fac0 - Vec2 * Fac3 + Vec3 * Fac4);
    tvec4<T> Inv2(Vec0 * Fac1 - Vec1 * Fac3 + Vec3 * Fac5);
    tvec4<T> Inv3(Vec0 * Fac2 - Vec1 * Fac4 + Vec2 * Fac5);
    tvec4<T> SignA(+1, -1, +1, -1), SignB(-1, +1, -1, +1);
    tmat4<T> Inverse(Inv0 * SignA, Inv1 * SignB, Inv2 * SignA, Inv3 * SignB);
    tvec4<T> Row0(Inverse[0][0], Inverse[1][0], Inverse[2][0], Inverse[3][0]);
    tvec4<T> Dot0(m[0] * Row0);
    T Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);
    T OneOverDeterminant = T(1) / Dot1;
    return Inverse * OneOverDeterminant;
}
template <typename T> tmat4<T> translate(const tmat4<T>& m, const tvec3<T>& v) {
    tmat4<T> R(m);
    R[3] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2] + m[3];
    return R;
}
template <typename T> tmat4<T> scale(const tmat4<T>& m, const tvec3<T>& v) {
    return tmat4<T>(m[0] * v[0], m[1] * v[1], m[2] * v[2], m[3]);
}
template <typename T> tmat4<T> eulerAngleZYX(T t1, T t2, T t3) {
    T c1 = std::cos(t1), s1 = std::sin(t1), c2 = std::cos(t2), s2 = std::sin(t2), c3 = std::cos(t3), s3 = std::sin(t3);
    tmat4<T> R;
    R[0][0] = c1 * c2;                R[0][1] = c2 * s1;                R[0][2] = -s2;     R[0][3] = 0;
    R[1][0] = c1 * s2 * s3 - c3 * s1; R[1][1] = c1 * c3 + s1 * s2 * s3; R[1][2] = c2 * s3; R[1][3] = 0;
    R[2][0] = s1 * s3 + c1 * c3 * s2; R[2][1] = c3 * s1 * s2 - c1 * s3; R[2][2] = c2 * c3; R[2][3] = 0;
    R[3][0] = 0; R[3][1] = 0; R[3][2] = 0; R[3][3] = 1;
    return R;
}

typedef tmat3<float> mat3;  typedef tmat3<double> dmat3;
typedef tmat4<float> mat4;  typedef tmat4<double> dmat4;

// gtc/type_aligned: same types with 16-byte alignment
struct alignas(8) aligned_vec2 : vec2 { using vec2::vec2; aligned_vec2() {} aligned_vec2(const vec2& v) : vec2(v) {} };
struct alignas(16) aligned_vec3 : vec3 { using vec3::vec3; aligned_vec3() {} aligned_vec3(const vec3& v) : vec3(v) {} };
struct alignas(16) aligned_vec4 : vec4 { using vec4::vec4; aligned_vec4() {} aligned_vec4(const vec4& v) : vec4(v) {} };
struct alignas(16) aligned_mat4 : mat4 { using mat4::mat4; aligned_mat4() {} aligned_mat4(const mat4& m) : mat4(m) {} };
struct alignas(16) aligned_mat3 : mat3 { using mat3::mat3; aligned_mat3() {} aligned_mat3(const mat3& m) : mat3(m) {} };

// packing
inline uint packSnorm2x16(const vec2& v) {
    const vec2 t = round(clamp(v, -1.0f, 1.0f) * 32767.0f);
    const int16_t a = (int16_t)t.x, b = (int16_t)t.y;
    return (uint)(uint16_t)a | ((uint)(uint16_t)b << 16);
}
inline vec2 unpackSnorm2x16(uint p) {
    const int16_t a = (int16_t)(p & 0xFFFFu), b = (int16_t)(p >> 16);
    return clamp(vec2((float)a, (float)b) * 3.0518509475997192297128208258309e-5f, -1.0f, 1.0f);
}
inline uint16_t shim_float_to_half(float f) {            // glm::detail::toFloat16: round to nearest, ties away from zero; overflow to infinity
    uint32_t i; std::memcpy(&i, &f, 4);
    int s = (i >> 16) & 0x00008000;
    int e = ((i >> 23) & 0x000000ff) - (127 - 15);
    int m = i & 0x007fffff;
    if (e <= 0) {
        if (e < -10) return (uint16_t)s;
        m = (m | 0x00800000) >> (1 - e);
        if (m & 0x00001000) m += 0x00002000;
        return (uint16_t)(s | (m >> 13));
    } else if (e == 0xff - (127 - 15)) {
        if (m == 0) return (uint16_t)(s | 0x7c00);
        m >>= 13;
        return (uint16_t)(s | 0x7c00 | m | (m == 0));
    } else {
        if (m & 0x00001000) { m += 0x00002000; if (m & 0x00800000) { m = 0; e += 1; } }
        if (e > 30) return (uint16_t)(s | 0x7c00);
        return (uint16_t)(s | (e << 10) | (m >> 13));
    }
}
inline uint packHalf2x16(const vec2& v) { return (uint)shim_float_to_half(v.x) | ((uint)shim_float_to_half(v.y) << 16); }

}  // namespace glm

// gtx/hash
namespace std {
template <> struct hash<glm::vec2> { size_t operator()(const glm::vec2& v) const { size_t s = 0; hash<float> h; for (int i = 0; i < 2; i++) s ^= h(v[i]) + 0x9e3779b9 + (s << 6) + (s >> 2); return s; } };
template <> struct hash<glm::vec3> { size_t operator()(const glm::vec3& v) const { size_t s = 0; hash<float> h; for (int i = 0; i < 3; i++) s ^= h(v[i]) + 0x9e3779b9 + (s << 6) + (s >> 2); return s; } };
template <> struct hash<glm::vec4> { size_t operator()(const glm::vec4& v) const { size_t s = 0; hash<float> h; for (int i = 0; i < 4; i++) s ^= h(v[i]) + 0x9e3779b9 + (s << 6) + (s >> 2); return s; } };
}
