// Minimal restatement of the part of glm 0.9.9 (pinned by the reference at commit d214fbaa, enginee/lib/init.sh:39-41;
// NOT vendored and absent from this container) that the reference's hot-path translation units use.
// TEST INFRASTRUCTURE ONLY (oracle/_ref).  Semantics follow the published 0.9.9 scalar (non-SIMD) code paths:
//   dot(a,b)      = a.x*b.x + a.y*b.y (+ a.z*b.z), summed left to right      (detail/func_geometric.inl compute_dot)
//   length(v)     = sqrt(dot(v,v));  length2(v) = dot(v,v)                    (gtx/norm.inl)
//   normalize(v)  = v * inversesqrt(dot(v,v)),  inversesqrt(x) = 1/sqrt(x)    (func_geometric.inl, func_exponential.inl)
//   distance(a,b) = length(b - a);  clamp(x,lo,hi) = min(max(x,lo),hi);  min(a,b) = b<a ? b : a;  max(a,b) = a<b ? b : a
//   vecN / scalar = component-wise division;  vec2(vec3) keeps (x, y)  (implicit unless GLM_FORCE_EXPLICIT_CTOR)
#pragma once
#include <cmath>
#include <cstddef>

namespace glm {

template <int N, typename T> struct vec { T v[N]; T& operator[](int i) { return v[i]; } const T& operator[](int i) const { return v[i]; } };

struct vec3;
struct vec4;

struct vec2 {
    float x, y;
    vec2(): x(0), y(0) {}
    explicit vec2(float s): x(s), y(s) {}
    vec2(float a, float b): x(a), y(b) {}
    template <typename A, typename B> vec2(A a, B b): x(static_cast<float>(a)), y(static_cast<float>(b)) {}
    vec2(const vec3& v);
    float& operator[](int i) { return i == 0 ? x : y; }
    const float& operator[](int i) const { return i == 0 ? x : y; }
    vec2& operator+=(const vec2& o) { x += o.x; y += o.y; return *this; }
    vec2& operator-=(const vec2& o) { x -= o.x; y -= o.y; return *this; }
    vec2& operator*=(float s) { x *= s; y *= s; return *this; }
    vec2& operator/=(float s) { x /= s; y /= s; return *this; }
    vec2& operator*=(const vec2& o) { x *= o.x; y *= o.y; return *this; }
};
inline vec2 operator+(const vec2& a, const vec2& b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(const vec2& a, const vec2& b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(const vec2& a, const vec2& b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator/(const vec2& a, const vec2& b) { return vec2(a.x / b.x, a.y / b.y); }
inline vec2 operator+(const vec2& a, float s) { return vec2(a.x + s, a.y + s); }
inline vec2 operator-(const vec2& a, float s) { return vec2(a.x - s, a.y - s); }
inline vec2 operator*(const vec2& a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator*(float s, const vec2& a) { return vec2(s * a.x, s * a.y); }
inline vec2 operator/(const vec2& a, float s) { return vec2(a.x / s, a.y / s); }
inline vec2 operator/(float s, const vec2& a) { return vec2(s / a.x, s / a.y); }
inline vec2 operator-(const vec2& a) { return vec2(-a.x, -a.y); }
inline bool operator==(const vec2& a, const vec2& b) { return a.x == b.x && a.y == b.y; }
inline bool operator!=(const vec2& a, const vec2& b) { return !(a == b); }

// GLM_FORCE_SWIZZLE member proxy used by comm.cpp:203,213 (`d.pos.xz`)
struct swizzle_xz {
    float a, skip_, b;
    operator vec2() const { return vec2(a, b); }
};
inline vec2 operator-(const swizzle_xz& s, const vec2& b) { return vec2(s) - b; }
inline vec2 operator+(const swizzle_xz& s, const vec2& b) { return vec2(s) + b; }

struct vec3 {
    union {
        struct { float x, y, z; };
        swizzle_xz xz;
    };
    vec3(): x(0), y(0), z(0) {}
    explicit vec3(float s): x(s), y(s), z(s) {}
    vec3(float a, float b, float c): x(a), y(b), z(c) {}
    template <typename A, typename B, typename C> vec3(A a, B b, C c): x(static_cast<float>(a)), y(static_cast<float>(b)), z(static_cast<float>(c)) {}
    vec3(const vec2& v, float c): x(v.x), y(v.y), z(c) {}
    vec3(const vec4& v);
    vec3(const vec3& o): x(o.x), y(o.y), z(o.z) {}
    vec3& operator=(const vec3& o) { x = o.x; y = o.y; z = o.z; return *this; }
    float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    const float& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    vec3& operator+=(const vec3& o) { x += o.x; y += o.y; z += o.z; return *this; }
    vec3& operator-=(const vec3& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
    vec3& operator*=(float s) { x *= s; y *= s; z *= s; return *this; }
    vec3& operator/=(float s) { x /= s; y /= s; z /= s; return *this; }
};
inline vec2::vec2(const vec3& v): x(v.x), y(v.y) {}
inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(const vec3& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(const vec3& a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }
inline bool operator==(const vec3& a, const vec3& b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline bool operator!=(const vec3& a, const vec3& b) { return !(a == b); }

struct vec4 {
    float x, y, z, w;
    vec4(): x(0), y(0), z(0), w(0) {}
    explicit vec4(float s): x(s), y(s), z(s), w(s) {}
    vec4(float a, float b, float c, float d): x(a), y(b), z(c), w(d) {}
    vec4(const vec3& v, float d): x(v.x), y(v.y), z(v.z), w(d) {}
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};
inline vec3::vec3(const vec4& v): x(v.x), y(v.y), z(v.z) {}
inline vec4 operator*(const vec4& a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator+(const vec4& a, const vec4& b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

struct quat { float x, y, z, w; quat(): x(0), y(0), z(0), w(1) {} };

// column-major, m[col][row]
struct mat2x2 {
    vec2 c[2];
    mat2x2() { c[0] = vec2(1, 0); c[1] = vec2(0, 1); }
    mat2x2(const vec2& a, const vec2& b) { c[0] = a; c[1] = b; }
    vec2& operator[](int i) { return c[i]; }
    const vec2& operator[](int i) const { return c[i]; }
};
typedef mat2x2 mat2;
inline mat2x2 transpose(const mat2x2& m) { return mat2x2(vec2(m[0][0], m[1][0]), vec2(m[0][1], m[1][1])); }
// detail/type_mat2x2.inl: m[0][0]*v.x + m[1][0]*v.y,  m[0][1]*v.x + m[1][1]*v.y
inline vec2 operator*(const mat2x2& m, const vec2& v) { return vec2(m[0][0] * v.x + m[1][0] * v.y, m[0][1] * v.x + m[1][1] * v.y); }

struct mat4 {
    vec4 c[4];
    mat4() { for (int i = 0; i < 4; ++i) { c[i] = vec4(0); c[i][i] = 1.f; } }
    explicit mat4(float s) { for (int i = 0; i < 4; ++i) { c[i] = vec4(0); c[i][i] = s; } }
    vec4& operator[](int i) { return c[i]; }
    const vec4& operator[](int i) const { return c[i]; }
};
typedef mat4 mat4x4;
inline mat4 operator*(const mat4& a, const mat4& b) {
    mat4 r(0.f);
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i) r[j][i] = a[0][i] * b[j][0] + a[1][i] * b[j][1] + a[2][i] * b[j][2] + a[3][i] * b[j][3];
    return r;
}
inline vec4 operator*(const mat4& m, const vec4& v) {
    return vec4(m[0][0] * v.x + m[1][0] * v.y + m[2][0] * v.z + m[3][0] * v.w, m[0][1] * v.x + m[1][1] * v.y + m[2][1] * v.z + m[3][1] * v.w,
                m[0][2] * v.x + m[1][2] * v.y + m[2][2] * v.z + m[3][2] * v.w, m[0][3] * v.x + m[1][3] * v.y + m[2][3] * v.z + m[3][3] * v.w);
}
mat4 inverse(const mat4& m);  // general 4x4 inverse (Transform::set_global_mat); defined in shim_impl.cpp
inline vec4 column(const mat4& m, int i) { return m[i]; }
inline mat4 column(const mat4& m, int i, const vec4& v) { mat4 r = m; r[i] = v; return r; }

inline float dot(const vec2& a, const vec2& b) { float tx = a.x * b.x, ty = a.y * b.y; return tx + ty; }
inline float dot(const vec3& a, const vec3& b) { float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z; return tx + ty + tz; }
inline float length2(const vec2& v) { return dot(v, v); }
inline float length2(const vec3& v) { return dot(v, v); }
inline float length(const vec2& v) { return std::sqrt(dot(v, v)); }
inline float length(const vec3& v) { return std::sqrt(dot(v, v)); }
inline float inversesqrt(float x) { return 1.0f / std::sqrt(x); }
inline vec2 normalize(const vec2& v) { return v * inversesqrt(dot(v, v)); }
inline vec3 normalize(const vec3& v) { return v * inversesqrt(dot(v, v)); }
inline float distance(const vec2& a, const vec2& b) { return length(b - a); }
inline float distance(const vec3& a, const vec3& b) { return length(b - a); }
template <typename T> inline T min(T a, T b) { return (b < a) ? b : a; }
template <typename T> inline T max(T a, T b) { return (a < b) ? b : a; }
template <typename T> inline T clamp(T x, T lo, T hi) { return min(max(x, lo), hi); }
inline vec2 clamp(const vec2& v, const vec2& lo, const vec2& hi) { return vec2(clamp(v.x, lo.x, hi.x), clamp(v.y, lo.y, hi.y)); }
inline vec2 clamp(const vec2& v, float lo, float hi) { return vec2(clamp(v.x, lo, hi), clamp(v.y, lo, hi)); }
inline float abs(float x) { return std::fabs(x); }
inline vec2 abs(const vec2& v) { return vec2(std::fabs(v.x), std::fabs(v.y)); }
// func_geometric.inl compute_cross
inline vec3 cross(const vec3& x, const vec3& y) { return vec3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
template <typename T> inline T pi() { return static_cast<T>(3.14159265358979323846264338327950288); }
template <typename T> inline T half_pi() { return static_cast<T>(1.57079632679489661923132169163975144); }
template <typename T> inline T one_over_pi() { return static_cast<T>(0.318309886183790671537767526745028724); }
template <typename T> inline T two_over_pi() { return static_cast<T>(0.636619772367581343075535053490057448); }
inline const float* value_ptr(const vec2& v) { return &v.x; }
inline const float* value_ptr(const vec3& v) { return &v.x; }
inline const float* value_ptr(const mat4& m) { return &m.c[0].x; }

}  // namespace glm
