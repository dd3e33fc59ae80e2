// Minimal from-scratch stand-in for the parts of GLM that the reference's BDPT
// CPU path touches (object.h / object.cpp / ray.h / cpu_bdpt.cpp).  TEST
// INFRASTRUCTURE ONLY: it exists so that the *unmodified* reference sources can
// be compiled into oracle/_ref (glm is an un-vendored, un-pinned dependency of
// the reference, CMakeLists.txt:55, and is absent from this image).
//
// The scalar evaluation order below restates upstream GLM's published scalar
// (non-SIMD) implementations so fp32 results are bit-identical to a real GLM
// build without -ffast-math:
//   dot(a,b)      : tmp = a*b; (tmp.x + tmp.y) + tmp.z           (func_geometric.inl, compute_dot<vec3>)
//   cross(x,y)    : (x.y*y.z - y.y*x.z, x.z*y.x - y.z*x.x, x.x*y.y - y.x*x.y)
//   normalize(v)  : v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x)
//   radians(d)    : d * 0.01745329251994329576923690768489
// Parity at this boundary is otherwise unpinned.
#pragma once
#include <cmath>

namespace glm {

struct vec3 {
    float x, y, z;
    vec3() = default;
    constexpr vec3(float s) : x(s), y(s), z(s) {}
    constexpr vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    template <class A, class B, class C>
    constexpr vec3(A a, B b, C c) : x(float(a)), y(float(b)), z(float(c)) {}
    float &operator[](int i) { return (&x)[i]; }
    const float &operator[](int i) const { return (&x)[i]; }
    vec3 &operator+=(const vec3 &o) { x += o.x; y += o.y; z += o.z; return *this; }
    vec3 &operator-=(const vec3 &o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
    vec3 &operator*=(float s) { x *= s; y *= s; z *= s; return *this; }
    vec3 &operator/=(float s) { x /= s; y /= s; z /= s; return *this; }
};

inline vec3 operator+(const vec3 &a, const vec3 &b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3 &a, const vec3 &b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(const vec3 &a, const vec3 &b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(const vec3 &a, const vec3 &b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator*(const vec3 &a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3 &a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(const vec3 &a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator+(const vec3 &a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3 operator-(const vec3 &a, float s) { return vec3(a.x - s, a.y - s, a.z - s); }
inline vec3 operator-(const vec3 &a) { return vec3(-a.x, -a.y, -a.z); }
inline bool operator==(const vec3 &a, const vec3 &b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

inline float dot(const vec3 &a, const vec3 &b) {
    vec3 tmp(a * b);
    return tmp.x + tmp.y + tmp.z;
}
inline vec3 cross(const vec3 &x, const vec3 &y) {
    return vec3(x.y * y.z - y.y * x.z,
                x.z * y.x - y.z * x.x,
                x.x * y.y - y.x * x.y);
}
inline float inversesqrt(float x) { return 1.0f / std::sqrt(x); }
inline float length(const vec3 &v) { return std::sqrt(dot(v, v)); }
inline vec3 normalize(const vec3 &v) { return v * inversesqrt(dot(v, v)); }

inline float min(float a, float b) { return (b < a) ? b : a; }
inline float max(float a, float b) { return (a < b) ? b : a; }
inline vec3 min(const vec3 &a, const vec3 &b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec3 max(const vec3 &a, const vec3 &b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline float clamp(float x, float lo, float hi) { return min(max(x, lo), hi); }
inline float radians(float deg) { return deg * 0.01745329251994329576923690768489f; }

}  // namespace glm
