// TEST INFRASTRUCTURE ONLY (oracle/). Minimal stand-in for the un-vendored, un-pinned
// glm dependency of the reference (CMakeLists.txt:13; includes at ray.h:4-7, triangles.h:6),
// so that the reference's own sources compile unmodified into oracle/_ref/.
//
// glm is absent from /root/reference and from this image, so last-ulp semantics of the vector
// helpers are fixed HERE, following glm 0.9.9's generic (non-SIMD) code paths:
//   dot(vec3) = x*x' + y*y' + z*z'            (left to right)
//   dot(vec4) = (x*x' + y*y') + (z*z' + w*w')
//   length    = sqrt(dot(v,v));  normalize = v * (1/sqrt(dot(v,v)))
//   cross     = (a.y*b.z - b.y*a.z, a.z*b.x - b.z*a.x, a.x*b.y - b.x*a.y)
//   reflect   = I - N * dot(N,I) * 2;   mix = x*(1-a) + y*a
// Only what the reference's call sites use is provided.
#ifndef PT_ORACLE_GLM_SHIM_HPP
#define PT_ORACLE_GLM_SHIM_HPP

#include <array>  // the reference gets <array> transitively (triangles.h:23 uses std::array without including it)
#include <cmath>

namespace glm {

struct vec4;
struct vec3;

struct vec2 {
    union { float x, r; };
    union { float y, g; };
    vec2() : x(0), y(0) {}
    vec2(float s) : x(s), y(s) {}
    vec2(float a, float b) : x(a), y(b) {}
};

struct ivec2 {
    int x, y;
    ivec2() : x(0), y(0) {}
    ivec2(int a, int b) : x(a), y(b) {}
    ivec2(const vec2 &v) : x(static_cast<int>(v.x)), y(static_cast<int>(v.y)) {}
};

struct vec3 {
    union { float x, r; };
    union { float y, g; };
    union { float z, b; };
    vec3() : x(0), y(0), z(0) {}
    vec3(float s) : x(s), y(s), z(s) {}
    vec3(int s) : x(static_cast<float>(s)), y(static_cast<float>(s)), z(static_cast<float>(s)) {}
    vec3(float a, float b_, float c) : x(a), y(b_), z(c) {}
    inline vec3(const vec4 &v);  // truncation, drops w
    vec3 &operator+=(const vec3 &o) { x += o.x; y += o.y; z += o.z; return *this; }
    vec3 &operator*=(const vec3 &o) { x *= o.x; y *= o.y; z *= o.z; return *this; }
};

struct vec4 {
    union { float x, r; };
    union { float y, g; };
    union { float z, b; };
    union { float w, a; };
    vec4() : x(0), y(0), z(0), w(0) {}
    vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(int s) : x(static_cast<float>(s)), y(static_cast<float>(s)), z(static_cast<float>(s)), w(static_cast<float>(s)) {}
    template <class A, class B, class C, class D>
    vec4(A a_, B b_, C c_, D d_)
        : x(static_cast<float>(a_)), y(static_cast<float>(b_)), z(static_cast<float>(c_)), w(static_cast<float>(d_)) {}
    template <class D>
    vec4(const vec3 &v, D d_) : x(v.x), y(v.y), z(v.z), w(static_cast<float>(d_)) {}
    template <class S>
    vec4 &operator*=(S s) { const float f = static_cast<float>(s); x *= f; y *= f; z *= f; w *= f; return *this; }
};

inline vec3::vec3(const vec4 &v) : x(v.x), y(v.y), z(v.z) {}

// ---- vec3 arithmetic ----
inline vec3 operator+(const vec3 &a, const vec3 &b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3 &a, const vec3 &b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(const vec3 &a, const vec3 &b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(const vec3 &a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3 &a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(const vec3 &a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline bool operator==(const vec3 &a, const vec3 &b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline bool operator!=(const vec3 &a, const vec3 &b) { return !(a == b); }

// ---- vec4 arithmetic ----
inline vec4 operator+(const vec4 &a, const vec4 &b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(const vec4 &a, const vec4 &b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline vec4 operator*(const vec4 &a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator*(float s, const vec4 &a) { return vec4(s * a.x, s * a.y, s * a.z, s * a.w); }

// ---- geometric ----
inline float dot(const vec3 &a, const vec3 &b) {
    const vec3 t(a * b);
    return t.x + t.y + t.z;
}
inline float dot(const vec4 &a, const vec4 &b) {
    const vec4 t(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w);
    return (t.x + t.y) + (t.z + t.w);
}
inline vec3 cross(const vec3 &a, const vec3 &b) {
    return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline float length(const vec3 &v) { return std::sqrt(dot(v, v)); }
inline float length(const vec4 &v) { return std::sqrt(dot(v, v)); }
inline float inversesqrt(float v) { return 1.0f / std::sqrt(v); }
inline vec3 normalize(const vec3 &v) { return v * inversesqrt(dot(v, v)); }
inline vec4 normalize(const vec4 &v) { return v * inversesqrt(dot(v, v)); }
inline vec4 reflect(const vec4 &I, const vec4 &N) { return I - N * dot(N, I) * 2.0f; }
inline vec3 reflect(const vec3 &I, const vec3 &N) { return I - N * dot(N, I) * 2.0f; }

// ---- common ----
inline vec3 mix(const vec3 &x, const vec3 &y, float a) { return x * (1.0f - a) + y * a; }
inline vec3 pow(const vec3 &b, const vec3 &e) { return vec3(std::pow(b.x, e.x), std::pow(b.y, e.y), std::pow(b.z, e.z)); }
inline vec3 round(const vec3 &v) { return vec3(std::round(v.x), std::round(v.y), std::round(v.z)); }

}  // namespace glm

#endif
