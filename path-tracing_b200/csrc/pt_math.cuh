// pt_math.cuh — fp32 arithmetic of the reference's hot path, in the reference's evaluation order.
//
// This translation unit is compiled with -fmad=false (no FMA contraction) and nvcc's default IEEE division
// and square root (-prec-div=true -prec-sqrt=true), so every expression below rounds exactly like the
// reference compiled by g++ for plain x86-64 (CMakeLists.txt:4-5, no -march, no -ffast-math).  Where an FMA is
// wanted (BVH slab tests, which only have to be conservative) it is written explicitly with __fmaf_rn.
//
// Vector helper semantics (the reference's glm dependency is unpinned and absent): dot(vec3) sums
// left to right, dot(vec4) = (x+y)+(z+w), length = sqrt(dot), normalize = v * (1/sqrt(dot)),
// cross = (a.y*b.z - b.y*a.z, a.z*b.x - b.z*a.x, a.x*b.y - b.x*a.y), reflect = I - N*dot(N,I)*2.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ptd {

struct V3 {
    float x, y, z;
};

__device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }

// glm::dot(vec3, vec3)
__device__ __forceinline__ float dot3(V3 a, V3 b) {
    const float x = a.x * b.x, y = a.y * b.y, z = a.z * b.z;
    return x + y + z;
}
// glm::dot(vec4, vec4) with both w == 0 (directions and normals, ray.h:13, triangles.h:38)
__device__ __forceinline__ float dot4w0(V3 a, V3 b) {
    const float x = a.x * b.x, y = a.y * b.y, z = a.z * b.z, w = 0.0f;
    return (x + y) + (z + w);
}
__device__ __forceinline__ V3 cross3(V3 a, V3 b) {
    return mk(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
__device__ __forceinline__ float length3(V3 v) { return sqrtf(dot3(v, v)); }
__device__ __forceinline__ V3 normalize3(V3 v) { return v * (1.0f / sqrtf(dot3(v, v))); }
// glm::normalize(vec4) with w == 0: Ray's constructor and Ray::Reflect (ray.h:23,47)
__device__ __forceinline__ V3 normalize4w0(V3 v) { return v * (1.0f / sqrtf(dot4w0(v, v))); }

// ---- Philox4x32-10 (Salmon et al. SC'11).  Stream contract (oracle/philox.h): key = (seed, 0),
// counter = (pixel, sample, slot, 0); slot 0 = camera jitter, slot 1+depth = bounce.
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                               uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}
// u in (0,1): top 24 bits * 2^-24, the zero pattern mapped to 2^-25 (a zero draw is undefined behaviour in the
// reference's Material::Process, material.h:43-48; see oracle/philox.h)
__device__ __forceinline__ float u01(uint32_t w) {
    return fmaxf((float)(w >> 8) * (1.0f / 16777216.0f), 1.0f / 33554432.0f);
}

// ---- sinf / cosf as the reference's libm computes them.
// material.h:93-94 calls std::cos/std::sin(float); the reference links glibc, whose sinf/cosf (since 2.28) is the
// double-precision polynomial scheme of ARM's optimized routines: fast reduction by pi/2 in double, a degree-7 /
// degree-8 minimax polynomial in double, one final rounding to float.  Restated here so that the diffuse direction
// matches the CPU reference bit for bit; tests compare it with the host libm over all 2^24 inputs the renderer can
// produce.  Valid for |y| < 120 (the renderer's argument is 2*pi*xi2 < 6.29).
// The two polynomial forms of glibc's sinf_poly: the sine form for an even quadrant count n, the cosine form (all
// coefficients times -1 when `neg`) for an odd one.
__device__ __forceinline__ float sin_form(double x, double x2) {
    const double x3 = x * x2;
    const double s1 = 0x1.1107605230bc4p-7 + x2 * -0x1.994eb3774cf24p-13;
    const double x7 = x3 * x2;
    const double s = x + x3 * -0x1.555545995a603p-3;
    return (float)(s + x7 * s1);
}
__device__ __forceinline__ float cos_form(double x2, bool neg) {
    const double sg = neg ? -1.0 : 1.0;
    const double x4 = x2 * x2;
    const double c2 = sg * -0x1.6c087e89a359dp-10 + x2 * (sg * 0x1.99343027bf8c3p-16);
    const double c1 = sg * 0x1p0 + x2 * (sg * -0x1.ffffffd0c621cp-2);
    const double x6 = x4 * x2;
    const double c = c1 + x4 * (sg * 0x1.55553e1068f19p-5);
    return (float)(c + x6 * c2);
}
// Branch-free: glibc takes |y| < pi/4 through the same two polynomials with n = 0 (its fast reduction gives exactly
// n = 0 and x unchanged there), and picks sine / cosine form by the parity of n — here both forms are evaluated once
// for all lanes and swapped by parity.  (As three nested branches the 32 lanes of a warp split 3 / 27 and then 15 / 12:
// ncu's source view of k_wf_bounce, profiles/r02_ncu_bounce_before.txt.)  Same operations, same bits: checked against
// the host libm on all 2^24 inputs the renderer can produce.
__device__ __forceinline__ void ref_sincosf(float y, float *sp, float *cp) {
    const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ffu;
    double x = (double)y;
    const double r = x * 0x1.45F306DC9C883p+23;
    const int n = ((int)r + 0x800000) >> 24;
    x = x - (double)n * 0x1.921FB54442D18p0;
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const bool neg = (n & 2) != 0;
    const float S = sin_form(x * sgn, x * x), C = cos_form(x * x, neg);
    const bool even = (n & 1) == 0;
    const bool tiny = top < ((0x39800000u >> 20) & 0x7ffu);  // |y| < 2^-12: glibc returns (y, 1)
    *sp = tiny ? y : (even ? S : C);
    *cp = tiny ? 1.0f : (even ? C : S);
}

// primary ray, main.cpp:126-129 + Ray's constructor ray.h:21-25.  The x / y expressions are evaluated in double
// (int + double, / int, -/+ 0.5f) and narrowed to float by the vec4 constructor.  g++ evaluates the constructor
// arguments right to left, so the first draw of the slot feeds the y jitter (oracle/philox.h).
__device__ __forceinline__ V3 primary_dir(int x, int y, int W, int H, uint32_t pixel, uint32_t sample, uint32_t seed) {
    const uint4 w = philox4x32_10(pixel, sample, 0u, 0u, seed, 0u);
    const double jy = -0.5 + (double)u01(w.x);
    const double jx = -0.5 + (double)u01(w.y);
    V3 d;
    d.x = (float)(((double)x + jx) / (double)W - 0.5);
    d.y = (float)(-((double)y + jy) / (double)H + 0.5);
    d.z = 1.0f;
    return normalize4w0(d);
}

}  // namespace ptd
