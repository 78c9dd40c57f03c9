// pt_bvh.cuh — SoA repack + GPU BVH build: Morton codes -> radix sort -> Karras 2012 binary hierarchy -> bottom-up fit
// -> collapse into an 8-wide tree with quantised child boxes (k_collapse, level by level) -> records in node order.
//
// The reference has no acceleration structure (scene.cpp:114-120 tests every triangle); the LBVH must therefore
// return exactly the brute-force answer of the reference's intersector.  That intersector (triangles.h:48-73) is
// NOT a geometric triangle test: it accepts a plane hit point P when  |PA x PB| + |PA x PC| + |PC x PB| <= S + eps
// with an ABSOLUTE eps, and the plane may come from a rounded `vn` (scene.cpp:102-104).  The accepted set is the
// convex "lens" { P : g(P) <= S + eps },  g = sum of the three parallelogram areas.  Its in-plane cross-section is
// the hexagon with vertices  v_k + (v_k - v_j) * eps/(2S)  (k != j)  — a triangle with 0.01-long edges is dilated by
// a quarter of its size at eps = 1e-4 — and its height over the geometric plane is at most
// sqrt(2 S eps + eps^2) / (|AB| + |BC| + |CA|).  Leaf boxes bound that set (tri_bounds below), not the triangle.
#pragma once
#include <cub/cub.cuh>
#include <float.h>
#include "pt_math.cuh"
#include "pt_scene.cuh"

namespace ptd {

// ---- Triangle constructor arithmetic, triangles.h:27-44 (+ `vn` override, scene.cpp:102-104) -------------------
__device__ __forceinline__ void triangle_setup(V3 v0, V3 v1, V3 v2, bool has_vn, V3 vn, float4 *plane, float *square) {
    const V3 AB = v1 - v0, AC = v2 - v0;
    const V3 c = cross3(AB, AC);
    V3 n = normalize3(c);     // SetNormal(cross(AB, AC))
    *square = length3(c);     // ParallelogramSquare(AB, AC)
    if (has_vn) n = normalize3(vn);
    *plane = make_float4(n.x, n.y, n.z, -dot3(n, v0));
}

__global__ void k_triangle_setup(const float *__restrict__ verts, const uint8_t *__restrict__ has_vn,
                                 const float *__restrict__ vn, int n, float4 *__restrict__ planes,
                                 float *__restrict__ squares) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *v = verts + 9 * (size_t)i;
    const bool hv = has_vn && vn && has_vn[i];
    const V3 nn = hv ? mk(vn[3 * (size_t)i], vn[3 * (size_t)i + 1], vn[3 * (size_t)i + 2]) : mk(0, 0, 0);
    float4 p;
    float s;
    triangle_setup(mk(v[0], v[1], v[2]), mk(v[3], v[4], v[5]), mk(v[6], v[7], v[8]), hv, nn, &p, &s);
    planes[i] = p;
    squares[i] = s;
}

// ---- ordered-int encoding so that float min/max can use integer atomics -------------------------------------
__device__ __forceinline__ int f2ord(float f) {
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

struct BuildScratch {  // scene vertex bounds (ordered ints) lo xyz, hi xyz, large-triangle count, quad units, unbounded count
    int lo[3], hi[3], n_large, n_quads, n_unbounded;
    unsigned direct_lo, direct_hi;  // k_large_pack: large records the watertight mode must test without the scan (SceneView)
    float amax;                     // k_large_pack: largest |coordinate| of the scene (at least 20)
};

__global__ void k_init_scratch(BuildScratch *s) {
    s->lo[0] = s->lo[1] = s->lo[2] = f2ord(FLT_MAX);
    s->hi[0] = s->hi[1] = s->hi[2] = f2ord(-FLT_MAX);
    s->n_large = 0;
    s->n_quads = 0;
    s->n_unbounded = 0;
    s->direct_lo = s->direct_hi = 0u;
    s->amax = 20.0f;
}

__global__ void k_vertex_bounds(const float *__restrict__ verts, int n, BuildScratch *s) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float *v = verts + 9 * (size_t)i;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const float f = v[k];
            if (f == f && fabsf(f) <= FLT_MAX) {
                lo[k % 3] = fminf(lo[k % 3], f);
                hi[k % 3] = fmaxf(hi[k % 3], f);
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        for (int o = 16; o; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(&s->lo[a], f2ord(lo[a]));
            atomicMax(&s->hi[a], f2ord(hi[a]));
        }
    }
}

// Effective eps of the bounds: what the reference's area-sum test can accept in fp32, not in exact arithmetic.
//  * the sums carry rounding proportional to S (4e-6 (S + eps));
//  * the hit point P and the differences P - v_k carry a few ulp of the scene's largest coordinate (amax), so each
//    cross product of two such differences of length r is off by up to ~ r (4 amax + 2 r) 2^-23.  For a triangle of
//    ordinary size this is nothing: the accepted points lie within eps / (2 L) of it.  For a SLIVER (shortest edge L of
//    1e-4 and below in a 30-unit scene) that distance is units, the noise approaches eps itself, and beyond it the three
//    differences become the same float and the computed areas vanish: the reference then accepts the whole plane.
//    robust_eps finds the reach by iteration.  A triangle whose reach runs away is UNBOUNDED: no box, no edge plane — it is tested exactly against every ray (it joins the
//    shared-memory list; a scene with more than kMaxLarge of them is rendered by brute force).
// (Found by test_deep_and_degenerate_trees: clusters of 1e-6-sized triangles, 292 of 40,000 hit IDs off.)
__device__ __forceinline__ float robust_eps(float S, float l_min, float l_max, float eps, float amax, bool *unbounded) {
    const float e0 = eps * 1.01f + 4e-6f * (S + eps);
    // reach r of the accepted set beyond the triangle and the noise at that reach feed each other: iterate
    // r = 1.05 (e0 + 2 noise(r)) / (2 l_min), noise(r) = 3 * 2^-22 * (r + l_max) (4 amax + 2 (r + l_max)); for an ordinary
    // triangle this settles after a step or two, for a sliver it runs away
    float r = 0.0f, noise = 0.0f;
    bool grows = false;
#pragma unroll 1
    for (int it = 0; it < 8; ++it) {
        const float f = r + l_max;
        noise = 3.0f * 2.4e-7f * f * (4.0f * amax + 2.0f * f);
        const float rn = 1.05f * (e0 + 2.0f * noise) / (2.0f * l_min);
        grows = !(rn <= 1.1f * r + 1e-7f);  // still growing by more than 10 % (or NaN / inf)
        r = rn;
        if (!(r <= 8.0f * amax)) break;
    }
    *unbounded = !(r <= 8.0f * amax) || grows;
    return e0 + 2.0f * noise;
}

// Conservative box of everything the reference's Intersect can accept for this triangle (see header comment).
// *unbounded: no finite box holds that set (a sliver, robust_eps): lo / hi are then the vertex box.
__device__ __forceinline__ void tri_bounds(V3 v0, V3 v1, V3 v2, float4 plane, float S, float eps, float pad, float amax,
                                           V3 *lo, V3 *hi, bool *unbounded) {
    const V3 v[3] = {v0, v1, v2};
    const float l0 = length3(v1 - v0), l1 = length3(v2 - v1), l2 = length3(v0 - v2);
    *unbounded = false;
    V3 l = mk(fminf(v0.x, fminf(v1.x, v2.x)), fminf(v0.y, fminf(v1.y, v2.y)), fminf(v0.z, fminf(v1.z, v2.z)));
    V3 h = mk(fmaxf(v0.x, fmaxf(v1.x, v2.x)), fmaxf(v0.y, fmaxf(v1.y, v2.y)), fmaxf(v0.z, fmaxf(v1.z, v2.z)));
    if (!(S > 0.0f) || !(S <= FLT_MAX) || !(plane.x == plane.x)) {  // degenerate: NaN plane, never accepted here
        *lo = l; *hi = h;
        return;
    }
    const float e = robust_eps(S, fminf(l0, fminf(l1, l2)), fmaxf(l0, fmaxf(l1, l2)), eps, amax, unbounded);
    if (*unbounded) {
        *lo = l; *hi = h;
        return;
    }
    const float lam = e / (2.0f * S);
    const V3 ng = normalize3(cross3(v1 - v0, v2 - v0));
    const V3 ns = mk(plane.x, plane.y, plane.z);
    const float nsng = dot3(ns, ng);
    float zmax = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            if (j == k) continue;
            const V3 q = v[k] + (v[k] - v[j]) * lam;
            l = mk(fminf(l.x, q.x), fminf(l.y, q.y), fminf(l.z, q.z));
            h = mk(fmaxf(h.x, q.x), fmaxf(h.y, q.y), fmaxf(h.z, q.z));
            // height over the geometric plane of the stored-plane point above q
            const float z = fabsf((dot3(ns, q) + plane.w) / nsng);
            zmax = fmaxf(zmax, z == z ? z : FLT_MAX);
        }
    const float per = l0 + l1 + l2;
    const float zlens = sqrtf(2.0f * S * e + e * e) / per;
    const float zp = fminf(zmax * 1.001f + 1e-7f, zlens * 1.001f);
    const V3 ex = mk(fabsf(ng.x) * zp + pad, fabsf(ng.y) * zp + pad, fabsf(ng.z) * zp + pad);
    *lo = l - ex;
    *hi = h + ex;
}

__device__ __forceinline__ uint64_t spread21(uint32_t v) {  // 21 bits -> every third bit of 63
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// per triangle: dilated box, 63-bit Morton key of its centre; bit 63 marks a "large" triangle (sorted last), the all-ones
// key an unbounded one (last of all)
__global__ void k_tri_boxes_keys(const float *__restrict__ verts, const float4 *__restrict__ planes,
                                 const float *__restrict__ squares, int n, float eps, BuildScratch *s,
                                 float4 *__restrict__ box_lo, float4 *__restrict__ box_hi,
                                 uint64_t *__restrict__ keys, int *__restrict__ vals, int all_large) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 slo = mk(ord2f(s->lo[0]), ord2f(s->lo[1]), ord2f(s->lo[2]));
    const V3 shi = mk(ord2f(s->hi[0]), ord2f(s->hi[1]), ord2f(s->hi[2]));
    const float amax = fmaxf(20.0f, fmaxf(fmaxf(fmaxf(fabsf(slo.x), fabsf(shi.x)), fmaxf(fabsf(slo.y), fabsf(shi.y))),
                                          fmaxf(fabsf(slo.z), fabsf(shi.z))));
    const float pad = amax * (32.0f * 1.1920929e-7f);  // 32 ulp of the largest coordinate in play
    const float *v = verts + 9 * (size_t)i;
    V3 lo, hi;
    bool unbounded;
    tri_bounds(mk(v[0], v[1], v[2]), mk(v[3], v[4], v[5]), mk(v[6], v[7], v[8]), planes[i], squares[i], eps, pad, amax, &lo,
               &hi, &unbounded);
    box_lo[i] = make_float4(lo.x, lo.y, lo.z, 0.0f);
    box_hi[i] = make_float4(hi.x, hi.y, hi.z, 0.0f);
    const V3 ext = shi - slo;
    const float emax = fmaxf(ext.x, fmaxf(ext.y, ext.z));
    const V3 c = (lo + hi) * 0.5f;
    const float sc = emax > 0.0f ? 2097151.0f / emax : 0.0f;
    const uint32_t qx = (uint32_t)fminf(fmaxf((c.x - slo.x) * sc, 0.0f), 2097151.0f);
    const uint32_t qy = (uint32_t)fminf(fmaxf((c.y - slo.y) * sc, 0.0f), 2097151.0f);
    const uint32_t qz = (uint32_t)fminf(fmaxf((c.z - slo.z) * sc, 0.0f), 2097151.0f);
    uint64_t key = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
    const V3 be = hi - lo;
    const bool large = all_large || unbounded || fmaxf(be.x, fmaxf(be.y, be.z)) > 0.25f * emax;
    if (large) {
        key |= 1ull << 63;
        atomicAdd(&s->n_large, 1);
    }
    if (unbounded) {  // sorted behind every other large triangle: the shared-memory list takes the LAST kMaxLarge
        key = ~0ull;
        atomicAdd(&s->n_unbounded, 1);
    }
    keys[i] = key;
    vals[i] = i;
}

// ---- Karras 2012: one thread per internal node over the sorted keys ------------------------------------------
__device__ __forceinline__ int lbvh_delta(const uint64_t *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ b));
}

__global__ void k_karras(const uint64_t *__restrict__ keys, int n, int2 *__restrict__ children,
                         int *__restrict__ parent, int2 *__restrict__ range) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lbvh_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    // children: internal node index k, or leaf encoded as ~k (k = sorted position)
    const int c0 = (first == gamma) ? ~gamma : gamma;
    const int c1 = (last == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    children[i] = make_int2(c0, c1);
    range[i] = make_int2(first, last);
    if (c0 >= 0) parent[c0] = i; else parent[n - 1 + gamma] = i;
    if (c1 >= 0) parent[c1] = i; else parent[n - 1 + gamma + 1] = i;
    if (i == 0) parent[0] = -1;
}

// ---- which subtrees become the children of a wide node: the cost-optimal collapse of Ylitie, Karras, Laine 2017 ------
// C(n, i) = smallest SAH cost of the binary subtree n represented as at most i roots of the wide tree (i = 1..7):
//   C(n, 1) = min( leaf: A_n P_n c_prim  (P_n <= kLeafMax triangles),  inner: A_n c_node + D(n, 8) )
//   C(n, i) = min( D(n, i), C(n, i - 1) ),   D(n, j) = min over 0 < k < j of C(left, k) + C(right, j - k)
// A = half surface area of the (dilated, padded) box.  Filled bottom-up together with the boxes; k_collapse reads the
// decisions back top-down.  dec[8 n + i]: i = 1: 0xff leaf, else the k of D(n, 8); i = 2..7: 0 = "as with i - 1", else
// the k of D(n, i); dec[8 n + 0]: the k of D(n, 8) whatever C(n, 1) chose (a root must be an inner node).
#ifndef PT_SAH_NODE
#define PT_SAH_NODE 1.0f   // cost of entering a wide node ...
#endif
#ifndef PT_SAH_PRIM
#define PT_SAH_PRIM 0.3f   // ... against scanning one triangle
#endif
struct CollapseDP {
    float *cost;         // 7 per inner binary node: C(n, 1..7)
    unsigned char *dec;  // 8 per inner binary node
};
__device__ __forceinline__ float half_area(float4 lo, float4 hi) {
    const float ex = hi.x - lo.x, ey = hi.y - lo.y, ez = hi.z - lo.z;
    return ex * ey + ey * ez + ez * ex;
}

// bottom-up fit: the second thread to reach a node merges its children's boxes and fills the node's cost table
__global__ void k_fit(const int *__restrict__ vals, const float4 *__restrict__ box_lo,
                      const float4 *__restrict__ box_hi, int n, const int2 *__restrict__ children,
                      const int2 *__restrict__ range, const int *__restrict__ parent, int *__restrict__ flags,
                      float4 *__restrict__ node_lo, float4 *__restrict__ node_hi, CollapseDP dp) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    // leaf k's box is box_*[vals[k]]
    int p = parent[n - 1 + k];
    while (p >= 0) {
        if (atomicAdd(&flags[p], 1) == 0) return;  // first arrival: the sibling is not done yet
        __threadfence();
        const int2 c = children[p];
        const float4 l0 = c.x >= 0 ? __ldcg(&node_lo[c.x]) : box_lo[vals[~c.x]];
        const float4 h0 = c.x >= 0 ? __ldcg(&node_hi[c.x]) : box_hi[vals[~c.x]];
        const float4 l1 = c.y >= 0 ? __ldcg(&node_lo[c.y]) : box_lo[vals[~c.y]];
        const float4 h1 = c.y >= 0 ? __ldcg(&node_hi[c.y]) : box_hi[vals[~c.y]];
        const float4 lo = make_float4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z), 0.0f);
        const float4 hi = make_float4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), 0.0f);
        node_lo[p] = lo;
        node_hi[p] = hi;
        {   // cost table of p from its children's (a single triangle costs its box times c_prim, whatever the budget)
            float cl[8], cr[8];
            const float al = half_area(l0, h0) * PT_SAH_PRIM, ar = half_area(l1, h1) * PT_SAH_PRIM;
            for (int i = 1; i <= 7; ++i) {
                cl[i] = c.x >= 0 ? __ldcg(&dp.cost[7 * (size_t)c.x + i - 1]) : al;
                cr[i] = c.y >= 0 ? __ldcg(&dp.cost[7 * (size_t)c.y + i - 1]) : ar;
            }
            float D[9];
            unsigned char K[9];
            for (int j = 2; j <= 8; ++j) {
                float best = FLT_MAX;
                int bk = 1;
                for (int kk = 1; kk < j; ++kk) {
                    if (kk > 7 || j - kk > 7) continue;
                    const float v = cl[kk] + cr[j - kk];
                    if (v < best) { best = v; bk = kk; }
                }
                D[j] = best;
                K[j] = (unsigned char)bk;
            }
            const float A = half_area(lo, hi);
            const int cnt = range[p].y - range[p].x + 1;
            const float c_leaf = cnt <= kLeafMax ? A * (float)cnt * PT_SAH_PRIM : FLT_MAX;
            const float c_inner = A * PT_SAH_NODE + D[8];
            unsigned char *dec = dp.dec + 8 * (size_t)p;
            float *cost = dp.cost + 7 * (size_t)p;
            dec[0] = K[8];
            float cprev = fminf(c_leaf, c_inner);
            dec[1] = c_leaf <= c_inner ? 0xff : K[8];
            cost[0] = cprev;
            for (int i = 2; i <= 7; ++i) {
                const bool split = D[i] < cprev;
                dec[i] = split ? K[i] : 0;
                cprev = split ? D[i] : cprev;
                cost[i - 1] = cprev;
            }
        }
        __threadfence();
        p = parent[p];
    }
}

// triangle records in sorted order (vals == nullptr: caller order)
__global__ void k_pack_tris(const float *__restrict__ verts, const float4 *__restrict__ planes,
                            const float *__restrict__ squares, const int *__restrict__ mat_ids,
                            const int *__restrict__ vals, int n, float4 *__restrict__ tris) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int i = vals ? vals[k] : k;
    const float *v = verts + 9 * (size_t)i;
    float4 *o = tris + 4 * (size_t)k;
    o[0] = planes[i];
    o[1] = make_float4(v[0], v[1], v[2], squares[i]);
    o[2] = make_float4(v[3], v[4], v[5], __int_as_float(i));
    o[3] = make_float4(v[6], v[7], v[8], __int_as_float(mat_ids[i]));
}

// ---- shared-memory image of the large triangles (layout: pt_scene.cuh) -----------------------------------------
// EDGE PLANES (m, w): m is the in-plane inward unit normal of an edge, and every point the reference's test can accept
// satisfies m . P + w >= 0.  w already contains the dilation of that edge, eps'/(2 L) (the accepted hexagon of the
// header comment is bounded by the three offset edge lines; a displacement along the triangle's normal does not
// change m . P), plus 64 ulp of the largest coordinate for the fp32 evaluation.
struct LargeTri {
    V3 p[3];
    float4 plane;
    float S;
    int idx, mat;
};

__device__ __forceinline__ float4 edge_plane(const LargeTri &t, int j, V3 ng, bool ok, float e, float slack) {
    const V3 a = t.p[j], b = t.p[(j + 1) % 3], c = t.p[(j + 2) % 3];
    const V3 ab = b - a;
    const float L = length3(ab);
    V3 m = normalize3(cross3(ng, ab));
    if (dot3(m, c - a) < 0.0f) m = m * -1.0f;
    const float w = -dot3(m, a) + 1.05f * e / (2.0f * L) + slack;
    // degenerate triangle: a plane that accepts everything (the exact test decides)
    return (ok && L > 0.0f && m.x == m.x) ? make_float4(m.x, m.y, m.z, w) : make_float4(0.f, 0.f, 0.f, 1.0f);
}
__device__ __forceinline__ bool same_point(V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
// index of the edge (p[j], p[j+1]) of `t` whose two end points are both vertices of `u`; -1 unless exactly two are shared
__device__ __forceinline__ int shared_edge(const LargeTri &t, const LargeTri &u) {
    bool in[3];
    int n = 0;
    for (int k = 0; k < 3; ++k) {
        in[k] = same_point(t.p[k], u.p[0]) || same_point(t.p[k], u.p[1]) || same_point(t.p[k], u.p[2]);
        n += in[k];
    }
    if (n != 2) return -1;
    return in[0] && in[1] ? 0 : in[1] && in[2] ? 1 : 2;
}
// effective eps of triangle t (robust_eps); *unbounded: its edge planes must accept everything
__device__ __forceinline__ float tri_e(const LargeTri &t, float eps, float amax, bool *unbounded) {
    const float l0 = length3(t.p[1] - t.p[0]), l1 = length3(t.p[2] - t.p[1]), l2 = length3(t.p[0] - t.p[2]);
    return robust_eps(t.S, fminf(l0, fminf(l1, l2)), fmaxf(l0, fmaxf(l1, l2)), eps, amax, unbounded);
}
// largest value of the plane function f over the region of t's plane the reference can accept (the dilated hexagon
// v_k + (v_k - v_j) e / (2 S)); with `absolute`, of |f|
__device__ __forceinline__ float max_over_hexagon(const LargeTri &t, float eps, float amax, float4 f, bool absolute) {
    bool unb;
    const float lam = 1.05f * tri_e(t, eps, amax, &unb) / (2.0f * t.S);
    if (unb) return FLT_MAX;
    float best = -FLT_MAX;
    for (int k = 0; k < 3; ++k)
        for (int j = 0; j < 3; ++j) {
            if (j == k) continue;
            const V3 q = t.p[k] + (t.p[k] - t.p[j]) * lam;
            float v = f.x * q.x + f.y * q.y + f.z * q.z + f.w;
            if (absolute) v = fabsf(v);
            best = fmaxf(best, v == v ? v : FLT_MAX);
        }
    return best;
}

// Can a and b be scanned as one unit?  They must share an edge, carry stored planes that are (nearly) their geometric
// planes and (nearly) each other's, and form a convex quadrilateral.
__device__ bool quad_eligible(const LargeTri &a, const LargeTri &b, float eps, float amax) {
    if (!(a.S > 0.0f && a.S <= FLT_MAX && b.S > 0.0f && b.S <= FLT_MAX)) return false;
    const int ja = shared_edge(a, b), jb = shared_edge(b, a);
    if (ja < 0 || jb < 0) return false;
    const V3 nga = normalize3(cross3(a.p[1] - a.p[0], a.p[2] - a.p[0])), ngb = normalize3(cross3(b.p[1] - b.p[0], b.p[2] - b.p[0]));
    const V3 nsa = mk(a.plane.x, a.plane.y, a.plane.z), nsb = mk(b.plane.x, b.plane.y, b.plane.z);
    if (!(fabsf(dot3(nsa, nga)) >= 0.999999f && fabsf(dot3(nsb, ngb)) >= 0.999999f && dot3(nsa, nsb) >= 0.999999f)) return false;
    if (!(max_over_hexagon(b, eps, amax, a.plane, true) <= 1e-4f && max_over_hexagon(a, eps, amax, b.plane, true) <= 1e-4f)) return false;
    // convex: the far vertex of each lies strictly inside the two outer edges of the other
    const V3 fa = a.p[(ja + 2) % 3], fb = b.p[(jb + 2) % 3];
    const float margin = 1e-5f * amax;
    for (int j = 0; j < 3; ++j) {
        if (j != ja) {
            const float4 pl = edge_plane(a, j, nga, true, 0.0f, 0.0f);
            if (!(pl.x * fb.x + pl.y * fb.y + pl.z * fb.z + pl.w > margin)) return false;
        }
        if (j != jb) {
            const float4 pl = edge_plane(b, j, ngb, true, 0.0f, 0.0f);
            if (!(pl.x * fa.x + pl.y * fa.y + pl.z * fa.z + pl.w > margin)) return false;
        }
    }
    return true;
}

// One block of kMaxLarge threads: pairs the large triangles (sorted positions [nb, nb + n_large) of `vals`) into quad
// scan units, REORDERS that tail of `vals` (quads first, partners adjacent: k_pack_tris runs afterwards and must see
// the same order), writes the shared-memory image and the number of quads (s->n_quads).
__global__ void __launch_bounds__(kMaxLarge)
k_large_pack(int *__restrict__ vals, const float *__restrict__ verts, const float4 *__restrict__ planes,
             const float *__restrict__ squares, const int *__restrict__ mat_ids, int nb, int n_large, float eps,
             BuildScratch *s, float4 *__restrict__ lrecs) {
    __shared__ LargeTri tri[kMaxLarge];
    __shared__ int order[kMaxLarge], partner[kMaxLarge], n_quads_s;
    const int k = threadIdx.x;
    const float amax = fmaxf(20.0f, fmaxf(fmaxf(fmaxf(fabsf(ord2f(s->lo[0])), fabsf(ord2f(s->hi[0]))),
                                                fmaxf(fabsf(ord2f(s->lo[1])), fabsf(ord2f(s->hi[1])))),
                                          fmaxf(fabsf(ord2f(s->lo[2])), fabsf(ord2f(s->hi[2])))));
    if (k < n_large) {
        const int i = vals[nb + k];
        const float *v = verts + 9 * (size_t)i;
        tri[k].p[0] = mk(v[0], v[1], v[2]); tri[k].p[1] = mk(v[3], v[4], v[5]); tri[k].p[2] = mk(v[6], v[7], v[8]);
        tri[k].plane = planes[i];
        tri[k].S = squares[i];
        tri[k].idx = i;
        tri[k].mat = mat_ids[i];
        partner[k] = -1;
    }
    __syncthreads();
    if (k == 0) {  // greedy matching, at most 64 x 64 / 2 cheap tests
        int nq = 0, pos = 0;
        for (int a = 0; a < n_large; ++a) {
            if (partner[a] >= 0) continue;
            for (int b = a + 1; b < n_large; ++b)
                if (partner[b] < 0 && quad_eligible(tri[a], tri[b], eps, amax)) { partner[a] = b; partner[b] = a; break; }
        }
        for (int a = 0; a < n_large; ++a)
            if (partner[a] > a) { order[pos++] = a; order[pos++] = partner[a]; ++nq; }
        for (int a = 0; a < n_large; ++a)
            if (partner[a] < 0) order[pos++] = a;
        n_quads_s = nq;
        s->n_quads = nq;
    }
    __syncthreads();
    const int nq = n_quads_s, n_single = n_large - 2 * nq;
    const float slack = amax * (64.0f * 1.1920929e-7f);
    if (k == 0) s->amax = amax;
    if (k < n_large) {  // record rows in the new order
        const LargeTri &t = tri[order[k]];
        {   // Watertight mode (pt_trace.cuh large_phase): the scan units bound what the REFERENCE accepts around the STORED
            // plane.  They also bound the geometric triangle (the region is a dilation of it) when the stored plane is
            // the geometric one: unit normals agree and every vertex lies within 1e-5 amax of the plane.  Otherwise
            // (a `vn` plane, a degenerate triangle) the record is tested directly.
            const V3 ng = normalize3(cross3(t.p[1] - t.p[0], t.p[2] - t.p[0])), ns = mk(t.plane.x, t.plane.y, t.plane.z);
            float dev = 0.0f;
            for (int j = 0; j < 3; ++j) dev = fmaxf(dev, fabsf(ns.x * t.p[j].x + ns.y * t.p[j].y + ns.z * t.p[j].z + t.plane.w));
            bool unb;
            tri_e(t, eps, amax, &unb);
            const bool geometric = t.S > 0.0f && t.S <= FLT_MAX && !unb && fabsf(dot3(ns, ng)) >= 0.999999f && dev <= 1e-5f * amax;
            if (!geometric) atomicOr(k < 32 ? &s->direct_lo : &s->direct_hi, 1u << (k & 31));
        }
        float4 *o = lrecs + 4 * (size_t)k;
        o[0] = t.plane;
        o[1] = make_float4(t.p[0].x, t.p[0].y, t.p[0].z, t.S);
        o[2] = make_float4(t.p[1].x, t.p[1].y, t.p[1].z, __int_as_float(t.idx));
        o[3] = make_float4(t.p[2].x, t.p[2].y, t.p[2].z, __int_as_float(t.mat));
        vals[nb + k] = t.idx;
    }
    if (k < nq) {  // quad unit k: records 2k (A) and 2k + 1 (B)
        const LargeTri &a = tri[order[2 * k]], &b = tri[order[2 * k + 1]];
        const int ja = shared_edge(a, b), jb = shared_edge(b, a);
        const V3 nga = normalize3(cross3(a.p[1] - a.p[0], a.p[2] - a.p[0])), ngb = normalize3(cross3(b.p[1] - b.p[0], b.p[2] - b.p[0]));
        bool ua, ub;  // (an unbounded triangle never pairs: quad_eligible, max_over_hexagon)
        const float ea = tri_e(a, eps, amax, &ua), eb = tri_e(b, eps, amax, &ub);
        float4 *o = lrecs + 4 * (size_t)n_large + kQuadRows * (size_t)k;
        o[0] = a.plane;
        o[1] = edge_plane(a, (ja + 1) % 3, nga, true, ea, slack);
        o[2] = edge_plane(a, (ja + 2) % 3, nga, true, ea, slack);
        o[3] = edge_plane(b, (jb + 1) % 3, ngb, true, eb, slack);
        o[4] = edge_plane(b, (jb + 2) % 3, ngb, true, eb, slack);
        const float4 diag = edge_plane(a, ja, nga, true, ea, slack);
        o[5] = diag;
        // B is scanned with A's plane: its own hit distance differs by at most (distance of B's hit point from A's
        // plane + the rounding of both numerators) / |d . n| and by |t| times the rounding of the denominators
        const float dev = max_over_hexagon(b, eps, amax, a.plane, true);
        o[6] = make_float4(1.5f * dev + amax * 1e-6f, 1e-6f, max_over_hexagon(b, eps, amax, diag, false) + slack, 0.0f);
    }
    if (k < n_single) {
        const LargeTri &t = tri[order[2 * nq + k]];
        const V3 ng = normalize3(cross3(t.p[1] - t.p[0], t.p[2] - t.p[0]));
        bool unb;
        const float e = tri_e(t, eps, amax, &unb);
        const bool ok = t.S > 0.0f && t.S <= FLT_MAX && ng.x == ng.x && !unb;
        float4 *o = lrecs + 4 * (size_t)n_large + kQuadRows * (size_t)nq + kSingleRows * (size_t)k;
        o[0] = t.plane;
        for (int j = 0; j < 3; ++j) o[1 + j] = edge_plane(t, j, ng, ok, e, slack);
    }
}
// ---- collapse of the binary tree into the 8-wide tree (layout: pt_scene.cuh) -----------------------------------------
// A binary ref is an inner node index (>= 0) or ~k for the leaf at sorted position k.
struct WideItem {
    int node;  // index of the wide node to emit
    int ref;   // binary ref whose subtree it covers
};
struct CollapseCounters {
    int n_nodes;    // wide nodes allocated so far
    int n_out;      // work items appended for the next level
    int tri_cursor; // triangle records placed so far
    int pad;
};

struct BinTree {
    const int2 *children, *range;
    const unsigned char *dec;  // CollapseDP::dec
    const int *vals;
    const float4 *box_lo, *box_hi, *node_lo, *node_hi;
    __device__ __forceinline__ int count(int ref) const { return ref >= 0 ? range[ref].y - range[ref].x + 1 : 1; }
    // does the subtree become a wide node of its own (the cost table's choice for one root), or a leaf child?
    __device__ __forceinline__ bool is_inner(int ref) const { return ref >= 0 && dec[8 * (size_t)ref + 1] != 0xff; }
    __device__ __forceinline__ int first(int ref) const { return ref >= 0 ? range[ref].x : ~ref; }
    __device__ __forceinline__ void box(int ref, V3 *lo, V3 *hi) const {
        const float4 l = ref >= 0 ? node_lo[ref] : box_lo[vals[~ref]], h = ref >= 0 ? node_hi[ref] : box_hi[vals[~ref]];
        *lo = mk(l.x, l.y, l.z);
        *hi = mk(h.x, h.y, h.z);
    }
    __device__ __forceinline__ float area(int ref) const {
        V3 lo, hi;
        box(ref, &lo, &hi);
        const V3 e = hi - lo;
        return e.x * e.y + e.y * e.z + e.z * e.x;
    }
};

// decode of a quantised plane as the builder checks it: p + q s with one rounding
__device__ __forceinline__ float q_decode(int q, float s, float p) { return __fmaf_rn((float)q, s, p); }

// One thread per wide node of the current level.  Its children are the cost-optimal cut of the binary subtree into at
// most 8 roots (the decision table k_fit filled); those the table made leaves (<= kLeafMax triangles, contiguous in the
// sorted order) become leaf children, the others inner children, numbered in slot order from child_base and queued for
// the next level.  Quantisation: q_lo = floor, q_hi = ceil on the grid p + q s, s = extent / 255, pushed outwards until
// the DECODED plane (q_decode) encloses the child box with a margin of 2^-7 of a grid step: node8_visit
// (pt_trace.cuh) evaluates (p + q s - o) / d in a different association (one FMA on a 2^15-scaled grid) whose rounding
// stays below 2^-9 of a step plus a few ulp of |p - o| / |d|; the latter are covered by the 32-ulp padding every
// triangle box already carries (k_tri_boxes_keys).
__global__ void __launch_bounds__(128)
k_collapse(const WideItem *__restrict__ in, int n_in, WideItem *__restrict__ out, CollapseCounters *ctr, BinTree bt,
           float4 *__restrict__ nodes, int *__restrict__ perm) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_in) return;
    const WideItem item = in[t];
    // The children: read the cost-optimal cut below item.ref off the decision table (k_fit).  A work list of
    // (subtree, budget of roots) starts with the two binary children sharing 8 slots as D(n, 8) found best.
    int refs[8];
    int n = 0;
    if (item.ref >= 0) {
        int st_ref[16], st_bud[16], sp = 0;
        const int2 c0 = bt.children[item.ref];
        const int k0 = bt.dec[8 * (size_t)item.ref];
        st_ref[sp] = c0.y; st_bud[sp++] = 8 - k0;
        st_ref[sp] = c0.x; st_bud[sp++] = k0;
        while (sp > 0 && n < 8) {
            const int m = st_ref[--sp];
            int j = st_bud[sp];
            if (m < 0) { refs[n++] = m; continue; }  // a single triangle
            const unsigned char *dec = bt.dec + 8 * (size_t)m;
            while (j > 1 && dec[j] == 0) --j;        // "as with one root fewer"
            if (j == 1) { refs[n++] = m; continue; } // m itself is a child: a leaf or an inner node (dec[1])
            const int2 c = bt.children[m];
            const int k = dec[j];
            st_ref[sp] = c.y; st_bud[sp++] = j - k;
            st_ref[sp] = c.x; st_bud[sp++] = k;
        }
    } else {  // a one-triangle tree
        refs[n++] = item.ref;
    }
    // child boxes, node box
    V3 clo[8], chi[8];
    V3 lo = mk(FLT_MAX, FLT_MAX, FLT_MAX), hi = mk(-FLT_MAX, -FLT_MAX, -FLT_MAX);
    for (int i = 0; i < n; ++i) {
        bt.box(refs[i], &clo[i], &chi[i]);
        lo = mk(fminf(lo.x, clo[i].x), fminf(lo.y, clo[i].y), fminf(lo.z, clo[i].z));
        hi = mk(fmaxf(hi.x, chi[i].x), fmaxf(hi.y, chi[i].y), fmaxf(hi.z, chi[i].z));
    }
    // slot assignment: greedy on dot(child centre - node centre, octant direction of the slot)
    int slot_of_child[8], child_in_slot[8];
    for (int i = 0; i < 8; ++i) { slot_of_child[i] = -1; child_in_slot[i] = -1; }
    const V3 ctr_n = (lo + hi) * 0.5f;
    for (int round = 0; round < n; ++round) {
        int bc = -1, bs = -1;
        float bv = -FLT_MAX;
        for (int i = 0; i < n; ++i) {
            if (slot_of_child[i] >= 0) continue;
            const V3 c = (clo[i] + chi[i]) * 0.5f - ctr_n;
            for (int sl = 0; sl < 8; ++sl) {
                if (child_in_slot[sl] >= 0) continue;
                const float v = ((sl & 1) ? c.x : -c.x) + ((sl & 2) ? c.y : -c.y) + ((sl & 4) ? c.z : -c.z);
                if (v > bv || bc < 0) { bv = v; bc = i; bs = sl; }
            }
        }
        slot_of_child[bc] = bs;
        child_in_slot[bs] = bc;
    }
    // grid: the origin one step below the node box and 253 steps across it, so that every child plane has a grid
    // plane strictly outside it on either side (the outward margins below)
    float p[3], sc[3];
    const float lo3[3] = {lo.x, lo.y, lo.z}, hi3[3] = {hi.x, hi.y, hi.z};
    for (int a = 0; a < 3; ++a) {
        float sa = fmaxf((hi3[a] - lo3[a]) * (1.0f / 252.0f), 1e-30f);
        p[a] = lo3[a] - sa;
        while (q_decode(255, sa, p[a]) < hi3[a] + sa * 0.0078125f) sa *= 1.0000005f;
        sc[a] = sa;
    }
    // inner / leaf children in slot order
    int n_inner = 0, n_tri = 0;
    for (int sl = 0; sl < 8; ++sl) {
        const int i = child_in_slot[sl];
        if (i < 0) continue;
        const int c = bt.count(refs[i]);
        if (bt.is_inner(refs[i])) ++n_inner; else n_tri += c;
    }
    const int child_base = n_inner ? atomicAdd(&ctr->n_nodes, n_inner) : 0;
    const int out_base = n_inner ? atomicAdd(&ctr->n_out, n_inner) : 0;
    const int tri_base = n_tri ? atomicAdd(&ctr->tri_cursor, n_tri) : 0;
    unsigned meta[2] = {0u, 0u}, imask = 0u;
    unsigned q[6][2] = {{0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}};  // qlo x y z, qhi x y z; slots 0..3, 4..7
    int k_inner = 0, k_tri = 0;
    for (int sl = 0; sl < 8; ++sl) {
        const int i = child_in_slot[sl];
        const int sh = 8 * (sl & 3), w = sl >> 2;
        if (i < 0) {  // empty slot: meta 0 contributes no hit bits whatever the box test says
            for (int a = 0; a < 3; ++a) q[a][w] |= 255u << sh;
            continue;
        }
        const int c = bt.count(refs[i]);
        if (bt.is_inner(refs[i])) {
            meta[w] |= (0x38u + (unsigned)sl) << sh;
            imask |= 1u << sl;
            out[out_base + k_inner] = WideItem{child_base + k_inner, refs[i]};
            ++k_inner;
        } else {
            meta[w] |= ((((1u << c) - 1u) << 5) | (unsigned)k_tri) << sh;
            const int f = bt.first(refs[i]);
            for (int j = 0; j < c; ++j) perm[tri_base + k_tri + j] = f + j;
            k_tri += c;
        }
        const float l3[3] = {clo[i].x, clo[i].y, clo[i].z}, h3[3] = {chi[i].x, chi[i].y, chi[i].z};
        for (int a = 0; a < 3; ++a) {
            const float margin = sc[a] * 0.0078125f;  // 2^-7 of a step
            int ql = (int)floorf((l3[a] - p[a]) / sc[a]);
            ql = max(0, min(255, ql));
            while (ql > 0 && q_decode(ql, sc[a], p[a]) > l3[a] - margin) --ql;
            int qh = (int)ceilf((h3[a] - p[a]) / sc[a]);
            qh = max(0, min(255, qh));
            while (qh < 255 && q_decode(qh, sc[a], p[a]) < h3[a] + margin) ++qh;
            q[a][w] |= (unsigned)ql << sh;
            q[3 + a][w] |= (unsigned)qh << sh;
        }
    }
    float4 *o = nodes + kNodeRows * (size_t)item.node;
    o[0] = make_float4(p[0], p[1], p[2], sc[0]);
    o[1] = make_float4(sc[1], sc[2], __int_as_float(child_base), __int_as_float(tri_base));
    o[2] = make_float4(__uint_as_float(meta[0]), __uint_as_float(meta[1]), __uint_as_float(imask), 0.0f);
    o[3] = make_float4(__uint_as_float(q[0][0]), __uint_as_float(q[0][1]), __uint_as_float(q[1][0]), __uint_as_float(q[1][1]));
    o[4] = make_float4(__uint_as_float(q[2][0]), __uint_as_float(q[2][1]), __uint_as_float(q[3][0]), __uint_as_float(q[3][1]));
    o[5] = make_float4(__uint_as_float(q[4][0]), __uint_as_float(q[4][1]), __uint_as_float(q[5][0]), __uint_as_float(q[5][1]));
}

// final order of the BVH triangles: wide-node order (perm over the sorted positions) composed with the sort
__global__ void k_compose_order(const int *__restrict__ perm, const int *__restrict__ vals, int nb, int *__restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < nb) out[k] = vals[perm[k]];
}

// scan records of the BVH triangles, in record order: the plane and the three edge planes of scan_singles (pt_trace.cuh)
__global__ void k_pack_scans(const float *__restrict__ verts, const float4 *__restrict__ planes,
                             const float *__restrict__ squares, const int *__restrict__ order, int nb, float eps,
                             const BuildScratch *__restrict__ s, float4 *__restrict__ scans) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nb) return;
    const float amax = fmaxf(20.0f, fmaxf(fmaxf(fmaxf(fabsf(ord2f(s->lo[0])), fabsf(ord2f(s->hi[0]))),
                                                fmaxf(fabsf(ord2f(s->lo[1])), fabsf(ord2f(s->hi[1])))),
                                          fmaxf(fabsf(ord2f(s->lo[2])), fabsf(ord2f(s->hi[2])))));
    const float slack = amax * (64.0f * 1.1920929e-7f);
    const int i = order[k];
    const float *v = verts + 9 * (size_t)i;
    LargeTri t;
    t.p[0] = mk(v[0], v[1], v[2]); t.p[1] = mk(v[3], v[4], v[5]); t.p[2] = mk(v[6], v[7], v[8]);
    t.plane = planes[i];
    t.S = squares[i];
    t.idx = i;
    t.mat = 0;
    const V3 ng = normalize3(cross3(t.p[1] - t.p[0], t.p[2] - t.p[0]));
    bool unb;  // (cannot be set for a triangle inside the BVH: pt_scene_create keeps unbounded ones out or goes brute force)
    const float e = tri_e(t, eps, amax, &unb);
    const bool ok = t.S > 0.0f && t.S <= FLT_MAX && ng.x == ng.x && !unb;
    float4 *o = scans + 4 * (size_t)k;
    o[0] = t.plane;
    for (int j = 0; j < 3; ++j) o[1 + j] = edge_plane(t, j, ng, ok, e, slack);
}

// box of the whole tree and the two child boxes of the binary root (node layout of the root test, pt_scene.cuh)
__global__ void k_root_box(const int *__restrict__ vals, const float4 *__restrict__ box_lo,
                           const float4 *__restrict__ box_hi, int nb, const int2 *__restrict__ children,
                           const float4 *__restrict__ node_lo, const float4 *__restrict__ node_hi,
                           float4 *__restrict__ out5) {
    float4 l[2], h[2];
    if (nb >= 2) {
        const int2 c = children[0];
        for (int k = 0; k < 2; ++k) {
            const int ck = k ? c.y : c.x;
            l[k] = ck >= 0 ? node_lo[ck] : box_lo[vals[~ck]];
            h[k] = ck >= 0 ? node_hi[ck] : box_hi[vals[~ck]];
        }
    } else {
        l[0] = l[1] = box_lo[vals[0]];
        h[0] = h[1] = box_hi[vals[0]];
    }
    out5[0] = make_float4(fminf(l[0].x, l[1].x), fminf(l[0].y, l[1].y), fminf(l[0].z, l[1].z), 0.0f);
    out5[1] = make_float4(fmaxf(h[0].x, h[1].x), fmaxf(h[0].y, h[1].y), fmaxf(h[0].z, h[1].z), 0.0f);
    out5[2] = make_float4(l[0].x, h[0].x, l[0].y, h[0].y);
    out5[3] = make_float4(l[1].x, h[1].x, l[1].y, h[1].y);
    out5[4] = make_float4(l[0].z, h[0].z, l[1].z, h[1].z);
}

// Factory (material.h:58-106) as data: which closures exist and the mirror's selection chance
__global__ void k_pack_materials(const float *__restrict__ mats10, const float *__restrict__ ext2, int m,
                                 float4 *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const float *p = mats10 + 10 * (size_t)i;
    const bool emissive = p[0] != 0.0f || p[1] != 0.0f || p[2] != 0.0f;
    const float Ns = p[9];
    const bool has_mirror = (Ns != 0.0f) && (p[6] != 0.0f || p[7] != 0.0f || p[8] != 0.0f);
    float mirror_chance = Ns / 1000;
    float diffuse_chance = 1 - Ns / 1000;
    const bool has_diffuse = diffuse_chance > 0;
    const float Ni = ext2 ? ext2[2 * (size_t)i] : 0.0f, Tr = ext2 ? ext2[2 * (size_t)i + 1] : 0.0f;
    const bool has_refract = Tr > 0.0f && Ni > 0.0f;  // extension (oracle/pt_oracle.c trace_ray)
    int kind;
    if (emissive) kind = MAT_EMISSIVE;
    else if (has_refract) kind = MAT_MULTI;
    else if (has_mirror && has_diffuse) kind = MAT_BOTH;
    else if (has_mirror) kind = MAT_MIRROR;
    else if (has_diffuse) kind = MAT_DIFFUSE;
    else kind = MAT_NONE;
    if (has_refract) {
        mirror_chance = Ns / 1000 * (1.0f - Tr);
        diffuse_chance = (1 - Ns / 1000) * (1.0f - Tr);
    }
    const int mask = (has_mirror ? 1 : 0) | (has_diffuse ? 2 : 0) | (has_refract ? 4 : 0);
    out[3 * i + 0] = make_float4(p[3], p[4], p[5], __int_as_float(kind));
    out[3 * i + 1] = make_float4(p[6], p[7], p[8], mirror_chance);
    out[3 * i + 2] = make_float4(Ni, Tr, __int_as_float(mask), diffuse_chance);
}

}  // namespace ptd
