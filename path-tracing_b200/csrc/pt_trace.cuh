// pt_trace.cuh — nearest hit (Scene::TraceRay's loop, scene.cpp:114-120) through the LBVH, and the reference's
// Triangle::Intersect (triangles.h:48-73) on the packed triangle records.
//
// Nearest-hit semantics to reproduce: the reference scans triangles in index order and accepts a triangle only when
// its plane distance is strictly smaller than the best so far, so the answer is "minimum t over every triangle that
// passes the inside test, ties to the lowest index".  Any traversal order gives that answer as long as (i) no box
// that can contain an accepted hit is culled — boxes are dilated for this (pt_bvh.cuh) and a box is kept while its
// entry distance is <= the best t, not <; (ii) ties are resolved on the caller's triangle index.
#pragma once
#include "pt_math.cuh"
#include "pt_scene.cuh"

namespace ptd {

struct Hit {
    float t;   // best distance so far (+inf = none)
    int idx;   // caller's triangle index of the best hit (-1 = none)
    int rec;   // record index of the best hit in SceneView::tris
};

// Triangle::Intersect (triangles.h:48-73; PlaneIntersect :10-13, ParallelogramSquare :15-17) on record `rec`
// (4 float4 at r).  Differences from the reference, both outside its defined behaviour: a NaN distance (0/0: ray
// in the plane) is rejected instead of accepted, and equal distances go to the lower caller index explicitly.
__device__ __forceinline__ void tri_intersect_ref(const float4 *__restrict__ r, int rec, V3 o, V3 d, float eps, Hit &h) {
    const float4 p = r[0];
    const float sd = d.x * p.x + d.y * p.y + d.z * p.z;
    const float nd = -(o.x * p.x + o.y * p.y + o.z * p.z + p.w) / sd;
    if (!(nd <= h.t) || nd < eps) return;
    const float4 a = r[1], b = r[2], c = r[3];
    const int idx = __float_as_int(b.w);
    if (nd == h.t && !(idx < h.idx)) return;  // also rejects nd == +inf against the initial (inf, -1)
    const float S = a.w;
    const V3 P = mk(o.x + d.x * nd, o.y + d.y * nd, o.z + d.z * nd);
    const V3 f0 = P - mk(a.x, a.y, a.z), f1 = P - mk(b.x, b.y, b.z), f2 = P - mk(c.x, c.y, c.z);
    const float s1 = length3(cross3(f0, f1));
    if (s1 > S + eps) return;
    const float s2 = length3(cross3(f0, f2));
    if (s1 + s2 > S + eps) return;
    const float s3 = length3(cross3(f2, f1));
    if (fabsf(S - s1 - s2 - s3) > eps) return;
    h.t = nd;
    h.idx = idx;
    h.rec = rec;
}

// Same test on a record staged in shared memory (the large triangles)
__device__ __forceinline__ void tri_intersect_ref_smem(const float4 *r, int rec, V3 o, V3 d, float eps, Hit &h) {
    const float4 p = r[0];
    const float sd = d.x * p.x + d.y * p.y + d.z * p.z;
    const float nd = -(o.x * p.x + o.y * p.y + o.z * p.z + p.w) / sd;
    if (!(nd <= h.t) || nd < eps) return;
    const float4 a = r[1], b = r[2], c = r[3];
    const int idx = __float_as_int(b.w);
    if (nd == h.t && !(idx < h.idx)) return;
    const float S = a.w;
    const V3 P = mk(o.x + d.x * nd, o.y + d.y * nd, o.z + d.z * nd);
    const V3 f0 = P - mk(a.x, a.y, a.z), f1 = P - mk(b.x, b.y, b.z), f2 = P - mk(c.x, c.y, c.z);
    const float s1 = length3(cross3(f0, f1));
    if (s1 > S + eps) return;
    const float s2 = length3(cross3(f0, f2));
    if (s1 + s2 > S + eps) return;
    const float s3 = length3(cross3(f2, f1));
    if (fabsf(S - s1 - s2 - s3) > eps) return;
    h.t = nd;
    h.idx = idx;
    h.rec = rec;
}

// Slab test of one child box: entry distance, or +inf when the ray [0, tmax] misses the box.  fminf/fmaxf drop a
// NaN operand (0 * inf when the origin lies on a slab plane of an axis-parallel ray), which keeps the test
// conservative together with the boxes' padding.
__device__ __forceinline__ float box_entry(float lox, float hix, float loy, float hiy, float loz, float hiz, V3 o,
                                           V3 inv, float tmax) {
    const float x0 = (lox - o.x) * inv.x, x1 = (hix - o.x) * inv.x;
    const float y0 = (loy - o.y) * inv.y, y1 = (hiy - o.y) * inv.y;
    const float z0 = (loz - o.z) * inv.z, z1 = (hiz - o.z) * inv.z;
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), tmax));
    // 2-ulp widening of the exit distance (Ize 2013) on top of the padded boxes
    return tn <= tf * 1.0000003f ? tn : __int_as_float(0x7f800000);
}

template <bool COUNT>
__device__ __forceinline__ Hit nearest_hit(const SceneView &sc, const float4 *s_large, V3 o, V3 d, float eps,
                                           unsigned long long &n_nodes, unsigned long long &n_tris) {
    Hit h;
    h.t = __int_as_float(0x7f800000);
    h.idx = -1;
    h.rec = -1;
    if (sc.brute) {  // scene.cpp:114-120 as written: every triangle, in order
        if (s_large) {
            for (int k = 0; k < sc.n_tri; ++k) tri_intersect_ref_smem(s_large + 4 * k, k, o, d, eps, h);
        } else {
            for (int k = 0; k < sc.n_tri; ++k) tri_intersect_ref(sc.tris + 4 * (size_t)k, k, o, d, eps, h);
        }
        if (COUNT) n_tris += sc.n_tri;
        return h;
    }
    // large triangles first: they bound t before the tree is walked
    for (int k = 0; k < sc.n_large; ++k) tri_intersect_ref_smem(s_large + 4 * k, sc.n_bvh + k, o, d, eps, h);
    if (COUNT) n_tris += sc.n_large;

    const V3 inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[kStackDepth];
    int sp = 0;
    int cur = sc.root;
    while (cur != kRefEmpty) {
        if (cur >= 0) {
            const float4 *nd = sc.nodes + 4 * (size_t)cur;
            const float4 n0 = __ldg(nd), n1 = __ldg(nd + 1), n2 = __ldg(nd + 2), n3 = __ldg(nd + 3);
            if (COUNT) ++n_nodes;
            const float t0 = box_entry(n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, o, inv, h.t);
            const float t1 = box_entry(n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, o, inv, h.t);
            const int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            // box_entry already clipped against h.t (<=, so that an equal-distance lower index is still found);
            // +inf means a miss
            const bool g0 = t0 < __int_as_float(0x7f800000), g1 = t1 < __int_as_float(0x7f800000);
            if (g0 && g1) {
                const bool swap = t1 < t0;
                cur = swap ? c1 : c0;
                if (sp < kStackDepth) stack[sp++] = swap ? c0 : c1;
            } else if (g0) {
                cur = c0;
            } else if (g1) {
                cur = c1;
            } else {
                cur = sp ? stack[--sp] : kRefEmpty;
            }
        } else {
            const int first = leaf_first(cur), cnt = leaf_count(cur);
            for (int k = 0; k < cnt; ++k) tri_intersect_ref(sc.tris + 4 * (size_t)(first + k), first + k, o, d, eps, h);
            if (COUNT) n_tris += cnt;
            cur = sp ? stack[--sp] : kRefEmpty;
        }
    }
    return h;
}

}  // namespace ptd
