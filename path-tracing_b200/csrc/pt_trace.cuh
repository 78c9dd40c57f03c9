// pt_trace.cuh — nearest hit (Scene::TraceRay's loop, scene.cpp:114-120) and the reference's Triangle::Intersect
// (triangles.h:48-73) on the packed triangle records.
//
// Nearest-hit semantics to reproduce: the reference scans triangles in index order and accepts a triangle only when
// its plane distance is strictly smaller than the best so far, so the answer is "minimum t over every triangle that
// passes the inside test, ties to the lowest index".  Any visiting order gives that answer as long as (i) no box
// that can contain an accepted hit is culled — boxes are dilated for this (pt_bvh.cuh) and a box is kept while its
// entry distance is <= the best t, not <; (ii) ties are resolved on the caller's triangle index.
//
// Two stages per ray:
//   large_phase  the few "large" triangles (walls) from shared memory, all lanes in step.  The cheap part (plane
//                distance, range test, hit point inside the triangle's dilated box) runs for every triangle; the
//                expensive inside test (3 cross products, 3 IEEE square roots) is DEFERRED: candidates are marked in
//                a per-lane bit mask and resolved together after the scan, so that lanes whose candidates sit at
//                different triangle indices still run the inside test in the same instructions.
//   traverse     stack-based BVH2 traversal over the LBVH for rays whose t-range still reaches the root box.
#pragma once
#include "pt_math.cuh"
#include "pt_scene.cuh"

namespace ptd {

// Hit::idx value "not fetched yet": k_wf_traverse takes a ray over with the record of its best hit so far and reads that
// record's caller index only if a candidate ties with it (one dependent fetch less per ray); caller indices are >= 0
constexpr int kIdxLazy = (int)0x80000000;

struct Hit {
    float t;   // best distance so far (+inf = none)
    int idx;   // caller's triangle index of the best hit (-1 = none, kIdxLazy = look it up in record `rec`)
    int rec;   // record index of the best hit in SceneView::tris
};

__device__ __forceinline__ Hit no_hit() {
    Hit h;
    h.t = __int_as_float(0x7f800000);
    h.idx = -1;
    h.rec = -1;
    return h;
}

// PlaneIntersect, triangles.h:10-13
__device__ __forceinline__ float plane_distance(float4 p, V3 o, V3 d) {
    const float sd = d.x * p.x + d.y * p.y + d.z * p.z;
    return -(o.x * p.x + o.y * p.y + o.z * p.z + p.w) / sd;
}

// the area-sum inside test of triangles.h:56-70 for the plane hit at distance nd; a = (v0, S), b = (v1, .), c = (v2, .)
// Compile-time variant -DPT_INSIDE_APPROX=1 (0 = off, the default build; NOT yet measured — prepared for the next GPU
// session, tools/ab_quick.py): the three areas first with MUFU.RSQ instead of three IEEE square roots.  The staged
// test below accepts exactly when |S - (s1 + s2 + s3)| <= eps (the early exits follow from it, the areas being
// non-negative), so a sum that misses the threshold by more than a margin covering every rounding on either side —
// rsqrt.approx is within 2^-22, the sums within a few 2^-24 of S + s1 + s2 + s3 — decides; anything closer to the
// threshold, and any NaN, goes through the exact test.
#ifndef PT_INSIDE_APPROX
#define PT_INSIDE_APPROX 0
#endif
#ifndef PT_TRI_PREREJECT
#define PT_TRI_PREREJECT 0
#endif
#ifndef PT_COOP_EXACT
#define PT_COOP_EXACT 0
#endif
// MUFU.RCP: at most 1 ulp from 1/x (PTX ISA, rcp.approx.ftz.f32); denormal inputs flush to zero -> +-inf
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ float rsqrt_approx(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ bool inside_test(float4 a, float4 b, float4 c, V3 o, V3 d, float nd, float eps) {
    const float S = a.w;
    const V3 P = mk(o.x + d.x * nd, o.y + d.y * nd, o.z + d.z * nd);
    const V3 f0 = P - mk(a.x, a.y, a.z), f1 = P - mk(b.x, b.y, b.z), f2 = P - mk(c.x, c.y, c.z);
#if PT_INSIDE_APPROX
    {
        const V3 c1 = cross3(f0, f1), c2 = cross3(f0, f2), c3 = cross3(f2, f1);
        const float q1 = dot3(c1, c1), q2 = dot3(c2, c2), q3 = dot3(c3, c3);
        const float sum = q1 * rsqrt_approx(q1) + q2 * rsqrt_approx(q2) + q3 * rsqrt_approx(q3);  // 0 * inf = NaN: exact path
        const float dev = fabsf(S - sum), margin = 8e-6f * (S + sum);
        if (dev > eps + margin) return false;
        if (dev < eps - margin) return true;
    }
#endif
    const float s1 = length3(cross3(f0, f1));
    if (s1 > S + eps) return false;
    const float s2 = length3(cross3(f0, f2));
    if (s1 + s2 > S + eps) return false;
    const float s3 = length3(cross3(f2, f1));
    return !(fabsf(S - s1 - s2 - s3) > eps);
}

// 256-bit read-only load (LDG.E.256, sm_100): two float4 rows with one request.  A divergent warp pays one L1 tag
// lookup per lane and instruction, and the traversal kernel is bound by exactly that (ncu: l1tex throughput 77 %,
// issue 56 %), so a 64-byte node or triangle record is fetched with two of these instead of four LDG.128.
// p must be 32-byte aligned (nodes and records are 64-byte aligned).
__device__ __forceinline__ void ldg256(const float4 *p, float4 &a, float4 &b) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
}

// Triangle::Intersect (triangles.h:48-73) on record `rec` (4 float4 at r).  Differences from the reference, both
// outside its defined behaviour: a NaN distance (0/0: ray in the plane) is rejected instead of accepted, and equal
// distances go to the lower caller index explicitly.
__device__ __forceinline__ void tri_intersect_ref(const float4 *r, int rec, V3 o, V3 d, float eps, Hit &h) {
    float4 pl, a, b, c;
    ldg256(r, pl, a);
    ldg256(r + 2, b, c);
#if PT_TRI_PREREJECT
    // Compile-time variant (0 = off, the default build; NOT yet measured — prepared for the next GPU session): most
    // candidates of a leaf fail the inside test, and they pay an IEEE division and up to three IEEE square roots to be
    // told so (ncu: a third of the kernel's warp instructions, 6.7 of 32 lanes active).  Here the hit point comes from
    // MUFU.RCP and the three areas from MUFU.RSQ; the candidate leaves when that area sum misses S by more than eps
    // plus a bound on everything that separates it from the exact evaluation:
    //   8e-6 (S + sum)     rsqrt (2^-22), the roundings of the sums on both sides
    //   per * dP           the hit point moves by dP <= 1e-6 (|t| + |o|_1) (quotient within 2^-21, roundings of o + d t
    //                      and of P - v), every area by at most dP times its edge; per <= 2 l1, l1 = sum of |P - v_k|_1
    //   1e-6 l1^2          the products of the cross products round independently on both sides
    // A NaN fails the comparison and goes on to the exact test.
    {
        const float sd = d.x * pl.x + d.y * pl.y + d.z * pl.z;
        const float num = -(o.x * pl.x + o.y * pl.y + o.z * pl.z + pl.w);
        const float tq = num * rcp_approx(sd);
        const V3 Pq = mk(o.x + d.x * tq, o.y + d.y * tq, o.z + d.z * tq);
        const V3 g0 = Pq - mk(a.x, a.y, a.z), g1 = Pq - mk(b.x, b.y, b.z), g2 = Pq - mk(c.x, c.y, c.z);
        const V3 c1 = cross3(g0, g1), c2 = cross3(g0, g2), c3 = cross3(g2, g1);
        const float q1 = dot3(c1, c1), q2 = dot3(c2, c2), q3 = dot3(c3, c3);
        const float sum = q1 * rsqrt_approx(q1) + q2 * rsqrt_approx(q2) + q3 * rsqrt_approx(q3);
        const float l1 = fabsf(g0.x) + fabsf(g0.y) + fabsf(g0.z) + fabsf(g1.x) + fabsf(g1.y) + fabsf(g1.z) + fabsf(g2.x) +
                         fabsf(g2.y) + fabsf(g2.z);
        const float dP = 1e-6f * (fabsf(tq) + fabsf(o.x) + fabsf(o.y) + fabsf(o.z));
        const float margin = 8e-6f * (a.w + sum) + 2.0f * l1 * dP + 1e-6f * l1 * l1;
        if (fabsf(a.w - sum) > eps + margin) return;
    }
#endif
    const float nd = plane_distance(pl, o, d);
    if (!(nd <= h.t) || nd < eps) return;
    const int idx = __float_as_int(b.w);
    if (nd == h.t) {  // a tie goes to the lower caller index; also rejects nd == +inf against the initial (inf, -1)
        if (h.idx == kIdxLazy) h.idx = __float_as_int(__ldg(r - 4 * (ptrdiff_t)rec + 4 * (ptrdiff_t)h.rec + 2).w);
        if (!(idx < h.idx)) return;
    }
    if (!inside_test(a, b, c, o, d, nd, eps)) return;
    h.t = nd;
    h.idx = idx;
    h.rec = rec;
}

// ---- large triangles from shared memory (image layout: pt_scene.cuh, built by k_large_pack) ------------------
__device__ __forceinline__ void stage_large(const SceneView &sc, float4 *s_large) {
    const int rows = large_rows(sc.n_large, sc.n_quads);
    for (int k = threadIdx.x; k < rows; k += blockDim.x) s_large[k] = sc.lrecs[k];
}

// Scan + resolve.
// The SCAN is one short, branch-free instruction sequence per unit, the same for every lane: the plane distance with
// an approximate reciprocal instead of the IEEE division, a range test and the edge planes of the unit, all with
// tolerances that make it CONSERVATIVE — it may mark a triangle the exact test rejects, never the reverse: numerator
// and denominator are the reference's own (bit-identical), so t' = num * rcp(sd) is within 2^-21 relative of the
// reference's quotient; the hit point P' = o + d t' is then within |t'| 2^-20 + a few ulp of the reference's, which
// is added to the edge planes' margin (k_large_pack).
// A QUAD unit covers two near-coplanar triangles A, B that share an edge with one plane evaluation: the plane is
// A's; B's own distance differs from it by at most tau = (c_abs + |t| c_rel) / |d . n| (k_large_pack), which widens
// the range test and every edge margin; the four outer edges bound the union, the shared edge's plane splits the
// marks between A and B (both near the diagonal).
// The RESOLVE runs the reference's exact test (IEEE division, area sums, tie rule) on each lane's marked triangles;
// lanes whose candidates sit at different indices execute it together.
// (tnear, rnear): smallest scanned distance among the marked units so far and the first record of that unit
__device__ __forceinline__ unsigned scan_quads(const float4 *units, int u0, int n, V3 o, V3 d, float t_lo, float t_hi,
                                               float &tnear, int &rnear) {
    unsigned mask = 0u;  // bit 2 (u - u0) = A of unit u, bit 2 (u - u0) + 1 = B
    for (int u = 0; u < n; ++u) {
        const float4 *r = units + kQuadRows * (u0 + u);
        const float4 p = r[0], meta = r[6];
        // contracted (FMA) dot products: they differ from the reference's by a few roundings of terms <= 2 amax, which
        // tau already covers for both triangles of the unit (k_large_pack: c_abs >= 1e-6 amax, c_rel = 1e-6)
        const float sd = __fmaf_rn(d.x, p.x, __fmaf_rn(d.y, p.y, d.z * p.z));
        const float num = -__fmaf_rn(o.x, p.x, __fmaf_rn(o.y, p.y, __fmaf_rn(o.z, p.z, p.w)));
        const float rc = rcp_approx(sd);
        const float t = num * rc;
        const float tau = __fmaf_rn(fabsf(t), meta.y, meta.x) * fabsf(rc);
        const float tol = __fmaf_rn(fabsf(t), 2e-6f, tau);
        const float Px = __fmaf_rn(d.x, t, o.x), Py = __fmaf_rn(d.y, t, o.y), Pz = __fmaf_rn(d.z, t, o.z);
        const float4 e0 = r[1], e1 = r[2], e2 = r[3], e3 = r[4], e4 = r[5];
        const float g0 = __fmaf_rn(e0.x, Px, __fmaf_rn(e0.y, Py, __fmaf_rn(e0.z, Pz, e0.w)));
        const float g1 = __fmaf_rn(e1.x, Px, __fmaf_rn(e1.y, Py, __fmaf_rn(e1.z, Pz, e1.w)));
        const float g2 = __fmaf_rn(e2.x, Px, __fmaf_rn(e2.y, Py, __fmaf_rn(e2.z, Pz, e2.w)));
        const float g3 = __fmaf_rn(e3.x, Px, __fmaf_rn(e3.y, Py, __fmaf_rn(e3.z, Pz, e3.w)));
        const float gd = __fmaf_rn(e4.x, Px, __fmaf_rn(e4.y, Py, __fmaf_rn(e4.z, Pz, e4.w)));
        // written so that a NaN anywhere (sd == 0, infinities) marks both triangles: the resolve decides
        const bool out = (t - tau > t_hi) | (t + tau < t_lo) | (fminf(fminf(g0, g1), fminf(g2, g3)) < -tol);
        const unsigned a = (out | (gd < -tol)) ? 0u : 1u, b = (out | (gd > meta.z + tol)) ? 0u : 2u;
        mask |= (a | b) << (2 * u);
        const bool nearer = (a | b) != 0u && t < tnear;
        tnear = nearer ? t : tnear;
        rnear = nearer ? 2 * (u0 + u) : rnear;
    }
    return mask;
}

// one 32-triangle chunk of single units [s0, s0 + n): returns their candidate mask
__device__ __forceinline__ unsigned scan_singles(const float4 *units, int s0, int n, int rec0, V3 o, V3 d, float t_lo,
                                                 float t_hi, float &tnear, int &rnear) {
    unsigned mask = 0u;
    for (int k = 0; k < n; ++k) {
        const float4 *r = units + kSingleRows * (s0 + k);
        const float4 p = r[0];
        const float sd = d.x * p.x + d.y * p.y + d.z * p.z;
        const float num = -(o.x * p.x + o.y * p.y + o.z * p.z + p.w);
        const float t = num * rcp_approx(sd);
        const float tol = fabsf(t) * 2e-6f;
        const float Px = __fmaf_rn(d.x, t, o.x), Py = __fmaf_rn(d.y, t, o.y), Pz = __fmaf_rn(d.z, t, o.z);
        const float4 e0 = r[1], e1 = r[2], e2 = r[3];
        const float g0 = __fmaf_rn(e0.x, Px, __fmaf_rn(e0.y, Py, __fmaf_rn(e0.z, Pz, e0.w)));
        const float g1 = __fmaf_rn(e1.x, Px, __fmaf_rn(e1.y, Py, __fmaf_rn(e1.z, Pz, e1.w)));
        const float g2 = __fmaf_rn(e2.x, Px, __fmaf_rn(e2.y, Py, __fmaf_rn(e2.z, Pz, e2.w)));
        // written so that a NaN anywhere (sd == 0, infinities) marks the triangle: the resolve decides
        const bool out = (t > t_hi) | (t < t_lo) | (fminf(fminf(g0, g1), g2) < -tol);
        mask |= (out ? 0u : 1u) << k;
        const bool nearer = !out && t < tnear;
        tnear = nearer ? t : tnear;
        rnear = nearer ? rec0 + s0 + k : rnear;
    }
    return mask;
}

// exact test on the marked records k0 + bit of the shared-memory record rows
__device__ __forceinline__ void large_resolve32(const float4 *s_large, int k0, unsigned mask, int rec_base, V3 o, V3 d,
                                                float eps, Hit &h) {
    while (mask) {
        const int k = k0 + __ffs(mask) - 1;
        mask &= mask - 1u;
        const float4 *r = s_large + 4 * k;
        const float nd = plane_distance(r[0], o, d);
        if (!(nd <= h.t) || nd < eps) continue;
        const float4 a = r[1], b = r[2], c = r[3];
        const int idx = __float_as_int(b.w);
        if (nd == h.t && !(idx < h.idx)) continue;
        if (!inside_test(a, b, c, o, d, nd, eps)) continue;
        h.t = nd;
        h.idx = idx;
        h.rec = rec_base + k;
    }
}

__device__ __forceinline__ void large_phase(const float4 *s_large, int n_large, int n_quads, int rec_base, V3 o, V3 d,
                                            float eps, Hit &h) {
    const float t_hi = h.t + fabsf(h.t) * 2e-6f, t_lo = eps - fabsf(eps) * 2e-6f;
    const float4 *quads = s_large + 4 * n_large, *singles = quads + kQuadRows * n_quads;
    const int n_single = n_large - 2 * n_quads;
    // records 2u, 2u + 1 belong to quad u; record 2 n_quads + s to single s
    float tnear = __int_as_float(0x7f800000);
    int rnear = -1;
    unsigned q0 = scan_quads(quads, 0, min(n_quads, 16), o, d, t_lo, t_hi, tnear, rnear);
    unsigned q1 = n_quads > 16 ? scan_quads(quads, 16, n_quads - 16, o, d, t_lo, t_hi, tnear, rnear) : 0u;
    unsigned s0 = scan_singles(singles, 0, min(n_single, 32), 2 * n_quads, o, d, t_lo, t_hi, tnear, rnear);
    unsigned s1 = n_single > 32 ? scan_singles(singles, 32, n_single - 32, 2 * n_quads, o, d, t_lo, t_hi, tnear, rnear) : 0u;
    // The unit with the smallest scanned distance first, all lanes together: it is the hit for most rays, and once
    // h.t is set the other candidates leave the exact test at its range check.  Any order gives the same answer
    // (header comment).
    if (rnear >= 0) {
        unsigned bits;
        if (rnear < 2 * n_quads) {
            unsigned &q = rnear < 32 ? q0 : q1;
            const int sh = rnear & 31;
            bits = (q >> sh) & 3u;
            q &= ~(3u << sh);
        } else {
            const int k = rnear - 2 * n_quads;
            unsigned &m = k < 32 ? s0 : s1;
            bits = 1u;
            m &= ~(1u << (k & 31));
        }
        large_resolve32(s_large, rnear, bits, rec_base, o, d, eps, h);
    }
    large_resolve32(s_large, 0, q0, rec_base, o, d, eps, h);
    if (q1) large_resolve32(s_large, 32, q1, rec_base, o, d, eps, h);
    large_resolve32(s_large, 2 * n_quads, s0, rec_base, o, d, eps, h);
    if (s1) large_resolve32(s_large, 2 * n_quads + 32, s1, rec_base, o, d, eps, h);
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

// Does the ray [0, tmax] have to walk the tree?  Tested against the root's two child boxes (tighter than the root
// box: on the 1M-triangle scene half of the rays that enter the root box leave it without entering a child).
__device__ __forceinline__ bool reaches_root(const SceneView &sc, V3 o, V3 d, float tmax) {
    if (sc.root == kRefEmpty) return false;
    const float inf = __int_as_float(0x7f800000);
    const V3 inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    if (sc.root < 0)  // the whole tree is one leaf
        return box_entry(sc.root_lo.x, sc.root_hi.x, sc.root_lo.y, sc.root_hi.y, sc.root_lo.z, sc.root_hi.z, o, inv, tmax) < inf;
    const float4 n0 = sc.root_n0, n1 = sc.root_n1, n2 = sc.root_n2;
    return box_entry(n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, o, inv, tmax) < inf ||
           box_entry(n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, o, inv, tmax) < inf;
}

// The triangles of one leaf: plane distances of all of them first (cheap, same instructions for every lane), then
// the reference's exact test on those whose distance can still win (up to kLeafMax records, contiguous).
template <bool COUNT>
__device__ __forceinline__ void leaf_step(const SceneView &sc, int ref, V3 o, V3 d, float eps, Hit &h,
                                          unsigned long long &n_tris) {
    const int first = leaf_first(ref), cnt = leaf_count(ref);
    unsigned mask = 0u;
    // all plane rows first (independent loads in flight together), then the arithmetic
    float4 pl[kLeafMax];
#pragma unroll
    for (int k = 0; k < kLeafMax; ++k)
        pl[k] = k < cnt ? __ldg(sc.tris + 4 * (size_t)(first + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
    // The plane pass with an approximate reciprocal instead of up to kLeafMax IEEE divisions: numerator and denominator
    // are the reference's own, MUFU.RCP is within 2^-21 of the IEEE quotient and the range is widened by 2e-6 relative
    // (as in scan_singles), so every triangle the exact test can accept is marked; the exact test repeats the division
    // for the marked ones.  A NaN marks.  (-3.7 % traversal time on the 1M-triangle scene: the leaf phase, with its
    // divisions, square roots and few active lanes, is where the walk is short of issue slots.)
    const float t_hi = h.t + fabsf(h.t) * 2e-6f, t_lo = eps - fabsf(eps) * 2e-6f;
#pragma unroll
    for (int k = 0; k < kLeafMax; ++k) {
        const float4 p = pl[k];
        const float sd = d.x * p.x + d.y * p.y + d.z * p.z;
        const float num = -(o.x * p.x + o.y * p.y + o.z * p.z + p.w);
        const float t = num * rcp_approx(sd);
        mask |= (unsigned)(k < cnt && !((t > t_hi) | (t < t_lo))) << k;
    }
    while (mask) {
        const int k = __ffs(mask) - 1;
        mask &= mask - 1u;
        tri_intersect_ref(sc.tris + 4 * (size_t)(first + k), first + k, o, d, eps, h);
    }
    if (COUNT) n_tris += cnt;
}

#if PT_COOP_EXACT
// Compile-time variant (0 = off, the default build; NOT yet measured — prepared for the next GPU session): the exact
// tests of a warp's leaf phase shared out over all 32 lanes.  As written above every lane runs its own candidates one
// after the other — ncu: a third of the kernel's warp instructions at 6.7 active lanes.  Here the (ray, candidate)
// pairs of the warp are numbered by a prefix sum over the lanes' candidate counts, lane i tests pair base + i with
// the owner's ray fetched by shuffles, and every owner merges the results of its pairs (minimum t, ties to the lower
// caller index: the nearest-hit rule does not depend on the order, pt_trace.cuh header).  All candidates of a leaf see
// the t the ray came in with, so a hit no longer prunes its neighbours' inside tests — they run side by side instead.

// the plane pass of leaf_step: candidate mask of leaf `ref` for this lane's ray
template <bool COUNT>
__device__ __forceinline__ unsigned leaf_candidates(const SceneView &sc, int ref, V3 o, V3 d, float eps, const Hit &h,
                                                    unsigned long long &n_tris) {
    const int first = leaf_first(ref), cnt = leaf_count(ref);
    unsigned mask = 0u;
    float4 pl[kLeafMax];
#pragma unroll
    for (int k = 0; k < kLeafMax; ++k)
        pl[k] = k < cnt ? __ldg(sc.tris + 4 * (size_t)(first + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float t_hi = h.t + fabsf(h.t) * 2e-6f, t_lo = eps - fabsf(eps) * 2e-6f;
#pragma unroll
    for (int k = 0; k < kLeafMax; ++k) {
        const float4 p = pl[k];
        const float sd = d.x * p.x + d.y * p.y + d.z * p.z;
        const float num = -(o.x * p.x + o.y * p.y + o.z * p.z + p.w);
        const float t = num * rcp_approx(sd);
        mask |= (unsigned)(k < cnt && !((t > t_hi) | (t < t_lo))) << k;
    }
    if (COUNT) n_tris += cnt;
    return mask;
}

// an accepted candidate (t, caller index, record) against the best hit so far
__device__ __forceinline__ void hit_merge(const float4 *tris, Hit &h, float t, int idx, int rec) {
    if (!(t <= h.t)) return;
    if (t == h.t) {
        if (h.idx == kIdxLazy) h.idx = __float_as_int(__ldg(tris + 4 * (size_t)h.rec + 2).w);
        if (!(idx < h.idx)) return;
    }
    h.t = t;
    h.idx = idx;
    h.rec = rec;
}

// Must be reached by all 32 lanes.  mask / first: this lane's candidates (bit k = record first + k), 0 = none.
__device__ __forceinline__ void coop_exact(const SceneView &sc, unsigned mask, int first, V3 o, V3 d, float eps, Hit &h) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int mycnt = __popc(mask);
    int incl = mycnt;  // inclusive prefix sum of the candidate counts
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(full, incl, off);
        if (lane >= off) incl += v;
    }
    const int start = incl - mycnt, total = __shfl_sync(full, incl, 31);
    for (int base = 0; base < total; base += 32) {  // the same trip count for every lane
        const int pair = base + lane;
        // owner of the pair: the first lane whose inclusive sum exceeds it
        int owner = 0;
#pragma unroll
        for (int step = 16; step; step >>= 1) {
            const int v = __shfl_sync(full, incl, owner + step - 1);
            if (v <= pair) owner += step;
        }
        const bool valid = pair < total;
        owner = min(owner, 31);
        int j = pair - __shfl_sync(full, start, owner);  // which of the owner's candidates
        unsigned om = __shfl_sync(full, mask, owner);
        const int ofirst = __shfl_sync(full, first, owner);
        const V3 oo = mk(__shfl_sync(full, o.x, owner), __shfl_sync(full, o.y, owner), __shfl_sync(full, o.z, owner));
        const V3 dd = mk(__shfl_sync(full, d.x, owner), __shfl_sync(full, d.y, owner), __shfl_sync(full, d.z, owner));
        Hit hh;
        hh.t = __shfl_sync(full, h.t, owner);
        hh.idx = __shfl_sync(full, h.idx, owner);
        hh.rec = __shfl_sync(full, h.rec, owner);
        int got = -1;  // record of the accepted candidate
        if (valid) {
            for (; j > 0; --j) om &= om - 1u;
            const int rec = ofirst + __ffs(om) - 1;
            tri_intersect_ref(sc.tris + 4 * (size_t)rec, rec, oo, dd, eps, hh);
            if (hh.rec == rec) got = rec;
        }
        // every owner collects the results of its pairs that were tested in this round
#pragma unroll
        for (int q = 0; q < kLeafMax; ++q) {
            const int src = start - base + q;
            const bool mine = q < mycnt && src >= 0 && src < 32;
            const int s = mine ? src : lane;
            const float rt = __shfl_sync(full, hh.t, s);
            const int ri = __shfl_sync(full, hh.idx, s), rr = __shfl_sync(full, got, s);
            if (mine && rr >= 0) hit_merge(sc.tris, h, rt, ri, rr);
        }
    }
}
#endif

// One step on the inner node `cur` (>= 0); returns the next ref: a node, a leaf (< 0) or kRefEmpty when the stack ran out.
template <bool COUNT>
__device__ __forceinline__ int node_step(const SceneView &sc, int cur, V3 o, V3 inv, const Hit &h, int *stack, int &sp,
                                         unsigned long long &n_nodes) {
    {
        const float4 *nd = sc.nodes + 4 * (size_t)cur;
        float4 n0, n1, n2, n3;
        ldg256(nd, n0, n1);
        ldg256(nd + 2, n2, n3);
        if (COUNT) ++n_nodes;
        // box_entry clips against h.t with <=, so that an equal-distance lower index is still found; +inf = miss
        const float t0 = box_entry(n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, o, inv, h.t);
        const float t1 = box_entry(n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, o, inv, h.t);
        const int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
        const bool g0 = t0 < __int_as_float(0x7f800000), g1 = t1 < __int_as_float(0x7f800000);
        if (g0 && g1) {
            const bool swap = t1 < t0;
            if (sp < kStackDepth) stack[sp++] = swap ? c0 : c1;
            return swap ? c1 : c0;
        }
        if (g0) return c0;
        if (g1) return c1;
        return sp ? stack[--sp] : kRefEmpty;
    }
}

// One traversal step on `cur` (a node or a leaf ref); returns the next ref (kRefEmpty when the stack is empty).
template <bool COUNT>
__device__ __forceinline__ int traverse_step(const SceneView &sc, int cur, V3 o, V3 d, V3 inv, float eps, Hit &h,
                                             int *stack, int &sp, unsigned long long &n_nodes,
                                             unsigned long long &n_tris) {
    if (cur >= 0) return node_step<COUNT>(sc, cur, o, inv, h, stack, sp, n_nodes);
    leaf_step<COUNT>(sc, cur, o, d, eps, h, n_tris);
    return sp ? stack[--sp] : kRefEmpty;
}

// scene.cpp:114-120 as written: every triangle, in order (the checker mode, PT_ACCEL_BRUTE)
template <bool COUNT>
__device__ __forceinline__ Hit brute_hit(const SceneView &sc, V3 o, V3 d, float eps, unsigned long long &n_tris) {
    Hit h = no_hit();
    for (int k = 0; k < sc.n_tri; ++k) tri_intersect_ref(sc.tris + 4 * (size_t)k, k, o, d, eps, h);
    if (COUNT) n_tris += sc.n_tri;
    return h;
}

// Whole nearest-hit query in one thread (pt_intersect, the megakernel).
template <bool COUNT>
__device__ __forceinline__ Hit nearest_hit(const SceneView &sc, const float4 *s_large, V3 o, V3 d, float eps,
                                           unsigned long long &n_nodes, unsigned long long &n_tris) {
    if (sc.brute) return brute_hit<COUNT>(sc, o, d, eps, n_tris);
    Hit h = no_hit();
    large_phase(s_large, sc.n_large, sc.n_quads, sc.n_bvh, o, d, eps, h);
    if (COUNT) n_tris += sc.n_large;
    if (!reaches_root(sc, o, d, h.t)) return h;
    const V3 inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[kStackDepth];
    int sp = 0;
    int cur = sc.root;
    while (cur != kRefEmpty) cur = traverse_step<COUNT>(sc, cur, o, d, inv, eps, h, stack, sp, n_nodes, n_tris);
    return h;
}

}  // namespace ptd
