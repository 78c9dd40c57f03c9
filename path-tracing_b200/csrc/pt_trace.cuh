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
//   traverse     stack-based walk of the 8-wide BVH (node8_visit, bvh8_traverse) for rays whose t-range still reaches
//                the root's boxes; a leaf triangle goes through its conservative scan record first (tri_test) and
//                through the reference's exact test only if that marks it.
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

// MUFU.RCP: at most 1 ulp from 1/x (PTX ISA, rcp.approx.ftz.f32); denormal inputs flush to zero -> +-inf
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// the area-sum inside test of triangles.h:56-70 for the plane hit at distance nd; a = (v0, S), b = (v1, .), c = (v2, .)
__device__ __forceinline__ bool inside_test(float4 a, float4 b, float4 c, V3 o, V3 d, float nd, float eps) {
    const float S = a.w;
    const V3 P = mk(o.x + d.x * nd, o.y + d.y * nd, o.z + d.z * nd);
    const V3 f0 = P - mk(a.x, a.y, a.z), f1 = P - mk(b.x, b.y, b.z), f2 = P - mk(c.x, c.y, c.z);
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

// ---- the throughput intersector (PtConfig.intersector = PT_ISECT_WATERTIGHT; north_star's "watertight
// Moller-Trumbore") -------------------------------------------------------------------------------------------------
// Woop, Benthin, Wald 2013, "Watertight ray/triangle intersection": the ray is sheared so that it runs along +z of its
// dominant axis, and the triangle is tested with three 2D edge functions whose shared-edge values are exactly
// antisymmetric between neighbours (plain products and differences — this TU never contracts to FMA), so a ray cannot
// slip between two triangles that share an edge.  Edge functions that come out exactly 0 are re-evaluated in double,
// as in the paper.  No IEEE division or square root: one MUFU.RCP for the shear, one for T / det.
// This is NOT the reference's test (triangles.h:48-73): it hits the geometric triangle — no eps dilation, no `vn`
// plane — so hit IDs can differ from the reference's on edge-grazing rays and differ wholesale where the stored plane
// is not the geometric one (Tor.obj); DESIGN.md carries the measured mismatch rates.  Nearest-hit rule, eps and the
// tie-break on the caller's index are the same; shading still uses the stored plane's normal (scene.cpp:123).
__device__ __forceinline__ float pick3(V3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }

__device__ __forceinline__ bool watertight_test(float4 a, float4 b, float4 c, V3 o, V3 d, float *t_out) {
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    const int kz = ax > ay ? (ax > az ? 0 : 2) : (ay > az ? 1 : 2);
    int kx = kz == 2 ? 0 : kz + 1, ky = kx == 2 ? 0 : kx + 1;
    const float dz = pick3(d, kz);
    if (dz < 0.0f) { const int tmp = kx; kx = ky; ky = tmp; }  // keeps the winding
    const float Sz = rcp_approx(dz), Sx = pick3(d, kx) * Sz, Sy = pick3(d, ky) * Sz;
    const V3 A = mk(a.x, a.y, a.z) - o, B = mk(b.x, b.y, b.z) - o, C = mk(c.x, c.y, c.z) - o;
    const float Akz = pick3(A, kz), Bkz = pick3(B, kz), Ckz = pick3(C, kz);
    const float Ax = pick3(A, kx) - Sx * Akz, Ay = pick3(A, ky) - Sy * Akz;
    const float Bx = pick3(B, kx) - Sx * Bkz, By = pick3(B, ky) - Sy * Bkz;
    const float Cx = pick3(C, kx) - Sx * Ckz, Cy = pick3(C, ky) - Sy * Ckz;
    float U = Cx * By - Cy * Bx, V = Ax * Cy - Ay * Cx, W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;  // two-sided
    const float det = U + V + W;
    if (det == 0.0f) return false;
    const float T = U * (Sz * Akz) + V * (Sz * Bkz) + W * (Sz * Ckz);
    *t_out = T * rcp_approx(det);
    return true;
}

__device__ __forceinline__ void tri_intersect_wt(float4 a, float4 b, float4 c, int rec, const float4 *tris, V3 o, V3 d,
                                                 float eps, Hit &h) {
    float t;
    if (!watertight_test(a, b, c, o, d, &t)) return;
    if (!(t <= h.t) || t < eps) return;
    const int idx = __float_as_int(b.w);
    if (t == h.t) {
        if (h.idx == kIdxLazy) h.idx = __float_as_int(__ldg(tris + 4 * (size_t)h.rec + 2).w);
        if (!(idx < h.idx)) return;
    }
    h.t = t;
    h.idx = idx;
    h.rec = rec;
}

// the intersector the kernel was instantiated for (WT = watertight), on global record `rec`
template <bool WT>
__device__ __forceinline__ void tri_intersect(const SceneView &sc, int rec, V3 o, V3 d, float eps, Hit &h) {
    PT_CHK(rec >= 0 && rec < sc.n_tri, CHK_RECORD);
    const float4 *r = sc.tris + 4 * (size_t)rec;
    if (WT) {
        float4 pl, a, b, c;
        ldg256(r, pl, a);
        ldg256(r + 2, b, c);
        tri_intersect_wt(a, b, c, rec, sc.tris, o, d, eps, h);
    } else {
        tri_intersect_ref(r, rec, o, d, eps, h);
    }
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
// (Measured alternative, round 2, profiles/r02_sweep_bounce.txt: a range-only first pass (14 instructions per unit) with
// edge planes and exact tests only for the units in range, nearest first.  On the convex 12-wall room it saves 6 % of
// k_wf_bounce, but the per-lane loops over "my in-range units" diverge: Cornell, where the nearest plane is often not
// the hit, fell from 15.1 to 7.7 Grays/s, and with the exact tests gathered behind the loop still to 10.9.  The
// branch-free scan of every unit, 60 instructions each, wins because all 32 lanes run it together.)
// (tnear, rnear): smallest scanned distance among the marked units so far and the first record of that unit
// WT (the watertight mode's use of the scan, large_phase): the tolerance grows by (wt_c + 2e-6 |t|) / |d . n| — the
// watertight test decides at the TRUE hit point, which the scanned one misses by the rounding of numerator and
// denominator and by the distance between the stored plane and the geometric triangle (<= 1e-5 amax, k_large_pack).
template <bool WT>
__device__ __forceinline__ unsigned scan_quads(const float4 *units, int u0, int n, V3 o, V3 d, float t_lo, float t_hi,
                                               float wt_c, float &tnear, int &rnear) {
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
        const float tau = __fmaf_rn(fabsf(t), WT ? meta.y + 2e-6f : meta.y, WT ? meta.x + wt_c : meta.x) * fabsf(rc);
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
template <bool WT>
__device__ __forceinline__ unsigned scan_singles(const float4 *units, int s0, int n, int rec0, V3 o, V3 d, float t_lo,
                                                 float t_hi, float wt_c, float &tnear, int &rnear) {
    unsigned mask = 0u;
    for (int k = 0; k < n; ++k) {
        const float4 *r = units + kSingleRows * (s0 + k);
        const float4 p = r[0];
        const float sd = d.x * p.x + d.y * p.y + d.z * p.z;
        const float num = -(o.x * p.x + o.y * p.y + o.z * p.z + p.w);
        const float rc = rcp_approx(sd);
        const float t = num * rc;
        const float tau = WT ? __fmaf_rn(fabsf(t), 2e-6f, wt_c) * fabsf(rc) : 0.0f;
        const float tol = WT ? __fmaf_rn(fabsf(t), 2e-6f, tau) : fabsf(t) * 2e-6f;
        const float Px = __fmaf_rn(d.x, t, o.x), Py = __fmaf_rn(d.y, t, o.y), Pz = __fmaf_rn(d.z, t, o.z);
        const float4 e0 = r[1], e1 = r[2], e2 = r[3];
        const float g0 = __fmaf_rn(e0.x, Px, __fmaf_rn(e0.y, Py, __fmaf_rn(e0.z, Pz, e0.w)));
        const float g1 = __fmaf_rn(e1.x, Px, __fmaf_rn(e1.y, Py, __fmaf_rn(e1.z, Pz, e1.w)));
        const float g2 = __fmaf_rn(e2.x, Px, __fmaf_rn(e2.y, Py, __fmaf_rn(e2.z, Pz, e2.w)));
        // written so that a NaN anywhere (sd == 0, infinities) marks the triangle: the resolve decides
        const bool out = (WT ? (t - tau > t_hi) | (t + tau < t_lo) : (t > t_hi) | (t < t_lo)) | (fminf(fminf(g0, g1), g2) < -tol);
        mask |= (out ? 0u : 1u) << k;
        const bool nearer = !out && t < tnear;
        tnear = nearer ? t : tnear;
        rnear = nearer ? rec0 + s0 + k : rnear;
    }
    return mask;
}

// exact test on the marked records k0 + bit of the shared-memory record rows
template <bool WT>
__device__ __forceinline__ void large_resolve32(const float4 *s_large, int k0, unsigned mask, int rec_base, V3 o, V3 d,
                                                float eps, Hit &h) {
    while (mask) {
        const int k = k0 + __ffs(mask) - 1;
        mask &= mask - 1u;
        const float4 *r = s_large + 4 * k;
        if (WT) {  // h.idx is never lazy here (the shared-memory phase starts every query)
            tri_intersect_wt(r[1], r[2], r[3], rec_base + k, nullptr, o, d, eps, h);
            continue;
        }
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

// direct: bit k = large record k goes to the watertight test whatever its scan unit says (SceneView::direct_lo / hi)
template <bool WT>
__device__ __forceinline__ void large_phase(const float4 *s_large, int n_large, int n_quads, int rec_base, V3 o, V3 d,
                                            float eps, Hit &h, unsigned long long direct = 0ull, float wt_c = 0.0f) {
    // Watertight mode: the same scan in front of the watertight test.  The units bound what the reference accepts
    // around the stored plane — a dilation of the geometric triangle when that plane is the geometric one; records whose
    // plane is not (`vn`) are in `direct`.  The tolerances grow as scan_quads describes; a false mark costs one test.
    // (Round 2 tested every shared-memory triangle directly in this mode: k_wf_bounce took twice the parity mode's time.)
    const float t_hi = h.t + fabsf(h.t) * 2e-6f, t_lo = eps - fabsf(eps) * 2e-6f;
    const float4 *quads = s_large + 4 * n_large, *singles = quads + kQuadRows * n_quads;
    const int n_single = n_large - 2 * n_quads;
    // records 2u, 2u + 1 belong to quad u; record 2 n_quads + s to single s
    float tnear = __int_as_float(0x7f800000);
    int rnear = -1;
    unsigned q0 = scan_quads<WT>(quads, 0, min(n_quads, 16), o, d, t_lo, t_hi, wt_c, tnear, rnear);
    unsigned q1 = n_quads > 16 ? scan_quads<WT>(quads, 16, n_quads - 16, o, d, t_lo, t_hi, wt_c, tnear, rnear) : 0u;
    unsigned s0 = scan_singles<WT>(singles, 0, min(n_single, 32), 2 * n_quads, o, d, t_lo, t_hi, wt_c, tnear, rnear);
    unsigned s1 = n_single > 32 ? scan_singles<WT>(singles, 32, n_single - 32, 2 * n_quads, o, d, t_lo, t_hi, wt_c, tnear, rnear) : 0u;
    if (WT && direct) {
        const int nq2 = 2 * n_quads;
        const unsigned long long dq = nq2 >= 64 ? direct : direct & ((1ull << nq2) - 1ull), dsg = nq2 >= 64 ? 0ull : direct >> nq2;
        q0 |= (unsigned)dq; q1 |= (unsigned)(dq >> 32);
        s0 |= (unsigned)dsg; s1 |= (unsigned)(dsg >> 32);
    }
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
        large_resolve32<WT>(s_large, rnear, bits, rec_base, o, d, eps, h);
    }
    large_resolve32<WT>(s_large, 0, q0, rec_base, o, d, eps, h);
    if (q1) large_resolve32<WT>(s_large, 32, q1, rec_base, o, d, eps, h);
    large_resolve32<WT>(s_large, 2 * n_quads, s0, rec_base, o, d, eps, h);
    if (s1) large_resolve32<WT>(s_large, 2 * n_quads + 32, s1, rec_base, o, d, eps, h);
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

// Does the ray [0, tmax] have to walk the tree?  Tested against the two child boxes of the binary tree's root (tighter
// than the root box: on the 1M-triangle scene half of the rays that enter the root box leave it without entering one).
__device__ __forceinline__ bool reaches_root(const SceneView &sc, V3 o, V3 d, float tmax) {
    if (sc.root == kRefEmpty) return false;
    const float inf = __int_as_float(0x7f800000);
    // MUFU.RCP (within 2^-22 of 1 / d, +-inf for d == 0 like the division): the boxes are padded by 32 ulp of the scene's
    // largest coordinate and the exit distance is widened below
    const V3 inv = mk(rcp_approx(d.x), rcp_approx(d.y), rcp_approx(d.z));
    const float4 n0 = sc.root_n0, n1 = sc.root_n1, n2 = sc.root_n2;
    return box_entry(n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, o, inv, tmax) < inf ||
           box_entry(n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, o, inv, tmax) < inf;
}

// ---- 8-wide BVH (node layout: pt_scene.cuh, builder: pt_bvh.cuh k_collapse) ---------------------------------------
__device__ __forceinline__ float fmin3(float a, float b, float c) {  // FMNMX3 (sm_100)
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}

struct Ray8 {
    V3 o, d;
    V3 idir;        // 1 / d with |d| clamped away from 0: the slab arithmetic below must stay finite
    unsigned oct4;  // octant of the direction (bit 0 x, 1 y, 2 z; set = non-negative), replicated into the 4 bytes
};
__device__ __forceinline__ float clamped_rcp(float d) {
    // a component below 1e-18 moves the ray by less than 1e-16 over any distance inside the scene: far below the
    // padding of the boxes
    return 1.0f / (fabsf(d) > 1e-18f ? d : copysignf(1e-18f, d));
}
__device__ __forceinline__ void ray8_setup(V3 o, V3 d, Ray8 &r) {
    r.o = o;
    r.d = d;
    r.idir = mk(clamped_rcp(d.x), clamped_rcp(d.y), clamped_rcp(d.z));
    const unsigned oct = (r.idir.x >= 0.0f ? 1u : 0u) | (r.idir.y >= 0.0f ? 2u : 0u) | (r.idir.z >= 0.0f ? 4u : 0u);
    r.oct4 = oct * 0x01010101u;
}

// 0xff in every byte of x whose top bit is set (PRMT's sign-replicate mode: selector nibbles 8 + i.  __byte_perm masks
// the selector to 3 bits per nibble, so this has to be the PTX instruction itself)
__device__ __forceinline__ unsigned sign_bytes(unsigned x) {
    unsigned r;
    asm("prmt.b32 %0, %1, 0, 0xba98;" : "=r"(r) : "r"(x));
    return r;
}

// byte k of w as the float 1 + q 2^-15 (q in mantissa bits 8..15): one PRMT instead of a conversion.  `one` is
// 0x3f800000 handed in through SceneView::one_bits, i.e. a value neither nvcc nor ptxas can fold: PRMT takes one
// immediate, and with the constant as the immediate the four selectors lived in registers that were re-materialised 45
// times per node visit (ncu source view: 40 IMAD.U32 R, RZ, RZ, UR per visit).
template <int K>
__device__ __forceinline__ float qfloat(unsigned w, unsigned one) {
    return __uint_as_float(__byte_perm(w, one, 0x7604 | (K << 4)));
}


// four children (one meta word, slots 4 W .. 4 W + 3) against the ray; returns their hit bits
template <int K>
__device__ __forceinline__ unsigned child_hit(unsigned nx, unsigned fx, unsigned ny, unsigned fy, unsigned nz, unsigned fz,
                                              float Ax, float Bx, float Ay, float By, float Az, float Bz, float tmax,
                                              unsigned child_bits4, unsigned bit_index4, unsigned one) {
    // slab distances on the 2^15-scaled grid: (p + q s - o) / d = (1 + q 2^-15) (2^15 s / d) + ((p - o) / d - 2^15 s / d)
    const float tlx = __fmaf_rn(qfloat<K>(nx, one), Ax, Bx), thx = __fmaf_rn(qfloat<K>(fx, one), Ax, Bx);
    const float tly = __fmaf_rn(qfloat<K>(ny, one), Ay, By), thy = __fmaf_rn(qfloat<K>(fy, one), Ay, By);
    const float tlz = __fmaf_rn(qfloat<K>(nz, one), Az, Bz), thz = __fmaf_rn(qfloat<K>(fz, one), Az, Bz);
    const float tn = fmaxf(fmax3(tlx, tly, tlz), 0.0f);
    const float tf = fminf(fmin3(thx, thy, thz), tmax);  // clipped against the best t with <=: equal distances are kept
    const unsigned bits = (child_bits4 >> (8 * K)) & 0xffu, idx = (bit_index4 >> (8 * K)) & 0xffu;
    return tn <= tf ? bits << idx : 0u;
}

// Visit of wide node `ni`: all 8 child boxes against the ray [0, tmax].  Returns the hit mask — bits 24..31 the inner
// children, at position 24 + (slot ^ octant) so that the highest bit is the child to enter first; bits 0..23 the
// triangles of the hit leaf children (record tri_base + bit) — and the node's child_base / tri_base / imask.
// Conservative by construction of the grid (k_collapse) and the padding of the triangle boxes.
__device__ __forceinline__ unsigned node8_visit(const float4 *__restrict__ nodes, unsigned one, int ni, const Ray8 &r,
                                                float tmax, unsigned &child_base, unsigned &tri_base, unsigned &imask) {
    const float4 *nd = nodes + kNodeRows * (size_t)ni;
    float4 r0, r1, r2, r3, r4, r5;
    ldg256(nd, r0, r1);
    ldg256(nd + 2, r2, r3);
    ldg256(nd + 4, r4, r5);
    child_base = __float_as_uint(r1.z);
    tri_base = __float_as_uint(r1.w);
    imask = __float_as_uint(r2.z);
    const float Ax = r0.w * r.idir.x * 32768.0f, Ay = r1.x * r.idir.y * 32768.0f, Az = r1.y * r.idir.z * 32768.0f;
    const float Bx = (r0.x - r.o.x) * r.idir.x - Ax, By = (r0.y - r.o.y) * r.idir.y - Ay, Bz = (r0.z - r.o.z) * r.idir.z - Az;
    // near / far planes per axis by the direction's sign
    const bool px = (r.oct4 & 1u) != 0u, py = (r.oct4 & 2u) != 0u, pz = (r.oct4 & 4u) != 0u;
    const unsigned lx0 = __float_as_uint(r3.x), lx1 = __float_as_uint(r3.y), ly0 = __float_as_uint(r3.z), ly1 = __float_as_uint(r3.w);
    const unsigned lz0 = __float_as_uint(r4.x), lz1 = __float_as_uint(r4.y), hx0 = __float_as_uint(r4.z), hx1 = __float_as_uint(r4.w);
    const unsigned hy0 = __float_as_uint(r5.x), hy1 = __float_as_uint(r5.y), hz0 = __float_as_uint(r5.z), hz1 = __float_as_uint(r5.w);
    unsigned hits = 0u;
#pragma unroll
    for (int w = 0; w < 2; ++w) {
        const unsigned meta4 = __float_as_uint(w ? r2.y : r2.x);
        const unsigned is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;           // low 5 bits >= 24
        const unsigned inner_mask4 = sign_bytes(is_inner4 << 3);                    // 0xff in the bytes of inner children
        const unsigned bit_index4 = (meta4 ^ (r.oct4 & inner_mask4)) & 0x1f1f1f1fu;  // inner: 24 + (slot ^ octant); leaf: offset
        const unsigned child_bits4 = (meta4 >> 5) & 0x07070707u;                    // inner: 1; leaf: unary triangle count
        const unsigned nx = w ? (px ? lx1 : hx1) : (px ? lx0 : hx0), fx = w ? (px ? hx1 : lx1) : (px ? hx0 : lx0);
        const unsigned ny = w ? (py ? ly1 : hy1) : (py ? ly0 : hy0), fy = w ? (py ? hy1 : ly1) : (py ? hy0 : ly0);
        const unsigned nz = w ? (pz ? lz1 : hz1) : (pz ? lz0 : hz0), fz = w ? (pz ? hz1 : lz1) : (pz ? hz0 : lz0);
        hits |= child_hit<0>(nx, fx, ny, fy, nz, fz, Ax, Bx, Ay, By, Az, Bz, tmax, child_bits4, bit_index4, one);
        hits |= child_hit<1>(nx, fx, ny, fy, nz, fz, Ax, Bx, Ay, By, Az, Bz, tmax, child_bits4, bit_index4, one);
        hits |= child_hit<2>(nx, fx, ny, fy, nz, fz, Ax, Bx, Ay, By, Az, Bz, tmax, child_bits4, bit_index4, one);
        hits |= child_hit<3>(nx, fx, ny, fy, nz, fz, Ax, Bx, Ay, By, Az, Bz, tmax, child_bits4, bit_index4, one);
    }
    return hits;
}

// The scan record of one BVH triangle (plane + three edge planes): the conservative test of scan_singles with the same
// tolerances — approximate reciprocal, range widened by 2e-6 relative, edge planes already dilated by the builder; a
// NaN anywhere marks.  true = the reference's exact test has to decide.
template <bool WT>
__device__ __forceinline__ bool scan_pass(const SceneView &sc, int rec, V3 o, V3 d, float eps, const Hit &h) {
    PT_CHK(rec >= 0 && rec < sc.n_bvh, CHK_RECORD);
    if (WT) {  // watertight mode: the scan record bounds the REFERENCE's acceptance set around the stored plane, which
                     // need not contain the geometric hit; the test itself is cheap enough to stand in for its own filter
        const float4 *r = sc.tris + 4 * (size_t)rec;
        float4 pl, a, b, c;
        ldg256(r, pl, a);
        ldg256(r + 2, b, c);
        float t;
        return watertight_test(a, b, c, o, d, &t) && t <= h.t && !(t < eps);
    }
    const float4 *s = sc.scans + 4 * (size_t)rec;
    float4 p, e0, e1, e2;
    ldg256(s, p, e0);
    ldg256(s + 2, e1, e2);
    const float t_hi = h.t + fabsf(h.t) * 2e-6f, t_lo = eps - fabsf(eps) * 2e-6f;
    const float sd = d.x * p.x + d.y * p.y + d.z * p.z;
    const float num = -(o.x * p.x + o.y * p.y + o.z * p.z + p.w);
    const float t = num * rcp_approx(sd);
    const float tol = fabsf(t) * 2e-6f;
    const float Px = __fmaf_rn(d.x, t, o.x), Py = __fmaf_rn(d.y, t, o.y), Pz = __fmaf_rn(d.z, t, o.z);
    const float g0 = __fmaf_rn(e0.x, Px, __fmaf_rn(e0.y, Py, __fmaf_rn(e0.z, Pz, e0.w)));
    const float g1 = __fmaf_rn(e1.x, Px, __fmaf_rn(e1.y, Py, __fmaf_rn(e1.z, Pz, e1.w)));
    const float g2 = __fmaf_rn(e2.x, Px, __fmaf_rn(e2.y, Py, __fmaf_rn(e2.z, Pz, e2.w)));
    const bool out = (t > t_hi) | (t < t_lo) | (fminf(fminf(g0, g1), g2) < -tol);
    return !out;
}

// One BVH triangle: scan record first, the reference's exact test if that marks it.  Returns whether the exact test ran.
template <bool WT>
__device__ __forceinline__ bool tri_test(const SceneView &sc, int rec, V3 o, V3 d, float eps, Hit &h) {
    if (!scan_pass<WT>(sc, rec, o, d, eps, h)) return false;
    tri_intersect<WT>(sc, rec, o, d, eps, h);
    return true;
}

// Traversal state of one ray: a group of inner children still to enter (x = child_base, y = hit bits 24..31 | imask)
// and a group of triangles still to test (x = tri_base, y = bits 0..23).  Stack entries are one or the other, told
// apart by y > 0x00ffffff.
__device__ __forceinline__ bool has_inner(uint2 g) { return g.y > 0x00ffffffu; }

// takes the nearest inner child out of node group `ng` (re-pushing the rest), visits it and leaves its hits in ng / tg
template <bool COUNT>
__device__ __forceinline__ void enter_child(const SceneView &sc, const Ray8 &r, float tmax, uint2 &ng, uint2 &tg,
                                            uint2 *stack, int &sp, unsigned long long &n_nodes) {
    const unsigned hits = ng.y;
    const int bit = 31 - __clz((int)hits);
    ng.y &= ~(1u << bit);
    if (has_inner(ng)) {
        PT_CHK(sp < kStackDepth, CHK_STACK);
        stack[sp++] = ng;
    }
    const unsigned slot = (unsigned)(bit - 24) ^ (r.oct4 & 7u);
    const int ni = (int)ng.x + __popc(hits & 0xffu & ~(0xffffffffu << slot));
    PT_CHK(ni >= 0 && ni < sc.n_nodes, CHK_NODE);
    unsigned cb, tb, im;
    const unsigned hm = node8_visit(sc.nodes, sc.one_bits, ni, r, tmax, cb, tb, im);
    if (COUNT) ++n_nodes;
    ng = make_uint2(cb, (hm & 0xff000000u) | im);
    tg = make_uint2(tb, hm & 0x00ffffffu);
}

// all triangles of group tg
template <bool COUNT, bool WT>
__device__ __forceinline__ void test_group(const SceneView &sc, const Ray8 &r, float eps, Hit &h, uint2 &tg,
                                           unsigned long long &n_tris, unsigned long long &n_exact) {
    while (tg.y) {
        const int j = 31 - __clz((int)tg.y);
        tg.y &= ~(1u << j);
        const bool exact = tri_test<WT>(sc, (int)tg.x + j, r.o, r.d, eps, h);
        if (COUNT) { ++n_tris; n_exact += exact ? 1u : 0u; }
    }
}

// the whole walk in one thread (pt_intersect, the megakernel)
template <bool COUNT, bool WT>
__device__ __forceinline__ void bvh8_traverse(const SceneView &sc, V3 o, V3 d, float eps, Hit &h,
                                              unsigned long long &n_nodes, unsigned long long &n_tris) {
    Ray8 r;
    ray8_setup(o, d, r);
    uint2 stack[kStackDepth];
    int sp = 0;
    uint2 ng = make_uint2(0u, 0x80000000u), tg = make_uint2(0u, 0u);  // the root as the only child of a virtual group
    unsigned long long n_exact = 0;
    while (true) {
        if (has_inner(ng)) {
            enter_child<COUNT>(sc, r, h.t, ng, tg, stack, sp, n_nodes);
        } else {  // a triangle group popped from the stack
            tg = ng;
            ng = make_uint2(0u, 0u);
        }
        test_group<COUNT, WT>(sc, r, eps, h, tg, n_tris, n_exact);
        if (!has_inner(ng)) {
            if (sp == 0) break;
            ng = stack[--sp];
        }
    }
}

// scene.cpp:114-120 as written: every triangle, in order (the checker mode, PT_ACCEL_BRUTE)
template <bool COUNT, bool WT>
__device__ __forceinline__ Hit brute_hit(const SceneView &sc, V3 o, V3 d, float eps, unsigned long long &n_tris) {
    Hit h = no_hit();
    for (int k = 0; k < sc.n_tri; ++k) tri_intersect<WT>(sc, k, o, d, eps, h);
    if (COUNT) n_tris += sc.n_tri;
    return h;
}

// Whole nearest-hit query in one thread (pt_intersect, the megakernel).
template <bool COUNT, bool WT>
__device__ __forceinline__ Hit nearest_hit(const SceneView &sc, const float4 *s_large, V3 o, V3 d, float eps,
                                           unsigned long long &n_nodes, unsigned long long &n_tris) {
    if (sc.brute) return brute_hit<COUNT, WT>(sc, o, d, eps, n_tris);
    Hit h = no_hit();
    large_phase<WT>(s_large, sc.n_large, sc.n_quads, sc.n_bvh, o, d, eps, h, ((unsigned long long)sc.direct_hi << 32) | sc.direct_lo, sc.wt_cabs);
    if (COUNT) n_tris += sc.n_large;
    if (!reaches_root(sc, o, d, h.t)) return h;
    bvh8_traverse<COUNT, WT>(sc, o, d, eps, h, n_nodes, n_tris);
    return h;
}

}  // namespace ptd
