// pt_scene.cuh — device-resident scene layout shared by the LBVH builder and the render kernels.
//
// HBM layout (all 16-byte aligned, read with LDG.128):
//   tris   4 x float4 per triangle, in BVH (Morton) order, the "large" triangles last:
//            [0] Triangle::plane_            (nx, ny, nz, w)            triangles.h:21,40-44
//            [1] vertices_[0], square        (x, y, z, S)               triangles.h:23-24
//            [2] vertices_[1], caller index  (x, y, z, bits(int))
//            [3] vertices_[2], material id   (x, y, z, bits(int))
//   nodes  4 x float4 per BVH2 node (both child boxes live in the parent, one 64-byte fetch per visit):
//            [0] child0 lo.x hi.x lo.y hi.y   [1] child1 lo.x hi.x lo.y hi.y
//            [2] child0 lo.z hi.z, child1 lo.z hi.z      [3] bits: child0 ref, child1 ref, 0, 0
//          child ref >= 0: node index; < 0: leaf, bits 0..27 = first triangle record, bits 28..30 = count-1
//   mats   3 x float4 per material: (Kd.xyz, bits(lobe kind)), (Ks.xyz, mirror chance), (Ni, Tr, bits(lobe mask),
//          diffuse chance) — the third row is read for MAT_MULTI (refractive extension) only
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ptd {

#ifndef PT_LEAF_MAX
#define PT_LEAF_MAX 4
#endif
constexpr int kLeafMax = PT_LEAF_MAX;  // triangles per LBVH leaf (<= 8: leaf_count)
constexpr int kMaxLarge = 64;        // triangles kept out of the LBVH and tested from shared memory first
constexpr int kAllLargeMax = 48;     // scenes up to this size keep every triangle in shared memory (no LBVH)
// Shared-memory image of the large triangles (k_large_pack), rows of float4:
//   [0, 4 n_large)            the 4 record rows of every large triangle (same layout as `tris`)
//   then n_quads x kQuadRows  scan units of two near-coplanar triangles sharing an edge (records 2u, 2u + 1):
//                             plane of the first, 4 outer edge planes, the shared edge's plane, (tolerances, threshold)
//   then n_single x kSingleRows  scan units of the remaining triangles (records 2 n_quads + s): plane, 3 edge planes
constexpr int kQuadRows = 7, kSingleRows = 4;
constexpr int kLargeRowsMax = 4 * kMaxLarge + kSingleRows * kMaxLarge;  // worst case: no pairs
constexpr int kStackDepth = 96;      // traversal stack entries (LBVH depth <= 63 key bits + index tie-break bits)
constexpr int kStatWords = 24;       // device statistics: 4 totals, queued rays, 16 log2 buckets of nodes per queued ray
constexpr int kRefEmpty = 0x7fffffff;  // child ref of an empty subtree / stack sentinel

// lobe kinds built by Factory (material.h:58-106)
// MAT_MULTI (extension): a material with the refractive lobe; its lobe set is the mask in row 2 (1 mirror, 2 diffuse, 4 refract)
enum : int { MAT_EMISSIVE = 0, MAT_NONE = 1, MAT_MIRROR = 2, MAT_DIFFUSE = 3, MAT_BOTH = 4, MAT_MULTI = 5 };

__host__ __device__ __forceinline__ int leaf_ref(int first, int count) {
    return (int)(0x80000000u | ((uint32_t)(count - 1) << 28) | (uint32_t)first);
}
__host__ __device__ __forceinline__ int leaf_first(int ref) { return ref & 0x0fffffff; }
__host__ __device__ __forceinline__ int leaf_count(int ref) { return ((ref >> 28) & 7) + 1; }

struct SceneView;
__host__ __device__ __forceinline__ int large_rows(int n_large, int n_quads) {
    return 4 * n_large + kQuadRows * n_quads + kSingleRows * (n_large - 2 * n_quads);
}

struct SceneView {
    const float4 *tris;
    const float4 *nodes;
    const float4 *mats;
    int n_tri;    // all triangle records
    int n_bvh;    // records [0, n_bvh) are in the LBVH
    int n_large;  // records [n_bvh, n_bvh + n_large) are tested first, from shared memory
    int n_quads;  // of which the first 2 n_quads form quad scan units
    int root;     // child ref of the LBVH root (kRefEmpty if n_bvh == 0)
    int brute;    // 1: test records [0, n_tri) in order from global memory, no LBVH (checker mode, PT_ACCEL_BRUTE)
    const float4 *lrecs;     // shared-memory image of the large triangles (layout above), large_rows() float4
    float4 root_lo, root_hi; // box of the LBVH root
    float4 root_n0, root_n1, root_n2;  // the root node's two child boxes (node layout rows 0..2) when root == 0
    const unsigned char *sky;  // equirect skybox, 24-bit BGR rows top-down (bitmap_image's layout); nullptr = none
    int sky_w, sky_h;
    unsigned long long *counters;  // [0] nodes visited, [1] triangles tested; nullptr = off
};

}  // namespace ptd
