// srt_types.h — device-resident scene layout shared by the host builder and the CUDA kernels.
//
// Everything the kernels read lives in ONE device allocation per scene; `DScene` is the table of
// typed pointers into it (passed to kernels by value, 200-odd bytes of kernel parameters, so every
// field is a uniform-register load).  Layout rules:
//   * 16-byte-aligned records read with one vector load: spheres float4, rects 2 x float4,
//     packed leaf triangles 3 x float4, wide BVH nodes 8 x float4;
//   * the flattened reference graph (entries / ops / groups / prims) is kept as it came through the
//     C ABI (include/srt_ffi.h) — it is small and L1/constant-cache resident.
#ifndef SRT_TYPES_H
#define SRT_TYPES_H

#include <stdint.h>

#include "srt_ffi.h"

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#else
struct alignas(16) float4 {
    float x, y, z, w;
};
struct alignas(16) int4 {
    int x, y, z, w;
};
#endif

// ---- wide BVH (4-ary, full-precision child boxes, SoA per node; 128 B = 8 x 16 B loads) --------
// child[i] >= 0            : index of an inner node
// child[i] <  0, count > 0 : leaf; first = ~child[i] is the first slot of the bvh's leaf arrays
// count[i] == 0 and child[i] < 0 : empty slot (its box is inverted, never hit)
struct alignas(16) WideNode4 {
    float4 lox, loy, loz;
    float4 hix, hiy, hiz;
    int4 child;
    int4 count;
};

// ---- the reference's own bvh_node tree (bvh.h:96-119), kept only as boxes ---------------------
// A primitive inside a reference bvh_node is reachable only if every ancestor box passes
// aabb::hit (aabb.h:33-49) with the entry (t_min, t_max); the reference boxes are un-padded, so
// this test decides real pixels.  The kernels find candidates with their own SAH
// tree and then replay the candidate's ancestor chain here.  Node over leaf ranks [lo, lo+n):
// n <= 2 -> no children (left = right = -1); else left covers [lo, lo+n/2), right the rest.
struct alignas(16) RefNode {
    float lo[3];
    int32_t left;  // node index or -1
    float hi[3];
    int32_t right;  // node index or -1
    int32_t mid;    // first rank of the right subtree
    int32_t _pad[3];
};

enum { BVH_KIND_GENERIC = 0, BVH_KIND_SPHERE = 1, BVH_KIND_TRIANGLE = 2 };

// One per SrtBvh.  Leaf slot s of this bvh (s in [0, item_count)) holds:
//   leaf_rank[leaf_base + s]   : rank of the item in the reference tree's in-order leaf sequence
//                                (== position in bvh_items; ties on t go to the higher rank)
//   kind SPHERE   : leaf_geom[geom_base + s]            = (cx, cy, cz, r)
//   kind TRIANGLE : leaf_geom[geom_base + 3*s .. +2]    = (p0, .), (p1, .), (p2, .)
//   kind GENERIC  : the item's group is groups[bvh_items[first_item + rank]]
struct DBvh {
    int32_t kind;
    int32_t first_item;
    int32_t item_count;
    int32_t node_base;  // into wide_nodes
    int32_t ref_base;   // into ref_nodes (root of the reference tree)
    int32_t leaf_base;  // into leaf_rank
    int32_t geom_base;  // into leaf_geom (float4 units)
    int32_t _pad;
};

struct DImage {
    const uint8_t* data;
    int32_t nx, ny;
};

struct DScene {
    const SrtEntry* entries;
    const SrtXformOp* ops;
    const SrtGroup* groups;
    const SrtPrim* prims;
    const int32_t* bvh_items;
    const SrtMedium* media;
    const DBvh* bvhs;
    const WideNode4* wide_nodes;
    const RefNode* ref_nodes;
    const int32_t* leaf_rank;
    const float4* leaf_geom;
    const float4* spheres;  // [n][1]
    const float* msph;      // [n][9]
    const float4* rects;    // [n][2] = (a0, a1, b0, b1), (k, 0, 0, 0)
    const float* tri_pos;   // [n][9]
    const float* tri_nrm;   // [n][9]
    const float* tri_uv;    // [n][6]
    const SrtLight* lights;
    const SrtMaterial* materials;
    const SrtTexture* textures;
    const DImage* images;
    const float* perlin_ranvec;  // [256][3]
    const int32_t* perlin_perm;  // [3][256]
    const double* const* merl;
    int32_t n_entries, n_prims, n_lights, n_media;
    int32_t n_materials, n_textures, n_bvhs, _pad;
    SrtCamera camera;
};

#endif  // SRT_TYPES_H
