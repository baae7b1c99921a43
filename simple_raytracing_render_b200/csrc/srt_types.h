// srt_types.h — device-resident scene layout shared by the host builder and the CUDA kernels.
//
// Everything the kernels read lives in ONE device allocation per scene; `DScene` is the table of
// typed pointers into it (passed to kernels by value, so every field is a uniform-register load).
// Layout rules:
//   * every record the inner loops touch is a whole number of 16-byte units read with vector
//     loads: DPrim 2 x 16 B (geometry + plane / type / material / flags inline: one load chain per
//     primitive instead of prims[] -> rects[] / spheres[]), DEntry 4 x 16 B (with conservative
//     local-frame bounds), DMaterial 3 x 16 B (constant textures folded in), wide BVH nodes
//     8 x 16 B, packed leaf triangles 3 x 16 B;
//   * the "hot" tables every ray walks (entries, instance ops, primitive records, lights) are
//     contiguous at the front of the blob so that a CTA can stage them into shared memory with one
//     bulk copy (srt_kernels.cuh) when they fit; everything else stays in global memory and is
//     L1 / L2 resident (no scene of BASELINE.json:configs C1-C4 exceeds the 126 MB L2).
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

// ---- kernel specialisation: which reference classes a scene instantiates ------------------------
// The render kernel is compiled for a few feature masks (srt_variants.h); a scene runs on the
// smallest variant whose mask covers its own.  Always present: rect / sphere groups, lambertian,
// brdfmaterial (samples like lambertian, material.h:217-236), diffuse_light, constant_texture,
// xz_rect lights.
enum : uint32_t {
    F_XFORM = 1u << 0,        // translate / rotate_x / rotate_y above a world entry
    F_BVH_SPHERE = 1u << 1,   // a bvh_node whose leaves are all spheres
    F_BVH_TRI = 1u << 2,      // a bvh_node whose leaves are all triangles
    F_BVH_GENERIC = 1u << 3,  // any other bvh_node
    F_TRI = 1u << 4,          // triangle outside a triangle-only bvh (lists, lights, generic bvh)
    F_MSPHERE = 1u << 5,      // moving_sphere
    F_ENV = 1u << 6,          // env_sphere as a world object
    F_MEDIA = 1u << 7,        // constant_medium (+ isotropic)
    F_ORENNAYAR = 1u << 8,
    F_BECKMANN = 1u << 9,
    F_METAL = 1u << 10,
    F_DIELECTRIC = 1u << 11,
    F_TEX_CHECKER = 1u << 12,
    F_TEX_NOISE = 1u << 13,
    F_TEX_IMAGE = 1u << 14,
    F_LIGHT_SPHERE = 1u << 15,
    F_LIGHT_TRI = 1u << 16,
    F_LIGHT_ENV = 1u << 17,
    F_LIGHT_NONE = 1u << 18,  // a light-list element without pdf_value / random (hitable.h:31-32)
    F_SMEM_TABLES = 1u << 24, // launch option, not a scene property: hot tables staged in smem
    F_CORRECTED = 1u << 25,   // launch option: corrected-transport mode (SRT_RENDER_CORRECTED, srt_shade.cuh)
    F_ALL_SCENE = (1u << 19) - 1u
};

// ---- flat primitive record: 32 B, index in the array == primitive id (SrtSceneDesc.prims) --------
// rect            : g = (a0, a1, b0, b1), k = plane
// sphere / env    : g = (cx, cy, cz, r)
// moving / triangle: aux = row in msph[] / tri_*[]
struct alignas(16) DPrim {
    float4 g;
    float k;
    int32_t type;      // SRT_PRIM_*
    int32_t material;  // index into materials[] or -1
    uint32_t flags_aux;  // bit 0 = SRT_PRIM_FLIP; bits 1.. = aux row (flags_aux >> 1)
};

// ---- world entry: 64 B --------------------------------------------------------------------------
enum : uint32_t {
    DENTRY_CULL = 1u,  // bounds are valid and finite: a ray missing them cannot hit the geometry
    DENTRY_TCULL = 2u  // every leaf below reports a parametric t and honours (t_min, t_max)
};
struct alignas(16) DEntry {
    int32_t first_prim, prim_count;  // group entries: range in dprims[]
    int32_t first_op, op_count;
    int32_t geom_kind, geom_index;   // SRT_GEOM_*; bvhs[] index for SRT_GEOM_BVH
    int32_t medium;
    uint32_t flags;                  // DENTRY_*
    float lo[3];                     // conservative bounds of the geometry in the entry's own frame
    float _p0;
    float hi[3];
    float _p1;
};

// ---- material with its texture head folded in: 48 B ---------------------------------------------
struct alignas(16) DMaterial {
    int32_t type;      // SRT_MAT_*
    int32_t texture;   // index into textures[] (-1: none)
    int32_t tex_type;  // SRT_TEX_* of that texture (-1: none); CONSTANT -> c is its colour
    int32_t _pad;
    float p[4];        // p0..p3 of SrtMaterial
    float c[3];        // constant-texture colour, or p0..p2 for brdf / metal albedo
    float _pad2;
};

// ---- wide BVH (4-ary, full-precision child boxes, SoA per node; 128 B = 8 x 16 B loads) --------
// child[i] >= 0 : index of an inner node; child[i] < 0 : leaf whose first slot in the bvh's leaf
// arrays is ~child[i] (builder bookkeeping; the kernels read `ref` only)
// ref[i]        : what traversal follows, ready to use: the inner node index (>= 0), or
//                 ~((first << 3) | count) for a leaf of `count` (1..7) slots, or 0x80000000 for an
//                 unused slot (its box is inverted as well, never hit)
struct alignas(16) WideNode4 {
    float4 lox, loy, loz;
    float4 hix, hiy, hiz;
    int4 child;
    int4 ref;
};

// ---- the reference's own bvh_node tree (bvh.h:96-119), kept only as boxes ---------------------
// A primitive inside a reference bvh_node is reachable only if every ancestor box passes
// aabb::hit (aabb.h:33-49) with the entry (t_min, t_max); the reference boxes are un-padded, so
// this test decides real pixels.  The kernels find the winner with their own SAH
// tree and then replay its ancestor chain here.  Node over leaf ranks [lo, lo+n):
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
//   kind TRIANGLE : leaf_geom[geom_base + 3*s .. +2]    = (p0, rank bits), (p1, primitive id bits), (p2, 0): the
//                   traversal reads neither leaf_rank nor bvh_items / groups for a triangle leaf
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
    // hot tables (contiguous in the blob, in this order; hot_bytes covers all four)
    const DEntry* entries;
    const SrtXformOp* ops;
    const DPrim* dprims;
    const DPrim* lights;  // the light_shape list as primitive records (material = -1)
    // the rest
    const SrtGroup* groups;
    const int32_t* bvh_items;
    const SrtMedium* media;
    const DBvh* bvhs;
    const WideNode4* wide_nodes;
    const RefNode* ref_nodes;
    const int32_t* leaf_rank;
    const float4* leaf_geom;
    const float* msph;      // [n][9]
    const float* tri_pos;   // [n][9]
    const float* tri_nrm;   // [n][9]
    const float* tri_uv;    // [n][6]
    const DMaterial* materials;
    const SrtTexture* textures;
    const DImage* images;
    const float* perlin_ranvec;  // [256][3]
    const int32_t* perlin_perm;  // [3][256]
    const double* const* merl;
    int32_t n_entries, n_prims, n_lights, n_media;
    int32_t n_materials, n_textures, n_bvhs, n_ops;
    unsigned long long* diag;  // [0]: wide-BVH stack overflows since the context was created (SrtCounters)
    uint32_t hot_bytes;  // bytes from `entries` to the end of `lights` (16-byte multiple)
    uint32_t features;   // F_* mask of the scene
    SrtCamera camera;
};

#endif  // SRT_TYPES_H
