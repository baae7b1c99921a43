// srt_variants.h — the render kernel is compiled once per feature mask, one translation unit per
// variant (srt_render_variant.cu compiled with -DSRT_VARIANT=<n>), so that the variants build in
// parallel and each stays small enough for the instruction cache (profiles/r01_notes.md: the
// all-features kernel was 22 600 SASS instructions and spent 51 % of its warp-cycles waiting for
// instruction fetch on a scene that executes 2 200 of them).
#ifndef SRT_VARIANTS_H
#define SRT_VARIANTS_H

#include <cuda_runtime.h>

#include "srt_types.h"

namespace srt {
struct RenderArgs;
}

struct RenderLaunch {
    const DScene* scene;
    const srt::RenderArgs* args;
    const uint32_t* sobol;
    long long* out;  // fixed-point accumulation buffer, 3 per pixel
    unsigned long long* counters;
    int grid;  // 0: only query (blocks_per_sm, registers)
    size_t smem_bytes;
    cudaStream_t stream;
};

struct RenderVariantInfo {
    uint32_t features;  // F_* the variant was compiled with (including F_SMEM_TABLES)
    int counting;       // 1: gathers SrtCounters
    int blocks_per_sm;  // filled by the launcher from the occupancy calculator
    int registers;
};

typedef cudaError_t (*RenderVariantFn)(const RenderLaunch& L, RenderVariantInfo* info);

// scene classes the variants are cut for
#define SRT_VM_LIST (F_XFORM | F_METAL | F_DIELECTRIC)
#define SRT_VM_SPHERES (SRT_VM_LIST | F_BVH_SPHERE | F_TEX_CHECKER | F_LIGHT_SPHERE)
#define SRT_VM_MESH                                                                                     \
    (SRT_VM_SPHERES | F_BVH_TRI | F_BVH_GENERIC | F_TRI | F_BECKMANN | F_ORENNAYAR | F_TEX_IMAGE | F_LIGHT_TRI | \
     F_LIGHT_NONE)
#define SRT_VM_FULL F_ALL_SCENE

// variant n -> (mask, counting)
// variant 7: corrected-transport mode (SRT_RENDER_CORRECTED, srt_shade.cuh), all features, tables in global memory
#define SRT_N_VARIANTS 8
#define SRT_VARIANT_MASK(n)                                  \
    ((n) == 0   ? (SRT_VM_LIST | F_SMEM_TABLES)              \
     : (n) == 1 ? (SRT_VM_SPHERES | F_SMEM_TABLES)           \
     : (n) == 2 ? (SRT_VM_MESH | F_SMEM_TABLES)              \
     : (n) == 3 ? (SRT_VM_FULL | F_SMEM_TABLES)              \
     : (n) == 4 ? (SRT_VM_MESH)                              \
     : (n) == 5 ? (SRT_VM_FULL)                              \
     : (n) == 6 ? (SRT_VM_FULL)                              \
                : (SRT_VM_FULL | F_CORRECTED))
#define SRT_VARIANT_COUNTING(n) ((n) == 6)
// Resident CTAs per SM the register allocator is bounded for (__launch_bounds__).  Measured on B200
// (profiles/r01_notes.md, capture C): the list and mesh kernels gain 30-40 % from 4 CTAs (64
// registers, < 200 bytes of spills) because their stalls are fixed-latency waits and instruction
// fetch that more warps hide; the sphere-BVH kernel spills ~1 KB at 64 registers and is best at 2.
#define SRT_VARIANT_MINBLOCKS(n) ((n) == 1 || (n) == 7 ? 2 : 4)

// ---- the sorted wavefront kernel (srt_wavefront.cuh), one translation unit per mask -------------------
struct WfLaunch {
    const DScene* scene;
    const srt::RenderArgs* args;
    const void* state;  // srt::WfState
    const uint32_t* sobol;
    long long* out;  // fixed-point accumulation buffer, 3 per pixel
    unsigned long long* counters;
    int per_thread;  // slots per thread of the CTA
    int grid;
    int threads;  // threads per CTA (<= the variant's SRT_WF_VARIANT_PERSIST_THREADS); 0: that maximum
    size_t smem_bytes;
    cudaStream_t stream;
};

struct WfVariantInfo {
    uint32_t features;
    int blocks_per_sm, registers, threads;
    int ctas_per_sm;  // CTAs the launch puts on one SM (SRT_WF_VARIANT_CTAS_PER_SM)
    int slot_bytes;  // bytes of one slot record (srt_wavefront.cuh: kSlotQuads x 16)
};

typedef cudaError_t (*WfVariantFn)(const WfLaunch& L, WfVariantInfo* info);

// variant 4 is the COUNTING build of k_wf_sorted (all features, tables in global memory): srt_render_counted
// replays a launch on it, so that the work counters come from the scheduler that was timed
// variant 5: corrected-transport mode, all features
// variants 6, 7: the two mesh classes cut to what a triangle-mesh scene actually instantiates — one bvh kind (the
// traversal's kind dispatch folds away), no loose triangles, sphere bvh, Oren-Nayar, checker or sphere / triangle
// lights: a textured microfacet mesh in a list of rects / boxes / spheres (C3, C4), and a mesh with a
// constant_medium (C5).  Less code per phase and fewer registers (srt_api.cu tries them before 2 and 3).
#define SRT_VM_TRIMESH (SRT_VM_LIST | F_BVH_TRI | F_BECKMANN | F_TEX_IMAGE)
#define SRT_VM_TRIVOLUME (SRT_VM_LIST | F_BVH_TRI | F_MEDIA)
#define SRT_N_WF_VARIANTS 8
#define SRT_WF_VARIANT_MASK(n)                                 \
    ((n) == 0   ? (SRT_VM_LIST | F_SMEM_TABLES)                \
     : (n) == 1 ? (SRT_VM_SPHERES | F_SMEM_TABLES)             \
     : (n) == 2 ? (SRT_VM_MESH)                                \
     : (n) == 5 ? (SRT_VM_FULL | F_CORRECTED)                  \
     : (n) == 6 ? (SRT_VM_TRIMESH)                             \
     : (n) == 7 ? (SRT_VM_TRIVOLUME)                           \
                : (SRT_VM_FULL))
// the order in which pick_wf_variant tries them: the smallest covering mask first
#define SRT_WF_VARIANT_ORDER {0, 1, 6, 2, 7, 3, 4, 5}
#define SRT_WF_VARIANT_COUNTING(n) ((n) == 4)
// threads of the one persistent CTA per SM (k_wf_sorted).  Measured (profiles/r01_notes.md,
// capture K): the sphere-BVH variant needs 107 registers and is best at 640 threads (96 registers, 20
// warps); the mesh and full variants spill there and are best at 512 (128 registers); the two lean mesh
// variants (6, 7) fit 96 registers with 32-96 bytes of spills and gain 1-3 % from the four extra warps.
#define SRT_WF_VARIANT_PERSIST_THREADS(n) ((n) == 7 ? 320 : ((n) <= 1 || (n) == 6 ? 640 : 512))
// CTAs per SM.  One everywhere but on the triangle-bvh + constant_medium variant (7), whose frames spend 36-39 % of
// their stall samples at the barrier that closes the trace phase (a group lasts as long as its longest traversal):
// two CTAs of 320 threads share the SM's slots, and one traces while the other waits.  Measured (profiles/
// r02_notes.md, capture AA): C5 at 10 M triangles -9.5 %, C5-like -0.6 %; the same split is +2.5 % on the sphere
// variant, -4 % on C3 but +13 % on C4 (both variant 6: two CTAs in different phases miss the instruction cache more).
#define SRT_WF_VARIANT_CTAS_PER_SM(n) ((n) == 7 ? 2 : 1)
// Variant 6 (list classes + triangle bvh + Beckmann + image textures) decides per scene, at launch: its 640-thread CTA is
// launched as two CTAs of 320 threads (same registers, same slots per SM) when the scene's wide BVHs hold at least this
// many nodes — deep traversals of very different lengths, a long wait at the trace barrier (C3: -4.4 %) — and as one
// CTA for small meshes among many materials, where two CTAs in different phases miss the instruction cache (C4: +13 %).
#define SRT_WF_SPLIT_CTA_MIN_NODES 65536

cudaError_t srt_wavefront_variant_0(const WfLaunch&, WfVariantInfo*);
cudaError_t srt_wavefront_variant_1(const WfLaunch&, WfVariantInfo*);
cudaError_t srt_wavefront_variant_2(const WfLaunch&, WfVariantInfo*);
cudaError_t srt_wavefront_variant_3(const WfLaunch&, WfVariantInfo*);
cudaError_t srt_wavefront_variant_4(const WfLaunch&, WfVariantInfo*);
cudaError_t srt_wavefront_variant_5(const WfLaunch&, WfVariantInfo*);
cudaError_t srt_wavefront_variant_6(const WfLaunch&, WfVariantInfo*);
cudaError_t srt_wavefront_variant_7(const WfLaunch&, WfVariantInfo*);

cudaError_t srt_render_variant_0(const RenderLaunch&, RenderVariantInfo*);
cudaError_t srt_render_variant_1(const RenderLaunch&, RenderVariantInfo*);
cudaError_t srt_render_variant_2(const RenderLaunch&, RenderVariantInfo*);
cudaError_t srt_render_variant_3(const RenderLaunch&, RenderVariantInfo*);
cudaError_t srt_render_variant_4(const RenderLaunch&, RenderVariantInfo*);
cudaError_t srt_render_variant_5(const RenderLaunch&, RenderVariantInfo*);
cudaError_t srt_render_variant_6(const RenderLaunch&, RenderVariantInfo*);
cudaError_t srt_render_variant_7(const RenderLaunch&, RenderVariantInfo*);

#endif
