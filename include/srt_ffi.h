/*
 * srt_ffi.h — C ABI of the B200-native path-tracing backend (libsrt_b200.so).
 *
 * The reference (truemeat001/Simple-Raytracing-Render) has no FFI: its surface is the set of
 * header-defined C++ classes and the render loop in Raytracing_n/Raytracing_n.cpp.  This header is
 * the boundary a binding of that loop attaches to.  Every entry point cites the reference
 * interface it stands in for (paths relative to the reference root; Raytracing_n.cpp line numbers
 * are those of its UTF-8 conversion).
 *
 * Conventions
 *   - plain C, caller-owned arrays in, opaque handles out;
 *   - every function returns 0 on success, a negative SRT_E_* code on failure, never exits and
 *     never throws; srt_last_error() returns the message of the calling thread's last failure;
 *   - a scene is immutable after srt_scene_create (as the reference's scene graph is after
 *     construction) and may be rendered from one host thread at a time per context;
 *   - device work is issued on the CUDA stream given to srt_context_create.
 *
 * The scene description (SrtSceneDesc) is the reference object graph flattened into arrays; the
 * flattening rules are in DESIGN.md §3 and are implemented twice: for the reference's own C++
 * classes in simple_raytracing_render_b200/host/srt_flatten.hpp and for the Python mirror of those
 * classes in simple_raytracing_render_b200/scene.py.
 */
#ifndef SRT_FFI_H
#define SRT_FFI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRT_ABI_VERSION 2

/* ---- status codes ------------------------------------------------------------------------ */
enum {
    SRT_OK = 0,
    SRT_E_INVALID = -1,     /* bad argument / malformed description */
    SRT_E_UNSUPPORTED = -2, /* object graph shape the flattener / backend does not cover */
    SRT_E_CUDA = -3,        /* CUDA runtime failure (message holds cudaGetErrorString) */
    SRT_E_NOMEM = -4,
    SRT_E_NODEVICE = -5     /* no CUDA device: there is no CPU fallback */
};

/* ---- primitives: the leaf hitables ------------------------------------------------------- */
enum {
    SRT_PRIM_SPHERE = 0,        /* sphere.h:17-30          data: spheres[4*i]   = cx cy cz r        */
    SRT_PRIM_MOVING_SPHERE = 1, /* moving_sphere.h:4-17    data: msph[9*i]      = c0 c1 t0 t1 r     */
    SRT_PRIM_XY_RECT = 2,       /* aarect.h:5-35           data: rects[5*i]     = a0 a1 b0 b1 k     */
    SRT_PRIM_XZ_RECT = 3,       /* aarect.h:37-64          (a,b) = (x,y) / (x,z) / (y,z)            */
    SRT_PRIM_YZ_RECT = 4,       /* aarect.h:66-94                                                   */
    SRT_PRIM_TRIANGLE = 5,      /* triangle.h:9-106        data: tri_pos[9*i], tri_nrm[9*i], tri_uv[6*i] */
    SRT_PRIM_ENV_SPHERE = 6,    /* env_sphere.h:7-20       data: spheres[4*i]                       */
    SRT_PRIM_NONE = -1          /* light list only: a shape without pdf_value/random (hitable.h:31-32) */
};

#define SRT_PRIM_FLIP 1u /* wrapped in an odd number of flip_normals (aarect.h:149-171) */

typedef struct SrtPrim {
    int32_t type;     /* SRT_PRIM_* */
    int32_t index;    /* row in the per-type data array */
    int32_t material; /* index into materials[] (hit_record::mat_ptr, hitable.h:24) */
    uint32_t flags;   /* SRT_PRIM_FLIP */
} SrtPrim;

/* A group is a contiguous run of prims hit with hitable_list semantics (hitable_list.h:21-33):
 * tested in order, t_max narrowed to the last accepted t.  A single primitive is a group of one;
 * `box` (box.h:18-33) is a group of six rects. */
typedef struct SrtGroup {
    int32_t first_prim;
    int32_t prim_count;
} SrtGroup;

/* A bvh is one reference `bvh_node` tree (bvh.h:8-19).  Its leaves are groups, listed in
 * bvh_items[first_item .. first_item+item_count) in the reference tree's own in-order leaf order:
 * that order fixes the reference topology (median split, bvh.h:104-114), its boxes
 * (surrounding_box, aabb.h:54-62) and the tie rule "right child wins on equal t" (bvh.h:69-76).
 * The backend builds its own SAH wide BVH over the same items and reproduces those semantics. */
typedef struct SrtBvh {
    int32_t first_item;
    int32_t item_count;
} SrtBvh;

/* ---- instance transforms (hitable.h:35-203, aarect.h:149-171) ----------------------------- */
enum {
    SRT_OP_TRANSLATE = 0, /* a,b,c = offset                    hitable.h:35-61   */
    SRT_OP_ROTATE_Y = 1,  /* a = sin_theta, b = cos_theta      hitable.h:65-132  */
    SRT_OP_ROTATE_X = 2,  /* a = sin_theta, b = cos_theta      hitable.h:135-203 */
    SRT_OP_FLIP = 3       /* flip_normals above a bvh / chain  aarect.h:149-171  */
};

typedef struct SrtXformOp {
    int32_t kind;
    float a, b, c;
} SrtXformOp;

enum { SRT_GEOM_GROUP = 0, SRT_GEOM_BVH = 1 };

/* One element of the world hitable_list (Raytracing_n.cpp:58 calls world->hit): a chain of
 * transforms, outermost first, ending in a group or a bvh; optionally the whole thing is the
 * boundary of a constant_medium (constant_medium.h:4-17). */
typedef struct SrtEntry {
    int32_t first_op;
    int32_t op_count;
    int32_t geom_kind;  /* SRT_GEOM_* */
    int32_t geom_index; /* groups[] or bvhs[] */
    int32_t medium;     /* -1, or index into media[] */
    int32_t _pad;
} SrtEntry;

typedef struct SrtMedium {
    float density;          /* constant_medium.h:15 */
    int32_t phase_material; /* the isotropic material made at constant_medium.h:8 */
} SrtMedium;

/* ---- lights: the `light_shape` hitable_list handed to color() (Raytracing_n.cpp:55, :74-84) -- */
typedef struct SrtLight {
    int32_t type;  /* SRT_PRIM_XZ_RECT / SPHERE / TRIANGLE / ENV_SPHERE, or SRT_PRIM_NONE */
    int32_t index; /* row in the per-type data array */
} SrtLight;

/* ---- materials (material.h) and textures (texture.h) ------------------------------------- */
enum {
    SRT_MAT_LAMBERTIAN = 0,    /* material.h:95-114   tex = albedo                          */
    SRT_MAT_ORENNAYAR = 1,     /* material.h:127-149  tex = albedo, p0 = A, p1 = B           */
    SRT_MAT_BECKMANN = 2,      /* material.h:151-199  tex = albedo, p0 = alphax, p1 = alphay */
    SRT_MAT_BRDF = 3,          /* material.h:201-241  p0..2 = albedo, p3 = merl table index (as float) */
    SRT_MAT_METAL = 4,         /* material.h:243-261  p0..2 = albedo, p3 = fuzz              */
    SRT_MAT_DIELECTRIC = 5,    /* material.h:282-324  p0 = ref_idx                           */
    SRT_MAT_DIFFUSE_LIGHT = 6, /* material.h:341-356  tex = emit                             */
    SRT_MAT_ISOTROPIC = 7      /* material.h:359-369  tex = albedo                           */
};

typedef struct SrtMaterial {
    int32_t type;
    int32_t texture; /* index into textures[] or -1 */
    float p[6];
} SrtMaterial;

enum {
    SRT_TEX_CONSTANT = 0, /* texture.h:25-33  c = colour            */
    SRT_TEX_CHECKER = 1,  /* texture.h:9-23   a = even, b = odd     */
    SRT_TEX_NOISE = 2,    /* texture.h:35-46  c[0] = scale          */
    SRT_TEX_IMAGE = 3     /* texture.h:48-70  a = index into images[] */
};

typedef struct SrtTexture {
    int32_t type;
    int32_t a, b;
    float c[3];
} SrtTexture;

typedef struct SrtImage {
    const uint8_t* data; /* RGB8, row-major, nx*ny*3 bytes (texture.h:66-68 assumes 3 channels) */
    int32_t nx, ny;
} SrtImage;

/* camera.h:61-67: the fields camera::get_ray reads. */
typedef struct SrtCamera {
    float origin[3];
    float lower_left_corner[3];
    float horizontal[3];
    float vertical[3];
    float u[3], v[3], w[3];
    float time0, time1;
    float lens_radius;
} SrtCamera;

/* The flattened object graph.  All arrays are caller-owned host memory, copied by
 * srt_scene_create; counts are element counts (rows), not floats. */
typedef struct SrtSceneDesc {
    int32_t abi_version; /* SRT_ABI_VERSION */

    /* world: Raytracing_n.cpp:58 */
    const SrtEntry* entries;
    int32_t n_entries;
    const SrtXformOp* ops;
    int32_t n_ops;
    const SrtGroup* groups;
    int32_t n_groups;
    const SrtBvh* bvhs;
    int32_t n_bvhs;
    const int32_t* bvh_items; /* group indices */
    int32_t n_bvh_items;
    /* [n_bvh_items][6] = min xyz, max xyz: each leaf's OWN reference box, i.e. what
     * leaf->bounding_box(0, 1, box) returns (every bvh_node in the reference is built with the
     * time range (0, 1): Raytracing_n.cpp:274-709).  The reference tree's node boxes are
     * surrounding_box unions of these (bvh.h:115-118, aabb.h:54-62); they are un-padded and
     * decide which leaves a ray can reach at all, so they cross the boundary
     * verbatim instead of being re-derived. */
    const float* bvh_item_boxes;
    const SrtMedium* media;
    int32_t n_media;

    const SrtPrim* prims; /* primitive id == index in this array */
    int32_t n_prims;

    const float* spheres; /* [n_spheres][4] */
    int32_t n_spheres;
    const float* msph; /* [n_msph][9] */
    int32_t n_msph;
    const float* rects; /* [n_rects][5] */
    int32_t n_rects;
    const float* tri_pos; /* [n_tris][9]  p0 p1 p2 */
    const float* tri_nrm; /* [n_tris][9]  n0 n1 n2 (triangle.h:101) */
    const float* tri_uv;  /* [n_tris][6]  uvs[k].x, uvs[k].y (triangle.h:104) */
    int32_t n_tris;

    /* lights: hitable_list::list_size entries are used (hitable_list.h:54-67) */
    const SrtLight* lights;
    int32_t n_lights;

    const SrtMaterial* materials;
    int32_t n_materials;
    const SrtTexture* textures;
    int32_t n_textures;
    const SrtImage* images;
    int32_t n_images;

    /* perlin.h:60-63 static tables (one set per process in the reference); may be NULL when no
     * noise_texture is present */
    const float* perlin_ranvec;  /* [256][3] */
    const int32_t* perlin_perm_x; /* [256] */
    const int32_t* perlin_perm_y;
    const int32_t* perlin_perm_z;

    /* MERL tables, brdf.h:156-188 layout: 3*90*90*180 doubles each */
    const double* const* merl_tables;
    int32_t n_merl;

    SrtCamera camera;
} SrtSceneDesc;

/* ---- render parameters: CRender(maxDepth,width,height,num_ray,thread_count,filePath)
 *      (Render.h:20) and the globals at Raytracing_n.cpp:33-43 --------------------------------- */
#define SRT_COMPAT_PIXEL_MOD_NY 1u /* i = curpixel % ny as at Raytracing_n.cpp:827 (square images only) */
#define SRT_RENDER_WAVEFRONT 2u    /* ask for the persistent sorted wavefront instead of the per-thread
                                      path loop; without the flag the backend picks it for scenes with an
                                      acceleration structure from 2^20 paths per call up.  Same image either way. */
#define SRT_RENDER_CORRECTED 4u    /* corrected-transport mode, explicitly OUTSIDE parity with the reference:
                                      cosine-weighted hemisphere sampling about the face-forward normal instead of
                                      pdf.h:10-18 / :47-56, next-event estimation with a traced shadow ray and
                                      balance-heuristic MIS instead of the `while (pdf_val == 0)` loop of
                                      Raytracing_n.cpp:75-84, the microfacet BRDF value and the MERL value
                                      (discarded at material.h:231-232) actually applied */

typedef struct SrtRenderParams {
    int32_t nx, ny;      /* Raytracing_n.cpp:39-40 */
    int32_t ns;          /* samples per pixel of the whole image, Raytracing_n.cpp:41 */
    int32_t s_begin;     /* this call renders Sobol indices [s_begin, s_end) of every pixel:      */
    int32_t s_end;       /*   the sample-range shard of one GPU                    */
    int32_t max_depth;   /* Raytracing_n.cpp:42 */
    uint32_t seed;       /* Philox key word; replaces the global drand48 stream (mathf.h:12-19)   */
    uint32_t flags;      /* SRT_COMPAT_* | SRT_RENDER_* */
} SrtRenderParams;

typedef struct SrtCounters {
    uint64_t paths;        /* color() calls from the sample loop, Raytracing_n.cpp:841 */
    uint64_t rays;         /* world->hit calls in color(), Raytracing_n.cpp:58 */
    uint64_t probe_rays;   /* hitable_pdf::value light probes, pdf.h:162-164 */
    uint64_t node_visits;  /* wide-BVH nodes fetched */
    uint64_t prim_tests;   /* primitive intersection tests */
    uint64_t reject_iters; /* iterations of the while(pdf_val==0) loop, Raytracing_n.cpp:79-83 */
    uint64_t kernel_launches; /* kernels launched by this library on this context so far */
    uint64_t culled_rays;  /* continuation rays the reference traces with an all-zero throughput
                              (they cannot change the sample); the backend does not trace them */
    uint64_t stack_overflows; /* wide-BVH traversals whose per-thread stack was full when a subtree had to be
                                 postponed, since the context was created (every kernel, not only the counted
                                 replay).  MUST be 0: a non-zero value means a closest hit may be wrong. */
} SrtCounters;

typedef struct SrtSceneInfo {
    int32_t n_prims, n_entries, n_bvhs, n_lights;
    int64_t bvh_nodes;       /* wide nodes over all bvhs */
    int64_t bvh_node_bytes;  /* bytes of node storage */
    int64_t device_bytes;    /* total device memory held by the scene */
    float sah_cost;          /* SAH cost of the largest bvh */
    float build_ms;          /* host build time */
    uint32_t features;       /* which reference classes the scene instantiates (srt_types.h F_*) */
    int32_t render_variant;  /* render-kernel variant of the last srt_render (-1: none yet) */
    int32_t render_registers;     /* registers per thread of that variant */
    int32_t render_blocks_per_sm; /* its resident blocks per SM */
    int64_t hot_table_bytes; /* entries + instance ops + primitive records + lights */
    int32_t bvh_max_depth;   /* deepest wide-BVH leaf over all bvhs (levels of inner nodes above it) */
    int32_t bvh_builder;     /* 0: host binned SAH (srt_bvh_build.cpp), 1: device LBVH (srt_lbvh.cuh) */
    float upload_ms;         /* host -> device copies of srt_scene_create */
    int32_t _pad;
} SrtSceneInfo;

typedef struct SrtContext SrtContext;
typedef struct SrtScene SrtScene;

/* Message of the last failing call on this thread ("" if none). */
const char* srt_last_error(void);

/* ABI version of the loaded library (== SRT_ABI_VERSION it was built with). */
int srt_abi_version(void);

/* One context per GPU / rank.  `cuda_stream` is a cudaStream_t (NULL = the legacy default
 * stream).  Replaces: the process-wide globals at Raytracing_n.cpp:33-45. */
int srt_context_create(int device, void* cuda_stream, SrtContext** out);
int srt_context_destroy(SrtContext* ctx);

/* Upload a flattened scene and build its acceleration structures.
 * Replaces: the scene functions' `*scene = new hitable_list(list, i)` / `new bvh_node(...)`
 * (Raytracing_n.cpp:216-711, bvh.h:96-119). */
int srt_scene_create(SrtContext* ctx, const SrtSceneDesc* desc, SrtScene** out);
int srt_scene_destroy(SrtScene* scene);
int srt_scene_info(const SrtScene* scene, SrtSceneInfo* out);

/* Closest hit of n rays against the world with the reference's semantics:
 * world->hit(r, 0.001, FLT_MAX, rec) (Raytracing_n.cpp:58; hitable_list.h:21-33, bvh.h:64-93).
 * rays: [n][7] = origin, direction, time (ray.h:6-19).  prim_id[i] = index into desc.prims or -1;
 * t[i] = hit_record::t (undefined on miss).  rec (optional, may be NULL): [n][8] = u v p normal
 * (hitable.h:17-25).  All pointers may be host or device memory (detected per pointer). */
int srt_trace_closest(SrtScene* scene, const float* rays, int64_t n, int32_t* prim_id, float* t,
                      float* rec);

/* The camera rays of the sample loop: for every pixel (in `curpixel` order) and every Sobol index s in
 * [s_begin, s_end): u = (sobol[s][0] + i) / nx, v = (sobol[s][1] + j) / ny, camera::get_ray(u, v)
 * (Raytracing_n.cpp:827-836, camera.h:51-59), generated by the SAME device routine the render kernels start
 * their paths with (same Philox counters: a ray here is the ray srt_render traces for that pixel and sample).
 * rays_out: [nx*ny*(s_end-s_begin)][7] = origin, direction, time; host or device pointer. */
int srt_primary_rays(SrtScene* scene, const SrtRenderParams* params, float* rays_out);

/* mixture-density probe used by the light-sampling branch: for each (origin, direction) pair
 * returns hitable_list::pdf_value over the light list (hitable_list.h:54-62, pdf.h:162-164).
 * od: [n][6]; out: [n]; host or device pointers. */
int srt_light_pdf(SrtScene* scene, const float* od, int64_t n, float* out);

/* Render Sobol sample indices [s_begin, s_end) of every pixel and ADD the per-pixel radiance sums
 * (de_nan applied per sample, Raytracing_n.cpp:47-53, :841) into accum_rgb, a DEVICE buffer of
 * nx*ny*3 floats indexed by the reference's `curpixel` (Raytracing_n.cpp:827-828, :863).
 * Replaces: renderthread's sample loop (Raytracing_n.cpp:815-848) and color() (:55-106).
 * Asynchronous on the context's stream. */
int srt_render(SrtScene* scene, const SrtRenderParams* params, float* accum_rgb_device);

/* mean -> sqrt -> int(255.99 c) -> clamp (Raytracing_n.cpp:848-862) of an accumulation buffer
 * holding n_samples per pixel.  accum is a device pointer; rgb8_out (n_pixels*3 bytes) may be a
 * host or a device pointer.  Synchronises the stream when rgb8_out is host memory. */
int srt_resolve(SrtContext* ctx, const float* accum_rgb_device, int64_t n_pixels, int32_t n_samples,
                uint8_t* rgb8_out);

/* The sharded form of srt_render + srt_resolve.  srt_render_fixed ADDS the radiance
 * sums of Sobol indices [s_begin, s_end) into fixed_device, a DEVICE buffer of nx*ny*3 int64 holding sums in
 * fixed point with 31 fractional bits (the representation the render kernels accumulate in: integer addition
 * is associative, so a pixel's sum does not depend on the order in which paths finish — nor on how the
 * sample range was split over GPUs: buffers of different shards add with an exact integer reduce, e.g.
 * ncclReduce(ncclInt64, ncclSum), and the image is bit-identical for any number of GPUs).
 * srt_resolve_fixed: mean -> sqrt -> int(255.99 c) -> clamp (Raytracing_n.cpp:848-862) of such a buffer
 * holding n_samples per pixel; rgb8_out: host or device pointer, or NULL; accum_out_device (optional):
 * the float sums srt_render would have produced; clear != 0 zeroes the buffer for the next frame. */
int srt_render_fixed(SrtScene* scene, const SrtRenderParams* params, int64_t* fixed_device);
int srt_resolve_fixed(SrtContext* ctx, int64_t* fixed_device, int64_t n_pixels, int32_t n_samples, uint8_t* rgb8_out,
                      float* accum_out_device, int clear);

/* The exchange step of the multi-GPU path as ONE kernel over NVLink peer memory, replacing the
 * ncclReduce + srt_resolve_fixed pair above for the ranks of one box.  Every rank creates its side (one device block:
 * two int64 accumulation buffers that alternate by frame, the 8-bit image, flags), publishes a 64-byte IPC handle,
 * and connects to the handles of all ranks (in rank order, its own included).  Per frame: srt_render_fixed into
 * srt_exchange_fixed(), then srt_exchange_run(): rank g waits INSIDE the kernel until every rank has finished
 * rendering, sums pixel tile g over the buffers of all ranks with peer loads, resolves it
 * (Raytracing_n.cpp:848-862) and stores the 8-bit tile into the image of every rank; when the call's work has run on
 * the stream the whole image is in srt_exchange_image() on every rank (copied to rgb8_host when that is non-NULL,
 * which synchronises).  The accumulation buffer of the NEXT frame is zero on return.  Every rank must make the same
 * sequence of calls; a rank that does not arrive within a few seconds makes the others return SRT_E_CUDA instead of
 * hanging.  n_ranks <= 16. */
typedef struct SrtExchange SrtExchange;
#define SRT_EXCHANGE_HANDLE_BYTES 64
int srt_exchange_create(SrtContext* ctx, int64_t n_pixels, int32_t n_ranks, int32_t rank, SrtExchange** out);
int srt_exchange_handle(SrtExchange* ex, uint8_t* handle_out);
int srt_exchange_connect(SrtExchange* ex, const uint8_t* handles_of_all_ranks);
int64_t* srt_exchange_fixed(SrtExchange* ex);
uint8_t* srt_exchange_image(SrtExchange* ex);
int srt_exchange_run(SrtExchange* ex, int32_t n_samples, uint8_t* rgb8_host);
int srt_exchange_destroy(SrtExchange* ex);

/* Whole-image convenience with HOST buffers: zero an internal accumulation buffer, srt_render
 * over [s_begin, s_end), srt_resolve, copy back.  rgb8_host: nx*ny*3 bytes; accum_host (optional,
 * may be NULL): nx*ny*3 floats of linear radiance sums.
 * Replaces: the body of CRender::Run (Render.h:23; Raytracing_n.cpp:926-946). */
int srt_render_image(SrtScene* scene, const SrtRenderParams* params, uint8_t* rgb8_host,
                     float* accum_host);

/* The allocation-free form of srt_render_image for small frames, device output.
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Renders [s_begin, s_end) of an nx x ny frame with the fused kernel (k_render) and writes the resolved 8-bit
 * image to rgb8_device (nx*ny*3 bytes of DEVICE memory, the caller's).  The kernel's scratch — the fixed-point
 * radiance sums and the Sobol table — is a file-scope device array of the library (6 MiB + 8 KiB, held by
 * every GPU that loads libsrt_b200.so until the process exits), so the call makes no cudaMalloc / cudaFree and
 * calls no allocating library routine.  Limits: nx*ny <= SRT_STATIC_MAX_PIXELS and s_end <=
 * SRT_STATIC_MAX_SAMPLES, else SRT_E_INVALID with a message naming the limit.  Same bytes as srt_render_image
 * for the same scene and params (SRT_RENDER_WAVEFRONT is ignored: the wavefront scheduler needs its slot pool).
 * The FIRST call on a device also looks the array up and uploads the Sobol table (synchronous, once per
 * device): make it outside graph capture.  Every call on a device shares the one array (it is cleared at the
 * start of each call), so calls that overlap on the GPU — two streams, two contexts, two threads — race:
 * serialise them.  Work is issued on the context's stream and not waited for. */
#define SRT_STATIC_MAX_PIXELS (512 * 512)
#define SRT_STATIC_MAX_SAMPLES 1024
int srt_render_image_static(SrtScene* scene, const SrtRenderParams* params, uint8_t* rgb8_device);

/* Counters of the last instrumented render (params.flags has no effect on them; they are
 * gathered by srt_render_counted, an untimed replay). */
int srt_render_counted(SrtScene* scene, const SrtRenderParams* params, float* accum_rgb_device,
                       SrtCounters* out);
int srt_counters_get(const SrtContext* ctx, SrtCounters* out);

/* brdf::lookup_brdf_val (brdf.h:190-214) on MERL table `table` of the scene for n angle
 * quadruples.  angles: [n][4] = theta_in, fi_in, theta_out, fi_out (doubles); rgb: [n][3].
 * Host or device pointers.  (brdfmaterial::scatter evaluates this and discards the result,
 * material.h:231-232; the entry point exists so the index arithmetic can be checked.) */
int srt_merl_lookup(SrtScene* scene, int32_t table, const double* angles, int64_t n, double* rgb);

/* Block the calling host thread until the context's stream is idle. */
int srt_context_synchronize(SrtContext* ctx);

/* Self-test of the backend's exact division: counts the operand pairs (a[i], b[i]) for which the
 * hoisted-reciprocal quotient the kernels use differs from IEEE round-to-nearest a[i] / b[i]
 * (must be 0).  a, b: device or host pointers to n floats. */
int srt_selftest_division(SrtContext* ctx, const float* a, const float* b, int64_t n, uint64_t* mismatches);

/* Measurement aid (bench.py): FP32 throughput of this GPU in TFLOP/s from a dependent-FMA-chain kernel
 * (16 chains per thread, 8 CTAs x 256 threads per SM, best of 6 launches) — the measured denominator of
 * the FP32 roofline fraction, since MEASURED_PEAKS.json carries HBM and tensor peaks only. */
int srt_measure_fp32_peak(SrtContext* ctx, double* tflops);

/* Measurement aid (bench.py): L2-resident streaming read bandwidth of this GPU in GB/s — every SM reads a
 * 32 MiB buffer (a quarter of the 126 MB L2) over and over with 128-bit loads, best of 6 launches after a
 * warm-up pass that brings the buffer into L2.  The measured denominator of the L2 roofline fraction. */
int srt_measure_l2_peak(SrtContext* ctx, double* gbps);

/* Host-side helpers (no device work): the reference's Sobol table (Raytracing_n.cpp:721-812,
 * D = 2; direction numbers of contents/sobol/new-joe-kuo-6.21201 line 2) as u32 numerators of
 * x / 2^32.  out: [n][2]. */
int srt_sobol_points_u32(uint32_t n, uint32_t* out);

/* Host-side helper (no device work) for the flatteners: the in-order leaf sequence of the reference's
 * `new bvh_node(l, n, t0, t1)` (bvh.h:96-119) over n items whose bounding-box minima are mins[n][3] — random
 * axis from the global drand48 stream (mathf.h:12-19; *seed is its state on entry and on return), qsort with
 * the never-equal comparators (bvh.h:21-55), median split.  order_out[k] = input index of the k-th leaf: the
 * order in which SrtSceneDesc.bvh_items must list them.  Multi-threaded (10 M items: a few seconds). */
int srt_ref_bvh_leaf_order(const float* mins, int64_t n, uint64_t* seed, int32_t* order_out);

/* Diagnostic: the bounding boxes of the reference's bvh_node tree (bvh.h:96-119) over n leaves given in their
 * in-order sequence, as srt_scene_create keeps them for the reachability replay.  item_boxes: [n][6] (min, max);
 * nodes_out: 12 x 32-bit words per node in depth-first order — lo[3], left, hi[3], right, mid, 3 x 0 (left / right:
 * node index or -1; mid: first rank of the right subtree) — room for 2 n nodes (srt_api.cu RefNode).  on_device != 0
 * runs the per-level kernels srt_scene_create uses for large trees, 0 the host recursion it uses for small ones:
 * the two must agree bit for bit (tests/test_gpu_features.py).  *count_out = nodes written. */
int srt_ref_tree_boxes(SrtContext* ctx, const float* item_boxes, int32_t n, int on_device, void* nodes_out,
                       int32_t* count_out);

#ifdef __cplusplus
}
#endif
#endif /* SRT_FFI_H */
