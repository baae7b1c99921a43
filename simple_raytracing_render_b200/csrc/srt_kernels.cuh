// srt_kernels.cuh — the render kernel (one instantiation per variant, srt_render_variant.cu).
#ifndef SRT_KERNELS_CUH
#define SRT_KERNELS_CUH

#include "srt_shade.cuh"

namespace srt {

struct RenderArgs {
    int nx, ny;
    int s_begin, s_end;
    int max_depth;
    uint32_t seed;
    int mod_ny;           // SRT_COMPAT_PIXEL_MOD_NY
    int chunk;            // Sobol indices per work item
    int n_chunks;         // work items per pixel
    unsigned long long n_items;  // n_pixels * n_chunks
    float fx_clamp;       // largest |item sum| the fixed-point accumulator takes (4e9 / samples per pixel)
    int regen_threshold;  // idle lanes in a warp that trigger a regeneration phase
    int grab;             // work items a warp takes from the global counter at a time
    int dbg_pixel, dbg_sample;  // counting build: print the bounces of this one path (-1: off)
    int cta_sync;         // 1: the warps of a CTA start every bounce together; 2: and every shade together
};

constexpr int kRenderThreads = 256;

// item number -> (pixel, chunk, first and one-past-last Sobol index): pixel-major, so that the lanes
// which start paths together (they take consecutive items) start them in the same pixel
SRT_DEV void decode_item(const RenderArgs& A, unsigned long long item, int& pixel, int& chunk_id, int& s_next, int& s_stop) {
    pixel = (int)(item / (unsigned)A.n_chunks);
    chunk_id = (int)(item - (unsigned long long)pixel * (unsigned)A.n_chunks);
    s_next = A.s_begin + chunk_id * A.chunk;
    s_stop = min(s_next + A.chunk, A.s_end);
}

// The image accumulates in 64-bit fixed point (31 fractional bits) with integer atomics: integer
// addition is associative, so the sum of a pixel does not depend on the order in which its items
// finish — the frame is bit-reproducible for ANY schedule, with 24 bytes per pixel instead of one
// float3 slot per work item, which is what lets a work item be a single path.  x * 2^31 is exact in
// float and its conversion is exact; contributions are clamped so that the sum cannot overflow.
SRT_DEV void item_commit(long long* __restrict__ fx, int pixel, V3 sum, float clampv) {
    unsigned long long* o = reinterpret_cast<unsigned long long*>(fx + 3 * (size_t)pixel);
    float x = fminf(fmaxf(sum.x, -clampv), clampv), y = fminf(fmaxf(sum.y, -clampv), clampv),
          z = fminf(fmaxf(sum.z, -clampv), clampv);
    atomicAdd(o + 0, (unsigned long long)__float2ll_rn(x * 2147483648.0f));
    atomicAdd(o + 1, (unsigned long long)__float2ll_rn(y * 2147483648.0f));
    atomicAdd(o + 2, (unsigned long long)__float2ll_rn(z * 2147483648.0f));
}

SRT_DEV View global_view(const DScene& S) { return View{S, S.entries, S.ops, S.dprims, S.lights}; }

// The start of one path: pixel mapping, Sobol jitter and camera::get_ray as the sample loop does them
// (Raytracing_n.cpp:827-836: `float u = float(samplepoints[s][0] + i) / float(nx)`), with the path's Philox
// stream initialised on (pixel, global Sobol index).  Every scheduler and srt_primary_rays call this one routine.
SRT_DEV Ray start_path(const SrtCamera& cam, const RenderArgs& A, const uint32_t* __restrict__ sobol, int pixel, int s,
                       Philox& rng) {
    const int i = pixel % (A.mod_ny ? A.ny : A.nx);
    const int j = A.ny - 1 - (pixel / A.nx);
    const double sx = (double)__ldg(&sobol[2 * s]) * 2.3283064365386963e-10;
    const double sy = (double)__ldg(&sobol[2 * s + 1]) * 2.3283064365386963e-10;
    const float u = xdiv((float)(sx + (double)i), (float)A.nx);
    const float v = xdiv((float)(sy + (double)j), (float)A.ny);
    philox_init(rng, A.seed, (uint32_t)pixel, (uint32_t)s);
    return camera_get_ray(cam, u, v, rng);
}

// The sample loop of renderthread (Raytracing_n.cpp:826-848) and color() (:55-106) for every pixel,
// as a persistent grid of autonomous warps:
//
//   * a WORK ITEM is `chunk` consecutive Sobol indices of one pixel; items are numbered pixel-major
//     (item = pixel * n_chunks + c) and handed out in increasing order, `grab` at a time per warp
//     from one global counter, so the lanes of a warp work on neighbouring pixels;
//   * every lane owns one item and one path at a time.  A warp iteration traces ONE ray and shades
//     ONE bounce for all its live paths; lanes whose path ended wait until `regen_threshold` lanes are
//     idle, then all of them start their next path together (or fetch their next item with one
//     ballot / popc rank — no atomics inside the warp);
//   * a finished item writes its radiance sum to its own slot — out[(c * n_pixels + pixel) * 3] —
//     so the image is bit-reproducible: no float atomics, one writer per address;
//     k_sum_partials then adds the n_chunks slots of a pixel in fixed order into the caller's
//     accumulation buffer (with n_chunks == 1 the kernel adds into it directly).
//
// The hot tables of the scene (entries, instance ops, primitive records, lights) are staged into
// shared memory when the variant has F_SMEM_TABLES: every lane of a warp reads the same record at
// the same time, which is a broadcast from shared memory.
// resident CTAs per SM the register allocation is bounded for (srt_variants.h picks it per variant)
#ifndef SRT_MINBLOCKS
#define SRT_MINBLOCKS 2
#endif

template <uint32_t F, bool COUNT>
__global__ void __launch_bounds__(kRenderThreads, SRT_MINBLOCKS)
    k_render(DScene S, RenderArgs A, const uint32_t* __restrict__ sobol, long long* __restrict__ out,
             unsigned long long* __restrict__ counters) {
    extern __shared__ float4 smem_hot[];
    const DEntry* entries = S.entries;
    const SrtXformOp* ops = S.ops;
    const DPrim* dprims = S.dprims;
    const DPrim* lights = S.lights;
    if (F & F_SMEM_TABLES) {
        const float4* src = reinterpret_cast<const float4*>(S.entries);
        const unsigned n16 = S.hot_bytes >> 4;
        for (unsigned i = threadIdx.x; i < n16; i += blockDim.x) smem_hot[i] = __ldg(src + i);
        __syncthreads();
        const char* base = reinterpret_cast<const char*>(smem_hot);
        const char* g0 = reinterpret_cast<const char*>(S.entries);
        entries = reinterpret_cast<const DEntry*>(base);
        ops = reinterpret_cast<const SrtXformOp*>(base + (reinterpret_cast<const char*>(S.ops) - g0));
        dprims = reinterpret_cast<const DPrim*>(base + (reinterpret_cast<const char*>(S.dprims) - g0));
        lights = reinterpret_cast<const DPrim*>(base + (reinterpret_cast<const char*>(S.lights) - g0));
    }
    View W{S, entries, ops, dprims, lights};

    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int n_pixels = A.nx * A.ny;
    unsigned long long* work_counter = counters + 7;

    Counters cnt = {0, 0, 0, 0, 0, 0, 0, 0};
    Tally<COUNT> tl{&cnt};
    unsigned long long paths = 0;

    // per-lane item and path state
    bool has_item = false, alive = false;
    int pixel = 0, chunk_id = 0, s_next = 0, s_stop = 0;
    V3 sum = v3(0.f, 0.f, 0.f);
    Path P;
    P.depth = 0;
    P.emit_w = 1.f;
    Philox rng;
    // corrected-transport mode (F_CORRECTED): the shadow ray of the last vertex, traced by the NEXT trip of the
    // bounce loop through the same world_hit call as every other ray (one copy of the closest-hit code)
    Nee nee;
    nee.valid = false;
    bool main_dead = false;
    // warp-uniform range of items
    unsigned long long w_cur = 0, w_end = 0;
    bool exhausted = false;

    // cta_sync: the kernel is several times the instruction cache, and warps at unrelated program
    // counters each stream it through on their own (profiles/r01_notes.md, capture A: 61 % of the
    // stall samples were `no_instruction`).  Starting every bounce (and every shade) of a CTA's warps
    // together makes one instruction-cache miss serve all of them.
    bool warp_done = false;
    for (;;) {
        if (A.cta_sync && !__syncthreads_or(!warp_done)) break;
        bool run = !warp_done;
        unsigned dead = __ballot_sync(FULL, !alive);
        if (run && (dead == FULL || __popc(dead) >= A.regen_threshold)) {
            // ---- regeneration phase ---------------------------------------------------------------
            bool item_done = !alive && has_item && s_next >= s_stop;
            if (item_done) {
                item_commit(out, pixel, sum, A.fx_clamp);
                has_item = false;
            }
            bool need_item = !alive && !has_item;
            unsigned m = __ballot_sync(FULL, need_item);
            if (m) {
                if (w_cur == w_end && !exhausted) {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(work_counter, (unsigned long long)A.grab);
                    base = __shfl_sync(FULL, base, 0);
                    if (base >= A.n_items) {
                        exhausted = true;
                    } else {
                        w_cur = base;
                        w_end = min(base + (unsigned long long)A.grab, A.n_items);
                    }
                }
                unsigned avail = (unsigned)(w_end - w_cur);
                unsigned rank = __popc(m & lt_mask);
                if (need_item && rank < avail) {
                    decode_item(A, w_cur + rank, pixel, chunk_id, s_next, s_stop);
                    sum = v3(0.f, 0.f, 0.f);
                    has_item = true;
                }
                w_cur += min((unsigned)__popc(m), avail);
            }
            if (!alive && has_item && s_next < s_stop) {
                int s = s_next++;
                P.r = start_path(S.camera, A, sobol, pixel, s, rng);
                P.thr = v3(1.f, 1.f, 1.f);
                P.L = v3(0.f, 0.f, 0.f);
                P.depth = 0;
                P.emit_w = 1.f;
                alive = true;
                if (COUNT) {
                    paths++;
                    cnt.dbg = (pixel == A.dbg_pixel && s == A.dbg_sample) ? 1 : 0;
                    if (cnt.dbg)
                        printf("[dbg] pixel %d sample %d: o=(%g %g %g) d=(%g %g %g)\n", pixel, s, P.r.o.x, P.r.o.y, P.r.o.z,
                               P.r.d.x, P.r.d.y, P.r.d.z);
                }
            }
            if (!__any_sync(FULL, alive)) {
                bool pending = __any_sync(FULL, has_item);
                if (!pending && exhausted && w_cur == w_end) {
                    if (!A.cta_sync) break;
                    warp_done = true;
                }
                run = false;
            }
        }
        // ---- one ray + one bounce for every live path ---------------------------------------------
        Hit h;
        const bool shadow = (F & F_CORRECTED) && nee.valid;  // this trip traces the pending shadow ray instead
        const Ray tr = shadow ? Ray{nee.o, nee.dir, nee.time} : P.r;
        if (run && alive) {
            if (COUNT) cnt.rays++;
            h = world_hit<F>(W, tr, 0.001f, kFltMax, rng, tl);
        }
        if (A.cta_sync >= 2) __syncthreads();
        if (run && alive) {
            if (COUNT && cnt.dbg)
                printf("[dbg] depth %d hit prim %d entry %d t %g thr=(%g %g %g) L=(%g %g %g)\n", P.depth, h.prim, h.entry, h.t,
                       P.thr.x, P.thr.y, P.thr.z, P.L.x, P.L.y, P.L.z);
            bool ended;
            if (F & F_CORRECTED) {
                if (shadow) {  // next-event estimation: the emitter's radiance if an emitter is what the ray reaches
                    V3 e = emitted_at<F>(W, tr, h);
                    P.L = v3(fmaf(nee.weight.x, e.x, P.L.x), fmaf(nee.weight.y, e.y, P.L.y), fmaf(nee.weight.z, e.z, P.L.z));
                    nee.valid = false;
                    ended = main_dead;
                } else {
                    main_dead = !shade_bounce_corrected<F>(W, P, h, A.max_depth, rng, nee, tl);
                    ended = main_dead && !nee.valid;
                }
            } else {
                ended = !shade_bounce<F>(W, P, h, A.max_depth, rng, tl);
            }
            if (ended) {
                V3 c = de_nan(P.L);  // col += de_nan(color(...)), Raytracing_n.cpp:841
                sum = v3(sum.x + c.x, sum.y + c.y, sum.z + c.z);
                alive = false;
                main_dead = false;
            }
        }
    }
    if (COUNT) {
        atomicAdd(&counters[0], paths);
        atomicAdd(&counters[1], cnt.rays);
        atomicAdd(&counters[2], cnt.probe_rays);
        atomicAdd(&counters[3], cnt.node_visits);
        atomicAdd(&counters[4], cnt.prim_tests);
        atomicAdd(&counters[5], cnt.reject_iters);
        atomicAdd(&counters[6], cnt.culled_rays);
    }
}

}  // namespace srt
#endif
