// srt_wavefront.cuh — the render loop as a two-stage wavefront over a fixed pool of path slots.
//
// Why (profiles/r01_notes.md, captures A-D): the fused kernel of srt_kernels.cuh is bound by
// instruction fetch — its hot code (37 KB for the sphere-BVH scene, 95 KB for the Beckmann mesh) is
// larger than the SM's instruction cache, 60 % of its stall samples are `no_instruction`, while the
// closest-hit code run on its own (k_trace_closest) sustains 10-27 Grays/s, five times the fused rate.
// Splitting the bounce into a TRACE kernel (world->hit only) and a SHADE kernel (color()'s material /
// pdf / light-sampling step, path regeneration and ray generation) gives each stage a hot set that
// fits the instruction cache, and gives the trace stage full warps: a slot whose path ended is
// refilled in the same shade pass, so every lane of every warp traces a live ray.
//
// State lives in global memory as 16-byte SoA records, 144 B per slot; the pool is sized to the
// resident threads of the GPU (a few 100 k slots, tens of MB), so it stays in the 126 MB L2 between
// the passes and HBM sees none of it.  One iteration = one k_wf_trace + one k_wf_shade launch; the
// host enqueues iterations in batches and polls the live-slot count of an earlier pass, never
// draining the stream (srt_api.cu: launch_wavefront).
//
// Results are the ones of the fused kernel: same work items, same Philox counters (the buffered
// outputs of a path's generator are part of the slot), same per-item partial sums.
#ifndef SRT_WAVEFRONT_CUH
#define SRT_WAVEFRONT_CUH

#include "srt_kernels.cuh"

// resident CTAs per SM the two kernels are register-bounded for (srt_variants.h picks per variant)
#ifndef SRT_WF_TRACE_MINBLOCKS
#define SRT_WF_TRACE_MINBLOCKS 3
#endif
#ifndef SRT_WF_SHADE_MINBLOCKS
#define SRT_WF_SHADE_MINBLOCKS 3
#endif

namespace srt {

struct WfState {
    float4* ray_o;  // o.xyz, time
    float4* ray_d;  // d.xyz, -
    float4* hit;    // t, prim (int bits), entry (int bits), -
    float4* thr;    // throughput.xyz, depth (int bits)
    float4* rad;    // radiance gathered so far .xyz, -
    float4* sum;    // radiance sum of the current work item .xyz, flags (int bits: 1 alive, 2 has_item)
    int4* item;     // pixel, chunk_id, s_next, s_stop
    uint4* rng;     // (c2 << 3) | have, out[0], out[1], out[2]
    int n_slots;    // multiple of 32
};

constexpr int kWfThreads = 256;
constexpr int kWfAlive = 1, kWfHasItem = 2;

SRT_DEV void wf_load_rng(const WfState& st, int slot, uint32_t seed, uint32_t pixel, uint32_t sample, Philox& p) {
    uint4 v = st.rng[slot];
    philox_init(p, seed, pixel, sample);
    p.c2 = v.x >> 3;
    p.have = (int)(v.x & 7u);
    p.out[0] = v.y;
    p.out[1] = v.z;
    p.out[2] = v.w;
    p.out[3] = 0;
}
SRT_DEV void wf_store_rng(const WfState& st, int slot, const Philox& p) {
    st.rng[slot] = make_uint4((p.c2 << 3) | (uint32_t)p.have, p.out[0], p.out[1], p.out[2]);
}

// Stage the hot tables of the scene into shared memory (F_SMEM_TABLES) and return the view.
template <uint32_t F>
SRT_DEV View wf_view(const DScene& S, float4* smem_hot) {
    if (F & F_SMEM_TABLES) {
        const float4* src = reinterpret_cast<const float4*>(S.entries);
        const unsigned n16 = S.hot_bytes >> 4;
        for (unsigned i = threadIdx.x; i < n16; i += blockDim.x) smem_hot[i] = __ldg(src + i);
        __syncthreads();
        const char* base = reinterpret_cast<const char*>(smem_hot);
        const char* g0 = reinterpret_cast<const char*>(S.entries);
        return View{S, reinterpret_cast<const DEntry*>(base),
                    reinterpret_cast<const SrtXformOp*>(base + (reinterpret_cast<const char*>(S.ops) - g0)),
                    reinterpret_cast<const DPrim*>(base + (reinterpret_cast<const char*>(S.dprims) - g0)),
                    reinterpret_cast<const DPrim*>(base + (reinterpret_cast<const char*>(S.lights) - g0))};
    }
    return global_view(S);
}

// ---- TRACE: world->hit(r, 0.001, FLT_MAX, rec) (Raytracing_n.cpp:58) for every live slot -----------
// `live[pass]` is the number of live slots the shade pass `pass` left behind; a trace / shade pass
// whose predecessor left none (and is not the first) has nothing to do and returns at once: the host
// may enqueue a few passes more than needed (it polls with a lag), they cost a launch each.
template <uint32_t F>
SRT_DEV void wf_trace_slot(const View& W, const RenderArgs& A, const WfState& st, int slot) {
    Tally<false> tl{nullptr};
    const int flags = __float_as_int(st.sum[slot].w);
    if (!(flags & kWfAlive)) return;
    float4 o = st.ray_o[slot], d = st.ray_d[slot];
    Ray r{v3(o.x, o.y, o.z), v3(d.x, d.y, d.z), o.w};
    Philox rng;
    rng.have = 0;
    if (F & F_MEDIA) {  // constant_medium::hit draws its free-flight distance (constant_medium.h:33)
        int4 it = st.item[slot];
        wf_load_rng(st, slot, A.seed, (uint32_t)it.x, (uint32_t)(it.z - 1), rng);
    }
    Hit h = world_hit<F>(W, r, 0.001f, kFltMax, rng, tl);
    st.hit[slot] = make_float4(h.t, __int_as_float(h.prim), __int_as_float(h.entry), 0.f);
    if (F & F_MEDIA) wf_store_rng(st, slot, rng);
}

template <uint32_t F>
__global__ void __launch_bounds__(kWfThreads, SRT_WF_TRACE_MINBLOCKS)
    k_wf_trace(DScene S, RenderArgs A, WfState st, const int* __restrict__ live, int pass) {
    extern __shared__ float4 smem_hot[];
    if (pass > 0 && live[pass - 1] == 0) return;
    View W = wf_view<F>(S, smem_hot);
    const int stride = gridDim.x * blockDim.x;
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < st.n_slots; slot += stride)
        wf_trace_slot<F>(W, A, st, slot);
}

// ---- SHADE: one level of color() for every live slot, then refill ----------------------------------
// For a live slot: shade_bounce on the hit the trace pass stored (emitted / scatter / mixture sampling,
// Raytracing_n.cpp:61-99); a path that ends adds de_nan(radiance) to its work item's sum (:841).
// For a slot without a live path: close the finished work item (write its sum to the item's own
// partial slot, as the fused kernel does), take the next item from the global counter (one atomic per
// warp), and start the next path: Sobol jitter + camera::get_ray (:827-836).
// counters[7] is the work-item counter; live[pass] receives the number of slots left alive.
// One slot of the shade pass; warp-collective (the refill takes work items with one ballot and one
// atomic per warp): every lane of the warp must call it, each with its own slot.  Returns whether
// the slot holds a live path afterwards.
template <uint32_t F>
SRT_DEV bool wf_shade_slot(const DScene& S, const View& W, const RenderArgs& A, const WfState& st,
                           const uint32_t* __restrict__ sobol, long long* __restrict__ out,
                           unsigned long long* __restrict__ work_counter, int slot, bool first,
                           volatile int* exhausted = nullptr) {
    Tally<false> tl{nullptr};
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int n_pixels = A.nx * A.ny;
    float4 sm = st.sum[slot];
    int flags = first ? 0 : __float_as_int(sm.w);
    V3 sum = first ? v3(0.f, 0.f, 0.f) : v3(sm.x, sm.y, sm.z);
    int4 it = first ? make_int4(0, 0, 0, 0) : st.item[slot];
    int pixel = it.x, chunk_id = it.y, s_next = it.z, s_stop = it.w;
    bool alive = (flags & kWfAlive) != 0, has_item = (flags & kWfHasItem) != 0;
    Philox rng;
    rng.have = 0;
    Path P;
    if (alive) {
        float4 o = st.ray_o[slot], d = st.ray_d[slot], hv = st.hit[slot], tv = st.thr[slot], lv = st.rad[slot];
        P.r = Ray{v3(o.x, o.y, o.z), v3(d.x, d.y, d.z), o.w};
        P.thr = v3(tv.x, tv.y, tv.z);
        P.depth = __float_as_int(tv.w);
        P.L = v3(lv.x, lv.y, lv.z);
        Hit h;
        h.t = hv.x;
        h.prim = __float_as_int(hv.y);
        h.entry = __float_as_int(hv.z);
        wf_load_rng(st, slot, A.seed, (uint32_t)pixel, (uint32_t)(s_next - 1), rng);
        if (!shade_bounce<F>(W, P, h, A.max_depth, rng, tl)) {
            V3 c = de_nan(P.L);
            sum = v3(sum.x + c.x, sum.y + c.y, sum.z + c.z);
            alive = false;
        }
    }
    // ---- refill (the regeneration phase of srt_kernels.cuh, per lane and without a threshold) ----
    if (!alive && has_item && s_next >= s_stop) {
        item_commit(out, pixel, sum, A.fx_clamp);
        has_item = false;
    }
    const bool need_item = !alive && !has_item;
    const unsigned m = __ballot_sync(FULL, need_item);
    // `exhausted` (shared memory, persistent kernel): once the counter has run past the last item
    // the draining warps stop hammering it
    if (m && !(exhausted && *exhausted)) {
        unsigned long long base = 0;
        if (lane == 0) {
            base = atomicAdd(work_counter, (unsigned long long)__popc(m));
            if (exhausted && base >= A.n_items) *exhausted = 1;
        }
        base = __shfl_sync(FULL, base, 0);
        const unsigned long long item = base + (unsigned)__popc(m & lt_mask);
        if (need_item && item < A.n_items) {
            decode_item(A, item, pixel, chunk_id, s_next, s_stop);
            sum = v3(0.f, 0.f, 0.f);
            has_item = true;
        }
    }
    if (!alive && has_item && s_next < s_stop) {
        int s = s_next++;
        P.r = start_path(S.camera, A, sobol, pixel, s, rng);
        P.thr = v3(1.f, 1.f, 1.f);
        P.L = v3(0.f, 0.f, 0.f);
        P.depth = 0;
        alive = true;
    }
    // ---- store ------------------------------------------------------------------------------------
    flags = (alive ? kWfAlive : 0) | (has_item ? kWfHasItem : 0);
    st.sum[slot] = make_float4(sum.x, sum.y, sum.z, __int_as_float(flags));
    st.item[slot] = make_int4(pixel, chunk_id, s_next, s_stop);
    if (alive) {
        st.ray_o[slot] = make_float4(P.r.o.x, P.r.o.y, P.r.o.z, P.r.time);
        st.ray_d[slot] = make_float4(P.r.d.x, P.r.d.y, P.r.d.z, 0.f);
        st.thr[slot] = make_float4(P.thr.x, P.thr.y, P.thr.z, __int_as_float(P.depth));
        st.rad[slot] = make_float4(P.L.x, P.L.y, P.L.z, 0.f);
        wf_store_rng(st, slot, rng);
        return true;
    }
    return false;
}

template <uint32_t F>
__global__ void __launch_bounds__(kWfThreads, SRT_WF_SHADE_MINBLOCKS)
    k_wf_shade(DScene S, RenderArgs A, WfState st, const uint32_t* __restrict__ sobol, long long* __restrict__ out,
               unsigned long long* __restrict__ counters, int* __restrict__ live, int pass) {
    extern __shared__ float4 smem_hot[];
    if (pass > 0 && live[pass - 1] == 0) return;
    View W = wf_view<F>(S, smem_hot);
    const int stride = gridDim.x * blockDim.x;
    int n_alive = 0;
    // n_slots is a multiple of 32 and blockDim of 32: the warps of this loop are always full
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < st.n_slots; slot += stride)
        n_alive += wf_shade_slot<F>(S, W, A, st, sobol, out, counters + 7, slot, pass == 0) ? 1 : 0;
    const unsigned total = __reduce_add_sync(0xffffffffu, (unsigned)n_alive);
    if ((threadIdx.x & 31) == 0 && total) atomicAdd(live + pass, (int)total);
}


// ---- PERSISTENT form: the same two stages as phases of ONE kernel ------------------------------------
// Every CTA owns a private, contiguous range of the slot pool (slot = base + k * blockDim + thread,
// k < per_thread: a slot is only ever touched by one thread, coalesced across the warp) and
// alternates   shade phase -> __syncthreads -> trace phase -> __syncthreads   until a shade phase
// leaves none of its slots alive.  The barriers are not needed for correctness; they keep every warp
// of the CTA in the same phase, so the SM's instruction cache holds one stage's code at a time
// (what the two-kernel form buys), without a launch, a table re-staging and a ramp-up/tail per
// stage and bounce: profiles/r02_notes.md measured ~14 us of those per launch, 1500 launches per
// C2 frame.  CTAs take work items from the same global counter, so load balance is dynamic and a
// CTA's tail is bounded by chunk x depth iterations over its own slots only.
#ifndef SRT_WFP_THREADS
#define SRT_WFP_THREADS 512
#endif
#ifndef SRT_WFP_LANE_SCHED
#define SRT_WFP_LANE_SCHED 0  // 1: wf_shade_phase (rejection loop scheduled per lane); measured slower, see notes
#endif
#ifndef SRT_WFP_DYNAMIC
#define SRT_WFP_DYNAMIC 1  // warps take 32-slot groups from a shared-memory cursor instead of a fixed mapping
#endif
#ifndef SRT_WFS_PREFETCH
#define SRT_WFS_PREFETCH 0  // 1: gather the next group's rays while the current group is traced; measured 2-4 % slower (8 more live registers spill)
#endif
#ifndef SRT_WFS_TAIL_GRAIN
#define SRT_WFS_TAIL_GRAIN 32  // 8 and 16 measured slower (most queues are short, they never leave the tail regime)
#endif
#ifndef SRT_WFS_TRACE_NEW_FIRST
#define SRT_WFS_TRACE_NEW_FIRST 0  // 0: continuing (incoherent, slow) rays first, so the phase ends on short groups
#endif
#ifndef SRT_WFP_MINBLOCKS
#define SRT_WFP_MINBLOCKS 1
#endif

// The shade phase of the persistent kernel, for one warp: wf_shade_slot's work, but scheduled per
// LANE.  A lane that reaches the `while (pdf_val == 0)` rejection loop of color() (:79-84) keeps its
// slot and runs ONE iteration per trip of the loop below, while the lanes that are done store their
// slot and fetch the next one from the CTA's cursor.  In the per-slot form the warp waits for its
// unluckiest lane: the loop ran at 3 of 32 lanes and held a quarter of the kernel's instructions
// (profiles/r01_notes.md, capture E).  Same paths, same Philox counters, same sums.
template <uint32_t F>
SRT_DEV int wf_shade_phase(const DScene& S, const View& W, const RenderArgs& A, const WfState& st,
                           const uint32_t* __restrict__ sobol, long long* __restrict__ out,
                           unsigned long long* __restrict__ work_counter, int cta_base, int n_cta_slots, int* cursor,
                           volatile int* exhausted, bool first) {
    Tally<false> tl{nullptr};
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int n_pixels = A.nx * A.ny;
    int n_alive = 0;
    bool sampling = false, more = true;
    int slot = 0;
    V3 sum = v3(0.f, 0.f, 0.f);
    int pixel = 0, chunk_id = 0, s_next = 0, s_stop = 0;
    bool has_item = false;
    Philox rng;
    rng.have = 0;
    Path P;
    SampleCtx C;
    for (;;) {
        // ---- 1. lanes without a pending vertex take the next slots of this CTA ------------------------
        const unsigned free_m = __ballot_sync(FULL, !sampling);
        bool got = false;
        if (more && free_m) {
            int base = 0;
            if (lane == 0) base = atomicAdd(cursor, __popc(free_m));
            base = __shfl_sync(FULL, base, 0);
            const int idx = base + __popc(free_m & lt_mask);
            got = !sampling && idx < n_cta_slots;
            slot = got ? cta_base + idx : slot;
            more = base + __popc(free_m) < n_cta_slots;
        }
        if (!__ballot_sync(FULL, got || sampling)) break;
        bool finishing = false, alive = false;
        if (got) {
            float4 sm = st.sum[slot];
            int flags = first ? 0 : __float_as_int(sm.w);
            sum = first ? v3(0.f, 0.f, 0.f) : v3(sm.x, sm.y, sm.z);
            int4 it = first ? make_int4(0, 0, 0, 0) : st.item[slot];
            pixel = it.x, chunk_id = it.y, s_next = it.z, s_stop = it.w;
            alive = (flags & kWfAlive) != 0;
            has_item = (flags & kWfHasItem) != 0;
            finishing = true;
            if (alive) {
                float4 o = st.ray_o[slot], d = st.ray_d[slot], hv = st.hit[slot], tv = st.thr[slot], lv = st.rad[slot];
                P.r = Ray{v3(o.x, o.y, o.z), v3(d.x, d.y, d.z), o.w};
                P.thr = v3(tv.x, tv.y, tv.z);
                P.depth = __float_as_int(tv.w);
                P.L = v3(lv.x, lv.y, lv.z);
                Hit h;
                h.t = hv.x;
                h.prim = __float_as_int(hv.y);
                h.entry = __float_as_int(hv.z);
                wf_load_rng(st, slot, A.seed, (uint32_t)pixel, (uint32_t)(s_next - 1), rng);
                ShadeStatus ss = shade_hit<F>(W, P, h, A.max_depth, rng, C, tl);
                if (ss == SHADE_SAMPLE) {
                    sampling = true;
                    finishing = false;
                } else if (ss == SHADE_ENDED) {
                    V3 c = de_nan(P.L);
                    sum = v3(sum.x + c.x, sum.y + c.y, sum.z + c.z);
                    alive = false;
                }
            }
        }
        // ---- 2. one iteration of the rejection loop for every lane that holds a diffuse vertex ----------
        if (sampling && shade_sample_iter<F>(W, C, rng, tl)) {
            alive = shade_finish<F>(P, C, tl);
            if (!alive) {
                V3 c = de_nan(P.L);
                sum = v3(sum.x + c.x, sum.y + c.y, sum.z + c.z);
            }
            sampling = false;
            finishing = true;
        }
        // ---- 3. lanes that are through: close / take a work item, start the next path, store ------------
        if (finishing && !alive && has_item && s_next >= s_stop) {
            item_commit(out, pixel, sum, A.fx_clamp);
            has_item = false;
        }
        const bool need_item = finishing && !alive && !has_item;
        const unsigned m = __ballot_sync(FULL, need_item);
        if (m && !*exhausted) {
            unsigned long long base = 0;
            if (lane == 0) {
                base = atomicAdd(work_counter, (unsigned long long)__popc(m));
                if (base >= A.n_items) *exhausted = 1;
            }
            base = __shfl_sync(FULL, base, 0);
            const unsigned long long item = base + (unsigned)__popc(m & lt_mask);
            if (need_item && item < A.n_items) {
                decode_item(A, item, pixel, chunk_id, s_next, s_stop);
                sum = v3(0.f, 0.f, 0.f);
                has_item = true;
            }
        }
        if (finishing) {
            if (!alive && has_item && s_next < s_stop) {
                int s = s_next++;
                P.r = start_path(S.camera, A, sobol, pixel, s, rng);
                P.thr = v3(1.f, 1.f, 1.f);
                P.L = v3(0.f, 0.f, 0.f);
                P.depth = 0;
                alive = true;
            }
            const int flags = (alive ? kWfAlive : 0) | (has_item ? kWfHasItem : 0);
            st.sum[slot] = make_float4(sum.x, sum.y, sum.z, __int_as_float(flags));
            st.item[slot] = make_int4(pixel, chunk_id, s_next, s_stop);
            if (alive) {
                st.ray_o[slot] = make_float4(P.r.o.x, P.r.o.y, P.r.o.z, P.r.time);
                st.ray_d[slot] = make_float4(P.r.d.x, P.r.d.y, P.r.d.z, 0.f);
                st.thr[slot] = make_float4(P.thr.x, P.thr.y, P.thr.z, __int_as_float(P.depth));
                st.rad[slot] = make_float4(P.L.x, P.L.y, P.L.z, 0.f);
                wf_store_rng(st, slot, rng);
                n_alive++;
            }
        }
    }
    return n_alive;
}

template <uint32_t F>
__global__ void __launch_bounds__(SRT_WFP_THREADS, SRT_WFP_MINBLOCKS)
    k_wf_persist(DScene S, RenderArgs A, WfState st, const uint32_t* __restrict__ sobol, long long* __restrict__ out,
                 unsigned long long* __restrict__ counters, int per_thread) {
    extern __shared__ float4 smem_hot[];
    __shared__ int s_exhausted;
    __shared__ int s_cursor[2];  // next slot of the shade phase / next 32-slot group of the trace phase
    if (threadIdx.x == 0) s_exhausted = s_cursor[0] = s_cursor[1] = 0;
    View W = wf_view<F>(S, smem_hot);
    __syncthreads();
    const int n_cta_slots = per_thread * blockDim.x;
    const int cta_base = blockIdx.x * n_cta_slots;
    const int lane = threadIdx.x & 31;
    for (int iter = 0;; iter++) {
#if SRT_WFP_LANE_SCHED
        const int n_alive = wf_shade_phase<F>(S, W, A, st, sobol, out, counters + 7, cta_base, n_cta_slots, &s_cursor[0],
                                              &s_exhausted, iter == 0);
#else
        int n_alive = 0;
#if SRT_WFP_DYNAMIC
        for (;;) {
            int g = 0;
            if (lane == 0) g = atomicAdd(&s_cursor[0], 32);
            g = __shfl_sync(0xffffffffu, g, 0);
            if (g >= n_cta_slots) break;
            n_alive += wf_shade_slot<F>(S, W, A, st, sobol, out, counters + 7, cta_base + g + lane, iter == 0, &s_exhausted)
                           ? 1
                           : 0;
        }
#else
        for (int k = 0; k < per_thread; k++)
            n_alive += wf_shade_slot<F>(S, W, A, st, sobol, out, counters + 7, cta_base + k * blockDim.x + threadIdx.x,
                                        iter == 0, &s_exhausted)
                           ? 1
                           : 0;
#endif
#endif
        const int any = __syncthreads_or(n_alive);
        if (threadIdx.x == 0) s_cursor[0] = 0;  // next read after the barrier below
        if (!any) break;
        // trace phase: warps take groups of 32 slots until the CTA's range is through
#if SRT_WFP_DYNAMIC
        for (;;) {
            int g = 0;
            if (lane == 0) g = atomicAdd(&s_cursor[1], 32);
            g = __shfl_sync(0xffffffffu, g, 0);
            if (g >= n_cta_slots) break;
            wf_trace_slot<F>(W, A, st, cta_base + g + lane);
        }
#else
        for (int k = 0; k < per_thread; k++) wf_trace_slot<F>(W, A, st, cta_base + k * blockDim.x + threadIdx.x);
#endif
        __syncthreads();
        if (threadIdx.x == 0) s_cursor[1] = 0;  // next read after the next barrier
    }
}


// ---- SORTED persistent form: material-sorted shade queues, generation-sorted trace queues ------------
// k_wf_persist walks its slots in storage order, so a warp shades whatever mix of materials its 32
// slots happen to hold (capture E: the shading code ran at 6 of 32 lanes) and traces fresh camera
// rays next to fifth-bounce rays.  Here the CTA keeps, in shared memory, index queues over its own
// slots (16-bit indices, `kSortCap` per queue):
//   * the TRACE phase appends every traced slot to the queue of what it hit — miss, light, diffuse,
//     Beckmann, metal, glass — and the SHADE phase drains those queues one after the other, 32
//     entries per warp: all lanes of a warp run the same material's code;
//   * the SHADE phase appends every continuing path to the `cont` queue and every path it starts to
//     the `new` queue (lanes that start paths together take consecutive work items: neighbouring
//     samples of one pixel), and the TRACE phase drains `new` first: coherent camera rays stay together.
// Appends are one ballot + one shared-memory atomic per warp.  A slot without a path is in no queue:
// the tail of the frame shrinks with the live set.  Requires one path per work item (chunk == 1).
constexpr int kSortCap = 8192;  // most slots a CTA may own (the queues live in dynamic shared memory)
enum { SC_DEAD = 0, SC_MISS, SC_LIGHT, SC_DIFFUSE, SC_BECKMANN, SC_METAL, SC_GLASS, SC_N, TC_NEW = SC_N, TC_CONT, Q_N };

struct SortQueues {
    unsigned short* idx;  // [Q_N][cap], dynamic shared memory behind the hot tables
    int cap;
    int count[Q_N];
    int cursor[Q_N];
    int exhausted;
};

SRT_DEV void q_push(SortQueues& Q, int q, bool pred, int local_slot) {
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (!m) return;
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == 0) base = atomicAdd(&Q.count[q], __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (pred) Q.idx[q * Q.cap + base + __popc(m & ((1u << lane) - 1u))] = (unsigned short)local_slot;
}

// append each valid lane to the queue `q` of its own: lanes with the same q find each other with one
// match, their lowest lane reserves the range with one shared-memory atomic
SRT_DEV void q_push_class(SortQueues& Q, bool valid, int q, int local_slot) {
    const int lane = threadIdx.x & 31;
    const unsigned peers = __match_any_sync(0xffffffffu, valid ? q : -1);
    const int leader = __ffs(peers) - 1;
    int base = 0;
    if (valid && lane == leader) base = atomicAdd(&Q.count[q], __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (valid) Q.idx[q * Q.cap + base + __popc(peers & ((1u << lane) - 1u))] = (unsigned short)local_slot;
}

// The next entries of queue q for this warp: 32 at a time, and SRT_WFS_TAIL_GRAIN at a time once fewer than
// one full round of the CTA's warps is left — a group lasts as long as its slowest lane, and the warps
// that are through wait at the phase barrier for the last group (capture H: 12.7 % of C3's stall samples),
// so the last entries are spread over all warps in short groups.  Returns false when the queue is drained.
SRT_DEV bool q_take(SortQueues& Q, int q, int& local_slot, bool& valid, int& pos) {
    const int lane = threadIdx.x & 31;
    const int n = Q.count[q];
    int g = 0, step = 32;
    if (lane == 0) {
        const int left = n - *(volatile int*)&Q.cursor[q];
        step = left > (int)blockDim.x ? 32 : SRT_WFS_TAIL_GRAIN;
        g = atomicAdd(&Q.cursor[q], step);
    }
    g = __shfl_sync(0xffffffffu, g, 0);
    step = __shfl_sync(0xffffffffu, step, 0);
    if (g >= n) return false;
    pos = g + lane;
    valid = lane < step && pos < n;
    local_slot = valid ? (int)Q.idx[q * Q.cap + pos] : 0;
    return true;
}

template <uint32_t F>
SRT_DEV int wf_hit_class(const View& W, const Hit& h) {
    const DScene& S = W.S;
    if (h.prim < 0) return SC_MISS;
    int mat_index;
    if ((F & F_MEDIA) && h.prim >= S.n_prims)
        mat_index = S.media[h.prim - S.n_prims].phase_material;
    else
        mat_index = load_prim(W.dprims + h.prim).material;
    if (mat_index < 0) return SC_MISS;
    const int type = __float_as_int(__ldg(reinterpret_cast<const float*>(S.materials + mat_index)));
    if (type == SRT_MAT_DIFFUSE_LIGHT) return SC_LIGHT;
    if (type == SRT_MAT_BECKMANN) return SC_BECKMANN;
    if (type == SRT_MAT_METAL) return SC_METAL;
    if (type == SRT_MAT_DIELECTRIC || type == SRT_MAT_ISOTROPIC) return SC_GLASS;
    return SC_DIFFUSE;
}

template <uint32_t F, bool COUNT>
__global__ void __launch_bounds__(SRT_WFP_THREADS, SRT_WFP_MINBLOCKS)
    k_wf_sorted(DScene S, RenderArgs A, WfState st, const uint32_t* __restrict__ sobol, long long* __restrict__ out,
                unsigned long long* __restrict__ counters, int per_thread) {
    extern __shared__ float4 smem_hot[];
    __shared__ SortQueues Q;
    __shared__ unsigned long long s_item_base;  // first work item of this pass's regeneration
    __shared__ int s_item_avail;                // how many of the queued slots get one
    View W = wf_view<F>(S, smem_hot);
    const int lane = threadIdx.x & 31;
    const int cap = per_thread * blockDim.x;  // <= kSortCap (srt_api.cu)
    const int cta_base = blockIdx.x * cap;
    unsigned long long* work_counter = counters + 7;
    // COUNT: the instrumented build of THIS kernel (srt_render_counted): same schedule, same paths, plus the
    // per-thread tallies of SrtCounters (the shipping kernel's own visit / test counters)
    Counters cnt = {0, 0, 0, 0, 0, 0, 0, 0};
    Tally<COUNT> tl{&cnt};
    if (threadIdx.x < Q_N) Q.count[threadIdx.x] = Q.cursor[threadIdx.x] = 0;
    if (threadIdx.x == 0) {
        Q.exhausted = 0;
        Q.cap = cap;
        Q.idx = reinterpret_cast<unsigned short*>(smem_hot + ((F & F_SMEM_TABLES) ? ((S.hot_bytes + 15) >> 4) : 0));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < cap; i += blockDim.x) Q.idx[SC_DEAD * cap + i] = (unsigned short)i;
    __syncthreads();
    if (threadIdx.x == 0) Q.count[SC_DEAD] = cap;  // every slot starts without a path
    for (;;) {
        // ---- REGENERATE: every slot whose path ended in the last shade phase (queue SC_DEAD) starts the
        // next path, Raytracing_n.cpp:827-836.  The CTA takes ALL the work items it needs with one atomic
        // on the global counter; queue position i gets item base + i, so the new camera rays of the CTA
        // are consecutive samples of one or two pixels, generated by full warps in converged code. --------
        // once the global counter has run out no slot can be restarted: the drain at the end of the frame
        // skips this pass and its two barriers.  The flag is read BEFORE the barrier: thread 0 sets it
        // between the two barriers below, so every thread of the CTA sees the same value here.
        const int exhausted_before = Q.exhausted;
        __syncthreads();
        if (!exhausted_before) {
            if (threadIdx.x == 0) {
                const int n = Q.count[SC_DEAD];
                unsigned long long base = 0;
                int avail = 0;
                if (n > 0 && !Q.exhausted) {
                    base = atomicAdd(work_counter, (unsigned long long)n);
                    if (base < A.n_items) avail = (int)min((unsigned long long)n, A.n_items - base);
                    if (base + (unsigned long long)n >= A.n_items) Q.exhausted = 1;
                }
                s_item_base = base;
                s_item_avail = avail;
            }
            __syncthreads();
            {
                const unsigned long long base = s_item_base;
                const int avail = s_item_avail;
                int ls, pos;
                bool valid;
                while (q_take(Q, SC_DEAD, ls, valid, pos)) {
                    if (pos - lane >= avail) break;  // the rest of the queue gets no item: those slots stay empty
                    const bool start = valid && pos < avail;
                    if (start) {
                        const int slot = cta_base + ls;
                        const unsigned long long item = base + (unsigned)pos;
                        const int pixel = (int)(item / (unsigned)A.n_chunks);
                        const int s = A.s_begin + (int)(item - (unsigned long long)pixel * (unsigned)A.n_chunks);
                        Philox rng;
                        const Ray r = start_path(S.camera, A, sobol, pixel, s, rng);
                        if (COUNT) cnt.paths++;
                        st.item[slot] = make_int4(pixel, s, 0, 0);
                        st.ray_o[slot] = make_float4(r.o.x, r.o.y, r.o.z, r.time);
                        st.ray_d[slot] = make_float4(r.d.x, r.d.y, r.d.z, 0.f);
                        st.thr[slot] = make_float4(1.f, 1.f, 1.f, __int_as_float(0));
                        st.rad[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
                        wf_store_rng(st, slot, rng);
                    }
                    q_push(Q, TC_NEW, start, ls);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x < SC_N) Q.count[threadIdx.x] = Q.cursor[threadIdx.x] = 0;  // all shade queues are consumed
        const int live = Q.count[TC_NEW] + Q.count[TC_CONT];
        __syncthreads();
        if (live == 0) break;
        // ---- TRACE: world->hit for every live slot; queue by what was hit ------------------------------
        // The ray of the NEXT group is fetched (a gather over the L2-resident pool) before the current
        // group is traced, so its latency is covered by the traversal (capture L: the first uses of the
        // gathered ray held 8 % of the stall samples).
        for (int ci = 0; ci < 2; ci++) {
            const int c = (SRT_WFS_TRACE_NEW_FIRST ? ci == 0 : ci == 1) ? TC_NEW : TC_CONT;
            if (Q.count[c] == 0) continue;
            int ls, pos, ls_next = 0;
            bool valid, valid_next = false;
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f), d = o, o_next = o, d_next = o;
            bool have = q_take(Q, c, ls, valid, pos);
            if (have && valid) {
                o = st.ray_o[cta_base + ls];
                d = st.ray_d[cta_base + ls];
            }
            while (have) {
                const bool have_next = SRT_WFS_PREFETCH ? q_take(Q, c, ls_next, valid_next, pos) : false;
                if (have_next && valid_next) {
                    o_next = st.ray_o[cta_base + ls_next];
                    d_next = st.ray_d[cta_base + ls_next];
                }
                const int slot = cta_base + ls;
                int cls = SC_MISS;
                if (valid) {
                    Ray r{v3(o.x, o.y, o.z), v3(d.x, d.y, d.z), o.w};
                    Philox rng;
                    rng.have = 0;
                    if (F & F_MEDIA) {  // constant_medium::hit draws its free-flight distance (constant_medium.h:33)
                        int4 it = st.item[slot];
                        wf_load_rng(st, slot, A.seed, (uint32_t)it.x, (uint32_t)it.y, rng);
                    }
                    if (COUNT) cnt.rays++;
                    Hit h = world_hit<F>(W, r, 0.001f, kFltMax, rng, tl);
                    st.hit[slot] = make_float4(h.t, __int_as_float(h.prim), __int_as_float(h.entry), 0.f);
                    if (F & F_MEDIA) wf_store_rng(st, slot, rng);
                    cls = wf_hit_class<F>(W, h);
                }
                q_push_class(Q, valid, cls, ls);
                if (SRT_WFS_PREFETCH) {
                    have = have_next;
                    ls = ls_next;
                    valid = valid_next;
                    o = o_next;
                    d = d_next;
                } else {
                    have = q_take(Q, c, ls, valid, pos);
                    if (have && valid) {
                        o = st.ray_o[cta_base + ls];
                        d = st.ray_d[cta_base + ls];
                    }
                }
            }
        }
        __syncthreads();
        if (threadIdx.x >= TC_NEW && threadIdx.x < Q_N) Q.count[threadIdx.x] = Q.cursor[threadIdx.x] = 0;
        __syncthreads();
        // ---- SHADE: one level of color() per queued slot, material by material -------------------------
        for (int c = SC_MISS; c < SC_N; c++) {
            if (Q.count[c] == 0) continue;  // nothing of this material: no trip to the cursor
            int ls, pos;
            bool valid;
            while (q_take(Q, c, ls, valid, pos)) {
                const int slot = cta_base + ls;
                bool alive = false;
                if (valid) {
                    float4 o = st.ray_o[slot], d = st.ray_d[slot], hv = st.hit[slot], tv = st.thr[slot], lv = st.rad[slot];
                    int4 it = st.item[slot];
                    Path P;
                    P.r = Ray{v3(o.x, o.y, o.z), v3(d.x, d.y, d.z), o.w};
                    P.thr = v3(tv.x, tv.y, tv.z);
                    P.depth = __float_as_int(tv.w);
                    P.L = v3(lv.x, lv.y, lv.z);
                    Hit h;
                    h.t = hv.x;
                    h.prim = __float_as_int(hv.y);
                    h.entry = __float_as_int(hv.z);
                    Philox rng;
                    wf_load_rng(st, slot, A.seed, (uint32_t)it.x, (uint32_t)it.y, rng);
                    alive = shade_bounce<F>(W, P, h, A.max_depth, rng, tl);
                    if (alive) {
                        st.ray_o[slot] = make_float4(P.r.o.x, P.r.o.y, P.r.o.z, P.r.time);
                        st.ray_d[slot] = make_float4(P.r.d.x, P.r.d.y, P.r.d.z, 0.f);
                        st.thr[slot] = make_float4(P.thr.x, P.thr.y, P.thr.z, __int_as_float(P.depth));
                        st.rad[slot] = make_float4(P.L.x, P.L.y, P.L.z, 0.f);
                        wf_store_rng(st, slot, rng);
                    } else {
                        item_commit(out, it.x, de_nan(P.L), A.fx_clamp);  // col += de_nan(color(...)), :841
                    }
                }
                q_push(Q, TC_CONT, alive, ls);
                q_push(Q, SC_DEAD, valid && !alive, ls);
            }
        }
    }
    if (COUNT) {
        atomicAdd(&counters[0], cnt.paths);
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
