// srt_wavefront.cuh — the render loop as a persistent, sorted wavefront over a fixed pool of path slots.
//
// Why a wavefront (profiles/r01_notes.md, captures A-G): the fused kernel of srt_kernels.cuh is bound by
// instruction fetch and divergence — its hot code is larger than the SM's instruction cache and its warps
// sit at unrelated program counters — while the closest-hit code run on its own sustains five times the
// fused rate.  Here a bounce is split into a TRACE phase (world->hit only) and a SHADE phase (color()'s
// material / pdf / light-sampling step), both inside ONE persistent kernel, one CTA per SM; barriers keep
// all warps of the CTA in the same phase, so the instruction cache holds one stage's code at a time.
// (Two smaller CTAs per SM where the wait at the trace barrier costs more than the second stage's code in the
// cache: srt_variants.h SRT_WF_VARIANT_CTAS_PER_SM / SRT_WF_SPLIT_CTA_MIN_NODES, srt_api.cu run_wavefront.)
//
// State: one record per path slot in global memory (the pool is sized to stay in L2 and carries a
// persisting access-policy window, srt_api.cu).  A record is kSlotQuads 16-byte words, array-of-structures:
// a gather touches 3 consecutive 32-byte sectors of one or two lines instead of one sector in each of
// seven column arrays (round-1 layout; capture L: the gathers held a quarter of the stall samples).
//
//   q0  ray origin .xyz, time      q2  hit t, primitive, entry, -      q4  radiance so far .xyz, Philox block counter
//   q1  ray direction .xyz, -      q3  throughput .xyz, depth          q5  pixel, Sobol index, -, -
//
// The generator state of a path between two consumers of random numbers is its BLOCK COUNTER alone: every
// consumer (camera ray, one specular scatter, one attempt of the sampling loop, one free-flight draw)
// starts a new Philox block (srt_math.cuh: philox_fresh), so left-over outputs never cross a phase.
//
// Every scheduler starts the same paths with the same Philox counters and accumulates in fixed point, so
// this kernel's image equals the fused kernel's up to float contraction differences between template
// instantiations (tests/test_gpu_kernels.py).
#ifndef SRT_WAVEFRONT_CUH
#define SRT_WAVEFRONT_CUH

#include "srt_kernels.cuh"

#ifndef SRT_WFP_THREADS
#define SRT_WFP_THREADS 512
#endif
#ifndef SRT_WFP_MINBLOCKS
#define SRT_WFP_MINBLOCKS 1
#endif
#ifndef SRT_WFS_TAIL_GRAIN
#define SRT_WFS_TAIL_GRAIN 32  // 8 and 16 measured slower (most queues are short, they never leave the tail regime)
#endif
#ifndef SRT_WFS_TRACE_NEW_FIRST
#define SRT_WFS_TRACE_NEW_FIRST 0  // 0: continuing (incoherent, slow) rays first, so the phase ends on short groups
#endif
#ifndef SRT_WF_SLOT_QUADS
#define SRT_WF_SLOT_QUADS 6  // 16-byte words per slot record (6 used; 8 = one 128-byte line per record)
#endif
#ifndef SRT_WF_SLOT_CG
#define SRT_WF_SLOT_CG 0  // 1: slot records bypass L1 (ld.global.cg / st.global.cg); measured 0-2.5 % slower (profiles/r02_notes.md)
#endif

namespace srt {

constexpr int kSlotQuads = SRT_WF_SLOT_QUADS;
// corrected-transport mode (F_CORRECTED, srt_shade.cuh) keeps the shadow ray of the last vertex in the record:
//   q1.w  MIS weight of the emission the main ray may find      q5.z  flags: kNeePending | kMainDead
//   q6    shadow-ray direction .xyz    q7  its weight .xyz       q8    its closest hit t, primitive, entry
// The trace phase traces both rays of a slot through the same world_hit call; the NEXT shade phase adds the
// emitter's radiance the shadow ray found, then shades the main hit.  A path whose BSDF sample failed but
// whose shadow ray is pending stays queued (kMainDead) for one more iteration.
constexpr int kSlotQuadsCorrected = 9;
template <uint32_t F>
struct SlotQuads {
    static constexpr int value = (F & F_CORRECTED) ? kSlotQuadsCorrected : kSlotQuads;
};
enum { SQ_O = 0, SQ_D, SQ_HIT, SQ_THR, SQ_RAD, SQ_ITEM, SQ_SDIR, SQ_SW, SQ_SHIT };
enum { kNeePending = 1, kMainDead = 2 };

struct WfState {
    float4* rec;  // [n_slots][kSlotQuads]
    int n_slots;  // multiple of 32
};

SRT_DEV float4 slot_ld(const float4* p) {
#if SRT_WF_SLOT_CG
    return __ldcg(p);
#else
    return *p;
#endif
}
SRT_DEV void slot_st(float4* p, float4 v) {
#if SRT_WF_SLOT_CG
    __stcg(p, v);
#else
    *p = v;
#endif
}
SRT_DEV void slot_st1(float* p, float v) {
#if SRT_WF_SLOT_CG
    __stcg(p, v);
#else
    *p = v;
#endif
}

SRT_DEV void wf_rng(Philox& p, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t block) {
    philox_init(p, seed, pixel, sample);
    p.c2 = block;
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

// ---- index queues over the CTA's own slots (16-bit indices in shared memory) -------------------------
//   * the TRACE phase appends every traced slot to the queue of what it hit — miss, light, diffuse,
//     Beckmann, metal, glass — and the SHADE phase drains those queues one after the other, 32
//     entries per warp: all lanes of a warp run the same material's code;
//   * the SHADE phase appends every continuing path to `cont` and every ended path's slot to `dead`;
//   * the REGENERATE pass starts new paths in the dead slots and appends them to `new` (consecutive work
//     items: neighbouring samples of one pixel), and the TRACE phase drains `cont` and `new` separately:
//     coherent camera rays stay together.
// Appends are one match + one shared-memory atomic per warp.  A slot without a path is in no queue: the
// tail of the frame shrinks with the live set.  Requires one path per work item (chunk == 1).
//
// The `cont` queue alternates between two queues by iteration: the shade phase fills the one the NEXT trace
// phase drains, so the drained one is reset without a barrier of its own.
//
// Measured and rejected (profiles/r02_notes.md): letting the warps that are through with the trace phase start
// shading full groups while the others still trace (the warps of a CTA are busy for only 78-91 % of the trace
// phase; the rest is the wait at its closing barrier) — 7-13 % SLOWER on every configuration, with or without
// polling: the two stages' code and data then share the SM's caches again, which is what the phases avoid.
// Re-queueing a slot whose sampling attempt was rejected (pdf_val == 0, Raytracing_n.cpp:79-84) instead of
// looping in the warp: -4 % on C4, +4 % on C2, +8 % on C3.  Filling the tail of the trace phase with the work of
// the next regenerate pass (the pass leaves 1/8 of the slots empty; a warp that finds the trace queues drained
// starts 32 new paths there and traces their camera rays at once, same code as the warps still tracing): the
// warps are then busy for 82-94 % of the phase, but the frame is 0-4 % slower — the emptier pool costs more
// than the shorter wait gives back.  Index queues in global memory (L2) instead of shared memory, which hands
// ~110 KB per SM back to L1: no difference on any configuration.
constexpr int kSortCap = 8192;  // most slots a CTA may own (the queues live in dynamic shared memory)
enum {
    SC_DEAD = 0, SC_MISS, SC_LIGHT, SC_DIFFUSE, SC_BECKMANN, SC_METAL, SC_GLASS, SC_N,
    TC_NEW = SC_N, TC_CONT_A, TC_CONT_B, Q_N
};

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
// one full round of the CTA's warps is left.  Returns false when the queue is drained.  Only for queues that
// nobody appends to at the same time.
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

// One group of the shade phase: one level of color() (Raytracing_n.cpp:55-106) for the slot of every valid
// lane, then the slot goes to the `cont` queue the NEXT trace phase drains, or to `dead`.
template <uint32_t F, bool COUNT>
SRT_DEV void wf_shade_group(const View& W, const RenderArgs& A, SortQueues& Q, float4* pool, long long* __restrict__ out,
                            int ls, bool valid, int q_cont, Tally<COUNT> tl) {
    float4* rec = pool + (size_t)ls * SlotQuads<F>::value;
    int dest = SC_DEAD;
    if (valid) {
        const float4 o = slot_ld(rec + SQ_O), d = slot_ld(rec + SQ_D), hv = slot_ld(rec + SQ_HIT), tv = slot_ld(rec + SQ_THR),
                     lv = slot_ld(rec + SQ_RAD), it = slot_ld(rec + SQ_ITEM);
        Path P;
        P.r = Ray{v3(o.x, o.y, o.z), v3(d.x, d.y, d.z), o.w};
        P.thr = v3(tv.x, tv.y, tv.z);
        P.depth = __float_as_int(tv.w);
        P.L = v3(lv.x, lv.y, lv.z);
        P.emit_w = d.w;
        Hit h;
        h.t = hv.x;
        h.prim = __float_as_int(hv.y);
        h.entry = __float_as_int(hv.z);
        const int pixel = __float_as_int(it.x);
        Philox rng;
        wf_rng(rng, A.seed, (uint32_t)pixel, __float_as_uint(it.y), __float_as_uint(lv.w));
        bool cont;
        int flags = 0;
        Nee nee;
        nee.valid = false;
        if (F & F_CORRECTED) {
            const int old_flags = __float_as_int(it.z);
            if (old_flags & kNeePending) {  // what the shadow ray of the previous vertex reached
                const float4 sd = slot_ld(rec + SQ_SDIR), sw = slot_ld(rec + SQ_SW), sh = slot_ld(rec + SQ_SHIT);
                Hit hs;
                hs.t = sh.x;
                hs.prim = __float_as_int(sh.y);
                hs.entry = __float_as_int(sh.z);
                const V3 e = emitted_at<F>(W, Ray{P.r.o, v3(sd.x, sd.y, sd.z), P.r.time}, hs);
                P.L = v3(fmaf(sw.x, e.x, P.L.x), fmaf(sw.y, e.y, P.L.y), fmaf(sw.z, e.z, P.L.z));
            }
            cont = !(old_flags & kMainDead) && shade_bounce_corrected<F>(W, P, h, A.max_depth, rng, nee, tl);
            flags = (nee.valid ? kNeePending : 0) | (cont ? 0 : kMainDead);
            if (!cont && nee.valid) P.r.o = nee.o;  // the shadow ray starts at the vertex
        } else {
            cont = shade_bounce<F>(W, P, h, A.max_depth, rng, tl);
        }
        if (cont || nee.valid) {
            slot_st(rec + SQ_O, make_float4(P.r.o.x, P.r.o.y, P.r.o.z, P.r.time));
            slot_st(rec + SQ_D, make_float4(P.r.d.x, P.r.d.y, P.r.d.z, P.emit_w));
            slot_st(rec + SQ_THR, make_float4(P.thr.x, P.thr.y, P.thr.z, __int_as_float(P.depth)));
            slot_st(rec + SQ_RAD, make_float4(P.L.x, P.L.y, P.L.z, __uint_as_float(rng.c2)));
            if (F & F_CORRECTED) {
                slot_st(rec + SQ_ITEM, make_float4(it.x, it.y, __int_as_float(flags), 0.f));
                if (nee.valid) {
                    slot_st(rec + SQ_SDIR, make_float4(nee.dir.x, nee.dir.y, nee.dir.z, 0.f));
                    slot_st(rec + SQ_SW, make_float4(nee.weight.x, nee.weight.y, nee.weight.z, 0.f));
                }
            }
            dest = q_cont;
        } else {
            item_commit(out, pixel, de_nan(P.L), A.fx_clamp);  // col += de_nan(color(...)), :841
        }
    }
    q_push_class(Q, valid, dest, ls);
}

// SRT_WFS_PHASE_CLOCKS: thread 0 of every CTA adds the clock64 ticks it spends in each phase (incl. the wait
// at the closing barrier) to counters[10..13] (regenerate, trace, shade, iterations), and every warp the
// clocks between its first take and its last group of a phase to counters[14..15]: a debug build for
// profiles/, not the shipped kernel.
#ifdef SRT_WFS_PHASE_CLOCKS
#define SRT_PHASE_TICK(i)                                            \
    do {                                                             \
        if (threadIdx.x == 0) {                                      \
            const long long now_ = clock64();                        \
            phase_ticks[i] += (unsigned long long)(now_ - phase_t0); \
            phase_t0 = now_;                                         \
        }                                                            \
    } while (0)
// busy clocks of THIS warp inside a phase (first take to last group done): sum over warps / (warps x phase
// clocks) = how full the SM is kept, i.e. what the closing barrier of the phase costs
#define SRT_WARP_BUSY_BEGIN() const long long busy_t0_ = clock64()
#define SRT_WARP_BUSY_END(i)                                                       \
    do {                                                                           \
        if (lane == 0) warp_busy[i] += (unsigned long long)(clock64() - busy_t0_); \
    } while (0)
#else
#define SRT_PHASE_TICK(i) \
    do {                  \
    } while (0)
#define SRT_WARP_BUSY_BEGIN() \
    do {                      \
    } while (0)
#define SRT_WARP_BUSY_END(i) \
    do {                     \
    } while (0)
#endif

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
    constexpr int NQ = SlotQuads<F>::value;
    float4* const pool = st.rec + (size_t)blockIdx.x * cap * NQ;  // this CTA's records
    unsigned long long* work_counter = counters + 7;
    // COUNT: the instrumented build of THIS kernel (srt_render_counted): same schedule, same paths, plus the
    // per-thread tallies of SrtCounters (the shipping kernel's own visit / test counters)
    Counters cnt = {0, 0, 0, 0, 0, 0, 0, 0};
    Tally<COUNT> tl{&cnt};
#ifdef SRT_WFS_PHASE_CLOCKS
    unsigned long long phase_ticks[4] = {0, 0, 0, 0};
    unsigned long long warp_busy[3] = {0, 0, 0};
    long long phase_t0 = clock64();
#endif
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
    // shade queues, the expensive classes first, so that the phase ends on the short groups (miss, light)
    // and the warps reach its closing barrier close together
    constexpr int kShadeOrder[SC_N - 1] = {SC_BECKMANN, SC_DIFFUSE, SC_GLASS, SC_METAL, SC_LIGHT, SC_MISS};
    for (int iter = 0;; iter++) {
        // `cont` queues alternate: this iteration's trace phase drains tc_cur, its shade phase fills tc_next
        const int tc_cur = TC_CONT_A + (iter & 1), tc_next = TC_CONT_A + ((iter & 1) ^ 1);
        // ---- REGENERATE: every slot whose path ended in the last shade phase (queue SC_DEAD) starts the
        // next path, Raytracing_n.cpp:827-836.  The CTA takes ALL the work items it needs with one atomic
        // on the global counter; queue position i gets item base + i, so the new camera rays of the CTA
        // are consecutive samples of one or two pixels, generated by full warps in converged code. --------
        // once the global counter has run out no slot can be restarted: the drain at the end of the frame
        // skips this pass and its two barriers.  The flag is read BEFORE the barrier: thread 0 sets it
        // between the two barriers below, so every thread of the CTA sees the same value here.
        const int exhausted_before = Q.exhausted;
        __syncthreads();
        SRT_PHASE_TICK(2);  // the shade phase ends at this barrier
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
                        float4* rec = pool + (size_t)ls * NQ;
                        const unsigned long long item = base + (unsigned)pos;
                        const int pixel = (int)(item / (unsigned)A.n_chunks);
                        const int s = A.s_begin + (int)(item - (unsigned long long)pixel * (unsigned)A.n_chunks);
                        Philox rng;
                        const Ray r = start_path(S.camera, A, sobol, pixel, s, rng);
                        if (COUNT) cnt.paths++;
                        slot_st(rec + SQ_O, make_float4(r.o.x, r.o.y, r.o.z, r.time));
                        slot_st(rec + SQ_D, make_float4(r.d.x, r.d.y, r.d.z, 1.f));
                        slot_st(rec + SQ_THR, make_float4(1.f, 1.f, 1.f, __int_as_float(0)));
                        slot_st(rec + SQ_RAD, make_float4(0.f, 0.f, 0.f, __uint_as_float(rng.c2)));
                        slot_st(rec + SQ_ITEM, make_float4(__int_as_float(pixel), __int_as_float(s), 0.f, 0.f));
                    }
                    q_push(Q, TC_NEW, start, ls);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x < SC_N) Q.count[threadIdx.x] = Q.cursor[threadIdx.x] = 0;  // all shade queues are consumed
        const int live = Q.count[TC_NEW] + Q.count[tc_cur];
        __syncthreads();
        SRT_PHASE_TICK(0);
        if (live == 0) break;
        // ---- TRACE: world->hit for every live slot; queue by what was hit ------------------------------
        {
            SRT_WARP_BUSY_BEGIN();
            for (int ci = 0; ci < 2; ci++) {
                const int c = (SRT_WFS_TRACE_NEW_FIRST ? ci == 0 : ci == 1) ? TC_NEW : tc_cur;
                if (Q.count[c] == 0) continue;
                int ls, pos;
                bool valid;
                while (q_take(Q, c, ls, valid, pos)) {
                    float4* rec = pool + (size_t)ls * NQ;
                    int cls = SC_MISS;
                    float4 o = make_float4(0.f, 0.f, 0.f, 0.f), d = o;
                    int flags = 0;
                    Philox rng;
                    rng.have = 0;
                    if (valid) {
                        o = slot_ld(rec + SQ_O);
                        d = slot_ld(rec + SQ_D);
                        if (F & (F_MEDIA | F_CORRECTED)) {
                            const float4 lv = slot_ld(rec + SQ_RAD), it = slot_ld(rec + SQ_ITEM);
                            // constant_medium::hit draws its free-flight distance (constant_medium.h:33)
                            if (F & F_MEDIA) wf_rng(rng, A.seed, __float_as_uint(it.x), __float_as_uint(it.y), __float_as_uint(lv.w));
                            if (F & F_CORRECTED) flags = __float_as_int(it.z);
                        }
                    }
                    // the main ray, then (corrected-transport mode) the shadow ray of the slot's last vertex: one
                    // world_hit call site for both
                    for (int k = 0; k < ((F & F_CORRECTED) ? 2 : 1); k++) {
                        const bool act = valid && (k == 0 ? !(flags & kMainDead) : (flags & kNeePending) != 0);
                        if ((F & F_CORRECTED) && !__any_sync(0xffffffffu, act)) continue;
                        if (act) {
                            if (k == 1) d = slot_ld(rec + SQ_SDIR);
                            Ray r{v3(o.x, o.y, o.z), v3(d.x, d.y, d.z), o.w};
                            if (COUNT) cnt.rays++;
                            Hit h = world_hit<F>(W, r, 0.001f, kFltMax, rng, tl);
                            slot_st(rec + (k == 0 ? SQ_HIT : SQ_SHIT),
                                    make_float4(h.t, __int_as_float(h.prim), __int_as_float(h.entry), 0.f));
                            if (k == 0) cls = wf_hit_class<F>(W, h);
                        }
                    }
                    if ((F & F_MEDIA) && valid) slot_st1(reinterpret_cast<float*>(rec + SQ_RAD) + 3, __uint_as_float(rng.c2));
                    q_push_class(Q, valid, cls, ls);
                }
            }
            SRT_WARP_BUSY_END(0);
        }
        __syncthreads();
        // the trace queues are drained; nobody appends to them before the next iteration
        if (threadIdx.x == TC_NEW || threadIdx.x == tc_cur) Q.count[threadIdx.x] = Q.cursor[threadIdx.x] = 0;
        SRT_PHASE_TICK(1);
        // ---- SHADE: one level of color() per queued slot, material by material ----------------------------
        {
            SRT_WARP_BUSY_BEGIN();
            for (int ci = 0; ci < SC_N - 1; ci++) {
                const int c = kShadeOrder[ci];
                if (Q.count[c] == 0) continue;  // nothing of this material: no trip to the cursor
                int ls, pos;
                bool valid;
                while (q_take(Q, c, ls, valid, pos)) wf_shade_group<F>(W, A, Q, pool, out, ls, valid, tc_next, tl);
            }
            SRT_WARP_BUSY_END(1);
        }
#ifdef SRT_WFS_PHASE_CLOCKS
        if (threadIdx.x == 0) phase_ticks[3]++;
#endif
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
#ifdef SRT_WFS_PHASE_CLOCKS
    if (threadIdx.x == 0)
        for (int i = 0; i < 4; i++) atomicAdd(&counters[10 + i], phase_ticks[i]);
    if (lane == 0) {
        atomicAdd(&counters[14], warp_busy[0]);
        atomicAdd(&counters[15], warp_busy[1]);
    }
#endif
}

}  // namespace srt
#endif
