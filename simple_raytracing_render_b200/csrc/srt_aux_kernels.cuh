// srt_aux_kernels.cuh — the small __global__ entry points around the render kernel: batch closest
// hit, light pdf, partial-sum reduction, resolve, MERL lookup, division self-test.  Included by
// srt_api.cu only.
#ifndef SRT_AUX_KERNELS_CUH
#define SRT_AUX_KERNELS_CUH

#include "srt_kernels.cuh"

namespace srt {

// Closest hit for a batch of rays: world->hit(r, 0.001, FLT_MAX, rec) (Raytracing_n.cpp:58).
template <uint32_t FM>
__global__ void __launch_bounds__(128) k_trace_closest(DScene S, const float* __restrict__ rays, long long n,
                                                       int32_t* __restrict__ prim, float* __restrict__ t,
                                                       float* __restrict__ rec, uint32_t seed) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    View W = global_view(S);
    const float* q = rays + 7 * i;
    Ray r{v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6]};
    Philox rng;
    philox_init(rng, seed, (uint32_t)i, (uint32_t)(i >> 32));
    Tally<false> tl{nullptr};
    Hit h = world_hit<FM>(W, r, 0.001f, kFltMax, rng, tl);
    if (h.prim < 0) {
        prim[i] = -1;
        t[i] = 0.f;
        if (rec)
            for (int k = 0; k < 8; k++) rec[8 * i + k] = 0.f;
        return;
    }
    prim[i] = h.prim;
    t[i] = h.t;
    if (rec) {
        HitRecord hr = make_record<FM>(W, r, h, true);
        float* o = rec + 8 * i;
        o[0] = hr.u;
        o[1] = hr.v;
        o[2] = hr.p.x;
        o[3] = hr.p.y;
        o[4] = hr.p.z;
        o[5] = hr.normal.x;
        o[6] = hr.normal.y;
        o[7] = hr.normal.z;
    }
}

// The camera rays of the sample loop (Raytracing_n.cpp:827-836, camera.h:51-59) through start_path, the routine
// every render scheduler starts its paths with.  One thread per (pixel, sample); rays: [n_pixels * count][7].
__global__ void __launch_bounds__(128) k_primary_rays(SrtCamera cam, RenderArgs A, const uint32_t* __restrict__ sobol,
                                                      float* __restrict__ rays) {
    const long long n = (long long)A.nx * A.ny * (A.s_end - A.s_begin);
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int count = A.s_end - A.s_begin;
    const int pixel = (int)(i / count);
    const int s = A.s_begin + (int)(i - (long long)pixel * count);
    Philox rng;
    const Ray r = start_path(cam, A, sobol, pixel, s, rng);
    float* o = rays + 7 * i;
    o[0] = r.o.x;
    o[1] = r.o.y;
    o[2] = r.o.z;
    o[3] = r.d.x;
    o[4] = r.d.y;
    o[5] = r.d.z;
    o[6] = r.time;
}

// hitable_list::pdf_value over the light list (hitable_list.h:54-62) for (origin, direction) pairs.
__global__ void __launch_bounds__(128) k_light_pdf(DScene S, const float* __restrict__ od, long long n,
                                                   float* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    View W = global_view(S);
    const float* q = od + 6 * i;
    Tally<false> tl{nullptr};
    out[i] = S.n_lights > 0 ? light_pdf_value<F_ALL_SCENE>(W, v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), tl) : 0.f;
}

// accum[i] += fixed[i] * 2^-31 (the render kernels' fixed-point sums, srt_kernels.cuh: item_commit), and
// fixed[i] is cleared for the next render (n_values = 3 * n_pixels).
__global__ void __launch_bounds__(256) k_fixed_to_accum(long long* __restrict__ fixed, long long n_values,
                                                        float* __restrict__ accum) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_values) return;
    accum[i] += (float)((double)fixed[i] * 4.656612873077392578125e-10);
    fixed[i] = 0;
}

// mean -> sqrt -> int(255.99 c) -> clamp (Raytracing_n.cpp:848-862)
__global__ void __launch_bounds__(256) k_resolve(const float* __restrict__ accum, long long n_values, float ns,
                                                 uint8_t* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_values) return;
    float k = (float)(1.0 / (double)ns);  // vec3::operator/=(float): k = 1.0 / t, then e *= k
    float c = xmul(accum[i], k);
    c = xsqrt(c);
    int v = (int)(255.99 * (double)c);
    v = v > 255 ? 255 : v;
    v = v < 0 ? 0 : v;
    out[i] = (uint8_t)v;
}

// The fixed-point sums of a whole render (all shards already added) straight to the 8-bit image: the two
// kernels above in one pass, same float operations in the same order (the float sum a caller of srt_render
// would hold is (float)(fixed * 2^-31) added to 0), so the bytes equal those of srt_render + srt_resolve.
// accum_out (optional) receives that float sum; clear != 0 leaves the buffer zero for the next render.
__global__ void __launch_bounds__(256) k_resolve_fixed(long long* __restrict__ fixed, long long n_values, float ns,
                                                       uint8_t* __restrict__ out, float* __restrict__ accum_out, int clear) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_values) return;
    float sum = (float)((double)fixed[i] * 4.656612873077392578125e-10);
    if (clear) fixed[i] = 0;
    if (accum_out) accum_out[i] = sum;
    if (out) {
        float k = (float)(1.0 / (double)ns);
        float c = xsqrt(xmul(sum, k));
        int v = (int)(255.99 * (double)c);
        v = v > 255 ? 255 : v;
        v = v < 0 ? 0 : v;
        out[i] = (uint8_t)v;
    }
}

// ---- the exchange step of the multi-GPU path over NVLink PEER MEMORY --------------------
// Every rank holds one block [fixed 0 | fixed 1 | rgb | flags] that all ranks of the box have mapped
// (cudaIpcOpenMemHandle).  Frame f accumulates in `fixed (f & 1)`.  k_exchange_peers on rank g:
//   1. ARRIVE: tells every rank "g has finished rendering frame f" (a 4-byte remote store per rank), then waits until
//      every rank has told it the same — the only synchronisation of the step, inside the kernel, no host, no NCCL;
//   2. for ITS pixel tile: loads the int64 sums of all ranks (local + peer loads over NVLink, 16 bytes per load), adds
//      them exactly, resolves (mean -> sqrt -> int(255.99 c) -> clamp, Raytracing_n.cpp:848-862) and stores the four
//      8-bit values of a thread as one word into the image of EVERY rank;
//   3. DONE: the last CTA tells every rank that tile g of frame f is complete in its image.
// k_exchange_wait then holds the stream until the tiles of all ranks have arrived.  The sums of frame f are cleared by
// their owner after ITS frame-(f+1) kernel has passed step 1: by then every rank has finished reading them.
// The waits are bounded (a rank that died must not hang the others' GPUs): a time-out raises a flag the host reads.
constexpr int kExMaxRanks = 16;
struct ExArgs {
    uint8_t* base[kExMaxRanks];  // the block of every rank, as mapped into THIS process
    int n_ranks, rank;
    long long tile_values, n_values;  // values per tile (a multiple of 4), values of the image
    long long off_fixed, off_rgb, off_flags;  // byte offsets inside a block; off_fixed is that of THIS frame's buffer
    float ns;
    unsigned frame;
};
constexpr long long kExSpinLimit = 1ll << 24;  // x (64 ns sleep + one load): several seconds

__device__ __forceinline__ bool ex_wait_all(const volatile unsigned* flags, int n, unsigned frame) {
    for (int r = 0; r < n; r++) {
        long long spins = 0;
        while ((int)(flags[r] - frame) < 0) {
            if (++spins > kExSpinLimit) return false;
            __nanosleep(64);
        }
    }
    return true;
}

__global__ void __launch_bounds__(256) k_exchange_peers(ExArgs a, unsigned* __restrict__ cta_counter) {
    __shared__ int s_ok;
    unsigned* my_flags = reinterpret_cast<unsigned*>(a.base[a.rank] + a.off_flags);  // [0..15] arrive, [16..31] done, [32] error
    if (threadIdx.x == 0) {
        // every CTA announces the arrival (idempotent): no CTA's wait depends on another CTA having been scheduled
        __threadfence_system();  // this rank's render is complete (stream order) and visible
        for (int r = 0; r < a.n_ranks; r++)
            *reinterpret_cast<volatile unsigned*>(a.base[r] + a.off_flags + 4 * a.rank) = a.frame;
        const bool ok = ex_wait_all(my_flags, a.n_ranks, a.frame);
        if (!ok) atomicExch(my_flags + 32, 1u);
        s_ok = ok ? 1 : 0;
        __threadfence_system();
    }
    __syncthreads();
    if (s_ok) {
        const long long v0 = a.rank * a.tile_values;
        const long long i = v0 + 4 * ((long long)blockIdx.x * blockDim.x + threadIdx.x);
        if (i < v0 + a.tile_values && i < a.n_values) {
            long long s[4] = {0, 0, 0, 0};
            for (int r = 0; r < a.n_ranks; r++) {
                const longlong2* src = reinterpret_cast<const longlong2*>(a.base[r] + a.off_fixed) + (i >> 1);
                const longlong2 x = __ldcv(src), y = __ldcv(src + 1);  // never from a stale cache line
                s[0] += x.x;
                s[1] += x.y;
                s[2] += y.x;
                s[3] += y.y;
            }
            const float k = (float)(1.0 / (double)a.ns);
            unsigned packed = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float sum = (float)((double)s[j] * 4.656612873077392578125e-10);
                const float c = xsqrt(xmul(sum, k));
                int v = (int)(255.99 * (double)c);
                v = v > 255 ? 255 : v;
                v = v < 0 ? 0 : v;
                packed |= (unsigned)v << (8 * j);
            }
            for (int r = 0; r < a.n_ranks; r++) *reinterpret_cast<unsigned*>(a.base[r] + a.off_rgb + i) = packed;
        }
    }
    // DONE: the last CTA of this rank, after every CTA's stores have been fenced
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned n = atomicAdd(cta_counter, 1u);
        if (n == gridDim.x - 1) {
            *cta_counter = 0;
            __threadfence_system();
            for (int r = 0; r < a.n_ranks; r++)
                *reinterpret_cast<volatile unsigned*>(a.base[r] + a.off_flags + 64 + 4 * a.rank) = a.frame;
        }
    }
}

__global__ void k_exchange_wait(unsigned* __restrict__ my_flags, int n_ranks, unsigned frame) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        if (!ex_wait_all(my_flags + 16, n_ranks, frame)) atomicExch(my_flags + 32, 1u);
        __threadfence_system();
    }
}

__global__ void k_merl_lookup(const double* __restrict__ table, const double* __restrict__ angles, long long n,
                              double* __restrict__ rgb) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* a = angles + 4 * i;
    merl_lookup(table, a[0], a[1], a[2], a[3], rgb + 3 * i);
}

// a / b with the hoisted-reciprocal path against the library's __fdiv_rn (tests only)
__global__ void k_check_xdiv(const float* __restrict__ a, const float* __restrict__ b, long long n,
                             unsigned long long* __restrict__ mismatches) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Rcp k = make_rcp(b[i]);
    float q1 = xdiv(a[i], k), q2 = __fdiv_rn(a[i], b[i]);
    if (__float_as_uint(q1) != __float_as_uint(q2) && !(q1 != q1 && q2 != q2)) atomicAdd(mismatches, 1ull);
}

// Measured FP32 roofline denominator (bench.py): 16 independent FMA chains per thread, enough
// resident warps to saturate both FP32 datapaths of every SM sub-partition.  2 flops per FMA.
__global__ void __launch_bounds__(256) k_fp32_peak(float* __restrict__ out, int iters, float seed) {
    float a[16];
#pragma unroll
    for (int k = 0; k < 16; k++) a[k] = seed + (float)(threadIdx.x + k);
    const float m = 1.0000001f, c = 1e-7f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) a[k] = fmaf(a[k], m, c);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 16; k++) s += a[k];
    if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;  // never true: keeps the chains alive
}

// Measured L2 roofline denominator (bench.py): every CTA streams the whole buffer (L2-resident: 32 MiB) with
// 128-bit loads, `passes` times, each CTA starting at a different offset so that the SMs do not march
// through the slices in lockstep.  n_vec must be a multiple of blockDim.x * 4.
__global__ void __launch_bounds__(512) k_l2_peak(const float4* __restrict__ buf, long long n_vec, int passes,
                                                 float* __restrict__ out) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const long long stride = (long long)blockDim.x;
    const long long start = ((long long)blockIdx.x * 7919 * stride * 4) % n_vec;
    for (int p = 0; p < passes; p++) {
        long long i = start + threadIdx.x;
        for (long long k = 0; k < n_vec; k += 4 * stride) {
            float4 a, b, c, d;
            // ld.global.cg: cache at L2 only (the point is the L2 -> SM path, not L1 hits)
            long long i0 = i, i1 = i + stride, i2 = i + 2 * stride, i3 = i + 3 * stride;
            i0 -= i0 >= n_vec ? n_vec : 0;
            i1 -= i1 >= n_vec ? n_vec : 0;
            i2 -= i2 >= n_vec ? n_vec : 0;
            i3 -= i3 >= n_vec ? n_vec : 0;
            a = __ldcg(buf + i0);
            b = __ldcg(buf + i1);
            c = __ldcg(buf + i2);
            d = __ldcg(buf + i3);
            acc.x += a.x + b.x + c.x + d.x;
            acc.y += a.y + b.y + c.y + d.y;
            acc.z += a.z + b.z + c.z + d.z;
            acc.w += a.w + b.w + c.w + d.w;
            i += 4 * stride;
            i -= i >= n_vec ? n_vec : 0;
        }
    }
    if (acc.x + acc.y + acc.z + acc.w == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x;
}

}  // namespace srt
#endif
