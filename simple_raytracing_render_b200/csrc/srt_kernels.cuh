// srt_kernels.cuh — the __global__ entry points.
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
    int lanes_per_pixel;  // power of two <= 32
    int n_tiles;          // warp tiles of 32 / lanes_per_pixel pixels
};

// Closest hit for a batch of rays: world->hit(r, 0.001, FLT_MAX, rec) (Raytracing_n.cpp:58).
__global__ void __launch_bounds__(128) k_trace_closest(DScene S, const float* __restrict__ rays, long long n,
                                                       int32_t* __restrict__ prim, float* __restrict__ t,
                                                       float* __restrict__ rec, uint32_t seed) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* q = rays + 7 * i;
    Ray r{v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), q[6]};
    Philox rng;
    philox_init(rng, seed, (uint32_t)i, (uint32_t)(i >> 32));
    Tally<false> tl{nullptr};
    Hit h = world_hit(S, r, 0.001f, kFltMax, rng, tl);
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
        HitRecord hr = make_record<true>(S, r, h);
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

// hitable_list::pdf_value over the light list (hitable_list.h:54-62) for (origin, direction) pairs.
__global__ void __launch_bounds__(128) k_light_pdf(DScene S, const float* __restrict__ od, long long n,
                                                   float* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* q = od + 6 * i;
    Tally<false> tl{nullptr};
    out[i] = S.n_lights > 0 ? light_pdf_value(S, v3(q[0], q[1], q[2]), v3(q[3], q[4], q[5]), tl) : 0.f;
}

// The sample loop of renderthread (Raytracing_n.cpp:826-848) for every pixel, as a persistent grid:
// each warp takes tiles of 32 / L consecutive pixels; the L lanes of a pixel split its Sobol
// indices [s_begin, s_end) (lane l takes s_begin + l, + L, ...), so a warp's primary rays are
// nearly identical (coherent traversal) and a pixel's sum is reduced by a fixed shuffle tree:
// bit-reproducible, no atomics.  accum[curpixel] += sum of de_nan(color(...)).
template <bool COUNT>
__global__ void __launch_bounds__(256) k_render(DScene S, RenderArgs A, const uint32_t* __restrict__ sobol,
                                               float* __restrict__ accum, unsigned long long* __restrict__ counters) {
    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x >> 5;
    const int warp_global = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
    const int total_warps = gridDim.x * warps_per_block;
    const int L = A.lanes_per_pixel;
    const int ppw = 32 / L;
    const int n_pixels = A.nx * A.ny;
    Counters cnt = {0, 0, 0, 0, 0, 0};
    Tally<COUNT> tl{&cnt};
    unsigned long long paths = 0;

    for (int tile = warp_global; tile < A.n_tiles; tile += total_warps) {
        int pixel = tile * ppw + lane / L;
        int sl = lane % L;
        V3 col = v3(0.f, 0.f, 0.f);
        if (pixel < n_pixels) {
            int i = pixel % (A.mod_ny ? A.ny : A.nx);  // Raytracing_n.cpp:827-828
            int j = A.ny - 1 - (pixel / A.nx);
            for (int s = A.s_begin + sl; s < A.s_end; s += L) {
                // float u = float(samplepoints[s][0] + i) / float(nx)   (Raytracing_n.cpp:834-835)
                double sx = (double)__ldg(&sobol[2 * s]) * 2.3283064365386963e-10;
                double sy = (double)__ldg(&sobol[2 * s + 1]) * 2.3283064365386963e-10;
                float u = xdiv((float)(sx + (double)i), (float)A.nx);
                float v = xdiv((float)(sy + (double)j), (float)A.ny);
                Philox rng;
                philox_init(rng, A.seed, (uint32_t)pixel, (uint32_t)s);
                Ray r = camera_get_ray(S.camera, u, v, rng);
                V3 c = de_nan(trace_path(S, r, A.max_depth, rng, tl));
                col = col + c;
                if (COUNT) paths++;
            }
        }
        // fixed-order tree over the L lanes of the pixel
        for (int off = L >> 1; off > 0; off >>= 1) {
            col.x = xadd(col.x, __shfl_down_sync(0xffffffffu, col.x, off));
            col.y = xadd(col.y, __shfl_down_sync(0xffffffffu, col.y, off));
            col.z = xadd(col.z, __shfl_down_sync(0xffffffffu, col.z, off));
        }
        if (sl == 0 && pixel < n_pixels) {
            float* a = accum + 3 * (size_t)pixel;
            a[0] = xadd(a[0], col.x);
            a[1] = xadd(a[1], col.y);
            a[2] = xadd(a[2], col.z);
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

__global__ void k_merl_lookup(const double* __restrict__ table, const double* __restrict__ angles, long long n,
                              double* __restrict__ rgb) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* a = angles + 4 * i;
    merl_lookup(table, a[0], a[1], a[2], a[3], rgb + 3 * i);
}

}  // namespace srt
#endif
