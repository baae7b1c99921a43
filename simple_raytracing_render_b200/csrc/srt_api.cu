// srt_api.cu — the C ABI of include/srt_ffi.h on top of the CUDA kernels.  No CPU path: every
// compute entry point fails with SRT_E_NODEVICE / SRT_E_CUDA when there is no usable GPU.
#include <algorithm>
#include <thread>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "srt_bvh_build.h"
#include "srt_lbvh.cuh"
#include "srt_ffi.h"
#include "srt_aux_kernels.cuh"
#include "srt_types.h"
#include "srt_variants.h"
#include "srt_wavefront.cuh"

namespace {

thread_local std::string g_error;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_error = buf;
    return code;
}

#define CUDA_TRY(expr)                                                                          \
    do {                                                                                        \
        cudaError_t e_ = (expr);                                                                \
        if (e_ != cudaSuccess) return fail(SRT_E_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
    } while (0)

}  // namespace

struct SrtContext {
    int device = 0;
    cudaStream_t stream = nullptr;
    int sm_count = 0;
    unsigned long long launches = 0;
    SrtCounters last = {};
    unsigned long long* d_counters = nullptr;  // 16 x u64: [0..6] SrtCounters of a counted replay, [7] work counter,
                                               // [8] wide-BVH stack overflows (DScene::diag), never reset
    uint8_t* d_rgb8 = nullptr;                 // resolve staging for host outputs
    size_t rgb8_cap = 0;
    uint8_t* h_rgb8 = nullptr;  // pinned
    size_t h_rgb8_cap = 0;
    // work buffers shared by every scene of the context (a context renders one call at a time):
    // kept across scenes so that re-creating a scene per frame costs no allocation
    uint32_t* d_sobol = nullptr;  // Sobol numerators, [sobol_n][2]
    uint32_t sobol_n = 0;
    float* d_accum = nullptr;  // srt_render_image's internal accumulation buffer
    size_t accum_cap = 0;
    long long* d_fixed = nullptr;  // [n_pixels][3] fixed-point radiance sums of the render in flight (zero between renders)
    size_t fixed_cap = 0;
    bool fixed_dirty = false;
    int variant_blocks[SRT_N_VARIANTS] = {};  // occupancy query results, 0 = not queried yet
    int variant_regs[SRT_N_VARIANTS] = {};
    size_t variant_smem[SRT_N_VARIANTS] = {};  // dynamic shared memory the query was made with
    // wavefront path (srt_wavefront.cuh): the slot pool
    uint8_t* d_wf = nullptr;
    size_t wf_cap = 0;
    WfVariantInfo wf_info[SRT_N_WF_VARIANTS] = {};
    size_t wf_info_smem[SRT_N_WF_VARIANTS] = {};
    // persisting-L2 window over the slot pool (run_wavefront)
    size_t l2_persist_max = 0, l2_window_max = 0;
    const void* l2_window_base = nullptr;
    size_t l2_window_bytes = 0;
    // pinned staging ring of staged_upload (large host -> device copies of srt_scene_create)
    uint8_t* h_stage[2] = {nullptr, nullptr};
    cudaEvent_t stage_ev[2] = {nullptr, nullptr};
};

struct SrtScene {
    SrtContext* ctx = nullptr;
    uint8_t* d_blob = nullptr;
    size_t blob_bytes = 0;
    DScene dev = {};
    SrtSceneInfo info = {};
    std::vector<double*> d_merl;
    std::vector<const double*> merl_ptrs_host;
    int variant = -1;        // render-kernel variant this scene runs on
    int variant_blocks = 0;  // its resident blocks per SM
    int variant_regs = 0;
    size_t variant_smem = 0;
};

namespace {

bool is_device_pointer(const void* p) {
    if (!p) return false;
    cudaPointerAttributes at;
    cudaError_t e = cudaPointerGetAttributes(&at, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

// A device view of `bytes` at p: p itself when it is device memory, else a temporary upload.
struct DevIn {
    const void* ptr = nullptr;
    void* owned = nullptr;
    int init(const void* p, size_t bytes, cudaStream_t st) {
        if (bytes == 0 || is_device_pointer(p)) {
            ptr = p;
            return 0;
        }
        CUDA_TRY(cudaMalloc(&owned, bytes));
        CUDA_TRY(cudaMemcpyAsync(owned, p, bytes, cudaMemcpyHostToDevice, st));
        ptr = owned;
        return 0;
    }
    ~DevIn() {
        if (owned) cudaFree(owned);
    }
};
struct DevOut {
    void* ptr = nullptr;
    void* owned = nullptr;
    void* host = nullptr;
    size_t bytes = 0;
    int init(void* p, size_t n) {
        bytes = n;
        if (n == 0 || p == nullptr || is_device_pointer(p)) {
            ptr = p;
            return 0;
        }
        host = p;
        CUDA_TRY(cudaMalloc(&owned, n));
        ptr = owned;
        return 0;
    }
    int finish(cudaStream_t st) {
        if (owned) {
            CUDA_TRY(cudaMemcpyAsync(host, owned, bytes, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
        }
        return 0;
    }
    ~DevOut() {
        if (owned) cudaFree(owned);
    }
};

struct Layout {
    size_t total = 0;
    size_t add(size_t bytes, size_t align = 256) {
        size_t off = (total + align - 1) & ~(align - 1);
        total = off + bytes;
        return off;
    }
};

void pad_box(float* b) {
    for (int a = 0; a < 3; a++) {
        float m = std::max(std::fabs(b[a]), std::fabs(b[3 + a]));
        float pad = 1e-4f * std::max(m, b[3 + a] - b[a]) + 1e-7f;
        b[a] -= pad;
        b[3 + a] += pad;
    }
}

// Geometric bounds of one primitive in its own frame (for OUR tree only; never decides results).
bool prim_bounds(const SrtSceneDesc& d, const SrtPrim& p, float* b) {
    switch (p.type) {
        case SRT_PRIM_SPHERE:
        case SRT_PRIM_ENV_SPHERE: {
            const float* s = d.spheres + 4 * (size_t)p.index;
            float r = std::fabs(s[3]);
            for (int a = 0; a < 3; a++) {
                b[a] = s[a] - r;
                b[3 + a] = s[a] + r;
            }
            return true;
        }
        case SRT_PRIM_MOVING_SPHERE: {
            const float* m = d.msph + 9 * (size_t)p.index;
            float r = std::fabs(m[8]);
            float times[6] = {0.f, 1.f, m[6], m[7], d.camera.time0, d.camera.time1};
            for (int a = 0; a < 3; a++) {
                b[a] = 3.4e38f;
                b[3 + a] = -3.4e38f;
            }
            for (float tm : times) {
                float k = (tm - m[6]) / (m[7] - m[6]);
                if (!(k == k) || std::isinf(k)) k = 0.f;
                for (int a = 0; a < 3; a++) {
                    float c = m[a] + k * (m[3 + a] - m[a]);
                    b[a] = std::min(b[a], c - r);
                    b[3 + a] = std::max(b[3 + a], c + r);
                }
            }
            return true;
        }
        case SRT_PRIM_XY_RECT:
        case SRT_PRIM_XZ_RECT:
        case SRT_PRIM_YZ_RECT: {
            const float* q = d.rects + 5 * (size_t)p.index;
            int ka = p.type == SRT_PRIM_XY_RECT ? 2 : (p.type == SRT_PRIM_XZ_RECT ? 1 : 0);
            int aa = p.type == SRT_PRIM_YZ_RECT ? 1 : 0;
            int ab = p.type == SRT_PRIM_XY_RECT ? 1 : 2;
            b[aa] = std::min(q[0], q[1]);
            b[3 + aa] = std::max(q[0], q[1]);
            b[ab] = std::min(q[2], q[3]);
            b[3 + ab] = std::max(q[2], q[3]);
            b[ka] = b[3 + ka] = q[4];
            return true;
        }
        case SRT_PRIM_TRIANGLE: {
            const float* t = d.tri_pos + 9 * (size_t)p.index;
            for (int a = 0; a < 3; a++) {
                b[a] = std::min(t[a], std::min(t[3 + a], t[6 + a]));
                b[3 + a] = std::max(t[a], std::max(t[3 + a], t[6 + a]));
            }
            return true;
        }
        default:
            return false;
    }
}

int validate(const SrtSceneDesc& d) {
    if (d.abi_version != SRT_ABI_VERSION) return fail(SRT_E_INVALID, "SrtSceneDesc.abi_version mismatch");
    if (d.n_entries < 0 || d.n_prims < 0 || d.n_groups < 0 || d.n_bvhs < 0) return fail(SRT_E_INVALID, "negative count");
    if (d.n_bvh_items > 0 && !d.bvh_item_boxes) return fail(SRT_E_INVALID, "bvh_item_boxes is required with bvh_items");
    for (int i = 0; i < d.n_prims; i++) {
        const SrtPrim& p = d.prims[i];
        int n = -1;
        switch (p.type) {
            case SRT_PRIM_SPHERE:
            case SRT_PRIM_ENV_SPHERE: n = d.n_spheres; break;
            case SRT_PRIM_MOVING_SPHERE: n = d.n_msph; break;
            case SRT_PRIM_XY_RECT:
            case SRT_PRIM_XZ_RECT:
            case SRT_PRIM_YZ_RECT: n = d.n_rects; break;
            case SRT_PRIM_TRIANGLE: n = d.n_tris; break;
            default: return fail(SRT_E_INVALID, "prims[%d]: unknown type %d", i, p.type);
        }
        if (p.index < 0 || p.index >= n) return fail(SRT_E_INVALID, "prims[%d]: data index out of range", i);
        if (p.material < -1 || p.material >= d.n_materials) return fail(SRT_E_INVALID, "prims[%d]: bad material", i);
    }
    for (int i = 0; i < d.n_groups; i++) {
        const SrtGroup& g = d.groups[i];
        if (g.first_prim < 0 || g.prim_count < 0 || g.first_prim + g.prim_count > d.n_prims)
            return fail(SRT_E_INVALID, "groups[%d]: prim range out of bounds", i);
    }
    for (int i = 0; i < d.n_bvhs; i++) {
        const SrtBvh& b = d.bvhs[i];
        if (b.first_item < 0 || b.item_count < 1 || b.first_item + b.item_count > d.n_bvh_items)
            return fail(SRT_E_INVALID, "bvhs[%d]: item range out of bounds", i);
    }
    for (int i = 0; i < d.n_bvh_items; i++)
        if (d.bvh_items[i] < 0 || d.bvh_items[i] >= d.n_groups) return fail(SRT_E_INVALID, "bvh_items[%d] out of range", i);
    for (int i = 0; i < d.n_entries; i++) {
        const SrtEntry& e = d.entries[i];
        if (e.first_op < 0 || e.op_count < 0 || e.first_op + e.op_count > d.n_ops)
            return fail(SRT_E_INVALID, "entries[%d]: op range out of bounds", i);
        int n = e.geom_kind == SRT_GEOM_GROUP ? d.n_groups : (e.geom_kind == SRT_GEOM_BVH ? d.n_bvhs : -1);
        if (e.geom_index < 0 || e.geom_index >= n) return fail(SRT_E_INVALID, "entries[%d]: bad geometry reference", i);
        if (e.medium < -1 || e.medium >= d.n_media) return fail(SRT_E_INVALID, "entries[%d]: bad medium", i);
    }
    for (int i = 0; i < d.n_media; i++)
        if (d.media[i].phase_material < 0 || d.media[i].phase_material >= d.n_materials)
            return fail(SRT_E_INVALID, "media[%d]: bad phase material", i);
    for (int i = 0; i < d.n_lights; i++) {
        const SrtLight& L = d.lights[i];
        int n = 0;
        switch (L.type) {
            case SRT_PRIM_NONE: continue;
            case SRT_PRIM_XZ_RECT: n = d.n_rects; break;
            case SRT_PRIM_SPHERE:
            case SRT_PRIM_ENV_SPHERE: n = d.n_spheres; break;
            case SRT_PRIM_TRIANGLE: n = d.n_tris; break;
            default: return fail(SRT_E_INVALID, "lights[%d]: type %d has no pdf_value/random in the reference", i, L.type);
        }
        if (L.index < 0 || L.index >= n) return fail(SRT_E_INVALID, "lights[%d]: data index out of range", i);
    }
    for (int i = 0; i < d.n_materials; i++) {
        const SrtMaterial& m = d.materials[i];
        if (m.type < SRT_MAT_LAMBERTIAN || m.type > SRT_MAT_ISOTROPIC) return fail(SRT_E_INVALID, "materials[%d]: unknown type", i);
        if (m.texture < -1 || m.texture >= d.n_textures) return fail(SRT_E_INVALID, "materials[%d]: bad texture", i);
        if (m.type == SRT_MAT_BRDF && ((int)m.p[3] < 0 || (int)m.p[3] >= d.n_merl))
            return fail(SRT_E_INVALID, "materials[%d]: bad MERL table index", i);
    }
    for (int i = 0; i < d.n_textures; i++) {
        const SrtTexture& t = d.textures[i];
        if (t.type == SRT_TEX_CHECKER && (t.a < 0 || t.a >= d.n_textures || t.b < 0 || t.b >= d.n_textures))
            return fail(SRT_E_INVALID, "textures[%d]: bad checker children", i);
        if (t.type == SRT_TEX_IMAGE && (t.a < 0 || t.a >= d.n_images)) return fail(SRT_E_INVALID, "textures[%d]: bad image", i);
        if (t.type == SRT_TEX_NOISE && !d.perlin_ranvec) return fail(SRT_E_INVALID, "noise_texture without perlin tables");
        if (t.type < SRT_TEX_CONSTANT || t.type > SRT_TEX_IMAGE) return fail(SRT_E_INVALID, "textures[%d]: unknown type", i);
    }
    return 0;
}

}  // namespace

// fn(lo, hi) over [0, n) in contiguous chunks, on the host threads when the range is worth a fork
template <class F>
static void parallel_chunks(long long n, F&& fn, long long kMin = 1 << 16) {
    unsigned hw = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
    long long parts = std::min<long long>(hw, n / kMin);
    if (parts <= 1) {
        fn(0ll, n);
        return;
    }
    std::vector<std::thread> th;
    for (long long t = 0; t < parts; t++) th.emplace_back([&, t] { fn(n * t / parts, n * (t + 1) / parts); });
    for (auto& x : th) x.join();
}

// Host -> device copy of a large pageable range: the host threads copy 32 MiB slices into a two-deep pinned ring and
// each slice goes out as one DMA, so a 10 M-triangle scene uploads at host-memcpy speed instead of the driver's
// single-threaded pageable staging.  Small ranges (and SRT_STAGED_UPLOAD=0) take cudaMemcpyAsync directly.
// Either way the source range is consumed when the call returns.
static cudaError_t staged_upload(SrtContext* c, void* dst, const void* src, size_t bytes) {
    constexpr size_t kSlice = (size_t)32 << 20;
    static const bool enabled = [] {
        const char* e = getenv("SRT_STAGED_UPLOAD");
        return !e || atoi(e) != 0;
    }();
    if (bytes < ((size_t)4 << 20) || !enabled) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream);
    for (int i = 0; i < 2; i++) {
        if (!c->h_stage[i]) {
            cudaError_t e = cudaHostAlloc((void**)&c->h_stage[i], kSlice, cudaHostAllocDefault);
            if (e != cudaSuccess) {
                c->h_stage[i] = nullptr;
                (void)cudaGetLastError();
                return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream);
            }
        }
        if (!c->stage_ev[i]) {
            cudaError_t e = cudaEventCreateWithFlags(&c->stage_ev[i], cudaEventDisableTiming);
            if (e != cudaSuccess) return e;
        }
    }
    int turn = 0;
    for (size_t at = 0; at < bytes; at += kSlice, turn ^= 1) {
        const size_t len = std::min(kSlice, bytes - at);
        cudaError_t e = cudaEventSynchronize(c->stage_ev[turn]);  // the DMA that last read this slot
        if (e != cudaSuccess) return e;
        uint8_t* stage = c->h_stage[turn];
        const uint8_t* from = (const uint8_t*)src + at;
        parallel_chunks((long long)len, [&](long long lo, long long hi) { std::memcpy(stage + lo, from + lo, (size_t)(hi - lo)); },
                        (long long)1 << 20);
        e = cudaMemcpyAsync((uint8_t*)dst + at, stage, len, cudaMemcpyHostToDevice, c->stream);
        if (e != cudaSuccess) return e;
        e = cudaEventRecord(c->stage_ev[turn], c->stream);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// ---- device-side reference tree (large BVHs) ---------------------------------------------------------------------------
// The reference bvh_node tree over n leaves in their in-order sequence is a pure function of n: the node at level L,
// position j (the bits of j, most significant first, say left/right) covers [lo, lo + cnt) with cnt halved floor/ceil on
// the way down (bvh.h:96-119), and its depth-first index adds 1 for a left step, 1 + nodes(left size) for a right step.
// A level holds two sizes at most, so nodes() is a two-entry table per level.  One launch per level, deepest first.
struct RefLevelTable {
    int size_lo[32];     // smaller subtree size at this level (the other is size_lo + 1)
    int nodes_lo[32];    // nodes of a subtree of size_lo leaves
    int nodes_hi[32];    // ... of size_lo + 1 leaves
};

__global__ void k_ref_tree_level(RefLevelTable tab, int level, int n, const float* __restrict__ item_boxes, RefNode* __restrict__ out) {
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= (1ll << level)) return;
    int lo = 0, cnt = n, idx = 0;
    for (int l = 0; l < level; l++) {
        if (cnt <= 2) return;  // this position has no node: an ancestor is a leaf pair
        const int nl = cnt / 2;
        if ((j >> (level - 1 - l)) & 1) {
            idx += 1 + (nl == tab.size_lo[l + 1] ? tab.nodes_lo[l + 1] : tab.nodes_hi[l + 1]);
            lo += nl;
            cnt -= nl;
        } else {
            idx += 1;
            cnt = nl;
        }
    }
    RefNode nd;
    nd._pad[0] = nd._pad[1] = nd._pad[2] = 0;
    if (cnt <= 2) {
        // n == 1: left = right = l[0]; n == 2: left = l[0], right = l[1] (bvh.h:104-110)
        const float* a = item_boxes + 6 * (size_t)lo;
        const float* b = item_boxes + 6 * (size_t)(lo + cnt - 1);
        for (int k = 0; k < 3; k++) {
            nd.lo[k] = a[k] < b[k] ? a[k] : b[k];                  // ffmin, aabb.h:8
            nd.hi[k] = a[3 + k] > b[3 + k] ? a[3 + k] : b[3 + k];  // ffmax, aabb.h:9
        }
        nd.left = nd.right = -1;
        nd.mid = lo;
    } else {
        const int nl = cnt / 2;
        const int l = idx + 1, r = idx + 1 + (nl == tab.size_lo[level + 1] ? tab.nodes_lo[level + 1] : tab.nodes_hi[level + 1]);
        const RefNode &L = out[l], &R = out[r];
        for (int k = 0; k < 3; k++) {
            nd.lo[k] = L.lo[k] < R.lo[k] ? L.lo[k] : R.lo[k];
            nd.hi[k] = L.hi[k] > R.hi[k] ? L.hi[k] : R.hi[k];
        }
        nd.left = l;
        nd.right = r;
        nd.mid = lo + nl;
    }
    out[idx] = nd;
}

// DPrim records of a large scene straight from the uploaded SrtPrim rows (same fields as make_dprim below)
__global__ void k_make_dprims(const SrtPrim* __restrict__ prims, int n, const float* __restrict__ spheres,
                              const float* __restrict__ rects, DPrim* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    SrtPrim p = prims[i];
    DPrim o;
    o.g = make_float4(0.f, 0.f, 0.f, 0.f);
    o.k = 0.f;
    o.type = p.type;
    o.material = p.material;
    o.flags_aux = p.flags & SRT_PRIM_FLIP;
    switch (p.type) {
        case SRT_PRIM_SPHERE:
        case SRT_PRIM_ENV_SPHERE: {
            const float* q = spheres + 4 * (size_t)p.index;
            o.g = make_float4(q[0], q[1], q[2], q[3]);
            break;
        }
        case SRT_PRIM_XY_RECT:
        case SRT_PRIM_XZ_RECT:
        case SRT_PRIM_YZ_RECT: {
            const float* q = rects + 5 * (size_t)p.index;
            o.g = make_float4(q[0], q[1], q[2], q[3]);
            o.k = q[4];
            break;
        }
        default:
            o.flags_aux |= (uint32_t)p.index << 1;
            break;
    }
    out[i] = o;
}

extern "C" {

const char* srt_last_error(void) { return g_error.c_str(); }
int srt_abi_version(void) { return SRT_ABI_VERSION; }

int srt_context_create(int device, void* cuda_stream, SrtContext** out) {
    if (!out) return fail(SRT_E_INVALID, "out is null");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(SRT_E_NODEVICE, "no CUDA device (%s); this backend has no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= n) return fail(SRT_E_INVALID, "device %d out of range (%d devices)", device, n);
    CUDA_TRY(cudaSetDevice(device));
    SrtContext* c = new SrtContext();
    c->device = device;
    c->stream = (cudaStream_t)cuda_stream;
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    c->l2_persist_max = (size_t)std::max(0, prop.persistingL2CacheMaxSize);
    c->l2_window_max = (size_t)std::max(0, prop.accessPolicyMaxWindowSize);
    CUDA_TRY(cudaMalloc(&c->d_counters, 16 * sizeof(unsigned long long)));
    CUDA_TRY(cudaMemset(c->d_counters, 0, 16 * sizeof(unsigned long long)));
    *out = c;
    return 0;
}

int srt_context_destroy(SrtContext* ctx) {
    if (!ctx) return 0;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaFree(ctx->d_counters);
    cudaFree(ctx->d_rgb8);
    cudaFree(ctx->d_sobol);
    cudaFree(ctx->d_accum);
    cudaFree(ctx->d_fixed);
    cudaFree(ctx->d_wf);
    if (ctx->h_rgb8) cudaFreeHost(ctx->h_rgb8);
    for (int i = 0; i < 2; i++) {
        if (ctx->h_stage[i]) cudaFreeHost(ctx->h_stage[i]);
        if (ctx->stage_ev[i]) cudaEventDestroy(ctx->stage_ev[i]);
    }
    delete ctx;
    return 0;
}

int srt_context_synchronize(SrtContext* ctx) {
    if (!ctx) return fail(SRT_E_INVALID, "ctx is null");
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int srt_scene_destroy(SrtScene* s) {
    if (!s) return 0;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    cudaFree(s->d_blob);
    for (double* p : s->d_merl) cudaFree(p);
    delete s;
    return 0;
}

// SRT_SCENE_TIMING=1: per-stage host times of srt_scene_create on stderr
static int env_int_early_value(const char* n, int fallback) {
    const char* v = getenv(n);
    return v && *v ? atoi(v) : fallback;
}

struct StageTimer {
    bool on;
    std::chrono::steady_clock::time_point t0;
    StageTimer() : on(env_int_early("SRT_SCENE_TIMING")), t0(std::chrono::steady_clock::now()) {}
    static bool env_int_early(const char* n) {
        const char* v = getenv(n);
        return v && *v && atoi(v) != 0;
    }
    void tick(const char* what) {
        if (!on) return;
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[srt_scene_create] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

int srt_scene_create(SrtContext* ctx, const SrtSceneDesc* desc, SrtScene** out) {
    StageTimer stage;
    if (!ctx || !desc || !out) return fail(SRT_E_INVALID, "null argument");
    *out = nullptr;
    const SrtSceneDesc& d = *desc;
    int rc = validate(d);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(ctx->device));
    auto t0 = std::chrono::steady_clock::now();
    stage.tick("validate");

    // ---- acceleration structures -----------------------------------------------------------
    // Per bvh the three arrays the traversal reads (wide nodes, leaf ranks, leaf geometry) come either from the
    // host builder (vectors) or from the device builder (device buffers): `Seg` lists keep them in bvh order, the
    // blob copy below takes each from where it is.
    struct Seg {
        const void* src;
        size_t bytes;
        bool on_device;
    };
    struct TempDev {  // device scratch of this call: freed on every exit path
        std::vector<void*> ptrs;
        std::vector<srt_lbvh::DeviceBvh> built;
        ~TempDev() {
            for (void* p : ptrs) cudaFree(p);
            for (auto& b : built) b.release();
        }
    } tmpdev;
    std::vector<DBvh> bvhs(d.n_bvhs);
    std::vector<Seg> seg_wide, seg_rank, seg_geom;
    std::vector<std::vector<WideNode4>> keep_wide(d.n_bvhs);
    std::vector<std::vector<int32_t>> keep_rank(d.n_bvhs);
    std::vector<std::vector<float4>> keep_geom(d.n_bvhs);
    std::vector<std::vector<float>> dev_root_box(d.n_bvhs);  // device-built: root bounds (for the entry record)
    std::vector<std::vector<RefNode>> keep_ref(d.n_bvhs);
    std::vector<Seg> seg_ref;
    size_t n_ref = 0;
    size_t n_wide = 0, n_rank = 0, n_geom = 0;
    float top_sah = 0.f;
    int top_items = 0, max_depth = 0, builder_used = 0;
    float* d_tri_tmp = nullptr;  // the triangle positions, uploaded early for the device builder
    const int dev_min = env_int_early_value("SRT_DEVICE_BVH_MIN", 32768);
    const bool dev_ok = env_int_early_value("SRT_DEVICE_BVH", 1) != 0;
    double ms_build_dev = 0.0, ms_build_top = 0.0;
    for (int bi = 0; bi < d.n_bvhs; bi++) {
        const SrtBvh& sb = d.bvhs[bi];
        int n = sb.item_count;
        std::atomic<bool> all_sphere{true}, all_tri{true};
        parallel_chunks(n, [&](long long lo, long long hi) {
            bool as = true, at = true;
            for (long long k = lo; k < hi && (as || at); k++) {
                const SrtGroup& g = d.groups[d.bvh_items[sb.first_item + k]];
                bool single = g.prim_count == 1;
                as = as && single && d.prims[g.first_prim].type == SRT_PRIM_SPHERE;
                at = at && single && d.prims[g.first_prim].type == SRT_PRIM_TRIANGLE;
            }
            if (!as) all_sphere = false;
            if (!at) all_tri = false;
        });
        DBvh& B = bvhs[bi];
        B.kind = all_sphere.load() ? BVH_KIND_SPHERE : (all_tri.load() ? BVH_KIND_TRIANGLE : BVH_KIND_GENERIC);
        B.first_item = sb.first_item;
        B.item_count = n;
        B.node_base = (int32_t)n_wide;
        B.ref_base = (int32_t)n_ref;
        B.leaf_base = (int32_t)n_rank;
        B.geom_base = (int32_t)n_geom;
        B._pad = 0;
        if (dev_ok && n >= dev_min) {
            // the reference tree on the device: boxes up, one launch per level from the leaves to the root
            RefLevelTable tab;
            const int levels = ref_tree_levels(n, tab.size_lo, tab.nodes_lo, tab.nodes_hi);
            const size_t count = (size_t)tab.nodes_lo[0];
            float* d_boxes = nullptr;
            RefNode* d_ref = nullptr;
            CUDA_TRY(cudaMalloc(&d_boxes, sizeof(float) * 6 * (size_t)n));
            tmpdev.ptrs.push_back(d_boxes);
            CUDA_TRY(cudaMalloc(&d_ref, sizeof(RefNode) * count));
            tmpdev.ptrs.push_back(d_ref);
            CUDA_TRY(staged_upload(ctx, d_boxes, d.bvh_item_boxes + 6 * (size_t)sb.first_item, sizeof(float) * 6 * (size_t)n));
            for (int L = levels - 1; L >= 0; L--) {
                const long long width = 1ll << L;
                k_ref_tree_level<<<(unsigned)((width + 255) / 256), 256, 0, ctx->stream>>>(tab, L, n, d_boxes, d_ref);
            }
            CUDA_TRY(cudaGetLastError());
            seg_ref.push_back(Seg{d_ref, sizeof(RefNode) * count, true});
            n_ref += count;
            stage.tick("reference tree (device)");
        } else {
            build_ref_tree(d.bvh_item_boxes + 6 * (size_t)sb.first_item, n, keep_ref[bi]);
            seg_ref.push_back(Seg{keep_ref[bi].data(), sizeof(RefNode) * keep_ref[bi].size(), false});
            n_ref += keep_ref[bi].size();
            stage.tick("build_ref_tree");
        }
        if (dev_ok && B.kind == BVH_KIND_TRIANGLE && n >= dev_min) {
            // ---- device build (srt_lbvh.cuh) --------------------------------------------------------------
            if (!d_tri_tmp) {
                CUDA_TRY(cudaMalloc(&d_tri_tmp, sizeof(float) * 9 * (size_t)d.n_tris));
                tmpdev.ptrs.push_back(d_tri_tmp);
                CUDA_TRY(staged_upload(ctx, d_tri_tmp, d.tri_pos, sizeof(float) * 9 * (size_t)d.n_tris));
            }
            std::unique_ptr<int32_t[]> rows(new int32_t[2 * (size_t)n]);  // first touched by the threads that fill it
            int32_t* item_prim = rows.get() + n;
            parallel_chunks(n, [&](long long lo, long long hi) {
                for (long long k = lo; k < hi; k++) {
                    const int32_t prim = d.groups[d.bvh_items[sb.first_item + k]].first_prim;
                    item_prim[(size_t)k] = prim;
                    rows[(size_t)k] = d.prims[prim].index;
                }
            });
            int32_t* d_rows = nullptr;
            CUDA_TRY(cudaMalloc(&d_rows, sizeof(int32_t) * 2 * (size_t)n));
            tmpdev.ptrs.push_back(d_rows);
            CUDA_TRY(staged_upload(ctx, d_rows, rows.get(), sizeof(int32_t) * 2 * (size_t)n));
            stage.tick("device build: uploads queued");
            srt_lbvh::DeviceBvh db;
            std::string err;
            if (!srt_lbvh::build(ctx->stream, d_tri_tmp, d_rows, d_rows + n, n, db, err)) return fail(SRT_E_CUDA, "%s", err.c_str());
            tmpdev.built.push_back(db);
            stage.tick("device build: srt_lbvh::build");
            if (stage.on)
                fprintf(stderr, "[srt_lbvh::build] n %d: device %.2f ms, host top SAH over %d clusters %.2f ms\n", n, db.ms_device,
                        db.n_clusters, db.ms_host_top);
            ms_build_dev += db.ms_device;
            ms_build_top += db.ms_host_top;
            seg_wide.push_back(Seg{db.nodes, sizeof(WideNode4) * (size_t)db.n_nodes, true});
            seg_rank.push_back(Seg{db.leaf_rank, sizeof(int32_t) * (size_t)n, true});
            seg_geom.push_back(Seg{db.leaf_geom, sizeof(float4) * 3 * (size_t)n, true});
            n_wide += (size_t)db.n_nodes;
            n_rank += (size_t)n;
            n_geom += 3 * (size_t)n;
            max_depth = std::max(max_depth, db.max_depth);
            builder_used = 1;
            // root bounds = union of the four child boxes of wide node 0
            WideNode4 rootn;
            CUDA_TRY(cudaMemcpy(&rootn, db.nodes, sizeof rootn, cudaMemcpyDeviceToHost));
            std::vector<float>& rb = dev_root_box[bi];
            rb.assign({3.4e38f, 3.4e38f, 3.4e38f, -3.4e38f, -3.4e38f, -3.4e38f});
            for (int i = 0; i < 4; i++) {
                if ((&rootn.ref.x)[i] == (int32_t)0x80000000) continue;
                rb[0] = std::min(rb[0], (&rootn.lox.x)[i]);
                rb[1] = std::min(rb[1], (&rootn.loy.x)[i]);
                rb[2] = std::min(rb[2], (&rootn.loz.x)[i]);
                rb[3] = std::max(rb[3], (&rootn.hix.x)[i]);
                rb[4] = std::max(rb[4], (&rootn.hiy.x)[i]);
                rb[5] = std::max(rb[5], (&rootn.hiz.x)[i]);
            }
            if (n > top_items) {
                top_items = n;
                top_sah = db.sah_cost;
            }
            continue;
        }
        // ---- host build (srt_bvh_build.cpp) -----------------------------------------------------------------
        std::vector<float> boxes(6 * (size_t)n);
        for (int k = 0; k < n; k++) {
            const SrtGroup& g = d.groups[d.bvh_items[sb.first_item + k]];
            float* b = &boxes[6 * (size_t)k];
            for (int a = 0; a < 3; a++) {
                b[a] = 3.4e38f;
                b[3 + a] = -3.4e38f;
            }
            for (int i = 0; i < g.prim_count; i++) {
                float pb[6];
                const SrtPrim& p = d.prims[g.first_prim + i];
                if (!prim_bounds(d, p, pb)) return fail(SRT_E_UNSUPPORTED, "primitive type %d inside a bvh", p.type);
                if (p.type == SRT_PRIM_ENV_SPHERE) return fail(SRT_E_UNSUPPORTED, "env_sphere inside a bvh_node");
                for (int a = 0; a < 3; a++) {
                    b[a] = std::min(b[a], pb[a]);
                    b[3 + a] = std::max(b[3 + a], pb[3 + a]);
                }
            }
            if (g.prim_count == 0) {  // an empty list never hits; give it a point box
                for (int a = 0; a < 6; a++) b[a] = 0.f;
            }
            pad_box(b);
        }
        WideBvh wb;
        stage.tick("item boxes");
        build_wide_bvh(boxes.data(), n, wb);
        stage.tick("build_wide_bvh");
        keep_wide[bi].swap(wb.nodes);
        keep_rank[bi].swap(wb.order);
        const std::vector<int32_t>& order = keep_rank[bi];
        std::vector<float4>& geom = keep_geom[bi];
        if (B.kind == BVH_KIND_SPHERE) {
            geom.reserve((size_t)n);
            for (int s = 0; s < n; s++) {
                const SrtPrim& p = d.prims[d.groups[d.bvh_items[sb.first_item + order[s]]].first_prim];
                const float* q = d.spheres + 4 * (size_t)p.index;
                geom.push_back(float4{q[0], q[1], q[2], q[3]});
            }
        } else if (B.kind == BVH_KIND_TRIANGLE) {
            geom.reserve(3 * (size_t)n);
            auto bits = [](int32_t v) {
                float f;
                std::memcpy(&f, &v, sizeof f);
                return f;
            };
            for (int s = 0; s < n; s++) {
                const int32_t prim = d.groups[d.bvh_items[sb.first_item + order[s]]].first_prim;
                const SrtPrim& p = d.prims[prim];
                const float* q = d.tri_pos + 9 * (size_t)p.index;
                // the spare words carry the rank and the primitive id (srt_types.h, DBvh)
                geom.push_back(float4{q[0], q[1], q[2], bits(order[s])});
                geom.push_back(float4{q[3], q[4], q[5], bits(prim)});
                geom.push_back(float4{q[6], q[7], q[8], 0.f});
            }
        }
        seg_wide.push_back(Seg{keep_wide[bi].data(), sizeof(WideNode4) * keep_wide[bi].size(), false});
        seg_rank.push_back(Seg{keep_rank[bi].data(), sizeof(int32_t) * keep_rank[bi].size(), false});
        seg_geom.push_back(Seg{geom.data(), sizeof(float4) * geom.size(), false});
        n_wide += keep_wide[bi].size();
        n_rank += keep_rank[bi].size();
        n_geom += geom.size();
        max_depth = std::max(max_depth, wb.max_depth);
        if (n > top_items) {
            top_items = n;
            top_sah = wb.sah_cost;
        }
    }

    stage.tick("leaf arrays");
    // ---- flat primitive records (index == primitive id) ------------------------------------------
    uint32_t feat = 0;
    auto make_dprim = [&](int type, int index, int material, uint32_t flags) {
        DPrim p;
        std::memset(&p, 0, sizeof p);
        p.type = type;
        p.material = material;
        p.flags_aux = flags & SRT_PRIM_FLIP;
        switch (type) {
            case SRT_PRIM_SPHERE:
            case SRT_PRIM_ENV_SPHERE: {
                const float* q = d.spheres + 4 * (size_t)index;
                p.g = float4{q[0], q[1], q[2], q[3]};
                break;
            }
            case SRT_PRIM_XY_RECT:
            case SRT_PRIM_XZ_RECT:
            case SRT_PRIM_YZ_RECT: {
                const float* q = d.rects + 5 * (size_t)index;
                p.g = float4{q[0], q[1], q[2], q[3]};
                p.k = q[4];
                break;
            }
            default:  // moving sphere / triangle: row of the per-type array
                p.flags_aux |= (uint32_t)index << 1;
                break;
        }
        return p;
    };
    std::vector<char> in_tri_bvh(d.n_prims, 0);
    for (int bi = 0; bi < d.n_bvhs; bi++) {
        if (bvhs[bi].kind != BVH_KIND_TRIANGLE) continue;
        parallel_chunks(d.bvhs[bi].item_count, [&](long long lo, long long hi) {
            for (long long k = lo; k < hi; k++) in_tri_bvh[d.groups[d.bvh_items[d.bvhs[bi].first_item + k]].first_prim] = 1;
        });
        feat |= F_BVH_TRI;
    }
    for (int bi = 0; bi < d.n_bvhs; bi++) {
        if (bvhs[bi].kind == BVH_KIND_SPHERE) feat |= F_BVH_SPHERE;
        if (bvhs[bi].kind == BVH_KIND_GENERIC) feat |= F_BVH_GENERIC;
    }
    // large scenes: the records are written on the device from the uploaded SrtPrim rows (k_make_dprims), the host
    // loop only collects the feature bits
    const bool dprims_on_device = dev_ok && d.n_prims >= dev_min;
    std::vector<DPrim> dprims(dprims_on_device ? 0 : d.n_prims);
    {
        std::atomic<uint32_t> prim_feat{0};
        parallel_chunks(d.n_prims, [&](long long lo, long long hi) {
            uint32_t f = 0;
            for (long long i = lo; i < hi; i++) {
                const SrtPrim& p = d.prims[i];
                if (!dprims_on_device) dprims[(size_t)i] = make_dprim(p.type, p.index, p.material, p.flags);
                if (p.type == SRT_PRIM_TRIANGLE && !in_tri_bvh[(size_t)i]) f |= F_TRI;
                if (p.type == SRT_PRIM_MOVING_SPHERE) f |= F_MSPHERE;
                if (p.type == SRT_PRIM_ENV_SPHERE) f |= F_ENV;
            }
            prim_feat.fetch_or(f);
        });
        feat |= prim_feat.load();
    }
    std::vector<DPrim> dlights(d.n_lights);
    for (int i = 0; i < d.n_lights; i++) {
        const SrtLight& L = d.lights[i];
        if (L.type == SRT_PRIM_NONE) {
            std::memset(&dlights[i], 0, sizeof(DPrim));
            dlights[i].type = SRT_PRIM_NONE;
            dlights[i].material = -1;
            feat |= F_LIGHT_NONE;
            continue;
        }
        dlights[i] = make_dprim(L.type, L.index, -1, 0);
        if (L.type == SRT_PRIM_SPHERE) feat |= F_LIGHT_SPHERE;
        if (L.type == SRT_PRIM_TRIANGLE) feat |= F_LIGHT_TRI;
        if (L.type == SRT_PRIM_ENV_SPHERE) feat |= F_LIGHT_ENV;
    }

    // ---- world entries with conservative local-frame bounds -------------------------------------
    std::vector<DEntry> dentries(d.n_entries);
    for (int i = 0; i < d.n_entries; i++) {
        const SrtEntry& e = d.entries[i];
        DEntry& o = dentries[i];
        std::memset(&o, 0, sizeof o);
        o.first_op = e.first_op;
        o.op_count = e.op_count;
        o.geom_kind = e.geom_kind;
        o.geom_index = e.geom_index;
        o.medium = e.medium;
        if (e.op_count > 0) feat |= F_XFORM;
        if (e.medium >= 0) feat |= F_MEDIA;
        float b[6] = {3.4e38f, 3.4e38f, 3.4e38f, -3.4e38f, -3.4e38f, -3.4e38f};
        bool boundable = true, honours_t = true;
        int n_inside = 0;
        auto add_group = [&](const SrtGroup& g) {
            for (int k = 0; k < g.prim_count; k++) {
                const SrtPrim& p = d.prims[g.first_prim + k];
                float pb[6];
                if (p.type == SRT_PRIM_ENV_SPHERE || !prim_bounds(d, p, pb)) {
                    boundable = false;
                    continue;
                }
                if (p.type == SRT_PRIM_TRIANGLE) honours_t = false;  // triangle::hit ignores (t0, t1)
                pad_box(pb);
                for (int a = 0; a < 3; a++) {
                    b[a] = std::min(b[a], pb[a]);
                    b[3 + a] = std::max(b[3 + a], pb[3 + a]);
                }
                n_inside++;
            }
        };
        if (e.geom_kind == SRT_GEOM_GROUP) {
            const SrtGroup& g = d.groups[e.geom_index];
            o.first_prim = g.first_prim;
            o.prim_count = g.prim_count;
            add_group(g);
        } else if (!dev_root_box[e.geom_index].empty()) {
            // device-built bvh: its root box IS the union of the padded item boxes (same padding, srt_lbvh.cuh)
            const std::vector<float>& rb = dev_root_box[e.geom_index];
            for (int a = 0; a < 6; a++) b[a] = rb[a];
            honours_t = false;  // triangle::hit ignores (t0, t1)
            n_inside = d.bvhs[e.geom_index].item_count;
        } else {
            const SrtBvh& sb = d.bvhs[e.geom_index];
            for (int k = 0; k < sb.item_count; k++) add_group(d.groups[d.bvh_items[sb.first_item + k]]);
        }
        bool finite = n_inside > 0;
        for (int a = 0; a < 6; a++) finite = finite && std::isfinite(b[a]) && std::fabs(b[a]) < 1e30f;
        // a bounds test costs about as much as two primitive tests: only worth it above a few
        if (boundable && finite && (e.geom_kind == SRT_GEOM_BVH || n_inside >= 4)) {
            o.flags |= DENTRY_CULL;
            if (honours_t) o.flags |= DENTRY_TCULL;
            for (int a = 0; a < 3; a++) {
                o.lo[a] = b[a];
                o.hi[a] = b[3 + a];
            }
        }
    }

    // ---- materials with their texture heads folded in ---------------------------------------------
    std::vector<DMaterial> dmats(d.n_materials);
    for (int i = 0; i < d.n_materials; i++) {
        const SrtMaterial& m = d.materials[i];
        DMaterial& o = dmats[i];
        std::memset(&o, 0, sizeof o);
        o.type = m.type;
        o.texture = m.texture;
        o.tex_type = m.texture >= 0 ? d.textures[m.texture].type : -1;
        for (int k = 0; k < 4; k++) o.p[k] = m.p[k];
        if (m.type == SRT_MAT_BRDF || m.type == SRT_MAT_METAL) {
            for (int k = 0; k < 3; k++) o.c[k] = m.p[k];
        } else if (o.tex_type == SRT_TEX_CONSTANT) {
            for (int k = 0; k < 3; k++) o.c[k] = d.textures[m.texture].c[k];
        }
        switch (m.type) {
            case SRT_MAT_ORENNAYAR: feat |= F_ORENNAYAR; break;
            case SRT_MAT_BECKMANN: feat |= F_BECKMANN; break;
            case SRT_MAT_METAL: feat |= F_METAL; break;
            case SRT_MAT_DIELECTRIC: feat |= F_DIELECTRIC; break;
            case SRT_MAT_ISOTROPIC: feat |= F_MEDIA; break;
            default: break;
        }
    }
    for (int i = 0; i < d.n_textures; i++) {
        if (d.textures[i].type == SRT_TEX_CHECKER) feat |= F_TEX_CHECKER;
        if (d.textures[i].type == SRT_TEX_NOISE) feat |= F_TEX_NOISE;
        if (d.textures[i].type == SRT_TEX_IMAGE) feat |= F_TEX_IMAGE;
    }

    std::vector<int32_t> perlin_perm;
    if (d.perlin_ranvec) {
        if (!d.perlin_perm_x || !d.perlin_perm_y || !d.perlin_perm_z) return fail(SRT_E_INVALID, "incomplete perlin tables");
        perlin_perm.insert(perlin_perm.end(), d.perlin_perm_x, d.perlin_perm_x + 256);
        perlin_perm.insert(perlin_perm.end(), d.perlin_perm_y, d.perlin_perm_y + 256);
        perlin_perm.insert(perlin_perm.end(), d.perlin_perm_z, d.perlin_perm_z + 256);
    }

    // ---- one device allocation; the hot tables first and contiguous (16-byte aligned pieces) -----
    struct Piece {
        const void* src;
        size_t bytes;
        size_t off;
        bool on_device;
    };
    std::vector<Piece> pieces;
    Layout lay;
    auto piece = [&](const void* src, size_t bytes, size_t align = 256) -> size_t {
        size_t off = lay.add(bytes, align);
        if (bytes) pieces.push_back(Piece{src, bytes, off, false});
        return off;
    };
    auto seg_pieces = [&](const std::vector<Seg>& segs, size_t total_bytes) -> size_t {
        size_t off = lay.add(total_bytes, 256), at = off;
        for (const Seg& g : segs) {
            if (g.bytes) pieces.push_back(Piece{g.src, g.bytes, at, g.on_device});
            at += g.bytes;
        }
        return off;
    };
    static_assert(sizeof(DEntry) == 64 && sizeof(DPrim) == 32 && sizeof(SrtXformOp) == 16 && sizeof(DMaterial) == 48,
                  "record sizes are part of the device layout");
    size_t o_entries = piece(dentries.data(), sizeof(DEntry) * dentries.size());
    size_t o_ops = piece(d.ops, sizeof(SrtXformOp) * (size_t)d.n_ops, 16);
    size_t o_dprims = dprims_on_device ? lay.add(sizeof(DPrim) * (size_t)d.n_prims, 16)
                                       : piece(dprims.data(), sizeof(DPrim) * dprims.size(), 16);
    size_t o_lights = piece(dlights.data(), sizeof(DPrim) * dlights.size(), 16);
    size_t hot_end = (lay.total + 15) & ~(size_t)15;
    size_t o_groups = piece(d.groups, sizeof(SrtGroup) * (size_t)d.n_groups);
    size_t o_items = piece(d.bvh_items, sizeof(int32_t) * (size_t)d.n_bvh_items);
    size_t o_media = piece(d.media, sizeof(SrtMedium) * (size_t)d.n_media);
    size_t o_bvhs = piece(bvhs.data(), sizeof(DBvh) * bvhs.size());
    size_t o_wide = seg_pieces(seg_wide, sizeof(WideNode4) * n_wide);
    size_t o_ref = seg_pieces(seg_ref, sizeof(RefNode) * n_ref);
    size_t o_rank = seg_pieces(seg_rank, sizeof(int32_t) * n_rank);
    size_t o_geom = seg_pieces(seg_geom, sizeof(float4) * n_geom);
    size_t o_msph = piece(d.msph, sizeof(float) * 9 * (size_t)d.n_msph);
    size_t o_tpos = lay.add(sizeof(float) * 9 * (size_t)d.n_tris, 256);
    if (d.n_tris)  // already on the device when the device builder ran
        pieces.push_back(Piece{d_tri_tmp ? (const void*)d_tri_tmp : (const void*)d.tri_pos, sizeof(float) * 9 * (size_t)d.n_tris,
                               o_tpos, d_tri_tmp != nullptr});
    size_t o_tnrm = piece(d.tri_nrm, sizeof(float) * 9 * (size_t)d.n_tris);
    size_t o_tuv = piece(d.tri_uv, sizeof(float) * 6 * (size_t)d.n_tris);
    size_t o_mats = piece(dmats.data(), sizeof(DMaterial) * dmats.size());
    size_t o_texs = piece(d.textures, sizeof(SrtTexture) * (size_t)d.n_textures);
    size_t o_pran = piece(d.perlin_ranvec, d.perlin_ranvec ? sizeof(float) * 768 : 0);
    size_t o_pperm = piece(perlin_perm.data(), sizeof(int32_t) * perlin_perm.size());
    std::vector<size_t> o_img(d.n_images);
    for (int i = 0; i < d.n_images; i++) {
        if (!d.images[i].data || d.images[i].nx <= 0 || d.images[i].ny <= 0) return fail(SRT_E_INVALID, "images[%d] is empty", i);
        o_img[i] = piece(d.images[i].data, 3 * (size_t)d.images[i].nx * d.images[i].ny);
    }
    size_t o_imgtab = lay.add(sizeof(DImage) * (size_t)d.n_images);
    size_t o_merltab = lay.add(sizeof(double*) * (size_t)d.n_merl);

    SrtScene* s = new SrtScene();
    s->ctx = ctx;
    s->blob_bytes = lay.total + 256;
    stage.tick("records, entries, layout");
    cudaError_t e = cudaMalloc(&s->d_blob, s->blob_bytes);
    if (e != cudaSuccess) {
        delete s;
        return fail(e == cudaErrorMemoryAllocation ? SRT_E_NOMEM : SRT_E_CUDA, "cudaMalloc(%zu): %s", lay.total,
                    cudaGetErrorString(e));
    }
#define SCENE_TRY(expr)                                                           \
    do {                                                                          \
        cudaError_t e2_ = (expr);                                                 \
        if (e2_ != cudaSuccess) {                                                 \
            srt_scene_destroy(s);                                                 \
            return fail(SRT_E_CUDA, "%s: %s", #expr, cudaGetErrorString(e2_));    \
        }                                                                         \
    } while (0)
    SCENE_TRY(cudaMemsetAsync(s->d_blob, 0, s->blob_bytes, ctx->stream));
    for (const Piece& p : pieces) {
        if (p.on_device) SCENE_TRY(cudaMemcpyAsync(s->d_blob + p.off, p.src, p.bytes, cudaMemcpyDeviceToDevice, ctx->stream));
        else SCENE_TRY(staged_upload(ctx, s->d_blob + p.off, p.src, p.bytes));
    }
    if (dprims_on_device) {
        SrtPrim* d_prims_in = nullptr;
        float *d_sph = nullptr, *d_rect = nullptr;
        SCENE_TRY(cudaMalloc(&d_prims_in, sizeof(SrtPrim) * (size_t)d.n_prims));
        tmpdev.ptrs.push_back(d_prims_in);
        SCENE_TRY(cudaMalloc(&d_sph, sizeof(float) * 4 * (size_t)std::max(d.n_spheres, 1)));
        tmpdev.ptrs.push_back(d_sph);
        SCENE_TRY(cudaMalloc(&d_rect, sizeof(float) * 5 * (size_t)std::max(d.n_rects, 1)));
        tmpdev.ptrs.push_back(d_rect);
        SCENE_TRY(staged_upload(ctx, d_prims_in, d.prims, sizeof(SrtPrim) * (size_t)d.n_prims));
        if (d.n_spheres) SCENE_TRY(staged_upload(ctx, d_sph, d.spheres, sizeof(float) * 4 * (size_t)d.n_spheres));
        if (d.n_rects) SCENE_TRY(staged_upload(ctx, d_rect, d.rects, sizeof(float) * 5 * (size_t)d.n_rects));
        k_make_dprims<<<(unsigned)((d.n_prims + 255) / 256), 256, 0, ctx->stream>>>(d_prims_in, d.n_prims, d_sph, d_rect,
                                                                                    (DPrim*)(s->d_blob + o_dprims));
        SCENE_TRY(cudaGetLastError());
    }
    std::vector<DImage> imgtab(d.n_images);
    for (int i = 0; i < d.n_images; i++) imgtab[i] = DImage{s->d_blob + o_img[i], d.images[i].nx, d.images[i].ny};
    if (d.n_images) SCENE_TRY(cudaMemcpyAsync(s->d_blob + o_imgtab, imgtab.data(), sizeof(DImage) * imgtab.size(), cudaMemcpyHostToDevice, ctx->stream));
    for (int i = 0; i < d.n_merl; i++) {
        double* p = nullptr;
        SCENE_TRY(cudaMalloc(&p, sizeof(double) * 3 * 90 * 90 * 180));
        s->d_merl.push_back(p);
        SCENE_TRY(cudaMemcpyAsync(p, d.merl_tables[i], sizeof(double) * 3 * 90 * 90 * 180, cudaMemcpyHostToDevice, ctx->stream));
    }
    stage.tick("cudaMalloc + copies enqueued");
    if (d.n_merl) SCENE_TRY(cudaMemcpyAsync(s->d_blob + o_merltab, s->d_merl.data(), sizeof(double*) * s->d_merl.size(), cudaMemcpyHostToDevice, ctx->stream));
    SCENE_TRY(cudaStreamSynchronize(ctx->stream));
    stage.tick("upload complete");
#undef SCENE_TRY

    DScene& D = s->dev;
    uint8_t* base = s->d_blob;
    D.entries = (const DEntry*)(base + o_entries);
    D.ops = (const SrtXformOp*)(base + o_ops);
    D.dprims = (const DPrim*)(base + o_dprims);
    D.lights = (const DPrim*)(base + o_lights);
    D.groups = (const SrtGroup*)(base + o_groups);
    D.bvh_items = (const int32_t*)(base + o_items);
    D.media = (const SrtMedium*)(base + o_media);
    D.bvhs = (const DBvh*)(base + o_bvhs);
    D.wide_nodes = (const WideNode4*)(base + o_wide);
    D.ref_nodes = (const RefNode*)(base + o_ref);
    D.leaf_rank = (const int32_t*)(base + o_rank);
    D.leaf_geom = (const float4*)(base + o_geom);
    D.msph = (const float*)(base + o_msph);
    D.tri_pos = (const float*)(base + o_tpos);
    D.tri_nrm = (const float*)(base + o_tnrm);
    D.tri_uv = (const float*)(base + o_tuv);
    D.materials = (const DMaterial*)(base + o_mats);
    D.textures = (const SrtTexture*)(base + o_texs);
    D.images = (const DImage*)(base + o_imgtab);
    D.perlin_ranvec = (const float*)(base + o_pran);
    D.perlin_perm = (const int32_t*)(base + o_pperm);
    D.merl = (const double* const*)(base + o_merltab);
    D.n_entries = d.n_entries;
    D.n_prims = d.n_prims;
    D.n_lights = d.n_lights;
    D.n_media = d.n_media;
    D.n_materials = d.n_materials;
    D.n_textures = d.n_textures;
    D.n_bvhs = d.n_bvhs;
    D.n_ops = d.n_ops;
    D.diag = ctx->d_counters + 8;
    D.hot_bytes = (uint32_t)(hot_end - o_entries);
    D.features = feat;
    D.camera = d.camera;
    (void)o_ops;
    (void)o_dprims;
    (void)o_lights;

    auto t1 = std::chrono::steady_clock::now();
    s->info.n_prims = d.n_prims;
    s->info.n_entries = d.n_entries;
    s->info.n_bvhs = d.n_bvhs;
    s->info.n_lights = d.n_lights;
    s->info.bvh_nodes = (int64_t)n_wide;
    s->info.bvh_node_bytes = (int64_t)(n_wide * sizeof(WideNode4));
    s->info.device_bytes = (int64_t)s->blob_bytes + (int64_t)d.n_merl * 8 * 3 * 90 * 90 * 180;
    s->info.sah_cost = top_sah;
    s->info.bvh_max_depth = max_depth;
    s->info.bvh_builder = builder_used;
    (void)ms_build_dev;
    (void)ms_build_top;
    s->info.build_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
    *out = s;
    return 0;
}

int srt_scene_info(const SrtScene* scene, SrtSceneInfo* out) {
    if (!scene || !out) return fail(SRT_E_INVALID, "null argument");
    *out = scene->info;
    out->features = scene->dev.features;
    out->render_variant = scene->variant;
    out->render_registers = scene->variant_regs;
    out->render_blocks_per_sm = scene->variant_blocks;
    out->hot_table_bytes = scene->dev.hot_bytes;
    return 0;
}

int srt_trace_closest(SrtScene* s, const float* rays, int64_t n, int32_t* prim_id, float* t, float* rec) {
    if (!s || (n > 0 && (!rays || !prim_id || !t))) return fail(SRT_E_INVALID, "null argument");
    if (n <= 0) return 0;
    SrtContext* c = s->ctx;
    CUDA_TRY(cudaSetDevice(c->device));
    DevIn in;
    DevOut o_id, o_t, o_rec;
    int rc;
    if ((rc = in.init(rays, sizeof(float) * 7 * (size_t)n, c->stream))) return rc;
    if ((rc = o_id.init(prim_id, sizeof(int32_t) * (size_t)n))) return rc;
    if ((rc = o_t.init(t, sizeof(float) * (size_t)n))) return rc;
    if ((rc = o_rec.init(rec, sizeof(float) * 8 * (size_t)n))) return rc;
    unsigned blocks = (unsigned)((n + 127) / 128);
    // the smallest of the feature classes of srt_variants.h that covers the scene
    const uint32_t need = s->dev.features & F_ALL_SCENE;
    auto kernel = srt::k_trace_closest<F_ALL_SCENE>;
    if ((need & ~(uint32_t)SRT_VM_LIST) == 0)
        kernel = srt::k_trace_closest<SRT_VM_LIST>;
    else if ((need & ~(uint32_t)SRT_VM_SPHERES) == 0)
        kernel = srt::k_trace_closest<SRT_VM_SPHERES>;
    else if ((need & ~(uint32_t)SRT_VM_MESH) == 0)
        kernel = srt::k_trace_closest<SRT_VM_MESH>;
    kernel<<<blocks, 128, 0, c->stream>>>(s->dev, (const float*)in.ptr, (long long)n, (int32_t*)o_id.ptr, (float*)o_t.ptr,
                                          (float*)o_rec.ptr, 0x5EEDu);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    if ((rc = o_id.finish(c->stream))) return rc;
    if ((rc = o_t.finish(c->stream))) return rc;
    if ((rc = o_rec.finish(c->stream))) return rc;
    if (in.owned) CUDA_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

int srt_light_pdf(SrtScene* s, const float* od, int64_t n, float* out) {
    if (!s || (n > 0 && (!od || !out))) return fail(SRT_E_INVALID, "null argument");
    if (n <= 0) return 0;
    SrtContext* c = s->ctx;
    CUDA_TRY(cudaSetDevice(c->device));
    DevIn in;
    DevOut o;
    int rc;
    if ((rc = in.init(od, sizeof(float) * 6 * (size_t)n, c->stream))) return rc;
    if ((rc = o.init(out, sizeof(float) * (size_t)n))) return rc;
    unsigned blocks = (unsigned)((n + 127) / 128);
    srt::k_light_pdf<<<blocks, 128, 0, c->stream>>>(s->dev, (const float*)in.ptr, (long long)n, (float*)o.ptr);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    if ((rc = o.finish(c->stream))) return rc;
    if (in.owned) CUDA_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

int srt_merl_lookup(SrtScene* s, int32_t table, const double* angles, int64_t n, double* rgb) {
    if (!s || (n > 0 && (!angles || !rgb))) return fail(SRT_E_INVALID, "null argument");
    if (table < 0 || table >= (int32_t)s->d_merl.size()) return fail(SRT_E_INVALID, "MERL table %d out of range", table);
    if (n <= 0) return 0;
    SrtContext* c = s->ctx;
    CUDA_TRY(cudaSetDevice(c->device));
    DevIn in;
    DevOut o;
    int rc;
    if ((rc = in.init(angles, sizeof(double) * 4 * (size_t)n, c->stream))) return rc;
    if ((rc = o.init(rgb, sizeof(double) * 3 * (size_t)n))) return rc;
    unsigned blocks = (unsigned)((n + 127) / 128);
    srt::k_merl_lookup<<<blocks, 128, 0, c->stream>>>(s->d_merl[table], (const double*)in.ptr, (long long)n, (double*)o.ptr);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    if ((rc = o.finish(c->stream))) return rc;
    if (in.owned) CUDA_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

int srt_selftest_division(SrtContext* c, const float* a, const float* b, int64_t n, uint64_t* mismatches) {
    if (!c || !mismatches || (n > 0 && (!a || !b))) return fail(SRT_E_INVALID, "null argument");
    *mismatches = 0;
    if (n <= 0) return 0;
    CUDA_TRY(cudaSetDevice(c->device));
    DevIn ia, ib;
    int rc;
    if ((rc = ia.init(a, sizeof(float) * (size_t)n, c->stream))) return rc;
    if ((rc = ib.init(b, sizeof(float) * (size_t)n, c->stream))) return rc;
    CUDA_TRY(cudaMemsetAsync(c->d_counters, 0, sizeof(unsigned long long), c->stream));
    srt::k_check_xdiv<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>((const float*)ia.ptr, (const float*)ib.ptr,
                                                                          (long long)n, c->d_counters);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    unsigned long long h = 0;
    CUDA_TRY(cudaMemcpyAsync(&h, c->d_counters, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    *mismatches = h;
    return 0;
}

int srt_measure_fp32_peak(SrtContext* c, double* tflops) {
    if (!c || !tflops) return fail(SRT_E_INVALID, "null argument");
    CUDA_TRY(cudaSetDevice(c->device));
    float* d = nullptr;
    const int blocks = c->sm_count * 8, iters = 4096;
    CUDA_TRY(cudaMalloc(&d, sizeof(float) * 256 * (size_t)blocks));
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {  // first repetitions warm the clocks up
        CUDA_TRY(cudaEventRecord(e0, c->stream));
        srt::k_fp32_peak<<<blocks, 256, 0, c->stream>>>(d, iters, 1.0f + rep);
        CUDA_TRY(cudaEventRecord(e1, c->stream));
        CUDA_TRY(cudaEventSynchronize(e1));
        c->launches++;
        float ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        double fl = 2.0 * 16.0 * (double)iters * 256.0 * (double)blocks;
        if (ms > 0.f) best = std::max(best, fl / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    CUDA_TRY(cudaGetLastError());
    *tflops = best;
    return 0;
}

int srt_measure_l2_peak(SrtContext* c, double* gbps) {
    if (!c || !gbps) return fail(SRT_E_INVALID, "null argument");
    CUDA_TRY(cudaSetDevice(c->device));
    const size_t bytes = 32u << 20;
    const long long n_vec = (long long)(bytes / sizeof(float4));
    const int threads = 512, blocks = c->sm_count * 2, passes = 4;
    float4* buf = nullptr;
    float* sink = nullptr;
    CUDA_TRY(cudaMalloc(&buf, bytes));
    CUDA_TRY(cudaMalloc(&sink, sizeof(float) * threads * (size_t)blocks));
    CUDA_TRY(cudaMemsetAsync(buf, 0, bytes, c->stream));
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 7; rep++) {  // the first repetition brings the buffer into L2
        CUDA_TRY(cudaEventRecord(e0, c->stream));
        srt::k_l2_peak<<<blocks, threads, 0, c->stream>>>(buf, n_vec, passes, sink);
        CUDA_TRY(cudaEventRecord(e1, c->stream));
        CUDA_TRY(cudaEventSynchronize(e1));
        c->launches++;
        float ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        double moved = (double)bytes * passes * blocks;
        if (rep > 0 && ms > 0.f) best = std::max(best, moved / (ms * 1e-3) / 1e9);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(buf);
    cudaFree(sink);
    CUDA_TRY(cudaGetLastError());
    *gbps = best;
    return 0;
}

int srt_ref_bvh_leaf_order(const float* mins, int64_t n, uint64_t* seed, int32_t* order_out) {
    if (n < 0 || n > 0x7fffffffLL) return fail(SRT_E_INVALID, "bad item count");
    if (n > 0 && (!mins || !order_out)) return fail(SRT_E_INVALID, "null argument");
    if (!seed) return fail(SRT_E_INVALID, "seed is null");
    unsigned long long s = *seed;
    ref_bvh_leaf_order(mins, (long long)n, &s, order_out);
    *seed = s;
    return 0;
}

int srt_ref_tree_boxes(SrtContext* ctx, const float* item_boxes, int32_t n, int on_device, void* nodes_out, int32_t* count_out) {
    if (!ctx || !item_boxes || !nodes_out || !count_out || n < 1) return fail(SRT_E_INVALID, "bad argument");
    if (!on_device) {
        std::vector<RefNode> t;
        build_ref_tree(item_boxes, n, t);
        std::memcpy(nodes_out, t.data(), sizeof(RefNode) * t.size());
        *count_out = (int32_t)t.size();
        return 0;
    }
    CUDA_TRY(cudaSetDevice(ctx->device));
    RefLevelTable tab;
    const int levels = ref_tree_levels(n, tab.size_lo, tab.nodes_lo, tab.nodes_hi);
    const size_t count = (size_t)tab.nodes_lo[0];
    float* d_boxes = nullptr;
    RefNode* d_ref = nullptr;
    CUDA_TRY(cudaMalloc(&d_boxes, sizeof(float) * 6 * (size_t)n));
    cudaError_t e = cudaMalloc(&d_ref, sizeof(RefNode) * count);
    if (e == cudaSuccess) e = staged_upload(ctx, d_boxes, item_boxes, sizeof(float) * 6 * (size_t)n);
    for (int L = levels - 1; L >= 0 && e == cudaSuccess; L--) {
        const long long width = 1ll << L;
        k_ref_tree_level<<<(unsigned)((width + 255) / 256), 256, 0, ctx->stream>>>(tab, L, n, d_boxes, d_ref);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(nodes_out, d_ref, sizeof(RefNode) * count, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_boxes);
    cudaFree(d_ref);
    if (e != cudaSuccess) return fail(SRT_E_CUDA, "srt_ref_tree_boxes: %s", cudaGetErrorString(e));
    *count_out = (int32_t)count;
    return 0;
}

int srt_sobol_points_u32(uint32_t N, uint32_t* out) {
    // Raytracing_n.cpp:721-812 with D = 2: gray-code order, dimension 1 all m = 1, dimension 2 from
    // line 2 of new-joe-kuo-6.21201 ("2 1 0 1": s = 1, a = 0, m1 = 1).
    if (N == 0) return 0;
    if (!out) return fail(SRT_E_INVALID, "out is null");
    uint32_t L = 0;
    while ((1ull << L) < N) L++;
    std::vector<uint32_t> V1(L + 2), V2(L + 2);
    for (uint32_t i = 1; i <= L; i++) V1[i] = 1u << (32 - i);
    const uint32_t sdeg = 1, m1 = 1;
    if (L <= sdeg) {
        for (uint32_t i = 1; i <= L; i++) V2[i] = m1 << (32 - i);
    } else {
        V2[1] = m1 << 31;
        for (uint32_t i = sdeg + 1; i <= L; i++) V2[i] = V2[i - sdeg] ^ (V2[i - sdeg] >> sdeg);
    }
    uint32_t x0 = 0, x1 = 0;
    out[0] = out[1] = 0;
    for (uint32_t i = 1; i < N; i++) {
        uint32_t c = 1, v = i - 1;
        while (v & 1) {
            v >>= 1;
            c++;
        }
        x0 ^= V1[c];
        x1 ^= V2[c];
        out[2 * i] = x0;
        out[2 * i + 1] = x1;
    }
    return 0;
}

}  // extern "C"

namespace {

int ensure_sobol(SrtContext* s, uint32_t n) {
    if (s->sobol_n >= n) return 0;
    uint32_t cap = std::max<uint32_t>(n, 1024);
    std::vector<uint32_t> pts(2 * (size_t)cap);
    srt_sobol_points_u32(cap, pts.data());
    if (s->d_sobol) {
        CUDA_TRY(cudaStreamSynchronize(s->stream));
        CUDA_TRY(cudaFree(s->d_sobol));
        s->d_sobol = nullptr;
        s->sobol_n = 0;
    }
    CUDA_TRY(cudaMalloc(&s->d_sobol, sizeof(uint32_t) * pts.size()));
    CUDA_TRY(cudaMemcpy(s->d_sobol, pts.data(), sizeof(uint32_t) * pts.size(), cudaMemcpyHostToDevice));
    s->sobol_n = cap;
    return 0;
}

int check_params(const SrtRenderParams* p) {
    if (!p) return fail(SRT_E_INVALID, "params is null");
    if (p->nx <= 0 || p->ny <= 0 || (long long)p->nx * p->ny > 0x7fffffffLL / 4) return fail(SRT_E_INVALID, "bad image size");
    if (p->s_begin < 0 || p->s_end < p->s_begin || p->s_end > p->ns) return fail(SRT_E_INVALID, "sample range must satisfy 0 <= s_begin <= s_end <= ns");
    if (p->max_depth < 0) return fail(SRT_E_INVALID, "negative max_depth");
    return 0;
}

const RenderVariantFn kVariants[SRT_N_VARIANTS] = {srt_render_variant_0, srt_render_variant_1, srt_render_variant_2,
                                                    srt_render_variant_3, srt_render_variant_4, srt_render_variant_5,
                                                    srt_render_variant_6, srt_render_variant_7};
constexpr uint32_t kVariantMask[SRT_N_VARIANTS] = {SRT_VARIANT_MASK(0), SRT_VARIANT_MASK(1), SRT_VARIANT_MASK(2),
                                                   SRT_VARIANT_MASK(3), SRT_VARIANT_MASK(4), SRT_VARIANT_MASK(5),
                                                   SRT_VARIANT_MASK(6), SRT_VARIANT_MASK(7)};
constexpr int kVariantCounting[SRT_N_VARIANTS] = {SRT_VARIANT_COUNTING(0), SRT_VARIANT_COUNTING(1), SRT_VARIANT_COUNTING(2),
                                                  SRT_VARIANT_COUNTING(3), SRT_VARIANT_COUNTING(4), SRT_VARIANT_COUNTING(5),
                                                  SRT_VARIANT_COUNTING(6), SRT_VARIANT_COUNTING(7)};
constexpr size_t kMaxHotSmem = 64 * 1024;  // above this the hot tables stay in global memory

int env_int(const char* name, int fallback) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : fallback;
}

// The smallest compiled variant whose mask covers the scene (first match in list order).
int pick_variant(const SrtScene* s, bool counting, bool corrected = false) {
    uint32_t need = (s->dev.features & F_ALL_SCENE) | (corrected ? (uint32_t)F_CORRECTED : 0u);
    bool smem_ok = s->dev.hot_bytes <= kMaxHotSmem && !env_int("SRT_NO_SMEM", 0);
    int forced = env_int("SRT_VARIANT", -1);
    for (int v = 0; v < SRT_N_VARIANTS; v++) {
        if (forced >= 0 && v != forced) continue;
        if (kVariantCounting[v] != (counting ? 1 : 0)) continue;
        if ((kVariantMask[v] & F_SMEM_TABLES) && !smem_ok) continue;
        if ((kVariantMask[v] & need) != need) continue;
        if (((kVariantMask[v] & F_CORRECTED) != 0) != corrected) continue;
        return v;
    }
    return -1;
}

const WfVariantFn kWfVariants[SRT_N_WF_VARIANTS] = {srt_wavefront_variant_0, srt_wavefront_variant_1, srt_wavefront_variant_2,
                                                     srt_wavefront_variant_3, srt_wavefront_variant_4, srt_wavefront_variant_5,
                                                     srt_wavefront_variant_6, srt_wavefront_variant_7};
constexpr uint32_t kWfMask[SRT_N_WF_VARIANTS] = {SRT_WF_VARIANT_MASK(0), SRT_WF_VARIANT_MASK(1), SRT_WF_VARIANT_MASK(2),
                                                 SRT_WF_VARIANT_MASK(3), SRT_WF_VARIANT_MASK(4), SRT_WF_VARIANT_MASK(5),
                                                 SRT_WF_VARIANT_MASK(6), SRT_WF_VARIANT_MASK(7)};
constexpr int kWfCounting[SRT_N_WF_VARIANTS] = {SRT_WF_VARIANT_COUNTING(0), SRT_WF_VARIANT_COUNTING(1),
                                                SRT_WF_VARIANT_COUNTING(2), SRT_WF_VARIANT_COUNTING(3),
                                                SRT_WF_VARIANT_COUNTING(4), SRT_WF_VARIANT_COUNTING(5),
                                                SRT_WF_VARIANT_COUNTING(6), SRT_WF_VARIANT_COUNTING(7)};
constexpr int kWfOrder[SRT_N_WF_VARIANTS] = SRT_WF_VARIANT_ORDER;

int pick_wf_variant(const SrtScene* s, bool counting, bool corrected = false) {
    uint32_t need = (s->dev.features & F_ALL_SCENE) | (corrected ? (uint32_t)F_CORRECTED : 0u);
    bool smem_ok = s->dev.hot_bytes <= kMaxHotSmem && !env_int("SRT_NO_SMEM", 0);
    int forced = counting ? -1 : env_int("SRT_WF_VARIANT", -1);
    for (int k = 0; k < SRT_N_WF_VARIANTS; k++) {
        const int v = kWfOrder[k];
        if (forced >= 0 && v != forced) continue;
        if (kWfCounting[v] != (counting ? 1 : 0)) continue;
        if ((kWfMask[v] & F_SMEM_TABLES) && !smem_ok) continue;
        if ((kWfMask[v] & need) != need) continue;
        if (((kWfMask[v] & F_CORRECTED) != 0) != corrected) continue;
        return v;
    }
    return -1;
}

// The wavefront form of one render (srt_wavefront.cuh): ONE launch of the persistent sorted kernel, one CTA
// per SM (two for some variants and scenes, below), every CTA owning per_thread x threads slots of the pool.
int run_wavefront(SrtScene* s, const srt::RenderArgs& A, long long* out, int v) {
    SrtContext* c = s->ctx;
    const size_t smem = (kWfMask[v] & F_SMEM_TABLES) ? s->dev.hot_bytes : 0;
    WfLaunch L = {};
    L.scene = &s->dev;
    L.args = &A;
    L.sobol = c->d_sobol;
    L.out = out;
    L.counters = c->d_counters;
    L.stream = c->stream;
    // k_wf_sorted keeps its index queues behind the hot tables: Q_N x kSortCap 16-bit entries at most
    const size_t smem_hot16 = (smem + 15) & ~(size_t)15;
    const size_t smem_max = smem_hot16 + (size_t)srt::Q_N * srt::kSortCap * sizeof(unsigned short);
    if (c->wf_info[v].blocks_per_sm == 0 || smem > c->wf_info_smem[v]) {
        L.smem_bytes = smem_max;  // opt-in ceiling
        cudaError_t e = kWfVariants[v](L, &c->wf_info[v]);
        if (e != cudaSuccess) return fail(SRT_E_CUDA, "wavefront variant %d: %s", v, cudaGetErrorString(e));
        c->wf_info_smem[v] = smem;
    }
    const WfVariantInfo& info = c->wf_info[v];
    const int slots_per_sm = std::max(256, env_int("SRT_WF_SLOTS_PER_SM", 6144)) & ~31;
    // CTAs per SM: several smaller CTAs share an SM's slots and one fills the other's barrier waits.  Fixed per variant
    // (srt_variants.h: SRT_WF_VARIANT_CTAS_PER_SM), except that variant 6 splits its CTA in two for scenes with large BVHs
    // (SRT_WF_SPLIT_CTA_MIN_NODES).  SRT_WF_CTAS_PER_SM overrides for experiments.  A launch may use fewer threads than
    // the variant's launch bound, never more; a CTA that does not fit beside the others simply runs after them.
    int ctas = std::max(1, info.ctas_per_sm);
    if (v == 6 && s->info.bvh_nodes >= SRT_WF_SPLIT_CTA_MIN_NODES) ctas = 2;
    ctas = std::max(1, std::min(8, env_int("SRT_WF_CTAS_PER_SM", ctas)));
    const int threads_per_sm = std::max(32, info.threads) * std::max(1, info.ctas_per_sm);
    const int threads = std::max(32, std::min(std::max(32, info.threads), (threads_per_sm / ctas) & ~31));
    const int per_thread = std::max(1, std::min(slots_per_sm / ctas / threads, srt::kSortCap / threads));
    const int grid = c->sm_count * ctas;
    const int n_slots = grid * per_thread * threads;
    const size_t need = (size_t)n_slots * (size_t)info.slot_bytes;
    if (c->wf_cap < need) {
        CUDA_TRY(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_wf);
        c->d_wf = nullptr;
        c->wf_cap = 0;
        CUDA_TRY(cudaMalloc(&c->d_wf, need));
        c->wf_cap = need;
    }
    // Persisting-L2 window over the slot pool.  The pool is rewritten a few hundred times per frame and is
    // sized to live in L2; without a policy L2 evicts (and writes back) part of it all frame long — measured
    // 6 GB (C2) to 200 GB (C5-like) of DRAM writes per frame that no algorithm needs (profiles/r01_notes.md,
    // captures L, N).  With hitProp = persisting the pool's lines stay resident in the set-aside part of L2;
    // everything else (scene, image) uses the rest as before.  SRT_L2_PERSIST=0 switches it off (A/B).
    if (env_int("SRT_L2_PERSIST", 1) && c->l2_persist_max > 0 && c->l2_window_max > 0 &&
        (c->l2_window_base != c->d_wf || c->l2_window_bytes != need)) {
        const size_t window = std::min(need, c->l2_window_max);
        const size_t setaside = std::min(window, c->l2_persist_max);
        // best effort: a stream or driver that refuses the policy renders exactly as before
        cudaError_t pe = cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, setaside);
        cudaStreamAttrValue at = {};
        at.accessPolicyWindow.base_ptr = c->d_wf;
        at.accessPolicyWindow.num_bytes = window;
        at.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)setaside / (double)window);
        at.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        at.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        if (pe == cudaSuccess) pe = cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &at);
        if (pe != cudaSuccess) cudaGetLastError();
        c->l2_window_base = c->d_wf;
        c->l2_window_bytes = need;
    }
    srt::WfState st;
    st.rec = reinterpret_cast<float4*>(c->d_wf);
    st.n_slots = n_slots;
    L.state = &st;
    L.per_thread = per_thread;
    L.threads = threads;
    L.smem_bytes = smem_hot16 + (size_t)srt::Q_N * per_thread * threads * sizeof(unsigned short);
    L.grid = grid;
    c->launches++;
    cudaError_t pe = kWfVariants[v](L, nullptr);
    if (pe != cudaSuccess) return fail(SRT_E_CUDA, "k_wf_sorted (variant %d): %s", v, cudaGetErrorString(pe));
#ifdef SRT_WFS_PHASE_CLOCKS
    {  // debug build (tools/build_alt.sh clocks -DSRT_WFS_PHASE_CLOCKS): where thread 0 of the CTAs spent its clocks
        unsigned long long t[6] = {0, 0, 0, 0, 0, 0};
        CUDA_TRY(cudaMemcpyAsync(t, c->d_counters + 10, sizeof t, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(cudaMemsetAsync(c->d_counters + 10, 0, sizeof t, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
        const double tot = (double)(t[0] + t[1] + t[2]);
        fprintf(stderr, "[k_wf_sorted phases] regenerate %.1f %%, trace %.1f %%, shade %.1f %%; %.1f iterations per CTA, %.3f Mclk per CTA\n",
                100.0 * t[0] / tot, 100.0 * t[1] / tot, 100.0 * t[2] / tot, (double)t[3] / grid, tot / grid / 1e6);
        const double warps = (double)(threads / 32);
        fprintf(stderr, "[k_wf_sorted phases] warps busy: %.1f %% of the trace phase, %.1f %% of the shade phase\n",
                100.0 * t[4] / (warps * t[1]), 100.0 * t[5] / (warps * t[2]));
    }
#endif
    return 0;
}

// Wavefront by default when the scene has an acceleration structure and the frame is a few paths per
// slot of the pool: measured on B200 (tools/chunk_check.py) the sorted wavefront wins from 1.9 M paths
// up on C2 / C3 (1.44 vs 1.96 ms at 1200x800 @ 2), which is also what a rank of an 8-GPU C2 frame renders;
// scenes that are a plain list of primitives (C1) stay on the fused kernel at every size (47 vs 72 ms
// at 2048x2048 @ 64).
bool auto_wavefront(const SrtScene* s, const SrtRenderParams* p) {
    long long paths = (long long)p->nx * p->ny * (p->s_end - p->s_begin);
    return s->dev.n_bvhs > 0 && paths >= (1LL << 20);
}

// accum != null: the radiance sums are ADDED to the caller's float buffer (through the context's fixed-point
// buffer and k_fixed_to_accum).  fixed_out != null: the kernels add their 2^-31 fixed-point sums straight
// into the caller's int64 buffer and nothing else is launched.
int launch_render(SrtScene* s, const SrtRenderParams* p, float* accum, bool counting, long long* fixed_out = nullptr) {
    SrtContext* c = s->ctx;
    int count = p->s_end - p->s_begin;
    if (count == 0) return 0;
    int rc = ensure_sobol(c, (uint32_t)p->s_end);
    if (rc) return rc;
    // scheduler: the fused persistent-warp kernel (srt_kernels.cuh) or the two-stage wavefront
    // (srt_wavefront.cuh).  SRT_RENDER_WAVEFRONT asks for the latter; without the flag the backend picks
    // (DESIGN.md 5.1): scenes made of a handful of primitives stay on the fused kernel, whose whole
    // frame is shorter than the wavefront's ramp-up and tail.
    int wf_env = env_int("SRT_WAVEFRONT", -1);
    // a counted replay runs on the instrumented build of the scheduler the plain call would pick
    bool wavefront = wf_env >= 0 ? wf_env != 0 : ((p->flags & SRT_RENDER_WAVEFRONT) != 0 || auto_wavefront(s, p));
    // corrected-transport mode has no instrumented build: a counted replay of it is refused below
    const bool corrected = (p->flags & SRT_RENDER_CORRECTED) != 0;
    if (corrected && counting) return fail(SRT_E_UNSUPPORTED, "srt_render_counted: no instrumented build of the corrected-transport mode");
    // the estimator needs a real density for every shape it samples: env_sphere::pdf_value is the constant 1
    // (env_sphere.h:48-50) and shapes without pdf_value / random return 0 / (1,0,0) (hitable.h:31-32)
    if (corrected && (s->dev.features & (F_LIGHT_ENV | F_LIGHT_NONE)))
        return fail(SRT_E_UNSUPPORTED, "corrected-transport mode: the light list holds a shape without a sampling density "
                                       "(env_sphere, or a class without pdf_value / random)");
    int wv = wavefront ? pick_wf_variant(s, counting, corrected) : -1;
    if (wv < 0) wavefront = false;
    // the queues of k_wf_sorted hold one path per work item: SRT_CHUNK > 1 renders on the fused kernel
    if (wavefront && std::min(count, env_int("SRT_CHUNK", 1)) > 1) wavefront = false;
    const int wf_mode = wavefront ? 3 : 0;  // SrtSceneInfo::render_variant = 100 x mode + variant (3: k_wf_sorted)
    int v = pick_variant(s, counting, corrected);
    if (v < 0) return fail(SRT_E_UNSUPPORTED, "no render-kernel variant covers feature mask 0x%x", s->dev.features);
    size_t smem = (kVariantMask[v] & F_SMEM_TABLES) ? s->dev.hot_bytes : 0;

    RenderLaunch L;
    L.scene = &s->dev;
    L.sobol = c->d_sobol;
    L.counters = c->d_counters;
    L.smem_bytes = smem;
    L.stream = c->stream;
    L.grid = 0;
    RenderVariantInfo info = {};
    if (wavefront) {
        info.blocks_per_sm = 1;
    } else if (c->variant_blocks[v] > 0 && smem <= c->variant_smem[v]) {
        info.blocks_per_sm = c->variant_blocks[v];
        info.registers = c->variant_regs[v];
    } else {
        L.args = nullptr;
        L.out = nullptr;
        cudaError_t e = kVariants[v](L, &info);
        if (e != cudaSuccess) return fail(SRT_E_CUDA, "render variant %d: %s", v, cudaGetErrorString(e));
        c->variant_blocks[v] = info.blocks_per_sm;
        c->variant_regs[v] = info.registers;
        c->variant_smem[v] = smem;
    }
    if (!counting && !wavefront) {
        s->variant = v;
        s->variant_blocks = info.blocks_per_sm;
        s->variant_regs = info.registers;
        s->variant_smem = smem;
    }
    if (wavefront && !counting) s->variant = 100 * wf_mode + wv;
    int blocks_per_sm = std::max(1, info.blocks_per_sm);

    // work decomposition (srt_kernels.cuh): an item is `chunk` consecutive Sobol indices of one pixel,
    // numbered pixel-major.  One path per item by default: the lanes of a warp that start paths
    // together then start them in the same pixel, and the drain at the end of the frame is one path
    // long (profiles/r01_notes.md, capture F).
    long long n_pixels = (long long)p->nx * p->ny;
    int chunk = std::max(1, std::min(count, env_int("SRT_CHUNK", 1)));
    int n_chunks = (count + chunk - 1) / chunk;

    srt::RenderArgs A;
    A.nx = p->nx;
    A.ny = p->ny;
    A.s_begin = p->s_begin;
    A.s_end = p->s_end;
    A.max_depth = p->max_depth;
    A.seed = p->seed;
    A.mod_ny = (p->flags & SRT_COMPAT_PIXEL_MOD_NY) ? 1 : 0;
    A.chunk = chunk;
    A.n_chunks = n_chunks;
    A.n_items = (unsigned long long)n_pixels * (unsigned long long)n_chunks;
    // per-sample clamp such that a whole pixel (all ns samples, over every shard) stays inside int64
    A.fx_clamp = 4.0e9f / (float)std::max(count, p->ns);
    A.regen_threshold = std::max(1, std::min(32, env_int("SRT_REGEN", 8)));
    A.grab = std::max(32, env_int("SRT_GRAB", 128));
    A.cta_sync = std::max(0, std::min(2, env_int("SRT_CTA_SYNC", 0)));
    A.dbg_pixel = counting ? env_int("SRT_DEBUG_PIXEL", -1) : -1;
    A.dbg_sample = counting ? env_int("SRT_DEBUG_SAMPLE", -1) : -1;

    // the kernels accumulate into the context's fixed-point buffer; k_fixed_to_accum adds it to the
    // caller's float buffer and leaves it zero for the next render
    if (!fixed_out) {
        size_t need = sizeof(long long) * 3 * (size_t)n_pixels;
        if (c->fixed_cap < need) {
            CUDA_TRY(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_fixed);
            c->d_fixed = nullptr;
            c->fixed_cap = 0;
            CUDA_TRY(cudaMalloc(&c->d_fixed, need));
            c->fixed_cap = need;
            c->fixed_dirty = true;
        }
        if (c->fixed_dirty) CUDA_TRY(cudaMemsetAsync(c->d_fixed, 0, c->fixed_cap, c->stream));
        c->fixed_dirty = true;
    }
    long long* out = fixed_out ? fixed_out : c->d_fixed;
    auto finish = [&]() -> int {
        if (fixed_out) return 0;
        long long nv = 3 * n_pixels;
        srt::k_fixed_to_accum<<<(unsigned)((nv + 255) / 256), 256, 0, c->stream>>>(c->d_fixed, nv, accum);
        c->launches++;
        CUDA_TRY(cudaGetLastError());
        c->fixed_dirty = false;
        return 0;
    };
    // counters[7] is the work counter
    CUDA_TRY(cudaMemsetAsync(c->d_counters + 7, 0, sizeof(unsigned long long), c->stream));
    if (wavefront) {
        int rc2 = run_wavefront(s, A, out, wv);
        if (rc2) return rc2;
        return finish();
    }
    long long warps_needed = ((long long)A.n_items + 31) / 32;
    long long blocks_needed = (warps_needed + (srt::kRenderThreads / 32) - 1) / (srt::kRenderThreads / 32);
    L.args = &A;
    L.out = out;
    L.grid = (int)std::max<long long>(1, std::min<long long>(blocks_needed, (long long)c->sm_count * blocks_per_sm));
    cudaError_t e = kVariants[v](L, nullptr);
    c->launches++;
    if (e != cudaSuccess) return fail(SRT_E_CUDA, "k_render (variant %d): %s", v, cudaGetErrorString(e));
    return finish();
}

// ---- srt_render_image_static: the fused kernel on scratch that is part of the library image ----------------------
// k_render's scratch — the [n_pixels][3] fixed-point radiance sums it accumulates into and the Sobol numerators
// it reads — lives in two file-scope device arrays instead of cudaMalloc'ed buffers of the context, so that the
// call allocates nothing.  Every GPU that loads the library holds them (6 MiB + 8 KiB) until the process exits.
// In host code the names below are host-side stand-ins, NOT device addresses: the addresses the kernels get
// come from cudaGetSymbolAddress, looked up once per device with that device current (static_scratch).
static __device__ __align__(256) long long g_static_fixed[3 * (size_t)SRT_STATIC_MAX_PIXELS];
static __device__ __align__(256) uint32_t g_static_sobol[2 * (size_t)SRT_STATIC_MAX_SAMPLES];

struct StaticScratch {
    long long* fixed = nullptr;  // device address of g_static_fixed on this device
    uint32_t* sobol = nullptr;   // device address of g_static_sobol on this device, filled
    bool ready = false;
};
constexpr int kStaticMaxDevices = 64;
StaticScratch g_static_dev[kStaticMaxDevices];
std::mutex g_static_mutex;

// Set-up of the static scratch of the CURRENT device (the caller has made c->device current): symbol look-up and
// the one upload of the Sobol table (the same numerators ensure_sobol uploads: the sequence is a prefix
// property of the generator, srt_sobol_points_u32).  Synchronous; runs on the first call per device only.
int static_scratch(SrtContext* c, StaticScratch** out) {
    if (c->device < 0 || c->device >= kStaticMaxDevices) return fail(SRT_E_INVALID, "device ordinal %d beyond %d", c->device, kStaticMaxDevices);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    StaticScratch& S = g_static_dev[c->device];
    if (!S.ready) {
        long long* fixed = nullptr;
        uint32_t* sobol = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void**)&fixed, g_static_fixed));
        CUDA_TRY(cudaGetSymbolAddress((void**)&sobol, g_static_sobol));
        if (!fixed || !sobol) return fail(SRT_E_CUDA, "cudaGetSymbolAddress returned a null address");
        // the numerators of srt_sobol_points_u32 (Raytracing_n.cpp:721-812, D = 2) with fixed-size direction tables:
        // no heap allocation on this path either.  A prefix of the table ensure_sobol uploads.
        static uint32_t pts[2 * (size_t)SRT_STATIC_MAX_SAMPLES];  // under g_static_mutex
        uint32_t V1[34], V2[34];
        uint32_t L = 0;
        while ((1ull << L) < (unsigned long long)SRT_STATIC_MAX_SAMPLES) L++;
        static_assert(SRT_STATIC_MAX_SAMPLES > 2 && SRT_STATIC_MAX_SAMPLES <= (1u << 31), "direction tables hold 32 bits");
        for (uint32_t i = 1; i <= L; i++) V1[i] = 1u << (32 - i);
        V2[1] = 1u << 31;
        for (uint32_t i = 2; i <= L; i++) V2[i] = V2[i - 1] ^ (V2[i - 1] >> 1);
        uint32_t x0 = 0, x1 = 0;
        pts[0] = pts[1] = 0;
        for (uint32_t i = 1; i < (uint32_t)SRT_STATIC_MAX_SAMPLES; i++) {
            uint32_t cbit = 1, v = i - 1;
            while (v & 1) {
                v >>= 1;
                cbit++;
            }
            x0 ^= V1[cbit];
            x1 ^= V2[cbit];
            pts[2 * i] = x0;
            pts[2 * i + 1] = x1;
        }
        CUDA_TRY(cudaMemcpy(sobol, pts, sizeof pts, cudaMemcpyHostToDevice));
        S.fixed = fixed;
        S.sobol = sobol;
        S.ready = true;
    }
    *out = &S;
    return 0;
}

// One frame on the fused kernel with the static scratch, then k_resolve_fixed into the caller's device buffer.
// Same variant pick, same RenderArgs and same grid as launch_render's fused branch: the two produce the same
// fixed-point sums (integer adds, any order), hence the same bytes.
int launch_render_static(SrtScene* s, const SrtRenderParams* p, uint8_t* rgb8_device) {
    SrtContext* c = s->ctx;
    const long long n_pixels = (long long)p->nx * p->ny;
    const int count = p->s_end - p->s_begin;
    if (n_pixels > (long long)SRT_STATIC_MAX_PIXELS)
        return fail(SRT_E_INVALID, "srt_render_image_static: %d x %d = %lld pixels exceed the static scratch of %d pixels "
                                   "(SRT_STATIC_MAX_PIXELS); use srt_render_image",
                    p->nx, p->ny, n_pixels, (int)SRT_STATIC_MAX_PIXELS);
    if (p->s_end > (int)SRT_STATIC_MAX_SAMPLES)
        return fail(SRT_E_INVALID, "srt_render_image_static: sample index %d exceeds the static Sobol table of %d points "
                                   "(SRT_STATIC_MAX_SAMPLES); use srt_render_image",
                    p->s_end, (int)SRT_STATIC_MAX_SAMPLES);
    if (count <= 0) return fail(SRT_E_INVALID, "srt_render_image_static: empty sample range");
    if (!is_device_pointer(rgb8_device)) return fail(SRT_E_INVALID, "srt_render_image_static: rgb8_device is not device memory");
    const bool corrected = (p->flags & SRT_RENDER_CORRECTED) != 0;
    if (corrected && (s->dev.features & (F_LIGHT_ENV | F_LIGHT_NONE)))
        return fail(SRT_E_UNSUPPORTED, "corrected-transport mode: the light list holds a shape without a sampling density "
                                       "(env_sphere, or a class without pdf_value / random)");
    StaticScratch* S = nullptr;
    int rc = static_scratch(c, &S);
    if (rc) return rc;

    const int v = pick_variant(s, false, corrected);
    if (v < 0) return fail(SRT_E_UNSUPPORTED, "no render-kernel variant covers feature mask 0x%x", s->dev.features);
    const size_t smem = (kVariantMask[v] & F_SMEM_TABLES) ? s->dev.hot_bytes : 0;
    RenderLaunch L;
    L.scene = &s->dev;
    L.sobol = S->sobol;
    L.counters = c->d_counters;
    L.smem_bytes = smem;
    L.stream = c->stream;
    L.grid = 0;
    int blocks_per_sm = c->variant_blocks[v];
    if (blocks_per_sm <= 0 || smem > c->variant_smem[v]) {  // occupancy query (no allocation), cached as launch_render does
        RenderVariantInfo info = {};
        L.args = nullptr;
        L.out = nullptr;
        cudaError_t e = kVariants[v](L, &info);
        if (e != cudaSuccess) return fail(SRT_E_CUDA, "render variant %d: %s", v, cudaGetErrorString(e));
        c->variant_blocks[v] = info.blocks_per_sm;
        c->variant_regs[v] = info.registers;
        c->variant_smem[v] = smem;
        blocks_per_sm = info.blocks_per_sm;
    }
    blocks_per_sm = std::max(1, blocks_per_sm);

    const int chunk = std::max(1, std::min(count, env_int("SRT_CHUNK", 1)));
    const int n_chunks = (count + chunk - 1) / chunk;
    srt::RenderArgs A;
    A.nx = p->nx;
    A.ny = p->ny;
    A.s_begin = p->s_begin;
    A.s_end = p->s_end;
    A.max_depth = p->max_depth;
    A.seed = p->seed;
    A.mod_ny = (p->flags & SRT_COMPAT_PIXEL_MOD_NY) ? 1 : 0;
    A.chunk = chunk;
    A.n_chunks = n_chunks;
    A.n_items = (unsigned long long)n_pixels * (unsigned long long)n_chunks;
    A.fx_clamp = 4.0e9f / (float)std::max(count, p->ns);
    A.regen_threshold = std::max(1, std::min(32, env_int("SRT_REGEN", 8)));
    A.grab = std::max(32, env_int("SRT_GRAB", 128));
    A.cta_sync = std::max(0, std::min(2, env_int("SRT_CTA_SYNC", 0)));
    A.dbg_pixel = -1;
    A.dbg_sample = -1;

    // the array keeps the previous call's sums (every call shares it): the part this frame adds into starts at zero
    const size_t nv = 3 * (size_t)n_pixels;
    CUDA_TRY(cudaMemsetAsync(S->fixed, 0, sizeof(long long) * nv, c->stream));
    CUDA_TRY(cudaMemsetAsync(c->d_counters + 7, 0, sizeof(unsigned long long), c->stream));
    const long long warps_needed = ((long long)A.n_items + 31) / 32;
    const long long blocks_needed = (warps_needed + (srt::kRenderThreads / 32) - 1) / (srt::kRenderThreads / 32);
    L.args = &A;
    L.out = S->fixed;
    L.grid = (int)std::max<long long>(1, std::min<long long>(blocks_needed, (long long)c->sm_count * blocks_per_sm));
    cudaError_t e = kVariants[v](L, nullptr);
    c->launches++;
    if (e != cudaSuccess) return fail(SRT_E_CUDA, "k_render (variant %d, static scratch): %s", v, cudaGetErrorString(e));
    srt::k_resolve_fixed<<<(unsigned)((nv + 255) / 256), 256, 0, c->stream>>>(S->fixed, (long long)nv, (float)count,
                                                                              rgb8_device, nullptr, 0);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

int srt_render(SrtScene* s, const SrtRenderParams* p, float* accum_rgb_device) {
    if (!s || !accum_rgb_device) return fail(SRT_E_INVALID, "null argument");
    int rc = check_params(p);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(s->ctx->device));
    return launch_render(s, p, accum_rgb_device, false);
}

int srt_render_fixed(SrtScene* s, const SrtRenderParams* p, int64_t* fixed_device) {
    if (!s || !fixed_device) return fail(SRT_E_INVALID, "null argument");
    int rc = check_params(p);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(s->ctx->device));
    static_assert(sizeof(long long) == sizeof(int64_t), "int64_t is long long on this ABI");
    return launch_render(s, p, nullptr, false, reinterpret_cast<long long*>(fixed_device));
}

int srt_resolve_fixed(SrtContext* c, int64_t* fixed_device, int64_t n_pixels, int32_t n_samples, uint8_t* rgb8_out,
                      float* accum_out_device, int clear) {
    if (!c || !fixed_device) return fail(SRT_E_INVALID, "null argument");
    if (n_pixels <= 0) return 0;
    CUDA_TRY(cudaSetDevice(c->device));
    size_t nv = 3 * (size_t)n_pixels;
    bool dev_out = rgb8_out == nullptr || is_device_pointer(rgb8_out);
    uint8_t* dst = rgb8_out;
    if (!dev_out) {
        if (c->rgb8_cap < nv) {
            CUDA_TRY(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_rgb8);
            c->d_rgb8 = nullptr;
            c->rgb8_cap = 0;
            CUDA_TRY(cudaMalloc(&c->d_rgb8, nv));
            c->rgb8_cap = nv;
        }
        dst = c->d_rgb8;
    }
    srt::k_resolve_fixed<<<(unsigned)((nv + 255) / 256), 256, 0, c->stream>>>(
        reinterpret_cast<long long*>(fixed_device), (long long)nv, (float)n_samples, dst, accum_out_device, clear);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    if (!dev_out) {
        CUDA_TRY(cudaMemcpyAsync(rgb8_out, dst, nv, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
    }
    return 0;
}


// ---- peer-memory exchange (srt_aux_kernels.cuh: k_exchange_peers) ------------------------------------------------
struct SrtExchange {
    SrtContext* ctx = nullptr;
    int n_ranks = 0, rank = 0;
    long long n_pixels = 0, tile_values = 0, n_values = 0, padded_values = 0;
    size_t off_fixed[2] = {0, 0}, off_rgb = 0, off_flags = 0, bytes = 0;
    uint8_t* block = nullptr;                   // this rank's block (cudaMalloc)
    uint8_t* base[srt::kExMaxRanks] = {};       // every rank's block as mapped here (base[rank] == block)
    unsigned* d_cta_counter = nullptr;
    unsigned frame = 0;
    bool connected = false;
};

int srt_exchange_create(SrtContext* ctx, int64_t n_pixels, int32_t n_ranks, int32_t rank, SrtExchange** out) {
    if (!ctx || !out) return fail(SRT_E_INVALID, "null argument");
    *out = nullptr;
    if (n_pixels <= 0 || n_ranks < 1 || n_ranks > srt::kExMaxRanks || rank < 0 || rank >= n_ranks)
        return fail(SRT_E_INVALID, "srt_exchange_create: bad sizes");
    CUDA_TRY(cudaSetDevice(ctx->device));
    SrtExchange* ex = new SrtExchange();
    ex->ctx = ctx;
    ex->n_ranks = n_ranks;
    ex->rank = rank;
    ex->n_pixels = n_pixels;
    ex->n_values = 3 * n_pixels;
    ex->tile_values = ((ex->n_values + n_ranks - 1) / n_ranks + 3) & ~3ll;  // a multiple of 4: one word of 8-bit values per thread
    ex->padded_values = ex->tile_values * n_ranks;
    const size_t fixed_bytes = ((size_t)ex->padded_values * 8 + 255) & ~(size_t)255;
    ex->off_fixed[0] = 0;
    ex->off_fixed[1] = fixed_bytes;
    ex->off_rgb = 2 * fixed_bytes;
    ex->off_flags = ex->off_rgb + (((size_t)ex->padded_values + 255) & ~(size_t)255);
    ex->bytes = ex->off_flags + 256;
    cudaError_t e = cudaMalloc(&ex->block, ex->bytes);
    if (e == cudaSuccess) e = cudaMalloc(&ex->d_cta_counter, sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMemsetAsync(ex->block, 0, ex->bytes, ctx->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(ex->d_cta_counter, 0, sizeof(unsigned), ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) {
        cudaFree(ex->block);
        cudaFree(ex->d_cta_counter);
        delete ex;
        return fail(SRT_E_CUDA, "srt_exchange_create: %s", cudaGetErrorString(e));
    }
    ex->base[rank] = ex->block;
    *out = ex;
    return 0;
}

int srt_exchange_handle(SrtExchange* ex, uint8_t* handle_out) {
    if (!ex || !handle_out) return fail(SRT_E_INVALID, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == SRT_EXCHANGE_HANDLE_BYTES, "IPC handle size");
    CUDA_TRY(cudaSetDevice(ex->ctx->device));
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, ex->block));
    std::memcpy(handle_out, &h, sizeof h);
    return 0;
}

int srt_exchange_connect(SrtExchange* ex, const uint8_t* handles) {
    if (!ex || !handles) return fail(SRT_E_INVALID, "null argument");
    CUDA_TRY(cudaSetDevice(ex->ctx->device));
    for (int r = 0; r < ex->n_ranks; r++) {
        if (r == ex->rank) continue;
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handles + (size_t)r * SRT_EXCHANGE_HANDLE_BYTES, sizeof h);
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return fail(SRT_E_CUDA, "srt_exchange_connect: rank %d: %s", r, cudaGetErrorString(e));
        ex->base[r] = static_cast<uint8_t*>(p);
    }
    ex->connected = true;
    return 0;
}

int64_t* srt_exchange_fixed(SrtExchange* ex) {
    return ex ? reinterpret_cast<int64_t*>(ex->block + ex->off_fixed[(ex->frame + 1) & 1]) : nullptr;  // the NEXT frame's buffer
}

uint8_t* srt_exchange_image(SrtExchange* ex) { return ex ? ex->block + ex->off_rgb : nullptr; }

int srt_exchange_run(SrtExchange* ex, int32_t n_samples, uint8_t* rgb8_host) {
    if (!ex) return fail(SRT_E_INVALID, "null argument");
    if (!ex->connected && ex->n_ranks > 1) return fail(SRT_E_INVALID, "srt_exchange_run before srt_exchange_connect");
    SrtContext* c = ex->ctx;
    CUDA_TRY(cudaSetDevice(c->device));
    ex->frame++;
    srt::ExArgs a = {};
    for (int r = 0; r < ex->n_ranks; r++) a.base[r] = ex->base[r];
    a.n_ranks = ex->n_ranks;
    a.rank = ex->rank;
    a.tile_values = ex->tile_values;
    a.n_values = ex->n_values;
    a.off_fixed = (long long)ex->off_fixed[ex->frame & 1];
    a.off_rgb = (long long)ex->off_rgb;
    a.off_flags = (long long)ex->off_flags;
    a.ns = (float)n_samples;
    a.frame = ex->frame;
    const unsigned grid = (unsigned)((ex->tile_values / 4 + 255) / 256);
    srt::k_exchange_peers<<<grid, 256, 0, c->stream>>>(a, ex->d_cta_counter);
    srt::k_exchange_wait<<<1, 32, 0, c->stream>>>(reinterpret_cast<unsigned*>(ex->block + ex->off_flags), ex->n_ranks, ex->frame);
    c->launches += 2;
    CUDA_TRY(cudaGetLastError());
    // the sums of the PREVIOUS frame: every rank has finished reading them (it arrived at this frame's barrier after its
    // previous exchange kernel), so their owner clears them for the next frame
    CUDA_TRY(cudaMemsetAsync(ex->block + ex->off_fixed[(ex->frame + 1) & 1], 0, (size_t)ex->padded_values * 8, c->stream));
    if (rgb8_host) {
        CUDA_TRY(cudaMemcpyAsync(rgb8_host, ex->block + ex->off_rgb, (size_t)ex->n_values, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
        unsigned err = 0;
        CUDA_TRY(cudaMemcpy(&err, ex->block + ex->off_flags + 128, sizeof err, cudaMemcpyDeviceToHost));
        if (err) return fail(SRT_E_CUDA, "srt_exchange_run: a rank did not arrive (time-out inside the exchange kernel)");
    }
    return 0;
}

int srt_exchange_destroy(SrtExchange* ex) {
    if (!ex) return 0;
    cudaSetDevice(ex->ctx->device);
    cudaStreamSynchronize(ex->ctx->stream);
    for (int r = 0; r < ex->n_ranks; r++)
        if (r != ex->rank && ex->base[r]) cudaIpcCloseMemHandle(ex->base[r]);
    cudaFree(ex->block);
    cudaFree(ex->d_cta_counter);
    delete ex;
    return 0;
}

int srt_primary_rays(SrtScene* s, const SrtRenderParams* p, float* rays_out) {
    if (!s || !rays_out) return fail(SRT_E_INVALID, "null argument");
    int rc = check_params(p);
    if (rc) return rc;
    SrtContext* c = s->ctx;
    CUDA_TRY(cudaSetDevice(c->device));
    const long long n = (long long)p->nx * p->ny * (p->s_end - p->s_begin);
    if (n <= 0) return 0;
    if ((rc = ensure_sobol(c, (uint32_t)p->s_end))) return rc;
    DevOut o;
    if ((rc = o.init(rays_out, sizeof(float) * 7 * (size_t)n))) return rc;
    srt::RenderArgs A = {};
    A.nx = p->nx;
    A.ny = p->ny;
    A.s_begin = p->s_begin;
    A.s_end = p->s_end;
    A.seed = p->seed;
    A.mod_ny = (p->flags & SRT_COMPAT_PIXEL_MOD_NY) ? 1 : 0;
    srt::k_primary_rays<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(s->dev.camera, A, c->d_sobol, (float*)o.ptr);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    return o.finish(c->stream);
}

int srt_render_counted(SrtScene* s, const SrtRenderParams* p, float* accum_rgb_device, SrtCounters* out) {
    if (!s || !accum_rgb_device) return fail(SRT_E_INVALID, "null argument");
    int rc = check_params(p);
    if (rc) return rc;
    SrtContext* c = s->ctx;
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaMemsetAsync(c->d_counters, 0, 8 * sizeof(unsigned long long), c->stream));
    rc = launch_render(s, p, accum_rgb_device, true);
    if (rc) return rc;
    unsigned long long h[9];
    CUDA_TRY(cudaMemcpyAsync(h, c->d_counters, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    c->last.paths = h[0];
    c->last.rays = h[1];
    c->last.probe_rays = h[2];
    c->last.node_visits = h[3];
    c->last.prim_tests = h[4];
    c->last.reject_iters = h[5];
    c->last.kernel_launches = c->launches;
    c->last.culled_rays = h[6];
    c->last.stack_overflows = h[8];
    if (out) *out = c->last;
    return 0;
}

int srt_counters_get(const SrtContext* ctx, SrtCounters* out) {
    if (!ctx || !out) return fail(SRT_E_INVALID, "null argument");
    *out = ctx->last;
    out->kernel_launches = ctx->launches;
    // the overflow tally is live (every kernel of the context adds to it): read it behind the queued work
    CUDA_TRY(cudaSetDevice(ctx->device));
    unsigned long long ov = 0;
    CUDA_TRY(cudaMemcpyAsync(&ov, ctx->d_counters + 8, sizeof ov, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    out->stack_overflows = ov;
    return 0;
}

int srt_resolve(SrtContext* c, const float* accum_rgb_device, int64_t n_pixels, int32_t n_samples, uint8_t* rgb8_out) {
    if (!c || !accum_rgb_device || !rgb8_out) return fail(SRT_E_INVALID, "null argument");
    if (n_pixels <= 0) return 0;
    CUDA_TRY(cudaSetDevice(c->device));
    size_t nv = 3 * (size_t)n_pixels;
    bool dev_out = is_device_pointer(rgb8_out);
    uint8_t* dst = rgb8_out;
    if (!dev_out) {
        if (c->rgb8_cap < nv) {
            CUDA_TRY(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_rgb8);
            c->d_rgb8 = nullptr;
            c->rgb8_cap = 0;
            CUDA_TRY(cudaMalloc(&c->d_rgb8, nv));
            c->rgb8_cap = nv;
        }
        dst = c->d_rgb8;
    }
    unsigned blocks = (unsigned)((nv + 255) / 256);
    srt::k_resolve<<<blocks, 256, 0, c->stream>>>(accum_rgb_device, (long long)nv, (float)n_samples, dst);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    if (!dev_out) {
        CUDA_TRY(cudaMemcpyAsync(rgb8_out, dst, nv, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
    }
    return 0;
}

int srt_render_image(SrtScene* s, const SrtRenderParams* p, uint8_t* rgb8_host, float* accum_host) {
    if (!s || !rgb8_host) return fail(SRT_E_INVALID, "null argument");
    int rc = check_params(p);
    if (rc) return rc;
    SrtContext* c = s->ctx;
    CUDA_TRY(cudaSetDevice(c->device));
    size_t nv = 3 * (size_t)p->nx * p->ny;
    // the kernels add into the context's fixed-point buffer (zero between renders); ONE more kernel turns it
    // into the 8-bit image (and the float sums when asked for) and clears it again
    {
        size_t need = sizeof(long long) * nv;
        if (c->fixed_cap < need) {
            CUDA_TRY(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_fixed);
            c->d_fixed = nullptr;
            c->fixed_cap = 0;
            CUDA_TRY(cudaMalloc(&c->d_fixed, need));
            c->fixed_cap = need;
            c->fixed_dirty = true;
        }
        if (c->fixed_dirty) CUDA_TRY(cudaMemsetAsync(c->d_fixed, 0, c->fixed_cap, c->stream));
        c->fixed_dirty = true;
    }
    float* d_acc = nullptr;
    if (accum_host) {
        if (c->accum_cap < nv) {
            CUDA_TRY(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_accum);
            c->d_accum = nullptr;
            c->accum_cap = 0;
            CUDA_TRY(cudaMalloc(&c->d_accum, sizeof(float) * nv));
            c->accum_cap = nv;
        }
        d_acc = c->d_accum;
    }
    rc = launch_render(s, p, nullptr, false, c->d_fixed);
    if (rc) return rc;
    rc = srt_resolve_fixed(c, reinterpret_cast<int64_t*>(c->d_fixed), (int64_t)p->nx * p->ny, p->s_end - p->s_begin,
                           rgb8_host, d_acc, 1);
    if (rc) return rc;
    c->fixed_dirty = false;
    if (accum_host) {
        CUDA_TRY(cudaMemcpyAsync(accum_host, c->d_accum, sizeof(float) * nv, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
    }
    return 0;
}

// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
int srt_render_image_static(SrtScene* s, const SrtRenderParams* p, uint8_t* rgb8_device) {
    if (!s || !rgb8_device) return fail(SRT_E_INVALID, "null argument");
    int rc = check_params(p);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(s->ctx->device));
    return launch_render_static(s, p, rgb8_device);
}

}  // extern "C"
