// srt_api.cu — the C ABI of include/srt_ffi.h on top of the CUDA kernels.  No CPU path: every
// compute entry point fails with SRT_E_NODEVICE / SRT_E_CUDA when there is no usable GPU.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "srt_bvh_build.h"
#include "srt_ffi.h"
#include "srt_kernels.cuh"
#include "srt_types.h"

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
    unsigned long long* d_counters = nullptr;  // 8 x u64
    uint8_t* d_rgb8 = nullptr;                 // resolve staging for host outputs
    size_t rgb8_cap = 0;
    uint8_t* h_rgb8 = nullptr;  // pinned
    size_t h_rgb8_cap = 0;
};

struct SrtScene {
    SrtContext* ctx = nullptr;
    uint8_t* d_blob = nullptr;
    size_t blob_bytes = 0;
    DScene dev = {};
    SrtSceneInfo info = {};
    std::vector<double*> d_merl;
    std::vector<const double*> merl_ptrs_host;
    uint32_t* d_sobol = nullptr;
    uint32_t sobol_n = 0;
    float* d_accum = nullptr;  // srt_render_image's internal accumulation buffer
    size_t accum_cap = 0;
    float* h_accum = nullptr;  // pinned staging
    size_t h_accum_cap = 0;
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
    size_t add(size_t bytes) {
        size_t off = (total + 255) & ~(size_t)255;
        total = off + bytes;
        return off;
    }
};

void pad_box(float* b) {
    for (int a = 0; a < 3; a++) {
        float m = std::max(std::fabs(b[a]), std::fabs(b[3 + a]));
        float pad = 1e-5f * std::max(m, b[3 + a] - b[a]) + 1e-7f;
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
    CUDA_TRY(cudaMalloc(&c->d_counters, 8 * sizeof(unsigned long long)));
    *out = c;
    return 0;
}

int srt_context_destroy(SrtContext* ctx) {
    if (!ctx) return 0;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaFree(ctx->d_counters);
    cudaFree(ctx->d_rgb8);
    if (ctx->h_rgb8) cudaFreeHost(ctx->h_rgb8);
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
    cudaFree(s->d_sobol);
    cudaFree(s->d_accum);
    if (s->h_accum) cudaFreeHost(s->h_accum);
    delete s;
    return 0;
}

int srt_scene_create(SrtContext* ctx, const SrtSceneDesc* desc, SrtScene** out) {
    if (!ctx || !desc || !out) return fail(SRT_E_INVALID, "null argument");
    *out = nullptr;
    const SrtSceneDesc& d = *desc;
    int rc = validate(d);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(ctx->device));
    auto t0 = std::chrono::steady_clock::now();

    // ---- acceleration structures -----------------------------------------------------------
    std::vector<DBvh> bvhs(d.n_bvhs);
    std::vector<WideNode4> wide;
    std::vector<RefNode> refn;
    std::vector<int32_t> leaf_rank;
    std::vector<float4> leaf_geom;
    float top_sah = 0.f;
    int top_items = 0;
    for (int bi = 0; bi < d.n_bvhs; bi++) {
        const SrtBvh& sb = d.bvhs[bi];
        int n = sb.item_count;
        std::vector<float> boxes(6 * (size_t)n);
        bool all_sphere = true, all_tri = true;
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
            bool single = g.prim_count == 1;
            all_sphere = all_sphere && single && d.prims[g.first_prim].type == SRT_PRIM_SPHERE;
            all_tri = all_tri && single && d.prims[g.first_prim].type == SRT_PRIM_TRIANGLE;
        }
        WideBvh wb;
        build_wide_bvh(boxes.data(), n, wb);
        std::vector<RefNode> rt;
        build_ref_tree(d.bvh_item_boxes + 6 * (size_t)sb.first_item, n, rt);
        DBvh& B = bvhs[bi];
        B.kind = all_sphere ? BVH_KIND_SPHERE : (all_tri ? BVH_KIND_TRIANGLE : BVH_KIND_GENERIC);
        B.first_item = sb.first_item;
        B.item_count = n;
        B.node_base = (int32_t)wide.size();
        B.ref_base = (int32_t)refn.size();
        B.leaf_base = (int32_t)leaf_rank.size();
        B.geom_base = (int32_t)leaf_geom.size();
        B._pad = 0;
        wide.insert(wide.end(), wb.nodes.begin(), wb.nodes.end());
        refn.insert(refn.end(), rt.begin(), rt.end());
        leaf_rank.insert(leaf_rank.end(), wb.order.begin(), wb.order.end());
        if (B.kind == BVH_KIND_SPHERE) {
            for (int s = 0; s < n; s++) {
                const SrtPrim& p = d.prims[d.groups[d.bvh_items[sb.first_item + wb.order[s]]].first_prim];
                const float* q = d.spheres + 4 * (size_t)p.index;
                leaf_geom.push_back(float4{q[0], q[1], q[2], q[3]});
            }
        } else if (B.kind == BVH_KIND_TRIANGLE) {
            leaf_geom.reserve(leaf_geom.size() + 3 * (size_t)n);
            for (int s = 0; s < n; s++) {
                const SrtPrim& p = d.prims[d.groups[d.bvh_items[sb.first_item + wb.order[s]]].first_prim];
                const float* q = d.tri_pos + 9 * (size_t)p.index;
                leaf_geom.push_back(float4{q[0], q[1], q[2], 0.f});
                leaf_geom.push_back(float4{q[3], q[4], q[5], 0.f});
                leaf_geom.push_back(float4{q[6], q[7], q[8], 0.f});
            }
        }
        if (n > top_items) {
            top_items = n;
            top_sah = wb.sah_cost;
        }
    }

    // ---- repack small records to 16-byte units ------------------------------------------------
    std::vector<float4> rects(2 * (size_t)d.n_rects);
    for (int i = 0; i < d.n_rects; i++) {
        const float* q = d.rects + 5 * (size_t)i;
        rects[2 * i] = float4{q[0], q[1], q[2], q[3]};
        rects[2 * i + 1] = float4{q[4], 0.f, 0.f, 0.f};
    }
    std::vector<int32_t> perlin_perm;
    if (d.perlin_ranvec) {
        if (!d.perlin_perm_x || !d.perlin_perm_y || !d.perlin_perm_z) return fail(SRT_E_INVALID, "incomplete perlin tables");
        perlin_perm.insert(perlin_perm.end(), d.perlin_perm_x, d.perlin_perm_x + 256);
        perlin_perm.insert(perlin_perm.end(), d.perlin_perm_y, d.perlin_perm_y + 256);
        perlin_perm.insert(perlin_perm.end(), d.perlin_perm_z, d.perlin_perm_z + 256);
    }

    // ---- one device allocation -----------------------------------------------------------------
    struct Piece {
        const void* src;
        size_t bytes;
        size_t off;
    };
    std::vector<Piece> pieces;
    Layout lay;
    auto piece = [&](const void* src, size_t bytes) -> size_t {
        size_t off = lay.add(bytes);
        if (bytes) pieces.push_back(Piece{src, bytes, off});
        return off;
    };
    size_t o_entries = piece(d.entries, sizeof(SrtEntry) * (size_t)d.n_entries);
    size_t o_ops = piece(d.ops, sizeof(SrtXformOp) * (size_t)d.n_ops);
    size_t o_groups = piece(d.groups, sizeof(SrtGroup) * (size_t)d.n_groups);
    size_t o_prims = piece(d.prims, sizeof(SrtPrim) * (size_t)d.n_prims);
    size_t o_items = piece(d.bvh_items, sizeof(int32_t) * (size_t)d.n_bvh_items);
    size_t o_media = piece(d.media, sizeof(SrtMedium) * (size_t)d.n_media);
    size_t o_bvhs = piece(bvhs.data(), sizeof(DBvh) * bvhs.size());
    size_t o_wide = piece(wide.data(), sizeof(WideNode4) * wide.size());
    size_t o_ref = piece(refn.data(), sizeof(RefNode) * refn.size());
    size_t o_rank = piece(leaf_rank.data(), sizeof(int32_t) * leaf_rank.size());
    size_t o_geom = piece(leaf_geom.data(), sizeof(float4) * leaf_geom.size());
    size_t o_spheres = piece(d.spheres, sizeof(float) * 4 * (size_t)d.n_spheres);
    size_t o_msph = piece(d.msph, sizeof(float) * 9 * (size_t)d.n_msph);
    size_t o_rects = piece(rects.data(), sizeof(float4) * rects.size());
    size_t o_tpos = piece(d.tri_pos, sizeof(float) * 9 * (size_t)d.n_tris);
    size_t o_tnrm = piece(d.tri_nrm, sizeof(float) * 9 * (size_t)d.n_tris);
    size_t o_tuv = piece(d.tri_uv, sizeof(float) * 6 * (size_t)d.n_tris);
    size_t o_lights = piece(d.lights, sizeof(SrtLight) * (size_t)d.n_lights);
    size_t o_mats = piece(d.materials, sizeof(SrtMaterial) * (size_t)d.n_materials);
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
    for (const Piece& p : pieces) SCENE_TRY(cudaMemcpyAsync(s->d_blob + p.off, p.src, p.bytes, cudaMemcpyHostToDevice, ctx->stream));
    std::vector<DImage> imgtab(d.n_images);
    for (int i = 0; i < d.n_images; i++) imgtab[i] = DImage{s->d_blob + o_img[i], d.images[i].nx, d.images[i].ny};
    if (d.n_images) SCENE_TRY(cudaMemcpyAsync(s->d_blob + o_imgtab, imgtab.data(), sizeof(DImage) * imgtab.size(), cudaMemcpyHostToDevice, ctx->stream));
    for (int i = 0; i < d.n_merl; i++) {
        double* p = nullptr;
        SCENE_TRY(cudaMalloc(&p, sizeof(double) * 3 * 90 * 90 * 180));
        s->d_merl.push_back(p);
        SCENE_TRY(cudaMemcpyAsync(p, d.merl_tables[i], sizeof(double) * 3 * 90 * 90 * 180, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (d.n_merl) SCENE_TRY(cudaMemcpyAsync(s->d_blob + o_merltab, s->d_merl.data(), sizeof(double*) * s->d_merl.size(), cudaMemcpyHostToDevice, ctx->stream));
    SCENE_TRY(cudaStreamSynchronize(ctx->stream));
#undef SCENE_TRY

    DScene& D = s->dev;
    uint8_t* base = s->d_blob;
    D.entries = (const SrtEntry*)(base + o_entries);
    D.ops = (const SrtXformOp*)(base + o_ops);
    D.groups = (const SrtGroup*)(base + o_groups);
    D.prims = (const SrtPrim*)(base + o_prims);
    D.bvh_items = (const int32_t*)(base + o_items);
    D.media = (const SrtMedium*)(base + o_media);
    D.bvhs = (const DBvh*)(base + o_bvhs);
    D.wide_nodes = (const WideNode4*)(base + o_wide);
    D.ref_nodes = (const RefNode*)(base + o_ref);
    D.leaf_rank = (const int32_t*)(base + o_rank);
    D.leaf_geom = (const float4*)(base + o_geom);
    D.spheres = (const float4*)(base + o_spheres);
    D.msph = (const float*)(base + o_msph);
    D.rects = (const float4*)(base + o_rects);
    D.tri_pos = (const float*)(base + o_tpos);
    D.tri_nrm = (const float*)(base + o_tnrm);
    D.tri_uv = (const float*)(base + o_tuv);
    D.lights = (const SrtLight*)(base + o_lights);
    D.materials = (const SrtMaterial*)(base + o_mats);
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
    D._pad = 0;
    D.camera = d.camera;

    auto t1 = std::chrono::steady_clock::now();
    s->info.n_prims = d.n_prims;
    s->info.n_entries = d.n_entries;
    s->info.n_bvhs = d.n_bvhs;
    s->info.n_lights = d.n_lights;
    s->info.bvh_nodes = (int64_t)wide.size();
    s->info.bvh_node_bytes = (int64_t)(wide.size() * sizeof(WideNode4));
    s->info.device_bytes = (int64_t)s->blob_bytes + (int64_t)d.n_merl * 8 * 3 * 90 * 90 * 180;
    s->info.sah_cost = top_sah;
    s->info.build_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
    *out = s;
    return 0;
}

int srt_scene_info(const SrtScene* scene, SrtSceneInfo* out) {
    if (!scene || !out) return fail(SRT_E_INVALID, "null argument");
    *out = scene->info;
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
    srt::k_trace_closest<<<blocks, 128, 0, c->stream>>>(s->dev, (const float*)in.ptr, (long long)n, (int32_t*)o_id.ptr,
                                                       (float*)o_t.ptr, (float*)o_rec.ptr, 0x5EEDu);
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

int ensure_sobol(SrtScene* s, uint32_t n) {
    if (s->sobol_n >= n) return 0;
    uint32_t cap = std::max<uint32_t>(n, 64);
    std::vector<uint32_t> pts(2 * (size_t)cap);
    srt_sobol_points_u32(cap, pts.data());
    if (s->d_sobol) {
        CUDA_TRY(cudaStreamSynchronize(s->ctx->stream));
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

template <bool COUNT>
int launch_render(SrtScene* s, const SrtRenderParams* p, float* accum) {
    SrtContext* c = s->ctx;
    int count = p->s_end - p->s_begin;
    if (count == 0) return 0;
    int rc = ensure_sobol(s, (uint32_t)p->s_end);
    if (rc) return rc;
    srt::RenderArgs A;
    A.nx = p->nx;
    A.ny = p->ny;
    A.s_begin = p->s_begin;
    A.s_end = p->s_end;
    A.max_depth = p->max_depth;
    A.seed = p->seed;
    A.mod_ny = (p->flags & SRT_COMPAT_PIXEL_MOD_NY) ? 1 : 0;
    int L = 1;
    while (L < count && L < 32) L <<= 1;
    A.lanes_per_pixel = L;
    int ppw = 32 / L;
    long long n_pixels = (long long)p->nx * p->ny;
    A.n_tiles = (int)((n_pixels + ppw - 1) / ppw);
    const int threads = 256;
    int blocks_per_sm = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, srt::k_render<COUNT>, threads, 0));
    if (blocks_per_sm < 1) blocks_per_sm = 1;
    long long want = (A.n_tiles + (threads / 32) - 1) / (threads / 32);
    int grid = (int)std::min<long long>(want, (long long)c->sm_count * blocks_per_sm);
    if (grid < 1) grid = 1;
    srt::k_render<COUNT><<<grid, threads, 0, c->stream>>>(s->dev, A, s->d_sobol, accum, c->d_counters);
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
    return launch_render<false>(s, p, accum_rgb_device);
}

int srt_render_counted(SrtScene* s, const SrtRenderParams* p, float* accum_rgb_device, SrtCounters* out) {
    if (!s || !accum_rgb_device) return fail(SRT_E_INVALID, "null argument");
    int rc = check_params(p);
    if (rc) return rc;
    SrtContext* c = s->ctx;
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaMemsetAsync(c->d_counters, 0, 8 * sizeof(unsigned long long), c->stream));
    rc = launch_render<true>(s, p, accum_rgb_device);
    if (rc) return rc;
    unsigned long long h[8];
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
    if (out) *out = c->last;
    return 0;
}

int srt_counters_get(const SrtContext* ctx, SrtCounters* out) {
    if (!ctx || !out) return fail(SRT_E_INVALID, "null argument");
    *out = ctx->last;
    out->kernel_launches = ctx->launches;
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
    if (s->accum_cap < nv) {
        CUDA_TRY(cudaStreamSynchronize(c->stream));
        cudaFree(s->d_accum);
        s->d_accum = nullptr;
        s->accum_cap = 0;
        CUDA_TRY(cudaMalloc(&s->d_accum, sizeof(float) * nv));
        s->accum_cap = nv;
    }
    CUDA_TRY(cudaMemsetAsync(s->d_accum, 0, sizeof(float) * nv, c->stream));
    rc = launch_render<false>(s, p, s->d_accum);
    if (rc) return rc;
    if (accum_host) CUDA_TRY(cudaMemcpyAsync(accum_host, s->d_accum, sizeof(float) * nv, cudaMemcpyDeviceToHost, c->stream));
    return srt_resolve(c, s->d_accum, (int64_t)p->nx * p->ny, p->s_end - p->s_begin, rgb8_host);
}

}  // extern "C"
