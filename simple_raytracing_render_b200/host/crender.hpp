// crender.hpp — host side of the drop-in boundary: the DEFINITION of the reference's CRender.
//
// The reference declares `class CRender` in Raytracing_n/Render.h:16-51 (constructor
// CRender(maxDepth, width, height, num_ray, thread_count, filePath) and void Run()) but never
// defines it; its free-function twin is main()/renderthread() in Raytracing_n.cpp:815-952.  This
// header supplies the missing definitions on top of the C ABI in include/srt_ffi.h, so that
//
//     CRender r(50, 512, 512, 16, 8, "out.ppm");  r.Run();
//
// renders on the GPU and writes the same ASCII P3 file main() writes (Raytracing_n.cpp:886,
// :873-876).  Like srt_flatten.hpp it is compiled in a TU that already includes the reference
// headers, and it contains no reference code.  The backend library is bound with dlopen so that
// the host program does not need the CUDA runtime at link time; there is no CPU fallback: if the
// library or a GPU is missing, Run() reports the error and renders nothing.
#ifndef SRT_CRENDER_HPP
#define SRT_CRENDER_HPP

#include <dlfcn.h>

#include <chrono>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "Render.h"  // the reference's declaration (found through -I <reference>/Raytracing_n)
#include "srt_ffi.h"
#include "srt_flatten.hpp"

namespace srt_host {

struct Backend {
    void* handle = nullptr;
    const char* (*last_error)() = nullptr;
    int (*context_create)(int, void*, SrtContext**) = nullptr;
    int (*context_destroy)(SrtContext*) = nullptr;
    int (*scene_create)(SrtContext*, const SrtSceneDesc*, SrtScene**) = nullptr;
    int (*scene_destroy)(SrtScene*) = nullptr;
    int (*render_image)(SrtScene*, const SrtRenderParams*, uint8_t*, float*) = nullptr;
};

inline Backend& backend() {
    static Backend b;
    return b;
}
inline std::string& error_text() {
    static std::string e;
    return e;
}
inline int& status() {
    static int s = 0;
    return s;
}
inline const char* backend_error() { return error_text().c_str(); }
inline int last_status() { return status(); }

// Bind libsrt_b200.so (path may be null: then the dynamic loader's search path is used).
inline int load_backend(const char* path) {
    Backend& b = backend();
    if (b.handle) return 0;
    b.handle = dlopen(path ? path : "libsrt_b200.so", RTLD_NOW | RTLD_LOCAL);
    if (!b.handle) {
        error_text() = std::string("cannot load the CUDA backend: ") + dlerror();
        return SRT_E_NODEVICE;
    }
#define SRT_BIND(field, sym)                                          \
    b.field = (decltype(b.field))dlsym(b.handle, sym);                \
    if (!b.field) {                                                   \
        error_text() = std::string("backend lacks symbol ") + sym;    \
        dlclose(b.handle);                                            \
        b.handle = nullptr;                                           \
        return SRT_E_INVALID;                                         \
    }
    SRT_BIND(last_error, "srt_last_error")
    SRT_BIND(context_create, "srt_context_create")
    SRT_BIND(context_destroy, "srt_context_destroy")
    SRT_BIND(scene_create, "srt_scene_create")
    SRT_BIND(scene_destroy, "srt_scene_destroy")
    SRT_BIND(render_image, "srt_render_image")
#undef SRT_BIND
    return 0;
}

// The scene CRender::Run renders: the (world, light list, camera) triple every reference scene
// function returns (e.g. Raytracing_n.cpp:216).  Render.h gives Run() no way to receive it (its
// scene functions are private and undefined), so the host program registers it here first.
struct SceneTriple {
    const hitable* world = nullptr;
    const hitable* hlist = nullptr;
    const camera* cam = nullptr;
};
inline SceneTriple& scene_triple() {
    static SceneTriple t;
    return t;
}
inline void set_scene(const hitable* world, const hitable* hlist, const camera* cam) {
    scene_triple().world = world;
    scene_triple().hlist = hlist;
    scene_triple().cam = cam;
}

}  // namespace srt_host

inline CRender::CRender()
    : colors(nullptr), curfinishedcount(0), filePath(nullptr), maxDepth(50), nx(1000), ny(1000), ns(50),
      thread_count(8), finishedPixel(0), isfinished(false) {}  // defaults: Raytracing_n.cpp:33-43

inline CRender::CRender(int maxDepth, int width, int height, int num_ray, int thread_count, char* filePath)
    : colors(nullptr), curfinishedcount(0), filePath(filePath), maxDepth(maxDepth), nx(width), ny(height),
      ns(num_ray), thread_count(thread_count), finishedPixel(0), isfinished(false) {}

inline CRender::~CRender() {}

inline int CRender::getpixel() { return finishedPixel++; }  // Raytracing_n.cpp:716-719

// Blocking, one host thread (the reference's thread_count only sized its CPU pool; the GPU path
// ignores it).  Mirrors main(): header, render, per-pixel "r g b" lines, elapsed ms.
inline void CRender::Run() {
    using namespace srt_host;
    status() = 0;
    Backend& b = backend();
    if (!b.handle && load_backend(nullptr) != 0) {
        status() = SRT_E_NODEVICE;
        std::cerr << "CRender::Run: " << backend_error() << "\n";
        return;
    }
    const SceneTriple& st = scene_triple();
    SrtFlatScene flat;
    int rc = srt_flatten_scene(st.world, st.hlist, st.cam, flat);
    if (rc != 0) {
        status() = rc;
        error_text() = "flatten: " + flat.error;
        std::cerr << "CRender::Run: " << backend_error() << "\n";
        return;
    }
    SrtSceneDesc desc = flat.desc();
    auto t0 = std::chrono::steady_clock::now();
    SrtContext* ctx = nullptr;
    SrtScene* scene = nullptr;
    std::vector<uint8_t> rgb((size_t)nx * ny * 3);
    rc = b.context_create(0, nullptr, &ctx);
    if (rc == 0) rc = b.scene_create(ctx, &desc, &scene);
    if (rc == 0) {
        SrtRenderParams p;
        p.nx = nx;
        p.ny = ny;
        p.ns = ns;
        p.s_begin = 0;
        p.s_end = ns;
        p.max_depth = maxDepth;
        p.seed = 0x5EEDu;
        p.flags = (nx == ny) ? SRT_COMPAT_PIXEL_MOD_NY : 0u;  // Raytracing_n.cpp:827 is only valid for squares
        rc = b.render_image(scene, &p, rgb.data(), nullptr);
    }
    if (rc != 0) {
        status() = rc;
        error_text() = b.last_error();
        std::cerr << "CRender::Run: " << backend_error() << "\n";
    } else {
        // the render is complete whether or not a file was asked for: a host that polls isfinished /
        // finishedPixel as the reference's progress display does (Raytracing_n.cpp:819-825) must see it
        finishedPixel = nx * ny;
        isfinished = true;
    }
    // `.pnm` / `.p6`: binary PPM (3 bytes per pixel instead of ~12 characters); anything else, e.g. the `.ppm` the
    // reference writes, is its own P3 text format
    const std::string path_s = filePath ? filePath : "";
    const bool binary = path_s.size() >= 4 && (path_s.compare(path_s.size() - 4, 4, ".pnm") == 0 ||
                                               path_s.compare(path_s.size() - 3, 3, ".p6") == 0);
    if (rc == 0 && filePath && binary) {
        outfile.open(filePath, std::ios_base::out | std::ios_base::binary);
        outfile << "P6\n" << nx << " " << ny << "\n255\n";
        outfile.write(reinterpret_cast<const char*>(rgb.data()), (std::streamsize)rgb.size());
        outfile.close();
    } else if (rc == 0 && filePath) {
        outfile.open(filePath, std::ios_base::out);
        outfile << "P3\n" << nx << " " << ny << "\n255\n";  // Raytracing_n.cpp:886
        std::string text;
        text.reserve((size_t)nx * ny * 12);
        char line[48];
        for (size_t i = 0; i < (size_t)nx * ny; i++) {  // Raytracing_n.cpp:873-876
            int n = snprintf(line, sizeof line, "%d %d %d\n", rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
            text.append(line, n);
        }
        outfile << text;
        outfile.close();
    }
    if (scene) b.scene_destroy(scene);
    if (ctx) b.context_destroy(ctx);
    auto t1 = std::chrono::steady_clock::now();
    std::cout << std::chrono::duration_cast<std::chrono::milliseconds>(t1 - t0).count() << "ms" << std::endl;  // :944-946
}

#endif  // SRT_CRENDER_HPP
