// oracle/ref_harness.cpp — the real reference as the oracle.  TEST INFRASTRUCTURE ONLY.
//
// This file #includes the UNMODIFIED reference translation unit (a UTF-8 conversion of
// /root/reference/Raytracing_n/Raytracing_n.cpp produced by oracle/Makefile into oracle/_ref/;
// the reference headers are used in place through -I) and exports a small C API around the
// reference's OWN functions and classes so that tests can
//   * read known-answer values out of the reference (drand48, PCG32, Sobol, camera),
//   * build scenes out of real reference objects (one extern "C" constructor per class),
//   * run world->hit / hitable_list::pdf_value / color() / renderthread of the reference on them,
//   * flatten them with the product's C++ flattener (host/srt_flatten.hpp) into SrtSceneDesc.
// Nothing here is product code; only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
// legs load the resulting oracle/_ref/libsrt_ref.so.
//
// Built by oracle/Makefile:  g++ -std=c++14 -O2 -ffp-contract=off -fpermissive -w -include shim.h
#define main ref_main
#include "Raytracing_n.utf8.cpp"
#undef main

#include "../simple_raytracing_render_b200/host/srt_flatten.hpp"
#include "../simple_raytracing_render_b200/host/crender.hpp"

// assimp is a Windows binary in the reference tree (external/assimp/lib): stub the three symbols
// the TU references.  Mesh scenes are built from triangle soups through ref_triangle() instead.
Assimp::Importer::Importer() {}
Assimp::Importer::~Importer() {}
const aiScene* Assimp::Importer::ReadFile(const char*, unsigned int) { return nullptr; }

namespace {

enum Kind { K_TEX = 1, K_MAT, K_HIT, K_CAM, K_SCENE };

struct Obj {
    int kind;
    void* p;
};

struct Scene {
    hitable* world;
    hitable* hlist;
    camera* cam;
    SrtFlatScene flat;
    SrtSceneDesc desc;
    bool flattened;
    int flatten_status;
};

std::vector<Obj> g_objs;
std::string g_err;
std::mutex g_api_lock;

// a material whose address is the primitive id carrier (hit_record::mat_ptr survives every copy)
class tag_material : public material {
   public:
    virtual bool scatter(const ray&, const hit_record&, scatter_record&) const { return false; }
};

int put(int kind, void* p) {
    g_objs.push_back(Obj{kind, p});
    return (int)g_objs.size() - 1;
}

template <class T>
T* get(int h, int kind) {
    if (h < 0) return nullptr;
    if (h >= (int)g_objs.size() || g_objs[h].kind != kind) {
        g_err = "bad handle";
        return nullptr;
    }
    return (T*)g_objs[h].p;
}

// restated fallback for Raytracing_n.cpp:721-812 with D = 2 (used only where the direction-number
// file is absent, i.e. on the GPU box; checked against the reference's own function by the tests)
void sobol2_fallback(unsigned N, double** pts) {
    unsigned x0 = 0, x1 = 0;
    pts[0][0] = pts[0][1] = 0;
    for (unsigned i = 1; i < N; i++) {
        unsigned c = 1, v = i - 1;
        while (v & 1) {
            v >>= 1;
            c++;
        }
        unsigned v0 = 1u << (32 - c);
        unsigned v1 = 1u << 31;
        for (unsigned k = 2; k <= c; k++) v1 = v1 ^ (v1 >> 1);
        x0 ^= v0;
        x1 ^= v1;
        pts[i][0] = (double)x0 / 4294967296.0;
        pts[i][1] = (double)x1 / 4294967296.0;
    }
}

std::string g_sobol_file = "/root/reference/contents/sobol/new-joe-kuo-6.21201";

double** make_sobol(unsigned N) {
    if (access(g_sobol_file.c_str(), R_OK) == 0 && N >= 2) return sobol_points(N, 2, g_sobol_file.c_str());
    double** pts = new double*[N];
    for (unsigned i = 0; i < N; i++) pts[i] = new double[2];
    sobol2_fallback(N, pts);
    return pts;
}

void free_sobol(double** pts, unsigned N) {
    for (unsigned i = 0; i < N; i++) delete[] pts[i];
    delete[] pts;
}

struct CoutSilencer {
    std::streambuf* old;
    std::ostringstream sink;
    CoutSilencer() { old = std::cout.rdbuf(sink.rdbuf()); }
    ~CoutSilencer() { std::cout.rdbuf(old); }
};

// the reference opens "..\results\20210709_balls.ppm" at static-init (Raytracing_n.cpp:45); on
// Linux that is a file with backslashes in its name in the cwd: close and remove it.
struct StaticInitCleanup {
    StaticInitCleanup() {
        outfile.close();
        std::remove("..\\results\\20210709_balls.ppm");
    }
} g_cleanup;

material** material_slot(hitable* h) {
    if (sphere* s = dynamic_cast<sphere*>(h)) return &s->mat_ptr;
    if (env_sphere* s = dynamic_cast<env_sphere*>(h)) return &s->mat_ptr;
    if (moving_sphere* s = dynamic_cast<moving_sphere*>(h)) return &s->mat_ptr;
    if (xy_rect* r = dynamic_cast<xy_rect*>(h)) return &r->mp;
    if (xz_rect* r = dynamic_cast<xz_rect*>(h)) return &r->mp;
    if (yz_rect* r = dynamic_cast<yz_rect*>(h)) return &r->mp;
    if (triangle* t = dynamic_cast<triangle*>(h)) return &t->mp;
    return nullptr;
}

// counts world->hit calls made by color() (Raytracing_n.cpp:58): the definition of a "ray"
class counting_world : public hitable {
   public:
    counting_world(hitable* w) : w(w), rays(0) {}
    virtual bool hit(const ray& r, float t_min, float t_max, hit_record& rec, bool is_medium = false) const {
        rays++;
        return w->hit(r, t_min, t_max, rec, is_medium);
    }
    virtual bool bounding_box(float t0, float t1, aabb& box) const { return w->bounding_box(t0, t1, box); }
    hitable* w;
    mutable uint64_t rays;
};

}  // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

int ref_has_reference_tree() { return access(g_sobol_file.c_str(), R_OK) == 0; }

void ref_set_sobol_file(const char* path) { g_sobol_file = path; }

// ---- RNG state (mathf.h:12, pdf.h:20) -----------------------------------------------------
void ref_set_seed(uint64_t s) { seed = s; }
uint64_t ref_get_seed() { return seed; }
void ref_reset_pcg() { s_rng = rng(); }

// ---- known-answer values straight out of the reference ------------------------------------
void ref_kat_drand48(uint64_t s, int n, double* out) {
    uint64_t keep = seed;
    seed = s;
    for (int i = 0; i < n; i++) out[i] = ref_drand48();
    seed = keep;
}

void ref_kat_pcg32(int n, float* out_float, uint32_t* out_u32) {
    rng a, b;
    for (int i = 0; i < n; i++) {
        out_float[i] = a.UniformFloat();
        out_u32[i] = b.UniformUInt32();
    }
}

// Raytracing_n.cpp:721-812 with the reference's own file; out: [N][2]
int ref_kat_sobol(unsigned N, double* out) {
    if (access(g_sobol_file.c_str(), R_OK) != 0) {
        g_err = "direction-number file not present";
        return -1;
    }
    double** p = sobol_points(N, 2, g_sobol_file.c_str());
    for (unsigned i = 0; i < N; i++) {
        out[2 * i] = p[i][0];
        out[2 * i + 1] = p[i][1];
    }
    free_sobol(p, N);
    return 0;
}

void ref_sobol_fallback(unsigned N, double* out) {
    double** p = new double*[N];
    for (unsigned i = 0; i < N; i++) p[i] = new double[2];
    sobol2_fallback(N, p);
    for (unsigned i = 0; i < N; i++) {
        out[2 * i] = p[i][0];
        out[2 * i + 1] = p[i][1];
    }
    free_sobol(p, N);
}

// What beckmann_pdf::value returns before any generate() in THIS process: *pdf_value is a malloc(4)
// block only generate() writes (pdf.h:121-135), so when the mixture picks the light branch first the
// reference reads what the allocator left there — glibc's safe-linking residue of the block the
// previous beckmann_pdf freed, (chunk address >> 12) ^ next.  Its value as a float depends on where
// the heap lives: a low brk heap (non-PIE host such as this Python) gives a denormal (~0: the sample
// counts in full), a PIE binary's 0x55.. heap gives ~1e13 (the sample counts for nothing).  The
// tests read this probe so that the stand-in the backend uses (0) is checked against
// what the oracle process really observes.
float ref_probe_beckmann_stash(int) {
    BeckmannDistribution d(0.1f, 0.1f, true);
    float v = 0;
    for (int i = 0; i < 3; i++) {
        beckmann_pdf* p = new beckmann_pdf(&d, vec3(0, 1, 0));
        v = p->value(vec3(0, -1, 0), vec3(0, 1, 0));
        delete p;
    }
    return v;
}
float ref_kat_erf(float x) { return Erf(x); }
float ref_kat_erfinv(float x) { return ErfInv(x); }
float ref_kat_schlick(float c, float ri) { return schlick(c, ri); }
float ref_kat_roughness_to_alpha(float r) { return BeckmannDistribution::RoughnessToAlpha(r); }

// ---- object constructors: one per reference class -------------------------------------------
int ref_tex_constant(float r, float g, float b) { return put(K_TEX, new constant_texture(vec3(r, g, b))); }
int ref_tex_checker(int t0, int t1) {
    return put(K_TEX, new checker_texture(get<texture>(t0, K_TEX), get<texture>(t1, K_TEX)));
}
int ref_tex_noise(float scale) { return put(K_TEX, new noise_texture(scale)); }
int ref_tex_image(const unsigned char* data, int nx_, int ny_) {
    unsigned char* copy = new unsigned char[(size_t)nx_ * ny_ * 3];
    memcpy(copy, data, (size_t)nx_ * ny_ * 3);
    return put(K_TEX, new image_texture(copy, nx_, ny_));
}

int ref_mat_lambertian(int tex) { return put(K_MAT, new lambertian(get<texture>(tex, K_TEX))); }
int ref_mat_orennayar(int tex, float sigma) { return put(K_MAT, new orennayar(get<texture>(tex, K_TEX), sigma)); }
int ref_mat_beckmann(int tex, float rx, float ry) { return put(K_MAT, new beckmann(get<texture>(tex, K_TEX), rx, ry)); }
int ref_mat_metal(float r, float g, float b, float fuzz) { return put(K_MAT, new metal(vec3(r, g, b), fuzz)); }
int ref_mat_dielectric(float ri) { return put(K_MAT, new dielectric(ri)); }
int ref_mat_diffuse_light(int tex) { return put(K_MAT, new diffuse_light(get<texture>(tex, K_TEX))); }
int ref_mat_isotropic(int tex) { return put(K_MAT, new isotropic(get<texture>(tex, K_TEX))); }
// brdfmaterial::scatter (material.h:214-236) returns a non-specular record WITHOUT setting
// srec.pdf_ptr (the assignment at :234 is commented out): color() then calls generate / value through
// an uninitialised pointer and deletes it — the reference crashes (or reuses a freed cosine_pdf) on
// the first hit.  To have a defined oracle the harness completes the record with the cosine_pdf every
// other diffuse material supplies (material.h:109); everything else, including the discarded MERL
// lookup, is the reference's own scatter().  The backend documents the same stand-in.
class brdfmaterial_defined : public brdfmaterial {
   public:
    brdfmaterial_defined(const char* file, const vec3& a) : brdfmaterial(file, a) {}
    virtual bool scatter(const ray& r_in, const hit_record& hrec, scatter_record& srec) const {
        bool ok = brdfmaterial::scatter(r_in, hrec, srec);
        srec.pdf_ptr = new cosine_pdf(hrec.normal);
        return ok;
    }
};

int ref_mat_brdf(const char* file, float r, float g, float b) {
    CoutSilencer quiet;
    brdfmaterial* m = new brdfmaterial_defined(file, vec3(r, g, b));
    return put(K_MAT, m);
}

// brdf::lookup_brdf_val (brdf.h:190-214) on the table a brdfmaterial loaded: angles [n][4] =
// theta_in, fi_in, theta_out, fi_out; rgb [n][3].
int ref_merl_lookup(int mat, const double* angles, int64_t n, double* rgb) {
    brdfmaterial* m = dynamic_cast<brdfmaterial*>(get<material>(mat, K_MAT));
    if (!m) return -1;
    CoutSilencer quiet;  // "Below horizon." (brdf.h:211-212)
    for (int64_t i = 0; i < n; i++)
        m->brdf_reader->lookup_brdf_val(m->brdf_data, angles[4 * i], angles[4 * i + 1], angles[4 * i + 2],
                                        angles[4 * i + 3], rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
    return 0;
}

int ref_sphere(float x, float y, float z, float r, int mat) {
    return put(K_HIT, new sphere(vec3(x, y, z), r, get<material>(mat, K_MAT)));
}
int ref_env_sphere(float x, float y, float z, float r, int mat) {
    return put(K_HIT, new env_sphere(vec3(x, y, z), r, get<material>(mat, K_MAT)));
}
int ref_moving_sphere(const float* c0, const float* c1, float t0, float t1, float r, int mat) {
    return put(K_HIT, new moving_sphere(vec3(c0), vec3(c1), t0, t1, r, get<material>(mat, K_MAT)));
}
int ref_xy_rect(float a0, float a1, float b0, float b1, float k, int mat) {
    return put(K_HIT, new xy_rect(a0, a1, b0, b1, k, get<material>(mat, K_MAT)));
}
int ref_xz_rect(float a0, float a1, float b0, float b1, float k, int mat) {
    return put(K_HIT, new xz_rect(a0, a1, b0, b1, k, get<material>(mat, K_MAT)));
}
int ref_yz_rect(float a0, float a1, float b0, float b1, float k, int mat) {
    return put(K_HIT, new yz_rect(a0, a1, b0, b1, k, get<material>(mat, K_MAT)));
}
// pos: p0 p1 p2 (9); uv (may be null): 6; nrm (may be null): 9.  With nrm == null the 4-arg /
// 7-arg constructor is used and n0..n2 are then set to the geometric normal through the public
// members (the constructor leaves them uninitialised, triangle.h:13-36).
int ref_triangle(const float* pos, const float* uv, const float* nrm, int mat) {
    vec3 p0(pos), p1(pos + 3), p2(pos + 6);
    material* m = get<material>(mat, K_MAT);
    vec3 u0(0.0f), u1(0.0f), u2(0.0f);
    if (uv) {
        u0 = vec3(uv[0], uv[1], 0);
        u1 = vec3(uv[2], uv[3], 0);
        u2 = vec3(uv[4], uv[5], 0);
    }
    triangle* t;
    if (nrm) {
        t = new triangle(p0, p1, p2, m, u0, u1, u2, vec3(nrm), vec3(nrm + 3), vec3(nrm + 6));
    } else {
        t = uv ? new triangle(p0, p1, p2, m, u0, u1, u2) : new triangle(p0, p1, p2, m);
        t->n0 = t->n1 = t->n2 = t->normal;
        t->normals[0] = t->normals[1] = t->normals[2] = t->normal;
    }
    return put(K_HIT, t);
}
int ref_box(const float* p0, const float* p1, int mat) {
    return put(K_HIT, new box(vec3(p0), vec3(p1), get<material>(mat, K_MAT)));
}
int ref_flip_normals(int h) { return put(K_HIT, new flip_normals(get<hitable>(h, K_HIT))); }
int ref_translate(int h, float x, float y, float z) {
    return put(K_HIT, new translate(get<hitable>(h, K_HIT), vec3(x, y, z)));
}
int ref_rotate_y(int h, float angle) { return put(K_HIT, new rotate_y(get<hitable>(h, K_HIT), angle)); }
int ref_rotate_x(int h, float angle) { return put(K_HIT, new rotate_x(get<hitable>(h, K_HIT), angle)); }
int ref_constant_medium(int h, float density, int tex) {
    return put(K_HIT, new constant_medium(get<hitable>(h, K_HIT), density, get<texture>(tex, K_TEX)));
}
int ref_hitable_list(const int* hs, int n, int list_size) {
    hitable** l = new hitable*[n > 0 ? n : 1];
    for (int i = 0; i < n; i++) l[i] = get<hitable>(hs[i], K_HIT);
    return put(K_HIT, new hitable_list(l, list_size));
}
// bvh.h:96-119: consumes the global drand48 stream and sorts its argument array in place
int ref_bvh_node(const int* hs, int n, float t0, float t1) {
    hitable** l = new hitable*[n > 0 ? n : 1];
    for (int i = 0; i < n; i++) l[i] = get<hitable>(hs[i], K_HIT);
    return put(K_HIT, new bvh_node(l, n, t0, t1));
}
// `new bvh_node(triangles, n, 0, 1)` over n real triangle objects built from arrays in one call (what
// RefBuilder does one ref_triangle() at a time; 10 M Python round trips are too slow).  pos [n][9],
// uv [n][6], nrm [n][9].  The objects are not registered one by one; prim ids come from the flattener.
int ref_triangle_soup_bvh(const float* pos, const float* uv, const float* nrm, int64_t n, int mat) {
    material* m = get<material>(mat, K_MAT);
    hitable** l = new hitable*[n > 0 ? n : 1];
    for (int64_t i = 0; i < n; i++) {
        const float* p = pos + 9 * i;
        const float* u = uv + 6 * i;
        const float* q = nrm + 9 * i;
        l[i] = new triangle(vec3(p), vec3(p + 3), vec3(p + 6), m, vec3(u[0], u[1], 0), vec3(u[2], u[3], 0),
                            vec3(u[4], u[5], 0), vec3(q), vec3(q + 3), vec3(q + 6));
    }
    return put(K_HIT, new bvh_node(l, (int)n, 0, 1));
}
// teapot.h:76-166 -> bvh_node over its triangles (Raytracing_n.cpp:348-350).  triangleCount is
// zeroed and the triangles' n0..n2 set to the face normal through public members first.
int ref_teapot_bvh(float scale, int mat, int* n_tris_out) {
    teapot* tp = new teapot(scale, get<material>(mat, K_MAT));
    tp->triangleCount = 0;
    hitable** l = tp->createPloyTeapot();
    int n = tp->getTriangleCount();
    for (int i = 0; i < n; i++) {
        triangle* t = (triangle*)l[i];
        t->n0 = t->n1 = t->n2 = t->normal;
        t->normals[0] = t->normals[1] = t->normals[2] = t->normal;
    }
    if (n_tris_out) *n_tris_out = n;
    return put(K_HIT, new bvh_node(l, n, 0, 1));
}
// the teapot's triangle soup in construction order (before any bvh sort): pos [n][9]
int ref_teapot_triangles(float scale, float* pos_out, int cap) {
    teapot* tp = new teapot(scale, nullptr);
    tp->triangleCount = 0;
    hitable** l = tp->createPloyTeapot();
    int n = tp->getTriangleCount();
    for (int i = 0; i < n && i < cap; i++) {
        triangle* t = (triangle*)l[i];
        srt_flatten_detail::put3(pos_out + 9 * i, t->p0);
        srt_flatten_detail::put3(pos_out + 9 * i + 3, t->p1);
        srt_flatten_detail::put3(pos_out + 9 * i + 6, t->p2);
    }
    return n;
}

int ref_camera(const float* lookfrom, const float* lookat, const float* vup, float vfov, float aspect,
               float aperture, float focus, float t0, float t1) {
    return put(K_CAM, new camera(vec3(lookfrom), vec3(lookat), vec3(vup), vfov, aspect, aperture, focus, t0, t1));
}

int ref_scene(int world, int hlist, int cam) {
    Scene* s = new Scene();
    s->world = get<hitable>(world, K_HIT);
    s->hlist = get<hitable>(hlist, K_HIT);
    s->cam = get<camera>(cam, K_CAM);
    s->flattened = false;
    s->flatten_status = 0;
    if (!s->world || !s->cam) {
        g_err = "ref_scene: null world or camera";
        return -1;
    }
    return put(K_SCENE, s);
}

// The reference's own scene functions (Raytracing_n.cpp:216-711) that need no model file.  They
// load "..\contents\environment_map\sky_2.png" by a Windows-literal relative path: the caller
// chdir()s into a scratch directory holding a file of exactly that name (tests create a symlink
// to the reference's own image there).  name: "ball_scenes" (:379, the compiled-in default,
// sceneid = 2), "ball_orennayar_scenes" (:427).
int ref_scene_named(const char* name, float aspect) {
    hitable *world = nullptr, *hl = nullptr;
    camera* cam = nullptr;
    std::string n(name);
    if (n == "ball_scenes")
        ball_scenes(&world, &cam, &hl, aspect);
    else if (n == "ball_orennayar_scenes")
        ball_orennayar_scenes(&world, &cam, &hl, aspect);
    else {
        g_err = "unknown scene function";
        return -1;
    }
    Scene* s = new Scene();
    s->world = world;
    s->hlist = hl;
    s->cam = cam;
    s->flattened = false;
    s->flatten_status = 0;
    return put(K_SCENE, s);
}

// Flatten with the product's C++ flattener; the returned pointer stays valid for the scene's life.
const SrtSceneDesc* ref_scene_desc(int scene) {
    Scene* s = get<Scene>(scene, K_SCENE);
    if (!s) return nullptr;
    if (!s->flattened) {
        s->flatten_status = srt_flatten_scene(s->world, s->hlist, s->cam, s->flat);
        s->desc = s->flat.desc();
        s->flattened = true;
    }
    if (s->flatten_status != 0) {
        g_err = s->flat.error;
        return nullptr;
    }
    return &s->desc;
}

// ---- camera rays: camera::get_ray (camera.h:51-59) over the sample loop's (u, v)
//      (Raytracing_n.cpp:827-836).  rays: [ny*nx*ns][7] in (curpixel, s) order. ------------------
int ref_primary_rays(int scene, int nx_, int ny_, int ns_, uint64_t rng_seed, int mod_ny, float* rays) {
    Scene* s = get<Scene>(scene, K_SCENE);
    if (!s) return -1;
    double** pts = make_sobol(ns_);
    seed = rng_seed;
    size_t k = 0;
    for (int cur = 0; cur < nx_ * ny_; cur++) {
        int i = cur % (mod_ny ? ny_ : nx_);
        int j = ny_ - 1 - (cur / nx_);
        for (int sm = 0; sm < ns_; sm++) {
            float u = float(pts[sm][0] + i) / float(nx_);
            float v = float(pts[sm][1] + j) / float(ny_);
            ray r = s->cam->get_ray(u, v);
            float* o = rays + 7 * k++;
            o[0] = r.A.e[0];
            o[1] = r.A.e[1];
            o[2] = r.A.e[2];
            o[3] = r.B.e[0];
            o[4] = r.B.e[1];
            o[5] = r.B.e[2];
            o[6] = r._time;
        }
    }
    free_sobol(pts, ns_);
    return 0;
}

// ---- closest hit: world->hit(r, 0.001, FLT_MAX, rec) exactly as color() calls it
//      (Raytracing_n.cpp:58).  ids: SrtSceneDesc.prims index, n_prims + medium index for a
//      constant_medium scatter, -1 on miss.  rec: [n][8] = u v p normal (may be null). ----------
int ref_trace(int scene, const float* rays, int64_t n, uint64_t rng_seed, int32_t* ids, float* t, float* rec) {
    Scene* s = get<Scene>(scene, K_SCENE);
    if (!s || !ref_scene_desc(scene)) return -1;
    SrtFlatScene& fs = s->flat;
    size_t np = fs.prim_objects.size(), nm = fs.medium_objects.size();
    std::vector<tag_material> tags(np + nm);
    std::vector<material*> saved(np + nm);
    // prims under a medium are never reported; tag them first so that a shared object ends up
    // carrying the id of its directly visible instance
    std::vector<char> under_medium(np, 0);
    for (const SrtEntry& e : fs.entries) {
        if (e.medium < 0) continue;
        if (e.geom_kind == SRT_GEOM_GROUP) {
            const SrtGroup& g = fs.groups[e.geom_index];
            for (int i = 0; i < g.prim_count; i++) under_medium[g.first_prim + i] = 1;
        } else {
            const SrtBvh& b = fs.bvhs[e.geom_index];
            for (int k = 0; k < b.item_count; k++) {
                const SrtGroup& g = fs.groups[fs.bvh_items[b.first_item + k]];
                for (int i = 0; i < g.prim_count; i++) under_medium[g.first_prim + i] = 1;
            }
        }
    }
    std::vector<size_t> order;
    for (size_t i = 0; i < np; i++)
        if (under_medium[i]) order.push_back(i);
    for (size_t i = 0; i < np; i++)
        if (!under_medium[i]) order.push_back(i);
    // save in forward order of `order`, restore in reverse, so shared objects get their own back
    for (size_t k = 0; k < np; k++) {
        size_t i = order[k];
        material** slot = material_slot(const_cast<hitable*>(fs.prim_objects[i]));
        saved[i] = *slot;
        *slot = &tags[i];
    }
    for (size_t m = 0; m < nm; m++) {
        constant_medium* cm = const_cast<constant_medium*>((const constant_medium*)fs.medium_objects[m]);
        saved[np + m] = cm->phase_function;
        cm->phase_function = &tags[np + m];
    }
    seed = rng_seed;
    for (int64_t i = 0; i < n; i++) {
        const float* o = rays + 7 * i;
        ray r(vec3(o[0], o[1], o[2]), vec3(o[3], o[4], o[5]), o[6]);
        hit_record hr;
        if (s->world->hit(r, 0.001, numeric_limits<float>::max(), hr)) {
            ids[i] = (int32_t)((tag_material*)hr.mat_ptr - tags.data());
            t[i] = hr.t;
            if (rec) {
                float* q = rec + 8 * i;
                q[0] = hr.u;
                q[1] = hr.v;
                q[2] = hr.p.e[0];
                q[3] = hr.p.e[1];
                q[4] = hr.p.e[2];
                q[5] = hr.normal.e[0];
                q[6] = hr.normal.e[1];
                q[7] = hr.normal.e[2];
            }
        } else {
            ids[i] = -1;
            t[i] = 0;
            if (rec) memset(rec + 8 * i, 0, 8 * sizeof(float));
        }
    }
    for (size_t m = nm; m-- > 0;) {
        constant_medium* cm = const_cast<constant_medium*>((const constant_medium*)fs.medium_objects[m]);
        cm->phase_function = saved[np + m];
    }
    for (size_t k = np; k-- > 0;) {
        size_t i = order[k];
        *material_slot(const_cast<hitable*>(fs.prim_objects[i])) = saved[i];
    }
    return 0;
}

// ---- light probe: hitable_pdf::value -> hitable_list::pdf_value (pdf.h:162-164,
//      hitable_list.h:54-62).  od: [n][6]. ----------------------------------------------------
int ref_light_pdf(int scene, const float* od, int64_t n, float* out) {
    Scene* s = get<Scene>(scene, K_SCENE);
    if (!s || !s->hlist) return -1;
    for (int64_t i = 0; i < n; i++) {
        const float* q = od + 6 * i;
        out[i] = s->hlist->pdf_value(vec3(q[0], q[1], q[2]), vec3(q[3], q[4], q[5]));
    }
    return 0;
}

// ---- light sampling: hitable_pdf::generate -> hitable_list::random (pdf.h:166-168, hitable_list.h:64-67)
//      with the reference's own drand48 stream.  o: [n][3]; out: [n][3] directions. -----------------
int ref_light_random(int scene, const float* o, int64_t n, uint64_t rng_seed, float* out) {
    Scene* s = get<Scene>(scene, K_SCENE);
    if (!s || !s->hlist) return -1;
    seed = rng_seed;
    for (int64_t i = 0; i < n; i++) {
        vec3 d = s->hlist->random(vec3(o[3 * i], o[3 * i + 1], o[3 * i + 2]));
        out[3 * i] = d.e[0];
        out[3 * i + 1] = d.e[1];
        out[3 * i + 2] = d.e[2];
    }
    return 0;
}

// ---- float image: the sample loop of renderthread (Raytracing_n.cpp:826-848) restated without
//      the 8-bit quantise, calling the reference's own color() / de_nan(); single-threaded.
//      out: [ny*nx][3] per-pixel SUMS over samples [s_begin, s_end) (divide by the count for the
//      mean).  rays_out (optional): number of world->hit calls made by color(). ----------------
int ref_render_float(int scene, int nx_, int ny_, int ns_, int s_begin, int s_end, int max_depth,
                     uint64_t rng_seed, int mod_ny, float* out, uint64_t* rays_out) {
    std::lock_guard<std::mutex> lk(g_api_lock);
    Scene* s = get<Scene>(scene, K_SCENE);
    if (!s) return -1;
    double** pts = make_sobol(ns_);
    int keep_depth = maxDepth;
    maxDepth = max_depth;
    seed = rng_seed;
    s_rng = rng();
    counting_world cw(s->world);
    for (int cur = 0; cur < nx_ * ny_; cur++) {
        vec3 col(0, 0, 0);
        int i = cur % (mod_ny ? ny_ : nx_);
        int j = ny_ - 1 - (cur / nx_);
        for (int sm = s_begin; sm < s_end; sm++) {
            float u = float(pts[sm][0] + i) / float(nx_);
            float v = float(pts[sm][1] + j) / float(ny_);
            ray r = s->cam->get_ray(u, v);
            int depth = 0;
            col += de_nan(color(r, &cw, s->hlist, &depth));
        }
        out[3 * cur] = col.e[0];
        out[3 * cur + 1] = col.e[1];
        out[3 * cur + 2] = col.e[2];
    }
    if (rays_out) *rays_out = cw.rays;
    maxDepth = keep_depth;
    free_sobol(pts, ns_);
    return 0;
}

// ---- the timed baseline: the reference's OWN renderthread (Raytracing_n.cpp:815-879) on
//      `threads` std::threads exactly as main() does (:926-946), wall clock around the pool.
//      rgb8_out (optional): [ny*nx][3] the 8-bit values it stores in `colors`. -------------------
double ref_render_timed(int scene, int nx_, int ny_, int ns_, int max_depth, int threads, uint64_t rng_seed,
                        int32_t* rgb_out) {
    std::lock_guard<std::mutex> lk(g_api_lock);
    Scene* s = get<Scene>(scene, K_SCENE);
    if (!s) return -1.0;
    CoutSilencer quiet;
    double** pts = make_sobol(ns_);
    int keep_nx = nx, keep_ny = ny, keep_ns = ns, keep_depth = maxDepth;
    nx = nx_;
    ny = ny_;
    ns = ns_;
    maxDepth = max_depth;
    seed = rng_seed;
    s_rng = rng();
    finishedPixel = 0;
    isfinished = true;  // suppresses the text dump at :869-878
    // renderthread tests `finishedPixel < nx*ny` outside the lock (:817), so up to threads-1 workers
    // take a curpixel >= nx*ny at the end and store colors[curpixel] past the array (:863): give the
    // array that much slack instead of touching the reference
    size_t n_slots = (size_t)nx_ * ny_ + (size_t)threads + 8;
    colors = new int*[n_slots];
    for (size_t i = 0; i < n_slots; i++) colors[i] = nullptr;
    auto t0 = std::chrono::steady_clock::now();
    {
        std::vector<std::thread> pool;
        for (int i = 0; i < threads; i++) pool.emplace_back(renderthread, s->world, s->hlist, s->cam, pts);
        for (auto& th : pool) th.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    for (size_t i = 0; i < (size_t)nx_ * ny_; i++) {
        if (colors[i]) {
            if (rgb_out) {
                rgb_out[3 * i] = colors[i][0];
                rgb_out[3 * i + 1] = colors[i][1];
                rgb_out[3 * i + 2] = colors[i][2];
            }
            delete[] colors[i];
        }
    }
    for (size_t i = (size_t)nx_ * ny_; i < n_slots; i++) delete[] colors[i];
    delete[] colors;
    colors = nullptr;
    nx = keep_nx;
    ny = keep_ny;
    ns = keep_ns;
    maxDepth = keep_depth;
    free_sobol(pts, ns_);
    return std::chrono::duration<double>(t1 - t0).count();
}

// ---- CRender (Render.h:16-51) defined by the product's host side (host/crender.hpp) on top of
//      libsrt_b200.so: construct + Run on one of the reference scenes.  Returns 0 / negative. ---
int ref_crender_run(int scene, int max_depth, int width, int height, int num_ray, const char* lib_path,
                    const char* file_path) {
    Scene* s = get<Scene>(scene, K_SCENE);
    if (!s) return -1;
    if (srt_host::load_backend(lib_path) != 0) {
        g_err = srt_host::backend_error();
        return -2;
    }
    CRender r(max_depth, width, height, num_ray, 1, const_cast<char*>(file_path));
    srt_host::set_scene(s->world, s->hlist, s->cam);
    r.Run();
    if (srt_host::last_status() != 0) {
        g_err = srt_host::backend_error();
        return -3;
    }
    return 0;
}

}  // extern "C"
