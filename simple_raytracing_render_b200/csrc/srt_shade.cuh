// srt_shade.cuh — textures, materials, light sampling and one bounce of color() (the reference's
// Raytracing_n.cpp:55-106), with the reference's sampling semantics kept as
// they are, quirks included.
//
// Numerics: nothing in this file decides WHICH primitive a ray hits or its t, and the image gate is
// statistical (relMSE against the reference's converged render), so shading uses fused / fast float
// arithmetic (namespace fm) and float constants where the reference promotes to double.  What is kept
// exactly is the STRUCTURE: which random numbers are drawn in which branch, the mixture density, the
// wrong-hemisphere cosine generator and the `while (pdf_val == 0)` rejection loop, one-sided
// emission, the specular branch that drops `emitted`, de_nan.
#ifndef SRT_SHADE_CUH
#define SRT_SHADE_CUH

#include "srt_intersect.cuh"
#include "srt_microfacet.cuh"

namespace srt {

// ---- textures (texture.h, perlin.h) -------------------------------------------------------------
// perlin.h:7-58
SRT_DEV float perlin_noise(const DScene& S, V3 p) {
    float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
    float u = p.x - fx, v = p.y - fy, w = p.z - fz;
    int i = (int)fx, j = (int)fy, k = (int)fz;
    float uu = u * u * (3.f - 2.f * u), vv = v * v * (3.f - 2.f * v), ww = w * w * (3.f - 2.f * w);
    float accum = 0.f;
#pragma unroll
    for (int di = 0; di < 2; di++)
#pragma unroll
        for (int dj = 0; dj < 2; dj++)
#pragma unroll
            for (int dk = 0; dk < 2; dk++) {
                int idx = S.perlin_perm[(i + di) & 255] ^ S.perlin_perm[256 + ((j + dj) & 255)] ^
                          S.perlin_perm[512 + ((k + dk) & 255)];
                const float* c = S.perlin_ranvec + 3 * idx;
                float wx = u - di, wy = v - dj, wz = w - dk;
                accum += (di * uu + (1 - di) * (1.f - uu)) * (dj * vv + (1 - dj) * (1.f - vv)) *
                         (dk * ww + (1 - dk) * (1.f - ww)) * (c[0] * wx + c[1] * wy + c[2] * wz);
            }
    return accum;
}

SRT_DEV float perlin_turb(const DScene& S, V3 p) {  // perlin.h:48-58, depth 7
    float accum = 0.f, weight = 1.f;
    for (int i = 0; i < 7; i++) {
        accum += weight * perlin_noise(S, p);
        weight *= 0.5f;
        p = fm::scale(p, 2.f);
    }
    return fabsf(accum);
}

template <uint32_t F>
SRT_DEV V3 texture_value(const DScene& S, int tex, float u, float v, V3 p) {
    for (int guard = 0; guard < 16 && tex >= 0; guard++) {
        SrtTexture t = S.textures[tex];
        if (t.type == SRT_TEX_CONSTANT) return v3(t.c[0], t.c[1], t.c[2]);  // texture.h:29-31
        if ((F & F_TEX_CHECKER) && t.type == SRT_TEX_CHECKER) {              // texture.h:13-19
            float sines = sinf(10.f * p.x) * sinf(10.f * p.y) * sinf(10.f * p.z);
            tex = sines < 0.f ? t.b : t.a;
            continue;
        }
        if ((F & F_TEX_NOISE) && t.type == SRT_TEX_NOISE) {  // texture.h:39-43
            float s = t.c[0];
            float val = 0.5f * (1.0f + sinf(s * p.z + 5.f * perlin_turb(S, fm::scale(p, s))));
            return v3(val, val, val);
        }
        if ((F & F_TEX_IMAGE) && t.type == SRT_TEX_IMAGE) {  // texture.h:58-70: nearest texel, v flipped, 3 channels
            DImage im = S.images[t.a];
            int i = (int)(u * (float)im.nx);
            int j = (int)((double)((1.f - v) * (float)im.ny) - 0.001);
            if (i < 0) i = 0;
            if (j < 0) j = 0;
            if (i > im.nx - 1) i = im.nx - 1;
            if (j > im.ny - 1) j = im.ny - 1;
            const uint8_t* px = im.data + 3 * (size_t)i + 3 * (size_t)im.nx * j;
            const float k = 1.0f / 255.0f;
            return v3((float)px[0] * k, (float)px[1] * k, (float)px[2] * k);
        }
        break;
    }
    return v3(0.f, 0.f, 0.f);
}

// albedo / emission of a material at a hit: the constant-texture case is folded into DMaterial
template <uint32_t F>
SRT_DEV V3 material_color(const DScene& S, const DMaterial& m, const HitRecord& rec) {
    if (m.tex_type == SRT_TEX_CONSTANT) return v3(m.c[0], m.c[1], m.c[2]);
    if (F & (F_TEX_CHECKER | F_TEX_NOISE | F_TEX_IMAGE)) return texture_value<F>(S, m.texture, rec.u, rec.v, rec.p);
    return v3(0.f, 0.f, 0.f);
}

template <uint32_t F>
SRT_DEV bool material_needs_uv(const DMaterial& m) {
    // a checker may hold an image (texture.h:9-23)
    return (F & F_TEX_IMAGE) && (m.tex_type == SRT_TEX_IMAGE || m.tex_type == SRT_TEX_CHECKER);
}

SRT_DEV DMaterial load_material(const DMaterial* p) {
    const float4* q = reinterpret_cast<const float4*>(p);
    float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    DMaterial m;
    m.type = __float_as_int(a.x);
    m.texture = __float_as_int(a.y);
    m.tex_type = __float_as_int(a.z);
    m._pad = 0;
    m.p[0] = b.x;
    m.p[1] = b.y;
    m.p[2] = b.z;
    m.p[3] = b.w;
    m.c[0] = c.x;
    m.c[1] = c.y;
    m.c[2] = c.z;
    m._pad2 = 0.f;
    return m;
}

// ---- random helpers (material.h:43-50, camera.h:8-14, pdf.h:10-18, sphere.h:7-15) ----------------
SRT_DEV V3 random_in_unit_sphere(Philox& rng) {
    V3 p;
    do {
        p = v3(2.0f * rand_f(rng) - 1.f, 2.0f * rand_f(rng) - 1.f, 2.0f * rand_f(rng) - 1.f);
    } while (fm::dot(p, p) >= 1.0f);
    return p;
}

SRT_DEV V3 random_cosine_direction(Philox& rng) {  // pdf.h:10-18, the `2 *` on x, y included
    float r1 = rand_f(rng), r2 = rand_f(rng);
    float z = sqrtf(1.f - r2);
    float sn, cs;
    __sincosf(fm::kTwoPi * r1, &sn, &cs);
    float sq = 2.f * sqrtf(r2);
    return v3(cs * sq, sn * sq, z);
}

SRT_DEV V3 random_to_sphere(float radius, float distance_squared, Philox& rng) {
    float r1 = rand_f(rng), r2 = rand_f(rng);
    float z = 1.f + r2 * (sqrtf(1.f - radius * radius / distance_squared) - 1.f);
    float sn, cs;
    __sincosf(fm::kTwoPi * r1, &sn, &cs);
    float s = sqrtf(1.f - z * z);
    return v3(cs * s, sn * s, z);
}

// onb.h:21-30 in fast arithmetic (shading frames only)
SRT_DEV Onb onb_fast(V3 n) {
    Onb b;
    b.w = fm::normalize(n);
    V3 a = fabsf(b.w.x) > 0.9f ? v3(0.f, 1.f, 0.f) : v3(1.f, 0.f, 0.f);
    b.v = fm::normalize(fm::cross(b.w, a));
    b.u = fm::cross(b.w, b.v);
    return b;
}
SRT_DEV V3 onb_local_fast(const Onb& b, V3 a) {
    return fm::madd(b.u, a.x, fm::madd(b.v, a.y, fm::scale(b.w, a.z)));
}

// ---- light list: hitable_list::random / pdf_value over hlist (hitable_list.h:54-67) ----------------
// CORR (corrected-transport mode): the triangle is sampled uniformly — triangle::random draws v = drand48() * (1 - u),
// which crowds the points towards p1 (density 1 / (1 - u), triangle.h:89-94) while pdf_value assumes 1 / area.
template <uint32_t F, bool CORR = false>
SRT_DEV V3 light_random(const View& W, V3 o, Philox& rng) {
    const DScene& S = W.S;
    int index = 0;
    float pick = rand_f(rng);  // hitable_list.h:65: int(drand48() * list_size), drawn even for one light
    if (S.n_lights > 1) {
        index = (int)(pick * (float)S.n_lights);
        if (index >= S.n_lights) index = S.n_lights - 1;
    }
    DPrim pr = load_prim(W.lights + index);
    if (pr.type == SRT_PRIM_XZ_RECT) {  // aarect.h:57-60
        float x = fmaf(rand_f(rng), pr.g.y - pr.g.x, pr.g.x);
        float z = fmaf(rand_f(rng), pr.g.w - pr.g.z, pr.g.z);
        return fm::sub(v3(x, pr.k, z), o);
    }
    if ((F & (F_LIGHT_SPHERE | F_LIGHT_ENV)) && (pr.type == SRT_PRIM_SPHERE || pr.type == SRT_PRIM_ENV_SPHERE)) {
        // sphere.h:80-86 / env_sphere.h:40-46 (direction = o - center for the env sphere)
        V3 c = v3(pr.g.x, pr.g.y, pr.g.z);
        V3 direction = pr.type == SRT_PRIM_SPHERE ? fm::sub(c, o) : fm::sub(o, c);
        float d2 = fm::dot(direction, direction);
        Onb uvw = onb_fast(direction);
        return onb_local_fast(uvw, random_to_sphere(pr.g.w, d2, rng));
    }
    if ((F & F_LIGHT_TRI) && pr.type == SRT_PRIM_TRIANGLE) {  // triangle.h:89-94
        const float* p = S.tri_pos + 9 * (size_t)(pr.flags_aux >> 1);
        float u = rand_f(rng);
        float v = rand_f(rng) * (1.f - u);
        if (CORR) {  // uniform: b1 = 1 - sqrt(u1) on p1, b2 = u2 sqrt(u1) on p2
            const float su = sqrtf(u);
            v = v / fmaxf(1.f - u, 1e-20f) * su;
            u = 1.f - su;
        }
        V3 pt = fm::madd(v3(p), 1.f - u - v, fm::madd(v3(p + 3), u, fm::scale(v3(p + 6), v)));
        return fm::sub(pt, o);
    }
    return v3(1.f, 0.f, 0.f);  // hitable.h:32
}

// hitable_pdf::value -> hitable_list::pdf_value (pdf.h:162-164, hitable_list.h:54-62): each light
// shape re-intersects the ray (o, v) with (0.001, FLT_MAX) — these ARE hit tests of the reference
// shapes and use the exact routines, so that srt_light_pdf reproduces the reference's zero / non-zero
// pattern bit for bit; the density arithmetic after the hit is in fast float.
// CORR (corrected-transport mode): triangle::hit reports t as a DISTANCE (triangle.h:128), so distance^2 is t^2; the
// reference multiplies it by |v|^2 once more (triangle.h:80), which is only right for unit directions.
template <uint32_t F, bool COUNT, bool CORR = false>
SRT_DEV float light_pdf_value(const View& W, V3 o, V3 v, Tally<COUNT> tl) {
    const DScene& S = W.S;
    float weight = 1.0f / (float)S.n_lights;
    float sum = 0.f;
    RayK q = make_rayk(Ray{o, v, 0.f});
    for (int i = 0; i < S.n_lights; i++) {
        DPrim pr = load_prim(W.lights + i);
        float pdf = 0.f;
        if (COUNT) tl.c->probe_rays++;
        if (pr.type == SRT_PRIM_XZ_RECT) {  // aarect.h:45-55
            float t;
            if (rect_hit_axes<1, 0, 2>(pr.g, pr.k, q, 0.001f, kFltMax, t)) {
                float area = (pr.g.y - pr.g.x) * (pr.g.w - pr.g.z);
                float vv = q.a;  // v.squared_length()
                float distance_square = t * t * vv;
                float cosine = fabsf(v.y) * rsqrtf(vv);  // |dot(v, (0,1,0))| / |v|
                pdf = distance_square / (cosine * area);
            }
        } else if ((F & F_LIGHT_SPHERE) && pr.type == SRT_PRIM_SPHERE) {  // sphere.h:69-78
            V3 c = v3(pr.g.x, pr.g.y, pr.g.z);
            float t;
            if (sphere_hit(c, pr.g.w, q, 0.001f, kFltMax, t)) {
                // 1 - cos_theta_max cancels badly for a small or distant sphere: rounded exactly as the
                // reference rounds it (float up to the product with the double 2 * M_PI), the rest is cheap
                V3 co = c - o;
                float cos_theta_max = xsqrt(xsub(1.f, xdiv(xmul(pr.g.w, pr.g.w), dot(co, co))));
                float solid_angle = (float)(2.0 * kPi * (double)xsub(1.f, cos_theta_max));
                pdf = xdiv(1.f, solid_angle);
            }
        } else if ((F & F_LIGHT_ENV) && pr.type == SRT_PRIM_ENV_SPHERE) {
            pdf = 1.f;  // env_sphere.h:48-50
        } else if ((F & F_LIGHT_TRI) && pr.type == SRT_PRIM_TRIANGLE) {  // triangle.h:70-87
            size_t row = pr.flags_aux >> 1;
            const float* p = S.tri_pos + 9 * row;
            const float* n = S.tri_nrm + 9 * row;
            V3 p0 = v3(p), p1 = v3(p + 3), p2 = v3(p + 6);
            float t, uu, vv;
            if (tri_hit(p0, p1, p2, o, tri_dir(q.r), false, t, uu, vv)) {
                // triangle.h:72-84 operation by operation (sin from 1 - cos^2 cancels for thin triangles)
                float w = xsub(xsub(1.0f, uu), vv);
                V3 nrm = unit_vector(w * v3(n) + uu * v3(n + 3) + vv * v3(n + 6));  // rec.normal, triangle.h:179-183
                V3 v01 = p1 - p0, v02 = p2 - p0;
                float l01 = length(v01), l02 = length(v02);
                float cos102 = dot(v01 / l01, v02 / l02);
                float sin102 = xsqrt(xsub(1.f, xmul(cos102, cos102)));
                float h = xmul(l02, sin102);
                float area = (float)(0.5 * (double)l01 * (double)h);
                float distance_square = CORR ? xmul(t, t) : xmul(xmul(t, t), q.a);
                float cosine = xdiv(fabsf(dot(v, nrm)), xsqrt(q.a));
                pdf = xdiv(distance_square, xmul(cosine, area));
            }
        }
        sum = fmaf(weight, pdf, sum);
    }
    return sum;
}

// ---- the non-specular BSDF samplers: cosine_pdf, onrennayar_pdf, beckmann_pdf (pdf.h:30-156) ----
struct BsdfPdf {
    int kind;  // SRT_MAT_LAMBERTIAN / ORENNAYAR / BECKMANN
    V3 n;
    Onb uvw;
    float a, b;   // Oren-Nayar A, B or Beckmann alphax, alphay
    float stash;  // beckmann_pdf::pdf_value (pdf.h:133-135, :145-150)
};

template <uint32_t F>
SRT_DEV V3 bsdf_generate(BsdfPdf& P, V3 wo, Philox& rng) {
    if ((F & F_BECKMANN) && P.kind == SRT_MAT_BECKMANN) return beckmann_generate(P.uvw, P.a, P.b, wo, rng, P.stash);
    V3 g = random_cosine_direction(rng);  // pdf.h:47-56 / :103-112
    if (fm::dot(wo, P.n) < 0.f) g.z = -g.z;  // dot(-wo, n) > 0: flipped INTO the surface
    return onb_local_fast(P.uvw, g);
}

template <uint32_t F>
SRT_DEV float bsdf_value(const BsdfPdf& P, V3 wo, V3 wi) {
    if ((F & F_BECKMANN) && P.kind == SRT_MAT_BECKMANN) return P.stash;
    if ((F & F_ORENNAYAR) && P.kind == SRT_MAT_ORENNAYAR) return orennayar_value(P.uvw, P.a, P.b, wo, wi);
    // pdf.h:40-45: cosO = dot(unit(wo), n), cosI = dot(unit(wi), n); only signs and |cosI| matter
    float dO = fm::dot(wo, P.n);
    float dI = fm::dot(wi, P.n);
    if (dI * dO < 0.f) return fabsf(dI) * fm::rlength(wi) * fm::kInvPi;
    return 0.f;
}

// ---- materials ----------------------------------------------------------------------------------
SRT_DEV float schlick(float cosine, float ref_idx) {  // material.h:14-19
    float r0 = (1.f - ref_idx) / (1.f + ref_idx);
    r0 = r0 * r0;
    float m = 1.f - cosine;
    float m2 = m * m;
    return r0 + (1.f - r0) * (m2 * m2 * m);
}
SRT_DEV V3 reflect(V3 v, V3 n) { return fm::madd(n, -2.f * fm::dot(v, n), v); }  // material.h:34-36
SRT_DEV bool refract(V3 v, V3 n, float ni_over_nt, V3& refracted) {              // material.h:21-32
    V3 uv = fm::normalize(v);
    float dt = fm::dot(uv, n);
    float disc = 1.0f - ni_over_nt * ni_over_nt * (1.f - dt * dt);
    if (disc > 0.f) {
        refracted = fm::sub(fm::scale(fm::madd(n, -dt, uv), ni_over_nt), fm::scale(n, sqrtf(disc)));
        return true;
    }
    return false;
}

SRT_DEV V3 de_nan(V3 c) {  // Raytracing_n.cpp:47-53
    if (!(c.x == c.x)) c.x = 0.f;
    if (!(c.y == c.y)) c.y = 0.f;
    if (!(c.z == c.z)) c.z = 0.f;
    return c;
}

constexpr int kMaxRejectIters = 4096;  // the reference loop is unbounded (Raytracing_n.cpp:79-83)

// State of one path between bounces.  color()'s recursion
//     return emitted + attenuation * scattering_pdf * color(scattered) / pdf_val
// unrolls to radiance = sum over vertices of (throughput so far) * emitted.
struct Path {
    Ray r;
    V3 thr;  // product of attenuation * scattering_pdf / pdf_val so far
    V3 L;    // radiance gathered so far
    int depth;
    float emit_w;  // corrected-transport mode only: MIS weight of the emission the current ray may find
};

// One level of color() (Raytracing_n.cpp:55-106) for a path whose ray has just been traced to `h`,
// in three parts so that a scheduler can run the `while (pdf_val == 0)` rejection loop (:79-84) one
// iteration at a time across different paths (srt_wavefront.cuh); shade_bounce below chains them.
enum ShadeStatus { SHADE_ENDED = 0, SHADE_CONTINUES = 1, SHADE_SAMPLE = 2 };

// What the sampling loop of a non-specular vertex reads and writes.
struct SampleCtx {
    V3 p;            // rec.p
    V3 wo;           // direction of the incoming ray (r.direction())
    float time;      // r.time()
    V3 attenuation;  // srec.attenuation
    BsdfPdf B;       // the material's pdf object (normal, onb, parameters, Beckmann stash)
    int mat_type;
    V3 sdir;         // out: the accepted direction
    float pdf_val;   // out: its mixture density
    int iters;
};

// Part 1 (:58-74): miss / emission / depth cap / specular scatter, or the set-up of the sampling loop.
template <uint32_t F, bool COUNT>
SRT_DEV ShadeStatus shade_hit(const View& W, Path& P, const Hit& h, int max_depth, Philox& rng, SampleCtx& C,
                              Tally<COUNT> tl) {
    const DScene& S = W.S;
    if (h.prim < 0) {
        // miss: color() returns vec3(0) — kept as `thr * 0` so that an inf / NaN throughput poisons
        // the sample exactly as in the recursive form (de_nan then zeroes those channels)
        P.L = fm::madd(P.thr, 0.f, P.L);
        return SHADE_ENDED;
    }
    int mat_index;
    if ((F & F_MEDIA) && h.prim >= S.n_prims)
        mat_index = S.media[h.prim - S.n_prims].phase_material;
    else
        mat_index = load_prim(W.dprims + h.prim).material;
    if (mat_index < 0) return SHADE_ENDED;  // no material: the reference would dereference null
    DMaterial m = load_material(S.materials + mat_index);
    HitRecord rec = make_record<F>(W, P.r, h, material_needs_uv<F>(m));
    const Ray r = P.r;
    // the texture of the material at the hit, evaluated at ONE place for emission / attenuation / albedo below (three
    // inlined copies of the texture code — image lookup, checker chain, seven octaves of perlin — were 5-10 % of the
    // variants that have them); brdfmaterial and metal keep their colour in the record itself
    const V3 mcol = (m.type == SRT_MAT_BRDF || m.type == SRT_MAT_METAL || m.type == SRT_MAT_DIELECTRIC)
                        ? v3(m.c[0], m.c[1], m.c[2])
                        : material_color<F>(S, m, rec);

    if (m.type == SRT_MAT_DIFFUSE_LIGHT) {  // material.h:348-354: one-sided; scatter() returns false
        if (fm::dot(rec.normal, r.d) < 0.f) {
            V3 e = mcol;
            P.L = v3(fmaf(P.thr.x, e.x, P.L.x), fmaf(P.thr.y, e.y, P.L.y), fmaf(P.thr.z, e.z, P.L.z));
        } else {
            P.L = fm::madd(P.thr, 0.f, P.L);
        }
        return SHADE_ENDED;
    }
    if (!(P.depth < max_depth)) {  // Raytracing_n.cpp:63: returns emitted (0 for every non-light)
        P.L = fm::madd(P.thr, 0.f, P.L);
        return SHADE_ENDED;
    }
    // specular materials: return attenuation * color(specular_ray) (emitted is dropped, :66-70)
    if ((F & F_METAL) && m.type == SRT_MAT_METAL) {  // material.h:248-256
        philox_fresh(rng);
        V3 reflected = reflect(fm::normalize(r.d), rec.normal);
        V3 dir = fm::madd(random_in_unit_sphere(rng), m.p[3], reflected);
        P.thr = fm::mul(P.thr, v3(m.c[0], m.c[1], m.c[2]));
        P.r = Ray{rec.p, dir, 0.f};
        P.depth++;
        return SHADE_CONTINUES;
    }
    if ((F & F_DIELECTRIC) && m.type == SRT_MAT_DIELECTRIC) {  // material.h:285-324
        float ref_idx = m.p[0];
        V3 outward_normal;
        V3 reflected = reflect(r.d, rec.normal);  // the un-normalised direction, as in the reference
        // Under total internal reflection the reference leaves `refracted` unset and reflect_prob = 1
        // (material.h:292, :306-312; vec3() {} sets nothing, vec3.h:14), and its float rand01() CAN return 1.0f
        // (material.h:37): `1.0f < 1.0f` (:314) then takes the unset vector (:320) — undefined in the reference,
        // 2^-25 of the TIR events.  Here that case reflects: an unset
        // register made three of the 2.1 G paths of C5-like differ from run to run (profiles/r02_notes.md, capture AB).
        V3 refracted = reflected;
        float ni_over_nt, cosine, reflect_prob;
        float dn = fm::dot(r.d, rec.normal);
        float rl = fm::rlength(r.d);
        if (dn > 0.f) {
            outward_normal = fm::neg(rec.normal);
            ni_over_nt = ref_idx;
            cosine = dn * rl;  // material.h:299: NOT multiplied by ref_idx
        } else {
            outward_normal = rec.normal;
            ni_over_nt = 1.0f / ref_idx;
            cosine = -dn * rl;
        }
        if (refract(r.d, outward_normal, ni_over_nt, refracted))
            reflect_prob = schlick(cosine, ref_idx);
        else
            reflect_prob = 1.0f;
        philox_fresh(rng);
        V3 dir = (rand_f(rng) < reflect_prob) ? reflected : refracted;
        P.r = Ray{rec.p, dir, 0.f};
        P.depth++;
        return SHADE_CONTINUES;
    }
    if ((F & F_MEDIA) && m.type == SRT_MAT_ISOTROPIC) {  // material.h:362-367
        philox_fresh(rng);
        V3 dir = random_in_unit_sphere(rng);
        P.thr = fm::mul(P.thr, mcol);
        P.r = Ray{rec.p, dir, 0.f};
        P.depth++;
        return SHADE_CONTINUES;
    }
    // non-specular: lambertian, orennayar, beckmann, brdfmaterial
    BsdfPdf& B = C.B;
    B.n = rec.normal;
    B.stash = kBeckmannStashInit;
    B.a = m.p[0];
    B.b = m.p[1];
    if (m.type == SRT_MAT_BRDF) {
        // material.h:217-236: the MERL value is looked up and discarded, attenuation = albedo, and
        // srec.pdf_ptr is left unset (the reference then calls through a wild pointer).  Stand-in
        // (DESIGN.md §5): the cosine_pdf every other diffuse material gets.  The discarded lookup has
        // no side effect and is not evaluated here; its arithmetic is exported for parity through
        // srt_merl_lookup (srt_microfacet.cuh: merl_lookup).
        C.attenuation = v3(m.c[0], m.c[1], m.c[2]);
        B.kind = SRT_MAT_LAMBERTIAN;
    } else {
        C.attenuation = mcol;
        B.kind = (F & (F_ORENNAYAR | F_BECKMANN)) ? m.type : SRT_MAT_LAMBERTIAN;
    }
    B.uvw = onb_fast(rec.normal);
    C.p = rec.p;
    C.wo = r.d;
    C.time = r.time;
    C.mat_type = m.type;
    C.sdir = v3(0.f, 0.f, 0.f);
    C.pdf_val = 0.f;
    C.iters = 0;
    return SHADE_SAMPLE;
}

// Part 2: ONE iteration of the loop at Raytracing_n.cpp:79-84 (mixture_pdf(hitable_pdf, bsdf pdf),
// pdf.h:173-193), or the single draw of the light-less branch (:85-89).  Returns true when the loop
// is over (a direction with non-zero density, or the iteration bound).
template <uint32_t F, bool COUNT>
SRT_DEV bool shade_sample_iter(const View& W, SampleCtx& C, Philox& rng, Tally<COUNT> tl) {
    if (W.S.n_lights > 0) {
        if (++C.iters > kMaxRejectIters) return true;
        if (COUNT) tl.c->reject_iters++;
        philox_fresh(rng);  // one attempt = one block: it draws at most four numbers (choice, light pick, two coordinates)
        if (rand_f(rng) < 0.5f)  // pdf.h:183-189
            C.sdir = light_random<F>(W, C.p, rng);
        else
            C.sdir = bsdf_generate<F>(C.B, C.wo, rng);
        float pl = light_pdf_value<F>(W, C.p, C.sdir, tl);
        float pb = bsdf_value<F>(C.B, C.wo, C.sdir);
        C.pdf_val = 0.5f * pl + 0.5f * pb;  // pdf.h:177-182
        return C.pdf_val != 0.f;
    }
    philox_fresh(rng);
    C.sdir = bsdf_generate<F>(C.B, C.wo, rng);
    C.pdf_val = bsdf_value<F>(C.B, C.wo, C.sdir);
    return true;
}

// Part 3 (:90-99): weight the throughput and set the continuation ray.
template <uint32_t F, bool COUNT>
SRT_DEV bool shade_finish(Path& P, const SampleCtx& C, Tally<COUNT> tl) {
    float spdf;
    if ((F & F_BECKMANN) && C.mat_type == SRT_MAT_BECKMANN) {
        spdf = beckmann_scattering_pdf(C.B.n, C.B.a, C.B.b, C.wo, C.sdir);
    } else {  // material.h:100-105, :133-137, :207-213
        float c = fm::dot(C.B.n, C.sdir) * fm::rlength(C.sdir);
        if (c < 0.f) c = 0.f;
        spdf = c * fm::kInvPi;
    }
    // emitted is 0 for every scattering material (material.h:89-92)
    float wgt = spdf / C.pdf_val;
    if (COUNT && tl.c->dbg)
        printf("[dbg]   mat %d n=(%g %g %g) p=(%g %g %g) sdir=(%g %g %g) spdf %g pdf_val %g stash %g wgt %g\n", C.mat_type,
               C.B.n.x, C.B.n.y, C.B.n.z, C.p.x, C.p.y, C.p.z, C.sdir.x, C.sdir.y, C.sdir.z, spdf, C.pdf_val, C.B.stash,
               wgt);
    P.thr = fm::mul(P.thr, fm::scale(C.attenuation, wgt));
    P.r = Ray{C.p, C.sdir, C.time};
    P.depth++;
    // a zero throughput cannot contribute any more: finite * 0 == 0 in the recursive form too
    if (P.thr.x == 0.f && P.thr.y == 0.f && P.thr.z == 0.f) {
        if (COUNT) tl.c->culled_rays++;
        return false;
    }
    return true;
}

// Returns true when the path continues (P.r is the next ray), false when it has ended (P.L is final).
template <uint32_t F, bool COUNT>
SRT_DEV bool shade_bounce(const View& W, Path& P, const Hit& h, int max_depth, Philox& rng, Tally<COUNT> tl) {
    SampleCtx C;
    ShadeStatus st = shade_hit<F>(W, P, h, max_depth, rng, C, tl);
    if (st != SHADE_SAMPLE) return st == SHADE_CONTINUES;
    while (!shade_sample_iter<F>(W, C, rng, tl)) {
    }
    return shade_finish<F>(P, C, tl);
}

// ---- corrected-transport mode (F_CORRECTED, SRT_RENDER_CORRECTED) -----------------------------
// Explicitly OUTSIDE parity with the reference: what its estimator gets wrong is put right, its scene semantics
// (closest hit, one-sided emission, depth cap, specular branches, the light list as the set of sampled shapes)
// stay.  Per non-specular vertex:
//   * next-event estimation: one direction to a point of a light shape (hitable_list::random, the reference's own
//     area / cone samplers, which are correct), its mixture density p_l (hitable_list::pdf_value), and a SHADOW RAY
//     the scheduler traces (closest hit along it: the emitter's radiance if an emitter is what it reaches first);
//   * BSDF sampling about the face-forward normal: a true cosine-weighted direction (pdf.h:10-18 scales x and y by 2
//     and pdf.h:47-56 flips the lobe INTO the surface) for lambertian / Oren-Nayar / MERL, the Beckmann normal
//     distribution for beckmann;
//   * the BSDF VALUE applied: rho/pi, the Oren-Nayar term, D G / (4 cos cos) with the albedo as specular colour, and
//     the MERL table value the reference looks up and discards (material.h:231-232);
//   * balance-heuristic MIS between the two: the shadow ray carries p_l / (p_l + p_b), the continuation ray carries
//     p_b / (p_b + p_l) for the emission it may find (Path::emit_w; 1 after a camera or specular vertex).
// One Philox block per consumer, as everywhere: one for the light sample, one for the BSDF sample.
struct Nee {
    bool valid;
    V3 o;       // the vertex
    float time;
    V3 dir;     // unnormalised direction from the vertex to the sampled light point
    V3 weight;  // throughput * f * cos * w_mis / p_l: multiplies the emitted radiance the shadow ray finds
};

struct LocalFrame {
    Onb b;  // b.w = face-forward shading normal
    SRT_DEV V3 to_local(V3 d) const { return v3(fm::dot(d, b.u), fm::dot(d, b.v), fm::dot(d, b.w)); }
    SRT_DEV V3 to_world(V3 a) const { return onb_local_fast(b, a); }
};

// Beckmann normal distribution sampling (the full distribution, not the visible normals: the reference's
// visible-normal sampler leans on its Erf, which ADDS exp(-x^2), common.h:44-46, and therefore does not match any
// density).  pbrt-v3 BeckmannDistribution::Sample_wh, sampleVisibleArea = false.
SRT_DEV V3 beckmann_sample_wh(float ax, float ay, float u1, float u2) {
    float logSample = logf(fmaxf(1.f - u1, 1e-30f));
    float tan2Theta, phi;
    if (ax == ay) {
        tan2Theta = -ax * ax * logSample;
        phi = u2 * fm::kTwoPi;
    } else {
        phi = atanf(ay / ax * tanf(fm::kTwoPi * u2 + 0.5f * 3.14159265f));
        if (u2 > 0.5f) phi += 3.14159265f;
        float sp, cp;
        __sincosf(phi, &sp, &cp);
        tan2Theta = -logSample / (cp * cp / (ax * ax) + sp * sp / (ay * ay));
    }
    float cosTheta = rsqrtf(1.f + tan2Theta);
    float sinTheta = sqrtf(fmaxf(0.f, 1.f - cosTheta * cosTheta));
    float sp, cp;
    __sincosf(phi, &sp, &cp);
    return v3(sinTheta * cp, sinTheta * sp, cosTheta);
}

// f(wo, wi) * cos(theta_i) and the solid-angle density of the sampler, both in the local frame (z = face-forward
// normal; wo points AWAY from the surface).  Zero outside the upper hemisphere.
template <uint32_t F>
SRT_DEV void corrected_eval(const DScene& S, const DMaterial& m, V3 albedo, V3 wo, V3 wi, V3& fcos, float& pdf) {
    fcos = v3(0.f, 0.f, 0.f);
    pdf = 0.f;
    if (!(wi.z > 0.f) || !(wo.z > 0.f)) return;
    if ((F & F_BECKMANN) && m.type == SRT_MAT_BECKMANN) {
        const float ax = fmaxf(m.p[0], 1e-3f), ay = fmaxf(m.p[1], 1e-3f);
        V3 wh = fm::normalize(fm::add(wo, wi));
        float d = BeckmannD(wh, ax, ay);
        float g = BeckmannG(wo, wi, ax, ay);
        float k = d * g / (4.f * wo.z);  // D G / (4 cos_o cos_i) * cos_i
        fcos = fm::scale(albedo, k);
        pdf = d * fabsf(wh.z) / (4.f * fmaxf(fm::dot(wo, wh), 1e-8f));
        return;
    }
    pdf = wi.z * fm::kInvPi;
    if ((F & F_ORENNAYAR) && m.type == SRT_MAT_ORENNAYAR) {  // pdf.h:64-101's term, in this frame
        float sinI = SinTheta(wi), sinO = SinTheta(wo), maxCos = 0.f;
        if (sinI > 1e-4f && sinO > 1e-4f) maxCos = fmaxf(0.f, CosPhi(wi) * CosPhi(wo) + SinPhi(wi) * SinPhi(wo));
        float sinAlpha, tanBeta;
        if (wi.z > wo.z) {
            sinAlpha = sinO;
            tanBeta = sinI / wi.z;
        } else {
            sinAlpha = sinI;
            tanBeta = sinO / wo.z;
        }
        fcos = fm::scale(albedo, wi.z * (m.p[0] + m.p[1] * maxCos * sinAlpha * tanBeta) * fm::kInvPi);
        return;
    }
    if (m.type == SRT_MAT_BRDF) {  // the measured value, brdf.h:190-214 (fp64, as the reference evaluates it)
        double rgb[3];
        merl_lookup(S.merl[(int)m.p[3]], acos((double)fminf(wi.z, 1.f)), atan2((double)wi.y, (double)wi.x),
                    acos((double)fminf(wo.z, 1.f)), atan2((double)wo.y, (double)wo.x), rgb);
        fcos = v3(fmaxf((float)rgb[0], 0.f) * wi.z, fmaxf((float)rgb[1], 0.f) * wi.z, fmaxf((float)rgb[2], 0.f) * wi.z);
        return;
    }
    fcos = fm::scale(albedo, wi.z * fm::kInvPi);  // lambertian
}

template <uint32_t F>
SRT_DEV V3 corrected_sample(const DMaterial& m, V3 wo, Philox& rng) {
    float u1 = rand_pcg_like(rng), u2 = rand_pcg_like(rng);
    if ((F & F_BECKMANN) && m.type == SRT_MAT_BECKMANN) {
        V3 wh = beckmann_sample_wh(fmaxf(m.p[0], 1e-3f), fmaxf(m.p[1], 1e-3f), u1, u2);
        return fm::sub(fm::scale(wh, 2.f * fm::dot(wo, wh)), wo);  // reflect wo about wh
    }
    float r = sqrtf(u1), sn, cs;
    __sincosf(fm::kTwoPi * u2, &sn, &cs);
    return v3(r * cs, r * sn, sqrtf(fmaxf(0.f, 1.f - u1)));
}

// The emitted radiance a ray finds at its closest hit `h` (diffuse_light::emitted, material.h:348-354: front face
// only), zero for a miss or a non-emitter.
template <uint32_t F>
SRT_DEV V3 emitted_at(const View& W, const Ray& r, const Hit& h) {
    const DScene& S = W.S;
    if (h.prim < 0 || ((F & F_MEDIA) && h.prim >= S.n_prims)) return v3(0.f, 0.f, 0.f);
    const int mat_index = load_prim(W.dprims + h.prim).material;
    if (mat_index < 0) return v3(0.f, 0.f, 0.f);
    DMaterial m = load_material(S.materials + mat_index);
    if (m.type != SRT_MAT_DIFFUSE_LIGHT) return v3(0.f, 0.f, 0.f);
    HitRecord rec = make_record<F>(W, r, h, material_needs_uv<F>(m));
    if (!(fm::dot(rec.normal, r.d) < 0.f)) return v3(0.f, 0.f, 0.f);
    return material_color<F>(S, m, rec);
}

// One level of the corrected estimator.  P.emit_w weights the emission found at THIS hit.  Returns true when the
// path continues (P.r is the next ray); `nee` is the shadow ray of this vertex (valid or not) either way.
template <uint32_t F, bool COUNT>
SRT_DEV bool shade_bounce_corrected(const View& W, Path& P, const Hit& h, int max_depth, Philox& rng, Nee& nee, Tally<COUNT> tl) {
    const DScene& S = W.S;
    nee.valid = false;
    nee.o = nee.dir = nee.weight = v3(0.f, 0.f, 0.f);
    nee.time = P.r.time;
    if (h.prim < 0) return false;
    int mat_index;
    if ((F & F_MEDIA) && h.prim >= S.n_prims)
        mat_index = S.media[h.prim - S.n_prims].phase_material;
    else
        mat_index = load_prim(W.dprims + h.prim).material;
    if (mat_index < 0) return false;
    DMaterial m = load_material(S.materials + mat_index);
    HitRecord rec = make_record<F>(W, P.r, h, material_needs_uv<F>(m));
    const Ray r = P.r;
    if (m.type == SRT_MAT_DIFFUSE_LIGHT) {
        if (fm::dot(rec.normal, r.d) < 0.f) {
            V3 e = fm::scale(material_color<F>(S, m, rec), P.emit_w);
            P.L = v3(fmaf(P.thr.x, e.x, P.L.x), fmaf(P.thr.y, e.y, P.L.y), fmaf(P.thr.z, e.z, P.L.z));
        }
        return false;
    }
    if (!(P.depth < max_depth)) return false;
    if (((F & F_METAL) && m.type == SRT_MAT_METAL) || ((F & F_DIELECTRIC) && m.type == SRT_MAT_DIELECTRIC) ||
        ((F & F_MEDIA) && m.type == SRT_MAT_ISOTROPIC)) {
        SampleCtx C;  // the specular branches are the reference's (shade_hit): nothing to correct
        ShadeStatus st = shade_hit<F>(W, P, h, max_depth, rng, C, tl);
        P.emit_w = 1.f;
        return st == SHADE_CONTINUES;
    }
    // ---- a non-specular vertex -----------------------------------------------------------------------------
    const V3 albedo = m.type == SRT_MAT_BRDF ? v3(1.f, 1.f, 1.f) : material_color<F>(S, m, rec);
    LocalFrame fr;
    fr.b = onb_fast(fm::dot(rec.normal, r.d) < 0.f ? rec.normal : fm::neg(rec.normal));
    const V3 wo = fr.to_local(fm::neg(fm::normalize(r.d)));
    if (S.n_lights > 0) {  // next-event estimation
        philox_fresh(rng);
        V3 ldir = light_random<F, true>(W, rec.p, rng);
        float pl = light_pdf_value<F, COUNT, true>(W, rec.p, ldir, tl);
        V3 wl = fr.to_local(fm::normalize(ldir));
        V3 fcos;
        float pb;
        corrected_eval<F>(S, m, albedo, wo, wl, fcos, pb);
        if (pl > 0.f && isfinite(pl) && (fcos.x > 0.f || fcos.y > 0.f || fcos.z > 0.f)) {
            nee.valid = true;
            nee.o = rec.p;
            nee.dir = ldir;
            nee.weight = fm::mul(P.thr, fm::scale(fcos, 1.f / (pl + pb)));  // f cos / p_l * p_l / (p_l + p_b)
        }
    }
    philox_fresh(rng);
    V3 wi = corrected_sample<F>(m, wo, rng);
    V3 fcos;
    float pb;
    corrected_eval<F>(S, m, albedo, wo, wi, fcos, pb);
    if (!(pb > 0.f) || !isfinite(pb)) return false;
    P.thr = fm::mul(P.thr, fm::scale(fcos, 1.f / pb));
    if (P.thr.x == 0.f && P.thr.y == 0.f && P.thr.z == 0.f) return false;
    V3 dir = fr.to_world(wi);
    float pl2 = S.n_lights > 0 ? light_pdf_value<F, COUNT, true>(W, rec.p, dir, tl) : 0.f;
    P.emit_w = (pl2 > 0.f && isfinite(pl2)) ? pb / (pb + pl2) : 1.f;
    P.r = Ray{rec.p, dir, r.time};
    P.depth++;
    return true;
}

// camera::get_ray (camera.h:51-59).  Exact arithmetic: these rays are what srt_trace_closest is
// compared on, and the reference normalises the direction with three divisions.
SRT_DEV Ray camera_get_ray(const SrtCamera& C, float s, float t, Philox& rng) {
    V3 offset = v3(0.f, 0.f, 0.f);
    if (C.lens_radius != 0.f) {
        V3 p;
        do {  // camera.h:8-14
            p = 2.0f * v3(rand_f(rng), rand_f(rng), 0.f) - v3(1.f, 1.f, 0.f);
        } while (dot(p, p) >= 1.0f);
        V3 rd = C.lens_radius * p;
        offset = v3(C.u) * rd.x + v3(C.v) * rd.y;
    }
    float time = C.time0;
    if (C.time1 != C.time0) time = fmaf(rand_f(rng), C.time1 - C.time0, C.time0);
    V3 origin = v3(C.origin);
    V3 dir = v3(C.lower_left_corner) + s * v3(C.horizontal) + t * v3(C.vertical) - origin - offset;
    dir = unit_vector(dir);
    return Ray{origin + offset, dir, time};
}

}  // namespace srt
#endif
