// srt_intersect.cuh — closest-hit with the reference's semantics: primitive tests, list / instance
// / medium containers, wide-BVH traversal with reference-tree replay, hit-record reconstruction.
//
// Each routine cites the reference code it restates.  All arithmetic that decides ids or t uses the
// exact (non-fused) operators of srt_math.cuh; conservative pre-tests (entry bounds, wide-BVH boxes,
// the quotient estimate in rect_hit) may only REJECT what the exact arithmetic would reject too.
//
// Everything is templated on the feature mask F (srt_types.h): code for reference classes a scene
// does not instantiate is not compiled into the variant that renders it, which is what keeps the
// render kernel inside the instruction cache (profiles/r01_notes.md).
#ifndef SRT_INTERSECT_CUH
#define SRT_INTERSECT_CUH

#include "srt_math.cuh"
#include "srt_types.h"

namespace srt {

struct Counters {  // per-thread tallies of an instrumented launch (SrtCounters)
    unsigned long long rays, probe_rays, node_visits, prim_tests, reject_iters, culled_rays, paths;
    int dbg;  // counting build only: this lane traces the path selected by SRT_DEBUG_PIXEL / _SAMPLE
};

template <bool COUNT>
struct Tally {
    Counters* c;
    SRT_DEV void node(int n = 1) const {
        if (COUNT) c->node_visits += n;
    }
    SRT_DEV void prim(int n = 1) const {
        if (COUNT) c->prim_tests += n;
    }
};

// The scene as a kernel sees it: DScene plus the (possibly shared-memory) hot tables.
struct View {
    const DScene& S;
    const DEntry* entries;
    const SrtXformOp* ops;
    const DPrim* dprims;
    const DPrim* lights;
};

struct Hit {
    float t;
    int32_t prim;   // index into prims[]; n_prims + m for a scatter inside medium m; -1 = miss
    int32_t entry;  // world entry that produced it
};

struct HitRecord {  // hitable.h:17-25
    float t, u, v;
    V3 p, normal;
    int32_t material;
};

// A ray in some frame plus the divisor-only halves of the divisions its tests perform.
struct RayK {
    Ray r;
    Rcp kx, ky, kz;  // 1 / direction components (rect tests)
    Rcp ka;          // 1 / dot(d, d)            (sphere tests)
    float a;         // dot(d, d)
};
SRT_DEV RayK make_rayk(const Ray& r) {
    RayK q;
    q.r = r;
    q.kx = make_rcp(r.d.x);
    q.ky = make_rcp(r.d.y);
    q.kz = make_rcp(r.d.z);
    q.a = dot(r.d, r.d);
    q.ka = make_rcp(q.a);
    return q;
}

// ---- primitives ------------------------------------------------------------------------------

// sphere.h:36-66 (and moving_sphere.h:24-51 with center(time)): returns t only
SRT_DEV bool sphere_hit(V3 center, float radius, const RayK& q, float t_min, float t_max, float& t) {
    const Ray& r = q.r;
    V3 oc = r.o - center;
    float a = q.a;
    float b = dot(oc, r.d);
    float c = xsub(dot(oc, oc), xmul(radius, radius));
    float disc = xsub(xmul(b, b), xmul(a, c));
    if (disc > 0.f) {
        float s = xsqrt(disc);
        float temp = xdiv(xsub(-b, s), q.ka);
        if (temp < t_max && temp > t_min) {
            t = temp;
            return true;
        }
        temp = xdiv(xadd(-b, s), q.ka);
        if (temp < t_max && temp > t_min) {
            t = temp;
            return true;
        }
    }
    return false;
}

// moving_sphere.h:19-21
SRT_DEV V3 msphere_center(const float* m, float time) {
    V3 c0 = v3(m), c1 = v3(m + 3);
    float k = xdiv(xsub(time, m[6]), xsub(m[7], m[6]));
    return c0 + k * (c1 - c0);
}

// aarect.h:96-147 for one orientation: KA = the constant axis, (AA, AB) the in-plane axes.
//     t = (k - o[KA]) / d[KA];  if (t < t0 || t > t1) return false;
//     a = o[AA] + t*d[AA]; b = o[AB] + t*d[AB];  if (a < a0 || a > a1 || b < b0 || b > b1) return false;
// The quotient estimate (within 2 ulp of the rounded quotient when the divisor is in the safe
// window) rejects planes that are clearly behind the ray or beyond the closest hit so far before
// the division is completed.
template <int KA, int AA, int AB>
SRT_DEV bool rect_hit_axes(float4 g, float k, const RayK& q, float t0, float t1, float& t) {
    const Ray& r = q.r;
    const Rcp& kd = KA == 0 ? q.kx : (KA == 1 ? q.ky : q.kz);
    float num = xsub(k, KA == 0 ? r.o.x : (KA == 1 ? r.o.y : r.o.z));
    if (kd.safe) {
        float est = xdiv_estimate(num, kd);
        // |est - quotient| <= 4e-7 |quotient|; NaN compares false and falls through
        if (est < t0 - 1e-6f * fabsf(t0) || est > t1 + 1e-6f * fabsf(t1)) return false;
    }
    float tt = xdiv(num, kd);
    if (tt < t0 || tt > t1) return false;
    float oa = AA == 0 ? r.o.x : (AA == 1 ? r.o.y : r.o.z), da = AA == 0 ? r.d.x : (AA == 1 ? r.d.y : r.d.z);
    float ob = AB == 0 ? r.o.x : (AB == 1 ? r.o.y : r.o.z), db = AB == 0 ? r.d.x : (AB == 1 ? r.d.y : r.d.z);
    float a = xadd(oa, xmul(tt, da));
    float b = xadd(ob, xmul(tt, db));
    if (a < g.x || a > g.y || b < g.z || b > g.w) return false;
    t = tt;
    return true;
}
SRT_DEV bool rect_hit(int type, float4 g, float k, const RayK& q, float t0, float t1, float& t) {
    if (type == SRT_PRIM_XY_RECT) return rect_hit_axes<2, 0, 1>(g, k, q, t0, t1, t);
    if (type == SRT_PRIM_XZ_RECT) return rect_hit_axes<1, 0, 2>(g, k, q, t0, t1, t);
    return rect_hit_axes<0, 1, 2>(g, k, q, t0, t1, t);
}

// triangle.h:108-188.  Ignores (t0, t1) as the reference does; t is in units of the NORMALISED
// direction (triangle.h:128).  `(double)x < 0.0001` == `x <= 0.0001f` for float x.
// `dir` = r.direction() / r.direction().length(), computed once per ray.
SRT_DEV bool tri_hit_face(bool front, V3 p0, V3 p1, V3 p2, V3 o, V3 dir, float& t, float& u, float& v) {
    V3 e1 = front ? p1 - p0 : p0 - p1;
    V3 e2 = front ? p2 - p0 : p2 - p1;
    V3 P = cross(dir, e2);
    float det = dot(e1, P);
    V3 T;
    if (det > 0.f) {
        T = o - p0;
    } else {
        T = p0 - o;
        det = -det;
    }
    if (det <= 0.0001f) return false;
    float uu = dot(T, P);
    if (uu < 0.0f || uu > det) return false;
    V3 Q = cross(T, e1);
    float vv = dot(dir, Q);
    if (vv < 0.0f || xadd(vv, uu) > det) return false;
    float tt = dot(e2, Q);
    float inv = xdiv(1.0f, det);
    tt = xmul(tt, inv);
    uu = xmul(uu, inv);
    vv = xmul(vv, inv);
    if (tt <= 0.0001f) return false;
    t = tt;
    u = uu;
    v = vv;
    return true;
}
SRT_DEV V3 tri_dir(const Ray& r) { return r.d / length(r.d); }
SRT_DEV bool tri_hit(V3 p0, V3 p1, V3 p2, V3 o, V3 dir, bool is_medium, float& t, float& u, float& v) {
    bool h = tri_hit_face(true, p0, p1, p2, o, dir, t, u, v);
    if (!h && is_medium) h = tri_hit_face(false, p0, p1, p2, o, dir, t, u, v);  // triangle.h:108-115
    return h;
}

// env_sphere.h:27-38: always "hits"
SRT_DEV float env_sphere_t(V3 center, float radius, const Ray& r) {
    V3 oc = center - r.o;
    float proj = dot(r.d, oc);
    float distance2 = xsub(dot(oc, oc), xmul(proj, proj));
    float a = xsqrt(xsub(xmul(radius, radius), distance2));
    return xadd(a, proj);
}

// One primitive record against a ray in the primitive's own frame.
template <uint32_t F, bool COUNT>
SRT_DEV bool prim_hit(const View& W, const DPrim& pr, const RayK& q, float t_min, float t_max, bool is_medium,
                      float& t, Tally<COUNT> tl) {
    tl.prim();
    int type = pr.type;
    if (type >= SRT_PRIM_XY_RECT && type <= SRT_PRIM_YZ_RECT) return rect_hit(type, pr.g, pr.k, q, t_min, t_max, t);
    if (type == SRT_PRIM_SPHERE) return sphere_hit(v3(pr.g.x, pr.g.y, pr.g.z), pr.g.w, q, t_min, t_max, t);
    if ((F & F_TRI) && type == SRT_PRIM_TRIANGLE) {
        const float* p = W.S.tri_pos + 9 * (size_t)(pr.flags_aux >> 1);
        float u, v;
        return tri_hit(v3(p), v3(p + 3), v3(p + 6), q.r.o, tri_dir(q.r), is_medium, t, u, v);
    }
    if ((F & F_MSPHERE) && type == SRT_PRIM_MOVING_SPHERE) {
        const float* m = W.S.msph + 9 * (size_t)(pr.flags_aux >> 1);
        return sphere_hit(msphere_center(m, q.r.time), m[8], q, t_min, t_max, t);
    }
    if ((F & F_ENV) && type == SRT_PRIM_ENV_SPHERE) {
        t = env_sphere_t(v3(pr.g.x, pr.g.y, pr.g.z), pr.g.w, q.r);
        return true;
    }
    return false;
}

SRT_DEV DPrim load_prim(const DPrim* p) {
    const float4* q = reinterpret_cast<const float4*>(p);
    float4 a = q[0], b = q[1];
    DPrim r;
    r.g = a;
    r.k = b.x;
    r.type = __float_as_int(b.y);
    r.material = __float_as_int(b.z);
    r.flags_aux = __float_as_uint(b.w);
    return r;
}

// hitable_list::hit (hitable_list.h:21-33) over prims [first, first + count): in order, t_max
// narrowed to the last accepted t (the double closest_so_far holds a float value, so the narrowing
// is exact).
template <uint32_t F, bool COUNT>
SRT_DEV bool group_hit(const View& W, int first, int count, const RayK& q, float t_min, float t_max, bool is_medium,
                       float& t, int32_t& prim, Tally<COUNT> tl) {
    bool any = false;
    float closest = t_max;
    for (int i = 0; i < count; i++) {
        DPrim pr = load_prim(W.dprims + first + i);
        float tt;
        if (prim_hit<F>(W, pr, q, t_min, closest, is_medium, tt, tl)) {
            any = true;
            closest = tt;
            prim = first + i;
        }
    }
    t = closest;
    return any;
}

// ---- the reference's aabb::hit (aabb.h:33-49), bit for bit -------------------------------------
struct RayInv {
    float ix, iy, iz;
};
SRT_DEV bool ref_aabb_hit(float4 lo, float4 hi, const Ray& r, RayInv inv, float tmin, float tmax) {
    {
        float t0 = xmul(xsub(lo.x, r.o.x), inv.ix), t1 = xmul(xsub(hi.x, r.o.x), inv.ix);
        if (inv.ix < 0.0f) {
            float s = t0;
            t0 = t1;
            t1 = s;
        }
        tmin = t0 > tmin ? t0 : tmin;
        tmax = t1 < tmax ? t1 : tmax;
        if (tmax <= tmin) return false;
    }
    {
        float t0 = xmul(xsub(lo.y, r.o.y), inv.iy), t1 = xmul(xsub(hi.y, r.o.y), inv.iy);
        if (inv.iy < 0.0f) {
            float s = t0;
            t0 = t1;
            t1 = s;
        }
        tmin = t0 > tmin ? t0 : tmin;
        tmax = t1 < tmax ? t1 : tmax;
        if (tmax <= tmin) return false;
    }
    {
        float t0 = xmul(xsub(lo.z, r.o.z), inv.iz), t1 = xmul(xsub(hi.z, r.o.z), inv.iz);
        if (inv.iz < 0.0f) {
            float s = t0;
            t0 = t1;
            t1 = s;
        }
        tmin = t0 > tmin ? t0 : tmin;
        tmax = t1 < tmax ? t1 : tmax;
        if (tmax <= tmin) return false;
    }
    return true;
}

// Is the leaf of in-order rank `rank` reachable in the reference tree?  Every bvh_node on the way
// down tests its own box with the (t_min, t_max) the root was entered with (bvh.h:64-68).
template <bool COUNT>
SRT_DEV bool ref_reachable(const RefNode* root, int rank, const Ray& r, RayInv inv, float t_min, float t_max,
                           Tally<COUNT> tl) {
    int idx = 0;
    for (;;) {
        const float4* q = reinterpret_cast<const float4*>(root + idx);
        float4 a = __ldg(q), b = __ldg(q + 1);
        tl.node();
        if (!ref_aabb_hit(a, b, r, inv, t_min, t_max)) return false;
        int left = __float_as_int(a.w), right = __float_as_int(b.w);
        if (left < 0) return true;
        int mid = __ldg(&root[idx].mid);
        idx = rank < mid ? left : right;
    }
}

// ---- wide-BVH traversal -------------------------------------------------------------------------
// Result of a reference `bvh_node::hit(r, t_min, t_max)` (bvh.h:64-93): over all REACHABLE leaves
// that hit, the smallest t; on equal t the leaf later in the in-order sequence (the right child
// wins every `left_rec.t < right_rec.t` tie).  Leaves are tested with the un-narrowed entry
// (t_min, t_max), as in the reference; OUR tree only orders the search and prunes conservatively.
//
// Reachability (the un-padded reference ancestor boxes) fails only for rays grazing a box face, so
// it is checked for the WINNER only: a pass finds the best candidate strictly after the exclusion
// bound (ex_t, ex_rank) in the order "smaller t first, then higher rank"; if that candidate turns
// out unreachable it becomes the new bound and the search repeats.  One pass in all but ~1e-6 of rays.
// The stack holds the postponed children of the current root-to-node path: at most 3 per level of OUR tree.
// A stack that fills up completely is a COUNTED condition (DScene::diag -> SrtCounters::stack_overflows, asserted 0 by the tests
// on every configuration including the 10 M-triangle one): the postponed subtree would be lost, so the
// caller must be able to see it.  srt_bvh_build.cpp reports the tree's depth (SrtSceneInfo::bvh_max_depth);
// near-first descent with cull-on-pop keeps the occupied depth far below 3 x depth in practice.
#ifndef SRT_BVH_STACK
#define SRT_BVH_STACK 48  // tools/build_alt.sh tiny -DSRT_BVH_STACK=4 builds the library the overflow counter is checked with
#endif
constexpr int kStackSize = SRT_BVH_STACK;

// Wide-node step helpers.  Children that pass the slab test are ordered near-to-far with a 5-comparator
// network on 32-bit keys held in registers: key = bits of max(lower bound, 0) with the two low
// mantissa bits replaced by the child slot (non-negative floats order like unsigned integers), all
// ones for a child that is not visited.  No dynamically indexed local arrays, no data-dependent loop.
constexpr uint32_t kKeyNone = 0xffffffffu;
constexpr int32_t kRefEmpty = (int32_t)0x80000000;  // WideNode4::ref of an unused child slot
constexpr int32_t kRefDone = (int32_t)0x80000001;   // traversal cursor: nothing left
SRT_DEV void key_sort2(uint32_t& a, uint32_t& b) {
    uint32_t lo = min(a, b), hi = max(a, b);
    a = lo;
    b = hi;
}
SRT_DEV int32_t ref_select(int4 r, uint32_t key) {
    uint32_t i = key & 3u;
    int32_t lo = (i & 1u) ? r.y : r.x, hi = (i & 1u) ? r.w : r.z;
    return (i & 2u) ? hi : lo;
}

template <uint32_t F, bool COUNT>
SRT_DEV bool bvh_hit(const View& W, const DBvh& B, const RayK& q, float t_min, float t_max, bool is_medium,
                     float& t_out, int32_t& prim_out, Tally<COUNT> tl) {
    const DScene& S = W.S;
    const Ray& r = q.r;
    // pruning scale: triangle t is a distance, box slabs are parametric
    float len = xsqrt(q.a);
    const int kind = (F & F_BVH_GENERIC) || ((F & F_BVH_SPHERE) && (F & F_BVH_TRI))
                         ? B.kind
                         : ((F & F_BVH_TRI) ? BVH_KIND_TRIANGLE : BVH_KIND_SPHERE);
    float tscale = kind == BVH_KIND_SPHERE ? 1.0f : (kind == BVH_KIND_TRIANGLE ? len : fminf(1.0f, len));
    if (!(tscale == tscale)) return false;  // NaN direction: nothing sensible to report
    const WideNode4* nodes = S.wide_nodes + B.node_base;
    const int32_t* ranks = S.leaf_rank + B.leaf_base;
    V3 tdir = v3(0.f, 0.f, 0.f);
    if ((F & (F_BVH_TRI | F_BVH_GENERIC)) && kind != BVH_KIND_SPHERE) tdir = tri_dir(r);

    // Slab planes of OUR tree in the form fma(plane, 1/d, -o/d): it only orders and prunes, so it needs
    // to be conservative, not exact.  1/d is clamped to +-1e30 (axis-parallel and denormal directions:
    // inside the slab the interval is huge, outside it is empty, on the face the margin decides for
    // "hit"); `em` covers the rounding of o/d, which the subtract-first form does not have.
    const float cix = fminf(fmaxf(q.kx.r, -1e30f), 1e30f), ciy = fminf(fmaxf(q.ky.r, -1e30f), 1e30f),
                ciz = fminf(fmaxf(q.kz.r, -1e30f), 1e30f);
    const float nox = -r.o.x * cix, noy = -r.o.y * ciy, noz = -r.o.z * ciz;
    const float em = 1e-6f * fmaxf(fmaxf(fabsf(nox), fabsf(noy)), fabsf(noz));
    const float f_floor = fminf(t_min, 0.0f);

    float ex_t = -CUDART_INF_F;
    int ex_rank = 0x7fffffff;
    for (;;) {  // passes (see above)
        float best_t = CUDART_INF_F;
        int best_rank = -1, best_slot = -1;
        int32_t best_prim = -1;
        uint2 stack[kStackSize];  // (key, ref)
        int sp = 0;
#ifndef SRT_NO_OVERFLOW_CHECK
        // canary in the last entry: a push only ever writes it when the stack becomes completely full, which
        // is the state in which the next postponed subtree would be dropped.  One local store here and one
        // load after the traversal; the node loop itself carries no check (an atomic inside it cost 45 % on
        // the C5 scene: profiles/r02_notes.md)
        stack[kStackSize - 1] = make_uint2(kKeyNone, 0u);
#endif
        int32_t cur = 0;  // >= 0 inner node; < 0 leaf reference ~(first << 3 | count); kRefDone
        for (;;) {
            // ---- inner nodes: descend until this lane holds a leaf (or has nothing left) -------------
            while (cur >= 0) {
                const float4* nq = reinterpret_cast<const float4*>(nodes + cur);
                float4 lox = __ldg(nq), loy = __ldg(nq + 1), loz = __ldg(nq + 2);
                float4 hix = __ldg(nq + 3), hiy = __ldg(nq + 4), hiz = __ldg(nq + 5);
                int4 refs = __ldg(reinterpret_cast<const int4*>(nq + 7));
                tl.node();
                const float LX[4] = {lox.x, lox.y, lox.z, lox.w}, LY[4] = {loy.x, loy.y, loy.z, loy.w};
                const float LZ[4] = {loz.x, loz.y, loz.z, loz.w}, HX[4] = {hix.x, hix.y, hix.z, hix.w};
                const float HY[4] = {hiy.x, hiy.y, hiy.z, hiy.w}, HZ[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
                const int RF[4] = {refs.x, refs.y, refs.z, refs.w};
                uint32_t key[4];
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    float ax = fmaf(LX[i], cix, nox), bx = fmaf(HX[i], cix, nox);
                    float ay = fmaf(LY[i], ciy, noy), by = fmaf(HY[i], ciy, noy);
                    float az = fmaf(LZ[i], ciz, noz), bz = fmaf(HZ[i], ciz, noz);
                    float n = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
                    float f = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
                    // conservative: widen the interval before deciding
                    n = fmaf(fabsf(n), -4e-6f, n) - em;
                    f = fmaf(fabsf(f), 4e-6f, f) + em;
                    // lower bound of any leaf t inside this box, in the leaves' own units
                    float lb = n > 0.0f ? n * tscale : (kind == BVH_KIND_TRIANGLE ? 0.0f : n);
                    bool h = RF[i] != kRefEmpty && n <= f && f >= f_floor && lb <= best_t;
                    key[i] = h ? ((__float_as_uint(fmaxf(lb, 0.0f)) & ~3u) | (uint32_t)i) : kKeyNone;
                }
                key_sort2(key[0], key[1]);
                key_sort2(key[2], key[3]);
                key_sort2(key[0], key[2]);
                key_sort2(key[1], key[3]);
                key_sort2(key[1], key[2]);
                // push far-to-near, continue with the nearest
                if (key[3] != kKeyNone && sp < kStackSize) stack[sp++] = make_uint2(key[3], (uint32_t)ref_select(refs, key[3]));
                if (key[2] != kKeyNone && sp < kStackSize) stack[sp++] = make_uint2(key[2], (uint32_t)ref_select(refs, key[2]));
                if (key[1] != kKeyNone && sp < kStackSize) stack[sp++] = make_uint2(key[1], (uint32_t)ref_select(refs, key[1]));
                if (key[0] != kKeyNone) {
                    cur = ref_select(refs, key[0]);
                } else {
                    cur = kRefDone;
                    while (sp > 0) {  // pop, skipping entries the current best has overtaken
                        uint2 e = stack[--sp];
                        float lb = __uint_as_float(e.x & ~3u);
                        if (lb == 0.0f || lb <= best_t) {
                            cur = (int32_t)e.y;
                            break;
                        }
                    }
                }
            }
            if (cur == kRefDone) break;
            // ---- leaf ---------------------------------------------------------------------------------
            {
                int packed = ~cur;
                int first = packed >> 3, cnt = packed & 7;
                for (int s = first; s < first + cnt; s++) {
                    int rank = 0;
                    if (kind != BVH_KIND_TRIANGLE) rank = __ldg(&ranks[s]);
                    float tt;
                    int32_t pp;
                    bool h;
                    if (kind == BVH_KIND_SPHERE) {
                        float4 sp4 = __ldg(&S.leaf_geom[B.geom_base + s]);
                        tl.prim();
                        h = sphere_hit(v3(sp4.x, sp4.y, sp4.z), sp4.w, q, t_min, t_max, tt);
                        pp = -2;  // resolved from the rank at the end
                    } else if (kind == BVH_KIND_TRIANGLE) {
                        const float4* g = S.leaf_geom + B.geom_base + 3 * (size_t)s;
                        float4 a = __ldg(g), b = __ldg(g + 1), c = __ldg(g + 2);
                        float u, v;
                        tl.prim();
                        h = tri_hit(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), r.o, tdir, is_medium, tt, u,
                                    v);
                        rank = __float_as_int(a.w);  // rank and primitive id ride in the spare words of the record
                        pp = __float_as_int(b.w);
                    } else {
                        SrtGroup g = S.groups[__ldg(&S.bvh_items[B.first_item + rank])];
                        h = group_hit<F>(W, g.first_prim, g.prim_count, q, t_min, t_max, is_medium, tt, pp, tl);
                    }
                    if (h && (tt < best_t || (tt == best_t && rank > best_rank)) &&
                        (tt > ex_t || (tt == ex_t && rank < ex_rank))) {
                        best_t = tt;
                        best_rank = rank;
                        best_prim = pp;
                        best_slot = s;
                    }
                }
            }
            cur = kRefDone;
            while (sp > 0) {
                uint2 e = stack[--sp];
                float lb = __uint_as_float(e.x & ~3u);
                if (lb == 0.0f || lb <= best_t) {
                    cur = (int32_t)e.y;
                    break;
                }
            }
            if (cur == kRefDone) break;
        }
#ifndef SRT_NO_OVERFLOW_CHECK
        if (stack[kStackSize - 1].x != kKeyNone) atomicAdd(S.diag, 1ull);  // counted, never silent
#endif
        if (best_rank < 0) return false;
        RayInv inv{xdiv(1.0f, q.kx), xdiv(1.0f, q.ky), xdiv(1.0f, q.kz)};
        // Reachability shortcut.  Every ancestor box of the reference tree is a union that contains the
        // winner's own bounding_box() (sphere.h:31-34, triangle.h:53-68; surrounding_box is exact min / max),
        // and aabb::hit is monotone in the box: float subtraction and the multiplication by the same
        // 1/d keep the order of the slab planes (the swap for 1/d < 0 keeps it too, a NaN product leaves
        // tmin / tmax untouched, i.e. only loosens the bigger box), so per axis tmin' <= tmin and
        // tmax' >= tmax.  A ray that passes aabb::hit on the winner's own box passes it on every
        // ancestor: no walk down the reference tree (a chain of dependent loads, ~log2 n deep).
        bool reachable = false;
        if (kind == BVH_KIND_SPHERE) {
            float4 g = __ldg(&S.leaf_geom[B.geom_base + best_slot]);
            reachable = ref_aabb_hit(make_float4(xsub(g.x, g.w), xsub(g.y, g.w), xsub(g.z, g.w), 0.f),
                                     make_float4(xadd(g.x, g.w), xadd(g.y, g.w), xadd(g.z, g.w), 0.f), r, inv, t_min, t_max);
        } else if (kind == BVH_KIND_TRIANGLE) {
            const float4* g = S.leaf_geom + B.geom_base + 3 * (size_t)best_slot;
            float4 a = __ldg(g), b = __ldg(g + 1), c = __ldg(g + 2);
            reachable = ref_aabb_hit(
                make_float4(fminf(fminf(a.x, b.x), c.x), fminf(fminf(a.y, b.y), c.y), fminf(fminf(a.z, b.z), c.z), 0.f),
                make_float4(fmaxf(fmaxf(a.x, b.x), c.x), fmaxf(fmaxf(a.y, b.y), c.y), fmaxf(fmaxf(a.z, b.z), c.z), 0.f), r,
                inv, t_min, t_max);
        }
        if (reachable || ref_reachable(S.ref_nodes + B.ref_base, best_rank, r, inv, t_min, t_max, tl)) {
            if (best_prim == -2) best_prim = S.groups[S.bvh_items[B.first_item + best_rank]].first_prim;
            t_out = best_t;
            prim_out = best_prim;
            return true;
        }
        ex_t = best_t;
        ex_rank = best_rank;
    }
}

// ---- instance transforms (hitable.h:44-52, :109-132, :180-203) -----------------------------------
SRT_DEV Ray xform_ray_in(const SrtXformOp& op, Ray r) {
    if (op.kind == SRT_OP_TRANSLATE) {
        r.o = r.o - v3(op.a, op.b, op.c);
    } else if (op.kind == SRT_OP_ROTATE_Y) {
        float s = op.a, c = op.b;
        V3 o = r.o, d = r.d;
        o.x = xsub(xmul(c, r.o.x), xmul(s, r.o.z));
        o.z = xadd(xmul(s, r.o.x), xmul(c, r.o.z));
        d.x = xsub(xmul(c, r.d.x), xmul(s, r.d.z));
        d.z = xadd(xmul(s, r.d.x), xmul(c, r.d.z));
        r.o = o;
        r.d = d;
    } else if (op.kind == SRT_OP_ROTATE_X) {
        float s = op.a, c = op.b;
        V3 o = r.o, d = r.d;
        o.y = xsub(xmul(c, r.o.y), xmul(s, r.o.z));
        o.z = xadd(xmul(s, r.o.y), xmul(c, r.o.z));
        d.y = xsub(xmul(c, r.d.y), xmul(s, r.d.z));
        d.z = xadd(xmul(s, r.d.y), xmul(c, r.d.z));
        r.o = o;
        r.d = d;
    }
    return r;
}

SRT_DEV void xform_rec_out(const SrtXformOp& op, V3& p, V3& n) {
    if (op.kind == SRT_OP_TRANSLATE) {
        p = p + v3(op.a, op.b, op.c);
    } else if (op.kind == SRT_OP_ROTATE_Y) {
        float s = op.a, c = op.b;
        V3 pp = p, nn = n;
        pp.x = xadd(xmul(c, p.x), xmul(s, p.z));
        pp.z = xadd(xmul(-s, p.x), xmul(c, p.z));
        nn.x = xadd(xmul(c, n.x), xmul(s, n.z));
        nn.z = xadd(xmul(-s, n.x), xmul(c, n.z));
        p = pp;
        n = nn;
    } else if (op.kind == SRT_OP_ROTATE_X) {
        float s = op.a, c = op.b;
        V3 pp = p, nn = n;
        pp.y = xadd(xmul(c, p.y), xmul(s, p.z));
        pp.z = xadd(xmul(-s, p.y), xmul(c, p.z));
        nn.y = xadd(xmul(c, n.y), xmul(s, n.z));
        nn.z = xadd(xmul(-s, n.y), xmul(c, n.z));
        p = pp;
        n = nn;
    } else if (op.kind == SRT_OP_FLIP) {
        n = -n;
    }
}

SRT_DEV SrtXformOp load_op(const SrtXformOp* p) {
    float4 v = *reinterpret_cast<const float4*>(p);
    SrtXformOp o;
    o.kind = __float_as_int(v.x);
    o.a = v.y;
    o.b = v.z;
    o.c = v.w;
    return o;
}

SRT_DEV DEntry load_entry_head(const DEntry* p) {  // the first 32 bytes: ranges, kind, medium, flags
    const int4* q = reinterpret_cast<const int4*>(p);
    int4 a = q[0], b = q[1];
    DEntry e;
    e.first_prim = a.x;
    e.prim_count = a.y;
    e.first_op = a.z;
    e.op_count = a.w;
    e.geom_kind = b.x;
    e.geom_index = b.y;
    e.medium = b.z;
    e.flags = (uint32_t)b.w;
    return e;
}

template <uint32_t F>
SRT_DEV Ray entry_local_ray(const View& W, int first_op, int op_count, Ray r) {
    if (F & F_XFORM)
        for (int i = 0; i < op_count; i++) r = xform_ray_in(load_op(W.ops + first_op + i), r);
    return r;
}

// Conservative slab test of a ray against an entry's bounds (local frame).  Returns false only
// when no primitive below can be hit: the bounds are padded (srt_api.cu: entry_bounds) and the
// interval is widened by a relative and an origin-scaled margin before the decision.
SRT_DEV bool entry_bounds_hit(const DEntry* ep, const RayK& q, float t_min, float t_max, bool tcull) {
    const float4* p = reinterpret_cast<const float4*>(ep);
    float4 lo = p[2], hi = p[3];
    const Ray& r = q.r;
    if (!(q.kx.safe && q.ky.safe && q.kz.safe)) return true;  // axis-parallel or denormal direction
    float ax = (lo.x - r.o.x) * q.kx.r, bx = (hi.x - r.o.x) * q.kx.r;
    float ay = (lo.y - r.o.y) * q.ky.r, by = (hi.y - r.o.y) * q.ky.r;
    float az = (lo.z - r.o.z) * q.kz.r, bz = (hi.z - r.o.z) * q.kz.r;
    float n = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    float f = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    float em = 2e-6f * fmaxf(fmaxf(fabsf(r.o.x * q.kx.r), fabsf(r.o.y * q.ky.r)), fabsf(r.o.z * q.kz.r));
    n = n - fabsf(n) * 4e-6f - em;
    f = f + fabsf(f) * 4e-6f + em;
    if (!(n <= f)) return false;
    if (f < fminf(t_min, 0.0f)) return false;
    if (tcull && n > t_max) return false;
    return true;
}

// The geometry below an entry's transform chain (group or bvh), without the medium wrapper.
// `qw` = make_rayk(world_ray): an entry without a transform chain reuses it (the second set of reciprocals was
// ~50 instructions per ray and entry)
template <uint32_t F, bool COUNT>
SRT_DEV bool entry_geom_hit(const View& W, const DEntry* ep, const DEntry& e, const Ray& world_ray, const RayK& qw,
                            float t_min, float t_max, bool is_medium, float& t, int32_t& prim, Tally<COUNT> tl) {
    RayK q = qw;
    if ((F & F_XFORM) && e.op_count > 0) q = make_rayk(entry_local_ray<F>(W, e.first_op, e.op_count, world_ray));
    if ((e.flags & DENTRY_CULL) && !is_medium && !entry_bounds_hit(ep, q, t_min, t_max, (e.flags & DENTRY_TCULL) != 0))
        return false;
    if (!(F & (F_BVH_SPHERE | F_BVH_TRI | F_BVH_GENERIC)) || e.geom_kind == SRT_GEOM_GROUP)
        return group_hit<F>(W, e.first_prim, e.prim_count, q, t_min, t_max, is_medium, t, prim, tl);
    DBvh B = W.S.bvhs[e.geom_index];
    return bvh_hit<F>(W, B, q, t_min, t_max, is_medium, t, prim, tl);
}

// constant_medium::hit (constant_medium.h:19-50): two boundary hits, then an exponential
// free-flight distance.  The dead debug draw at :20 is not reproduced (counter RNG).
template <uint32_t F, bool COUNT>
SRT_DEV bool medium_hit(const View& W, const DEntry* ep, const DEntry& e, const Ray& r, const RayK& qw, float t_min,
                        float t_max, Philox& rng, float& t, Tally<COUNT> tl) {
    // the two boundary hits (constant_medium.h:22-25) through ONE call site: two inlined copies of the list walk and
    // of the traversal were a fifth of the code of the variants with media
    float t1 = 0.f, t2 = 0.f;
    float lo = -kFltMax;
#pragma unroll 1
    for (int k = 0; k < 2; k++) {
        float tk;
        int32_t pk;
        if (!entry_geom_hit<F>(W, ep, e, r, qw, lo, kFltMax, true, tk, pk, tl)) return false;
        if (k == 0) {
            t1 = tk;
            lo = (float)((double)t1 + 0.0001);
        } else {
            t2 = tk;
        }
    }
    if (t1 < t_min) t1 = t_min;
    if (t2 > t_max) t2 = t_max;
    if (t1 >= t2) return false;
    if (t1 < 0.f) t1 = 0.f;
    float len = length(r.d);
    float distance_inside = xmul(xsub(t2, t1), len);
    float density = W.S.media[e.medium].density;
    philox_fresh(rng);  // one free-flight draw = one block (srt_math.cuh)
    float hit_distance = (float)(-(double)xdiv(1.0f, density) * log(rand_d(rng)));
    if (hit_distance < distance_inside) {
        t = xadd(t1, xdiv(hit_distance, len));
        return true;
    }
    return false;
}

// world->hit(r, t_min, t_max, rec) (Raytracing_n.cpp:58): the top-level hitable_list.
template <uint32_t F, bool COUNT>
SRT_DEV Hit world_hit(const View& W, const Ray& r, float t_min, float t_max, Philox& rng, Tally<COUNT> tl) {
    Hit h;
    h.prim = -1;
    h.entry = -1;
    h.t = t_max;
    float closest = t_max;
    RayK qw = make_rayk(r);  // for entries without a transform chain
    const int n = W.S.n_entries;
    for (int ei = 0; ei < n; ei++) {
        const DEntry* ep = W.entries + ei;
        DEntry e = load_entry_head(ep);
        float t;
        int32_t prim = -1;
        bool got;
        if ((F & F_MEDIA) && e.medium >= 0) {
            got = medium_hit<F>(W, ep, e, r, qw, t_min, closest, rng, t, tl);
            prim = W.S.n_prims + e.medium;
        } else {
            // one call site for every kind of entry (an untransformed one reuses qw inside): a second inlined copy
            // of the list walk with all its primitive tests was ~3 % of every variant's code
            got = entry_geom_hit<F>(W, ep, e, r, qw, t_min, closest, false, t, prim, tl);
        }
        if (got) {
            closest = t;
            h.t = t;
            h.prim = prim;
            h.entry = ei;
        }
    }
    return h;
}

// ---- hit-record reconstruction ----------------------------------------------------------------
// hitable.h:10-15
SRT_DEV void get_sphere_uv(V3 p, float& u, float& v) {
    float phi = atan2f(p.z, p.x);
    float theta = asinf(p.y);
    u = (float)(1.0 - ((double)phi + kPi) / (2.0 * kPi));
    v = (float)(((double)theta + kPi / 2.0) / kPi);
}

// Re-runs the winning primitive's arithmetic to fill u, v, p, normal exactly as its hit() does,
// then unwinds flip_normals and the instance chain (aarect.h:152-159, hitable.h:44-52, :109-132).
// `want_uv` is a run-time flag (one copy of this routine per kernel, not one per value: srt_shade.cuh asked for both).
template <uint32_t F>
SRT_DEV HitRecord make_record(const View& W, const Ray& world_ray, const Hit& h, bool want_uv) {
    const DScene& S = W.S;
    HitRecord rec;
    rec.t = h.t;
    rec.u = rec.v = 0.f;
    rec.material = -1;
    if ((F & F_MEDIA) && h.prim >= S.n_prims) {  // constant_medium.h:39-44
        rec.p = point_at(world_ray, h.t);
        rec.normal = v3(1.f, 0.f, 0.f);
        rec.material = S.media[h.prim - S.n_prims].phase_material;
        return rec;
    }
    DEntry e = load_entry_head(W.entries + h.entry);
    Ray r = entry_local_ray<F>(W, e.first_op, e.op_count, world_ray);
    DPrim pr = load_prim(W.dprims + h.prim);
    rec.material = pr.material;
    int type = pr.type;
    if (type >= SRT_PRIM_XY_RECT && type <= SRT_PRIM_YZ_RECT) {
        int aa = type == SRT_PRIM_YZ_RECT ? 1 : 0;
        int ab = type == SRT_PRIM_XY_RECT ? 1 : 2;
        if (want_uv) {
            float a = xadd(comp(r.o, aa), xmul(h.t, comp(r.d, aa)));
            float b = xadd(comp(r.o, ab), xmul(h.t, comp(r.d, ab)));
            rec.u = xdiv(xsub(a, pr.g.x), xsub(pr.g.y, pr.g.x));
            rec.v = xdiv(xsub(b, pr.g.z), xsub(pr.g.w, pr.g.z));
        }
        rec.p = point_at(r, h.t);
        rec.normal = type == SRT_PRIM_XY_RECT ? v3(0.f, 0.f, 1.f)
                                              : (type == SRT_PRIM_XZ_RECT ? v3(0.f, 1.f, 0.f) : v3(1.f, 0.f, 0.f));
    } else if (type == SRT_PRIM_SPHERE || ((F & F_ENV) && type == SRT_PRIM_ENV_SPHERE)) {
        V3 c = v3(pr.g.x, pr.g.y, pr.g.z);
        rec.p = point_at(r, h.t);
        V3 qn = (rec.p - c) / pr.g.w;
        if (want_uv) get_sphere_uv(qn, rec.u, rec.v);
        rec.normal = type == SRT_PRIM_SPHERE ? qn : (c - rec.p) / pr.g.w;
    } else if ((F & (F_TRI | F_BVH_TRI)) && type == SRT_PRIM_TRIANGLE) {
        size_t row = pr.flags_aux >> 1;
        const float* p = S.tri_pos + 9 * row;
        const float* n = S.tri_nrm + 9 * row;
        V3 p0 = v3(p), p1 = v3(p + 3), p2 = v3(p + 6);
        float t, u = 0.f, v = 0.f;
        // a hit found through a medium boundary may be the back-face variant
        tri_hit(p0, p1, p2, r.o, tri_dir(r), true, t, u, v);
        float w = xsub(xsub(1.0f, u), v);
        if (want_uv) {
            const float* uv = S.tri_uv + 6 * row;
            rec.u = xadd(xadd(xmul(w, uv[0]), xmul(u, uv[2])), xmul(v, uv[4]));
            rec.v = xadd(xadd(xmul(w, uv[1]), xmul(u, uv[3])), xmul(v, uv[5]));
        }
        rec.normal = unit_vector(w * v3(n) + u * v3(n + 3) + v * v3(n + 6));  // FLAT_NORMAL 1, triangle.h:179-183
        rec.p = w * p0 + u * p1 + v * p2;
    } else if ((F & F_MSPHERE) && type == SRT_PRIM_MOVING_SPHERE) {
        const float* m = S.msph + 9 * (size_t)(pr.flags_aux >> 1);
        rec.p = point_at(r, h.t);
        rec.normal = (rec.p - msphere_center(m, r.time)) / m[8];
    } else {
        rec.p = point_at(r, h.t);
        rec.normal = v3(0.f, 0.f, 0.f);
    }
    if (pr.flags_aux & SRT_PRIM_FLIP) rec.normal = -rec.normal;
    if (F & F_XFORM)
        for (int i = e.op_count - 1; i >= 0; i--) xform_rec_out(load_op(W.ops + e.first_op + i), rec.p, rec.normal);
    return rec;
}

}  // namespace srt
#endif
