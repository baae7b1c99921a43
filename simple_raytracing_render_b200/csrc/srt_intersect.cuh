// srt_intersect.cuh — closest-hit with the reference's semantics: primitive tests, list / instance
// / medium containers, wide-BVH traversal with reference-tree replay, hit-record reconstruction.
//
// Each routine cites the reference code it restates.  All arithmetic that decides ids or t uses the
// exact (non-fused) operators of srt_math.cuh.
#ifndef SRT_INTERSECT_CUH
#define SRT_INTERSECT_CUH

#include "srt_math.cuh"
#include "srt_types.h"

namespace srt {

struct Counters {  // per-thread tallies of an instrumented launch (SrtCounters)
    unsigned long long rays, probe_rays, node_visits, prim_tests, reject_iters, culled_rays;
};

template <bool COUNT>
struct Tally {
    Counters* c;
    SRT_DEV void node(int n = 1) {
        if (COUNT) c->node_visits += n;
    }
    SRT_DEV void prim(int n = 1) {
        if (COUNT) c->prim_tests += n;
    }
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

// ---- primitives ------------------------------------------------------------------------------

// sphere.h:36-66 (and moving_sphere.h:24-51 with center(time)): returns t only
SRT_DEV bool sphere_hit(V3 center, float radius, const Ray& r, float t_min, float t_max, float& t) {
    V3 oc = r.o - center;
    float a = dot(r.d, r.d);
    float b = dot(oc, r.d);
    float c = xsub(dot(oc, oc), xmul(radius, radius));
    float disc = xsub(xmul(b, b), xmul(a, c));
    if (disc > 0.f) {
        float s = xsqrt(disc);
        float temp = xdiv(xsub(-b, s), a);
        if (temp < t_max && temp > t_min) {
            t = temp;
            return true;
        }
        temp = xdiv(xadd(-b, s), a);
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

// aarect.h:96-147.  axis triple (ka, aa, ab): the constant axis and the two in-plane axes.
SRT_DEV bool rect_hit(int type, float4 q0, float k, const Ray& r, float t0, float t1, float& t, float& a_out,
                      float& b_out) {
    int ka = type == SRT_PRIM_XY_RECT ? 2 : (type == SRT_PRIM_XZ_RECT ? 1 : 0);
    int aa = type == SRT_PRIM_YZ_RECT ? 1 : 0;
    int ab = type == SRT_PRIM_XY_RECT ? 1 : 2;
    float tt = xdiv(xsub(k, comp(r.o, ka)), comp(r.d, ka));
    if (tt < t0 || tt > t1) return false;
    float a = xadd(comp(r.o, aa), xmul(tt, comp(r.d, aa)));
    float b = xadd(comp(r.o, ab), xmul(tt, comp(r.d, ab)));
    if (a < q0.x || a > q0.y || b < q0.z || b > q0.w) return false;
    t = tt;
    a_out = a;
    b_out = b;
    return true;
}

// triangle.h:108-188.  Ignores (t0, t1) as the reference does; t is in units of the NORMALISED
// direction (triangle.h:128).  `(double)x < 0.0001` == `x <= 0.0001f` for float x.
SRT_DEV bool tri_hit_face(bool front, V3 p0, V3 p1, V3 p2, const Ray& r, float& t, float& u, float& v) {
    V3 e1 = front ? p1 - p0 : p0 - p1;
    V3 e2 = front ? p2 - p0 : p2 - p1;
    V3 dir = r.d / length(r.d);
    V3 P = cross(dir, e2);
    float det = dot(e1, P);
    V3 T;
    if (det > 0.f) {
        T = r.o - p0;
    } else {
        T = p0 - r.o;
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
SRT_DEV bool tri_hit(V3 p0, V3 p1, V3 p2, const Ray& r, bool is_medium, float& t, float& u, float& v) {
    bool h = tri_hit_face(true, p0, p1, p2, r, t, u, v);
    if (!h && is_medium) h = tri_hit_face(false, p0, p1, p2, r, t, u, v);  // triangle.h:108-115
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

// One primitive of prims[] against a ray in the primitive's own frame.
template <bool COUNT>
SRT_DEV bool prim_hit(const DScene& S, const SrtPrim& pr, const Ray& r, float t_min, float t_max, bool is_medium,
                      float& t, Tally<COUNT> tl) {
    tl.prim();
    switch (pr.type) {
        case SRT_PRIM_SPHERE: {
            float4 s = __ldg(&S.spheres[pr.index]);
            return sphere_hit(v3(s.x, s.y, s.z), s.w, r, t_min, t_max, t);
        }
        case SRT_PRIM_MOVING_SPHERE: {
            const float* m = S.msph + 9 * (size_t)pr.index;
            return sphere_hit(msphere_center(m, r.time), m[8], r, t_min, t_max, t);
        }
        case SRT_PRIM_XY_RECT:
        case SRT_PRIM_XZ_RECT:
        case SRT_PRIM_YZ_RECT: {
            float4 q0 = __ldg(&S.rects[2 * pr.index]);
            float k = __ldg(&S.rects[2 * pr.index + 1]).x;
            float a, b;
            return rect_hit(pr.type, q0, k, r, t_min, t_max, t, a, b);
        }
        case SRT_PRIM_TRIANGLE: {
            const float* p = S.tri_pos + 9 * (size_t)pr.index;
            float u, v;
            return tri_hit(v3(p), v3(p + 3), v3(p + 6), r, is_medium, t, u, v);
        }
        case SRT_PRIM_ENV_SPHERE: {
            float4 s = __ldg(&S.spheres[pr.index]);
            t = env_sphere_t(v3(s.x, s.y, s.z), s.w, r);
            return true;
        }
        default:
            return false;
    }
}

// hitable_list::hit (hitable_list.h:21-33) over one group: in order, t_max narrowed to the last
// accepted t (the double closest_so_far holds a float value, so the narrowing is exact).
template <bool COUNT>
SRT_DEV bool group_hit(const DScene& S, int group, const Ray& r, float t_min, float t_max, bool is_medium, float& t,
                       int32_t& prim, Tally<COUNT> tl) {
    SrtGroup g = S.groups[group];
    bool any = false;
    float closest = t_max;
    for (int i = 0; i < g.prim_count; i++) {
        SrtPrim pr = S.prims[g.first_prim + i];
        float tt;
        if (prim_hit(S, pr, r, t_min, closest, is_medium, tt, tl)) {
            any = true;
            closest = tt;
            prim = g.first_prim + i;
        }
    }
    t = closest;
    return any;
}

// ---- the reference's aabb::hit (aabb.h:33-49), bit for bit -------------------------------------
struct RayInv {
    float ix, iy, iz;
};
SRT_DEV bool ref_aabb_hit(const float* lo, const float* hi, const Ray& r, RayInv inv, float tmin, float tmax) {
    {
        float t0 = xmul(xsub(lo[0], r.o.x), inv.ix), t1 = xmul(xsub(hi[0], r.o.x), inv.ix);
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
        float t0 = xmul(xsub(lo[1], r.o.y), inv.iy), t1 = xmul(xsub(hi[1], r.o.y), inv.iy);
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
        float t0 = xmul(xsub(lo[2], r.o.z), inv.iz), t1 = xmul(xsub(hi[2], r.o.z), inv.iz);
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
        float lo[3] = {a.x, a.y, a.z}, hi[3] = {b.x, b.y, b.z};
        if (!ref_aabb_hit(lo, hi, r, inv, t_min, t_max)) return false;
        int left = __float_as_int(a.w), right = __float_as_int(b.w);
        if (left < 0) return true;
        int mid = __ldg(&root[idx].mid);
        idx = rank < mid ? left : right;
    }
}

// ---- wide-BVH traversal -------------------------------------------------------------------------
// Result of a reference `bvh_node::hit(r, t_min, t_max)` (bvh.h:64-93): over all reachable leaves
// that hit, the smallest t; on equal t the leaf later in the in-order sequence (the right child
// wins every `left_rec.t < right_rec.t` tie).  Leaves are tested with the un-narrowed entry
// (t_min, t_max), as in the reference; OUR tree only orders the search and prunes conservatively.
constexpr int kStackSize = 48;

template <bool COUNT>
SRT_DEV bool bvh_hit(const DScene& S, const DBvh& B, const Ray& r, float t_min, float t_max, bool is_medium,
                     float& t_out, int32_t& prim_out, Tally<COUNT> tl) {
    RayInv inv{xdiv(1.0f, r.d.x), xdiv(1.0f, r.d.y), xdiv(1.0f, r.d.z)};
    // pruning scale: triangle t is a distance, box slabs are parametric
    float len = length(r.d);
    float tscale = B.kind == BVH_KIND_SPHERE ? 1.0f : (B.kind == BVH_KIND_TRIANGLE ? len : fminf(1.0f, len));
    if (!(tscale == tscale)) return false;  // NaN direction: nothing sensible to report
    const WideNode4* nodes = S.wide_nodes + B.node_base;
    const RefNode* ref = S.ref_nodes + B.ref_base;
    const int32_t* ranks = S.leaf_rank + B.leaf_base;
    float best_t = CUDART_INF_F;
    int best_rank = -1;
    int32_t best_prim = -1;

    int32_t stack[kStackSize];
    int sp = 0;
    int32_t cur = 0;  // >= 0 inner node; < 0 leaf reference packed as ~(first << 3 | count)
    for (;;) {
        if (cur >= 0) {
            const float4* q = reinterpret_cast<const float4*>(nodes + cur);
            float4 lox = __ldg(q), loy = __ldg(q + 1), loz = __ldg(q + 2);
            float4 hix = __ldg(q + 3), hiy = __ldg(q + 4), hiz = __ldg(q + 5);
            int4 child = __ldg(reinterpret_cast<const int4*>(q + 6));
            int4 count = __ldg(reinterpret_cast<const int4*>(q + 7));
            tl.node();
            float tn[4];
            int32_t ref4[4];
            const float* LX = &lox.x;
            const float* LY = &loy.x;
            const float* LZ = &loz.x;
            const float* HX = &hix.x;
            const float* HY = &hiy.x;
            const float* HZ = &hiz.x;
            const int* CH = &child.x;
            const int* CN = &count.x;
            int nhit = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                float ax = (LX[i] - r.o.x) * inv.ix, bx = (HX[i] - r.o.x) * inv.ix;
                float ay = (LY[i] - r.o.y) * inv.iy, by = (HY[i] - r.o.y) * inv.iy;
                float az = (LZ[i] - r.o.z) * inv.iz, bz = (HZ[i] - r.o.z) * inv.iz;
                float n = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
                float f = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
                // conservative: widen the interval by a few ulps before deciding
                n = n - fabsf(n) * 4e-6f;
                f = f + fabsf(f) * 4e-6f;
                bool empty = CH[i] < 0 && CN[i] == 0;
                // lower bound of any leaf t inside this box, in the leaves' own units
                float lb = n > 0.0f ? n * tscale : (B.kind == BVH_KIND_TRIANGLE ? 0.0f : n);
                bool h = !empty && n <= f && f >= fminf(t_min, 0.0f) && lb <= best_t;
                if (h) {
                    tn[nhit] = n;
                    ref4[nhit] = CH[i] >= 0 ? CH[i] : ~(((~CH[i]) << 3) | CN[i]);
                    nhit++;
                }
            }
            if (nhit == 0) {
                if (sp == 0) break;
                cur = stack[--sp];
                continue;
            }
            // insertion sort by entry distance (nearest first), push the rest far-to-near
            for (int i = 1; i < nhit; i++) {
                float kt = tn[i];
                int32_t kr = ref4[i];
                int j = i - 1;
                while (j >= 0 && tn[j] > kt) {
                    tn[j + 1] = tn[j];
                    ref4[j + 1] = ref4[j];
                    j--;
                }
                tn[j + 1] = kt;
                ref4[j + 1] = kr;
            }
            for (int i = nhit - 1; i >= 1; i--)
                if (sp < kStackSize) stack[sp++] = ref4[i];
            cur = ref4[0];
            continue;
        }
        // leaf
        {
            int packed = ~cur;
            int first = packed >> 3, cnt = packed & 7;
            for (int s = first; s < first + cnt; s++) {
                int rank = __ldg(&ranks[s]);
                float tt;
                int32_t pp;
                bool h;
                if (B.kind == BVH_KIND_SPHERE) {
                    float4 sp4 = __ldg(&S.leaf_geom[B.geom_base + s]);
                    tl.prim();
                    h = sphere_hit(v3(sp4.x, sp4.y, sp4.z), sp4.w, r, t_min, t_max, tt);
                    pp = -2;  // resolved from the rank at the end
                } else if (B.kind == BVH_KIND_TRIANGLE) {
                    const float4* g = S.leaf_geom + B.geom_base + 3 * (size_t)s;
                    float4 a = __ldg(g), b = __ldg(g + 1), c = __ldg(g + 2);
                    float u, v;
                    tl.prim();
                    h = tri_hit(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), r, is_medium, tt, u, v);
                    pp = -2;
                } else {
                    int group = __ldg(&S.bvh_items[B.first_item + rank]);
                    h = group_hit(S, group, r, t_min, t_max, is_medium, tt, pp, tl);
                }
                if (h && (tt < best_t || (tt == best_t && rank > best_rank))) {
                    if (ref_reachable(ref, rank, r, inv, t_min, t_max, tl)) {
                        best_t = tt;
                        best_rank = rank;
                        best_prim = pp;
                    }
                }
            }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    if (best_rank < 0) return false;
    if (best_prim == -2) best_prim = S.groups[S.bvh_items[B.first_item + best_rank]].first_prim;
    t_out = best_t;
    prim_out = best_prim;
    return true;
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

SRT_DEV Ray entry_local_ray(const DScene& S, const SrtEntry& e, Ray r) {
    for (int i = 0; i < e.op_count; i++) r = xform_ray_in(S.ops[e.first_op + i], r);
    return r;
}

// The geometry below an entry's transform chain (group or bvh), without the medium wrapper.
template <bool COUNT>
SRT_DEV bool entry_geom_hit(const DScene& S, const SrtEntry& e, const Ray& world_ray, float t_min, float t_max,
                            bool is_medium, float& t, int32_t& prim, Tally<COUNT> tl) {
    Ray r = entry_local_ray(S, e, world_ray);
    if (e.geom_kind == SRT_GEOM_GROUP) return group_hit(S, e.geom_index, r, t_min, t_max, is_medium, t, prim, tl);
    DBvh B = S.bvhs[e.geom_index];
    return bvh_hit(S, B, r, t_min, t_max, is_medium, t, prim, tl);
}

// constant_medium::hit (constant_medium.h:19-50): two boundary hits, then an exponential
// free-flight distance.  The dead debug draw at :20 is not reproduced (counter RNG).
template <bool COUNT>
SRT_DEV bool medium_hit(const DScene& S, const SrtEntry& e, const Ray& r, float t_min, float t_max, Philox& rng,
                        float& t, Tally<COUNT> tl) {
    float t1, t2;
    int32_t p1, p2;
    if (!entry_geom_hit(S, e, r, -kFltMax, kFltMax, true, t1, p1, tl)) return false;
    float t_min2 = (float)((double)t1 + 0.0001);
    if (!entry_geom_hit(S, e, r, t_min2, kFltMax, true, t2, p2, tl)) return false;
    if (t1 < t_min) t1 = t_min;
    if (t2 > t_max) t2 = t_max;
    if (t1 >= t2) return false;
    if (t1 < 0.f) t1 = 0.f;
    float len = length(r.d);
    float distance_inside = xmul(xsub(t2, t1), len);
    float density = S.media[e.medium].density;
    float hit_distance = (float)(-(double)xdiv(1.0f, density) * log(rand_d(rng)));
    if (hit_distance < distance_inside) {
        t = xadd(t1, xdiv(hit_distance, len));
        return true;
    }
    return false;
}

// world->hit(r, t_min, t_max, rec) (Raytracing_n.cpp:58): the top-level hitable_list.
template <bool COUNT>
SRT_DEV Hit world_hit(const DScene& S, const Ray& r, float t_min, float t_max, Philox& rng, Tally<COUNT> tl) {
    Hit h;
    h.prim = -1;
    h.entry = -1;
    h.t = t_max;
    float closest = t_max;
    for (int ei = 0; ei < S.n_entries; ei++) {
        SrtEntry e = S.entries[ei];
        float t;
        int32_t prim = -1;
        bool got;
        if (e.medium >= 0) {
            got = medium_hit(S, e, r, t_min, closest, rng, t, tl);
            prim = S.n_prims + e.medium;
        } else {
            got = entry_geom_hit(S, e, r, t_min, closest, false, t, prim, tl);
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

// Re-runs the winning primitive's test to fill u, v, p, normal exactly as its hit() does, then
// unwinds flip_normals and the instance chain (aarect.h:152-159, hitable.h:44-52, :109-132).
template <bool WANT_UV>
SRT_DEV HitRecord make_record(const DScene& S, const Ray& world_ray, const Hit& h) {
    HitRecord rec;
    rec.t = h.t;
    rec.u = rec.v = 0.f;
    rec.material = -1;
    if (h.prim >= S.n_prims) {  // constant_medium.h:39-44
        rec.p = point_at(world_ray, h.t);
        rec.normal = v3(1.f, 0.f, 0.f);
        rec.material = S.media[h.prim - S.n_prims].phase_material;
        return rec;
    }
    SrtEntry e = S.entries[h.entry];
    Ray r = entry_local_ray(S, e, world_ray);
    SrtPrim pr = S.prims[h.prim];
    rec.material = pr.material;
    switch (pr.type) {
        case SRT_PRIM_SPHERE:
        case SRT_PRIM_ENV_SPHERE: {
            float4 s = __ldg(&S.spheres[pr.index]);
            V3 c = v3(s.x, s.y, s.z);
            rec.p = point_at(r, h.t);
            V3 q = (rec.p - c) / s.w;
            if (WANT_UV) get_sphere_uv(q, rec.u, rec.v);
            rec.normal = pr.type == SRT_PRIM_SPHERE ? q : (c - rec.p) / s.w;
            break;
        }
        case SRT_PRIM_MOVING_SPHERE: {
            const float* m = S.msph + 9 * (size_t)pr.index;
            rec.p = point_at(r, h.t);
            rec.normal = (rec.p - msphere_center(m, r.time)) / m[8];
            break;
        }
        case SRT_PRIM_XY_RECT:
        case SRT_PRIM_XZ_RECT:
        case SRT_PRIM_YZ_RECT: {
            float4 q0 = __ldg(&S.rects[2 * pr.index]);
            int aa = pr.type == SRT_PRIM_YZ_RECT ? 1 : 0;
            int ab = pr.type == SRT_PRIM_XY_RECT ? 1 : 2;
            float a = xadd(comp(r.o, aa), xmul(h.t, comp(r.d, aa)));
            float b = xadd(comp(r.o, ab), xmul(h.t, comp(r.d, ab)));
            rec.u = xdiv(xsub(a, q0.x), xsub(q0.y, q0.x));
            rec.v = xdiv(xsub(b, q0.z), xsub(q0.w, q0.z));
            rec.p = point_at(r, h.t);
            rec.normal = pr.type == SRT_PRIM_XY_RECT ? v3(0.f, 0.f, 1.f)
                                                     : (pr.type == SRT_PRIM_XZ_RECT ? v3(0.f, 1.f, 0.f) : v3(1.f, 0.f, 0.f));
            break;
        }
        case SRT_PRIM_TRIANGLE: {
            const float* p = S.tri_pos + 9 * (size_t)pr.index;
            const float* n = S.tri_nrm + 9 * (size_t)pr.index;
            const float* uv = S.tri_uv + 6 * (size_t)pr.index;
            V3 p0 = v3(p), p1 = v3(p + 3), p2 = v3(p + 6);
            float t, u = 0.f, v = 0.f;
            // a hit found through a medium boundary may be the back-face variant
            tri_hit(p0, p1, p2, r, true, t, u, v);
            float w = xsub(xsub(1.0f, u), v);
            rec.u = xadd(xadd(xmul(w, uv[0]), xmul(u, uv[2])), xmul(v, uv[4]));
            rec.v = xadd(xadd(xmul(w, uv[1]), xmul(u, uv[3])), xmul(v, uv[5]));
            rec.normal = unit_vector(w * v3(n) + u * v3(n + 3) + v * v3(n + 6));  // FLAT_NORMAL 1, triangle.h:179-183
            rec.p = w * p0 + u * p1 + v * p2;
            break;
        }
        default:
            rec.p = point_at(r, h.t);
            rec.normal = v3(0.f, 0.f, 0.f);
            break;
    }
    if (pr.flags & SRT_PRIM_FLIP) rec.normal = -rec.normal;
    for (int i = e.op_count - 1; i >= 0; i--) xform_rec_out(S.ops[e.first_op + i], rec.p, rec.normal);
    return rec;
}

}  // namespace srt
#endif
