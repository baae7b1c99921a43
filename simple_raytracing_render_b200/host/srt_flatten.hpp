// srt_flatten.hpp — host side of the drop-in boundary, C++ half.
//
// Walks a scene built with the reference's OWN classes (hitable / hitable_list / bvh_node /
// sphere / moving_sphere / x?_rect / box / triangle / flip_normals / translate / rotate_x /
// rotate_y / constant_medium, material.h, texture.h, camera.h) and flattens it into the arrays of
// SrtSceneDesc (include/srt_ffi.h).  It must be compiled in a translation unit that has already
// included the reference headers (it names their classes and reads their public members); it
// contains no reference code.  Every field it needs is public in the reference.
//
// Include order: the reference's mathf.h:6-8 defines the macros __m/__c/__a, which break any
// standard header included after it, so include <vector>, <unordered_map>, <string>, <cstring>
// and "srt_ffi.h" BEFORE the reference headers (INTEGRATION.md shows the exact preamble).
//
// Flattening rules:
//   world hitable_list element        -> one SrtEntry (nested hitable_lists are inlined: a list
//                                        inside a list hits exactly like its elements in place,
//                                        hitable_list.h:21-33)
//   translate / rotate_y / rotate_x   -> SrtXformOp, outermost first (hitable.h:44-52,109-132,180-203)
//   flip_normals                      -> folded into SrtPrim.flags when the chain ends in a group,
//                                        SRT_OP_FLIP when it ends in a bvh (aarect.h:149-171)
//   box                               -> group of its six rects (box.h:18-33)
//   bvh_node                          -> SrtBvh: leaves in the tree's in-order order (bvh.h:96-119)
//   constant_medium                   -> SrtEntry.medium, boundary flattened below it
//   hlist                             -> SrtLight per element, flip_normals looked through
#ifndef SRT_FLATTEN_HPP
#define SRT_FLATTEN_HPP

#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "srt_ffi.h"

struct SrtFlatScene {
    std::vector<SrtEntry> entries;
    std::vector<SrtXformOp> ops;
    std::vector<SrtGroup> groups;
    std::vector<SrtBvh> bvhs;
    std::vector<int32_t> bvh_items;
    std::vector<float> bvh_item_boxes;  // [n_bvh_items][6]
    std::vector<SrtMedium> media;
    std::vector<SrtPrim> prims;
    std::vector<float> spheres, msph, rects, tri_pos, tri_nrm, tri_uv;
    std::vector<SrtLight> lights;
    std::vector<SrtMaterial> materials;
    std::vector<SrtTexture> textures;
    std::vector<SrtImage> images;
    std::vector<const double*> merl;
    std::vector<float> perlin_ranvec;
    std::vector<int32_t> perlin_px, perlin_py, perlin_pz;
    SrtCamera camera;
    std::string error;

    // prim id -> the reference object it came from (used by the oracle harness to report ids)
    std::vector<const hitable*> prim_objects;
    std::vector<const hitable*> medium_objects;  // media[] index -> constant_medium

    std::unordered_map<const material*, int32_t> mat_index;
    std::unordered_map<const texture*, int32_t> tex_index;
    std::unordered_map<const unsigned char*, int32_t> img_index;
    std::unordered_map<const double*, int32_t> merl_index;

    SrtFlatScene() { std::memset(&camera, 0, sizeof(camera)); }

    int32_t n_spheres() const { return (int32_t)(spheres.size() / 4); }
    int32_t n_msph() const { return (int32_t)(msph.size() / 9); }
    int32_t n_rects() const { return (int32_t)(rects.size() / 5); }
    int32_t n_tris() const { return (int32_t)(tri_pos.size() / 9); }

    // The description points into this object's vectors: keep it alive while the desc is in use.
    SrtSceneDesc desc() const {
        SrtSceneDesc d;
        std::memset(&d, 0, sizeof(d));
        d.abi_version = SRT_ABI_VERSION;
        d.entries = entries.data();
        d.n_entries = (int32_t)entries.size();
        d.ops = ops.data();
        d.n_ops = (int32_t)ops.size();
        d.groups = groups.data();
        d.n_groups = (int32_t)groups.size();
        d.bvhs = bvhs.data();
        d.n_bvhs = (int32_t)bvhs.size();
        d.bvh_items = bvh_items.data();
        d.n_bvh_items = (int32_t)bvh_items.size();
        d.bvh_item_boxes = bvh_item_boxes.data();
        d.media = media.data();
        d.n_media = (int32_t)media.size();
        d.prims = prims.data();
        d.n_prims = (int32_t)prims.size();
        d.spheres = spheres.data();
        d.n_spheres = n_spheres();
        d.msph = msph.data();
        d.n_msph = n_msph();
        d.rects = rects.data();
        d.n_rects = n_rects();
        d.tri_pos = tri_pos.data();
        d.tri_nrm = tri_nrm.data();
        d.tri_uv = tri_uv.data();
        d.n_tris = n_tris();
        d.lights = lights.data();
        d.n_lights = (int32_t)lights.size();
        d.materials = materials.data();
        d.n_materials = (int32_t)materials.size();
        d.textures = textures.data();
        d.n_textures = (int32_t)textures.size();
        d.images = images.data();
        d.n_images = (int32_t)images.size();
        if (!perlin_ranvec.empty()) {
            d.perlin_ranvec = perlin_ranvec.data();
            d.perlin_perm_x = perlin_px.data();
            d.perlin_perm_y = perlin_py.data();
            d.perlin_perm_z = perlin_pz.data();
        }
        d.merl_tables = merl.empty() ? nullptr : merl.data();
        d.n_merl = (int32_t)merl.size();
        d.camera = camera;
        return d;
    }
};

namespace srt_flatten_detail {

inline void put3(float* dst, const vec3& v) {
    dst[0] = v.e[0];
    dst[1] = v.e[1];
    dst[2] = v.e[2];
}

inline bool fail(SrtFlatScene& fs, const std::string& why) {
    if (fs.error.empty()) fs.error = why;
    return false;
}

inline int32_t add_texture(SrtFlatScene& fs, const texture* t);

inline int32_t add_image(SrtFlatScene& fs, const image_texture* it) {
    auto f = fs.img_index.find(it->data);
    if (f != fs.img_index.end()) return f->second;
    SrtImage im;
    im.data = it->data;
    im.nx = it->nx;
    im.ny = it->ny;
    int32_t id = (int32_t)fs.images.size();
    fs.images.push_back(im);
    fs.img_index[it->data] = id;
    return id;
}

inline int32_t add_texture(SrtFlatScene& fs, const texture* t) {
    if (!t) return -1;
    auto f = fs.tex_index.find(t);
    if (f != fs.tex_index.end()) return f->second;
    SrtTexture st;
    std::memset(&st, 0, sizeof(st));
    st.a = st.b = -1;
    if (const constant_texture* c = dynamic_cast<const constant_texture*>(t)) {
        st.type = SRT_TEX_CONSTANT;
        put3(st.c, c->color);
    } else if (const checker_texture* c = dynamic_cast<const checker_texture*>(t)) {
        st.type = SRT_TEX_CHECKER;
        st.a = add_texture(fs, c->even);
        st.b = add_texture(fs, c->odd);
    } else if (const noise_texture* n = dynamic_cast<const noise_texture*>(t)) {
        st.type = SRT_TEX_NOISE;
        st.c[0] = n->scale;
        if (fs.perlin_ranvec.empty()) {  // perlin.h:60-63, :94-97: process-wide static tables
            fs.perlin_ranvec.resize(256 * 3);
            fs.perlin_px.resize(256);
            fs.perlin_py.resize(256);
            fs.perlin_pz.resize(256);
            for (int i = 0; i < 256; i++) {
                put3(&fs.perlin_ranvec[3 * i], perlin::ranvec[i]);
                fs.perlin_px[i] = perlin::perm_x[i];
                fs.perlin_py[i] = perlin::perm_y[i];
                fs.perlin_pz[i] = perlin::perm_z[i];
            }
        }
    } else if (const image_texture* im = dynamic_cast<const image_texture*>(t)) {
        st.type = SRT_TEX_IMAGE;
        st.a = add_image(fs, im);
    } else {
        fail(fs, "unknown texture class");
        return -1;
    }
    int32_t id = (int32_t)fs.textures.size();
    fs.textures.push_back(st);
    fs.tex_index[t] = id;
    return id;
}

inline int32_t add_material(SrtFlatScene& fs, const material* m) {
    if (!m) return -1;
    auto f = fs.mat_index.find(m);
    if (f != fs.mat_index.end()) return f->second;
    SrtMaterial sm;
    std::memset(&sm, 0, sizeof(sm));
    sm.texture = -1;
    if (const lambertian* l = dynamic_cast<const lambertian*>(m)) {
        sm.type = SRT_MAT_LAMBERTIAN;
        sm.texture = add_texture(fs, l->albedo);
    } else if (const orennayar* o = dynamic_cast<const orennayar*>(m)) {
        sm.type = SRT_MAT_ORENNAYAR;
        sm.texture = add_texture(fs, o->albedo);
        sm.p[0] = o->A;
        sm.p[1] = o->B;
    } else if (const beckmann* b = dynamic_cast<const beckmann*>(m)) {
        sm.type = SRT_MAT_BECKMANN;
        sm.texture = add_texture(fs, b->albedo);
        // the distribution's alphas are private (microfacet_distribution.h:150-152); recompute
        // them exactly as the constructor does (material.h:153-157)
        sm.p[0] = BeckmannDistribution::RoughnessToAlpha(b->roughx);
        sm.p[1] = BeckmannDistribution::RoughnessToAlpha(b->roughy);
    } else if (const brdfmaterial* bm = dynamic_cast<const brdfmaterial*>(m)) {
        sm.type = SRT_MAT_BRDF;
        put3(sm.p, bm->albedo);
        int32_t ti;
        auto mf = fs.merl_index.find(bm->brdf_data);
        if (mf != fs.merl_index.end()) {
            ti = mf->second;
        } else {
            ti = (int32_t)fs.merl.size();
            fs.merl.push_back(bm->brdf_data);
            fs.merl_index[bm->brdf_data] = ti;
        }
        sm.p[3] = (float)ti;
    } else if (const metal* me = dynamic_cast<const metal*>(m)) {
        sm.type = SRT_MAT_METAL;
        put3(sm.p, me->albedo);
        sm.p[3] = me->fuzz;
    } else if (const dielectric* d = dynamic_cast<const dielectric*>(m)) {
        sm.type = SRT_MAT_DIELECTRIC;
        sm.p[0] = d->ref_idx;
    } else if (const diffuse_light* dl = dynamic_cast<const diffuse_light*>(m)) {
        sm.type = SRT_MAT_DIFFUSE_LIGHT;
        sm.texture = add_texture(fs, dl->emit);
    } else if (const isotropic* is = dynamic_cast<const isotropic*>(m)) {
        sm.type = SRT_MAT_ISOTROPIC;
        sm.texture = add_texture(fs, is->albedo);
    } else {
        fail(fs, "unknown material class");
        return -1;
    }
    int32_t id = (int32_t)fs.materials.size();
    fs.materials.push_back(sm);
    fs.mat_index[m] = id;
    return id;
}

// Append the data row of a leaf primitive; returns false if `h` is not a leaf primitive.
inline bool add_prim_data(SrtFlatScene& fs, const hitable* h, int32_t& type, int32_t& index,
                          const material*& mat) {
    if (const sphere* s = dynamic_cast<const sphere*>(h)) {
        type = SRT_PRIM_SPHERE;
        index = fs.n_spheres();
        float row[4] = {s->center.e[0], s->center.e[1], s->center.e[2], s->radius};
        fs.spheres.insert(fs.spheres.end(), row, row + 4);
        mat = s->mat_ptr;
        return true;
    }
    if (const env_sphere* s = dynamic_cast<const env_sphere*>(h)) {
        type = SRT_PRIM_ENV_SPHERE;
        index = fs.n_spheres();
        float row[4] = {s->center.e[0], s->center.e[1], s->center.e[2], s->radius};
        fs.spheres.insert(fs.spheres.end(), row, row + 4);
        mat = s->mat_ptr;
        return true;
    }
    if (const moving_sphere* s = dynamic_cast<const moving_sphere*>(h)) {
        type = SRT_PRIM_MOVING_SPHERE;
        index = fs.n_msph();
        float row[9];
        put3(row, s->center0);
        put3(row + 3, s->center1);
        row[6] = s->time0;
        row[7] = s->time1;
        row[8] = s->radius;
        fs.msph.insert(fs.msph.end(), row, row + 9);
        mat = s->mat_ptr;
        return true;
    }
    if (const xy_rect* r = dynamic_cast<const xy_rect*>(h)) {
        type = SRT_PRIM_XY_RECT;
        index = fs.n_rects();
        float row[5] = {r->x0, r->x1, r->y0, r->y1, r->k};
        fs.rects.insert(fs.rects.end(), row, row + 5);
        mat = r->mp;
        return true;
    }
    if (const xz_rect* r = dynamic_cast<const xz_rect*>(h)) {
        type = SRT_PRIM_XZ_RECT;
        index = fs.n_rects();
        float row[5] = {r->x0, r->x1, r->z0, r->z1, r->k};
        fs.rects.insert(fs.rects.end(), row, row + 5);
        mat = r->mp;
        return true;
    }
    if (const yz_rect* r = dynamic_cast<const yz_rect*>(h)) {
        type = SRT_PRIM_YZ_RECT;
        index = fs.n_rects();
        float row[5] = {r->y0, r->y1, r->z0, r->z1, r->k};
        fs.rects.insert(fs.rects.end(), row, row + 5);
        mat = r->mp;
        return true;
    }
    if (const triangle* t = dynamic_cast<const triangle*>(h)) {
        type = SRT_PRIM_TRIANGLE;
        index = fs.n_tris();
        float pos[9], nrm[9], uv[6];
        put3(pos, t->p0);
        put3(pos + 3, t->p1);
        put3(pos + 6, t->p2);
        put3(nrm, t->n0);
        put3(nrm + 3, t->n1);
        put3(nrm + 6, t->n2);
        for (int k = 0; k < 3; k++) {
            uv[2 * k] = t->uvs[k].e[0];
            uv[2 * k + 1] = t->uvs[k].e[1];
        }
        fs.tri_pos.insert(fs.tri_pos.end(), pos, pos + 9);
        fs.tri_nrm.insert(fs.tri_nrm.end(), nrm, nrm + 9);
        fs.tri_uv.insert(fs.tri_uv.end(), uv, uv + 6);
        mat = t->mp;
        return true;
    }
    return false;
}

// Append the prims of one list-semantics leaf (a primitive, flip_normals(primitive), a box, or a
// hitable_list of those) to fs.prims; `flip` carries the parity of enclosing flip_normals.
inline bool add_leaf_prims(SrtFlatScene& fs, const hitable* h, bool flip) {
    if (const flip_normals* f = dynamic_cast<const flip_normals*>(h))
        return add_leaf_prims(fs, f->ptr, !flip);
    if (const box* b = dynamic_cast<const box*>(h)) return add_leaf_prims(fs, b->list_ptr, flip);
    if (const hitable_list* l = dynamic_cast<const hitable_list*>(h)) {
        for (int i = 0; i < l->list_size; i++)
            if (!add_leaf_prims(fs, l->list[i], flip)) return false;
        return true;
    }
    int32_t type, index;
    const material* mat = nullptr;
    if (!add_prim_data(fs, h, type, index, mat))
        return fail(fs, std::string("unsupported leaf hitable: ") + typeid(*h).name());
    SrtPrim p;
    p.type = type;
    p.index = index;
    p.material = add_material(fs, mat);
    p.flags = flip ? SRT_PRIM_FLIP : 0u;
    fs.prims.push_back(p);
    fs.prim_objects.push_back(h);
    return fs.error.empty();
}

inline int32_t add_group(SrtFlatScene& fs, const hitable* h, bool flip) {
    SrtGroup g;
    g.first_prim = (int32_t)fs.prims.size();
    if (!add_leaf_prims(fs, h, flip)) return -1;
    g.prim_count = (int32_t)fs.prims.size() - g.first_prim;
    fs.groups.push_back(g);
    return (int32_t)fs.groups.size() - 1;
}

// In-order leaves of a reference bvh_node tree (left before right; a node whose two children are
// the same object, bvh.h:104-106, contributes that object once).
inline void collect_bvh_leaves(const bvh_node* n, std::vector<const hitable*>& out) {
    const bvh_node* l = dynamic_cast<const bvh_node*>(n->left);
    const bvh_node* r = dynamic_cast<const bvh_node*>(n->right);
    if (l)
        collect_bvh_leaves(l, out);
    else
        out.push_back(n->left);
    if (n->right == n->left) return;
    if (r)
        collect_bvh_leaves(r, out);
    else
        out.push_back(n->right);
}

inline bool is_groupable(const hitable* h) {
    if (const flip_normals* f = dynamic_cast<const flip_normals*>(h)) return is_groupable(f->ptr);
    return dynamic_cast<const bvh_node*>(h) == nullptr && dynamic_cast<const translate*>(h) == nullptr &&
           dynamic_cast<const rotate_y*>(h) == nullptr && dynamic_cast<const rotate_x*>(h) == nullptr &&
           dynamic_cast<const constant_medium*>(h) == nullptr;
}

inline bool flatten_entry(SrtFlatScene& fs, const hitable* h, std::vector<SrtXformOp> chain,
                          int32_t medium) {
    for (;;) {
        SrtXformOp op;
        std::memset(&op, 0, sizeof(op));
        if (const translate* t = dynamic_cast<const translate*>(h)) {
            op.kind = SRT_OP_TRANSLATE;
            op.a = t->offset.e[0];
            op.b = t->offset.e[1];
            op.c = t->offset.e[2];
            chain.push_back(op);
            h = t->ptr;
        } else if (const rotate_y* r = dynamic_cast<const rotate_y*>(h)) {
            op.kind = SRT_OP_ROTATE_Y;
            op.a = r->sin_theta;
            op.b = r->cos_theta;
            chain.push_back(op);
            h = r->ptr;
        } else if (const rotate_x* r = dynamic_cast<const rotate_x*>(h)) {
            op.kind = SRT_OP_ROTATE_X;
            op.a = r->sin_theta;
            op.b = r->cos_theta;
            chain.push_back(op);
            h = r->ptr;
        } else if (const flip_normals* f = dynamic_cast<const flip_normals*>(h)) {
            op.kind = SRT_OP_FLIP;
            chain.push_back(op);
            h = f->ptr;
        } else if (const constant_medium* cm = dynamic_cast<const constant_medium*>(h)) {
            if (!chain.empty() || medium >= 0)
                return fail(fs, "constant_medium below a transform or another medium is not supported");
            SrtMedium m;
            m.density = cm->density;
            m.phase_material = add_material(fs, cm->phase_function);
            medium = (int32_t)fs.media.size();
            fs.media.push_back(m);
            fs.medium_objects.push_back(cm);
            h = cm->boundary;
        } else {
            break;
        }
    }
    if (const bvh_node* root = dynamic_cast<const bvh_node*>(h)) {
        std::vector<const hitable*> leaves;
        collect_bvh_leaves(root, leaves);
        SrtBvh b;
        b.first_item = (int32_t)fs.bvh_items.size();
        b.item_count = (int32_t)leaves.size();
        for (const hitable* leaf : leaves) {
            if (!is_groupable(leaf))
                return fail(fs, std::string("unsupported bvh_node leaf: ") + typeid(*leaf).name());
            int32_t g = add_group(fs, leaf, false);
            if (g < 0) return false;
            fs.bvh_items.push_back(g);
            aabb lb;  // the leaf's own box as the reference tree saw it (bvh.h:115-117)
            if (!leaf->bounding_box(0, 1, lb)) return fail(fs, "bvh_node leaf without a bounding box");
            float row[6];
            put3(row, lb._min);
            put3(row + 3, lb._max);
            fs.bvh_item_boxes.insert(fs.bvh_item_boxes.end(), row, row + 6);
        }
        {  // every reference bvh_node is built over the time range (0, 1); a root box that is not
           // the union of the leaves' (0, 1) boxes means another range was used
            float u[6] = {3.4e38f, 3.4e38f, 3.4e38f, -3.4e38f, -3.4e38f, -3.4e38f};
            const float* bx = fs.bvh_item_boxes.data() + 6 * (size_t)b.first_item;
            for (int k = 0; k < b.item_count; k++)
                for (int a = 0; a < 3; a++) {
                    if (bx[6 * k + a] < u[a]) u[a] = bx[6 * k + a];
                    if (bx[6 * k + 3 + a] > u[3 + a]) u[3 + a] = bx[6 * k + 3 + a];
                }
            for (int a = 0; a < 3; a++)
                if (u[a] != root->box._min.e[a] || u[3 + a] != root->box._max.e[a])
                    return fail(fs, "bvh_node was not built over the time range (0, 1)");
        }
        fs.bvhs.push_back(b);
        SrtEntry e;
        std::memset(&e, 0, sizeof(e));
        e.first_op = (int32_t)fs.ops.size();
        e.op_count = (int32_t)chain.size();
        fs.ops.insert(fs.ops.end(), chain.begin(), chain.end());
        e.geom_kind = SRT_GEOM_BVH;
        e.geom_index = (int32_t)fs.bvhs.size() - 1;
        e.medium = medium;
        fs.entries.push_back(e);
        return true;
    }
    if (const hitable_list* l = dynamic_cast<const hitable_list*>(h)) {
        // a list whose elements are all leaf primitives stays one group; otherwise inline
        bool all_leaf = true;
        for (int i = 0; i < l->list_size; i++) all_leaf = all_leaf && is_groupable(l->list[i]);
        if (!all_leaf) {
            if (medium >= 0) return fail(fs, "constant_medium over a heterogeneous list is not supported");
            for (int i = 0; i < l->list_size; i++)
                if (!flatten_entry(fs, l->list[i], chain, -1)) return false;
            return true;
        }
    }
    // group terminal: fold flips into the prim flags, keep the other ops
    bool flip = false;
    std::vector<SrtXformOp> kept;
    for (const SrtXformOp& op : chain) {
        if (op.kind == SRT_OP_FLIP)
            flip = !flip;
        else
            kept.push_back(op);
    }
    int32_t g = add_group(fs, h, flip);
    if (g < 0) return false;
    SrtEntry e;
    std::memset(&e, 0, sizeof(e));
    e.first_op = (int32_t)fs.ops.size();
    e.op_count = (int32_t)kept.size();
    fs.ops.insert(fs.ops.end(), kept.begin(), kept.end());
    e.geom_kind = SRT_GEOM_GROUP;
    e.geom_index = g;
    e.medium = medium;
    fs.entries.push_back(e);
    return true;
}

inline bool flatten_light(SrtFlatScene& fs, const hitable* h) {
    while (const flip_normals* f = dynamic_cast<const flip_normals*>(h)) h = f->ptr;  // aarect.h:163-169
    SrtLight L;
    L.type = SRT_PRIM_NONE;
    L.index = -1;
    if (dynamic_cast<const xz_rect*>(h) || dynamic_cast<const sphere*>(h) || dynamic_cast<const triangle*>(h) ||
        dynamic_cast<const env_sphere*>(h)) {
        const material* unused = nullptr;
        add_prim_data(fs, h, L.type, L.index, unused);
    } else if (dynamic_cast<const hitable_list*>(h)) {
        return fail(fs, "nested hitable_list in the light list is not supported");
    }
    // every other class inherits pdf_value = 0, random = (1,0,0) (hitable.h:31-32)
    fs.lights.push_back(L);
    return true;
}

}  // namespace srt_flatten_detail

// Flatten (world, light list, camera) — the three objects every reference scene function hands
// back (e.g. Raytracing_n.cpp:216) — into `fs`.  Returns SRT_OK or a negative SRT_E_* code with
// fs.error set.
inline int srt_flatten_scene(const hitable* world, const hitable* hlist, const camera* cam, SrtFlatScene& fs) {
    using namespace srt_flatten_detail;
    if (!world || !cam) {
        fs.error = "null world or camera";
        return SRT_E_INVALID;
    }
    if (const hitable_list* wl = dynamic_cast<const hitable_list*>(world)) {
        for (int i = 0; i < wl->list_size; i++)
            if (!flatten_entry(fs, wl->list[i], std::vector<SrtXformOp>(), -1)) return SRT_E_UNSUPPORTED;
    } else if (!flatten_entry(fs, world, std::vector<SrtXformOp>(), -1)) {
        return SRT_E_UNSUPPORTED;
    }
    if (hlist) {
        // color() casts light_shape to hitable_list unconditionally (Raytracing_n.cpp:74)
        const hitable_list* ll = dynamic_cast<const hitable_list*>(hlist);
        if (!ll) {
            fs.error = "light_shape must be a hitable_list";
            return SRT_E_UNSUPPORTED;
        }
        for (int i = 0; i < ll->list_size; i++)
            if (!flatten_light(fs, ll->list[i])) return SRT_E_UNSUPPORTED;
    }
    put3(fs.camera.origin, cam->origin);
    put3(fs.camera.lower_left_corner, cam->lower_left_corner);
    put3(fs.camera.horizontal, cam->horizontal);
    put3(fs.camera.vertical, cam->vertical);
    put3(fs.camera.u, cam->u);
    put3(fs.camera.v, cam->v);
    put3(fs.camera.w, cam->w);
    fs.camera.time0 = cam->time0;
    fs.camera.time1 = cam->time1;
    fs.camera.lens_radius = cam->lens_radius;
    return fs.error.empty() ? SRT_OK : SRT_E_UNSUPPORTED;
}

#endif  // SRT_FLATTEN_HPP
