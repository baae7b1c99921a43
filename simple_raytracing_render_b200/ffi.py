"""ctypes mirror of include/srt_ffi.h and the loader of the CUDA backend (libsrt_b200.so).

The structures here are field-for-field the C ones; `FlatScene` is the Python container of a
flattened scene (numpy arrays with the C structs' dtypes) that converts to and from `SrtSceneDesc`.

There is no CPU fallback: `load_backend()` raises if the shared library is missing, and the
library's entry points return SRT_E_NODEVICE / SRT_E_CUDA when no GPU is present.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

SRT_ABI_VERSION = 2

# status codes
SRT_OK, SRT_E_INVALID, SRT_E_UNSUPPORTED, SRT_E_CUDA, SRT_E_NOMEM, SRT_E_NODEVICE = 0, -1, -2, -3, -4, -5

# primitive types
PRIM_SPHERE, PRIM_MOVING_SPHERE, PRIM_XY_RECT, PRIM_XZ_RECT, PRIM_YZ_RECT, PRIM_TRIANGLE, PRIM_ENV_SPHERE = range(7)
PRIM_NONE = -1
PRIM_FLIP = 1

OP_TRANSLATE, OP_ROTATE_Y, OP_ROTATE_X, OP_FLIP = range(4)
GEOM_GROUP, GEOM_BVH = 0, 1

(MAT_LAMBERTIAN, MAT_ORENNAYAR, MAT_BECKMANN, MAT_BRDF, MAT_METAL, MAT_DIELECTRIC, MAT_DIFFUSE_LIGHT,
 MAT_ISOTROPIC) = range(8)
TEX_CONSTANT, TEX_CHECKER, TEX_NOISE, TEX_IMAGE = range(4)

COMPAT_PIXEL_MOD_NY = 1
RENDER_WAVEFRONT = 2
EXCHANGE_HANDLE_BYTES = 64
STATIC_MAX_PIXELS = 512 * 512  # srt_ffi.h SRT_STATIC_MAX_PIXELS (srt_render_image_static)
STATIC_MAX_SAMPLES = 1024      # srt_ffi.h SRT_STATIC_MAX_SAMPLES
RENDER_CORRECTED = 4  # corrected-transport mode (outside parity; include/srt_ffi.h)

MERL_DOUBLES = 3 * 90 * 90 * 180

# numpy dtypes of the C structs (all naturally aligned, no padding surprises)
PRIM_DT = np.dtype([("type", "<i4"), ("index", "<i4"), ("material", "<i4"), ("flags", "<u4")])
GROUP_DT = np.dtype([("first_prim", "<i4"), ("prim_count", "<i4")])
BVH_DT = np.dtype([("first_item", "<i4"), ("item_count", "<i4")])
OP_DT = np.dtype([("kind", "<i4"), ("a", "<f4"), ("b", "<f4"), ("c", "<f4")])
ENTRY_DT = np.dtype([("first_op", "<i4"), ("op_count", "<i4"), ("geom_kind", "<i4"), ("geom_index", "<i4"),
                     ("medium", "<i4"), ("_pad", "<i4")])
MEDIUM_DT = np.dtype([("density", "<f4"), ("phase_material", "<i4")])
LIGHT_DT = np.dtype([("type", "<i4"), ("index", "<i4")])
MATERIAL_DT = np.dtype([("type", "<i4"), ("texture", "<i4"), ("p", "<f4", (6,))])
TEXTURE_DT = np.dtype([("type", "<i4"), ("a", "<i4"), ("b", "<i4"), ("c", "<f4", (3,))])
CAMERA_DT = np.dtype([("origin", "<f4", (3,)), ("lower_left_corner", "<f4", (3,)), ("horizontal", "<f4", (3,)),
                      ("vertical", "<f4", (3,)), ("u", "<f4", (3,)), ("v", "<f4", (3,)), ("w", "<f4", (3,)),
                      ("time0", "<f4"), ("time1", "<f4"), ("lens_radius", "<f4")])


class SrtImage(C.Structure):
    _fields_ = [("data", C.c_void_p), ("nx", C.c_int32), ("ny", C.c_int32)]


class SrtCamera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("lower_left_corner", C.c_float * 3), ("horizontal", C.c_float * 3),
                ("vertical", C.c_float * 3), ("u", C.c_float * 3), ("v", C.c_float * 3), ("w", C.c_float * 3),
                ("time0", C.c_float), ("time1", C.c_float), ("lens_radius", C.c_float)]


class SrtSceneDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32),
        ("entries", C.c_void_p), ("n_entries", C.c_int32),
        ("ops", C.c_void_p), ("n_ops", C.c_int32),
        ("groups", C.c_void_p), ("n_groups", C.c_int32),
        ("bvhs", C.c_void_p), ("n_bvhs", C.c_int32),
        ("bvh_items", C.c_void_p), ("n_bvh_items", C.c_int32),
        ("bvh_item_boxes", C.c_void_p),
        ("media", C.c_void_p), ("n_media", C.c_int32),
        ("prims", C.c_void_p), ("n_prims", C.c_int32),
        ("spheres", C.c_void_p), ("n_spheres", C.c_int32),
        ("msph", C.c_void_p), ("n_msph", C.c_int32),
        ("rects", C.c_void_p), ("n_rects", C.c_int32),
        ("tri_pos", C.c_void_p), ("tri_nrm", C.c_void_p), ("tri_uv", C.c_void_p), ("n_tris", C.c_int32),
        ("lights", C.c_void_p), ("n_lights", C.c_int32),
        ("materials", C.c_void_p), ("n_materials", C.c_int32),
        ("textures", C.c_void_p), ("n_textures", C.c_int32),
        ("images", C.POINTER(SrtImage)), ("n_images", C.c_int32),
        ("perlin_ranvec", C.c_void_p), ("perlin_perm_x", C.c_void_p), ("perlin_perm_y", C.c_void_p),
        ("perlin_perm_z", C.c_void_p),
        ("merl_tables", C.POINTER(C.c_void_p)), ("n_merl", C.c_int32),
        ("camera", SrtCamera),
    ]


class SrtRenderParams(C.Structure):
    _fields_ = [("nx", C.c_int32), ("ny", C.c_int32), ("ns", C.c_int32), ("s_begin", C.c_int32),
                ("s_end", C.c_int32), ("max_depth", C.c_int32), ("seed", C.c_uint32), ("flags", C.c_uint32)]


class SrtCounters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("probe_rays", C.c_uint64),
                ("node_visits", C.c_uint64), ("prim_tests", C.c_uint64), ("reject_iters", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("culled_rays", C.c_uint64), ("stack_overflows", C.c_uint64)]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_ if not n.startswith("_")}


class SrtSceneInfo(C.Structure):
    _fields_ = [("n_prims", C.c_int32), ("n_entries", C.c_int32), ("n_bvhs", C.c_int32), ("n_lights", C.c_int32),
                ("bvh_nodes", C.c_int64), ("bvh_node_bytes", C.c_int64), ("device_bytes", C.c_int64),
                ("sah_cost", C.c_float), ("build_ms", C.c_float), ("features", C.c_uint32),
                ("render_variant", C.c_int32), ("render_registers", C.c_int32), ("render_blocks_per_sm", C.c_int32),
                ("hot_table_bytes", C.c_int64), ("bvh_max_depth", C.c_int32), ("bvh_builder", C.c_int32),
                ("upload_ms", C.c_float), ("_pad", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if not n.startswith("_")}


def _arr(a, dt, shape_tail=()):
    a = np.ascontiguousarray(np.asarray(a, dtype=dt))
    if shape_tail:
        a = a.reshape((-1,) + tuple(shape_tail))
    return a


@dataclass
class FlatScene:
    """A flattened scene: the arrays of SrtSceneDesc as numpy arrays (see include/srt_ffi.h)."""
    entries: np.ndarray = field(default_factory=lambda: np.zeros(0, ENTRY_DT))
    ops: np.ndarray = field(default_factory=lambda: np.zeros(0, OP_DT))
    groups: np.ndarray = field(default_factory=lambda: np.zeros(0, GROUP_DT))
    bvhs: np.ndarray = field(default_factory=lambda: np.zeros(0, BVH_DT))
    bvh_items: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    bvh_item_boxes: np.ndarray = field(default_factory=lambda: np.zeros((0, 6), np.float32))
    media: np.ndarray = field(default_factory=lambda: np.zeros(0, MEDIUM_DT))
    prims: np.ndarray = field(default_factory=lambda: np.zeros(0, PRIM_DT))
    spheres: np.ndarray = field(default_factory=lambda: np.zeros((0, 4), np.float32))
    msph: np.ndarray = field(default_factory=lambda: np.zeros((0, 9), np.float32))
    rects: np.ndarray = field(default_factory=lambda: np.zeros((0, 5), np.float32))
    tri_pos: np.ndarray = field(default_factory=lambda: np.zeros((0, 9), np.float32))
    tri_nrm: np.ndarray = field(default_factory=lambda: np.zeros((0, 9), np.float32))
    tri_uv: np.ndarray = field(default_factory=lambda: np.zeros((0, 6), np.float32))
    lights: np.ndarray = field(default_factory=lambda: np.zeros(0, LIGHT_DT))
    materials: np.ndarray = field(default_factory=lambda: np.zeros(0, MATERIAL_DT))
    textures: np.ndarray = field(default_factory=lambda: np.zeros(0, TEXTURE_DT))
    images: List[np.ndarray] = field(default_factory=list)  # each (ny, nx, 3) uint8
    perlin_ranvec: Optional[np.ndarray] = None  # (256, 3) f4
    perlin_perm: Optional[np.ndarray] = None  # (3, 256) i4  (x, y, z)
    merl: List[np.ndarray] = field(default_factory=list)  # each (MERL_DOUBLES,) f8
    camera: np.ndarray = field(default_factory=lambda: np.zeros((), CAMERA_DT))

    # ---- C view ---------------------------------------------------------------------------
    def to_desc(self):
        """Returns (SrtSceneDesc, keepalive): the struct points into this object's arrays."""
        keep = []

        def ptr(a):
            keep.append(a)
            return a.ctypes.data if a.size else None

        self.normalize()
        d = SrtSceneDesc()
        d.abi_version = SRT_ABI_VERSION
        d.entries, d.n_entries = ptr(self.entries), len(self.entries)
        d.ops, d.n_ops = ptr(self.ops), len(self.ops)
        d.groups, d.n_groups = ptr(self.groups), len(self.groups)
        d.bvhs, d.n_bvhs = ptr(self.bvhs), len(self.bvhs)
        d.bvh_items, d.n_bvh_items = ptr(self.bvh_items), len(self.bvh_items)
        if len(self.bvh_item_boxes) != len(self.bvh_items):
            raise ValueError("bvh_item_boxes must hold one box per bvh item")
        d.bvh_item_boxes = ptr(self.bvh_item_boxes)
        d.media, d.n_media = ptr(self.media), len(self.media)
        d.prims, d.n_prims = ptr(self.prims), len(self.prims)
        d.spheres, d.n_spheres = ptr(self.spheres), len(self.spheres)
        d.msph, d.n_msph = ptr(self.msph), len(self.msph)
        d.rects, d.n_rects = ptr(self.rects), len(self.rects)
        d.tri_pos, d.tri_nrm, d.tri_uv = ptr(self.tri_pos), ptr(self.tri_nrm), ptr(self.tri_uv)
        d.n_tris = len(self.tri_pos)
        d.lights, d.n_lights = ptr(self.lights), len(self.lights)
        d.materials, d.n_materials = ptr(self.materials), len(self.materials)
        d.textures, d.n_textures = ptr(self.textures), len(self.textures)
        if self.images:
            imgs = (SrtImage * len(self.images))()
            for i, im in enumerate(self.images):
                im = np.ascontiguousarray(im, dtype=np.uint8)
                self.images[i] = im
                keep.append(im)
                imgs[i].data = im.ctypes.data
                imgs[i].ny, imgs[i].nx = im.shape[0], im.shape[1]
            keep.append(imgs)
            d.images = imgs
        d.n_images = len(self.images)
        if self.perlin_ranvec is not None:
            d.perlin_ranvec = ptr(self.perlin_ranvec)
            px, py, pz = (np.ascontiguousarray(self.perlin_perm[k]) for k in range(3))
            d.perlin_perm_x, d.perlin_perm_y, d.perlin_perm_z = ptr(px), ptr(py), ptr(pz)
        if self.merl:
            tabs = (C.c_void_p * len(self.merl))()
            for i, t in enumerate(self.merl):
                t = np.ascontiguousarray(t, dtype=np.float64)
                self.merl[i] = t
                keep.append(t)
                tabs[i] = t.ctypes.data
            keep.append(tabs)
            d.merl_tables = tabs
        d.n_merl = len(self.merl)
        C.memmove(C.addressof(d.camera), self.camera.ctypes.data, C.sizeof(SrtCamera))
        keep.append(self)
        return d, keep

    def normalize(self):
        self.entries = _arr(self.entries, ENTRY_DT)
        self.ops = _arr(self.ops, OP_DT)
        self.groups = _arr(self.groups, GROUP_DT)
        self.bvhs = _arr(self.bvhs, BVH_DT)
        self.bvh_items = _arr(self.bvh_items, np.int32)
        self.bvh_item_boxes = _arr(self.bvh_item_boxes, np.float32, (6,))
        self.media = _arr(self.media, MEDIUM_DT)
        self.prims = _arr(self.prims, PRIM_DT)
        self.spheres = _arr(self.spheres, np.float32, (4,))
        self.msph = _arr(self.msph, np.float32, (9,))
        self.rects = _arr(self.rects, np.float32, (5,))
        self.tri_pos = _arr(self.tri_pos, np.float32, (9,))
        self.tri_nrm = _arr(self.tri_nrm, np.float32, (9,))
        self.tri_uv = _arr(self.tri_uv, np.float32, (6,))
        self.lights = _arr(self.lights, LIGHT_DT)
        self.materials = _arr(self.materials, MATERIAL_DT)
        self.textures = _arr(self.textures, TEXTURE_DT)
        self.camera = np.ascontiguousarray(np.asarray(self.camera, dtype=CAMERA_DT))
        if self.perlin_ranvec is not None:
            self.perlin_ranvec = _arr(self.perlin_ranvec, np.float32, (3,))
            self.perlin_perm = np.ascontiguousarray(np.asarray(self.perlin_perm, dtype=np.int32).reshape(3, 256))
        return self

    @staticmethod
    def from_desc(desc_ptr) -> "FlatScene":
        """Deep-copies the arrays behind a `const SrtSceneDesc*` (ctypes pointer or address)."""
        if isinstance(desc_ptr, int):
            desc_ptr = C.cast(desc_ptr, C.POINTER(SrtSceneDesc))
        d = desc_ptr.contents if hasattr(desc_ptr, "contents") else desc_ptr

        def rd(addr, n, dt, tail=()):
            count = int(n) * int(np.prod(tail, dtype=np.int64)) if tail else int(n)
            if not addr or count == 0:
                return np.zeros((0,) + tuple(tail), dt)
            nbytes = count * np.dtype(dt).itemsize
            buf = C.string_at(addr, nbytes)
            return np.frombuffer(buf, dtype=dt).reshape((-1,) + tuple(tail)).copy()

        fs = FlatScene()
        fs.entries = rd(d.entries, d.n_entries, ENTRY_DT)
        fs.ops = rd(d.ops, d.n_ops, OP_DT)
        fs.groups = rd(d.groups, d.n_groups, GROUP_DT)
        fs.bvhs = rd(d.bvhs, d.n_bvhs, BVH_DT)
        fs.bvh_items = rd(d.bvh_items, d.n_bvh_items, np.int32)
        fs.bvh_item_boxes = rd(d.bvh_item_boxes, d.n_bvh_items, np.float32, (6,))
        fs.media = rd(d.media, d.n_media, MEDIUM_DT)
        fs.prims = rd(d.prims, d.n_prims, PRIM_DT)
        fs.spheres = rd(d.spheres, d.n_spheres, np.float32, (4,))
        fs.msph = rd(d.msph, d.n_msph, np.float32, (9,))
        fs.rects = rd(d.rects, d.n_rects, np.float32, (5,))
        fs.tri_pos = rd(d.tri_pos, d.n_tris, np.float32, (9,))
        fs.tri_nrm = rd(d.tri_nrm, d.n_tris, np.float32, (9,))
        fs.tri_uv = rd(d.tri_uv, d.n_tris, np.float32, (6,))
        fs.lights = rd(d.lights, d.n_lights, LIGHT_DT)
        fs.materials = rd(d.materials, d.n_materials, MATERIAL_DT)
        fs.textures = rd(d.textures, d.n_textures, TEXTURE_DT)
        for i in range(d.n_images):
            im = d.images[i]
            fs.images.append(rd(im.data, im.nx * im.ny * 3, np.uint8).reshape(im.ny, im.nx, 3))
        if d.perlin_ranvec:
            fs.perlin_ranvec = rd(d.perlin_ranvec, 256, np.float32, (3,))
            fs.perlin_perm = np.stack([rd(p, 256, np.int32) for p in
                                       (d.perlin_perm_x, d.perlin_perm_y, d.perlin_perm_z)])
        for i in range(d.n_merl):
            fs.merl.append(rd(d.merl_tables[i], MERL_DOUBLES, np.float64))
        cam = np.zeros((), CAMERA_DT)
        C.memmove(cam.ctypes.data, C.addressof(d.camera), C.sizeof(SrtCamera))
        fs.camera = cam
        return fs

    # ---- persistence (tests/golden) -----------------------------------------------------------
    _ARRAYS = ("entries", "ops", "groups", "bvhs", "bvh_items", "bvh_item_boxes", "media", "prims", "spheres", "msph", "rects",
               "tri_pos", "tri_nrm", "tri_uv", "lights", "materials", "textures", "camera")

    def save(self, path, with_images=True):
        self.normalize()
        out = {k: getattr(self, k) for k in self._ARRAYS}
        out["n_images"] = np.int32(len(self.images))
        for i, im in enumerate(self.images):
            out[f"image_{i}"] = im if with_images else np.zeros((1, 1, 3), np.uint8)
        if self.perlin_ranvec is not None:
            out["perlin_ranvec"], out["perlin_perm"] = self.perlin_ranvec, self.perlin_perm
        out["n_merl"] = np.int32(len(self.merl))
        for i, t in enumerate(self.merl):
            out[f"merl_{i}"] = t
        np.savez_compressed(path, **out)

    @staticmethod
    def load(path) -> "FlatScene":
        z = np.load(path)
        fs = FlatScene()
        for k in FlatScene._ARRAYS:
            setattr(fs, k, z[k])
        fs.images = [z[f"image_{i}"] for i in range(int(z["n_images"]))]
        if "perlin_ranvec" in z:
            fs.perlin_ranvec, fs.perlin_perm = z["perlin_ranvec"], z["perlin_perm"]
        fs.merl = [z[f"merl_{i}"] for i in range(int(z["n_merl"]))]
        return fs.normalize()

    def equal(self, other: "FlatScene", images=True) -> List[str]:
        """Names of the arrays that differ bit-for-bit (empty list == identical)."""
        self.normalize(), other.normalize()
        bad = []
        for k in self._ARRAYS:
            a, b = getattr(self, k), getattr(other, k)
            if a.shape != b.shape or a.tobytes() != b.tobytes():
                bad.append(k)
        if images:
            if len(self.images) != len(other.images) or any(
                    x.shape != y.shape or not np.array_equal(x, y) for x, y in zip(self.images, other.images)):
                bad.append("images")
        if (self.perlin_ranvec is None) != (other.perlin_ranvec is None):
            bad.append("perlin")
        elif self.perlin_ranvec is not None and (
                self.perlin_ranvec.tobytes() != other.perlin_ranvec.tobytes()
                or self.perlin_perm.tobytes() != other.perlin_perm.tobytes()):
            bad.append("perlin")
        return bad


# ---- backend loader ---------------------------------------------------------------------------
_HERE = os.path.dirname(os.path.abspath(__file__))
BACKEND_PATH = os.path.join(_HERE, "csrc", "libsrt_b200.so")
_backend = None

_SIGNATURES = {
    "srt_last_error": (C.c_char_p, []),
    "srt_abi_version": (C.c_int, []),
    "srt_context_create": (C.c_int, [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    "srt_context_destroy": (C.c_int, [C.c_void_p]),
    "srt_context_synchronize": (C.c_int, [C.c_void_p]),
    "srt_scene_create": (C.c_int, [C.c_void_p, C.POINTER(SrtSceneDesc), C.POINTER(C.c_void_p)]),
    "srt_scene_destroy": (C.c_int, [C.c_void_p]),
    "srt_scene_info": (C.c_int, [C.c_void_p, C.POINTER(SrtSceneInfo)]),
    "srt_trace_closest": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "srt_primary_rays": (C.c_int, [C.c_void_p, C.POINTER(SrtRenderParams), C.c_void_p]),
    "srt_light_pdf": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "srt_render": (C.c_int, [C.c_void_p, C.POINTER(SrtRenderParams), C.c_void_p]),
    "srt_render_fixed": (C.c_int, [C.c_void_p, C.POINTER(SrtRenderParams), C.c_void_p]),
    "srt_resolve_fixed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_int]),
    "srt_exchange_create": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "srt_exchange_handle": (C.c_int, [C.c_void_p, C.c_void_p]),
    "srt_exchange_connect": (C.c_int, [C.c_void_p, C.c_void_p]),
    "srt_exchange_fixed": (C.c_void_p, [C.c_void_p]),
    "srt_exchange_image": (C.c_void_p, [C.c_void_p]),
    "srt_exchange_run": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p]),
    "srt_exchange_destroy": (C.c_int, [C.c_void_p]),
    "srt_resolve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    "srt_render_image": (C.c_int, [C.c_void_p, C.POINTER(SrtRenderParams), C.c_void_p, C.c_void_p]),
    "srt_render_image_static": (C.c_int, [C.c_void_p, C.POINTER(SrtRenderParams), C.c_void_p]),
    "srt_render_counted": (C.c_int, [C.c_void_p, C.POINTER(SrtRenderParams), C.c_void_p, C.POINTER(SrtCounters)]),
    "srt_counters_get": (C.c_int, [C.c_void_p, C.POINTER(SrtCounters)]),
    "srt_sobol_points_u32": (C.c_int, [C.c_uint32, C.c_void_p]),
    "srt_ref_bvh_leaf_order": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_uint64), C.c_void_p]),
    "srt_ref_tree_boxes": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int, C.c_void_p, C.POINTER(C.c_int32)]),
    "srt_merl_lookup": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "srt_measure_fp32_peak": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "srt_measure_l2_peak": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "srt_selftest_division": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_uint64)]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


class BackendError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"srt error {code}: {msg}")
        self.code = code


def load_backend(path: Optional[str] = None):
    """Loads libsrt_b200.so (built in-tree by __graft_entry__.build()).  Raises if it is missing:
    the product has no CPU path to fall back to."""
    global _backend
    if _backend is not None and path is None:
        return _backend
    p = path or os.environ.get("SRT_B200_LIB", BACKEND_PATH)
    if not os.path.exists(p):
        raise FileNotFoundError(
            f"{p} not found: the CUDA backend is not built (run `python -c 'import __graft_entry__ as g; "
            f"g.build()'`); there is no CPU fallback")
    lib = C.CDLL(p, mode=C.RTLD_GLOBAL)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if lib.srt_abi_version() != SRT_ABI_VERSION:
        raise RuntimeError("libsrt_b200.so ABI version mismatch")
    if path is None:
        _backend = lib
    return lib


def check(lib, rc):
    if rc != 0:
        raise BackendError(rc, (lib.srt_last_error() or b"").decode(errors="replace"))
