"""Python mirror of the reference's scene-authoring API — the host side of the drop-in boundary
for Python callers.

Same class names, constructor arguments and meaning as the reference's header-defined classes
(sphere.h, moving_sphere.h, aarect.h, box.h, triangle.h, hitable.h, hitable_list.h, bvh.h,
constant_medium.h, env_sphere.h, material.h, texture.h, camera.h), so a scene function of
Raytracing_n.cpp:108-711 transliterates line by line.  Objects only record what the reference
objects store (constructor-time arithmetic is reproduced in float32/float64 exactly as the C++
evaluates it); all per-ray work happens in the CUDA backend.  `flatten(world, hlist, cam)` emits
the arrays of SrtSceneDesc with the rules of host/srt_flatten.hpp, array for array, so a scene
built here and the same scene built from real reference objects flatten to identical bytes
(tests/test_flatten.py).

`drand48()` is the reference's global LCG (mathf.h:12-19): scene functions that place objects at
random (Raytracing_n.cpp:108-147, :475-533) and the bvh_node constructor's split-axis draws
(bvh.h:97) consume it in the same order as the reference.
"""
from __future__ import annotations

import ctypes
import ctypes.util
import math
import os
import tempfile
from typing import List, Optional, Sequence

import numpy as np

from . import ffi

f32 = np.float32
_libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
for _n in ("tanf", "logf", "sinf", "cosf"):
    getattr(_libm, _n).restype = ctypes.c_float
    getattr(_libm, _n).argtypes = [ctypes.c_float]


def _tanf(x):
    return f32(_libm.tanf(float(x)))


def _logf(x):
    return f32(_libm.logf(float(x)))


# ---- mathf.h:12-19 ------------------------------------------------------------------------------
_seed = 1


def set_seed(s: int):
    global _seed
    _seed = int(s) & 0xFFFFFFFFFFFF


def get_seed() -> int:
    return _seed


def drand48() -> float:
    """mathf.h:14-19: 48-bit LCG, returns (state >> 16) / 2^32 as a double."""
    global _seed
    _seed = (0x5DEECE66D * _seed + 0xB16) & 0xFFFFFFFFFFFF
    return float(_seed >> 16) / 4294967296.0


# ---- vec3.h ---------------------------------------------------------------------------------------
def vec3(*a) -> np.ndarray:
    if len(a) == 1:
        if np.ndim(a[0]) == 0:
            return np.full(3, f32(a[0]), f32)  # vec3(float e0), vec3.h:15
        return np.asarray(a[0], f32).reshape(3).copy()
    return np.array([f32(a[0]), f32(a[1]), f32(a[2])], f32)


def _length(v) -> np.float32:  # vec3.h:38-41
    return np.sqrt(f32(f32(v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]))


def unit_vector(v) -> np.ndarray:  # vec3.h:170-173
    l = _length(v)
    return np.array([v[0] / l, v[1] / l, v[2] / l], f32)


def cross(a, b) -> np.ndarray:  # vec3.h:112-117
    return np.array([f32(a[1] * b[2]) - f32(a[2] * b[1]),
                     f32(a[2] * b[0]) - f32(a[0] * b[2]),
                     f32(a[0] * b[1]) - f32(a[1] * b[0])], f32)


# ---- texture.h ------------------------------------------------------------------------------------
class texture:
    pass


class constant_texture(texture):  # texture.h:25-33
    def __init__(self, c):
        self.color = vec3(c)


class checker_texture(texture):  # texture.h:9-23
    def __init__(self, t0: texture, t1: texture):
        self.even, self.odd = t0, t1


class noise_texture(texture):  # texture.h:35-46
    def __init__(self, sc):
        self.scale = f32(sc)


class image_texture(texture):  # texture.h:48-70
    def __init__(self, pixels, A: Optional[int] = None, B: Optional[int] = None):
        px = np.ascontiguousarray(pixels, np.uint8)
        if px.ndim == 1:
            px = px.reshape(B, A, 3)
        self.data = px
        self.ny, self.nx = px.shape[0], px.shape[1]


_perlin_tables = None


def perlin_tables():
    """perlin.h:67-97: the static tables, built from the first 1533 draws of the drand48 stream
    that starts at seed = 1 (static initialisation runs before main)."""
    global _perlin_tables
    if _perlin_tables is not None:
        return _perlin_tables
    keep = get_seed()
    set_seed(1)
    ran = np.zeros((256, 3), f32)
    for i in range(256):
        # perlin.h:70 `vec3(-1 + 2 * drand48(), -1 + 2 * drand48(), -1 + 2 * drand48())`: the order of the three
        # draws is unspecified in C++; g++ (the oracle) and MSVC (the reference's compiler) both evaluate
        # constructor arguments right to left, so z takes the first draw
        z = f32(-1 + 2 * drand48())
        y = f32(-1 + 2 * drand48())
        x = f32(-1 + 2 * drand48())
        ran[i] = unit_vector(np.array([x, y, z], f32))
    perms = []
    for _ in range(3):
        p = list(range(256))
        for i in range(255, 0, -1):
            target = int(drand48() * (i + 1))
            p[i], p[target] = p[target], p[i]
        perms.append(p)
    set_seed(keep)
    _perlin_tables = (ran, np.asarray(perms, np.int32))
    return _perlin_tables


# ---- material.h -----------------------------------------------------------------------------------
class material:
    pass


class lambertian(material):  # material.h:95-114
    def __init__(self, a: texture):
        self.albedo = a


class orennayar(material):  # material.h:127-149
    def __init__(self, a: texture, sigma):
        self.albedo = a
        self.sigma = f32(sigma)
        s = f32(float(f32(self.sigma / f32(180))) * math.pi)  # sigma / 180 * M_PI, stored back to float
        ss = float(f32(s * s))
        self.A = f32(1 - 0.5 * float(s) * float(s) / (ss + 0.33))
        self.B = f32(0.45 * float(s) * float(s) / (ss + 0.09))


def roughness_to_alpha(roughness) -> np.float32:  # microfacet_distribution.h:139-144
    r = f32(max(float(f32(roughness)), float(f32(1e-3))))
    x = _logf(r)
    acc = f32(f32(1.62162) + f32(f32(0.819955) * x))
    acc = f32(acc + f32(f32(f32(0.1734) * x) * x))
    acc = f32(acc + f32(f32(f32(f32(0.0171201) * x) * x) * x))
    acc = f32(acc + f32(f32(f32(f32(f32(0.000640711) * x) * x) * x) * x))
    return acc


class beckmann(material):  # material.h:151-199
    def __init__(self, a: texture, roughx, roughy):
        self.albedo = a
        self.roughx, self.roughy = f32(roughx), f32(roughy)
        self.alphax, self.alphay = roughness_to_alpha(self.roughx), roughness_to_alpha(self.roughy)


class brdfmaterial(material):  # material.h:201-241
    """`table` is the MERL payload (brdf.h:171-184): 3*90*90*180 doubles, or a path to a .binary file."""

    def __init__(self, table, a):
        self.albedo = vec3(a)
        if isinstance(table, (str, os.PathLike)):
            with open(table, "rb") as fh:
                dims = np.frombuffer(fh.read(12), np.int32)
                if int(dims[0]) * int(dims[1]) * int(dims[2]) != ffi.MERL_DOUBLES // 3:
                    raise ValueError("Dimensions don't match")  # brdf.h:177-181
                self.table = np.frombuffer(fh.read(8 * ffi.MERL_DOUBLES), np.float64).copy()
        else:
            self.table = np.ascontiguousarray(table, np.float64).reshape(-1)
        if self.table.size != ffi.MERL_DOUBLES:
            raise ValueError("MERL table must hold 3*90*90*180 doubles")
        self._file = None

    def write_binary_file(self) -> str:
        if self._file is None:
            fd, path = tempfile.mkstemp(suffix=".binary", prefix="srt_merl_")
            with os.fdopen(fd, "wb") as fh:
                fh.write(np.array([90, 90, 180], np.int32).tobytes())
                fh.write(self.table.tobytes())
            self._file = path
        return self._file


class metal(material):  # material.h:243-261
    def __init__(self, a, f):
        self.albedo = vec3(a)
        self.fuzz_arg = f32(f)
        self.fuzz = self.fuzz_arg if self.fuzz_arg < 1 else f32(1)


class dielectric(material):  # material.h:282-324
    def __init__(self, ri):
        self.ref_idx = f32(ri)


class diffuse_light(material):  # material.h:341-356
    def __init__(self, a: texture):
        self.emit = a


class isotropic(material):  # material.h:359-369
    def __init__(self, a: texture):
        self.albedo = a


class FlattenError(ValueError):
    """An object graph shape the flattener / backend does not cover (SRT_E_UNSUPPORTED)."""


# ---- hitables -------------------------------------------------------------------------------------
class hitable:
    def box_min(self) -> np.ndarray:
        """bounding_box(0, 0, box).Min() — the bvh_node constructor's sort key (bvh.h:21-55)."""
        raise FlattenError(f"unsupported bvh_node leaf: {type(self).__name__}")

    def ref_box(self) -> np.ndarray:
        """bounding_box(0, 1, box) as [min xyz, max xyz] float32 — the leaf box the reference
        bvh_node unions into its node boxes (bvh.h:115-118)."""
        raise FlattenError(f"unsupported bvh_node leaf: {type(self).__name__}")


class sphere(hitable):  # sphere.h:17-30
    def __init__(self, cen, r, m: Optional[material]):
        self.center, self.radius, self.mat_ptr = vec3(cen), f32(r), m

    def box_min(self):  # sphere.h:32-35
        return self.center - self.radius

    def ref_box(self):
        return np.concatenate([self.center - self.radius, self.center + self.radius]).astype(f32)


class env_sphere(sphere):  # env_sphere.h:7-25
    pass


class moving_sphere(hitable):  # moving_sphere.h:4-17
    def __init__(self, cen0, cen1, t0, t1, r, m):
        self.center0, self.center1 = vec3(cen0), vec3(cen1)
        self.time0, self.time1, self.radius, self.mat_ptr = f32(t0), f32(t1), f32(r), m

    def center(self, time):  # moving_sphere.h:19-21
        k = f32(f32(f32(time) - self.time0) / f32(self.time1 - self.time0))
        return self.center0 + k * (self.center1 - self.center0)

    def box_min(self):  # moving_sphere.h:53-58 at t0 = t1 = 0
        return self.center(f32(0)) - self.radius

    def ref_box(self):  # moving_sphere.h:53-58 at (t0, t1) = (0, 1)
        c0, c1 = self.center(f32(0)), self.center(f32(1))
        lo = np.minimum(c0 - self.radius, c1 - self.radius)
        hi = np.maximum(c0 + self.radius, c1 + self.radius)
        return np.concatenate([lo, hi]).astype(f32)


def _pad(k, sign):  # aarect.h:12 `k - 0.0001` is evaluated in double and narrowed by vec3's constructor
    return f32(float(k) + sign * 0.0001)


class xy_rect(hitable):  # aarect.h:5-35
    def __init__(self, x0, x1, y0, y1, k, mat):
        self.x0, self.x1, self.y0, self.y1, self.k, self.mp = f32(x0), f32(x1), f32(y0), f32(y1), f32(k), mat

    def box_min(self):
        return np.array([self.x0, self.y0, _pad(self.k, -1)], f32)

    def ref_box(self):
        return np.array([self.x0, self.y0, _pad(self.k, -1), self.x1, self.y1, _pad(self.k, +1)], f32)

    def row(self):
        return [self.x0, self.x1, self.y0, self.y1, self.k]


class xz_rect(hitable):  # aarect.h:37-64
    def __init__(self, x0, x1, z0, z1, k, mat):
        self.x0, self.x1, self.z0, self.z1, self.k, self.mp = f32(x0), f32(x1), f32(z0), f32(z1), f32(k), mat

    def box_min(self):
        return np.array([self.x0, _pad(self.k, -1), self.z0], f32)

    def ref_box(self):
        return np.array([self.x0, _pad(self.k, -1), self.z0, self.x1, _pad(self.k, +1), self.z1], f32)

    def row(self):
        return [self.x0, self.x1, self.z0, self.z1, self.k]


class yz_rect(hitable):  # aarect.h:66-94
    def __init__(self, y0, y1, z0, z1, k, mat):
        self.y0, self.y1, self.z0, self.z1, self.k, self.mp = f32(y0), f32(y1), f32(z0), f32(z1), f32(k), mat

    def box_min(self):
        return np.array([_pad(self.k, -1), self.y0, self.z0], f32)

    def ref_box(self):
        return np.array([_pad(self.k, -1), self.y0, self.z0, _pad(self.k, +1), self.y1, self.z1], f32)

    def row(self):
        return [self.y0, self.y1, self.z0, self.z1, self.k]


class triangle(hitable):  # triangle.h:9-106
    """uvs / normals: optional (3, 2|3) arrays.  Without normals the reference's 4- and 7-argument
    constructors leave n0..n2 uninitialised (triangle.h:13-36); this mirror (like the oracle
    harness, through the public members) sets them to the unit face normal."""

    def __init__(self, p0, p1, p2, mat, uv0=None, uv1=None, uv2=None, n0=None, n1=None, n2=None):
        self.p0, self.p1, self.p2, self.mp = vec3(p0), vec3(p1), vec3(p2), mat
        self.uvs = None if uv0 is None else np.stack([vec3(_uv3(uv0)), vec3(_uv3(uv1)), vec3(_uv3(uv2))])
        self.normal = unit_vector(cross(self.p1 - self.p0, self.p2 - self.p0))  # triangle.h:21-22
        self.normals = None if n0 is None else np.stack([vec3(n0), vec3(n1), vec3(n2)])

    def shading_normals(self):
        return self.normals if self.normals is not None else np.stack([self.normal] * 3)

    def box_min(self):  # triangle.h:53-68
        return np.minimum(np.minimum(self.p0, self.p1), self.p2)

    def ref_box(self):
        return np.concatenate([np.minimum(np.minimum(self.p0, self.p1), self.p2),
                               np.maximum(np.maximum(self.p0, self.p1), self.p2)]).astype(f32)


def _uv3(uv):
    uv = list(np.asarray(uv, f32).reshape(-1))
    return uv + [f32(0)] * (3 - len(uv))


class flip_normals(hitable):  # aarect.h:149-171
    def __init__(self, p: hitable):
        self.ptr = p

    def box_min(self):
        return self.ptr.box_min()

    def ref_box(self):
        return self.ptr.ref_box()


class box(hitable):  # box.h:5-33
    def __init__(self, p0, p1, ptr: material):
        self.pmin, self.pmax, self.mat = vec3(p0), vec3(p1), ptr
        a, b = self.pmin, self.pmax
        self.list_ptr = hitable_list([
            xy_rect(a[0], b[0], a[1], b[1], b[2], ptr),
            flip_normals(xy_rect(a[0], b[0], a[1], b[1], a[2], ptr)),
            xz_rect(a[0], b[0], a[2], b[2], b[1], ptr),
            flip_normals(xz_rect(a[0], b[0], a[2], b[2], a[1], ptr)),
            yz_rect(a[1], b[1], a[2], b[2], b[0], ptr),
            flip_normals(yz_rect(a[1], b[1], a[2], b[2], a[0], ptr)),
        ], 6)

    def box_min(self):  # box.h:10-13
        return self.pmin

    def ref_box(self):
        return np.concatenate([self.pmin, self.pmax]).astype(f32)


class translate(hitable):  # hitable.h:35-61
    def __init__(self, p: hitable, displacement):
        self.ptr, self.offset = p, vec3(displacement)


class rotate_y(hitable):  # hitable.h:65-107
    def __init__(self, p: hitable, angle):
        self.ptr, self.angle = p, f32(angle)
        radians = float(f32((math.pi / 180.0) * float(self.angle)))  # `float radians = (M_PI / 180.) * angle;`
        self.sin_theta, self.cos_theta = f32(math.sin(radians)), f32(math.cos(radians))


class rotate_x(hitable):  # hitable.h:135-178
    def __init__(self, p: hitable, angle):
        self.ptr, self.angle = p, f32(angle)
        radians = float(f32((math.pi / 180.0) * float(self.angle)))  # `float radians = (M_PI / 180.) * angle;`
        self.sin_theta, self.cos_theta = f32(math.sin(radians)), f32(math.cos(radians))


class constant_medium(hitable):  # constant_medium.h:4-17
    def __init__(self, b: hitable, d, a: texture):
        self.boundary, self.density, self.albedo = b, f32(d), a
        self.phase_function = isotropic(a)


class hitable_list(hitable):  # hitable_list.h:7-19
    def __init__(self, l: Sequence[hitable], n: Optional[int] = None):
        self.list = list(l)
        self.list_size = len(self.list) if n is None else int(n)

    def box_min(self):  # hitable_list.h:35-52 unions list[0] with itself list_size times
        return self.list[0].box_min()

    def ref_box(self):
        return self.list[0].ref_box()


def _ref_sorted_order(keys: np.ndarray) -> np.ndarray:
    """Order glibc's qsort (merge sort; on `cmp > 0` the right run's element goes first) produces
    with the reference's comparators, which return -1 for a < b and +1 otherwise, never 0
    (bvh.h:21-55): ascending keys, equal keys in REVERSE input order."""
    n = len(keys)
    rev = np.argsort(keys[::-1], kind="stable")
    return (n - 1 - rev).astype(np.int64)


def _ref_bvh_leaf_order(mins: np.ndarray) -> np.ndarray:
    """In-order leaf sequence of `new bvh_node(l, n, t0, t1)` (bvh.h:96-119): at every node draw
    axis = int(3*drand48()), qsort the node's slice by box-min on that axis, split at n/2."""
    if len(mins) >= 4096:
        # large meshes: the same replay in the backend's host code (srt_ref_bvh_leaf_order, multi-threaded);
        # tests/test_flatten.py checks both forms against the live reference
        lib = ffi.load_backend()
        m = np.ascontiguousarray(mins, f32)
        out = np.empty(len(m), np.int32)
        sd = ctypes.c_uint64(get_seed())
        ffi.check(lib, lib.srt_ref_bvh_leaf_order(m.ctypes.data, len(m), ctypes.byref(sd), out.ctypes.data))
        set_seed(int(sd.value))
        return out.astype(np.int64)
    order = np.arange(len(mins), dtype=np.int64)

    def build(lo: int, n: int):
        axis = int(3 * drand48())
        sl = order[lo:lo + n]
        order[lo:lo + n] = sl[_ref_sorted_order(mins[sl, axis])]
        if n > 2:
            build(lo, n // 2)
            build(lo + n // 2, n - n // 2)

    import sys
    if len(mins):
        old = sys.getrecursionlimit()
        sys.setrecursionlimit(max(old, 10000))
        try:
            build(0, len(mins))
        finally:
            sys.setrecursionlimit(old)
    return order


class bvh_node(hitable):  # bvh.h:8-19, :96-119
    """Records the items and reproduces the reference constructor's effect on (a) the global
    drand48 stream and (b) the in-order leaf sequence, which fixes the reference tree, its boxes
    and its tie rule; the backend builds its own SAH wide BVH over the same leaves."""

    def __init__(self, l: Sequence[hitable], n: Optional[int] = None, time0=0.0, time1=1.0):
        n = len(l) if n is None else int(n)
        self.items = list(l)[:n]
        self.time0, self.time1 = f32(time0), f32(time1)
        self.seed_at_build = get_seed()
        mins = np.stack([it.box_min() for it in self.items]).astype(f32) if n else np.zeros((0, 3), f32)
        self.leaf_order = _ref_bvh_leaf_order(mins)

    def leaves(self) -> List[hitable]:
        return [self.items[i] for i in self.leaf_order]


_TEAPOT_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "utah_teapot.npz")


def teapot_triangles(scale, divs: int = 100) -> np.ndarray:
    """teapot.h:76-166 createPloyTeapot: 32 bicubic Bezier patches x divs x divs quads x 2 triangles,
    returned as [n][9] float32 in construction order.  Arithmetic follows teapot.h:18-37."""
    z = np.load(_TEAPOT_DATA)
    verts, patches = z["vertices"].astype(f32), z["patches"].astype(np.int64)
    scale = f32(scale)
    t = (np.arange(divs + 1, dtype=f32) / f32(divs)).astype(f32)  # (float)i / (float)divs
    one = f32(1)
    omt = (one - t).astype(f32)
    b0 = (omt * omt * omt).astype(f32)
    b1 = (f32(3) * t * omt * omt).astype(f32)
    b2 = (f32(3) * t * t * omt).astype(f32)
    b3 = (t * t * t).astype(f32)
    B = np.stack([b0, b1, b2, b3])  # [4][divs+1]
    out = []
    j, i = np.meshgrid(np.arange(divs), np.arange(divs), indexing="ij")
    v0 = ((divs + 1) * j + i).reshape(-1)
    v1 = v0 + 1
    v2 = ((divs + 1) * (j + 1) + i + 1).reshape(-1)
    v3 = v2 - 1
    for np_ in range(patches.shape[0]):
        cp = (verts[patches[np_] - 1] * scale).astype(f32)  # [16][3]
        # uCurve[i] = evalBezierCurve(cp + 4 i, u): p0*b0 + p1*b1 + p2*b2 + p3*b3, left to right
        ucurve = []
        for r in range(4):
            acc = (cp[4 * r + 0][None, :] * B[0][:, None]).astype(f32)
            for k in (1, 2, 3):
                acc = (acc + (cp[4 * r + k][None, :] * B[k][:, None]).astype(f32)).astype(f32)
            ucurve.append(acc)  # [divs+1 (u)][3]
        # P(u, v) = evalBezierCurve(uCurve, v)
        P = None
        for k in range(4):
            term = (ucurve[k][None, :, :] * B[k][:, None, None]).astype(f32)  # [v][u][3]
            P = term if P is None else (P + term).astype(f32)
        P = P.reshape(-1, 3)  # k = j*(divs+1) + i with j <-> v, i <-> u
        tri = np.empty((divs * divs, 2, 9), f32)
        tri[:, 0, 0:3], tri[:, 0, 3:6], tri[:, 0, 6:9] = P[v0], P[v1], P[v2]
        tri[:, 1, 0:3], tri[:, 1, 3:6], tri[:, 1, 6:9] = P[v0], P[v2], P[v3]
        out.append(tri.reshape(-1, 9))
    return np.concatenate(out)


class triangle_soup(hitable):
    """A bvh_node over many triangles kept as arrays (the per-object form would need one Python
    object per triangle): what `new bvh_node(mesh_triangles, n, 0, 1)` is in the reference
    (Raytracing_n.cpp:286, :349).  pos [n][9]; uv [n][6] / nrm [n][9] optional."""

    def __init__(self, pos, mat, uv=None, nrm=None):
        self.pos = np.ascontiguousarray(pos, f32).reshape(-1, 9)
        self.mat = mat
        n = len(self.pos)
        self.uv = np.zeros((n, 6), f32) if uv is None else np.ascontiguousarray(uv, f32).reshape(n, 6)
        if nrm is None:  # unit face normal, triangle.h:21-22
            e1 = (self.pos[:, 3:6] - self.pos[:, 0:3]).astype(f32)
            e2 = (self.pos[:, 6:9] - self.pos[:, 0:3]).astype(f32)
            c = np.stack([(e1[:, 1] * e2[:, 2]).astype(f32) - (e1[:, 2] * e2[:, 1]).astype(f32),
                          (e1[:, 2] * e2[:, 0]).astype(f32) - (e1[:, 0] * e2[:, 2]).astype(f32),
                          (e1[:, 0] * e2[:, 1]).astype(f32) - (e1[:, 1] * e2[:, 0]).astype(f32)], 1).astype(f32)
            l = np.sqrt(((c[:, 0] * c[:, 0]).astype(f32) + (c[:, 1] * c[:, 1]).astype(f32)).astype(f32)
                        + (c[:, 2] * c[:, 2]).astype(f32)).astype(f32)
            with np.errstate(invalid="ignore", divide="ignore"):
                nn = (c / l[:, None]).astype(f32)
            self.nrm = np.concatenate([nn, nn, nn], 1)
        else:
            self.nrm = np.ascontiguousarray(nrm, f32).reshape(n, 9)
        self.seed_at_build = get_seed()
        mins = np.minimum(np.minimum(self.pos[:, 0:3], self.pos[:, 3:6]), self.pos[:, 6:9])
        self.leaf_order = _ref_bvh_leaf_order(mins)


class teapot_bvh(triangle_soup):
    """`new bvh_node(teapot(scale, mat).createPloyTeapot(), 640000, 0, 1)` (Raytracing_n.cpp:348-350)."""

    def __init__(self, scale, mat, divs: int = 100):
        self.scale, self.divs = f32(scale), int(divs)
        super().__init__(teapot_triangles(scale, divs), mat)


# ---- camera.h -------------------------------------------------------------------------------------
class camera:  # camera.h:33-48
    def __init__(self, lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist, t0=0.0, t1=0.0):
        self.lookfrom, self.lookat, self.vup = vec3(lookfrom), vec3(lookat), vec3(vup)
        self.vfov, self.aspect, self.aperture, self.focus_dist = f32(vfov), f32(aspect), f32(aperture), f32(focus_dist)
        self.t0, self.t1 = f32(t0), f32(t1)
        self.time0, self.time1 = self.t0, self.t1
        self.lens_radius = f32(self.aperture / f32(2))
        theta = f32(float(self.vfov) * math.pi / 180)
        half_height = _tanf(f32(theta / f32(2)))
        half_width = f32(self.aspect * half_height)
        fd = self.focus_dist
        self.origin = self.lookfrom.copy()
        self.w = unit_vector(self.lookfrom - self.lookat)
        self.u = unit_vector(cross(self.vup, self.w))
        self.v = cross(self.w, self.u)
        hw, hh = f32(half_width * fd), f32(half_height * fd)
        self.lower_left_corner = (((self.origin - hw * self.u).astype(f32) - hh * self.v).astype(f32)
                                  - fd * self.w).astype(f32)
        self.horizontal = (f32(f32(f32(2) * half_width) * fd) * self.u).astype(f32)
        self.vertical = (f32(f32(f32(2) * half_height) * fd) * self.v).astype(f32)

    def record(self) -> np.ndarray:
        c = np.zeros((), ffi.CAMERA_DT)
        for k in ("origin", "lower_left_corner", "horizontal", "vertical", "u", "v", "w"):
            c[k] = getattr(self, k)
        c["time0"], c["time1"], c["lens_radius"] = self.time0, self.time1, self.lens_radius
        return c


# ---- flattening (same rules, same emission order as host/srt_flatten.hpp) ---------------------------

class _Flattener:
    def __init__(self):
        self.entries, self.ops, self.bvhs, self.media = [], [], [], []
        self.lights, self.materials, self.textures = [], [], []
        self.spheres, self.msph, self.rects = [], [], []
        # prims / groups / bvh items / triangles are kept as chunks (numpy blocks) with running
        # counts, so that array-form meshes (triangle_soup) are appended in one piece
        self.prim_chunks, self.group_chunks, self.item_chunks, self.item_box_chunks = [], [], [], []
        self.n_prims = self.n_groups = self.n_items = self.n_tris = 0
        self.tri_pos, self.tri_nrm, self.tri_uv = [], [], []
        self.images, self.merl = [], []
        self.mat_index, self.tex_index, self.img_index, self.merl_index = {}, {}, {}, {}
        self.use_perlin = False

    # textures / materials: index in order of first encounter, children before parents
    def add_texture(self, t):
        if t is None:
            return -1
        if id(t) in self.tex_index:
            return self.tex_index[id(t)]
        if isinstance(t, constant_texture):
            row = (ffi.TEX_CONSTANT, -1, -1, tuple(t.color))
        elif isinstance(t, checker_texture):
            a = self.add_texture(t.even)
            b = self.add_texture(t.odd)
            row = (ffi.TEX_CHECKER, a, b, (0.0, 0.0, 0.0))
        elif isinstance(t, noise_texture):
            row = (ffi.TEX_NOISE, -1, -1, (t.scale, 0.0, 0.0))
            self.use_perlin = True
        elif isinstance(t, image_texture):
            key = id(t.data)
            if key not in self.img_index:
                self.img_index[key] = len(self.images)
                self.images.append(t.data)
            row = (ffi.TEX_IMAGE, self.img_index[key], -1, (0.0, 0.0, 0.0))
        else:
            raise FlattenError("unknown texture class")
        self.textures.append(row)
        self.tex_index[id(t)] = len(self.textures) - 1
        return len(self.textures) - 1

    def add_material(self, m):
        if m is None:
            return -1
        if id(m) in self.mat_index:
            return self.mat_index[id(m)]
        p = [0.0] * 6
        tex = -1
        if isinstance(m, lambertian):
            ty, tex = ffi.MAT_LAMBERTIAN, self.add_texture(m.albedo)
        elif isinstance(m, orennayar):
            ty, tex = ffi.MAT_ORENNAYAR, self.add_texture(m.albedo)
            p[0], p[1] = m.A, m.B
        elif isinstance(m, beckmann):
            ty, tex = ffi.MAT_BECKMANN, self.add_texture(m.albedo)
            p[0], p[1] = m.alphax, m.alphay
        elif isinstance(m, brdfmaterial):
            ty = ffi.MAT_BRDF
            p[0:3] = list(m.albedo)
            key = id(m.table)
            if key not in self.merl_index:
                self.merl_index[key] = len(self.merl)
                self.merl.append(m.table)
            p[3] = float(self.merl_index[key])
        elif isinstance(m, metal):
            ty = ffi.MAT_METAL
            p[0:3] = list(m.albedo)
            p[3] = m.fuzz
        elif isinstance(m, dielectric):
            ty = ffi.MAT_DIELECTRIC
            p[0] = m.ref_idx
        elif isinstance(m, diffuse_light):
            ty, tex = ffi.MAT_DIFFUSE_LIGHT, self.add_texture(m.emit)
        elif isinstance(m, isotropic):
            ty, tex = ffi.MAT_ISOTROPIC, self.add_texture(m.albedo)
        else:
            raise FlattenError("unknown material class")
        self.materials.append((ty, tex, tuple(p)))
        self.mat_index[id(m)] = len(self.materials) - 1
        return len(self.materials) - 1

    def add_prim_data(self, h):
        """-> (type, index, material object) or None if `h` is not a leaf primitive."""
        if isinstance(h, env_sphere):
            self.spheres.append([*h.center, h.radius])
            return ffi.PRIM_ENV_SPHERE, len(self.spheres) - 1, h.mat_ptr
        if isinstance(h, sphere):
            self.spheres.append([*h.center, h.radius])
            return ffi.PRIM_SPHERE, len(self.spheres) - 1, h.mat_ptr
        if isinstance(h, moving_sphere):
            self.msph.append([*h.center0, *h.center1, h.time0, h.time1, h.radius])
            return ffi.PRIM_MOVING_SPHERE, len(self.msph) - 1, h.mat_ptr
        for cls, ty in ((xy_rect, ffi.PRIM_XY_RECT), (xz_rect, ffi.PRIM_XZ_RECT), (yz_rect, ffi.PRIM_YZ_RECT)):
            if isinstance(h, cls):
                self.rects.append(h.row())
                return ty, len(self.rects) - 1, h.mp
        if isinstance(h, triangle):
            self.tri_pos.append(np.concatenate([h.p0, h.p1, h.p2])[None, :])
            self.tri_nrm.append(h.shading_normals().reshape(1, 9))
            uv = np.zeros((1, 6), f32) if h.uvs is None else h.uvs[:, :2].reshape(1, 6)
            self.tri_uv.append(uv)
            self.n_tris += 1
            return ffi.PRIM_TRIANGLE, self.n_tris - 1, h.mp
        return None

    def add_leaf_prims(self, h, flip):
        if isinstance(h, flip_normals):
            return self.add_leaf_prims(h.ptr, not flip)
        if isinstance(h, box):
            return self.add_leaf_prims(h.list_ptr, flip)
        if isinstance(h, hitable_list):
            for x in h.list[:h.list_size]:
                self.add_leaf_prims(x, flip)
            return
        r = self.add_prim_data(h)
        if r is None:
            raise FlattenError(f"unsupported leaf hitable: {type(h).__name__}")
        ty, idx, mat = r
        self.prim_chunks.append(np.array([(ty, idx, self.add_material(mat), ffi.PRIM_FLIP if flip else 0)],
                                         ffi.PRIM_DT))
        self.n_prims += 1

    def add_group(self, h, flip):
        first = self.n_prims
        self.add_leaf_prims(h, flip)
        self.group_chunks.append(np.array([(first, self.n_prims - first)], ffi.GROUP_DT))
        self.n_groups += 1
        return self.n_groups - 1

    def add_soup(self, h):
        """A triangle_soup is n single-triangle groups in the reference tree's leaf order."""
        n = len(h.pos)
        order = h.leaf_order
        mat = self.add_material(h.mat)
        pr = np.zeros(n, ffi.PRIM_DT)
        pr["type"], pr["index"], pr["material"] = ffi.PRIM_TRIANGLE, np.arange(self.n_tris, self.n_tris + n), mat
        gr = np.zeros(n, ffi.GROUP_DT)
        gr["first_prim"], gr["prim_count"] = np.arange(self.n_prims, self.n_prims + n), 1
        self.item_chunks.append(np.arange(self.n_groups, self.n_groups + n, dtype=np.int32))
        ps = h.pos[order]
        self.item_box_chunks.append(np.concatenate(
            [np.minimum(np.minimum(ps[:, 0:3], ps[:, 3:6]), ps[:, 6:9]),
             np.maximum(np.maximum(ps[:, 0:3], ps[:, 3:6]), ps[:, 6:9])], 1).astype(f32))
        self.prim_chunks.append(pr)
        self.group_chunks.append(gr)
        self.tri_pos.append(h.pos[order])
        self.tri_nrm.append(h.nrm[order])
        self.tri_uv.append(h.uv[order])
        self.n_tris += n
        self.n_prims += n
        self.n_groups += n
        self.n_items += n
        return n

    @staticmethod
    def is_groupable(h):
        while isinstance(h, flip_normals):
            h = h.ptr
        return not isinstance(h, (bvh_node, triangle_soup, translate, rotate_y, rotate_x, constant_medium))

    def flatten_entry(self, h, chain, medium):
        chain = list(chain)
        while True:
            if isinstance(h, translate):
                chain.append((ffi.OP_TRANSLATE, h.offset[0], h.offset[1], h.offset[2]))
                h = h.ptr
            elif isinstance(h, rotate_y):
                chain.append((ffi.OP_ROTATE_Y, h.sin_theta, h.cos_theta, 0.0))
                h = h.ptr
            elif isinstance(h, rotate_x):
                chain.append((ffi.OP_ROTATE_X, h.sin_theta, h.cos_theta, 0.0))
                h = h.ptr
            elif isinstance(h, flip_normals):
                chain.append((ffi.OP_FLIP, 0.0, 0.0, 0.0))
                h = h.ptr
            elif isinstance(h, constant_medium):
                if chain or medium >= 0:
                    raise FlattenError("constant_medium below a transform or another medium is not supported")
                self.media.append((h.density, self.add_material(h.phase_function)))
                medium = len(self.media) - 1
                h = h.boundary
            else:
                break
        if isinstance(h, (bvh_node, triangle_soup)):
            first = self.n_items
            if isinstance(h, triangle_soup):
                count = self.add_soup(h)
            else:
                leaves = h.leaves()
                for leaf in leaves:
                    if not self.is_groupable(leaf):
                        raise FlattenError(f"unsupported bvh_node leaf: {type(leaf).__name__}")
                    self.item_chunks.append(np.array([self.add_group(leaf, False)], np.int32))
                    self.item_box_chunks.append(leaf.ref_box().reshape(1, 6))
                    self.n_items += 1
                count = len(leaves)
            self.bvhs.append((first, count))
            self.entries.append((len(self.ops), len(chain), ffi.GEOM_BVH, len(self.bvhs) - 1, medium, 0))
            self.ops.extend(chain)
            return
        if isinstance(h, hitable_list):
            if not all(self.is_groupable(x) for x in h.list[:h.list_size]):
                if medium >= 0:
                    raise FlattenError("constant_medium over a heterogeneous list is not supported")
                for x in h.list[:h.list_size]:
                    self.flatten_entry(x, chain, -1)
                return
        flip = False
        kept = []
        for op in chain:
            if op[0] == ffi.OP_FLIP:
                flip = not flip
            else:
                kept.append(op)
        g = self.add_group(h, flip)
        self.entries.append((len(self.ops), len(kept), ffi.GEOM_GROUP, g, medium, 0))
        self.ops.extend(kept)

    def flatten_light(self, h):
        while isinstance(h, flip_normals):
            h = h.ptr
        if isinstance(h, (xz_rect, sphere, triangle)):  # env_sphere is a sphere subclass here
            ty, idx, _ = self.add_prim_data(h)
            self.lights.append((ty, idx))
        elif isinstance(h, hitable_list):
            raise FlattenError("nested hitable_list in the light list is not supported")
        else:
            self.lights.append((ffi.PRIM_NONE, -1))


def _cat(chunks, dt):
    return np.concatenate(chunks).astype(dt) if chunks else np.zeros(0, dt)


def flatten(world: hitable, hlist: Optional[hitable], cam: camera) -> ffi.FlatScene:
    """(world, light list, camera) -> FlatScene; raises FlattenError for unsupported graph shapes."""
    fl = _Flattener()
    if isinstance(world, hitable_list):
        for x in world.list[:world.list_size]:
            fl.flatten_entry(x, [], -1)
    else:
        fl.flatten_entry(world, [], -1)
    if hlist is not None:
        if not isinstance(hlist, hitable_list):
            raise FlattenError("light_shape must be a hitable_list")  # Raytracing_n.cpp:74
        for x in hlist.list[:hlist.list_size]:
            fl.flatten_light(x)

    def rows(lst, dt):
        return np.array(lst, dt) if lst else np.zeros(0, dt)

    fs = ffi.FlatScene()
    fs.entries = rows(fl.entries, ffi.ENTRY_DT)
    fs.ops = rows(fl.ops, ffi.OP_DT)
    fs.prims = _cat(fl.prim_chunks, ffi.PRIM_DT)
    fs.groups = _cat(fl.group_chunks, ffi.GROUP_DT)
    fs.bvh_items = _cat(fl.item_chunks, np.int32)
    fs.bvh_item_boxes = (np.concatenate(fl.item_box_chunks).astype(f32) if fl.item_box_chunks
                         else np.zeros((0, 6), f32))
    fs.bvhs = rows(fl.bvhs, ffi.BVH_DT)
    fs.media = rows(fl.media, ffi.MEDIUM_DT)
    fs.spheres = np.array(fl.spheres, f32).reshape(-1, 4)
    fs.msph = np.array(fl.msph, f32).reshape(-1, 9)
    fs.rects = np.array(fl.rects, f32).reshape(-1, 5)
    if fl.tri_pos:
        fs.tri_pos = np.concatenate(fl.tri_pos).astype(f32)
        fs.tri_nrm = np.concatenate(fl.tri_nrm).astype(f32)
        fs.tri_uv = np.concatenate(fl.tri_uv).astype(f32)
    fs.lights = rows(fl.lights, ffi.LIGHT_DT)
    fs.materials = rows(fl.materials, ffi.MATERIAL_DT)
    fs.textures = rows(fl.textures, ffi.TEXTURE_DT)
    fs.images = list(fl.images)
    fs.merl = list(fl.merl)
    if fl.use_perlin:
        fs.perlin_ranvec, fs.perlin_perm = perlin_tables()
    fs.camera = cam.record()
    return fs.normalize()
