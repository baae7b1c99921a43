"""ctypes wrapper of oracle/_ref/libsrt_ref.so — the UNMODIFIED reference compiled as the oracle.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs, never by the product package.  The library is built by `make -C oracle ref` in a container
that has /root/reference; on the GPU box only the prebuilt file is present (it travels with the
repo snapshot).  `available()` says whether it can be loaded.

`RefBuilder.realize()` constructs REAL reference objects (one extern "C" constructor per reference
class, oracle/ref_harness.cpp) from a scene written with the product's Python mirror of the
reference scene API (simple_raytracing_render_b200/scene.py), so the same scene source can be
run through the reference and through the CUDA backend.
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libsrt_ref.so")
REFERENCE_ROOT = "/root/reference"

_lib = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def has_reference_tree() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "Raytracing_n"))


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not available():
        raise FileNotFoundError(f"{LIB_PATH} not built (make -C oracle ref needs /root/reference)")
    # the reference opens "..\results\20210709_balls.ppm" relative to the cwd at static-init time
    # (Raytracing_n.cpp:45); load from a scratch cwd so the stray file never lands in the repo
    cwd = os.getcwd()
    scratch = tempfile.mkdtemp(prefix="srt_ref_")
    try:
        os.chdir(scratch)
        L = C.CDLL(LIB_PATH)
    finally:
        os.chdir(cwd)
    f32p = C.POINTER(C.c_float)
    i32p = C.POINTER(C.c_int32)
    sig = {
        "ref_last_error": (C.c_char_p, []),
        "ref_has_reference_tree": (C.c_int, []),
        "ref_set_sobol_file": (None, [C.c_char_p]),
        "ref_set_seed": (None, [C.c_uint64]),
        "ref_get_seed": (C.c_uint64, []),
        "ref_reset_pcg": (None, []),
        "ref_probe_beckmann_stash": (C.c_float, [C.c_int]),
        "ref_merl_lookup": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_void_p]),
        "ref_kat_drand48": (None, [C.c_uint64, C.c_int, C.c_void_p]),
        "ref_kat_pcg32": (None, [C.c_int, C.c_void_p, C.c_void_p]),
        "ref_kat_sobol": (C.c_int, [C.c_uint, C.c_void_p]),
        "ref_sobol_fallback": (None, [C.c_uint, C.c_void_p]),
        "ref_kat_erf": (C.c_float, [C.c_float]),
        "ref_kat_erfinv": (C.c_float, [C.c_float]),
        "ref_kat_schlick": (C.c_float, [C.c_float, C.c_float]),
        "ref_kat_roughness_to_alpha": (C.c_float, [C.c_float]),
        "ref_tex_constant": (C.c_int, [C.c_float] * 3),
        "ref_tex_checker": (C.c_int, [C.c_int, C.c_int]),
        "ref_tex_noise": (C.c_int, [C.c_float]),
        "ref_tex_image": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
        "ref_mat_lambertian": (C.c_int, [C.c_int]),
        "ref_mat_orennayar": (C.c_int, [C.c_int, C.c_float]),
        "ref_mat_beckmann": (C.c_int, [C.c_int, C.c_float, C.c_float]),
        "ref_mat_metal": (C.c_int, [C.c_float] * 4),
        "ref_mat_dielectric": (C.c_int, [C.c_float]),
        "ref_mat_diffuse_light": (C.c_int, [C.c_int]),
        "ref_mat_isotropic": (C.c_int, [C.c_int]),
        "ref_mat_brdf": (C.c_int, [C.c_char_p, C.c_float, C.c_float, C.c_float]),
        "ref_sphere": (C.c_int, [C.c_float] * 4 + [C.c_int]),
        "ref_env_sphere": (C.c_int, [C.c_float] * 4 + [C.c_int]),
        "ref_moving_sphere": (C.c_int, [f32p, f32p, C.c_float, C.c_float, C.c_float, C.c_int]),
        "ref_xy_rect": (C.c_int, [C.c_float] * 5 + [C.c_int]),
        "ref_xz_rect": (C.c_int, [C.c_float] * 5 + [C.c_int]),
        "ref_yz_rect": (C.c_int, [C.c_float] * 5 + [C.c_int]),
        "ref_triangle": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
        "ref_box": (C.c_int, [f32p, f32p, C.c_int]),
        "ref_flip_normals": (C.c_int, [C.c_int]),
        "ref_translate": (C.c_int, [C.c_int, C.c_float, C.c_float, C.c_float]),
        "ref_rotate_y": (C.c_int, [C.c_int, C.c_float]),
        "ref_rotate_x": (C.c_int, [C.c_int, C.c_float]),
        "ref_constant_medium": (C.c_int, [C.c_int, C.c_float, C.c_int]),
        "ref_hitable_list": (C.c_int, [i32p, C.c_int, C.c_int]),
        "ref_bvh_node": (C.c_int, [i32p, C.c_int, C.c_float, C.c_float]),
        "ref_teapot_bvh": (C.c_int, [C.c_float, C.c_int, i32p]),
        "ref_teapot_triangles": (C.c_int, [C.c_float, C.c_void_p, C.c_int]),
        "ref_camera": (C.c_int, [f32p, f32p, f32p] + [C.c_float] * 6),
        "ref_scene": (C.c_int, [C.c_int, C.c_int, C.c_int]),
        "ref_scene_named": (C.c_int, [C.c_char_p, C.c_float]),
        "ref_scene_desc": (C.c_void_p, [C.c_int]),
        "ref_primary_rays": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p]),
        "ref_trace": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
        "ref_light_pdf": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_void_p]),
        "ref_light_random": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_uint64, C.c_void_p]),
        "ref_triangle_soup_bvh": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int]),
        "ref_render_float": (C.c_int, [C.c_int] * 7 + [C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]),
        "ref_render_timed": (C.c_double, [C.c_int] * 6 + [C.c_uint64, C.c_void_p]),
        "ref_crender_run": (C.c_int, [C.c_int] * 5 + [C.c_char_p, C.c_char_p]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _err():
    return (lib().ref_last_error() or b"").decode(errors="replace")


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


# ---- known-answer values ------------------------------------------------------------------------
def kat_drand48(seed: int, n: int) -> np.ndarray:
    out = np.zeros(n, np.float64)
    lib().ref_kat_drand48(seed, n, out.ctypes.data)
    return out


def kat_pcg32(n: int):
    f, u = np.zeros(n, np.float32), np.zeros(n, np.uint32)
    lib().ref_kat_pcg32(n, f.ctypes.data, u.ctypes.data)
    return f, u


def kat_sobol(n: int) -> np.ndarray:
    out = np.zeros((n, 2), np.float64)
    if lib().ref_kat_sobol(n, out.ctypes.data) != 0:
        raise RuntimeError(_err())
    return out


def sobol_fallback(n: int) -> np.ndarray:
    out = np.zeros((n, 2), np.float64)
    lib().ref_sobol_fallback(n, out.ctypes.data)
    return out


class RefScene:
    """A (world, light list, camera) triple of real reference objects."""

    def __init__(self, handle: int):
        if handle < 0:
            raise RuntimeError(_err())
        self.h = handle

    def desc_address(self) -> int:
        """Address of the SrtSceneDesc produced by the product's C++ flattener (srt_flatten.hpp)."""
        p = lib().ref_scene_desc(self.h)
        if not p:
            raise RuntimeError("flatten failed: " + _err())
        return p

    def flat(self):
        from simple_raytracing_render_b200.ffi import FlatScene
        return FlatScene.from_desc(self.desc_address())

    def primary_rays(self, nx, ny, ns, seed=1, mod_ny=False) -> np.ndarray:
        rays = np.zeros((nx * ny * ns, 7), np.float32)
        if lib().ref_primary_rays(self.h, nx, ny, ns, seed, int(mod_ny), rays.ctypes.data) != 0:
            raise RuntimeError(_err())
        return rays

    def trace(self, rays: np.ndarray, seed=1, want_rec=True):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 7)
        n = len(rays)
        ids, t = np.zeros(n, np.int32), np.zeros(n, np.float32)
        rec = np.zeros((n, 8), np.float32) if want_rec else None
        rc = lib().ref_trace(self.h, rays.ctypes.data, n, seed, ids.ctypes.data, t.ctypes.data,
                             rec.ctypes.data if want_rec else None)
        if rc != 0:
            raise RuntimeError(_err())
        return ids, t, rec

    def light_pdf(self, od: np.ndarray) -> np.ndarray:
        od = np.ascontiguousarray(od, np.float32).reshape(-1, 6)
        out = np.zeros(len(od), np.float32)
        if lib().ref_light_pdf(self.h, od.ctypes.data, len(od), out.ctypes.data) != 0:
            raise RuntimeError(_err())
        return out

    def light_random(self, o: np.ndarray, seed=1) -> np.ndarray:
        """hitable_list::random over the light list for origins o [n][3] (the reference's own sampler)."""
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3)
        out = np.zeros_like(o)
        if lib().ref_light_random(self.h, o.ctypes.data, len(o), seed, out.ctypes.data) != 0:
            raise RuntimeError(_err())
        return out

    def render_float(self, nx, ny, ns, s_begin=0, s_end=None, max_depth=50, seed=1, mod_ny=False):
        """Per-pixel radiance SUMS [ny*nx, 3] over Sobol indices [s_begin, s_end) and the ray count."""
        s_end = ns if s_end is None else s_end
        out = np.zeros((nx * ny, 3), np.float32)
        rays = C.c_uint64(0)
        rc = lib().ref_render_float(self.h, nx, ny, ns, s_begin, s_end, max_depth, seed, int(mod_ny),
                                    out.ctypes.data, C.byref(rays))
        if rc != 0:
            raise RuntimeError(_err())
        return out, int(rays.value)

    def render_timed(self, nx, ny, ns, threads, max_depth=50, seed=1, want_rgb=False):
        """Seconds the reference's own renderthread pool took (Raytracing_n.cpp:926-946)."""
        rgb = np.zeros((nx * ny, 3), np.int32) if want_rgb else None
        sec = lib().ref_render_timed(self.h, nx, ny, ns, max_depth, threads, seed,
                                     rgb.ctypes.data if want_rgb else None)
        if sec < 0:
            raise RuntimeError(_err())
        return (sec, rgb) if want_rgb else sec

    def crender_run(self, max_depth, width, height, num_ray, lib_path, file_path) -> int:
        rc = lib().ref_crender_run(self.h, max_depth, width, height, num_ray, lib_path.encode(), file_path.encode())
        if rc != 0:
            raise RuntimeError(f"CRender::Run failed ({rc}): {_err()}")
        return rc


def merl_lookup(mat, angles: np.ndarray) -> np.ndarray:
    """The reference's brdf::lookup_brdf_val on the table of a mirror `brdfmaterial` (scene.py)."""
    b = RefBuilder()
    h = b._mat(mat)
    angles = np.ascontiguousarray(angles, np.float64).reshape(-1, 4)
    out = np.zeros((len(angles), 3), np.float64)
    if lib().ref_merl_lookup(h, angles.ctypes.data, len(angles), out.ctypes.data) != 0:
        raise RuntimeError("not a brdfmaterial")
    return out


def named_scene(name: str, aspect: float) -> RefScene:
    """One of the reference's OWN scene functions (Raytracing_n.cpp:379 ball_scenes, :427
    ball_orennayar_scenes).  They stbi_load a Windows-literal relative path; run them from a scratch
    cwd holding a symlink of exactly that name to the reference's image."""
    src = os.path.join(REFERENCE_ROOT, "contents", "environment_map", "sky_2.png")
    if not os.path.exists(src):
        raise FileNotFoundError(src)
    cwd = os.getcwd()
    scratch = tempfile.mkdtemp(prefix="srt_ref_scene_")
    try:
        os.chdir(scratch)
        os.symlink(src, "..\\contents\\environment_map\\sky_2.png")
        return RefScene(lib().ref_scene_named(name.encode(), aspect))
    finally:
        os.chdir(cwd)


class RefBuilder:
    """Realises a scene written with simple_raytracing_render_b200.scene classes as reference objects."""

    def __init__(self):
        self.L = lib()
        self._memo = {}

    def _tex(self, t):
        from simple_raytracing_render_b200 import scene as S
        if t is None:
            return -1
        if id(t) in self._memo:
            return self._memo[id(t)]
        L = self.L
        if isinstance(t, S.constant_texture):
            h = L.ref_tex_constant(*[float(x) for x in t.color])
        elif isinstance(t, S.checker_texture):
            h = L.ref_tex_checker(self._tex(t.even), self._tex(t.odd))
        elif isinstance(t, S.noise_texture):
            h = L.ref_tex_noise(float(t.scale))
        elif isinstance(t, S.image_texture):
            px = np.ascontiguousarray(t.data, np.uint8)
            h = L.ref_tex_image(px.ctypes.data, t.nx, t.ny)
        else:
            raise TypeError(type(t))
        self._memo[id(t)] = h
        return h

    def _mat(self, m):
        from simple_raytracing_render_b200 import scene as S
        if m is None:
            return -1
        if id(m) in self._memo:
            return self._memo[id(m)]
        L = self.L
        if isinstance(m, S.lambertian):
            h = L.ref_mat_lambertian(self._tex(m.albedo))
        elif isinstance(m, S.orennayar):
            h = L.ref_mat_orennayar(self._tex(m.albedo), float(m.sigma))
        elif isinstance(m, S.beckmann):
            h = L.ref_mat_beckmann(self._tex(m.albedo), float(m.roughx), float(m.roughy))
        elif isinstance(m, S.metal):
            h = L.ref_mat_metal(*[float(x) for x in m.albedo], float(m.fuzz_arg))
        elif isinstance(m, S.dielectric):
            h = L.ref_mat_dielectric(float(m.ref_idx))
        elif isinstance(m, S.diffuse_light):
            h = L.ref_mat_diffuse_light(self._tex(m.emit))
        elif isinstance(m, S.isotropic):
            h = L.ref_mat_isotropic(self._tex(m.albedo))
        elif isinstance(m, S.brdfmaterial):
            path = m.write_binary_file()
            h = L.ref_mat_brdf(path.encode(), *[float(x) for x in m.albedo])
        else:
            raise TypeError(type(m))
        self._memo[id(m)] = h
        return h

    def _hit(self, o):
        from simple_raytracing_render_b200 import scene as S
        if o is None:
            return -1
        if id(o) in self._memo:
            return self._memo[id(o)]
        L = self.L
        if isinstance(o, S.env_sphere):
            h = L.ref_env_sphere(*[float(x) for x in o.center], float(o.radius), self._mat(o.mat_ptr))
        elif isinstance(o, S.sphere):
            h = L.ref_sphere(*[float(x) for x in o.center], float(o.radius), self._mat(o.mat_ptr))
        elif isinstance(o, S.moving_sphere):
            h = L.ref_moving_sphere(_f3(o.center0), _f3(o.center1), float(o.time0), float(o.time1),
                                    float(o.radius), self._mat(o.mat_ptr))
        elif isinstance(o, S.xy_rect):
            h = L.ref_xy_rect(float(o.x0), float(o.x1), float(o.y0), float(o.y1), float(o.k), self._mat(o.mp))
        elif isinstance(o, S.xz_rect):
            h = L.ref_xz_rect(float(o.x0), float(o.x1), float(o.z0), float(o.z1), float(o.k), self._mat(o.mp))
        elif isinstance(o, S.yz_rect):
            h = L.ref_yz_rect(float(o.y0), float(o.y1), float(o.z0), float(o.z1), float(o.k), self._mat(o.mp))
        elif isinstance(o, S.triangle):
            pos = np.ascontiguousarray(np.concatenate([o.p0, o.p1, o.p2]), np.float32)
            uv = None if o.uvs is None else np.ascontiguousarray(np.asarray(o.uvs)[:, :2].reshape(-1), np.float32)
            nrm = None if o.normals is None else np.ascontiguousarray(np.asarray(o.normals).reshape(-1), np.float32)
            h = L.ref_triangle(pos.ctypes.data, None if uv is None else uv.ctypes.data,
                               None if nrm is None else nrm.ctypes.data, self._mat(o.mp))
        elif isinstance(o, S.box):
            h = L.ref_box(_f3(o.pmin), _f3(o.pmax), self._mat(o.mat))
        elif isinstance(o, S.flip_normals):
            h = L.ref_flip_normals(self._hit(o.ptr))
        elif isinstance(o, S.translate):
            h = L.ref_translate(self._hit(o.ptr), *[float(x) for x in o.offset])
        elif isinstance(o, S.rotate_y):
            h = L.ref_rotate_y(self._hit(o.ptr), float(o.angle))
        elif isinstance(o, S.rotate_x):
            h = L.ref_rotate_x(self._hit(o.ptr), float(o.angle))
        elif isinstance(o, S.constant_medium):
            h = L.ref_constant_medium(self._hit(o.boundary), float(o.density), self._tex(o.albedo))
        elif isinstance(o, S.hitable_list):
            hs = (C.c_int32 * max(1, len(o.list)))(*[self._hit(x) for x in o.list])
            h = L.ref_hitable_list(hs, len(o.list), o.list_size)
        elif isinstance(o, S.teapot_bvh) and o.divs == 100:  # teapot.h:77 hard-codes divs = 100
            n = C.c_int32(0)
            L.ref_set_seed(int(o.seed_at_build))
            h = L.ref_teapot_bvh(float(o.scale), self._mat(o.mat), C.byref(n))
        elif isinstance(o, S.triangle_soup):
            # `new bvh_node(triangles, n, 0, 1)` over real triangle objects (Raytracing_n.cpp:286, :349)
            m = self._mat(o.mat)
            pos, uv, nrm = (np.ascontiguousarray(a, np.float32) for a in (o.pos, o.uv, o.nrm))
            L.ref_set_seed(int(o.seed_at_build))
            h = L.ref_triangle_soup_bvh(pos.ctypes.data, uv.ctypes.data, nrm.ctypes.data, len(pos), m)
        elif isinstance(o, S.bvh_node):
            hs = (C.c_int32 * max(1, len(o.items)))(*[self._hit(x) for x in o.items])
            # the reference's constructor draws its split axes from the global drand48 stream
            # (bvh.h:97): start it where the mirror's stream stood when it built this node
            L.ref_set_seed(int(o.seed_at_build))
            h = L.ref_bvh_node(hs, len(o.items), float(o.time0), float(o.time1))
        else:
            raise TypeError(type(o))
        if h < 0:
            raise RuntimeError(_err())
        self._memo[id(o)] = h
        return h

    def realize(self, world, hlist, cam) -> RefScene:
        w = self._hit(world)
        hl = self._hit(hlist) if hlist is not None else -1
        c = self.L.ref_camera(_f3(cam.lookfrom), _f3(cam.lookat), _f3(cam.vup), float(cam.vfov), float(cam.aspect),
                              float(cam.aperture), float(cam.focus_dist), float(cam.t0), float(cam.t1))
        return RefScene(self.L.ref_scene(w, hl, c))
