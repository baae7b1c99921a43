"""Host-side mirror of the reference's render entry for Python callers, on top of the C ABI
(include/srt_ffi.h, loaded by ffi.load_backend()).

`CRender(maxDepth, width, height, num_ray, thread_count, filePath)` + `Run()` is the class the
reference declares in Raytracing_n/Render.h:16-51 (and never defines); `Context` / `Scene` expose
the finer-grained entry points (closest-hit, light pdf, sample-range render into a device
accumulation buffer, resolve) that the parity tests, the benchmark and the multi-GPU driver use.

There is no CPU path here: every call goes through libsrt_b200.so and fails loudly without it or
without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
import time
from typing import Optional

import numpy as np

from . import ffi
from . import scene as S


def _ptr(x) -> Optional[int]:
    """Address of a numpy array, a torch tensor (host or device), or a raw integer address."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if isinstance(x, np.ndarray):
        if not x.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        if hasattr(x, "is_contiguous") and not x.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return int(x.data_ptr())
    raise TypeError(type(x))


class Context:
    """One per GPU / rank (srt_context_create).  `stream` is a cudaStream_t address or None."""

    def __init__(self, device: int = 0, stream: Optional[int] = None):
        self.lib = ffi.load_backend()
        h = C.c_void_p()
        ffi.check(self.lib, self.lib.srt_context_create(int(device), stream, C.byref(h)))
        self.h = h
        self.device = int(device)

    def synchronize(self):
        ffi.check(self.lib, self.lib.srt_context_synchronize(self.h))

    def counters(self) -> dict:
        c = ffi.SrtCounters()
        ffi.check(self.lib, self.lib.srt_counters_get(self.h, C.byref(c)))
        return c.as_dict()

    def measure_fp32_peak(self) -> float:
        """TFLOP/s of an FMA-chain kernel on this GPU (srt_measure_fp32_peak)."""
        v = C.c_double(0.0)
        ffi.check(self.lib, self.lib.srt_measure_fp32_peak(self.h, C.byref(v)))
        return float(v.value)

    def measure_l2_peak(self) -> float:
        """GB/s of an L2-resident streaming read on this GPU (srt_measure_l2_peak)."""
        v = C.c_double(0.0)
        ffi.check(self.lib, self.lib.srt_measure_l2_peak(self.h, C.byref(v)))
        return float(v.value)

    def ref_tree_boxes(self, item_boxes, on_device: bool):
        """The reference bvh_node tree over leaves in in-order sequence as 12-word records (srt_ref_tree_boxes)."""
        import numpy as np
        boxes = np.ascontiguousarray(item_boxes, dtype=np.float32).reshape(-1, 6)
        out = np.zeros((2 * len(boxes), 12), dtype=np.int32)  # a tree over n leaves has fewer than 2 n nodes
        count = C.c_int32(0)
        ffi.check(self.lib, self.lib.srt_ref_tree_boxes(self.h, boxes.ctypes.data, len(boxes), 1 if on_device else 0,
                                                        out.ctypes.data, C.byref(count)))
        return out[:count.value]

    def resolve(self, accum_dev, n_pixels: int, n_samples: int, rgb8_out):
        ffi.check(self.lib, self.lib.srt_resolve(self.h, _ptr(accum_dev), int(n_pixels), int(n_samples), _ptr(rgb8_out)))

    def resolve_fixed(self, fixed_dev, n_pixels: int, n_samples: int, rgb8_out=None, accum_out_dev=None, clear=True):
        """int64 fixed-point sums (srt_render_fixed, all shards added) -> 8-bit image (srt_resolve_fixed)."""
        ffi.check(self.lib, self.lib.srt_resolve_fixed(self.h, _ptr(fixed_dev), int(n_pixels), int(n_samples),
                                                       _ptr(rgb8_out), _ptr(accum_out_dev), 1 if clear else 0))

    def close(self):
        if self.h:
            self.lib.srt_context_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def render_params(nx, ny, ns, s_begin=0, s_end=None, max_depth=50, seed=0x5EED, flags=None) -> ffi.SrtRenderParams:
    p = ffi.SrtRenderParams()
    p.nx, p.ny, p.ns = int(nx), int(ny), int(ns)
    p.s_begin, p.s_end = int(s_begin), int(ns if s_end is None else s_end)
    p.max_depth, p.seed = int(max_depth), int(seed) & 0xFFFFFFFF
    # Raytracing_n.cpp:827 (`% ny`) is only meaningful for square images
    p.flags = (ffi.COMPAT_PIXEL_MOD_NY if nx == ny else 0) if flags is None else int(flags)
    return p


class Scene:
    """A flattened scene uploaded to one GPU (srt_scene_create)."""

    def __init__(self, ctx: Context, flat: ffi.FlatScene):
        self.ctx, self.lib = ctx, ctx.lib
        desc, keep = flat.to_desc()
        h = C.c_void_p()
        ffi.check(self.lib, self.lib.srt_scene_create(ctx.h, C.byref(desc), C.byref(h)))
        del keep  # the backend copied everything
        self.h = h
        self.n_prims = len(flat.prims)

    @classmethod
    def from_objects(cls, ctx: Context, world, hlist, cam) -> "Scene":
        return cls(ctx, S.flatten(world, hlist, cam))

    @classmethod
    def from_desc_address(cls, ctx: Context, addr: int) -> "Scene":
        """From a `const SrtSceneDesc*` produced elsewhere (the C++ flattener, host/srt_flatten.hpp)."""
        self = cls.__new__(cls)
        self.ctx, self.lib = ctx, ctx.lib
        h = C.c_void_p()
        ffi.check(self.lib, self.lib.srt_scene_create(ctx.h, C.cast(addr, C.POINTER(ffi.SrtSceneDesc)), C.byref(h)))
        self.h = h
        self.n_prims = self.info()["n_prims"]
        return self

    def info(self) -> dict:
        i = ffi.SrtSceneInfo()
        ffi.check(self.lib, self.lib.srt_scene_info(self.h, C.byref(i)))
        return i.as_dict()

    def trace_closest(self, rays: np.ndarray, want_rec=True):
        """world->hit(r, 0.001, FLT_MAX, rec) for rays [n][7] -> (prim ids, t, rec [n][8] | None)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 7)
        n = len(rays)
        ids, t = np.zeros(n, np.int32), np.zeros(n, np.float32)
        rec = np.zeros((n, 8), np.float32) if want_rec else None
        ffi.check(self.lib, self.lib.srt_trace_closest(self.h, rays.ctypes.data, n, ids.ctypes.data, t.ctypes.data,
                                                       rec.ctypes.data if want_rec else None))
        return ids, t, rec

    def primary_rays(self, params: ffi.SrtRenderParams) -> np.ndarray:
        """The camera rays the render kernels start their paths with: [ny*nx*(s_end-s_begin)][7] in
        (curpixel, s) order (Raytracing_n.cpp:827-836, camera.h:51-59)."""
        n = params.nx * params.ny * (params.s_end - params.s_begin)
        rays = np.zeros((n, 7), np.float32)
        ffi.check(self.lib, self.lib.srt_primary_rays(self.h, C.byref(params), rays.ctypes.data))
        return rays

    def light_pdf(self, od: np.ndarray) -> np.ndarray:
        od = np.ascontiguousarray(od, np.float32).reshape(-1, 6)
        out = np.zeros(len(od), np.float32)
        ffi.check(self.lib, self.lib.srt_light_pdf(self.h, od.ctypes.data, len(od), out.ctypes.data))
        return out

    def merl_lookup(self, table: int, angles: np.ndarray) -> np.ndarray:
        angles = np.ascontiguousarray(angles, np.float64).reshape(-1, 4)
        out = np.zeros((len(angles), 3), np.float64)
        ffi.check(self.lib, self.lib.srt_merl_lookup(self.h, int(table), angles.ctypes.data, len(angles), out.ctypes.data))
        return out

    def render(self, params: ffi.SrtRenderParams, accum_dev):
        """ADD the radiance sums of Sobol indices [s_begin, s_end) into a DEVICE buffer (async)."""
        ffi.check(self.lib, self.lib.srt_render(self.h, C.byref(params), _ptr(accum_dev)))

    def render_fixed(self, params: ffi.SrtRenderParams, fixed_dev):
        """ADD the 2^-31 fixed-point radiance sums of Sobol indices [s_begin, s_end) into a DEVICE int64
        buffer (async): the shard form, reduced across GPUs with an exact integer sum."""
        ffi.check(self.lib, self.lib.srt_render_fixed(self.h, C.byref(params), _ptr(fixed_dev)))

    def render_counted(self, params: ffi.SrtRenderParams, accum_dev) -> dict:
        c = ffi.SrtCounters()
        ffi.check(self.lib, self.lib.srt_render_counted(self.h, C.byref(params), _ptr(accum_dev), C.byref(c)))
        return c.as_dict()

    def render_image(self, params: ffi.SrtRenderParams, want_accum=False, rgb8: Optional[np.ndarray] = None,
                     accum: Optional[np.ndarray] = None):
        """Whole image with HOST buffers -> rgb8 [ny*nx][3] (and the linear sums [ny*nx][3])."""
        n = params.nx * params.ny
        if rgb8 is None:
            rgb8 = np.zeros((n, 3), np.uint8)
        if want_accum and accum is None:
            accum = np.zeros((n, 3), np.float32)
        ffi.check(self.lib, self.lib.srt_render_image(self.h, C.byref(params), _ptr(rgb8), _ptr(accum)))
        return (rgb8, accum) if want_accum else rgb8

    def render_image_static(self, params: ffi.SrtRenderParams, rgb8_dev):
        """srt_render_image_static: the frame into a DEVICE uint8 buffer of nx*ny*3 bytes (async), on the library's
        static scratch — no allocation inside the call; at most ffi.STATIC_MAX_PIXELS pixels and
        ffi.STATIC_MAX_SAMPLES samples.  No counterpart in the reference."""
        ffi.check(self.lib, self.lib.srt_render_image_static(self.h, C.byref(params), _ptr(rgb8_dev)))

    def close(self):
        if getattr(self, "h", None):
            self.lib.srt_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def write_ppm_p3(path: str, rgb8: np.ndarray, nx: int, ny: int):
    """The reference's output format: "P3\\nnx ny\\n255\\n" + one "r g b" line per pixel in curpixel
    order (Raytracing_n.cpp:886, :873-876)."""
    rgb8 = np.asarray(rgb8, np.uint8).reshape(-1, 3)
    with open(path, "w") as fh:
        fh.write(f"P3\n{nx} {ny}\n255\n")
        fh.write("\n".join(f"{r} {g} {b}" for r, g, b in rgb8.tolist()))
        fh.write("\n")


def write_ppm_p6(path: str, rgb8: np.ndarray, nx: int, ny: int):
    """Binary PPM: the same pixels as the P3 dump (Raytracing_n.cpp:873-876) in 3 bytes per pixel instead of
    ~12 characters."""
    rgb8 = np.ascontiguousarray(np.asarray(rgb8, np.uint8).reshape(-1))
    if rgb8.size != nx * ny * 3:
        raise ValueError("image size does not match nx * ny * 3")
    with open(path, "wb") as fh:
        fh.write(f"P6\n{nx} {ny}\n255\n".encode())
        fh.write(rgb8.tobytes())


def write_png(path: str, rgb8: np.ndarray, nx: int, ny: int):
    """8-bit RGB PNG (one IDAT, filter type 0 on every row, zlib level 6)."""
    import struct
    import zlib
    rows = np.ascontiguousarray(np.asarray(rgb8, np.uint8).reshape(ny, nx * 3))
    raw = np.concatenate([np.zeros((ny, 1), np.uint8), rows], axis=1).tobytes()

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as fh:
        fh.write(b"\x89PNG\r\n\x1a\n")
        fh.write(chunk(b"IHDR", struct.pack(">IIBBBBB", nx, ny, 8, 2, 0, 0, 0)))
        fh.write(chunk(b"IDAT", zlib.compress(raw, 6)))
        fh.write(chunk(b"IEND", b""))


def write_image(path: str, rgb8: np.ndarray, nx: int, ny: int):
    """By extension: .png -> PNG, .pnm / .p6 -> binary PPM, anything else (the reference writes `.ppm`) -> the
    reference's own P3 text format."""
    ext = os.path.splitext(path)[1].lower()
    if ext == ".png":
        write_png(path, rgb8, nx, ny)
    elif ext in (".pnm", ".p6"):
        write_ppm_p6(path, rgb8, nx, ny)
    else:
        write_ppm_p3(path, rgb8, nx, ny)


class CRender:
    """Render.h:16-51.  `Run()` renders the registered scene on the GPU and writes the P3 file the
    reference's main() writes; `thread_count` only sized the reference's CPU pool and is ignored."""

    def __init__(self, maxDepth: int = 50, width: int = 1000, height: int = 1000, num_ray: int = 50,
                 thread_count: int = 8, filePath: Optional[str] = None, device: int = 0):
        self.maxDepth, self.nx, self.ny, self.ns = maxDepth, width, height, num_ray
        self.thread_count, self.filePath, self.device = thread_count, filePath, device
        self.finishedPixel, self.isfinished = 0, False
        self._scene = None
        self.colors = None
        self.elapsed_ms = None

    def set_scene(self, world, hlist, cam):
        """The (world, light list, camera) triple every reference scene function returns
        (e.g. Raytracing_n.cpp:216).  Render.h gives Run() no way to receive it."""
        self._scene = (world, hlist, cam)

    def Run(self):
        if self._scene is None:
            raise RuntimeError("CRender.Run: no scene registered (set_scene)")
        t0 = time.time()
        ctx = Context(self.device)
        try:
            sc = Scene.from_objects(ctx, *self._scene)
            rgb8 = sc.render_image(render_params(self.nx, self.ny, self.ns, max_depth=self.maxDepth))
            sc.close()
        finally:
            ctx.close()
        self.colors = rgb8
        if self.filePath:
            write_image(self.filePath, rgb8, self.nx, self.ny)  # P3 text for `.ppm`, as main() writes it
        self.finishedPixel, self.isfinished = self.nx * self.ny, True
        self.elapsed_ms = int((time.time() - t0) * 1000)
        print(f"{self.elapsed_ms}ms")  # Raytracing_n.cpp:944-946
        return rgb8
