"""The C-ABI library loads and exports every symbol include/srt_ffi.h declares; without a GPU its
compute entry points fail loudly (there is no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT
from simple_raytracing_render_b200 import ffi


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "srt_ffi.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(srt_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(backend_lib):
    names = _declared_symbols()
    assert len(names) >= 16
    for n in names:
        assert hasattr(backend_lib, n), n
    assert sorted(ffi.EXPORTED_SYMBOLS) == names


def test_abi_version(backend_lib):
    assert backend_lib.srt_abi_version() == ffi.SRT_ABI_VERSION


def test_struct_sizes_match_c_layout():
    # field-for-field mirrors of the C structs (x86-64 SysV)
    assert ffi.PRIM_DT.itemsize == 16 and ffi.ENTRY_DT.itemsize == 24 and ffi.OP_DT.itemsize == 16
    assert ffi.MATERIAL_DT.itemsize == 32 and ffi.TEXTURE_DT.itemsize == 24 and ffi.CAMERA_DT.itemsize == 96
    assert C.sizeof(ffi.SrtCamera) == 96 and C.sizeof(ffi.SrtRenderParams) == 32 and C.sizeof(ffi.SrtCounters) == 72
    assert C.sizeof(ffi.SrtSceneInfo) == 88


def test_no_cpu_fallback_without_gpu(backend_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = backend_lib.srt_context_create(0, None, C.byref(h))
    assert rc == ffi.SRT_E_NODEVICE
    assert b"no CPU fallback" in backend_lib.srt_last_error()


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "simple_raytracing_render_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "from oracle" not in text and "import oracle" not in text and "libsrt_ref" not in text, f


def test_header_constants_match_the_python_mirror():
    src = open(os.path.join(ROOT, "include", "srt_ffi.h")).read()

    def macro(name):
        m = re.search(rf"#define\s+{name}\s+(.+)", src)
        assert m, name
        return eval(m.group(1).split("/*")[0].strip().rstrip("u"), {"__builtins__": {}})

    assert macro("SRT_STATIC_MAX_PIXELS") == ffi.STATIC_MAX_PIXELS
    assert macro("SRT_STATIC_MAX_SAMPLES") == ffi.STATIC_MAX_SAMPLES
    assert macro("SRT_EXCHANGE_HANDLE_BYTES") == ffi.EXCHANGE_HANDLE_BYTES
    assert macro("SRT_RENDER_CORRECTED") == ffi.RENDER_CORRECTED


def test_static_entry_point_allocates_nothing():
    """srt_render_image_static promises no cudaMalloc / cudaFree inside the call: the functions it is made of
    (launch_render_static, static_scratch in srt_api.cu) name neither, and its scratch is a static device array
    whose address comes from cudaGetSymbolAddress."""
    src = open(os.path.join(ROOT, "simple_raytracing_render_b200", "csrc", "srt_api.cu")).read()
    a = src.index("static __device__ __align__(256) long long g_static_fixed")
    b = src.index("}  // namespace", a)
    body = re.sub(r"//.*", "", src[a:b])
    assert "cudaGetSymbolAddress((void**)&fixed, g_static_fixed)" in body
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "std::vector", "new "):
        assert banned not in body, banned
    entry = src[src.index("int srt_render_image_static("):]
    entry = entry[:entry.index("\n}\n")]
    assert "launch_render_static" in entry and "cudaMalloc" not in entry
