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
