#!/usr/bin/env python
"""GPU box: throughput of the stand-alone closest-hit kernel (srt_trace_closest, device pointers) on
the reference camera's primary rays of a config — how fast traversal alone runs when it does not share
the instruction cache with shading.   tools/trace_bench.py <scene> [spp]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import get_scene  # noqa: E402
from oracle import ref  # noqa: E402
from simple_raytracing_render_b200 import ffi, render, scenes  # noqa: E402

name = sys.argv[1]
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 4
nx, ny, _ = scenes.CONFIG_SIZES[name]
triple = get_scene(name)
rs = ref.RefBuilder().realize(*triple)
rays = rs.primary_rays(nx, ny, spp, seed=3, mod_ny=nx == ny)
ctx = render.Context(0)
sc = render.Scene.from_objects(ctx, *triple)
ids, t, rec = sc.trace_closest(rays[:200000])
hit = ids >= 0
# second-bounce-like rays: from the hit points towards the first light (un-normalised)
n = len(rays)
d_rays = torch.from_numpy(rays).cuda()
d_ids = torch.empty(n, dtype=torch.int32, device="cuda")
d_t = torch.empty(n, dtype=torch.float32, device="cuda")
for label, buf in (("primary", d_rays),):
    for _ in range(2):
        ffi.check(sc.lib, sc.lib.srt_trace_closest(sc.h, buf.data_ptr(), n, d_ids.data_ptr(), d_t.data_ptr(), None))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    reps = 5
    for _ in range(reps):
        ffi.check(sc.lib, sc.lib.srt_trace_closest(sc.h, buf.data_ptr(), n, d_ids.data_ptr(), d_t.data_ptr(), None))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{name} {label}: {n} rays in {ms:.3f} ms = {n / ms / 1e3:.1f} Mrays/s (hit fraction {hit.mean():.2f})")
