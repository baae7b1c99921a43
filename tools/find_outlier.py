#!/usr/bin/env python
"""Debug helper for the GPU box: finds the Sobol index that puts the largest single-sample value into a
pixel of a golden-size render and replays that one path with the counting build's bounce trace.
    tools/find_outlier.py <scene> <pixel> [spp]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import GOLDEN, get_scene  # noqa: E402
from simple_raytracing_render_b200 import render  # noqa: E402

name, pixel = sys.argv[1], int(sys.argv[2])
g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
nx, ny = int(g["img_nx"]), int(g["img_ny"])
ns = int(sys.argv[3]) if len(sys.argv) > 3 else 4 * int(g["img_ns"])
ctx = render.Context(0)
sc = render.Scene.from_objects(ctx, *get_scene(name))


def pix(a, b):
    buf = torch.zeros(nx * ny * 3, dtype=torch.float32, device="cuda")
    sc.render(render.render_params(nx, ny, ns, a, b), buf)
    ctx.synchronize()
    return buf.view(-1, 3)[pixel].cpu().numpy()


lo, hi = 0, ns
while hi - lo > 1:
    mid = (lo + hi) // 2
    a, b = pix(lo, mid), pix(mid, hi)
    lo, hi = (lo, mid) if a.max() / max(mid - lo, 1) >= b.max() / max(hi - mid, 1) else (mid, hi)
print("pixel", pixel, "sample", lo, "value", pix(lo, lo + 1))
os.environ["SRT_DEBUG_PIXEL"], os.environ["SRT_DEBUG_SAMPLE"] = str(pixel), str(lo)
buf = torch.zeros(nx * ny * 3, dtype=torch.float32, device="cuda")
sc.render_counted(render.render_params(nx, ny, ns, lo, lo + 1), buf)
ctx.synchronize()
