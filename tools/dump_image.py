#!/usr/bin/env python
"""Debug helper for the GPU box: renders a named test scene at the golden image size and writes the
linear float means to gpurun_out/<name>_img.npy.   tools/dump_image.py <name> [spp]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import GOLDEN, get_scene  # noqa: E402
from simple_raytracing_render_b200 import render  # noqa: E402

name = sys.argv[1]
g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
nx, ny = int(g["img_nx"]), int(g["img_ny"])
ns = int(sys.argv[2]) if len(sys.argv) > 2 else 4 * int(g["img_ns"])
ctx = render.Context(0)
sc = render.Scene.from_objects(ctx, *get_scene(name))
_, acc = sc.render_image(render.render_params(nx, ny, ns), want_accum=True)
x = acc / ns
np.save(os.path.join(ROOT, "gpurun_out", f"{name}_img.npy"), x)
r = g["img"]
d = x - r
print(name, "mean", x.mean(), r.mean(), "max", x.max(), "relmse", float(np.mean(d.astype(np.float64) ** 2 / (r.astype(np.float64) ** 2 + 1e-2))))
idx = np.argsort(-np.abs(d).max(axis=1))[:10]
for i in idx:
    print(" pixel", int(i), "(col", int(i % nx), "row", int(i // nx), ") gpu", x[i], "ref", r[i])
