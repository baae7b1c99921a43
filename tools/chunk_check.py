#!/usr/bin/env python
"""GPU box: the same frame under different work decompositions (SRT_CHUNKS) and schedulers
(SRT_WAVEFRONT) must be the same image up to float summation order; prints kernel time and the
difference to the first setting.   tools/chunk_check.py <scene> nx ny ns  ENV=VAL[,ENV=VAL] ..."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import get_scene  # noqa: E402
from simple_raytracing_render_b200 import render  # noqa: E402

name, nx, ny, ns = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
ctx = render.Context(0)
sc = render.Scene.from_objects(ctx, *get_scene(name))
ref = None
for setting in sys.argv[5:]:
    envs = dict(kv.split("=") for kv in setting.split(",") if kv)
    for k, v in envs.items():
        os.environ[k] = v
    buf = torch.zeros(nx * ny * 3, dtype=torch.float32, device="cuda")
    for rep in range(2):
        buf.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sc.render(render.render_params(nx, ny, ns, flags=0), buf)
        ctx.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
    img = buf.cpu().numpy()
    for k in envs:
        os.environ.pop(k, None)
    if ref is None:
        ref = img
    close = np.isclose(img, ref, rtol=1e-3, atol=1e-4).mean()
    print(f"{name} {nx}x{ny}@{ns} [{setting or 'default'}]: {ms:8.2f} ms  variant {sc.info()['render_variant']}  "
          f"mean {img.mean():.6f}  close-to-first {close:.5f}")
